#!/bin/bash
# Two tiny rounds of every algorithm on synthetic ABCD-shape volumes with the sm_100a engine (one GPU, < 1 minute in total).
export OMP_NUM_THREADS=1
COMMON="--model 3DCNN --dataset ABCD --client_num_in_total 4 --synthetic_subjects 40 --batch_size 8 --epochs 1 --comm_round 2 --lr 0.01"
python -m neuroimagedisttraining_b200.cli.main_sailentgrads $COMMON --frac 1 --dense_ratio 0.5 "$@" | tail -1
python -m neuroimagedisttraining_b200.cli.main_fedavg $COMMON --frac 1 "$@" | tail -1
python -m neuroimagedisttraining_b200.cli.main_dispfl $COMMON --dense_ratio 0.5 "$@" | tail -1
for ALGO in ditto dpsgd fedfomo local subavg turboaggregate; do
  python -m neuroimagedisttraining_b200.cli.main_others $ALGO $COMMON "$@" | tail -1
done
