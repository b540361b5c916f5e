#!/bin/bash
# Two tiny rounds of every algorithm on synthetic ABCD-shape volumes with the sm_100a engine (< 1 minute in total on one GPU).
# GPUS=N runs every algorithm client-parallel, one process per GPU (all nine support it); DEVICE=cpu MODEL=tiny3DCNN SHAPE=32,32,32
# gives a CPU smoke run (gloo).
export OMP_NUM_THREADS=1
GPUS=${GPUS:-1}
RUN="python -m"
if [ "$GPUS" -gt 1 ]; then
  RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$GPUS --master-addr 127.0.0.1 --master-port ${PORT:-29533} -m"
fi
COMMON="--model ${MODEL:-3DCNN} --dataset ABCD --in_shape ${SHAPE:-121,145,121} --client_num_in_total ${CLIENTS:-4} --synthetic_subjects $(( ${CLIENTS:-4} * 10 )) \
  --batch_size 8 --epochs 1 --comm_round 2 --lr 0.01 ${DEVICE:+--device $DEVICE}"
$RUN neuroimagedisttraining_b200.cli.main_sailentgrads $COMMON --frac 1 --dense_ratio 0.5 "$@" | tail -1
$RUN neuroimagedisttraining_b200.cli.main_fedavg $COMMON --frac 1 "$@" | tail -1
$RUN neuroimagedisttraining_b200.cli.main_dispfl $COMMON --dense_ratio 0.5 "$@" | tail -1
for ALGO in ditto dpsgd fedfomo local subavg turboaggregate; do
  $RUN neuroimagedisttraining_b200.cli.main_others $ALGO $COMMON "$@" | tail -1
done
