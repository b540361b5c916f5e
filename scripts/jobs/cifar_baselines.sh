#!/bin/bash
# The reference's CIFAR job family (DisPFL / Ditto / D-PSGD / FedFomo / Local / Sub-FedAvg):
# resnet18, 100 clients, frac 0.1, 500 rounds, lr 0.1, decay 0.998, 5 epochs, seed 2022.
export OMP_NUM_THREADS=1
COMMON="--model resnet18 --dataset ${DATASET:-cifar10} ${DATA_DIR:+--data_dir $DATA_DIR} --partition_method ${PART:-dir} --partition_alpha ${ALPHA:-0.3} \
  --client_num_in_total 100 --frac 0.1 --comm_round 500 --batch_size 128 --lr 0.1 --lr_decay 0.998 --epochs 5 --seed 2022"
python -m neuroimagedisttraining_b200.cli.main_dispfl $COMMON --dense_ratio 0.5 --cs random "$@"
for ALGO in ditto dpsgd fedfomo local subavg; do
  python -m neuroimagedisttraining_b200.cli.main_others $ALGO $COMMON "$@"
done
