#!/bin/bash
# FedAvg on ABCD (reference: fedml_experiments/standalone/fedavg/*.sh; defaults lr 0.001, 5 epochs, frac 0.9)
export OMP_NUM_THREADS=1
GPUS=${GPUS:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$GPUS" --master-addr 127.0.0.1 --master-port 29501 \
  -m neuroimagedisttraining_b200.cli.main_fedavg --model 3DCNN --dataset ABCD ${DATA_DIR:+--data_dir "$DATA_DIR"} \
  --client_num_in_total 21 --frac 0.9 --batch_size 16 --lr 0.001 --epochs 5 --comm_round 200 --seed 0 "$@"
