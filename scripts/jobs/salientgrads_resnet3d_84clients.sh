#!/bin/bash
# BASELINE.json config 4: SalientGrads, 3-D ResNet-18, 90% sparsity (dense_ratio 0.1), 84 clients (rescale partition)
export OMP_NUM_THREADS=1
GPUS=${GPUS:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$GPUS" --master-addr 127.0.0.1 --master-port 29502 \
  -m neuroimagedisttraining_b200.cli.main_sailentgrads --model 3DResNet --dataset ABCD ${DATA_DIR:+--data_dir "$DATA_DIR"} \
  --rescale --client_num_in_total 84 --frac 1 --batch_size 8 --dense_ratio 0.1 --epochs 2 --comm_round 100 "$@"
