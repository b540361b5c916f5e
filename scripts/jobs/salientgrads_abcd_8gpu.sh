#!/bin/bash
#SBATCH -N 1
#SBATCH --gres=gpu:8
#SBATCH -t 1-00:00
#SBATCH -J salientgrads_abcd
# SalientGrads on ABCD, 21 site clients, one process per GPU (counterpart of the reference's
# fedml_experiments/standalone/sailentgrads/Jobs/sailentgradsjob.sh, which ran 1x V100).
export OMP_NUM_THREADS=1
GPUS=${GPUS:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$GPUS" --master-addr 127.0.0.1 --master-port 29500 \
  -m neuroimagedisttraining_b200.cli.main_sailentgrads \
  --model 3DCNN --dataset ABCD ${DATA_DIR:+--data_dir "$DATA_DIR"} --client_num_in_total 21 --frac 1 \
  --batch_size 16 --lr 0.01 --lr_decay 0.998 --epochs 2 --comm_round 200 --dense_ratio 0.5 --seed 1024 \
  --checkpoint_path ckpt/salientgrads_abcd.pt --checkpoint_every 5 "$@"
