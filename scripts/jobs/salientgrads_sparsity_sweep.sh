#!/bin/bash
# Sparsity sweep dense_ratio in {0.5,0.3,0.2,0.1,0.05} and IterSNIP sweep {1,20,50,100}
# (the reference's sailentgrads/Jobs/salientgradssparsitywith*iteration*sps.sh family: CIFAR-10, resnet18,
# 100 clients, frac 0.1, 500 rounds, lr 0.1, decay 0.998, 5 epochs, seed 2022).
export OMP_NUM_THREADS=1
for DR in 0.5 0.3 0.2 0.1 0.05; do
  for IT in ${ITERS:-1 20 50 100}; do
    python -m neuroimagedisttraining_b200.cli.main_sailentgrads --model resnet18 --dataset cifar10 ${DATA_DIR:+--data_dir "$DATA_DIR"} \
      --partition_method dir --partition_alpha 0.3 --client_num_in_total 100 --frac 0.1 --comm_round 500 --batch_size 128 \
      --lr 0.1 --lr_decay 0.998 --epochs 5 --seed 2022 --dense_ratio "$DR" --itersnip_iteration "$IT" "$@"
  done
done
