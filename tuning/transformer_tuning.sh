#!/usr/bin/env bash
# Grid over learning rate x weight decay, 5 epochs each on the 10 % subset.
set -euo pipefail
cd "$(dirname "$0")"
NGPU=${NGPU:-$(python -c "import torch; print(max(torch.cuda.device_count(), 1))")}
for lr in 1e-5 5e-5 1e-4; do
  for wd in 0 5e-5 5e-4; do
    echo "lr=${lr} --weight_decay=${wd}"
    python -m torch.distributed.run --master-addr 127.0.0.1 --master-port 12355 --nnodes=1 --nproc-per-node=${NGPU} \
      ./transformer_tuning.py --workers 4 --batch_size 64 --distributed --mode ddp --ngd --epoch 5 --lr ${lr} --weight_decay ${wd} "$@"
  done
done
