#!/usr/bin/env bash
# Grid over mixup alpha x StepLR gamma, 5 epochs each on the 10 % subset.
set -euo pipefail
cd "$(dirname "$0")"
NGPU=${NGPU:-$(python -c "import torch; print(max(torch.cuda.device_count(), 1))")}
for alpha in 0.2 0.5 0.99; do
  for gamma in 0.5 0.75 0.9; do
    echo "alpha=${alpha} gamma=${gamma}"
    python -m torch.distributed.run --master-addr 127.0.0.1 --master-port 12355 --nnodes=1 --nproc-per-node=${NGPU} \
      ./resnet50_tuning.py --workers 4 --bs 256 --distributed --epoch 5 --alpha ${alpha} --gamma ${gamma} "$@"
  done
done
