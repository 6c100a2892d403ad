#!/usr/bin/env bash
# One node, one process per GPU (NGPU defaults to every visible GPU).  Same two jobs and flags as
# the reference's run_distributed.sh, but world-size agnostic and rendezvousing on 127.0.0.1.
set -euo pipefail
NGPU=${NGPU:-$(python -c "import torch; print(max(torch.cuda.device_count(), 1))")}
PORT=${MASTER_PORT:-12355}
export OMP_NUM_THREADS=${OMP_NUM_THREADS:-12}
RUN="python -m torch.distributed.run --master-addr 127.0.0.1 --master-port ${PORT} --nnodes=1 --nproc-per-node=${NGPU}"
${RUN} ./resnet50_test.py --workers 4 --bs 256 --distributed --meta_learning --ngd --lr 0.01 "$@"
${RUN} ./transformer_test.py --workers 4 --batch_size 64 --distributed --ngd "$@"
