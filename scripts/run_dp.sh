#!/usr/bin/env bash
# Launch the data-parallel example on N ranks of this machine (the reference's run.sh spawned background
# python processes with a hand-written rank argument: lab/tutorial_1b/DP/gradient_aggr/run.sh:4-8).
set -euo pipefail
N="${1:-3}"; shift || true
cd "$(dirname "$0")/.."
SECONDS=0
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "${PORT:-29500}" \
  examples/dp_train.py "$@"
echo "elapsed: ${SECONDS}s"
