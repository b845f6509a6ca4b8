#!/usr/bin/env bash
# Launch the pipeline-parallel example: N ranks = pipelines x stages (lab/tutorial_1b/PP/1F1B/run*.sh).
set -euo pipefail
N="${1:-3}"; shift || true
cd "$(dirname "$0")/.."
SECONDS=0
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "${PORT:-29501}" \
  examples/pp_train.py "$@"
echo "elapsed: ${SECONDS}s"
