#!/usr/bin/env bash
# Re-measure every number quoted in BASELINE.md sections 4 - 4c on this box.
#   bash scripts/run_all_benchmarks.sh [N_GPUS] [OUT_DIR]
# One rank per GPU; every script times on the device (CUDA events, max over ranks) and prints one JSON object.
set -u
N=${1:-1}
OUT=${2:-bench_out}
mkdir -p "$OUT"
cd "$(dirname "$0")/.."
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29700
run() {            # run <name> <command...>: stdout's JSON lines -> $OUT/<name>.json, stderr -> $OUT/<name>.err
  local name=$1; shift
  port=$((port + 1))
  echo "== $name"
  if "$@" 2> "$OUT/$name.err" | grep '^{' > "$OUT/$name.json"; then tail -c 400 "$OUT/$name.json"; echo; else echo "   FAILED (see $OUT/$name.err)"; fi
}
python -m ddl25spring_b200.build > "$OUT/build.log" 2>&1 || { echo "build failed"; exit 1; }

# single-GPU kernels and steps
run aggregation_1gpu   python benchmarks/bench_aggregation.py
run resnet_step        python benchmarks/bench_resnet_kernels.py
run llama_step         python benchmarks/bench_llama_step.py
run dcgan_step         python benchmarks/bench_dcgan_step.py
run bench_1gpu         python bench.py
run bench_reference    python bench.py --impl reference --steps 3 --warmup 1

if [ "$N" -gt 1 ]; then
  run bench_${N}gpu      $T --master-port $port bench.py --gpus "$N"
  run aggregation_${N}gpu $T --master-port $port benchmarks/bench_aggregation.py
  run dp_llama_${N}gpu   $T --master-port $port benchmarks/bench_dp_llama.py --arms fused_in_graph:32:16,fused_in_graph_nvls:32:16,nccl_graph:0:16
  run vfl_${N}gpu        $T --master-port $port benchmarks/bench_vfl.py
  run fed_dcgan_${N}gpu  $T --master-port $port benchmarks/bench_fed_dcgan.py
  run nccl_baselines_${N}gpu $T --master-port $port baseline/nccl_baselines.py
  T2="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
  run pp_llama_2gpu      $T2 --master-port $port benchmarks/bench_pp_llama.py
fi
echo "results in $OUT/"
