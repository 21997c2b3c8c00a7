#!/bin/bash
# Multi-GPU check under `gpurun --gpus N`: dist parity tests, the C2 bench line and the C5 (FEM fp64) line at each world size.
#   tools/gpu_check_multi.sh <tag> <N...>         e.g. tools/gpu_check_multi.sh r02u 2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=$1; shift
T0=$SECONDS
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q -x -p no:cacheprovider > gpurun_out/${TAG}_dist_tests.log 2>&1
echo "[$((SECONDS-T0))s] dist tests: $(tail -1 gpurun_out/${TAG}_dist_tests.log)"
for N in "$@"; do
  timeout 400 $TR --nproc-per-node $N --master-port $((29600+N)) bench.py --gpus $N --steps 10 --warmup 3 \
      > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
  echo "[$((SECONDS-T0))s] C2 N=$N rc=$?: $(python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_n$N.json"))
    print(round(d["value"], 1), "it/s  e2e", round(d["e2e"]["value"], 1), " spmv/GPU", round(d["roofline"]["achieved"], 1), "GB/s  parity", d.get("parity_checked"))
except Exception as e:
    print("no line:", e)
PY
)"
  timeout 400 $TR --nproc-per-node $N --master-port $((29700+N)) tools/bench_fem_dist.py \
      > gpurun_out/${TAG}_fem_n$N.json 2> gpurun_out/${TAG}_fem_n$N.err
  echo "[$((SECONDS-T0))s] C5 N=$N rc=$?: $(cut -c1-400 gpurun_out/${TAG}_fem_n$N.json)"
done
