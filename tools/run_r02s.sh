#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -m gpu -q -x -k "spmv or cg or c2 or c5" -p no:cacheprovider 2>&1 | tail -3
for dyn in 1 0; do
  QDSP_SPMV_DYNAMIC=$dyn timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02s_bench_dyn$dyn.json 2> gpurun_out/r02s_bench_dyn$dyn.err
  echo "[$((SECONDS-T0))s] dyn=$dyn: $(python -c "import json; d=json.load(open('gpurun_out/r02s_bench_dyn$dyn.json')); print(round(d['value'],1),'it/s  spmv', round(d['roofline']['ms_per_launch']*1e3,2),'us', round(d['roofline']['achieved'],1),'GB/s frac', round(d['roofline']['frac'],4), ' e2e', round(d['e2e']['value'],1))")"
done
