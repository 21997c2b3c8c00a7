#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
timeout 300 python tools/debug_fused.py > gpurun_out/r02e_debug.log 2>&1
echo "[$((SECONDS-T0))s] debug exit $?"; cat gpurun_out/r02e_debug.log | tail -50
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02e_bench_fused.json 2> gpurun_out/r02e_bench_fused.err
echo "[$((SECONDS-T0))s] bench fused exit $?"; cut -c1-200 gpurun_out/r02e_bench_fused.json
timeout 120 python tools/bench_paths.py scan > gpurun_out/r02e_scan.jsonl 2>&1; cat gpurun_out/r02e_scan.jsonl
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "scan" -p no:cacheprovider 2>&1 | tail -3
