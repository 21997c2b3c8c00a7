#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "scan or fused" -p no:cacheprovider 2>&1 | tail -3
timeout 120 python tools/bench_paths.py scan > gpurun_out/r02h_scan.jsonl 2>&1; cat gpurun_out/r02h_scan.jsonl
echo "[$((SECONDS-T0))s]"
