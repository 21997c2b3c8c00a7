#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
echo -n "rounds   "; timeout 100 python tools/bench_paths.py scan 2>&1 | cut -c1-120
echo -n "fallback "; QDSP_SCAN_ROUNDS=0 timeout 100 python tools/bench_paths.py scan 2>&1 | cut -c1-120
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_algorithms.py -m gpu -q -x -k "scan or g4 or ops" -p no:cacheprovider 2>&1 | tail -2
QDSP_SCAN_ROUNDS=0 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "scan" -p no:cacheprovider 2>&1 | tail -2
