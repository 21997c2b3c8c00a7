#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for v in 1 2; do
  QDSP_RS_VARIANT=$v timeout 60 python tools/sort_probe.py 26 3 kv32,k64v32,k32 >> gpurun_out/r01p_variants.jsonl 2>> gpurun_out/r01p_variants.err
done
cat gpurun_out/r01p_variants.jsonl
QDSP_RS_VARIANT=2 timeout 150 python -m pytest tests/test_gpu_parity.py tests/test_gpu_algorithms.py -q -p no:cacheprovider -k "(sort or build) and not 20m" > gpurun_out/r01p_parity_v3.log 2>&1
echo "parity under variant 2 exit $?"; tail -3 gpurun_out/r01p_parity_v3.log
