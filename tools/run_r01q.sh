#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for it in 16 12 8; do
  QDSP_RS_ITEMS=$it timeout 60 python tools/sort_probe.py 26 3 kv32,k32 | sed "s/\"default\"/\"items$it\"/" >> gpurun_out/r01q_variants.jsonl 2>> gpurun_out/r01q_variants.err
done
timeout 60 python tools/sort_probe.py 26 3 k64v32 >> gpurun_out/r01q_variants.jsonl 2>> gpurun_out/r01q_variants.err
cat gpurun_out/r01q_variants.jsonl
timeout 150 python -m pytest tests/test_gpu_parity.py tests/test_gpu_algorithms.py -q -p no:cacheprovider -k "(sort or build) and not 20m" > gpurun_out/r01q_parity.log 2>&1
echo "parity exit $?"; tail -3 gpurun_out/r01q_parity.log
QDSP_RS_ITEMS=8 timeout 150 python -m pytest tests/test_gpu_parity.py -q -p no:cacheprovider -k "sort and not 2p28" > gpurun_out/r01q_parity8.log 2>&1
echo "parity items=8 exit $?"; tail -3 gpurun_out/r01q_parity8.log
