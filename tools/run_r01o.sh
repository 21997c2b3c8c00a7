#!/bin/bash
# onesweep v2 (instruction-trimmed) against the current kernel: timing + the sort / build parity tests under v2
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for v in 0 1; do
  QDSP_RS_VARIANT=$v timeout 60 python tools/sort_probe.py 26 3 kv32,k64v32,k32 >> gpurun_out/r01o_variants.jsonl 2>> gpurun_out/r01o_variants.err
done
cat gpurun_out/r01o_variants.jsonl
QDSP_RS_VARIANT=1 timeout 150 python -m pytest tests/test_gpu_parity.py tests/test_gpu_algorithms.py -q -p no:cacheprovider -k "(sort or build) and not 20m" > gpurun_out/r01o_parity_v2.log 2>&1
echo "parity under v2 exit $?"; tail -3 gpurun_out/r01o_parity_v2.log
