#!/bin/bash
# Sort ranking variants (CUDA events) + one ncu --set full capture of the onesweep kernel with source.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for v in 0 1 2; do
  QDSP_RS_RANK=$v timeout 60 python tools/sort_probe.py 26 3 kv32,k64v32,k32 >> gpurun_out/r01n_variants.jsonl 2>> gpurun_out/r01n_variants.err
done
cat gpurun_out/r01n_variants.jsonl
QDSP_RS_RANK=0 timeout 60 python tools/sort_probe.py 25 1 kv32 > gpurun_out/r01n_plain.log 2>&1 &&
timeout 200 ncu --set full --clock-control none --import-source on -k regex:rs_onesweep -s 8 -c 1 -f -o gpurun_out/r01n_onesweep \
    python tools/sort_probe.py 25 1 kv32 > gpurun_out/r01n_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/r01n_ncu.log
