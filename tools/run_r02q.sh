#!/bin/bash
# 8 GPUs: bench fused vs three-kernel, + phase trace of the fused kernel
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
for mode in 1 0; do
  QDSP_CG_FUSED=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2960$mode \
      bench.py --gpus 8 --steps 20 --warmup 3 2> gpurun_out/r02q_bench8_fused$mode.err | grep '^{' > gpurun_out/r02q_bench8_fused$mode.json
  echo "[$((SECONDS-T0))s] bench 8 GPUs fused=$mode exit $?"; cut -c1-140 gpurun_out/r02q_bench8_fused$mode.json
done
QDSP_CG_TRACE=2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29613 \
      bench.py --gpus 8 --steps 3 --warmup 3 2>&1 | grep "fused trace" | tail -16 > gpurun_out/r02q_trace8.txt
cat gpurun_out/r02q_trace8.txt | cut -c1-250
QDSP_CG_TRACE=1 QDSP_CG_FUSED=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29614 \
      bench.py --gpus 8 --steps 3 --warmup 3 2>&1 | grep "qdsp trace" | tail -8 > gpurun_out/r02q_trace8_classic.txt
cat gpurun_out/r02q_trace8_classic.txt | cut -c1-250
echo "[$((SECONDS-T0))s] done"
