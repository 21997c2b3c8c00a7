#!/bin/bash
# 2 GPUs: multi-GPU parity worker (new cases: changing x, one-way ghosts, FEM) + bench fused vs three-kernel
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r02n_dist_tests.log 2>&1
echo "[$((SECONDS-T0))s] dist pytest exit $?"; tail -25 gpurun_out/r02n_dist_tests.log | cut -c1-300
for mode in 1 0; do
  QDSP_CG_FUSED=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2950$mode \
      bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02n_bench2_fused$mode.json 2> gpurun_out/r02n_bench2_fused$mode.err
  echo "[$((SECONDS-T0))s] bench 2 GPUs fused=$mode exit $?"; cut -c1-160 gpurun_out/r02n_bench2_fused$mode.json; tail -2 gpurun_out/r02n_bench2_fused$mode.err
done
