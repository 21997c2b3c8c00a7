#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
for mode in 1 0; do
  QDSP_CG_FUSED=$mode timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02o_bench1_fused$mode.json 2> gpurun_out/r02o_bench1_fused$mode.err
  echo "[$((SECONDS-T0))s] bench 1 GPU fused=$mode exit $?"; cut -c1-140 gpurun_out/r02o_bench1_fused$mode.json
done
for mode in 1 0; do
  QDSP_CG_FUSED=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2950$mode \
      bench.py --gpus 2 --steps 10 --warmup 3 2> gpurun_out/r02o_bench2_fused$mode.err | grep '^{' > gpurun_out/r02o_bench2_fused$mode.json
  echo "[$((SECONDS-T0))s] bench 2 GPUs fused=$mode exit $?"; cut -c1-140 gpurun_out/r02o_bench2_fused$mode.json
done
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r02o_dist_tests.log 2>&1
echo "[$((SECONDS-T0))s] dist pytest exit $?"; tail -5 gpurun_out/r02o_dist_tests.log | cut -c1-300; grep -n "QdspError\|AssertionError\|differs" gpurun_out/r02o_dist_tests.log | head
