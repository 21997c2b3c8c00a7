#!/bin/bash
# First GPU call of the next round (1 GPU, ~4 min of box time): everything that round 1 ended without.
#   gpurun --timeout 420 -- 'bash tools/run_r02_first.sh'
# 1. the whole -m gpu suite with per-test durations (round 1 only ran subsets after the last sort change)
# 2. the bench line + its ncu launch list (the build part of the step changed: bulk append, new sort sweep)
# 3. the secondary configs (C3 sort/scan at 2^28, C4 build, C5 FEM, the algorithm library)
# 4. one ncu --set full capture of the final sort sweep (profiles/r01n_kernel_onesweep.txt is the kernel BEFORE the trimming)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
timeout 300 python -m pytest tests -m gpu -q -x --durations=15 -p no:cacheprovider > gpurun_out/r02a_tests.log 2>&1
echo "[$((SECONDS-T0))s] pytest -m gpu exit $?"; tail -25 gpurun_out/r02a_tests.log
timeout 120 python bench.py --steps 10 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err &&
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02a_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02a_ncu_bench.log 2>&1
echo "[$((SECONDS-T0))s] bench + launch list exit $?"; cut -c1-300 gpurun_out/r02a_bench.json
timeout 240 python tools/bench_paths.py > gpurun_out/r02a_paths.jsonl 2> gpurun_out/r02a_paths.err
echo "[$((SECONDS-T0))s] bench_paths exit $?"; cut -c1-220 gpurun_out/r02a_paths.jsonl
for it in 16 20 24; do   # larger tiles: the open direction of profiles/r01_sort_variants.jsonl
  QDSP_RS_ITEMS=$it timeout 60 python tools/sort_probe.py 26 3 kv32,k32 | sed "s/\"default\"/\"items$it\"/" >> gpurun_out/r02a_sort_items.jsonl
done
cat gpurun_out/r02a_sort_items.jsonl
timeout 60 python tools/sort_probe.py 25 1 kv32 > gpurun_out/r02a_sort_plain.log 2>&1 &&
timeout 200 ncu --set full --clock-control none --import-source on -k regex:rs_onesweep -s 8 -c 1 -f -o gpurun_out/r02a_onesweep \
    python tools/sort_probe.py 25 1 kv32 > gpurun_out/r02a_ncu_sort.log 2>&1
echo "[$((SECONDS-T0))s] sort ncu exit $?"
# then, here:  python profiles/summarize.py launches gpurun_out/r02a_launches.csv > profiles/r02a_launches.txt
#              python profiles/summarize.py kernel gpurun_out/r02a_onesweep.ncu-rep > profiles/r02a_kernel_onesweep.txt
# multi-GPU (separate call, gpurun --gpus 8):  bash tools/run8.sh ; and for C5:
#   for N in 2 4 8; do python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
#       --master-port $((29540+N)) tools/bench_fem_dist.py; done
