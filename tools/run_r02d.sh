#!/bin/bash
# Round 2, call d (1 GPU): the -m gpu suite with the persistent CG kernel as the default solver, then the bench line with
# the persistent kernel and with the three-kernel iteration (QDSP_CG_FUSED=0) on the same box.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
timeout 900 python -m pytest tests -m gpu -q -x --durations=12 -p no:cacheprovider > gpurun_out/r02d_tests.log 2>&1
echo "[$((SECONDS-T0))s] pytest -m gpu exit $?"; tail -22 gpurun_out/r02d_tests.log
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_fused.json 2> gpurun_out/r02d_bench_fused.err
echo "[$((SECONDS-T0))s] bench fused exit $?"; cut -c1-400 gpurun_out/r02d_bench_fused.json; tail -3 gpurun_out/r02d_bench_fused.err
QDSP_CG_FUSED=0 timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_classic.json 2> gpurun_out/r02d_bench_classic.err
echo "[$((SECONDS-T0))s] bench classic exit $?"; cut -c1-400 gpurun_out/r02d_bench_classic.json
