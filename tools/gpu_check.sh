#!/bin/bash
# The standard single-GPU check on a B200 box (under gpurun): GPU test suite, the bench line, the secondary configs.
#   tools/gpu_check.sh <tag> [tests|bench|paths|ncu-launches]...      (default: tests bench paths)
# Logs go to gpurun_out/<tag>_*.{log,json}; copy what is to be judged into profiles/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-check}; shift
WHAT=${@:-tests bench paths}
T0=$SECONDS
for w in $WHAT; do
  case $w in
    tests) timeout 1500 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/${TAG}_tests.log 2>&1
           echo "[$((SECONDS-T0))s] tests: $(tail -1 gpurun_out/${TAG}_tests.log)";;
    bench) timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
           echo "[$((SECONDS-T0))s] bench rc=$?: $(head -c 600 gpurun_out/${TAG}_bench.json)";;
    paths) timeout 900 python tools/bench_paths.py > gpurun_out/${TAG}_paths.jsonl 2> gpurun_out/${TAG}_paths.err
           echo "[$((SECONDS-T0))s] paths rc=$?"; cut -c1-200 gpurun_out/${TAG}_paths.jsonl;;
    ncu-launches) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
             --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
             --no-kernel-to-beat --no-live-traffic > gpurun_out/${TAG}_ncu.log 2>&1
           echo "[$((SECONDS-T0))s] ncu launches rc=$?";;
  esac
done
