#!/bin/bash
# last call of round 1 (about a minute of budget): smoke, the bench line and the sort figures on the final kernels
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 25 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01s_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/r01s_smoke.log
timeout 25 python bench.py --steps 5 --warmup 3 > gpurun_out/r01s_bench.json 2> gpurun_out/r01s_bench.err; echo "bench exit $?"; cut -c1-200 gpurun_out/r01s_bench.json
timeout 12 python tools/sort_probe.py 26 3 kv32,k32 > gpurun_out/r01s_sort.jsonl 2>&1; cat gpurun_out/r01s_sort.jsonl
