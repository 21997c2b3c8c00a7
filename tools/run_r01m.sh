#!/bin/bash
# One short GPU call (round 1, last ~10 GPU-minutes): new N1/N2 tests group by group (separate processes, so a CUDA
# fault in one group does not poison the rest), the bench line, the parity tests that touch the edited kernels, and the
# N1/N2 micro-bench.  Everything lands in gpurun_out/r01m_*.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
first=150
for k in "reduce and not by_key" exclusive_scan select reduce_by_key sort "builder or errors"; do
  name=$(echo "$k" | tr ' ' '_')
  timeout $first python -m pytest tests/test_gpu_algorithms.py -q -x -k "$k" -p no:cacheprovider > gpurun_out/r01m_algos_$name.log 2>&1
  echo "[$((SECONDS-T0))s] algos -k '$k' exit $?" | tee -a gpurun_out/r01m_summary.txt
  tail -3 gpurun_out/r01m_algos_$name.log | tee -a gpurun_out/r01m_summary.txt
  first=75
done
timeout 170 python bench.py --steps 5 --warmup 3 > gpurun_out/r01m_bench.json 2> gpurun_out/r01m_bench.err
echo "[$((SECONDS-T0))s] bench exit $?" | tee -a gpurun_out/r01m_summary.txt
cut -c1-400 gpurun_out/r01m_bench.json | tee -a gpurun_out/r01m_summary.txt
timeout 150 python -m pytest tests/test_gpu_parity.py -q -p no:cacheprovider -k "(sort or scan or build or error or g4 or spmv_g1) and not 2p28 and not 20m" > gpurun_out/r01m_parity.log 2>&1
echo "[$((SECONDS-T0))s] parity subset exit $?" | tee -a gpurun_out/r01m_summary.txt
tail -3 gpurun_out/r01m_parity.log | tee -a gpurun_out/r01m_summary.txt
timeout 90 python tools/bench_paths.py algos > gpurun_out/r01m_algos_bench.jsonl 2> gpurun_out/r01m_algos_bench.err
echo "[$((SECONDS-T0))s] algos bench exit $?" | tee -a gpurun_out/r01m_summary.txt
cat gpurun_out/r01m_algos_bench.jsonl | cut -c1-200 | tee -a gpurun_out/r01m_summary.txt
