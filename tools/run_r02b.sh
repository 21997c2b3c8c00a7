#!/bin/bash
# Round 2, call b (1 GPU): ncu --set full of the kernels that had no committed capture:
# scan, collapse / histogram / pack of the CSR build, the stream-path SpMV on the C5 FEM matrix, the two CG vector kernels.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 120 $NCU -k regex:scan_inclusive -s 3 -c 1 -o gpurun_out/r02b_scan python tools/bench_paths.py scan > gpurun_out/r02b_scan.log 2>&1
echo "[$((SECONDS-T0))s] scan ncu exit $?"
timeout 200 $NCU -k 'regex:collapse_kernel|rs_histogram|pack_triplets' -s 3 -c 3 -o gpurun_out/r02b_build python tools/bench_paths.py build > gpurun_out/r02b_build.log 2>&1
echo "[$((SECONDS-T0))s] build ncu exit $?"
timeout 200 $NCU -k 'regex:spmv_stream|spmv_dot_stream' -s 3 -c 2 -o gpurun_out/r02b_fem \
    --metrics lts__t_sectors_srcunit_tex.sum,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum \
    python tools/bench_paths.py fem > gpurun_out/r02b_fem.log 2>&1
echo "[$((SECONDS-T0))s] fem ncu exit $?"
timeout 200 $NCU -k 'regex:cg_residual|cg_direction' -s 20 -c 2 -o gpurun_out/r02b_cgvec python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02b_cgvec.log 2>&1
echo "[$((SECONDS-T0))s] cg vec ncu exit $?"
ls -la gpurun_out/*.ncu-rep
