cd /root/repo
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -2
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
timeout 400 $TR --master-port 29704 tools/bench_fem_dist.py > gpurun_out/r02z_fem_n2.json 2> gpurun_out/r02z_fem_n2.err; tail -1 gpurun_out/r02z_fem_n2.json | cut -c1-300
QDSP_CG_TRACE=3 timeout 400 $TR --master-port 29705 tools/bench_fem_dist.py 2>&1 >/dev/null | grep -o "\[qdsp timeline.*" | tail -2
