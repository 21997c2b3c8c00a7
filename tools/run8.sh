# 8-GPU run: correctness at 8 ranks, then the strong-scaling series of bench.py and a per-kernel trace
TR() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + $1)) "${@:2}"; }
val() { grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' '; }
timeout 400 python -m pytest tests/test_dist_gpu.py -m gpu -x -q 2>&1 | tail -4
for N in 8 4 2; do
  echo "== N=$N"; TR $N bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tee gpurun_out/r1k_bench_n$N.json | val; echo
done
echo "== N=8 trace"; QDSP_CG_TRACE=1 TR 8 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "qdsp trace" | tail -8 | sort
