TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
B="bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline"
val() { grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' '; }
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q 2>&1 | tail -15
echo "== N=2"; timeout 300 $TR $B 2>/dev/null | val; echo
echo "== N=2 trace"; QDSP_CG_TRACE=1 timeout 300 $TR $B 2>&1 | grep "qdsp trace" | tail -2
