TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512"
B4="bench.py --gpus 4 --steps 4 --warmup 3 --no-cpu-baseline"
val() { grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' '; }
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q 2>&1 | tail -3
echo "== N=4 classic"; QDSP_CG_VARIANT=classic QDSP_HALO_ORDER=0 timeout 200 $TR4 $B4 2>/dev/null | val; echo
echo "== N=4 classic, ghost blocks last"; QDSP_CG_VARIANT=classic timeout 200 $TR4 $B4 2>/dev/null | val; echo
echo "== N=4 classic trace"; QDSP_CG_TRACE=1 QDSP_CG_VARIANT=classic QDSP_HALO_ORDER=0 timeout 200 $TR4 $B4 2>&1 | grep "qdsp trace" | tail -4
