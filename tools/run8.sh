# 8-GPU measurement script (one gpurun --gpus 8 call): where does the per-iteration time go at 8 ranks?
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512"
B8="bench.py --gpus 8 --steps 4 --warmup 3 --no-cpu-baseline"
B4="bench.py --gpus 4 --steps 4 --warmup 3 --no-cpu-baseline"
val() { grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' '; }
echo "== N=8 classic, natural order";  QDSP_CG_VARIANT=classic QDSP_HALO_ORDER=0 timeout 200 $TR8 $B8 2>/dev/null | val; echo
echo "== N=8 classic, ghost blocks last"; QDSP_CG_VARIANT=classic timeout 200 $TR8 $B8 2>/dev/null | val; echo
echo "== N=8 merged, ghost blocks last"; QDSP_CG_VARIANT=merged timeout 200 $TR8 $B8 2>/dev/null | val; echo
echo "== N=8 classic trace (no graph)"; QDSP_CG_TRACE=1 QDSP_CG_VARIANT=classic timeout 200 $TR8 $B8 2>&1 | grep "qdsp trace" | tail -8 | sort
echo "== N=8 merged trace (no graph)"; QDSP_CG_TRACE=1 QDSP_CG_VARIANT=merged timeout 200 $TR8 $B8 2>&1 | grep "qdsp trace" | tail -8 | sort
echo "== N=8 classic, NCCL all-reduce"; QDSP_ALLREDUCE=nccl QDSP_CG_VARIANT=classic timeout 200 $TR8 $B8 2>/dev/null | val; echo
echo "== N=4 classic"; QDSP_CG_VARIANT=classic timeout 200 $TR4 $B4 2>/dev/null | val; echo
echo "== N=4 classic trace"; QDSP_CG_TRACE=1 QDSP_CG_VARIANT=classic timeout 200 $TR4 $B4 2>&1 | grep "qdsp trace" | tail -4
