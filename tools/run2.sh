set -x
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q 2>&1 | tail -5
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
B="bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline"
timeout 300 $TR $B 2>&1 | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' '; echo " <- merged+order"
QDSP_CG_VARIANT=classic timeout 300 $TR $B 2>&1 | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' '; echo " <- classic+order"
QDSP_CG_VARIANT=classic QDSP_HALO_ORDER=0 timeout 300 $TR $B 2>&1 | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' '; echo " <- classic noorder"
QDSP_CG_VARIANT=merged timeout 200 python tools/probe_slab.py 32 256 2>&1 | grep -v "^$" | tail -4
