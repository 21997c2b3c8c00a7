#!/bin/bash
# 8 GPUs: knobs of the three-kernel iteration
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
T0=$SECONDS
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
      bench.py --gpus 8 --steps 20 --warmup 3 2> gpurun_out/r02r_$name.err | grep '^{' > gpurun_out/r02r_$name.json
  echo "[$((SECONDS-T0))s] $name: $(python -c "import json,sys; d=json.load(open('gpurun_out/r02r_$name.json')); print(round(d['value'],1), 'it/s', round(d['ms_per_step']*10,2), 'us/iter')" 2>&1 | tail -1)"
}
run base QDSP_X=0
run vec2 QDSP_CG_VEC_CTAS=2
run vec4 QDSP_CG_VEC_CTAS=4
run pdl QDSP_PDL=1
run pdl_vec2 QDSP_PDL=1 QDSP_CG_VEC_CTAS=2
