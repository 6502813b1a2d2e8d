#!/bin/bash
# round 2 scaling runs on ONE box: NGPU ranks, the driver's own call (20 steps) and a 100-step run; 1 GPU on the same box beside them
mkdir -p gpurun_out
N=${NGPU:-8}
for steps in 20 100; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps $steps --warmup 5 --no-cpu-baseline --no-encoder --timeline gpurun_out/r2_timeline_${N}gpu_peer_${steps}steps.txt > gpurun_out/r2_scale_${N}gpu_${steps}steps.json 2> gpurun_out/scale_${N}_${steps}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2_scale_${N}gpu_${steps}steps.json")); print("n=$N steps=$steps", d["value"], d["ms_per_step"], d["clocks"]["reasons"], d["roofline"].get("whole_step_sustained",{}).get("positions_per_sec_per_gpu"))
except Exception as e: print("n=$N steps=$steps failed", e)
PY
done
[ -n "$SKIP1" ] && exit 0
for steps in 20 100; do
  python bench.py --steps $steps --warmup 5 --no-cpu-baseline --no-encoder > gpurun_out/r2_scale_1gpu_samebox${N}_${steps}steps.json 2>> gpurun_out/scale_1.err
  python -c "
import json; d=json.load(open('gpurun_out/r2_scale_1gpu_samebox${N}_${steps}steps.json')); print('n=1 steps=$steps', d['value'], d['ms_per_step'])"
done
