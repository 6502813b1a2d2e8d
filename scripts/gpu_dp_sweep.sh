#!/bin/bash
mkdir -p gpurun_out
N=${NGPU:-2}
timeout 600 python -m pytest tests/test_dp_gpu.py -x -q -p no:cacheprovider > gpurun_out/pytest_dp.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_dp.log
tail -3 gpurun_out/pytest_dp.log
CONFIGS=${CONFIGS:-8,4 4,4 2,4 0,4 4,3 4,6}
for cfg in $CONFIGS; do
  r=${cfg%,*}; b=${cfg#*,}
  AG_DP_XCHG=${XCHG:-peer} AG_DP_XCHG_CTAS=${CTAS:-32} AG_DP_RESERVE_SMS=$r AG_DP_BUCKETS=$b timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps ${STEPS:-100} --warmup 10 --no-cpu-baseline --no-encoder --no-sustained --timeline gpurun_out/r2_timeline_${N}gpu_${XCHG:-peer}_r${r}_b${b}.txt > gpurun_out/dp${N}_r${r}_b${b}.json 2> gpurun_out/dp${N}_r${r}_b${b}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/dp${N}_r${r}_b${b}.json")); print("n=$N reserve=$r buckets=$b", d["value"], d["ms_per_step"], d["clocks"]["reasons"])
except Exception as e: print("reserve=$r buckets=$b failed", e)
PY
done
