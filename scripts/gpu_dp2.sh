#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dp_gpu.py -x -q -p no:cacheprovider > gpurun_out/pytest_dp.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_dp.log
tail -5 gpurun_out/pytest_dp.log
N=${NGPU:-2}
for flags in 1 0; do
  for b in 4 6; do
  AG_DP_FLAGS=$flags AG_DP_BUCKETS=$b timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 --no-cpu-baseline --no-encoder --no-sustained > gpurun_out/dp${N}_f${flags}_b${b}.json 2> gpurun_out/dp${N}_f${flags}_b${b}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/dp${N}_f${flags}_b${b}.json")); print("flags=$flags buckets=$b", d["value"], d["ms_per_step"])
except Exception as e: print("flags=$flags buckets=$b failed", e)
PY
  done
done
timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-encoder --no-sustained > gpurun_out/dp1.json 2>gpurun_out/dp1.err; python -c "
import json; d=json.load(open('gpurun_out/dp1.json')); print('1 gpu', d['value'], d['ms_per_step'])"
