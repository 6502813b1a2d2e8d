#!/bin/bash
# what the 8-GPU step spends its extra time on: exchange variants + per-stream timelines of rank 0
mkdir -p gpurun_out
N=${NGPU:-8}
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps ${STEPS:-50} --warmup 5 --no-cpu-baseline --no-encoder --no-sustained --timeline gpurun_out/r2_timeline_${N}gpu_$name.txt > gpurun_out/dp${N}_$name.json 2> gpurun_out/dp${N}_$name.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/dp${N}_$name.json")); print("n=$N $name", d["value"], d["ms_per_step"])
except Exception as e: print("n=$N $name failed", e)
PY
  tail -3 gpurun_out/r2_timeline_${N}gpu_$name.txt 2>/dev/null
}
if [ -n "$MATRIX2" ]; then
run nvls_launchcut_b2_r8 AG_DP_FLAGS=0 AG_DP_BUCKETS=2 AG_DP_RESERVE_SMS=8 NCCL_ALGO=NVLS
run nvls_gated_b3_r16 AG_DP_FLAGS=1 AG_DP_BUCKETS=3 AG_DP_RESERVE_SMS=16 NCCL_ALGO=NVLS
run nvls_gated_b6_r8 AG_DP_FLAGS=1 AG_DP_BUCKETS=6 AG_DP_RESERVE_SMS=8 NCCL_ALGO=NVLS
run simple_launchcut_b2_r8 AG_DP_FLAGS=0 AG_DP_BUCKETS=2 AG_DP_RESERVE_SMS=8 NCCL_PROTO=Simple
run gated_b3_r24 AG_DP_FLAGS=1 AG_DP_BUCKETS=3 AG_DP_RESERVE_SMS=24
exit 0
fi
run launchcut_b2_r8 AG_DP_FLAGS=0 AG_DP_BUCKETS=2 AG_DP_RESERVE_SMS=8
run gated_b6_r4 AG_DP_FLAGS=1 AG_DP_BUCKETS=6 AG_DP_RESERVE_SMS=4
run gated_b6_r16 AG_DP_FLAGS=1 AG_DP_BUCKETS=6 AG_DP_RESERVE_SMS=16
run gated_b6_r4_cta4 AG_DP_FLAGS=1 AG_DP_BUCKETS=6 AG_DP_RESERVE_SMS=4 NCCL_MAX_CTAS=4
run gated_b3_r8_cta8 AG_DP_FLAGS=1 AG_DP_BUCKETS=3 AG_DP_RESERVE_SMS=8 NCCL_MAX_CTAS=8
run gated_b4_r16_cta16 AG_DP_FLAGS=1 AG_DP_BUCKETS=4 AG_DP_RESERVE_SMS=16 NCCL_MAX_CTAS=16
