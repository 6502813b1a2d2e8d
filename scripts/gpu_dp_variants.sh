#!/bin/bash
# N-GPU data-parallel step time under a few settings (env passed through to all ranks); usage: gpu_dp_variants.sh N
N=${1:-2}
mkdir -p gpurun_out
run() {
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 5 --no-encoder --no-cpu-baseline 2> gpurun_out/dp.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', d['value'], d['ms_per_step'])"
}
run AG_DP_BUCKETS=2
run AG_DP_BUCKETS=3
run AG_DP_BUCKETS=2 AG_DP_RESERVE_SMS=16
