#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): the 2-rank DP parity test, then the train bench at 1..N GPUs under torchrun.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus.txt 2>&1
timeout 600 python -m pytest tests/test_dp_gpu.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_dp.log 2>&1
echo "pytest dp exit $?" >> gpurun_out/pytest_dp.log
tail -3 gpurun_out/pytest_dp.log
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    if [ $n -eq 1 ]; then
      timeout 600 python bench.py --gpus 1 --steps 100 --warmup 5 --no-encoder --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    else
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 100 --warmup 5 --no-encoder --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    fi
    echo "n=$n exit $?"; cut -c1-400 gpurun_out/scale_$n.json; tail -2 gpurun_out/scale_$n.err
  fi
done
