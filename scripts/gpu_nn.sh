#!/bin/bash
# GPU-box visit for the conv stack: NN parity tests (verbose), a slice of the encoder tests, the benches,
# then the ncu launch list of a short train bench (only if everything before it exited 0).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_nn_gpu.py tests/test_policy_gpu.py -m gpu -v --maxfail=30 -p no:cacheprovider > gpurun_out/pytest_nn.log 2>&1
NN=$?
echo "pytest nn exit $NN" >> gpurun_out/pytest_nn.log
timeout 600 python -m pytest tests/test_encoder_gpu.py -m gpu -q --maxfail=30 -p no:cacheprovider -k "golden or edge or subsets" > gpurun_out/pytest_enc.log 2>&1
echo "pytest enc exit $?" >> gpurun_out/pytest_enc.log
timeout 900 python bench.py > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err
echo "bench train exit $?" >> gpurun_out/bench_train.err
if [ $NN -eq 0 ]; then
  B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-encoder"
  timeout 600 $B > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_train.csv $B > gpurun_out/ncu_list.log 2>&1
fi
grep -E "PASSED|FAILED|ERROR|passed|failed" gpurun_out/pytest_nn.log | tail -25
tail -3 gpurun_out/pytest_enc.log
cat gpurun_out/bench_train.json; tail -5 gpurun_out/bench_train.err
