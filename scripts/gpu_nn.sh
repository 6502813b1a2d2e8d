#!/bin/bash
# GPU-box visit for the conv stack: NN parity tests (verbose), encoder tests, then either a
# sanitizer pass on the smallest failing NN case or the benches.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_nn_gpu.py -m gpu -v --maxfail=30 -p no:cacheprovider -x > gpurun_out/pytest_nn.log 2>&1
NN=$?
echo "pytest nn exit $NN" >> gpurun_out/pytest_nn.log
timeout 600 python -m pytest tests/test_encoder_gpu.py -m gpu -q --maxfail=30 -p no:cacheprovider > gpurun_out/pytest_enc.log 2>&1
echo "pytest enc exit $?" >> gpurun_out/pytest_enc.log
if [ $NN -ne 0 ]; then
  timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_nn_gpu.py -m gpu -q -x -k "forward_matches_oracle and 13" -p no:cacheprovider > gpurun_out/sanitizer_nn.log 2>&1
else
  timeout 600 python bench.py --steps 100 > gpurun_out/bench.json 2> gpurun_out/bench.err
fi
tail -25 gpurun_out/pytest_nn.log
tail -3 gpurun_out/pytest_enc.log
