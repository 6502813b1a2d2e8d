#!/bin/bash
# Full GPU visit: the whole -m gpu suite (as the driver runs it), the last test file again alone in a fresh
# process, smoke, the default bench, and an ncu capture of the named kernel (only after a clean plain run).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1
ALL=$?
echo "pytest all exit $ALL" >> gpurun_out/pytest_all.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_zz_alone.log 2>&1
echo "pytest zz alone exit $?" >> gpurun_out/pytest_zz_alone.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err
echo "bench train exit $?" >> gpurun_out/bench_train.err
if [ $ALL -eq 0 ] && [ -n "$NCU_KERNEL" ]; then
  B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-encoder"
  timeout 600 $B > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$NCU_KERNEL" -s ${NCU_SKIP:-30} -c 2 -f -o gpurun_out/prof_${NCU_NAME:-kernel} $B > gpurun_out/ncu_full.log 2>&1
fi
tail -4 gpurun_out/pytest_all.log; tail -3 gpurun_out/pytest_zz_alone.log; tail -2 gpurun_out/smoke.log
cat gpurun_out/bench_train.json | cut -c1-2500; tail -3 gpurun_out/bench_train.err
