#!/bin/bash
# Round-end GPU visit: the whole -m gpu suite as the driver runs it, the last test file alone in a fresh process, smoke,
# the default benches, then (only after those exited 0 without ncu) the ncu launch list of the train bench and one
# full capture of the chained weight-gradient kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1
ALL=$?
echo "pytest all exit $ALL" >> gpurun_out/pytest_all.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_zz_alone.log 2>&1
echo "pytest zz alone exit $?" >> gpurun_out/pytest_zz_alone.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err
echo "bench train exit $?" >> gpurun_out/bench_train.err
timeout 300 python bench.py --workload encode > gpurun_out/bench_encode.json 2> gpurun_out/bench_encode.err
echo "bench encode exit $?" >> gpurun_out/bench_encode.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
echo "bench reference exit $?" >> gpurun_out/bench_reference.err
if [ $ALL -eq 0 ]; then
  B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-encoder"
  timeout 300 $B > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/train_launches.csv $B > gpurun_out/ncu_list.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:wgrad_chain_kernel -s 4 -c 2 -f -o gpurun_out/prof_wgrad_chain $B > gpurun_out/ncu_full.log 2>&1
fi
tail -3 gpurun_out/pytest_all.log; tail -2 gpurun_out/pytest_zz_alone.log; tail -2 gpurun_out/smoke.log
cut -c1-300 gpurun_out/bench_train.json; tail -2 gpurun_out/bench_train.err; cut -c1-200 gpurun_out/bench_encode.json; cut -c1-300 gpurun_out/bench_reference.json
tail -2 gpurun_out/ncu_list.log; tail -2 gpurun_out/ncu_full.log
