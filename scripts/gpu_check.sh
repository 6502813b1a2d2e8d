#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, then ncu (only if everything before it exited 0).
# Writes everything to gpurun_out/.  (compute-sanitizer is not available on this pool: bad accesses are hunted with the
# kernels' own bounded waits and traps, small cases and the CPU oracle.)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 -p no:cacheprovider > gpurun_out/pytest.log 2>&1
PT=$?
echo "pytest exit $PT" >> gpurun_out/pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
BENCH="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
if [ $PT -eq 0 ]; then
  timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
  echo "bench exit $?" >> gpurun_out/bench.err
  timeout 300 $BENCH > gpurun_out/bench_short.json 2>> gpurun_out/bench.err &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_list.log 2>&1
  timeout 300 $BENCH > /dev/null 2>> gpurun_out/bench.err &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNEL:-encode_kernel}" -s 4 -c 2 -f -o gpurun_out/prof_${NCU_NAME:-encode} $BENCH > gpurun_out/ncu_full.log 2>&1
fi
tail -5 gpurun_out/pytest.log
cat gpurun_out/bench.json 2>/dev/null
