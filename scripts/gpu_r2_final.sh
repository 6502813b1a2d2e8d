#!/bin/bash
# round 2, final visit: whole GPU suite, smoke, both bench workloads, reference arm, encoder sweep, self-play bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=40 -p no:cacheprovider --durations=8 -s > gpurun_out/pytest_final.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_final.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_final.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke_final.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench20_final.json 2> gpurun_out/bench_final.err
timeout 600 python bench.py > gpurun_out/bench_final.json 2>> gpurun_out/bench_final.err
timeout 600 python bench.py --workload encode > gpurun_out/bench_encode_final.json 2>> gpurun_out/bench_final.err
for n in 256 1024 4096 16384; do python scripts/enc_only.py $n 48 6; python scripts/enc_only.py $n 46 6; done > gpurun_out/r2_encode_sweep_final.txt 2>&1
timeout 600 python scripts/selfplay_bench.py 2>&1 | tail -2 > gpurun_out/r2_selfplay_bench_final.json
grep -v "Warning\|warn" gpurun_out/pytest_final.log | tail -16
cat gpurun_out/smoke_final.log
