#!/bin/bash
# round 2: whole GPU suite, smoke, bench (both arms), launch list, full ncu captures of the encoder kernels and the forward chain
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=40 -p no:cacheprovider --durations=10 -s > gpurun_out/pytest.log 2>&1
PT=$?
echo "pytest exit $PT" >> gpurun_out/pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench20.json 2> gpurun_out/bench.err
timeout 600 python bench.py > gpurun_out/bench.json 2>> gpurun_out/bench.err
timeout 600 python bench.py --workload encode > gpurun_out/bench_encode.json 2>> gpurun_out/bench.err
timeout 600 python bench.py --workload encode --planes 46 --no-cpu-baseline > gpurun_out/bench_encode46.json 2>> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
BENCH="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-sustained"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2_train_launches.csv $BENCH > gpurun_out/ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"encode_kernel|ladder_resolve" -s 6 -c 2 -f -o gpurun_out/r2_encode python scripts/enc_only.py 16384 48 6 > gpurun_out/ncu_enc.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:"conv_chain|wgrad_chain" -s 6 -c 3 -f -o gpurun_out/r2_chains $BENCH > gpurun_out/ncu_chain.log 2>&1
for n in 256 1024 4096 16384; do python scripts/enc_only.py $n 48 6; python scripts/enc_only.py $n 46 6; done > gpurun_out/r2_encode_sweep.txt 2>&1
grep -v "Warning\|warn" gpurun_out/pytest.log | tail -25
cat gpurun_out/smoke.log
