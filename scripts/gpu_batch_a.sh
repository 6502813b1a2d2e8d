#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_selfplay_gpu.py tests/test_rl_trainer_gpu.py tests/test_trainer_gpu.py tests/test_nn_gpu.py tests/test_bench_shape_gpu.py -x -q -s -p no:cacheprovider 2>&1 | grep -v Warn | tail -15 > gpurun_out/batch_a_pytest.log
timeout 600 python scripts/selfplay_bench.py 2>&1 | tail -3 > gpurun_out/r2_selfplay_bench.json
for c in 1 0 1 0; do AG_CHAIN_COOP=$c python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-encoder --no-sustained | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(\"coop=$c\", d[\"value\"], d[\"ms_per_step\"], d[\"roofline\"][\"in_step\"] if \"in_step\" in d[\"roofline\"] else \"\")"; done > gpurun_out/coop_ab.txt 2>&1
cat gpurun_out/batch_a_pytest.log gpurun_out/r2_selfplay_bench.json gpurun_out/coop_ab.txt
