#!/bin/bash
# One GPU visit that refreshes everything under profiles/ that comes from a single GPU: the full -m gpu suite, smoke, the
# default bench (train + encoder sub-object + CPU baseline), the encoder bench and its N sweep, the eval sweep (config 4),
# and the ncu launch lists (only after the plain runs exited 0).  Outputs land in gpurun_out/ and are copied by hand.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1
echo "pytest all exit $?" >> gpurun_out/pytest_all.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_zz_alone.log 2>&1
echo "pytest zz alone exit $?" >> gpurun_out/pytest_zz_alone.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err
T=$?
timeout 600 python bench.py --workload encode > gpurun_out/bench_encode.json 2> gpurun_out/bench_encode.err
E=$?
: > gpurun_out/encode_sweep.jsonl
for n in 256 512 1024 2048 4096 8192 16384; do
  timeout 300 python bench.py --workload encode --positions $n --no-cpu-baseline >> gpurun_out/encode_sweep.jsonl 2>> gpurun_out/bench_encode.err
done
timeout 300 python bench.py --workload encode --planes 46 --no-cpu-baseline > gpurun_out/bench_encode46.json 2>> gpurun_out/bench_encode.err
timeout 600 python scripts/eval_sweep.py > gpurun_out/eval_sweep.json 2> gpurun_out/eval_sweep.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
if [ $T -eq 0 ]; then
  B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-encoder"
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file gpurun_out/launches_train.csv $B > gpurun_out/ncu_list.log 2>&1
fi
if [ $E -eq 0 ]; then
  B="python bench.py --workload encode --steps 3 --warmup 3 --no-cpu-baseline"
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_encode.csv $B > gpurun_out/ncu_list2.log 2>&1
fi
tail -3 gpurun_out/pytest_all.log; tail -2 gpurun_out/pytest_zz_alone.log; tail -2 gpurun_out/smoke.log
cut -c1-600 gpurun_out/bench_train.json; echo; cut -c1-300 gpurun_out/bench_encode.json; echo
python - <<'PY'
import json
for l in open("gpurun_out/encode_sweep.jsonl"):
    d = json.loads(l); print(d["config"]["positions_per_step"], d["value"], d["ms_per_step"], d["roofline"]["frac"])
d = json.load(open("gpurun_out/bench_encode46.json")); print("46 planes", d["value"], d["ms_per_step"], d["roofline"]["frac"])
d = json.load(open("gpurun_out/eval_sweep.json"))
for k, rows in d.items(): print(k, [(r["batch"], r["latency_ms"], r["positions_per_s"]) for r in rows if r["batch"] in (1, 16, 256, 4096)])
print(open("gpurun_out/bench_reference.json").read()[:400])
PY
