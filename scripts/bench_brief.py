"""Run the default train bench (no encoder, no CPU baseline) and print value, ms/step and the per-kernel times briefly."""
import json
import subprocess
import sys

out = subprocess.run([sys.executable, "bench.py", "--no-encoder", "--no-cpu-baseline"] + sys.argv[1:], capture_output=True, text=True)
d = json.loads(out.stdout.strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], {k.split("(")[0] + k.split("(")[1][:8]: v["ms"] for k, v in d["roofline"]["kernels"].items()})
