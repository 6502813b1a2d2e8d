"""GPU timeline of a few training steps (torch.profiler / CUPTI): per-kernel durations inside the running step and the
idle gaps between consecutive kernels.  Prints a table; measurement aid, not a benchmark."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

from alphago_b200 import nn
from alphago_b200.training import SLTrainer

b = 256
net = nn.ConvNet(19, 48, [192] * 12, [5] + [3] * 11, max_batch=b)
net.init_uniform(1)
tr = SLTrainer(net, lr=0.003, decay=0.0001)
rng = np.random.RandomState(0)
x = torch.from_numpy((rng.rand(b, 48, 19, 19) < 0.3).astype(np.uint8)).cuda()
y = torch.from_numpy(rng.randint(0, 361, b).astype(np.int32)).cuda()
for _ in range(30):
    tr.step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10):
        tr.step(x, y)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type is not None and "cuda" in str(e.device_type).lower() and e.time_range is not None]
ev = sorted(ev, key=lambda e: e.time_range.start)
rows = [(e.name.split("(")[0][-40:], e.time_range.start, e.time_range.end) for e in ev]
# one step = from a conv_chain (forward) to the next sgd_all
names = [r[0] for r in rows]
idx = [i for i, n in enumerate(names) if "sgd_all" in n]
a, z = idx[4] + 1, idx[5] + 1
prev_end = rows[a - 1][2]
tot_busy = tot_gap = 0.0
for n, s, e in rows[a:z]:
    gap = s - prev_end
    print("%-42s dur %8.1f us   gap before %6.1f us" % (n, e - s, gap))
    tot_busy += e - s
    tot_gap += max(gap, 0.0)
    prev_end = max(prev_end, e)
print("step: busy %.1f us, idle between kernels %.1f us, span %.1f us" % (tot_busy, tot_gap, rows[z - 1][2] - rows[a - 1][2]))
