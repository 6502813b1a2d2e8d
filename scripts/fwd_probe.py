"""Forward-only timing of the policy net (192 filters, 48 planes) at a few batch sizes; prints one line.
Used on the GPU box to compare launch strategies (AG_CONV_CHAIN=0 per-layer kernels, 1 one chained launch)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from alphago_b200 import nn

net = nn.ConvNet(19, 48, [192] * 12, [5] + [3] * 11, max_batch=1024)
net.init_uniform(1)
out = {}
for b in (1, 16, 256, 1024):
    x = torch.from_numpy((np.random.RandomState(b).rand(b, 48, 19, 19) < 0.3).astype(np.uint8)).cuda()
    for _ in range(5):
        net.forward(x)
    reps = 50
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        net.forward(x)
    e1.record()
    torch.cuda.synchronize()
    out[b] = round(e0.elapsed_time(e1) / reps * 1e3, 1)
print("chain=%s forward us by batch:" % os.environ.get("AG_CONV_CHAIN", "default"), out)
