"""Per-kernel timing of the three GEMM kernels of a middle 3x3 layer (batch 256, 192 filters); prints one line.
Used on the GPU box to compare kernel variants (e.g. AG_CONV_STAGES=4 vs 5)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from alphago_b200 import nn

net = nn.ConvNet(19, 48, [192] * 12, [5] + [3] * 11, max_batch=256)
net.init_uniform(1)
x = torch.from_numpy((np.random.RandomState(0).rand(256, 48, 19, 19) < 0.3).astype(np.uint8)).cuda()
y = torch.zeros(256, dtype=torch.int32, device="cuda")
acc = torch.zeros(2, device="cuda")
net.forward_loss(x, y, acc)
for l in range(11, -1, -1):
    net.backward_layer(l, 256)
out = {}
flop = 256 * 2 * 361 * 9 * 192 * 192
for mode, name in ((0, "fwd"), (1, "dgrad"), (2, "wgrad")):
    for _ in range(5):
        net.run_layer(6, mode, 256)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        net.run_layer(6, mode, 256)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    out[name] = "%.1f us %.0f TF/s" % (ms * 1e3, flop / ms / 1e9)
print(os.environ.get("AG_CONV_STAGES", "default"), out)
