"""Time per UMMA of the forward conv kernel for different filter counts (N of the UMMA): checks how the tensor-core
cost scales with N.  AG_CONV_DEBUG=7 removes memory traffic so the MMA issue rate is what is measured."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from alphago_b200 import nn

for f in (64, 128, 192, 256):
    net = nn.ConvNet(19, 48, [f] * 4, [5, 3, 3, 3], max_batch=256)
    net.init_uniform(1)
    x = torch.from_numpy((np.random.RandomState(0).rand(256, 48, 19, 19) < 0.3).astype(np.uint8)).cuda()
    net.forward(x)
    for _ in range(5):
        net.run_layer(2, 0, 256)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        net.run_layer(2, 0, 256)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    kchunks = (f + 63) // 64
    mmas = 6 * 9 * kchunks * 4  # busiest SM: 6 tiles x taps x k-chunks x 4
    flop = 256 * 2 * 361 * 9 * f * f
    print("N=%d: %.1f us, %.0f TF/s, %.1f ns per UMMA (%d per SM)" % (f, ms * 1e3, flop / ms / 1e9, ms * 1e6 / mmas, mmas))
    net.close()
