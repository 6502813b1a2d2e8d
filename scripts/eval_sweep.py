"""BASELINE configs[4]: batched eval throughput/latency sweep, batch 1..4096, for the value net (49 planes, 13 convs + FC
tanh) and the policy net (48 planes), 192 filters.  Device-timed forward only (planes resident in HBM)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from alphago_b200 import nn

out = {}
for name, planes, head in (("value_192f_49planes", 49, nn.VALUE_HEAD), ("policy_192f_48planes", 48, nn.POLICY_HEAD)):
    net = nn.ConvNet(19, planes, [192] * 12, [5] + [3] * 11, head=head, max_batch=4096)
    net.init_uniform(1)
    rows = []
    for b in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096):
        x = torch.from_numpy((np.random.RandomState(b).rand(b, planes, 19, 19) < 0.3).astype(np.uint8)).cuda()
        for _ in range(5):
            net.forward(x)
        reps = 50 if b <= 256 else 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            net.forward(x)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        rows.append({"batch": b, "latency_ms": round(ms, 4), "positions_per_s": round(b / ms * 1e3, 1)})
    out[name] = rows
    net.close()
print(json.dumps(out))
