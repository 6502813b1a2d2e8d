"""Stress the chained forward: many back-to-back forwards at a few batch sizes (value net, as scripts/eval_sweep.py runs it).
A dependency that is ever missed shows up as the kernel's own timeout (printed by the device) or as a changed output."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from alphago_b200 import nn

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
net = nn.ConvNet(19, 49, [192] * 12, [5] + [3] * 11, head=nn.VALUE_HEAD, max_batch=4096)
net.init_uniform(1)
for b in (128, 64, 256, 100, 128):
    x = torch.from_numpy((np.random.RandomState(b).rand(b, 49, 19, 19) < 0.3).astype(np.uint8)).cuda()
    first = net.forward(x).clone()
    bad = 0
    for i in range(reps):
        out = net.forward(x)
        if i % 100 == 99:
            bad += int(not torch.equal(out, first))
    torch.cuda.synchronize()
    print("batch", b, "reps", reps, "mismatching checks", bad, flush=True)
print("done")

# ---- training soak: the chained data gradients and weight gradients, thousands of unsynchronised steps, two batch sizes
from alphago_b200.training import SLTrainer

for b in (256, 96):
    tnet = nn.ConvNet(19, 48, [192] * 12, [5] + [3] * 11, max_batch=256)
    tnet.init_uniform(1)
    tr = SLTrainer(tnet, lr=0.003, decay=0.0001)
    rng = np.random.RandomState(b)
    x = torch.from_numpy((rng.rand(b, 48, 19, 19) < 0.3).astype(np.uint8)).cuda()
    y = torch.from_numpy(rng.randint(0, 361, b).astype(np.int32)).cuda()
    for i in range(reps):
        tr.step(x, y)
    torch.cuda.synchronize()
    loss, acc = tr.pop_metrics(b * reps)
    print("train batch", b, "steps", reps, "mean loss %.4f acc %.4f" % (loss, acc), flush=True)
    assert np.isfinite(loss)
    tnet.close()
print("training soak done")
