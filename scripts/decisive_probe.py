"""Measurement aid: how many CUDA SGD steps on the 1033 real positions of the reference's HDF5 fixture make the 12 x 192
policy net DECISIVE (so that top-1 agreement between the bf16 CUDA forward and the fp32 oracle is not a coin flip between
near-equal moves), and what the agreement is then.  Prints one line per setting."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from alphago_b200 import go, nn
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES
from alphago_b200.training import SLTrainer
from oracle import nn_oracle as no

g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "ref_hdf5_fixture.npz"))
actions, offsets = g["actions"], g["offsets"]
states = []
for start, count in offsets:
    st = go.GameState(19)
    for i in range(start, start + count):
        states.append(st.copy())
        st.do_move((int(actions[i][0]), int(actions[i][1])))
pp = Preprocess(DEFAULT_FEATURES)
planes = pp.states_to_tensor(states)
d_act = torch.from_numpy(np.ascontiguousarray(actions)).cuda()
n = len(states)
print("positions", n, flush=True)
filters, widths = [192] * 12, [5] + [3] * 11
for lr, mom, steps, headx in [(0.03, 0.0, 300, 1), (0.03, 0.9, 300, 1), (0.01, 0.9, 600, 1), (0.03, 0.9, 600, 20), (0.003, 0.9, 600, 20)]:
    params = no.make_params(48, filters, widths, seed=1)
    params["head_w"] = (params["head_w"] * headx).astype(np.float32)
    net = nn.ConvNet(19, 48, filters, widths, max_batch=256)
    for l, (w, b) in enumerate(params["convs"]):
        net.set_conv(l, w, b)
    net.set_head(0, params["head_w"], params["head_b"])
    tr = SLTrainer(net, lr=lr, decay=0.0, momentum=mom)
    rng = np.random.RandomState(0)
    perm = rng.permutation(n)
    hist = []
    for s in range(steps):
        idx = perm[(np.arange(256) + s * 256) % n]
        x, y = tr.assemble(planes, d_act, idx, rng.randint(0, 8, 256))
        tr.step(x, y)
        if s % 50 == 49:
            hist.append("%.3f/%.2f" % tr.pop_metrics(50 * 256))
    torch.cuda.synchronize()
    w = {"convs": [net.get_conv(l) for l in range(12)], "head_w": net.get_head(0)[0], "head_b": net.get_head(0)[1], "pos_bias": None}
    sel = perm[:256]
    xs = planes[torch.from_numpy(sel).cuda()].contiguous()
    probs = net.forward(xs).cpu().numpy()
    t = time.time()
    fp = no.forward(w, xs.cpu().numpy())
    fpp = fp["probs"].numpy()
    logit_err = None
    same = (probs.argmax(1) == fpp.argmax(1)).mean()
    top2 = np.sort(fpp, 1)[:, -2:]
    print("lr %g mom %g steps %d headx %d: loss/acc %s | top1 agree %.4f, max prob err %.3g (max prob %.3g, median top prob %.3g, median margin %.3g) oracle %.1fs"
          % (lr, mom, steps, headx, " ".join(hist), same, np.abs(probs - fpp).max(), fpp.max(), np.median(top2[:, 1]),
             np.median(top2[:, 1] - top2[:, 0]), time.time() - t), flush=True)
    net.close()
