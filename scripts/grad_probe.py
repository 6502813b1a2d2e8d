"""Measurement aid: per-tensor gradient error of the 12 x 192 net at a given batch against the bf16-emulating oracle and the
plain fp32 oracle, and the distance between the two oracles (what bf16 rounding alone does).  Env AG_CONV_CHAIN / AG_WGRAD_CHAIN
select the launch plan."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch

from alphago_b200 import go
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES
from oracle import nn_oracle as no
from test_nn_gpu import cosine, grads_from_flat, make_net, rel_err

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
headx = float(sys.argv[2]) if len(sys.argv) > 2 else 20
F, W = [192] * 12, [5] + [3] * 11
b, a, pl, ko = go.playout_positions(2016, batch, size=19, max_moves=400, per_game=16)
planes = Preprocess(DEFAULT_FEATURES).encode_packed(*[torch.from_numpy(x).cuda() for x in (b, a, pl, ko)], 19)
x = planes.cpu().numpy()
rng = np.random.RandomState(3)
legal = x[:, 20:28].sum(1).reshape(batch, 361) > 0
labels = np.asarray([rng.choice(np.flatnonzero(l)) if l.any() else 0 for l in legal], np.int32)
params = no.make_params(48, F, W, seed=1, bias_scale=0.02)
params["head_w"] = (params["head_w"] * headx).astype(np.float32)
net = make_net(19, 48, F, W, params, max_batch=batch)
acc = torch.zeros(2, device="cuda")
net.forward_loss(planes, torch.from_numpy(labels).cuda(), acc)
net.backward_range(11, 0, batch)
torch.cuda.synchronize()
print("plan", net.launch_plan())
got = grads_from_flat(net, net.grad_tensor().cpu().numpy())
_, _, em, oem = no.loss_and_grads(params, x, labels, emulate_bf16=True)
_, _, fp, ofp = no.loss_and_grads(params, x, labels, emulate_bf16=False)
print("batch", batch, "headx", headx, "max prob", float(ofp["probs"].max()))
for l in range(11, -1, -1):
    gpre = net.debug_read(l, 1, batch) if l > 0 else None
    print("conv%-2d.w  cuda-vs-emul rel %.4f cos %.5f | cuda-vs-fp32 rel %.4f | emul-vs-fp32 rel %.4f   | bias: cuda-vs-emul %.4f emul-vs-fp32 %.4f"
          % (l, rel_err(got[2 * l], em[2 * l]), cosine(got[2 * l], em[2 * l]), rel_err(got[2 * l], fp[2 * l]), rel_err(em[2 * l], fp[2 * l]),
             rel_err(got[2 * l + 1], em[2 * l + 1]), rel_err(em[2 * l + 1], fp[2 * l + 1])))
