"""Measurement aid: (A) gradient parity of the 12 x 192 net at batch 256 on TRAINED weights (600 CUDA SGD steps on the 1033
real positions), against the bf16-emulating and the fp32 oracle; (B) kernel-level parity at that shape: each layer's weight
gradient and data gradient recomputed on the CPU from the CUDA path's OWN inputs (its activations and pre-activation
gradients, read back), which isolates one kernel from the conditioning of the whole backward pass."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
import torch.nn.functional as Fn

from alphago_b200 import go, nn
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES
from alphago_b200.training import SLTrainer
from oracle import nn_oracle as no
from test_nn_gpu import cosine, grads_from_flat, rel_err

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 600
g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "ref_hdf5_fixture.npz"))
actions, offsets = g["actions"], g["offsets"]
states = []
for start, count in offsets:
    st = go.GameState(19)
    for i in range(start, start + count):
        states.append(st.copy())
        st.do_move((int(actions[i][0]), int(actions[i][1])))
planes = Preprocess(DEFAULT_FEATURES).states_to_tensor(states)
d_act = torch.from_numpy(np.ascontiguousarray(actions)).cuda()
n = len(states)
F, W, B = [192] * 12, [5] + [3] * 11, 256
net = nn.ConvNet(19, 48, F, W, max_batch=B)
net.init_uniform(1)
tr = SLTrainer(net, lr=0.01, decay=0.0, momentum=0.9)
rng = np.random.RandomState(0)
perm = rng.permutation(n)
for s in range(steps):
    idx = perm[(np.arange(B) + s * B) % n]
    x, y = tr.assemble(planes, d_act, idx, rng.randint(0, 8, B))
    tr.step(x, y)
print("trained %d steps: loss/acc %s" % (steps, tr.pop_metrics(max(1, steps) * B)))
params = {"convs": [net.get_conv(l) for l in range(12)], "head_w": net.get_head(0)[0], "head_b": net.get_head(0)[1], "pos_bias": None}
sel = perm[:B]
xs = planes[torch.from_numpy(sel).cuda()].contiguous()
lab = torch.from_numpy((actions[sel, 0].astype(np.int32) * 19 + actions[sel, 1]).astype(np.int32)).cuda()
acc = torch.zeros(2, device="cuda")
probs = torch.empty(B, 361, device="cuda")
net.forward_loss(xs, lab, acc, probs)
net.backward_range(11, 0, B)
torch.cuda.synchronize()
got = grads_from_flat(net, net.grad_tensor().cpu().numpy())
xh, lh = xs.cpu().numpy(), lab.cpu().numpy()
t = time.time()
le, ae, em, oem = no.loss_and_grads(params, xh, lh, emulate_bf16=True)
lf, af, fp, ofp = no.loss_and_grads(params, xh, lh, emulate_bf16=False)
print("oracle %.1fs; loss cuda %.5f emul %.5f fp32 %.5f; correct cuda %d emul %d" % (time.time() - t, float(acc[0]) / B, le, lf, int(acc[1]), round(ae * B)))
pf = ofp["probs"].detach().numpy()
pc = probs.cpu().numpy()
logit_c = np.log(np.maximum(pc, 1e-30)); logit_f = np.log(np.maximum(pf, 1e-30))
print("top1 agree %.4f; max abs prob err %.4g (max prob %.3g); max abs log-prob err on p>1e-3: %.4g"
      % ((pc.argmax(1) == pf.argmax(1)).mean(), np.abs(pc - pf).max(), pf.max(), np.abs(logit_c - logit_f)[pf > 1e-3].max()))
for l in range(11, -1, -1):
    print("(A) conv%-2d.w cuda-vs-emul rel %.4f cos %.5f | cuda-vs-fp32 %.4f | emul-vs-fp32 %.4f | bias cuda-vs-emul %.4f"
          % (l, rel_err(got[2 * l], em[2 * l]), cosine(got[2 * l], em[2 * l]), rel_err(got[2 * l], fp[2 * l]), rel_err(em[2 * l], fp[2 * l]),
             rel_err(got[2 * l + 1], em[2 * l + 1])))
print("(A) head.w cuda-vs-emul %.4f" % rel_err(got[24], em[24]))
# (B) kernel level, from the CUDA path's own tensors
bf = lambda a: torch.from_numpy(a).to(torch.bfloat16).to(torch.float32)
for l in range(11, -1, -1):
    G = torch.from_numpy(net.debug_read(l, 1, B))                         # pre-activation gradient of layer l (bf16 values)
    X = torch.from_numpy(net.debug_read(l - 1, 0, B)) if l > 0 else torch.from_numpy(xh.astype(np.float32))
    k = W[l]
    w = torch.zeros(F[l], X.shape[1], k, k, requires_grad=True)
    out = Fn.conv2d(X, w, padding=k // 2)
    (out * G).sum().backward()
    dw = w.grad.numpy()
    db = G.sum((0, 2, 3)).numpy()
    line = "(B) conv%-2d wgrad-from-own-inputs rel %.5f, bias %.5f" % (l, rel_err(got[2 * l], dw), rel_err(got[2 * l + 1], db))
    if l > 0:
        wl = bf(params["convs"][l][0])
        dx = Fn.conv_transpose2d(G, wl, padding=k // 2)                      # data gradient into layer l-1's output
        act = torch.from_numpy(net.debug_read(l - 1, 0, B))
        want = (dx * (act > 0)).to(torch.bfloat16).to(torch.float32).numpy()
        have = net.debug_read(l - 1, 1, B)
        scale = np.abs(want).max()
        line += " | dgrad into layer %d: max err %.3g of max %.3g (%.2f bf16 ulps of max), rel fro %.5f" % (
            l - 1, np.abs(have - want).max(), scale, np.abs(have - want).max() / (scale * 2.0 ** -8), rel_err(have, want))
    print(line, flush=True)
