"""The chained launches (conv_chain_kernel, wgrad_chain_kernel) hold CTAs that wait for each other, so the whole grid must
be on the device at once.  They are launched cooperatively: the runtime refuses a grid that cannot be co-resident (the net
then falls back to its per-layer kernels) and the device starts the grid only as a whole.  Here a kernel on a second
stream holds 8 SMs for 3 s -- longer than the 2 s after which a chained kernel gives up on a dependency and traps -- while
a forward pass and a training step are queued next to it: they must come out clean and identical to the undisturbed run.
Runs in a child process so that a trap (a dead CUDA context) could not take the test session with it."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import sys, time
sys.path.insert(0, %r)
import numpy as np, torch
from alphago_b200 import nn, _lib
n = 64
net = nn.ConvNet(19, 48, [192] * 12, [5] + [3] * 11, max_batch=n)
net.init_uniform(1)
rng = np.random.RandomState(0)
x = torch.from_numpy((rng.rand(n, 48, 19, 19) < 0.3).astype(np.uint8)).cuda()
y = torch.from_numpy(rng.randint(0, 361, n).astype(np.int32)).cuda()
acc = torch.zeros(2, device="cuda")
ref_p = net.forward(x).clone()
net.forward_loss(x, y, acc)
net.backward_range(11, 0, n)
torch.cuda.synchronize()
ref_g = net.grad_tensor().clone()
plan = net.launch_plan()
assert plan["fwd"] == list(range(12)) and plan["wgrad"] == list(range(1, 12)) and plan["refused"] == 0, plan
side = torch.cuda.Stream()
hold_us = 3000000
t0 = time.time()
_lib.check(_lib.lib().ag_debug_hold_sms(8, hold_us, side.cuda_stream), "hold")
time.sleep(0.05)                      # the holders are on their SMs before the chain is queued
p = net.forward(x)
acc.zero_()
net.forward_loss(x, y, acc)
net.backward_range(11, 0, n)
torch.cuda.synchronize()
dt = time.time() - t0
assert torch.equal(p, ref_p), "forward differs next to the SM holder"
assert torch.equal(net.grad_tensor(), ref_g), "gradients differ next to the SM holder"
print("OK waited %%.2f s, plan %%s" %% (dt, net.launch_plan()))
''' % ROOT


def test_chained_launches_next_to_a_kernel_that_holds_sms():
    r = subprocess.run([sys.executable, "-c", CHILD], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK waited" in r.stdout, "child failed:\n%s\n%s" % (r.stdout[-2000:], r.stderr[-4000:])
    waited = float(r.stdout.split("OK waited")[1].split()[0])
    # the 148-CTA grids cannot start while 8 SMs are held: they waited for the holder instead of starting piecemeal
    assert waited >= 2.9, r.stdout
