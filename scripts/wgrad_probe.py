"""Device time of the stand-alone weight-gradient op (ag_wgrad: wgrad_kernel + wgrad_reduce_kernel) on one 3x3, 192 -> 192
layer at batch 256 (the training step's middle layers), CUDA events on the launching stream, 50 launches back to back.
AG_WGRAD_DEBUG selects the measurement-only modes of wgrad_kernel (wrong results, see wgrad.cu).  One JSON line."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from alphago_b200 import _lib

n, S, width, ch = int(os.environ.get("PROBE_N", "256")), 19, 3, 192
rs = S + 1
rows = n * rs * rs
L = _lib.lib()
g = torch.randn((rows, ch), device="cuda").to(torch.bfloat16)
x = (torch.rand((rows, ch), device="cuda") < 0.4).to(torch.bfloat16)
dw = torch.empty((9, ch, ch), dtype=torch.float32, device="cuda")
nbytes = L.ag_wgrad_scratch_bytes(width, ch, ch, rows)
scratch = torch.empty(nbytes // 4, dtype=torch.float32, device="cuda")


def run():
    rc = L.ag_wgrad(C.c_void_p(g.data_ptr()), C.c_void_p(x.data_ptr()), rows, ch, ch, width, rs, C.c_void_p(scratch.data_ptr()), nbytes,
                    C.c_void_p(dw.data_ptr()), _lib.current_stream())
    assert rc == 0, _lib.last_error()


for _ in range(5):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 50
e0.record()
for _ in range(reps):
    run()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / reps * 1e3
flop = 2.0 * n * S * S * 9 * ch * ch
print(json.dumps({"n": n, "debug": os.environ.get("AG_WGRAD_DEBUG", "0"), "us_per_layer": round(us, 2), "tflops": round(flop / us / 1e6, 1),
                  "splits": nbytes // L.ag_wgrad_partial_bytes(width, ch, ch)}))
