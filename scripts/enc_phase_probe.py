"""Measurement aid (library built with AG_NVCC_EXTRA=-DAG_ENC_PROFILE): where a warp of encode_kernel spends its cycles,
phase by phase, summed over all positions; plus kernel times for 46 / 48 planes."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from alphago_b200 import _lib, go
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
b, a, pl, ko = go.playout_positions(1000, n, size=19, max_moves=400, per_game=16)
d = [torch.from_numpy(x).cuda() for x in (b, a, pl, ko)]
F46 = [f for f in DEFAULT_FEATURES if not f.startswith("ladder")]
L = _lib.lib()
names = ["load+codes0", "union-find", "counts/liblists", "codes", "ladder mark", "expand+store"]
for feats in (F46, DEFAULT_FEATURES):
    pp = Preprocess(feats)
    out = torch.empty((n, pp.output_dim, 19, 19), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        pp.encode_packed(*d, 19, out=out)
    torch.cuda.synchronize()
    buf = (C.c_ulonglong * 8)()
    if hasattr(L, "ag_debug_enc_phases"):
        L.ag_debug_enc_phases(buf, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        pp.encode_packed(*d, 19, out=out)
    e1.record()
    torch.cuda.synchronize()
    print("%d planes, %d positions: %.3f ms per launch" % (pp.output_dim, n, e0.elapsed_time(e1) / 10))
    if hasattr(L, "ag_debug_enc_phases"):
        L.ag_debug_enc_phases(buf, 1)
        tot = sum(buf[i] for i in range(6))
        for i in range(6):
            print("   %-16s %6.1f %%   %8.0f cycles / position" % (names[i], 100.0 * buf[i] / tot, buf[i] / (10.0 * n)))
