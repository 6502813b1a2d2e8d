"""Runs the encoder alone (for ncu captures): python scripts/enc_only.py [positions] [46|48] [launches]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from alphago_b200 import go
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
planes = int(sys.argv[2]) if len(sys.argv) > 2 else 48
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
feats = DEFAULT_FEATURES if planes == 48 else [f for f in DEFAULT_FEATURES if not f.startswith("ladder")]
b, a, pl, ko = go.playout_positions(1000, n, size=19, max_moves=400, per_game=16)
d = [torch.from_numpy(x).cuda() for x in (b, a, pl, ko)]
pp = Preprocess(feats)
out = torch.empty((n, pp.output_dim, 19, 19), dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(reps):
    if i == reps - 1:
        e0.record()
    pp.encode_packed(*d, 19, out=out)
e1.record()
torch.cuda.synchronize()
print("%d positions, %d planes: last launch %.3f ms" % (n, pp.output_dim, e0.elapsed_time(e1)))
