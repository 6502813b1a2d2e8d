"""Leaves per second through alphago_b200.leaf_queue.LeafEvaluator (pack on the host, one encoder launch, policy + value
forward, results back on the host) for a few batch sizes; 192-filter policy (48 planes) and value (49 planes) nets.
Prints one JSON line."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from alphago_b200 import go
from alphago_b200.leaf_queue import LeafEvaluator
from alphago_b200.models import CNNPolicy, CNNValue
from alphago_b200.preprocessing import DEFAULT_FEATURES

policy = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, seed=1, max_batch=1024)
value = CNNValue(DEFAULT_FEATURES + ["color"], layers=12, filters_per_layer=192, seed=2, max_batch=1024)
rng = np.random.RandomState(0)
states = []
st = go.GameState(19)
while len(states) < 1024:
    moves = st.get_legal_moves(include_eyes=False)
    if not moves or len(st.history) > 300:
        st = go.GameState(19)
        continue
    st.do_move(moves[rng.randint(len(moves))])
    if len(st.history) > 20:
        states.append(st.copy())
rows = []
for b in (1, 8, 64, 256, 1024):
    q = LeafEvaluator(policy, value, max_batch=b)
    moves = [s.get_legal_moves() for s in states[:b]]
    q.evaluate(states[:b], moves)
    reps = max(5, 256 // b)
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(reps):
        q.evaluate(states[:b], moves)
    dt = (time.time() - t0) / reps
    for s_ in states[:b]:
        q.submit(s_, want_moves=False)
    q.flush_arrays()
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(reps):
        for s_ in states[:b]:
            q.submit(s_, want_moves=False)
        q.flush_arrays()
    da = (time.time() - t0) / reps
    rows.append({"batch": b, "ms_per_flush": round(dt * 1e3, 3), "leaves_per_s": round(b / dt, 1),
                 "arrays_ms_per_flush": round(da * 1e3, 3), "arrays_leaves_per_s": round(b / da, 1)})
print(json.dumps({"what": "LeafEvaluator.evaluate: host pack + encode (49 planes, shared) + policy and value forward + results on host",
                  "rows": rows}))
