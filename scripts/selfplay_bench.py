"""Self-play throughput of alphago_b200/reinforcement_policy_trainer.make_training_pairs: G games of 19x19 played in lock
step for a fixed number of plies by a 12-layer 128-filter policy (the reference's default CNNPolicy) against itself, then
one REINFORCE update per game.  Prints one JSON line (plies = moves played over all games)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from alphago_b200.ai import ProbabilisticPolicyPlayer
from alphago_b200.models import SGD, CNNPolicy
from alphago_b200.reinforcement_policy_trainer import make_training_pairs, train_batch

FEATS = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after", "sensibleness", "zeros"]
rows = []
for games in (16, 128):
    pol = CNNPolicy(FEATS, layers=12, filters_per_layer=128, seed=1, max_batch=256)
    pl = ProbabilisticPolicyPlayer(pol)
    pol.model.compile(loss="binary_crossentropy", optimizer=SGD(lr=0.001))
    np.random.seed(1)
    make_training_pairs(pl, pl, FEATS, games, max_plies=8)
    plies = 80
    torch.cuda.synchronize()
    t0 = time.time()
    X, y, w = make_training_pairs(pl, pl, FEATS, games, max_plies=plies)
    torch.cuda.synchronize()
    t1 = time.time()
    train_batch(pl, X, y, w, 0.001)
    torch.cuda.synchronize()
    t2 = time.time()
    make_training_pairs(pl, pl, FEATS, games, max_plies=8, fast=True)
    torch.cuda.synchronize()
    t3 = time.time()
    make_training_pairs(pl, pl, FEATS, games, max_plies=plies, fast=True)
    torch.cuda.synchronize()
    t4 = time.time()
    rows.append({"games": games, "plies_per_game": plies, "selfplay_plies_per_s": round(games * plies / (t1 - t0), 1),
                 "selfplay_s": round(t1 - t0, 3), "fast_selfplay_plies_per_s": round(games * plies / (t4 - t3), 1), "update_s": round(t2 - t1, 3), "training_pairs": int(sum(len(a) for a in X))})
# device-resident games (alphago_b200/selfplay.py): a ply is encode -> forward -> ag_selfplay_step, no host round trip
from alphago_b200.selfplay import make_training_pairs_device
dev_rows = []
for games in (128, 1024, 4096):
    pol = CNNPolicy(FEATS, layers=12, filters_per_layer=128, seed=1, max_batch=games)
    pl = ProbabilisticPolicyPlayer(pol)
    pol.model.compile(loss="binary_crossentropy", optimizer=SGD(lr=0.001))
    make_training_pairs_device(pl, pl, FEATS, games, max_plies=8)
    plies = 80
    torch.cuda.synchronize()
    t0 = time.time()
    X, y, w = make_training_pairs_device(pl, pl, FEATS, games, max_plies=plies)
    torch.cuda.synchronize()
    t1 = time.time()
    if games <= 128:
        train_batch(pl, X, y, w, 0.001)
    torch.cuda.synchronize()
    t2 = time.time()
    dev_rows.append({"games": games, "plies_per_game": plies, "device_selfplay_plies_per_s": round(games * plies / (t1 - t0), 1),
                     "selfplay_s": round(t1 - t0, 3), "update_s": round(t2 - t1, 3), "training_pairs": int(sum(len(a) for a in y))})
    del pol, pl
print(json.dumps({"what": "device-resident self-play, 19x19, 12 x 128-filter policy vs itself (encode + forward + ag_selfplay_step per ply)", "rows": dev_rows}))
print(json.dumps({"what": "make_training_pairs + train_batch, 19x19, 12 x 128-filter policy vs itself", "rows": rows}))
