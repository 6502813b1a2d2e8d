"""Throughput of the SGF -> dataset converter (alphago_b200/game_converter.py) on the six game records under
tests/golden/sgf, repeated to a few thousand positions: host replay + packing alone, the CUDA encoding alone, and the
whole sgfs_to_hdf5 call; beside it the CPU oracle encoding the same packed positions (the reference's own converter needs
the `sgf` and `h5py` packages and cannot run here; its Preprocess alone takes ~8.4 ms per position).
Prints one JSON line."""
import glob
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from alphago_b200.game_converter import game_converter
from oracle import go_oracle as og

feats = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
         "ladder_capture", "ladder_escape", "sensibleness", "zeros"]
import shutil
_src = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "sgf", "*.sgf")))
_copies = tempfile.mkdtemp()   # eight copies under names of their own (one name per file in `file_offsets`, as in h5py)
files = []
for rep in range(8):
    for f in _src:
        files.append(shutil.copy(f, os.path.join(_copies, "%d-%s" % (rep, os.path.basename(f)))))
conv = game_converter(feats)
t0 = time.time()
packed = []
for f in files:
    p, a, e = conv.pack_game(f, 19)
    packed += p
t_pack = time.time() - t0
conv.encode_packed_positions(packed[:512], 19)
torch.cuda.synchronize()
t0 = time.time()
planes = conv.encode_packed_positions(packed, 19)
t_enc = time.time() - t0
t0 = time.time()
conv.encode_packed_positions(packed, 19, bits=True)
t_bits = time.time() - t0
with tempfile.TemporaryDirectory() as d:
    t0 = time.time()
    conv.sgfs_to_hdf5(files, os.path.join(d, "out.h5"), bd_size=19)   # the reference's HDF5 layout (default)
    t_all = time.time() - t0
    hdf5_bytes = os.path.getsize(os.path.join(d, "out.h5"))
    t0 = time.time()
    conv.sgfs_to_hdf5(files, os.path.join(d, "out.npz"), bd_size=19, container="npz")   # bit-packed archive
    t_npz = time.time() - t0
n = len(packed)
s = min(n, 2048)
board = np.stack([p[0] for p in packed[:s]]); age = np.stack([p[1] for p in packed[:s]])
player = np.asarray([p[2] for p in packed[:s]], np.int8); ko = np.asarray([p[3] for p in packed[:s]], np.int16)
cores = os.cpu_count() or 1
t0 = time.time()
ref = og.encode_batch(board, age, player, ko, 19, feats, threads=cores)
t_cpu = time.time() - t0
assert np.array_equal(ref, planes[:s])
print(json.dumps({"positions": n, "planes": int(planes.shape[1]),
                  "host_replay_pack_positions_per_s": round(n / t_pack, 1),
                  "encode_incl_copies_positions_per_s": round(n / t_enc, 1),
                  "encode_bitpacked_incl_copies_positions_per_s": round(n / t_bits, 1),
                  "sgfs_to_hdf5_positions_per_s": round(n / t_all, 1), "hdf5_file_bytes": hdf5_bytes,
                  "sgfs_to_npz_archive_positions_per_s": round(n / t_npz, 1),
                  "cpu_oracle_encode_positions_per_s": round(s / t_cpu, 1), "cpu_cores": cores,
                  "bit_exact_vs_oracle_on": s}))
