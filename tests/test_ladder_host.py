"""CPU check of the ladder search the CUDA encoder runs: csrc/ladder.cuh compiled for the host (qualifiers
stubbed, tests/native/ladder_host_check.cpp) against the C oracle on seeded random-playout positions."""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle import go_oracle as og

HERE = os.path.dirname(os.path.abspath(__file__))


def _build():
    src = os.path.join(HERE, "native", "ladder_host_check.cpp")
    out = os.path.join(HERE, "native", "_ladder_host_check.so")
    hdr = os.path.join(os.path.dirname(HERE), "alphago_b200", "csrc", "ladder.cuh")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", out, src])
    return C.CDLL(out)


def test_device_ladder_logic_matches_oracle_on_cpu():
    L = _build()
    total_c = total_e = 0
    for size, n, seed in ((19, 160, 3), (9, 300, 4), (7, 200, 5), (13, 80, 6)):
        b, a, pl, ko = og.playout_positions(seed, n, size=size, max_moves=size * size + 30, per_game=8)
        want = og.encode_batch(b, a, pl, ko, size, ["ladder_capture", "ladder_escape"], threads=os.cpu_count() or 1)
        pts = size * size
        for i in range(n):
            cap = np.zeros(pts, np.uint8); esc = np.zeros(pts, np.uint8)
            board = np.ascontiguousarray(b[i])
            rc = L.ladder_planes_host(board.ctypes.data_as(C.c_void_p), size, int(pl[i]), int(ko[i]),
                                      cap.ctypes.data_as(C.c_void_p), esc.ctypes.data_as(C.c_void_p))
            assert rc == 0, (size, i, rc)
            assert np.array_equal(cap, want[i, 0].reshape(-1)), (size, i, "capture")
            assert np.array_equal(esc, want[i, 1].reshape(-1)), (size, i, "escape")
            total_c += int(cap.sum()); total_e += int(esc.sum())
    assert total_c > 50 and total_e > 50
