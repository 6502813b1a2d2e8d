"""`Preprocess`: GameState -> one-hot feature planes, drop-in for
AlphaGo/preprocessing/preprocessing.py:217-245, computed by the CUDA encoder (csrc/encode.cu).

`state_to_tensor(state)` keeps the reference contract (float64 ndarray (1, F, size, size)).
`states_to_tensor(states)` / `encode_packed(...)` are the batched paths the network consumes:
uint8 planes that stay on the device.  Accepts any GameState-like object (this package's or the
reference's): only `size`, `board`, `history`, `current_player` and `ko` are read.
"""
import ctypes as C

import numpy as np

from . import _lib

FEATURE_IDS = {
    "board": 0, "ones": 1, "turns_since": 2, "liberties": 3, "capture_size": 4,
    "self_atari_size": 5, "liberties_after": 6, "ladder_capture": 7, "ladder_escape": 8,
    "sensibleness": 9, "zeros": 10, "color": 11,
}
FEATURE_SIZES = {"board": 3, "ones": 1, "turns_since": 8, "liberties": 8, "capture_size": 8,
                 "self_atari_size": 8, "liberties_after": 8, "ladder_capture": 1, "ladder_escape": 1,
                 "sensibleness": 1, "zeros": 1, "color": 1}

DEFAULT_FEATURES = [
    "board", "ones", "turns_since", "liberties", "capture_size",
    "self_atari_size", "liberties_after", "ladder_capture", "ladder_escape",
    "sensibleness", "zeros"]


def pack_state(state):
    """(board int8[n], age uint8[n], player, ko) of a GameState-like object.  The age array is the
    literal history walk of preprocessing.py:20-43 over (history, board) -- a copied reference state
    can carry history entries for moves never played on it (go.py:167), which still age the stones."""
    if hasattr(state, "pack"):
        return state.pack()
    size = state.size
    board = np.ascontiguousarray(np.asarray(state.board).reshape(-1), dtype=np.int8)
    hist = np.asarray([-1 if m is None else m[0] * size + m[1] for m in state.history] or [0], dtype=np.int16)
    age = np.empty(size * size, np.uint8)
    _lib.lib().ag_age_from_history(size, _lib.np_ptr(board), _lib.np_ptr(hist), len(state.history), _lib.np_ptr(age))
    ko = -1 if state.ko is None else state.ko[0] * size + state.ko[1]
    return board, age, int(state.current_player), int(ko)


class Preprocess(object):
    """convert GameState objects to tensors of one-hot features (preprocessing.py:217-236)"""

    def __init__(self, feature_list=DEFAULT_FEATURES):
        self.output_dim = 0
        self.feature_list = feature_list
        self.processors = [None] * len(feature_list)
        ids = []
        for i in range(len(feature_list)):
            feat = feature_list[i].lower()
            if feat in FEATURE_IDS:
                self.processors[i] = feat
                self.output_dim += FEATURE_SIZES[feat]
                ids.append(FEATURE_IDS[feat])
            else:
                raise ValueError("uknown feature: %s" % feat)  # sic, preprocessing.py:236
        self._ids = np.asarray(ids, dtype=np.int32)

    # ---------------------------------------------------------------- device path
    def encode_packed(self, board, age, player, ko, size, out=None):
        """Packed positions already on the device -> uint8 planes (N, F, size, size) on the device.

        board int8 (N, size*size), age uint8 (N, size*size), player int8 (N,), ko int16 (N,) are
        CUDA torch tensors.  Asynchronous on the current stream."""
        import torch
        n = board.shape[0]
        for t, dt in ((board, torch.int8), (age, torch.uint8), (player, torch.int8), (ko, torch.int16)):
            if not t.is_cuda or t.dtype != dt or not t.is_contiguous():
                raise ValueError("encode_packed needs contiguous CUDA tensors int8/uint8/int8/int16")
        if out is None:
            out = torch.empty((n, self.output_dim, size, size), dtype=torch.uint8, device=board.device)
        rc = _lib.lib().ag_encode(C.c_void_p(board.data_ptr()), C.c_void_p(age.data_ptr()), C.c_void_p(player.data_ptr()),
                                  C.c_void_p(ko.data_ptr()), n, size, _lib.np_ptr(self._ids), len(self._ids),
                                  C.c_void_p(out.data_ptr()), _lib.current_stream())
        _lib.check(rc, "ag_encode")
        return out

    def encode_host(self, board, age, player, ko, size, device="cuda"):
        """Packed positions in HOST numpy arrays -> uint8 planes in a HOST numpy array; the
        host->device and device->host copies are part of the call (the end-to-end path)."""
        import torch
        n = len(player)
        hb = torch.from_numpy(np.ascontiguousarray(board, dtype=np.int8).reshape(n, -1))
        ha = torch.from_numpy(np.ascontiguousarray(age, dtype=np.uint8).reshape(n, -1))
        hp = torch.from_numpy(np.ascontiguousarray(player, dtype=np.int8).reshape(n))
        hk = torch.from_numpy(np.ascontiguousarray(ko, dtype=np.int16).reshape(n))
        dev = torch.device(device)
        d = [t.to(dev, non_blocking=True) for t in (hb, ha, hp, hk)]
        out = self.encode_packed(d[0], d[1], d[2], d[3], size)
        return out.cpu().numpy()

    def states_to_tensor(self, states, device="cuda"):
        """list of GameStates (one board size) -> uint8 device tensor (N, F, size, size)"""
        import torch
        if len(states) == 0:
            return torch.empty((0, self.output_dim, 0, 0), dtype=torch.uint8, device=device)
        size = states[0].size
        if not all(st.size == size for st in states):
            raise ValueError("all states must have the same size")
        packed = [pack_state(s) for s in states]
        board = torch.from_numpy(np.stack([p[0] for p in packed])).to(device, non_blocking=True)
        age = torch.from_numpy(np.stack([p[1] for p in packed])).to(device, non_blocking=True)
        player = torch.tensor([p[2] for p in packed], dtype=torch.int8).to(device, non_blocking=True)
        ko = torch.tensor([p[3] for p in packed], dtype=torch.int16).to(device, non_blocking=True)
        return self.encode_packed(board, age, player, ko, size)

    # ---------------------------------------------------------------- reference contract
    def state_to_tensor(self, state):
        """Convert a GameState to a (1, F, size, size) float64 ndarray (preprocessing.py:238-245)"""
        planes = self.states_to_tensor([state])
        return planes.cpu().numpy().astype(np.float64)
