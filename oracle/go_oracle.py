"""ctypes binding of oracle/go_oracle.c (CPU ORACLE -- test infrastructure, not product).

`build()` compiles the C file with gcc into oracle/_ref/libgo_oracle.so (git-ignored; it
travels to the GPU box with the snapshot).  `OState` mirrors the subset of the reference
`GameState` API (AlphaGo/go.py:9-349) the parity tests need.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "go_oracle.c")
_SO = os.path.join(_HERE, "_ref", "libgo_oracle.so")

FEATURE_IDS = {
    "board": 0, "ones": 1, "turns_since": 2, "liberties": 3, "capture_size": 4,
    "self_atari_size": 5, "liberties_after": 6, "ladder_capture": 7, "ladder_escape": 8,
    "sensibleness": 9, "zeros": 10, "color": 11,
}
DEFAULT_FEATURES = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
                    "liberties_after", "ladder_capture", "ladder_escape", "sensibleness", "zeros"]


def build(force=False):
    if not force and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(_SRC):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", "-o", _SO, _SRC])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        vp, i8p, u8p, i16p, i32p = C.c_void_p, C.POINTER(C.c_int8), C.POINTER(C.c_uint8), \
            C.POINTER(C.c_int16), C.POINTER(C.c_int32)
        L.og_new.restype = vp; L.og_new.argtypes = [C.c_int]
        L.og_copy.restype = vp; L.og_copy.argtypes = [vp]
        L.og_free.argtypes = [vp]
        for name in ("og_size", "og_player", "og_ko", "og_nhist", "og_is_end"):
            getattr(L, name).argtypes = [vp]
        L.og_hist.argtypes = [vp, C.c_int]
        L.og_set_player.argtypes = [vp, C.c_int]
        L.og_do_move.argtypes = [vp, C.c_int, C.c_int]
        L.og_is_legal.argtypes = [vp, C.c_int]
        L.og_is_eye.argtypes = [vp, C.c_int, C.c_int]
        L.og_is_eyeish.argtypes = [vp, C.c_int, C.c_int]
        L.og_is_ladder_capture.argtypes = [vp, C.c_int]
        L.og_is_ladder_escape.argtypes = [vp, C.c_int]
        L.og_legal_moves.argtypes = [vp, C.c_int, i16p]
        L.og_board.argtypes = [vp, i8p]
        L.og_liberty_counts.argtypes = [vp, i32p]
        L.og_group_sizes.argtypes = [vp, i32p]
        L.og_prisoners.argtypes = [vp, i32p, i32p]
        L.og_pack.argtypes = [vp, i8p, u8p, i8p, i16p]
        L.og_age_from_history.argtypes = [C.c_int, i8p, i16p, C.c_int, u8p]
        L.og_encode.argtypes = [i8p, u8p, C.c_int, C.c_int, C.c_int, i32p, C.c_int, u8p]
        L.og_encode_batch.argtypes = [i8p, u8p, i8p, i16p, C.c_int, C.c_int, i32p, C.c_int, u8p, C.c_int]
        L.og_playout_positions.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, i8p, u8p, i8p, i16p]
        L.og_feature_size.argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def feature_ids(feature_list):
    return np.asarray([FEATURE_IDS[f.lower()] for f in feature_list], dtype=np.int32)


def output_dim(feature_list):
    return int(sum(lib().og_feature_size(int(i)) for i in feature_ids(feature_list)))


def encode_batch(board, age, player, ko, size, feature_list=DEFAULT_FEATURES, threads=1):
    """Packed positions -> uint8 planes (N, F, size, size)."""
    board = np.ascontiguousarray(board, dtype=np.int8).reshape(-1, size * size)
    age = np.ascontiguousarray(age, dtype=np.uint8).reshape(-1, size * size)
    player = np.ascontiguousarray(player, dtype=np.int8).reshape(-1)
    ko = np.ascontiguousarray(ko, dtype=np.int16).reshape(-1)
    n = board.shape[0]
    ids = feature_ids(feature_list)
    F = output_dim(feature_list)
    out = np.empty((n, F, size, size), dtype=np.uint8)
    r = lib().og_encode_batch(_p(board, C.c_int8), _p(age, C.c_uint8), _p(player, C.c_int8), _p(ko, C.c_int16),
                              n, size, _p(ids, C.c_int32), len(ids), _p(out, C.c_uint8), int(threads))
    if r < 0:
        raise ValueError("bad feature id")
    return out


def playout_positions(seed, npos, size=19, max_moves=400, per_game=8):
    """Synthetic random-playout positions as packed encoder inputs."""
    n = size * size
    board = np.zeros((npos, n), np.int8); age = np.zeros((npos, n), np.uint8)
    player = np.zeros(npos, np.int8); ko = np.zeros(npos, np.int16)
    lib().og_playout_positions(int(seed), npos, size, max_moves, per_game, _p(board, C.c_int8), _p(age, C.c_uint8),
                               _p(player, C.c_int8), _p(ko, C.c_int16))
    return board, age, player, ko


class IllegalMove(Exception):
    pass


class OState(object):
    """Oracle game state with the reference's GameState surface (go.py:9-349)."""

    def __init__(self, size=19, _h=None):
        self._h = _h if _h is not None else lib().og_new(size)
        self.size = lib().og_size(self._h)

    def __del__(self):
        try:
            lib().og_free(self._h)
        except Exception:
            pass

    def _idx(self, a):
        return -1 if a is None else a[0] * self.size + a[1]

    def copy(self):
        return OState(_h=lib().og_copy(self._h))

    @property
    def current_player(self):
        return lib().og_player(self._h)

    @current_player.setter
    def current_player(self, v):
        lib().og_set_player(self._h, int(v))

    @property
    def ko(self):
        k = lib().og_ko(self._h)
        return None if k < 0 else divmod(k, self.size)

    @property
    def history(self):
        out = []
        for i in range(lib().og_nhist(self._h)):
            m = lib().og_hist(self._h, i)
            out.append(None if m < 0 else divmod(m, self.size))
        return out

    @property
    def is_end_of_game(self):
        return bool(lib().og_is_end(self._h))

    @property
    def board(self):
        b = np.zeros(self.size * self.size, np.int8)
        lib().og_board(self._h, _p(b, C.c_int8))
        return b.reshape(self.size, self.size)

    @property
    def liberty_counts(self):
        b = np.zeros(self.size * self.size, np.int32)
        lib().og_liberty_counts(self._h, _p(b, C.c_int32))
        return b.reshape(self.size, self.size)

    @property
    def group_sizes(self):
        b = np.zeros(self.size * self.size, np.int32)
        lib().og_group_sizes(self._h, _p(b, C.c_int32))
        return b.reshape(self.size, self.size)

    @property
    def prisoners(self):
        b, w = C.c_int32(), C.c_int32()
        lib().og_prisoners(self._h, C.byref(b), C.byref(w))
        return b.value, w.value

    def do_move(self, action, color=None):
        if lib().og_do_move(self._h, self._idx(action), int(color or 0)) != 0:
            raise IllegalMove(str(action))
        return self.is_end_of_game

    def is_legal(self, action):
        return bool(lib().og_is_legal(self._h, self._idx(action)))

    def is_eye(self, pos, owner):
        return bool(lib().og_is_eye(self._h, self._idx(pos), int(owner)))

    def is_eyeish(self, pos, owner):
        return bool(lib().og_is_eyeish(self._h, self._idx(pos), int(owner)))

    def is_ladder_capture(self, action):
        return bool(lib().og_is_ladder_capture(self._h, self._idx(action)))

    def is_ladder_escape(self, action):
        return bool(lib().og_is_ladder_escape(self._h, self._idx(action)))

    def get_legal_moves(self, include_eyes=True):
        buf = np.zeros(self.size * self.size, np.int16)
        c = lib().og_legal_moves(self._h, int(include_eyes), _p(buf, C.c_int16))
        return [divmod(int(m), self.size) for m in buf[:c]]

    def pack(self):
        n = self.size * self.size
        b = np.zeros(n, np.int8); a = np.zeros(n, np.uint8)
        pl = C.c_int8(); ko = C.c_int16()
        lib().og_pack(self._h, _p(b, C.c_int8), _p(a, C.c_uint8), C.byref(pl), C.byref(ko))
        return b, a, pl.value, ko.value

    def encode(self, feature_list=DEFAULT_FEATURES):
        b, a, pl, ko = self.pack()
        return encode_batch(b[None], a[None], [pl], [ko], self.size, feature_list)


def pack_gamestate(gs):
    """Pack any GameState-like object (reference or product) into encoder inputs,
    running the literal turns_since walk (preprocessing.py:20-43) over (history, board)."""
    size = gs.size
    board = np.asarray(gs.board).astype(np.int8).reshape(-1)
    hist = np.asarray([-1 if m is None else m[0] * size + m[1] for m in gs.history], dtype=np.int16)
    age = np.zeros(size * size, np.uint8)
    hp = _p(hist, C.c_int16) if len(hist) else C.POINTER(C.c_int16)()
    lib().og_age_from_history(size, _p(board, C.c_int8), hp, len(hist), _p(age, C.c_uint8))
    ko = -1 if gs.ko is None else gs.ko[0] * size + gs.ko[1]
    return board, age, int(gs.current_player), int(ko)
