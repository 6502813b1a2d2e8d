"""ctypes binding of libalphago_b200.so (include/alphago_b200.h).

There is no CPU fallback: if the library is missing or a CUDA entry point fails, callers get an
exception, never a silently different implementation.
"""
import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libalphago_b200.so")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "alphago_b200.h")


class AlphaGoB200Error(RuntimeError):
    pass


_lib = None


def header_symbols():
    """every function name declared in include/alphago_b200.h"""
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ag_[a-z0-9_]+)\s*\(", txt)))


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AlphaGoB200Error(
            "libalphago_b200.so is not built (%s). Run `python -m alphago_b200.build`; "
            "there is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i8p, u8p, i16p, i32p = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p
    L.ag_last_error.restype = C.c_char_p
    L.ag_launch_count.restype = C.c_longlong
    L.ag_feature_planes.argtypes = [C.c_int]
    L.ag_encode.argtypes = [i8p, u8p, i8p, i16p, C.c_int, C.c_int, i32p, C.c_int, u8p, vp]
    L.ag_state_new.restype = vp; L.ag_state_new.argtypes = [C.c_int]
    L.ag_state_copy.restype = vp; L.ag_state_copy.argtypes = [vp]
    L.ag_state_free.argtypes = [vp]; L.ag_state_free.restype = None
    L.ag_state_do_move.argtypes = [vp, C.c_int, C.c_int]
    L.ag_state_is_legal.argtypes = [vp, C.c_int]
    L.ag_state_is_eyeish.argtypes = [vp, C.c_int, C.c_int]
    L.ag_state_is_eye.argtypes = [vp, C.c_int, C.c_int]
    L.ag_state_legal_moves.argtypes = [vp, C.c_int, i16p]
    L.ag_state_winner.argtypes = [vp, C.c_double]
    L.ag_state_get.argtypes = [vp, C.c_int]
    L.ag_state_set_player.argtypes = [vp, C.c_int]
    L.ag_state_set_ko.argtypes = [vp, C.c_int]
    L.ag_state_board.argtypes = [vp, i8p]; L.ag_state_board.restype = None
    L.ag_state_history.argtypes = [vp, i16p]; L.ag_state_history.restype = None
    L.ag_state_groups.argtypes = [vp, i16p, i16p, i16p]; L.ag_state_groups.restype = None
    L.ag_state_pack.argtypes = [vp, i8p, u8p, i8p, i16p]; L.ag_state_pack.restype = None
    L.ag_state_replay_pack.argtypes = [vp, C.c_int, i16p, i8p, i8p, u8p, i8p, i16p, C.POINTER(C.c_int)]
    L.ag_age_from_history.argtypes = [C.c_int, i8p, i16p, C.c_int, u8p]; L.ag_age_from_history.restype = None
    L.ag_playout_positions.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, i8p, u8p, i8p, i16p]
    L.ag_lzf_decompress.restype = C.c_longlong
    L.ag_lzf_decompress.argtypes = [C.c_char_p, C.c_size_t, vp, C.c_size_t]
    L.ag_lzf_compress.restype = C.c_longlong
    L.ag_lzf_compress.argtypes = [vp, C.c_size_t, vp, C.c_size_t]
    _bind_optional(L)
    _lib = L
    return L


def _bind_optional(L):
    """entry points added by later build stages bind themselves here (see nn bindings)"""
    from . import _nn_bind
    _nn_bind.bind(L)


def check(rc, what=""):
    if rc != 0:
        msg = lib().ag_last_error().decode("utf-8", "replace")
        raise AlphaGoB200Error("%s failed (%d): %s" % (what or "libalphago_b200 call", rc, msg))


def last_error():
    return lib().ag_last_error().decode("utf-8", "replace")


def np_ptr(a):
    return C.c_void_p(a.ctypes.data)


def current_stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
