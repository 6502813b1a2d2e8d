"""Import the UNMODIFIED reference (go.py, preprocessing.py) under Python 3.12.

Build-container only: /root/reference does not exist on the GPU box, so nothing that runs
there may import this module.  Used by oracle/gen_golden.py (fixture generation) and by the
CPU tests that pin the oracle against the live reference (skipped when it is absent).

Three namespace shims, applied here and not by editing the reference:
  np.int / np.float  -- removed numpy aliases used by go.py:40
  AlphaGo.go.filter  -- Python 2 `filter` returned a list; go.py:242 calls len() on it
"""
import builtins
import os
import sys

import numpy as np

REFERENCE_ROOT = os.environ.get("ALPHAGO_REFERENCE", "/root/reference")


def available():
    return os.path.exists(os.path.join(REFERENCE_ROOT, "AlphaGo", "go.py"))


def load():
    """returns (go module, preprocessing module)"""
    if not available():
        raise ImportError("reference tree not present at %s" % REFERENCE_ROOT)
    if not hasattr(np, "int"):
        np.int = int
    if not hasattr(np, "float"):
        np.float = float
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import AlphaGo.go as go
    go.filter = lambda f, it: list(builtins.filter(f, it))
    import AlphaGo.preprocessing.preprocessing as pp
    return go, pp


def dealias_copy(gs):
    """GameState.copy() aliases `history` (go.py:167); give the copy its own list."""
    c = gs.copy()
    c.history = list(gs.history)
    return c
