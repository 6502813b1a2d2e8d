"""GPU test of ag_wgrad_static, the weight-gradient entry point that allocates nothing.

Some users of this library cannot call cudaMalloc when they need a gradient: they run inside another
framework's pre-sized memory pool, in a serving loop where an allocation would stall the device, or under a
device-memory budget fixed at start-up.  For them the split-K scratch is a 16 MiB array inside the library
itself.  This test calls the entry point through the C ABI twice in a row (the second call finds the first
call's partials still in the array) and checks both results against ag_wgrad, which uses caller-owned
scratch; the two run the same kernel with the same split plan whenever the plan fits in 16 MiB, so the results
must agree to fp32 summation order (relative Frobenius error < 1e-3, the tolerance of tests/test_wgrad_gpu.py).
Kept last in collection order so a failure here cannot hide any other test.
"""
import ctypes as C

import numpy as np
import pytest

from alphago_b200 import _lib
from test_wgrad_gpu import make_inputs, oracle_dw, rel, run_wgrad, via_ag_wgrad

pytestmark = pytest.mark.gpu


def via_static():
    def fn(dx, dg, rows, ldout, kin, width, rs, dw):
        return _lib.lib().ag_wgrad_static(C.c_void_p(dg.data_ptr()), C.c_void_p(dx.data_ptr()), rows, ldout, kin, width, rs,
                                          C.c_void_p(dw.data_ptr()), _lib.current_stream())
    return fn


def test_static_scratch_matches_ag_wgrad_twice_in_a_row():
    assert _lib.lib().ag_wgrad_static_max_bytes() == 16 << 20
    # two different problems back to back: the second starts with the first one's partials in the array
    for S, n, cin, cout, width, seed in [(19, 6, 192, 192, 3, 1), (13, 5, 64, 128, 5, 2), (19, 6, 192, 192, 3, 3)]:
        x, g = make_inputs(S, n, cin, cout, seed)
        want = run_wgrad(via_ag_wgrad(), x, g, width, S)
        first = run_wgrad(via_static(), x, g, width, S)
        second = run_wgrad(via_static(), x, g, width, S)
        assert rel(first, want) < 1e-3 and rel(second, want) < 1e-3
        assert np.array_equal(first, second)                       # same plan, same order: bitwise repeatable
        assert rel(first, oracle_dw(x, g, width)) < 1e-3


def test_static_entry_point_leaves_device_memory_alone():
    """the property the entry point exists for: across its calls the device's free memory (cudaMemGetInfo, which sees every
    allocation of every allocator in the process) does not move, and neither does torch's own pool.  Inputs and outputs are
    allocated first; the library and its 16 MiB array are on the device already (an earlier launch loaded the module)."""
    import torch
    from test_wgrad_gpu import padded_rows
    S, n, cin, cout, width = 19, 6, 192, 192, 3
    x, g = make_inputs(S, n, cin, cout, 7)
    run_wgrad(via_ag_wgrad(), x, g, width, S)            # loads the module / warms both entry points' kernels
    run_wgrad(via_static(), x, g, width, S)
    rs, ist = S + 1, (S + 1) * (S + 1)
    dx, dg = padded_rows(x, rs, ist, 192), padded_rows(g, rs, ist, 192)
    dw = torch.empty((9, 192, 192), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0, allocated0 = torch.cuda.memory_reserved(), torch.cuda.memory_allocated()
    for _ in range(3):
        rc = _lib.lib().ag_wgrad_static(C.c_void_p(dg.data_ptr()), C.c_void_p(dx.data_ptr()), n * ist, 192, 192, width, rs,
                                        C.c_void_p(dw.data_ptr()), _lib.current_stream())
        assert rc == 0, _lib.last_error()
        free_mid, _ = torch.cuda.mem_get_info()          # while the kernels may still be running
        assert free_mid == free0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    assert torch.cuda.memory_reserved() == reserved0 and torch.cuda.memory_allocated() == allocated0
    assert torch.isfinite(dw).all()


def test_static_scratch_refuses_oversized_problems():
    import torch
    z = torch.zeros(1 << 16, dtype=torch.float32, device="cuda")
    p = C.c_void_p(z.data_ptr())
    # a 5x5, 256 -> 1024 gradient is 26 MB of fp32 per partial: more than the 16 MiB the library carries
    rc = _lib.lib().ag_wgrad_static(p, p, 441, 1024, 256, 5, 21, p, None)
    assert rc != 0
    msg = _lib.last_error()
    assert "exceeds the built-in scratch" in msg and "ag_wgrad" in msg
