"""GPU parity test of the stand-alone weight-gradient op (ag_wgrad) through the C ABI against the CPU
oracle (torch-CPU autograd of the same cross-correlation, fp32).  Inputs are bf16-representable, so the
only difference is fp32 summation order: relative Frobenius error < 1e-3."""
import ctypes as C

import numpy as np
import pytest

from alphago_b200 import _lib

pytestmark = pytest.mark.gpu


def padded_rows(t, row_stride, img_stride, ld):
    """torch (N, C, S, S) float -> bf16 padded-pixel-row matrix [N*img_stride][ld] on the GPU"""
    import torch
    n, c, s, _ = t.shape
    out = torch.zeros((n, img_stride, ld), dtype=torch.float32)
    idx = (torch.arange(s)[:, None] * row_stride + torch.arange(s)[None, :]).reshape(-1)
    out[:, idx, :c] = t.permute(0, 2, 3, 1).reshape(n, s * s, c)
    return out.reshape(n * img_stride, ld).to(torch.bfloat16).cuda()


def oracle_dw(x, g, width):
    import torch
    import torch.nn.functional as F
    w = torch.zeros((g.shape[1], x.shape[1], width, width), requires_grad=True)
    y = F.conv2d(x, w, padding=width // 2)
    (y * g).sum().backward()
    return w.grad.numpy()  # [co][ci][kx][ky]


def run_wgrad(fn, x, g, width, S, extra_scratch=None):
    import torch
    n, cin, cout = x.shape[0], x.shape[1], g.shape[1]
    pad = width // 2
    rs, ist = S + pad, (S + pad) * (S + pad)
    kin, ldout = (cin + 63) // 64 * 64, (cout + 63) // 64 * 64
    dx, dg = padded_rows(x, rs, ist, kin), padded_rows(g, rs, ist, ldout)
    dw = torch.full((width * width, ldout, kin), float("nan"), dtype=torch.float32, device="cuda")
    rc = fn(dx, dg, n * ist, ldout, kin, width, rs, dw)
    torch.cuda.synchronize()
    assert rc == 0, _lib.last_error()
    out = dw.cpu().numpy()
    assert np.isfinite(out).all()
    return out[:, :cout, :cin].transpose(1, 2, 0).reshape(cout, cin, width, width)


def via_ag_wgrad(scratch_splits=None):
    import torch

    def fn(dx, dg, rows, ldout, kin, width, rs, dw):
        L = _lib.lib()
        nbytes = L.ag_wgrad_scratch_bytes(width, ldout, kin, rows)
        assert nbytes > 0
        if scratch_splits is not None:
            nbytes = scratch_splits * L.ag_wgrad_partial_bytes(width, ldout, kin)
        scratch = torch.empty(nbytes // 4, dtype=torch.float32, device="cuda")
        return L.ag_wgrad(C.c_void_p(dg.data_ptr()), C.c_void_p(dx.data_ptr()), rows, ldout, kin, width, rs,
                          C.c_void_p(scratch.data_ptr()), nbytes, C.c_void_p(dw.data_ptr()), _lib.current_stream())
    return fn


def rel(a, b):
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


CASES = [(19, 6, 192, 192, 3), (19, 3, 48, 192, 5), (13, 5, 64, 64, 3), (7, 9, 20, 100, 1), (19, 2, 128, 256, 3)]


def make_inputs(S, n, cin, cout, seed):
    import torch
    g0 = torch.Generator().manual_seed(seed)
    x = (torch.rand((n, cin, S, S), generator=g0) < 0.4).float() * torch.randint(1, 4, (n, cin, S, S), generator=g0).float()
    g = torch.randn((n, cout, S, S), generator=g0).to(torch.bfloat16).float()
    return x, g


@pytest.mark.parametrize("S,n,cin,cout,width", CASES)
def test_wgrad_matches_oracle(S, n, cin, cout, width):
    x, g = make_inputs(S, n, cin, cout, 1)
    want = oracle_dw(x, g, width)
    got = run_wgrad(via_ag_wgrad(), x, g, width, S)
    assert rel(got, want) < 1e-3
    # a scratch that only fits one split gives the same sums in a different association
    got1 = run_wgrad(via_ag_wgrad(scratch_splits=1), x, g, width, S)
    assert rel(got1, want) < 1e-3


def test_wgrad_rejects_bad_arguments():
    import torch
    L = _lib.lib()
    z = torch.zeros(64 * 64, dtype=torch.float32, device="cuda")
    p = C.c_void_p(z.data_ptr())
    assert L.ag_wgrad(p, p, 400, 64, 64, 3, 20, p, 16, p, None) != 0     # scratch smaller than one partial
    assert "scratch" in _lib.last_error()
    assert L.ag_wgrad(p, p, 400, 60, 64, 3, 20, p, 1 << 20, p, None) != 0  # unpadded channels
    assert L.ag_wgrad(p, p, 400, 64, 64, 4, 20, p, 1 << 20, p, None) != 0  # even width
