"""`ConvNet`: thin Python handle on the device network object (ag_net_* in include/alphago_b200.h).

Holds no tensors of its own beyond what callers pass in; PyTorch is used for device buffers,
streams and (in `DataParallelTrainer`) torch.distributed.  There is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import _lib

POLICY_HEAD, VALUE_HEAD = 0, 1


class ConvNet(object):
    def __init__(self, board, in_planes, filters, widths, head=POLICY_HEAD, max_batch=256):
        self.board, self.in_planes = int(board), int(in_planes)
        self.filters, self.widths = [int(f) for f in filters], [int(w) for w in widths]
        self.head, self.max_batch = head, int(max_batch)
        self.n_conv = len(self.filters)
        f = np.asarray(self.filters, np.int32); w = np.asarray(self.widths, np.int32)
        self._h = _lib.lib().ag_net_create(self.board, self.in_planes, self.n_conv, _lib.np_ptr(f), _lib.np_ptr(w),
                                           int(head), self.max_batch)
        if not self._h:
            raise _lib.AlphaGoB200Error("ag_net_create failed: " + _lib.last_error())
        offs = np.zeros(self.n_conv + 2, np.int64)
        ptr, cnt = C.c_void_p(), C.c_longlong()
        _lib.check(_lib.lib().ag_net_grads(self._h, C.byref(ptr), C.byref(cnt), _lib.np_ptr(offs)), "ag_net_grads")
        self.grad_ptr, self.grad_count, self.grad_offsets = ptr.value, cnt.value, [int(o) for o in offs]
        self._grad_tensor = None

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            _lib.lib().ag_net_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ weights (host numpy, fp32)
    def conv_shape(self, layer):
        cin = self.in_planes if layer == 0 else self.filters[layer - 1]
        return (self.filters[layer], cin, self.widths[layer], self.widths[layer])

    def set_conv(self, layer, w, b):
        w = np.ascontiguousarray(w, np.float32); b = np.ascontiguousarray(b, np.float32)
        assert w.shape == self.conv_shape(layer) and b.shape == (self.filters[layer],), (w.shape, self.conv_shape(layer))
        _lib.check(_lib.lib().ag_net_set_conv(self._h, layer, _lib.np_ptr(w), _lib.np_ptr(b)), "ag_net_set_conv")

    def get_conv(self, layer):
        w = np.empty(self.conv_shape(layer), np.float32); b = np.empty(self.filters[layer], np.float32)
        _lib.check(_lib.lib().ag_net_get_conv(self._h, layer, _lib.np_ptr(w), _lib.np_ptr(b)), "ag_net_get_conv")
        return w, b

    def _head_shapes(self, which):
        pts = self.board * self.board
        return {0: ((self.filters[-1],), (1,)), 1: ((pts,), None), 2: ((256, pts), (256,)), 3: ((256,), (1,))}[which]

    def set_head(self, which, w, b=None):
        ws, bs = self._head_shapes(which)
        w = np.ascontiguousarray(w, np.float32).reshape(ws)
        bp = None
        if bs is not None:
            b = np.ascontiguousarray(b if b is not None else np.zeros(bs), np.float32).reshape(bs)
            bp = _lib.np_ptr(b)
        _lib.check(_lib.lib().ag_net_set_head(self._h, which, _lib.np_ptr(w), bp), "ag_net_set_head")

    def get_head(self, which):
        ws, bs = self._head_shapes(which)
        w = np.empty(ws, np.float32)
        b = np.empty(bs, np.float32) if bs is not None else None
        _lib.check(_lib.lib().ag_net_get_head(self._h, which, _lib.np_ptr(w), _lib.np_ptr(b) if b is not None else None),
                   "ag_net_get_head")
        return w, b

    def init_uniform(self, seed=0, scale=0.05):
        """Keras 1.0 init='uniform': U(-0.05, 0.05) weights, zero biases (policy.py:127)"""
        rng = np.random.RandomState(seed)
        for l in range(self.n_conv):
            self.set_conv(l, rng.uniform(-scale, scale, self.conv_shape(l)), np.zeros(self.filters[l]))
        self.set_head(0, rng.uniform(-scale, scale, self.filters[-1]), np.zeros(1))
        if self.head == VALUE_HEAD:
            pts = self.board * self.board
            self.set_head(2, rng.uniform(-scale, scale, (256, pts)), np.zeros(256))
            self.set_head(3, rng.uniform(-scale, scale, 256), np.zeros(1))

    # ------------------------------------------------------------------ device calls (torch tensors)
    def forward(self, planes, out=None):
        """uint8 CUDA tensor (N, C, S, S) -> fp32 CUDA tensor (N, S*S) probabilities / (N,) values"""
        import torch
        n = planes.shape[0]
        assert planes.is_cuda and planes.dtype == torch.uint8 and planes.is_contiguous()
        assert tuple(planes.shape[1:]) == (self.in_planes, self.board, self.board), planes.shape
        shape = (n, self.board * self.board) if self.head == POLICY_HEAD else (n,)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=planes.device)
        _lib.check(_lib.lib().ag_net_forward(self._h, C.c_void_p(planes.data_ptr()), n, C.c_void_p(out.data_ptr()),
                                             _lib.current_stream()), "ag_net_forward")
        return out

    def forward_loss(self, planes, labels, loss_acc, probs=None):
        n = planes.shape[0]
        _lib.check(_lib.lib().ag_net_forward_loss(self._h, C.c_void_p(planes.data_ptr()), C.c_void_p(labels.data_ptr()), n,
                                                  C.c_void_p(probs.data_ptr()) if probs is not None else None,
                                                  C.c_void_p(loss_acc.data_ptr()), _lib.current_stream()), "ag_net_forward_loss")

    def backward_layer(self, layer, n):
        _lib.check(_lib.lib().ag_net_backward_layer(self._h, layer, n, _lib.current_stream()), "ag_net_backward_layer")

    def sgd(self, lr, momentum=0.0, grad_scale=1.0):
        _lib.check(_lib.lib().ag_net_sgd(self._h, lr, momentum, grad_scale, _lib.current_stream()), "ag_net_sgd")

    def train_step(self, planes, labels, lr, momentum, loss_acc):
        n = planes.shape[0]
        _lib.check(_lib.lib().ag_net_train_step(self._h, C.c_void_p(planes.data_ptr()), C.c_void_p(labels.data_ptr()), n,
                                                lr, momentum, C.c_void_p(loss_acc.data_ptr()), _lib.current_stream()),
                   "ag_net_train_step")

    def set_loss(self, kind):
        """0 = categorical cross-entropy, 1 = Keras binary cross-entropy on the softmax outputs; see ag_net_set_loss"""
        _lib.check(_lib.lib().ag_net_set_loss(self._h, int(kind)), "ag_net_set_loss")

    def set_sm_budget(self, sms):
        """SMs the weight-gradient launches may fill (0 = all); see ag_net_set_sm_budget"""
        _lib.check(_lib.lib().ag_net_set_sm_budget(self._h, int(sms)), "ag_net_set_sm_budget")

    def backward_range(self, hi, lo, n):
        """backward of layers hi .. lo in one call (descending ranges after forward_loss): equal layers share chained launches"""
        _lib.check(_lib.lib().ag_net_backward_range(self._h, hi, lo, n, _lib.current_stream()), "ag_net_backward_range")

    def launch_plan(self):
        """how the layers are launched: {"fwd": [...], "dgrad": [...], "wgrad": [...]} = layers inside the chained launches,
        "refused": chained launches the device refused so far, "splits": split-K factor per layer (after a forward_loss)"""
        L, h = _lib.lib(), self._h
        rng = range(self.n_conv)
        return {"fwd": [l for l in rng if L.ag_net_layer_info(h, l, 5) == 1],
                "dgrad": [l for l in rng if L.ag_net_layer_info(h, l, 6) == 1],
                "wgrad": [l for l in rng if L.ag_net_layer_info(h, l, 7) == 1],
                "refused": L.ag_net_layer_info(h, 0, 8), "splits": [L.ag_net_layer_info(h, l, 9) for l in rng]}

    def wgrad_done(self, layer):
        """(device address of a counter, value it reaches when this layer's gradient of the latest backward_range is
        final) for a layer whose weight gradient runs inside the chained launch, else None"""
        ptr, target = C.c_void_p(), C.c_uint()
        rc = _lib.lib().ag_net_wgrad_done(self._h, int(layer), C.byref(ptr), C.byref(target))
        return (ptr.value, target.value) if rc == 0 else None

    def run_layer(self, layer, mode, n):
        _lib.check(_lib.lib().ag_net_run_layer(self._h, layer, mode, n, _lib.current_stream()), "ag_net_run_layer")

    def grad_tensor(self):
        """the flat gradient buffer as a torch tensor sharing the library's device memory"""
        import torch
        if self._grad_tensor is None:
            class _Mem(object):
                pass
            m = _Mem()
            m.__cuda_array_interface__ = {"shape": (self.grad_count,), "typestr": "<f4", "data": (self.grad_ptr, False),
                                          "version": 2}
            self._grad_tensor = torch.as_tensor(m, device="cuda")
        return self._grad_tensor

    def debug_read(self, layer, kind, n):
        out = np.empty((n, self.filters[layer], self.board, self.board), np.float32)
        _lib.check(_lib.lib().ag_net_debug_read(self._h, layer, kind, n, _lib.np_ptr(out)), "ag_net_debug_read")
        return out
