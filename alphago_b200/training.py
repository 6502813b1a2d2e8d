"""SL training on the device network: the data-parallel step and the `run_training` entry point.

`SLTrainer` is what AlphaGo/training/supervised_policy_trainer.py gets from Keras'
`model.compile(SGD, categorical_crossentropy) + fit_generator` (spt.py:199-213): minibatch assembly
with the 8 board symmetries (spt.py:18-40,71-80) as a device gather, forward, softmax cross-entropy,
backward and the SGD update with Keras' `lr / (1 + decay * iterations)` schedule.  With
torch.distributed initialised (one process per GPU, NCCL over NVLink) every rank takes its share of
the global minibatch and the per-layer gradient slices are all-reduced while backward is still
running on the layers below.
"""
import ctypes as C

import numpy as np

from . import _lib, nn


def plan_buckets(n_conv, offsets, buckets=4):
    """Gradient exchange plan: [(layer_lo, layer_hi, start, end)] over the flat gradient buffer, in the order the
    slices become final during backward (last layers first).  `offsets` = n_conv+2 element offsets (layer starts,
    head start, end); the first bucket also carries the head's slice (ready right after the loss).  The LAST bucket is
    the first layer alone: nothing is left to overlap its all-reduce with, so it should be the smallest."""
    buckets = max(1, min(buckets, n_conv))
    tail = 1 if (buckets >= 2 and n_conv >= 3) else 0
    body = buckets - tail
    per = max(1, (n_conv - tail + body - 1) // body)
    plan, hi = [], n_conv
    while hi > tail:
        lo = max(tail, hi - per)
        end = offsets[-1] if not plan else offsets[hi]
        plan.append((lo, hi, offsets[lo], end))
        hi = lo
    if tail:
        plan.append((0, tail, offsets[0], offsets[tail] if plan else offsets[-1]))
    return plan


def plan_exchange(n_conv, offsets, buckets=5):
    """Exchange plan for the counter-gated path: [(layer_lo, layer_hi, start, end)] in the order the slices become final
    when the first layer's weight gradient is queued ahead of the chained launch: the first layer alone, then runs of
    layers top down (the first of them also carries the head's slice), and layers 1-2 last."""
    if n_conv < 4:
        return plan_buckets(n_conv, offsets, min(buckets, 2))
    # Every exchange kernel costs three cross-GPU barriers whatever its size (25-70 us from 2 to 8 GPUs): few buckets, and
    # the last one -- gated by the end of the chained launch, so fully exposed -- is the two lowest chained layers together
    # rather than two kernels of one layer each.
    plan = [(0, 1, offsets[0], offsets[1])]
    body = max(1, min(buckets - 2, n_conv - 3))          # buckets between the first layer's and the last two layers'
    per = (n_conv - 3 + body - 1) // body
    hi = n_conv
    while hi > 3:
        lo = max(3, hi - per)
        plan.append((lo, hi, offsets[lo], offsets[-1] if hi == n_conv else offsets[hi]))
        hi = lo
    plan.append((1, 3, offsets[1], offsets[3]))
    return plan


def allreduce_in_buckets(dist, flat, plan, group=None, after_bucket=None):
    """sum-all-reduce `flat` slice by slice following `plan`; `after_bucket(i)` (optional) runs before bucket i is
    issued -- SLTrainer enqueues that bucket's backward kernels there.  Returns the async work handles."""
    works = []
    for i, (lo, hi, start, end) in enumerate(plan):
        if after_bucket is not None:
            after_bucket(i, lo, hi)
        works.append(dist.all_reduce(flat[start:end], op=dist.ReduceOp.SUM, group=group, async_op=True))
    return works


class PeerExchange(object):
    """Sum all-reduce of ranges of a net's gradient buffer over NVLink peer memory (csrc/xchg.cu) for the ranks of one node:
    what SLTrainer uses instead of NCCL's ring when every rank can map every other rank's buffer.  Collective to build
    (the IPC handles travel through torch.distributed); raises AlphaGoB200Error when the peers cannot be mapped."""

    def __init__(self, net, dist, group=None, ctas=32):
        import torch
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        rec = (C.c_ubyte * 144)()
        self._h = _lib.lib().ag_xchg_create(self.rank, self.world, C.c_void_p(net.grad_ptr), net.grad_count, int(ctas), rec)
        ok = torch.tensor([1 if self._h else 0], device="cuda")
        mine = torch.frombuffer(bytearray(bytes(rec)), dtype=torch.uint8).cuda()
        every = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(every, mine, group=group)
        if self._h:
            table = b"".join(bytes(t.cpu().numpy().tobytes()) for t in every)
            if _lib.lib().ag_xchg_connect(self._h, table) != 0:
                ok[0] = 0
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)   # all ranks use it, or none
        if int(ok.item()) == 0:
            err = _lib.last_error()
            self.close()
            raise _lib.AlphaGoB200Error("peer gradient exchange unavailable: " + err)
        dist.barrier(group=group)

    def allreduce(self, start, end, stream):
        """in place over grads[start:end] (4-float aligned); asynchronous on `stream` (a torch.cuda.Stream)"""
        _lib.check(_lib.lib().ag_xchg_allreduce(self._h, int(start), int(end - start), C.c_void_p(stream.cuda_stream)), "ag_xchg_allreduce")

    def wait_readers(self, stream):
        """makes `stream` wait until every rank has finished reading this rank's buffer in the exchanges issued so far: to be
        queued on the stream that overwrites the buffer next, before it does (SLTrainer.step does so at its start)"""
        _lib.check(_lib.lib().ag_xchg_wait_readers(self._h, C.c_void_p(stream.cuda_stream)), "ag_xchg_wait_readers")

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            _lib.lib().ag_xchg_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SLTrainer(object):
    def __init__(self, net, lr=0.03, decay=0.0001, momentum=0.0, buckets=5, process_group=None):
        import torch
        self.torch = torch
        self.net, self.lr0, self.decay, self.momentum = net, float(lr), float(decay), float(momentum)
        self.iterations = 0
        self.loss_acc = torch.zeros(2, dtype=torch.float32, device="cuda")
        self.dist = None
        self.world = 1
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size(process_group) > 1:
                self.dist, self.world, self.group = dist, dist.get_world_size(process_group), process_group
        except ImportError:
            pass
        if self.dist is not None:
            # the all-reduce of bucket k runs on a few SMs while bucket k+1's weight gradients are computed: leave it room
            import os
            import torch as _t
            sms = _t.cuda.get_device_properties(_t.cuda.current_device()).multi_processor_count
            # AG_DP_XCHG: "peer" (default) = the library's two-shot all-reduce over mapped peer memory, NCCL if the peers cannot
            # be mapped; "nccl" = torch.distributed's all_reduce
            self.xchg = None
            if os.environ.get("AG_DP_XCHG", "peer") == "peer":
                try:
                    self.xchg = PeerExchange(net, self.dist, self.group, ctas=int(os.environ.get("AG_DP_XCHG_CTAS", "32")))
                except _lib.AlphaGoB200Error:
                    self.xchg = None
            # SMs kept out of the weight-gradient grids for the exchange running beside them: NCCL's kernels need SMs of their
            # own (4 measured best on 2 GPUs); the peer exchange has no shared memory and shares SMs with the chain
            reserve = os.environ.get("AG_DP_RESERVE_SMS", "0" if self.xchg is not None else "4")
            net.set_sm_budget(max(1, sms - int(reserve)) if int(reserve) > 0 else 0)
        # gradient buckets: contiguous runs of layers, last layers first (they finish first in backward)
        import os as _os
        self.plan = plan_buckets(net.n_conv, net.grad_offsets, int(_os.environ.get("AG_DP_BUCKETS", buckets)))
        # 1: each bucket's all-reduce is gated by the weight-gradient kernel's own per-layer "gradient final" counters
        # (cuStreamWaitValue32 on a side stream), so the exchange starts inside the one chained launch; 0: buckets are cut
        # at launch boundaries and gated by stream order
        self.flag_gated = _os.environ.get("AG_DP_FLAGS", "1") != "0"
        if self.dist is not None and self.flag_gated and getattr(self, "xchg", None) is not None:
            self.plan = plan_exchange(net.n_conv, net.grad_offsets, int(_os.environ.get("AG_DP_BUCKETS", buckets)))
        self._side = None
        self._out = None
        self._labels = None

    def current_lr(self):
        """Keras 1.0 SGD: lr * 1 / (1 + decay * iterations)"""
        return self.lr0 / (1.0 + self.decay * self.iterations)

    # ------------------------------------------------------------------ minibatch assembly
    def assemble(self, d_states, d_actions, index, transform):
        """dataset tensors resident on the device + host index/transform arrays -> (planes, labels)"""
        torch = self.torch
        n = len(index)
        planes, size = d_states.shape[1], d_states.shape[-1]
        if self._out is None or self._out.shape[0] != n:
            self._out = torch.empty((n, planes, size, size), dtype=torch.uint8, device="cuda")
            self._labels = torch.empty(n, dtype=torch.int32, device="cuda")
        # gather indices (int64) and symmetry codes (uint8) travel in ONE pinned staging buffer / one copy per step; two
        # staging buffers alternate so that the copy queued for step k is not overwritten while step k+1 is prepared
        if getattr(self, "_stage_n", None) != n:
            self._stage = [torch.empty(9 * n, dtype=torch.uint8).pin_memory() for _ in range(2)]
            self._stage_dev = [torch.empty(9 * n, dtype=torch.uint8, device="cuda") for _ in range(2)]
            self._stage_ev = [None, None]
            self._stage_n, self._stage_k = n, 0
        k = self._stage_k = self._stage_k ^ 1
        if self._stage_ev[k] is not None:
            self._stage_ev[k].synchronize()
        h = self._stage[k].numpy()
        h[:8 * n].view(np.int64)[:] = index
        h[8 * n:] = transform
        self._stage_dev[k].copy_(self._stage[k], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._stage_ev[k] = ev
        idx = self._stage_dev[k][:8 * n].view(torch.int64)
        tf = self._stage_dev[k][8 * n:]
        _lib.check(_lib.lib().ag_augment(C.c_void_p(d_states.data_ptr()), C.c_void_p(d_actions.data_ptr()),
                                         C.c_void_p(idx.data_ptr()), C.c_void_p(tf.data_ptr()), n, planes, size,
                                         C.c_void_p(self._out.data_ptr()), C.c_void_p(self._labels.data_ptr()),
                                         _lib.current_stream()), "ag_augment")
        return self._out, self._labels

    # ------------------------------------------------------------------ one optimisation step
    def step(self, planes, labels):
        """one SGD step on this rank's (planes uint8 (n,C,S,S), labels int32 (n,)); asynchronous.
        Loss sum / correct count of the step are accumulated into self.loss_acc."""
        net, n = self.net, planes.shape[0]
        lr = self.current_lr()
        if self.dist is None:
            net.train_step(planes, labels, lr, self.momentum, self.loss_acc)
        elif not self.flag_gated:
            net.forward_loss(planes, labels, self.loss_acc)

            def backward_bucket(i, lo, hi):
                net.backward_range(hi - 1, lo, n)
            # each bucket is reduced as soon as its layers' gradients are enqueued: NCCL's stream picks up after the
            # kernels issued so far, so the exchange overlaps the backward of the layers below
            works = allreduce_in_buckets(self.dist, net.grad_tensor(), self.plan, self.group, backward_bucket)
            for w in works:
                w.wait()
            net.sgd(lr, self.momentum, 1.0 / self.world)
        else:
            # The whole backward pass is queued at once (one data-gradient launch, one weight-gradient launch for the equal
            # layers, the first layer's own kernels).  The weight-gradient kernel counts, per layer, the CTAs that have
            # finished their share: a side stream waits for the count of a bucket's LOWEST layer (layers finish top down)
            # and NCCL's stream waits for the side stream, so bucket k is on the wire while the same launch is still
            # computing the layers below it.
            torch = self.torch
            if self.xchg is not None:   # the peers have finished reading last step's gradients before this step overwrites them
                self.xchg.wait_readers(torch.cuda.current_stream())
            net.forward_loss(planes, labels, self.loss_acc)
            net.backward_range(net.n_conv - 1, 0, n)
            flat = net.grad_tensor()
            if self._side is None:
                self._side = torch.cuda.Stream()
            works = []
            main = torch.cuda.current_stream()
            for lo, hi, start, end in self.plan:
                gate = net.wgrad_done(lo)
                if self.xchg is not None:
                    start4, end4 = start, end                      # layer slices and the buffer's end are whole 16-byte groups
                    if gate is None:
                        # the first layer: queued ahead of the chain, the library marks the point; else final when the main
                        # stream gets here
                        if not (lo == 0 and hi == 1 and _lib.lib().ag_net_wait_first_layer(net._h, C.c_void_p(self._side.cuda_stream)) == 0):
                            ev = torch.cuda.Event()
                            ev.record(main)
                            self._side.wait_event(ev)
                    else:
                        _lib.check(_lib.lib().ag_stream_wait_value(C.c_void_p(self._side.cuda_stream), C.c_void_p(gate[0]), gate[1]),
                                   "ag_stream_wait_value")
                    self.xchg.allreduce(start4, end4, self._side)
                    continue
                if gate is None:   # not a chained layer (the first layer, a refused launch): final when the stream gets there
                    works.append(self.dist.all_reduce(flat[start:end], op=self.dist.ReduceOp.SUM, group=self.group, async_op=True))
                    continue
                with torch.cuda.stream(self._side):
                    _lib.check(_lib.lib().ag_stream_wait_value(C.c_void_p(self._side.cuda_stream), C.c_void_p(gate[0]), gate[1]),
                               "ag_stream_wait_value")
                    works.append(self.dist.all_reduce(flat[start:end], op=self.dist.ReduceOp.SUM, group=self.group, async_op=True))
            for w in works:
                w.wait()
            if self.xchg is not None:
                ev = torch.cuda.Event()
                ev.record(self._side)
                main.wait_event(ev)
            net.sgd(lr, self.momentum, 1.0 / self.world)
        self.iterations += 1

    def evaluate(self, planes, labels):
        """loss / accuracy only (validation): accumulates into self.loss_acc, no update"""
        self.net.forward_loss(planes, labels, self.loss_acc)

    def pop_metrics(self, samples):
        """(mean loss, accuracy) over `samples` local samples since the last call; synchronises"""
        t = self.loss_acc.clone()
        self.loss_acc.zero_()
        cnt = self.torch.tensor([float(samples)], device="cuda")
        if self.dist is not None:
            self.dist.all_reduce(t, group=self.group)
            self.dist.all_reduce(cnt, group=self.group)
        t = t.cpu().numpy()
        c = float(cnt.cpu()[0])
        return float(t[0]) / max(c, 1.0), float(t[1]) / max(c, 1.0)


class HostFeeder(object):
    """Feeds an SLTrainer from a dataset that lives in HOST memory (numpy arrays, or HDF5 datasets read into them):
    the trainer-side half of the reference's `shuffled_hdf5_batch_generator` + `fit_generator` loop (spt.py:18-40,
    207-213) when the data does not fit, or is not wanted, on the device.

    Per step: gather the minibatch into one of two pinned staging buffers, copy it to the device on a copy stream,
    augment + train on the compute stream, and copy the running (loss sum, correct count) back.  Nothing waits for the
    step just queued: the copy of batch k overlaps the kernels of step k-1, and `step()` returns the metrics of the
    PREVIOUS step (None on the first call); `flush()` returns the last ones.

    `h_states` / `h_actions` are numpy arrays, or datasets that read rows from a file on demand (anything with
    `take_rows(index, out)`, e.g. hdf5_lite.Dataset: only the chunks holding the minibatch are decoded).  `stage(index)`
    does the gather of the NEXT step ahead of time -- host work only, safe on a worker thread while this thread queues
    the current step -- which is how `run_training` hides file reads behind the GPU.
    """

    def __init__(self, trainer, h_states, h_actions, batch, io_threads=0):
        import torch
        self.torch, self.trainer = torch, trainer
        self.h_states, self.h_actions, self.batch = h_states, h_actions, int(batch)
        shape = (self.batch,) + tuple(h_states.shape[1:])
        self.pin_x = [torch.empty(shape, dtype=torch.uint8).pin_memory() for _ in range(2)]
        self.pin_a = [torch.empty((self.batch, 2), dtype=torch.uint8).pin_memory() for _ in range(2)]
        self.pin_m = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.dx = [torch.empty(shape, dtype=torch.uint8, device="cuda") for _ in range(2)]
        self.da = [torch.empty((self.batch, 2), dtype=torch.uint8, device="cuda") for _ in range(2)]
        self.copy_stream = torch.cuda.Stream()
        self.ev_copied = [torch.cuda.Event() for _ in range(2)]
        self.ev_consumed = [torch.cuda.Event() for _ in range(2)]
        self.ev_metrics = [torch.cuda.Event() for _ in range(2)]
        self.ident = np.arange(self.batch)
        self.k = 0
        self.staged = 0      # minibatches gathered so far (at most one ahead of self.k)
        self.pool = None
        if io_threads > 0:
            from concurrent.futures import ThreadPoolExecutor
            self.pool = ThreadPoolExecutor(max_workers=io_threads)
        self.h2d_bytes = self.pin_x[0].numel() + self.pin_a[0].numel()
        self.d2h_bytes = 8

    def _metrics(self, slot):
        self.ev_metrics[slot].synchronize()
        return float(self.pin_m[slot][0]), float(self.pin_m[slot][1])

    def _gather(self, src, index, out):
        if hasattr(src, "take_rows"):
            src.take_rows(index, out, self.pool)
        elif isinstance(src, np.ndarray):
            np.take(src, index, axis=0, out=out)
        else:   # an h5py dataset: fancy indexing wants increasing, unique rows
            uniq, inverse = np.unique(np.asarray(index), return_inverse=True)
            out[...] = np.asarray(src[uniq.tolist()])[inverse]

    def stage(self, index):
        """gather minibatch number `self.staged` into its pinned staging buffers (host work only)"""
        if len(index) != self.batch:
            raise ValueError("HostFeeder was built for minibatches of %d, got %d" % (self.batch, len(index)))
        if self.staged > self.k + 1:
            raise RuntimeError("HostFeeder.stage: at most one minibatch ahead of step()")
        slot = self.staged & 1
        if self.staged >= 2:
            self.ev_copied[slot].synchronize()      # the staging buffer's previous copy has left the host
        self._gather(self.h_states, index, self.pin_x[slot].numpy())
        self._gather(self.h_actions, index, self.pin_a[slot].numpy())
        self.staged += 1

    def step(self, index, transform, train=True):
        """queue one step on minibatch `index` (already staged, or gathered now); train=False only evaluates it"""
        torch = self.torch
        slot = self.k & 1
        if len(index) != self.batch:
            raise ValueError("HostFeeder was built for minibatches of %d, got %d" % (self.batch, len(index)))
        if self.staged == self.k:
            self.stage(index)
        main = torch.cuda.current_stream()
        with torch.cuda.stream(self.copy_stream):
            if self.k >= 2:
                self.copy_stream.wait_event(self.ev_consumed[slot])  # the step that read dx[slot] two steps ago
            self.dx[slot].copy_(self.pin_x[slot], non_blocking=True)
            self.da[slot].copy_(self.pin_a[slot], non_blocking=True)
            self.ev_copied[slot].record(self.copy_stream)
        main.wait_event(self.ev_copied[slot])
        x, y = self.trainer.assemble(self.dx[slot], self.da[slot], self.ident, transform)
        self.ev_consumed[slot].record(main)
        if train:
            self.trainer.step(x, y)
        else:
            self.trainer.evaluate(x, y)
        self.pin_m[slot].copy_(self.trainer.loss_acc, non_blocking=True)
        self.ev_metrics[slot].record(main)
        self.k += 1
        return self._metrics(slot ^ 1) if self.k >= 2 else None

    def flush(self):
        """metrics after the last queued step (synchronises with it)"""
        return self._metrics((self.k - 1) & 1) if self.k else None
