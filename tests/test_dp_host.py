"""CPU tests of the data-parallel exchange logic (world_size 2, gloo): the bucket plan covers the flat
gradient buffer exactly once in backward order, and reducing bucket by bucket then scaling by 1/world
equals the mean of the ranks' gradients -- what SLTrainer.step does between backward and the SGD kernel."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from alphago_b200.training import plan_buckets, allreduce_in_buckets


def _offsets(n_conv):
    offs, o = [], 0
    for l in range(n_conv):
        offs.append(o)
        o += (25 if l == 0 else 9) * 192 * (64 if l == 0 else 192) + 192
    offs.append(o)
    offs.append(o + 625)
    return offs


def test_bucket_plan_covers_buffer_once():
    for n_conv in (1, 2, 5, 12, 13):
        for buckets in (1, 3, 4, 20):
            offs = _offsets(n_conv)
            plan = plan_buckets(n_conv, offs, buckets)
            cover = np.zeros(offs[-1], np.int32)
            layers = []
            for lo, hi, s, e in plan:
                cover[s:e] += 1
                layers += list(range(hi - 1, lo - 1, -1))
            assert (cover == 1).all()
            assert layers == list(range(n_conv - 1, -1, -1))       # backward order
            assert plan[0][3] == offs[-1] and plan[0][1] == n_conv  # head slice rides with the first bucket
            assert len(plan) <= max(1, min(buckets, n_conv))


def _worker(rank, world, port, n_conv, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    offs = _offsets(n_conv)
    rng = np.random.RandomState(100 + rank)
    flat = torch.from_numpy(rng.randn(offs[-1]).astype(np.float32))
    ref = flat.clone()
    order = []
    works = allreduce_in_buckets(dist, flat, plan_buckets(n_conv, offs, 4), None, lambda i, lo, hi: order.append((lo, hi)))
    for w in works:
        w.wait()
    dist.all_reduce(ref)
    ok = bool(torch.equal(flat, ref)) and order == [(lo, hi) for lo, hi, _, _ in plan_buckets(n_conv, offs, 4)]
    mean = flat / world
    q.put((rank, ok, float(mean.abs().sum())))
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_allreduce_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 5, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert abs(res[0][2] - res[1][2]) < 1e-3  # both ranks hold the same averaged gradient


def test_exchange_plan_for_the_counter_gated_path():
    """plan_exchange: the first layer's bucket first (its gradient is queued ahead of the chained launch), the chain's layers
    top down with the head's slice in the first of them, layers 1-2 together last; every float covered exactly once"""
    from alphago_b200.training import plan_exchange
    for n_conv in (3, 4, 5, 12, 13):
        for buckets in (2, 3, 5, 6, 20):
            offs = _offsets(n_conv)
            plan = plan_exchange(n_conv, offs, buckets)
            cover = np.zeros(offs[-1], np.int32)
            for lo, hi, s, e in plan:
                assert 0 <= lo < hi <= n_conv and s == offs[lo]
                cover[s:e] += 1
            assert (cover == 1).all(), (n_conv, buckets)
            if n_conv >= 4:
                assert plan[0][:2] == (0, 1) and plan[-1][:2] == (1, 3)
                assert plan[1][1] == n_conv and plan[1][3] == offs[-1]          # the head's slice rides with the top layers
                tops = [hi for lo, hi, _, _ in plan[1:]]
                assert tops == sorted(tops, reverse=True)                        # top down: the order the chain finishes them
                assert len(plan) <= max(3, buckets)
