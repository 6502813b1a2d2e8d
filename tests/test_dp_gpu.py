"""Multi-GPU test of data-parallel SL training (needs >= 2 GPUs; skipped otherwise): two ranks, each stepping on
half of a global minibatch with the bucketed NCCL gradient all-reduce, must end with the same weights as one rank
stepping on the whole minibatch (mean-of-means = global mean for equal shares).  Tolerance: the two runs sum the
same bf16 products in a different order -> relative Frobenius error of the weight UPDATE < 2e-2 per tensor."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


SHAPES = {"small": ([64, 64, 64], [5, 3, 3]), "bench": ([192] * 12, [5] + [3] * 11)}   # "bench": BASELINE configs[2]/[3]


def _setup_net(max_batch, shape="small"):
    from alphago_b200 import nn
    from oracle import nn_oracle as no
    filters, widths = SHAPES[shape]
    params = no.make_params(48, filters, widths, seed=21)
    params["head_w"] = (params["head_w"] * 20).astype(np.float32)
    net = nn.ConvNet(19, 48, filters, widths, max_batch=max_batch)
    for l, (w, b) in enumerate(params["convs"]):
        net.set_conv(l, w, b)
    net.set_head(0, params["head_w"], params["head_b"])
    return net, params


def _data(n):
    rng = np.random.RandomState(5)
    return (rng.rand(n, 48, 19, 19) < 0.3).astype(np.uint8), rng.randint(0, 361, n).astype(np.int32)


def _worker(rank, world, port, out, shape):
    import torch
    import torch.distributed as dist
    from alphago_b200.training import SLTrainer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    n = 32
    x, y = _data(n)
    share = n // world
    net, _ = _setup_net(share, shape)
    tr = SLTrainer(net, lr=0.05, decay=0.0, momentum=0.9)
    assert tr.world == world and tr.flag_gated
    # the peer-memory exchange on its own first: every rank holds (rank + 1) * pattern, the sum is world (world + 1) / 2 * pattern
    peer = tr.xchg is not None
    if peer:
        flat = net.grad_tensor()
        pattern = (torch.arange(flat.numel(), device="cuda", dtype=torch.float32) % 1021) - 510.0
        side = torch.cuda.Stream()
        for start, end in ((0, flat.numel()), (net.grad_offsets[1], net.grad_offsets[2]), (net.grad_offsets[-2], flat.numel())):
            tr.xchg.wait_readers(torch.cuda.current_stream())   # the peers are done with the previous round's values
            flat.copy_(pattern * (rank + 1))
            torch.cuda.synchronize()
            dist.barrier()
            tr.xchg.allreduce(start, end, side)
            side.synchronize()
            want = pattern * (rank + 1)
            want[start:end] = pattern[start:end] * (world * (world + 1) / 2)
            assert torch.equal(flat, want), (rank, start, end)
        tr.xchg.wait_readers(torch.cuda.current_stream())
        flat.zero_()
        torch.cuda.synchronize()
        dist.barrier()
    sl = slice(rank * share, (rank + 1) * share)
    for _ in range(2):
        tr.step(torch.from_numpy(x[sl]).cuda(), torch.from_numpy(y[sl]).cuda())
    loss, acc = tr.pop_metrics(2 * share)
    torch.cuda.synchronize()
    if rank == 0:
        # the exchange of the chained layers was gated by their "gradient final" counters, not by launch boundaries
        gated = [net.wgrad_done(l) is not None for l in range(net.n_conv)]
        np.savez(out, loss=loss, gated=np.asarray(gated), peer=np.asarray(peer), **{"w%d" % l: net.get_conv(l)[0] for l in range(net.n_conv)},
                 hw=net.get_head(0)[0])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("shape", ["small", "bench"])
def test_two_rank_step_equals_single_rank_step(tmp_path, shape):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from alphago_b200.training import SLTrainer
    out = os.path.join(str(tmp_path), "dp.npz")
    port = 29700 + os.getpid() % 200
    mp.spawn(_worker, args=(2, port, out, shape), nprocs=2, join=True)
    dp = np.load(out)
    n = 32
    x, y = _data(n)
    net, params = _setup_net(n, shape)
    tr = SLTrainer(net, lr=0.05, decay=0.0, momentum=0.9)
    for _ in range(2):
        tr.step(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
    loss, _ = tr.pop_metrics(2 * n)
    assert abs(loss - float(dp["loss"])) < 2e-3 * max(1.0, abs(loss))
    assert bool(dp["peer"])   # the ranks of one node exchange gradients through mapped peer memory (csrc/xchg.cu), not NCCL
    if shape == "bench":   # every hidden layer ran chained and gave a counter to wait for
        assert dp["gated"][1:].all() and not dp["gated"][0], dp["gated"]
    for l in range(net.n_conv):
        w = net.get_conv(l)[0]
        w0 = params["convs"][l][0]
        err = np.linalg.norm((w - w0) - (dp["w%d" % l] - w0)) / np.linalg.norm(w - w0)
        # the 12-layer net's gradient is ill-conditioned on a random init (tests/test_bench_shape_gpu.py): the second step
        # starts from weights that differ in the last bits and its gradient moves by a few per cent
        assert err < (5e-2 if shape == "bench" else 2e-2), (l, err)
    hw = net.get_head(0)[0]
    assert np.linalg.norm(hw - dp["hw"]) / np.linalg.norm(hw - params["head_w"]) < 2e-2
