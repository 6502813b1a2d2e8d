"""GPU parity tests AT THE SHAPE bench.py TIMES (BASELINE configs[2]): the SL policy net with 12 KxK convs x 192 filters
+ the 1x1 head, 48 planes, batch 256, positions from the CUDA encoder -- i.e. the chained forward launch, the 11-layer
data-gradient chain, the 11-layer chained weight gradients with their cross-layer split reduction, the taps-in-N
first-layer weight gradient at K = 256 * 441 rows and the one-launch SGD, all against oracle/nn_oracle.py
(supervised_policy_trainer.py:199-213 on policy.py:119-154).

Tolerances (bf16 operands, fp32 accumulation; the oracle rounds to bf16 at the same points, so what is left is summation
order): loss within 2e-3 relative, correct-count exact up to ties (<= 1), every gradient tensor relative Frobenius error
< 3e-2 and cosine > 0.999, the SGD update the same against -lr * oracle gradient.
"""
import numpy as np
import pytest

from alphago_b200 import go, nn
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES
from oracle import nn_oracle as no
from test_nn_gpu import cosine, grads_from_flat, make_net, rel_err

pytestmark = pytest.mark.gpu

FILTERS, WIDTHS, BATCH = [192] * 12, [5] + [3] * 11, 256


@pytest.fixture(scope="module")
def problem():
    """256 random-playout positions through the CUDA encoder, labels uniform over the legal moves, the shared
    U(-0.05, 0.05) init with a sharpened head, and the oracle's loss / gradients for them (computed once)"""
    import torch
    b, a, pl, ko = go.playout_positions(2016, BATCH, size=19, max_moves=400, per_game=16)
    planes = Preprocess(DEFAULT_FEATURES).encode_packed(*[torch.from_numpy(x).cuda() for x in (b, a, pl, ko)], 19)
    x = planes.cpu().numpy()
    rng = np.random.RandomState(3)
    legal = x[:, 20:28].sum(1).reshape(BATCH, 361) > 0
    labels = np.asarray([rng.choice(np.flatnonzero(l)) if l.any() else 0 for l in legal], np.int32)
    params = no.make_params(48, FILTERS, WIDTHS, seed=1, bias_scale=0.02)
    params["head_w"] = (params["head_w"] * 20).astype(np.float32)
    loss, acc, grads, out = no.loss_and_grads(params, x, labels, emulate_bf16=True)
    return {"planes": planes, "x": x, "labels": labels, "params": params, "loss": loss, "acc": acc, "grads": grads,
            "probs": out["probs"].detach().numpy()}


def names():
    out = []
    for l in range(len(FILTERS)):
        out += ["conv%d.w" % l, "conv%d.b" % l]
    return out + ["head.w", "head.b"]


def check_grads(net, problem):
    got = grads_from_flat(net, net.grad_tensor().cpu().numpy())
    worst = (0.0, "")
    for name, g, want in zip(names(), got, problem["grads"]):
        assert g.shape == want.shape, name
        if name == "head.b":   # sum of (p - y) over all moves: zero up to rounding
            assert abs(float(g[0])) < 1e-5 and abs(float(want[0])) < 1e-5
            continue
        r, c = rel_err(g, want), cosine(g, want)
        worst = max(worst, (r, name))
        assert r < 3e-2 and c > 0.999, "%s: rel %g cos %g" % (name, r, c)
    print("worst gradient tensor: %s rel %.4g" % (worst[1], worst[0]))


def check_loss(acc, problem):
    got_loss, got_correct = acc.cpu().numpy()
    assert abs(got_loss / BATCH - problem["loss"]) <= 2e-3 * max(1.0, abs(problem["loss"]))
    assert abs(got_correct - round(problem["acc"] * BATCH)) <= 1


def test_train_step_at_the_benchmarked_shape(problem):
    import torch
    from alphago_b200 import _lib
    net = make_net(19, 48, FILTERS, WIDTHS, problem["params"], max_batch=BATCH)
    labels = torch.from_numpy(problem["labels"]).cuda()
    acc = torch.zeros(2, device="cuda")
    probs = torch.empty(BATCH, 361, device="cuda")
    l0 = _lib.lib().ag_launch_count()
    net.forward_loss(problem["planes"], labels, acc, probs)
    net.backward_range(len(FILTERS) - 1, 0, BATCH)
    launches = _lib.lib().ag_launch_count() - l0
    torch.cuda.synchronize()
    # what ran is what the benchmark runs: every layer inside the chained launches, 8 launches before the SGD kernel
    plan = net.launch_plan()
    assert plan["fwd"] == list(range(12)) and plan["dgrad"] == list(range(1, 12)) and plan["wgrad"] == list(range(1, 12)), plan
    assert plan["refused"] == 0 and launches == 8, (plan, launches)
    check_loss(acc, problem)
    p = probs.cpu().numpy()
    assert np.abs(p - problem["probs"]).max() <= 2e-3 * problem["probs"].max()
    check_grads(net, problem)
    # the one-launch SGD over all 3.88 M parameters: w' - w = -lr * g (Keras SGD, momentum 0), and with momentum
    grads = grads_from_flat(net, net.grad_tensor().cpu().numpy())
    lr = 0.05
    net.sgd(lr, 0.9, 1.0)
    torch.cuda.synchronize()
    for l in range(12):
        w, b = net.get_conv(l)
        w0, b0 = problem["params"]["convs"][l]
        assert rel_err(w - w0, -lr * problem["grads"][2 * l]) < 3e-2, l
        assert rel_err(b - b0, -lr * problem["grads"][2 * l + 1]) < 3e-2, l
        assert np.abs((w - w0) + lr * grads[2 * l]).max() <= 1e-6 + 1e-5 * np.abs(lr * grads[2 * l]).max(), l   # exact fp32 update of ITS gradient
    hw, _ = net.get_head(0)
    assert rel_err(hw - problem["params"]["head_w"], -lr * problem["grads"][24]) < 3e-2
    # a second step uses the velocity: v = 0.9 v - lr g (Keras 1.0 SGD.get_updates)
    w_before = [net.get_conv(l)[0] for l in (0, 6, 11)]
    acc.zero_()
    net.forward_loss(problem["planes"], labels, acc)
    net.backward_range(11, 0, BATCH)
    g2 = grads_from_flat(net, net.grad_tensor().cpu().numpy())
    net.sgd(lr, 0.9, 1.0)
    torch.cuda.synchronize()
    for k, l in enumerate((0, 6, 11)):
        want = 0.9 * (-lr * grads[2 * l]) - lr * g2[2 * l]
        assert rel_err(net.get_conv(l)[0] - w_before[k], want) < 1e-3, l
    net.close()


def test_train_step_entry_point_equals_its_parts(problem):
    """ag_net_train_step (what bench.py's single-GPU step calls) = forward_loss + backward_range + sgd, bit for bit"""
    import torch
    labels = torch.from_numpy(problem["labels"]).cuda()
    a = make_net(19, 48, FILTERS, WIDTHS, problem["params"], max_batch=BATCH)
    b = make_net(19, 48, FILTERS, WIDTHS, problem["params"], max_batch=BATCH)
    acc_a, acc_b = torch.zeros(2, device="cuda"), torch.zeros(2, device="cuda")
    for _ in range(2):
        a.train_step(problem["planes"], labels, 0.02, 0.0, acc_a)
        b.forward_loss(problem["planes"], labels, acc_b)
        b.backward_range(11, 0, BATCH)
        b.sgd(0.02, 0.0, 1.0)
    torch.cuda.synchronize()
    assert torch.equal(acc_a, acc_b)
    for l in (0, 1, 5, 11):
        assert np.array_equal(a.get_conv(l)[0], b.get_conv(l)[0]), l
    a.close(); b.close()


def test_data_parallel_split_plan_at_the_benchmarked_shape(problem):
    """the multi-GPU step's launch plan on one rank: 140 of the SMs for the weight gradients (a different split-K plan)
    and the backward pass issued bucket by bucket as SLTrainer does before each all-reduce"""
    import torch
    from alphago_b200.training import plan_buckets
    net = make_net(19, 48, FILTERS, WIDTHS, problem["params"], max_batch=BATCH)
    full_splits = None
    labels = torch.from_numpy(problem["labels"]).cuda()
    acc = torch.zeros(2, device="cuda")
    net.forward_loss(problem["planes"], labels, acc)
    full_splits = net.launch_plan()["splits"]
    net.set_sm_budget(140)
    acc.zero_()
    net.forward_loss(problem["planes"], labels, acc)
    for buckets in (2, 4):
        acc.zero_()
        net.forward_loss(problem["planes"], labels, acc)
        for lo, hi, start, end in plan_buckets(net.n_conv, net.grad_offsets, buckets):
            net.backward_range(hi - 1, lo, BATCH)
        torch.cuda.synchronize()
        plan = net.launch_plan()
        assert plan["wgrad"] == list(range(1, 12)) and plan["refused"] == 0, plan
        assert plan["splits"] != full_splits, (plan["splits"], full_splits)
        check_loss(acc, problem)
        check_grads(net, problem)
    net.close()


def test_value_net_forward_at_batch_256(problem):
    """CNNValue stack (49 planes, 12 convs x 192 + 1x1 + Dense(256) + Dense(1, tanh), value.py:14-28) at batch 256"""
    import torch
    params = no.make_params(49, FILTERS, WIDTHS, seed=5, bias_scale=0.02, value_head=True)
    params["head_w"] = (params["head_w"] * 20).astype(np.float32)
    colour = np.zeros((BATCH, 1, 19, 19), np.uint8)
    colour[::2] = 1
    x = np.concatenate([problem["x"], colour], axis=1)
    net = make_net(19, 49, FILTERS, WIDTHS, params, max_batch=BATCH, head=nn.VALUE_HEAD)
    v = net.forward(torch.from_numpy(x).cuda()).cpu().numpy()
    assert net.launch_plan()["fwd"] == list(range(12))
    want = no.forward(params, x, emulate_bf16=True)["value"].numpy()
    fp = no.forward(params, x)["value"].numpy()
    assert v.shape == (BATCH,) and np.abs(v - want).max() < 2e-3 and np.abs(v - fp).max() < 2e-2
    assert np.std(want) > 1e-3   # the outputs are not all one value
    net.close()
