"""GPU parity tests AT THE SHAPE bench.py TIMES (BASELINE configs[2]): the SL policy net with 12 KxK convs x 192 filters
+ the 1x1 head, 48 planes, batch 256, positions from the CUDA encoder -- i.e. the chained forward launch, the 11-layer
data-gradient chain, the 11-layer chained weight gradients with their cross-layer split reduction, the taps-in-N
first-layer weight gradient at K = 256 * 441 rows and the one-launch SGD, all against oracle/nn_oracle.py
(supervised_policy_trainer.py:199-213 on policy.py:119-154).

Two kinds of check, because of what the first GPU run of these tests showed (scripts/grad_probe*.py):

* KERNEL level, from the CUDA path's own inputs.  Every layer's activations and pre-activation gradients are read back and
  each kernel's output is recomputed on the CPU from exactly the tensors that kernel read.  What is left is fp32 summation
  order: weight and bias gradients < 1e-3 relative Frobenius error (measured 1e-5), data gradients and activations within
  2 bf16 ulps of the layer's largest value (measured 0.2-1.1 ulp).  This is the sharp test of the chained launches.
* END TO END against oracle/nn_oracle.py.  Loss within 2e-3 relative, correct-count within one tie.  The GRADIENT of a
  12-layer net on the shared U(-0.05, 0.05) init is ill-conditioned: its deep activations are nearly constant over the
  board while dlogits = p - y sums to zero, so every weight gradient is the small remainder of a large cancellation and
  moves by 5-20 % when nothing but the bf16 rounding of the activations changes -- the oracle that rounds to bf16 and the
  plain fp32 oracle are that far apart from EACH OTHER (measured: 0.11-0.19 per tensor at batch 256).  The bar is
  therefore: the CUDA gradient is at least as close to the bf16-emulating oracle as that oracle is to fp32 (+ 1e-2), with
  cosine > 0.99; on 3-layer nets, where the problem is well conditioned, tests/test_nn_gpu.py keeps 3e-2 / 0.999.
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
    _, _, grads_fp32, _ = no.loss_and_grads(params, x, labels, emulate_bf16=False)
    return {"planes": planes, "x": x, "labels": labels, "params": params, "loss": loss, "acc": acc, "grads": grads,
            "grads_fp32": grads_fp32, "probs": out["probs"].detach().numpy()}


def names():
    out = []
    for l in range(len(FILTERS)):
        out += ["conv%d.w" % l, "conv%d.b" % l]
    return out + ["head.w", "head.b"]


def check_grads(net, problem):
    got = grads_from_flat(net, net.grad_tensor().cpu().numpy())
    worst = (0.0, "")
    for name, g, want, fp in zip(names(), got, problem["grads"], problem["grads_fp32"]):
        assert g.shape == want.shape, name
        if name == "head.b":   # sum of (p - y) over all moves: zero up to rounding
            assert abs(float(g[0])) < 1e-5 and abs(float(want[0])) < 1e-5
            continue
        r, c, cond = rel_err(g, want), cosine(g, want), rel_err(want, fp)
        worst = max(worst, (r, name, cond))
        bar = 3e-2 if name == "head.w" else max(3e-2, cond + 1e-2)   # see the module docstring
        assert r < bar and c > 0.99, "%s: rel %g cos %g (the two oracles are %g apart)" % (name, r, c, cond)
    print("worst gradient tensor: %s rel %.4g (bf16-emulating vs fp32 oracle: %.4g)" % (worst[1], worst[0], worst[2]))


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
    assert np.abs(p - problem["probs"]).max() <= 1e-2 * problem["probs"].max()   # as tests/test_nn_gpu.py states for 12 layers
    check_grads(net, problem)
    # the one-launch SGD over all 3.88 M parameters: w' - w = -lr * g (Keras SGD, momentum 0), and with momentum
    grads = grads_from_flat(net, net.grad_tensor().cpu().numpy())
    lr = 0.05
    net.sgd(lr, 0.9, 1.0)
    torch.cuda.synchronize()
    for l in range(12):
        w, b = net.get_conv(l)
        w0, b0 = problem["params"]["convs"][l]
        # the fp32 update of the gradient just checked (check_grads compared that gradient with the oracle)
        assert np.abs((w - w0) + lr * grads[2 * l]).max() <= 1e-6 + 1e-5 * np.abs(lr * grads[2 * l]).max(), l
        assert np.abs((b - b0) + lr * grads[2 * l + 1]).max() <= 1e-6 + 1e-5 * np.abs(lr * grads[2 * l + 1]).max(), l
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
    # the loss sum is accumulated with float atomics (one per image, any order): equal up to the last bits, not bitwise
    assert torch.allclose(acc_a, acc_b, rtol=1e-5, atol=0) and acc_a[1] == acc_b[1]
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


def test_every_kernel_of_the_step_from_its_own_inputs(problem):
    """kernel-level parity at the benchmarked shape (module docstring): forward chain, data-gradient chain, chained weight
    gradients, the first layer's taps-in-N weight gradient, each recomputed from the tensors the CUDA kernel itself read"""
    import torch
    import torch.nn.functional as Fn
    net = make_net(19, 48, FILTERS, WIDTHS, problem["params"], max_batch=BATCH)
    labels = torch.from_numpy(problem["labels"]).cuda()
    acc = torch.zeros(2, device="cuda")
    net.forward_loss(problem["planes"], labels, acc)
    net.backward_range(len(FILTERS) - 1, 0, BATCH)
    torch.cuda.synchronize()
    plan = net.launch_plan()
    assert plan["fwd"] == list(range(12)) and plan["dgrad"] == list(range(1, 12)) and plan["wgrad"] == list(range(1, 12)), plan
    got = grads_from_flat(net, net.grad_tensor().cpu().numpy())
    bf = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(torch.bfloat16).to(torch.float32)
    acts = [torch.from_numpy(net.debug_read(l, 0, BATCH)) for l in range(12)]
    x0 = torch.from_numpy(problem["x"].astype(np.float32))
    ulp = 2.0 ** -8
    for l in range(12):
        w, b = problem["params"]["convs"][l]
        k = WIDTHS[l]
        xin = x0 if l == 0 else acts[l - 1]
        # forward: relu(conv(x, bf16(w)) + b) rounded to bf16
        want = torch.relu(Fn.conv2d(xin, bf(w), torch.from_numpy(b), padding=k // 2)).to(torch.bfloat16).to(torch.float32)
        err = float((acts[l] - want).abs().max())
        assert err <= 2 * ulp * max(1.0, float(want.abs().max())), "forward layer %d: %g" % (l, err)
        # weight / bias gradient from this layer's own G and input
        G = torch.from_numpy(net.debug_read(l, 1, BATCH))
        wv = torch.zeros(FILTERS[l], xin.shape[1], k, k, requires_grad=True)
        (Fn.conv2d(xin, wv, padding=k // 2) * G).sum().backward()
        assert rel_err(got[2 * l], wv.grad.numpy()) < 1e-3, "weight gradient of layer %d: %g" % (l, rel_err(got[2 * l], wv.grad.numpy()))
        assert rel_err(got[2 * l + 1], G.sum((0, 2, 3)).numpy()) < 1e-3, "bias gradient of layer %d" % l
        if l > 0:   # data gradient into the layer below, gated by that layer's ReLU, stored in bf16
            dx = Fn.conv_transpose2d(G, bf(w), padding=k // 2)
            want = (dx * (acts[l - 1] > 0)).to(torch.bfloat16).to(torch.float32).numpy()
            have = net.debug_read(l - 1, 1, BATCH)
            assert np.abs(have - want).max() <= 2 * ulp * np.abs(want).max(), "data gradient into layer %d" % (l - 1)
            assert rel_err(have, want) < 1e-3, "data gradient into layer %d" % (l - 1)
    net.close()


def test_top1_agreement_on_trained_weights(golden_dir):
    """BASELINE north_star: ">= 99 % top-1 move agreement ... on the reference's own trained weights or a shared random
    init".  The reference tree holds no trained weights (README.md:18 points to an external data repository), and a
    random-init 13-layer net is flat (largest probability ~ 1.5 / 361), so its arg-max is decided far below bf16
    resolution.  The test therefore MAKES decisive weights: 600 CUDA SGD steps (batch 256, the 8 symmetries, lr 0.003,
    momentum 0.9) on the 1033 real positions of the reference's own HDF5 fixture, from the shared U(-0.05, 0.05) init
    -- about 0.6 s on a B200 -- and then compares the bf16 CUDA forward with the fp32 oracle on those weights over all
    1033 positions.  Bars: raw top-1 agreement >= 99 %; max-abs probability error <= 0.06 (fp32 oracle) and <= 0.03
    (bf16-emulating oracle); max-abs log-probability ("logit" up to the per-position normaliser) error over moves with
    p >= 1e-3 <= 0.5."""
    import os
    import torch
    from alphago_b200.training import SLTrainer
    from test_oracle import load_hdf5_fixture
    _, actions, offsets, _ = load_hdf5_fixture(golden_dir)
    states = []
    for start, count in offsets:
        st = go.GameState(19)
        for i in range(start, start + count):
            states.append(st.copy())
            st.do_move((int(actions[i][0]), int(actions[i][1])))
    planes = Preprocess(DEFAULT_FEATURES).states_to_tensor(states)
    d_act = torch.from_numpy(np.ascontiguousarray(actions)).cuda()
    n = len(states)
    params = no.make_params(48, FILTERS, WIDTHS, seed=1)
    params["head_w"] = (params["head_w"] * 20).astype(np.float32)
    net = make_net(19, 48, FILTERS, WIDTHS, params, max_batch=BATCH)
    tr = SLTrainer(net, lr=0.003, decay=0.0, momentum=0.9)
    rng = np.random.RandomState(0)
    perm = rng.permutation(n)
    first = last = None
    for s in range(600):
        idx = perm[(np.arange(BATCH) + s * BATCH) % n]
        x, y = tr.assemble(planes, d_act, idx, rng.randint(0, 8, BATCH))
        tr.step(x, y)
        if s % 50 == 49:
            m = tr.pop_metrics(50 * BATCH)
            first = first or m
            last = m
    print("training: loss/acc %.3f/%.2f -> %.3f/%.2f" % (first + last))
    assert last[0] < 0.5 * first[0] and last[1] > 0.5, (first, last)      # the net has become decisive
    trained = {"convs": [net.get_conv(l) for l in range(12)], "head_w": net.get_head(0)[0], "head_b": net.get_head(0)[1],
               "pos_bias": None}
    got = np.concatenate([net.forward(planes[s:s + BATCH].contiguous()).cpu().numpy() for s in range(0, n, BATCH)])
    x = planes.cpu().numpy()
    fp = no.forward(trained, x)["probs"].numpy()
    em = no.forward(trained, x, emulate_bf16=True)["probs"].numpy()
    agree, agree_em = (got.argmax(1) == fp.argmax(1)).mean(), (got.argmax(1) == em.argmax(1)).mean()
    err, err_em = np.abs(got - fp).max(), np.abs(got - em).max()
    sel = fp >= 1e-3
    logit_err = np.abs(np.log(np.maximum(got[sel], 1e-30)) - np.log(fp[sel])).max()
    print("top-1 agreement over %d positions: %.4f vs fp32 oracle, %.4f vs bf16-emulating oracle; max-abs prob err %.4g / %.4g; "
          "max-abs log-prob err (p >= 1e-3) %.4g; median top probability %.3f"
          % (n, agree, agree_em, err, err_em, logit_err, np.median(fp.max(1))))
    assert agree >= 0.99 and agree_em >= 0.99
    assert err <= 0.06 and err_em <= 0.03 and logit_err <= 0.5
    net.close()
