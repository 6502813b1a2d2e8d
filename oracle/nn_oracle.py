"""CPU ORACLE for the conv stack (test infrastructure only; never imported by alphago_b200/).

Restates the reference's network arithmetic -- AlphaGo/models/policy.py:119-154 (KxK 'same' ReLU
convs, 1x1 conv to one plane, Flatten in x*size+y order, softmax), AlphaGo/models/value.py:14-28,
loss/optimiser of AlphaGo/training/supervised_policy_trainer.py:199-200 -- on torch CPU tensors
(`conv2d` used strictly as the checker).  PARITY UNPINNED against the reference itself: Keras 1.0.0 /
Theano 0.8.1 are third-party, not vendored and not installable here, and the tree holds no golden
logits or trained weights.  What is pinned: architecture, index order, loss, update rule.

Convention: weights are CROSS-CORRELATION filters w[co][ci][kx][ky] (what torch conv2d and the
CUDA kernels use).  Theano's conv2d flips filters, so a Keras/Theano-trained array equals
np.flip(w, (2, 3)) of this one; `keras_to_xcorr` converts.

`emulate_bf16=True` rounds weights and every layer's activations to bf16 exactly where the CUDA
path does, so what is left between the two is fp32 accumulation order.
"""
import numpy as np
import torch
import torch.nn.functional as F


def keras_to_xcorr(w):
    return np.ascontiguousarray(np.flip(np.asarray(w), (2, 3)))


def _r(t, emulate):
    """round to bf16 (straight-through for autograd)"""
    if not emulate:
        return t
    return t + (t.detach().to(torch.bfloat16).to(t.dtype) - t.detach())


class _RoundGrad(torch.autograd.Function):
    """identity forward; rounds the incoming gradient to bf16 in backward.  The CUDA path stores every
    layer's pre-activation gradient in bf16 before its weight, bias and data gradients are formed."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(g.dtype)


def make_params(in_planes, filters, widths, seed=0, scale=0.05, bias_scale=0.0, board=19, value_head=False):
    """init='uniform' of Keras 1.0: U(-0.05, 0.05), zero biases (policy.py:127).  bias_scale>0
    gives non-zero biases so tests exercise the bias path."""
    rng = np.random.RandomState(seed)
    convs, cin = [], in_planes
    for f, k in zip(filters, widths):
        convs.append((rng.uniform(-scale, scale, (f, cin, k, k)).astype(np.float32),
                      rng.uniform(-bias_scale, bias_scale, f).astype(np.float32)))
        cin = f
    p = {"convs": convs, "head_w": rng.uniform(-scale, scale, cin).astype(np.float32),
         "head_b": rng.uniform(-bias_scale, bias_scale, 1).astype(np.float32), "pos_bias": None}
    if value_head:
        pts = board * board
        p["d1w"] = rng.uniform(-scale, scale, (256, pts)).astype(np.float32)
        p["d1b"] = rng.uniform(-bias_scale, bias_scale, 256).astype(np.float32)
        p["d2w"] = rng.uniform(-scale, scale, 256).astype(np.float32)
        p["d2b"] = rng.uniform(-bias_scale, bias_scale, 1).astype(np.float32)
    return p


def forward(params, planes, emulate_bf16=False, dtype=torch.float32, want_grad=False):
    """planes: uint8/any array (N, C, S, S).  returns dict(acts=[...], logits, probs, tensors)"""
    x = torch.as_tensor(np.asarray(planes)).to(dtype)
    tensors, acts = [], []
    for (w, b) in params["convs"]:
        wt = torch.tensor(w, dtype=dtype, requires_grad=want_grad)
        bt = torch.tensor(b, dtype=dtype, requires_grad=want_grad)
        tensors += [wt, bt]
        pre = F.conv2d(x, _r(wt, emulate_bf16), bt, padding=w.shape[2] // 2)
        if emulate_bf16 and want_grad:
            pre = _RoundGrad.apply(pre)
        x = _r(F.relu(pre), emulate_bf16)
        acts.append(x)
    hw = torch.tensor(params["head_w"], dtype=dtype, requires_grad=want_grad)
    hb = torch.tensor(params["head_b"], dtype=dtype, requires_grad=want_grad)
    tensors += [hw, hb]
    plane = (x * hw[None, :, None, None]).sum(1) + hb  # 1x1 conv to one plane (policy.py:144-150)
    logits = plane.flatten(1)                            # Flatten: index x*size+y (util.py:6-8)
    out = {"acts": acts, "tensors": tensors}
    if "d1w" in params:                                  # value.py:26-28
        h = logits @ torch.tensor(params["d1w"], dtype=dtype).T + torch.tensor(params["d1b"], dtype=dtype)
        v = torch.tanh(h @ torch.tensor(params["d2w"], dtype=dtype) + torch.tensor(params["d2b"], dtype=dtype)[0])
        out["value"] = v
        return out
    if params.get("pos_bias") is not None:
        pb = torch.tensor(params["pos_bias"], dtype=dtype, requires_grad=want_grad)
        tensors.append(pb)
        logits = logits + pb[None, :]
    out["logits"] = logits
    out["probs"] = torch.softmax(logits, 1)
    return out


def loss_and_grads(params, planes, labels, emulate_bf16=False, dtype=torch.float32, loss_kind="categorical"):
    """categorical cross-entropy (mean over the batch, Keras clips probabilities to [1e-7, 1-1e-7])
    + accuracy + gradients of every parameter (spt.py:199-200).  loss_kind="binary": Keras 1.0's binary_crossentropy on
    the softmax outputs against the one-hot move (reinforcement_policy_trainer.py:109): per sample the mean over the
    outputs of -(y log p + (1-y) log(1-p)) with p clipped to [1e-7, 1-1e-7], mean over the batch."""
    out = forward(params, planes, emulate_bf16, dtype, want_grad=True)
    lab = torch.as_tensor(np.asarray(labels), dtype=torch.long)
    if loss_kind == "binary":
        pr = torch.softmax(out["logits"], dim=1)
        pc = pr.clamp(1e-7, 1 - 1e-7)
        y = F.one_hot(lab, pr.shape[1]).to(pr.dtype)
        loss = (-(y * torch.log(pc) + (1 - y) * torch.log(1 - pc))).mean(1).mean()
        loss.backward()
        acc = (out["probs"].argmax(1) == lab).double().mean()
        grads = [t.grad.numpy() for t in out["tensors"]]
        return float(loss.detach()), float(acc), grads, out
    p = out["probs"].gather(1, lab[:, None])[:, 0]
    loss = -torch.log(p.clamp(1e-7, 1 - 1e-7)).mean()
    # gradient of the unclipped softmax-CE, which is what Theano's fused op yields
    F.cross_entropy(out["logits"], lab).backward()
    acc = (out["probs"].argmax(1) == lab).double().mean()
    grads = [t.grad.numpy() for t in out["tensors"]]
    return float(loss.detach()), float(acc), grads, out


def sgd_update(params, grads, lr, momentum=0.0, velocity=None):
    """Keras 1.0 SGD.get_updates: v = momentum*v - lr*g; p = p + v"""
    flat = [a for wb in params["convs"] for a in wb] + [params["head_w"], params["head_b"]]
    if params.get("pos_bias") is not None:
        flat.append(params["pos_bias"])
    velocity = velocity or [np.zeros_like(p) for p in flat]
    new = []
    for p, g, v in zip(flat, grads, velocity):
        v[...] = momentum * v - lr * g
        new.append((p + v).astype(np.float32))
    n = len(params["convs"])
    out = dict(params)
    out["convs"] = [(new[2 * i], new[2 * i + 1]) for i in range(n)]
    out["head_w"], out["head_b"] = new[2 * n], new[2 * n + 1]
    if params.get("pos_bias") is not None:
        out["pos_bias"] = new[2 * n + 2]
    return out, velocity


# supervised_policy_trainer.py:71-80 BOARD_TRANSFORMATIONS, restated (same order, arrays indexed [x][y])
BOARD_TRANSFORMATIONS = [
    lambda f: f,
    lambda f: np.rot90(f, 1),
    lambda f: np.rot90(f, 2),
    lambda f: np.rot90(f, 3),
    lambda f: np.fliplr(f),
    lambda f: np.flipud(f),
    lambda f: np.transpose(f),
    lambda f: np.fliplr(np.rot90(f, 1)),
]


def augment_batch(states, actions, index, transform):
    """shuffled_hdf5_batch_generator's per-sample work (spt.py:27-36): transform every plane and the
    one-hot action; returns uint8 (B, F, S, S) and the flat label of the transformed one-hot."""
    size = states.shape[-1]
    X = np.zeros((len(index),) + states.shape[1:], np.uint8)
    Y = np.zeros(len(index), np.int32)
    for b, (i, t) in enumerate(zip(index, transform)):
        tf = BOARD_TRANSFORMATIONS[t]
        X[b] = np.array([tf(plane) for plane in states[i]])
        onehot = np.zeros((size, size))
        onehot[tuple(actions[i])] = 1  # one_hot_action, spt.py:10-15
        Y[b] = int(np.argmax(tf(onehot).flatten()))
    return X, Y
