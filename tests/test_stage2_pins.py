"""Stage-2 code of the reference that runs without Keras, pinned by fixtures the UNMODIFIED reference produced
(oracle/gen_golden.py `stage2`): CNNPolicy's move-selection glue (policy.py:44-91), the training generator with its
8 board transforms (supervised_policy_trainer.py:10-40,71-80) and MetadataWriterCallback (spt.py:43-68).

Host tests check this package's host logic and the oracle's restatement against the fixtures; the GPU tests run the
CUDA encoder / `ag_augment` / the device renormalisation against the same fixtures.
"""
import json
import os

import numpy as np
import pytest

from alphago_b200 import go
from alphago_b200.models import CNNPolicy
from alphago_b200.supervised_policy_trainer import MetadataWriter, _index_stream

F3 = ["board", "ones", "turns_since"]


@pytest.fixture(scope="module")
def glue(golden_dir):
    return np.load(os.path.join(golden_dir, "ref_policy_glue.npz"))


@pytest.fixture(scope="module")
def gen(golden_dir):
    return np.load(os.path.join(golden_dir, "ref_generator.npz"))


def replay(glue, i):
    size = int(glue["size"])
    st = go.GameState(size)
    for m in glue["moves_%d" % i]:
        st.do_move(None if m < 0 else (int(m) // size, int(m) % size))
    return st


class _FakeOutput(object):
    """what SequentialModel.forward_device returns, as far as CNNPolicy uses it: `.cpu().numpy()`"""

    def __init__(self, a):
        self.a = a

    def cpu(self):
        return self

    def numpy(self):
        return self.a


def glue_policy(glue, encoder=None):
    """a CNNPolicy whose network is the fixture's lookup table, like the `forward` the reference ran with"""
    policy = object.__new__(CNNPolicy)
    net_out = glue["net_out"]

    class Pre(object):
        def states_to_tensor(self, states):
            if encoder is not None:   # the real CUDA encoder runs; its planes are not what is pinned here
                planes = encoder.states_to_tensor(states)
                assert tuple(planes.shape) == (len(states), 12, 19, 19)
            return len(states)

    class Model(object):
        row = 0

        def forward_device(self, n):
            return _FakeOutput(net_out[:n] if n > 1 else net_out[self.row:self.row + 1])
    policy.preprocessor, policy.model = Pre(), Model()
    return policy


def pairs(res, size):
    res = list(res)
    return np.asarray([m[0] * size + m[1] for m, _ in res], np.int32), np.asarray([p for _, p in res], np.float64)


def check_glue(glue, policy):
    n, size = int(glue["n_states"]), int(glue["size"])
    states = [replay(glue, i) for i in range(n)]
    res = policy.batch_eval_state(states)
    for i in range(n):
        idx, p = pairs(res[i], size)
        assert np.array_equal(idx, glue["batch_default_idx_%d" % i]), i        # default = legal moves, row-major
        assert np.array_equal(p, glue["batch_default_p_%d" % i]), i            # same float32 arithmetic, bit for bit
    asked = [[(int(m) // size, int(m) % size) for m in glue["asked_%d" % i]] for i in range(n)]
    res = policy.batch_eval_state(states, asked)
    for i in range(n):
        idx, p = pairs(res[i], size)
        assert np.array_equal(idx, glue["batch_asked_idx_%d" % i]) and np.array_equal(p, glue["batch_asked_p_%d" % i]), i
    assert len(res[1]) == 0                                                    # an empty move list -> [] (policy.py:48-49)
    for i in range(n):
        policy.model.row = i
        idx, p = pairs(policy.eval_state(states[i]), size)
        assert np.array_equal(idx, glue["eval_default_idx_%d" % i]) and np.array_equal(p, glue["eval_default_p_%d" % i])
        # eval_state(moves=[]) falls back to the legal moves (`moves or ...`, policy.py:90)
        idx, p = pairs(policy.eval_state(states[i], asked[i]), size)
        assert np.array_equal(idx, glue["eval_asked_idx_%d" % i]) and np.array_equal(p, glue["eval_asked_p_%d" % i])
    assert int(glue["empty_batch_is_empty_list"]) == 1 and policy.batch_eval_state([]) == []
    with pytest.raises(ValueError) as e:
        policy.batch_eval_state([go.GameState(19), go.GameState(13)])
    assert str(e.value) == str(glue["mixed_size_error"])


def test_policy_glue_matches_reference(glue):
    """_select_moves_and_normalize / batch_eval_state / eval_state on the host rules engine"""
    check_glue(glue, glue_policy(glue))


def test_metadata_writer_matches_reference_callback(gen, tmp_path):
    record = json.loads(str(gen["metadata_json"]))
    for name, rec in record.items():
        w = MetadataWriter(os.path.join(str(tmp_path), name + ".json"))
        for logs, after in zip(rec["logs"], rec["after"]):
            w.on_epoch_end(dict(logs))
            assert json.load(open(w.file)) == after, name


def test_index_stream_is_the_generators_order(gen):
    """the reference generator walks `indices` cyclically and cuts batches wherever they fall, so a batch may straddle the
    end of the list (spt.py:27-40)"""
    batch = int(gen["gen_batch"])
    used = gen["gen_used_index"]
    stream = _index_stream(gen["gen_indices"], batch)
    for k in range(len(used) // batch):
        assert np.array_equal(next(stream), used[k * batch:(k + 1) * batch]), k
    assert len(gen["gen_indices"]) % batch != 0


def test_oracle_transforms_match_reference(gen):
    """oracle/nn_oracle.py's restatement of BOARD_TRANSFORMATIONS / one_hot_action, transform by transform"""
    from oracle import nn_oracle as no
    for size in (19, 7):
        plane = gen["plane_%d" % size]
        for t in range(8):
            assert np.array_equal(no.BOARD_TRANSFORMATIONS[t](plane), gen["plane_tf_%d" % size][t]), (size, t)
        acts = gen["acts_%d" % size]
        states = np.zeros((len(acts), 1, size, size), np.uint8)
        for t in range(8):
            _, y = no.augment_batch(states, acts, np.arange(len(acts)), np.full(len(acts), t))
            assert np.array_equal(y, gen["acts_tf_%d" % size][t]), (size, t)
    X, Y = no.augment_batch(gen["gen_states"], gen["gen_actions"], gen["gen_used_index"], gen["gen_used_transform"])
    assert np.array_equal(X.reshape(gen["gen_X"].shape), gen["gen_X"])
    assert np.array_equal(Y, np.argmax(gen["gen_Y"].reshape(len(Y), -1), axis=1))
    assert (gen["gen_Y"].reshape(len(Y), -1).sum(1) == 1).all()


# ---------------------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_policy_glue_on_the_cuda_encoder(glue):
    """the same pins with the CUDA encoder in the loop (the packed states go through ag_encode)"""
    from alphago_b200.preprocessing import Preprocess
    check_glue(glue, glue_policy(glue, encoder=Preprocess(F3)))


def _assemble(states, actions, index, transform):
    import ctypes as C
    import torch
    from alphago_b200 import _lib
    n, planes, size = len(index), states.shape[1], states.shape[-1]
    ds, da = torch.from_numpy(states).cuda(), torch.from_numpy(np.ascontiguousarray(actions)).cuda()
    di = torch.from_numpy(np.asarray(index, np.int64)).cuda()
    dt = torch.from_numpy(np.asarray(transform, np.uint8)).cuda()
    out = torch.empty((n, planes, size, size), dtype=torch.uint8, device="cuda")
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    _lib.check(_lib.lib().ag_augment(C.c_void_p(ds.data_ptr()), C.c_void_p(da.data_ptr()), C.c_void_p(di.data_ptr()),
                                     C.c_void_p(dt.data_ptr()), n, planes, size, C.c_void_p(out.data_ptr()),
                                     C.c_void_p(lab.data_ptr()), _lib.current_stream()), "ag_augment")
    torch.cuda.synchronize()
    return out.cpu().numpy(), lab.cpu().numpy()


@pytest.mark.gpu
def test_ag_augment_equals_the_reference_generator(gen):
    """ag_augment on the generator's own (index, transform) choices must give the (X, Y) batches the reference yielded"""
    X, Y = _assemble(gen["gen_states"], gen["gen_actions"], gen["gen_used_index"], gen["gen_used_transform"])
    assert np.array_equal(X.reshape(gen["gen_X"].shape), gen["gen_X"])
    want = gen["gen_Y"].reshape(len(Y), -1)
    assert np.array_equal(Y, np.argmax(want, axis=1)) and (want.sum(1) == 1).all()
    for size in (19, 7):   # every transform alone, on an asymmetric plane of byte values and on corner / centre actions
        plane, acts = gen["plane_%d" % size], gen["acts_%d" % size]
        states = np.ascontiguousarray(np.broadcast_to(plane, (len(acts), 1, size, size)))
        for t in range(8):
            x, y = _assemble(states, acts, np.arange(len(acts)), np.full(len(acts), t))
            assert np.array_equal(x[0, 0], gen["plane_tf_%d" % size][t]), (size, t)
            assert np.array_equal(y, gen["acts_tf_%d" % size][t]), (size, t)


@pytest.mark.gpu
def test_leaf_queue_device_renormalisation_matches_reference_glue(glue):
    """LeafEvaluator.flush_arrays renormalises over the legal moves on the device; the reference does it per state in
    numpy (policy.py:44-54).  Same numbers on the fixture's states when the net output is the fixture's table."""
    import torch
    from alphago_b200.leaf_queue import LeafEvaluator
    from alphago_b200.preprocessing import Preprocess
    n, size = int(glue["n_states"]), int(glue["size"])
    policy = object.__new__(CNNPolicy)
    policy.preprocessor = Preprocess(F3 + ["capture_size"])
    net_out = torch.from_numpy(glue["net_out"]).cuda()

    class Model(object):
        def forward_device(self, planes):
            return net_out[:planes.shape[0]]
    policy.model = Model()
    q = LeafEvaluator(policy, None)
    for i in range(n):
        q.submit(replay(glue, i), want_moves=False)
    P, _ = q.flush_arrays()
    for i in range(n):
        want = np.zeros(size * size)
        want[glue["batch_default_idx_%d" % i]] = glue["batch_default_p_%d" % i]
        assert np.array_equal(P[i] > 0, want > 0), i
        assert np.abs(P[i] - want).max() <= 2e-7 * max(1.0, want.max()), i   # float32 sum order differs on the device
