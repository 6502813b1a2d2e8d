"""GPU tests of the CNNPolicy / CNNValue object surface: the reference's tests/test_policy.py
(shapes, eval_state / batch_eval_state contracts, save/load weight equality, players) restated
against alphago_b200.models, plus numeric checks against the CPU oracle."""
import json
import os

import numpy as np
import pytest

from alphago_b200 import go
from alphago_b200.models import CNNPolicy, CNNValue, SequentialModel
from oracle import go_oracle as og
from oracle import nn_oracle as no

pytestmark = pytest.mark.gpu
F3 = ["board", "ones", "turns_since"]


def test_default_policy_eval_state_and_batch():  # tests/test_policy.py:12-21
    policy = CNNPolicy(F3, layers=3, filters_per_layer=32)
    gs = go.GameState()
    for m in [(3, 3), (15, 15), (3, 15)]:
        gs.do_move(m)
    out = policy.eval_state(gs)
    assert len(out) == len(gs.get_legal_moves()) == 358
    assert abs(sum(p for _, p in out) - 1.0) < 1e-5 and all(isinstance(m, tuple) for m, _ in out)
    results = policy.batch_eval_state([go.GameState(), go.GameState()])
    assert len(results) == 2 and len(results[0]) == 361 and len(results[1]) == 361
    assert policy.batch_eval_state([]) == []
    assert policy.eval_state(gs, moves=[(0, 0), (1, 1)])[0][0] == (0, 0)
    with pytest.raises(ValueError):
        policy.batch_eval_state([go.GameState(19), go.GameState(13)])


def test_output_size():  # tests/test_policy.py:23-30
    p19 = CNNPolicy(F3, board=19, layers=2, filters_per_layer=16)
    assert p19.forward(p19.preprocessor.state_to_tensor(go.GameState(19))).shape == (1, 361)
    p13 = CNNPolicy(F3, board=13, layers=2, filters_per_layer=16)
    assert p13.forward(p13.preprocessor.state_to_tensor(go.GameState(13))).shape == (1, 169)
    assert p13.model.input_shape == (None, 12, 13, 13)  # reinforcement_policy_trainer.py:51 reads input_shape[-1]


def test_save_load_and_keras_weight_convention(tmp_path):  # tests/test_policy.py:32-57
    policy = CNNPolicy(F3 + ["liberties"], layers=3, filters_per_layer=16)
    model_file = os.path.join(str(tmp_path), "model.json")
    weights_file = os.path.join(str(tmp_path), "weights.hdf5")
    policy.save_model(model_file, weights_file)
    spec = json.load(open(model_file))
    assert set(spec) == {"keras_model", "feature_list", "weights_file"}
    copy = CNNPolicy.load_model(model_file)
    w0, w1 = policy.model.get_weights(), copy.model.get_weights()
    assert len(w0) == 8 and w0[0].shape == (16, 20, 5, 5) and w0[-2].shape == (1, 16, 1, 1)
    assert all(np.array_equal(a, b) for a, b in zip(w0, w1))
    gs = go.GameState()
    a, b = policy.eval_state(gs), copy.eval_state(gs)
    assert all(abs(x[1] - y[1]) < 1e-7 for x, y in zip(a, b))
    # get_weights returns Keras/Theano (true-convolution) filters: the net equals the oracle on the flipped arrays
    params = {"convs": [(no.keras_to_xcorr(w0[2 * i]), w0[2 * i + 1]) for i in range(3)], "head_w": w0[6].reshape(-1),
              "head_b": w0[7], "pos_bias": None}
    st = go.GameState()
    for m in [(2, 3), (16, 4), (10, 10), (3, 16)]:
        st.do_move(m)
    b, a, pl, ko = st.pack()
    planes = og.encode_batch(b[None], a[None], [pl], [ko], 19, F3 + ["liberties"])
    want = no.forward(params, planes, emulate_bf16=True)["probs"].numpy()[0]
    got = dict(policy.eval_state(st))
    legal = st.get_legal_moves()
    wsum = sum(want[x * 19 + y] for x, y in legal)
    assert max(abs(got[m] - want[m[0] * 19 + m[1]] / wsum) for m in legal) < 2e-3 * want.max()
    # set_weights round trip
    policy.model.set_weights([w * 0.5 for w in w0])
    assert np.allclose(policy.model.get_weights()[0], w0[0] * 0.5)


def test_players_style_usage_and_batch_growth():  # tests/test_policy.py:60-102 (players only need eval_state)
    policy = CNNPolicy(F3, layers=2, filters_per_layer=16, max_batch=4)
    gs = go.GameState(size=19)
    for _ in range(6):
        moves = gs.get_legal_moves(include_eyes=False)
        dist = policy.eval_state(gs, moves)
        gs.do_move(max(dist, key=lambda mp: mp[1])[0])
    states = [go.GameState() for _ in range(9)]  # > max_batch: buffers grow, weights survive
    before = policy.model.get_weights()
    res = policy.batch_eval_state(states)
    assert len(res) == 9 and all(np.array_equal(a, b) for a, b in zip(before, policy.model.get_weights()))


def test_value_net_surface():
    feats = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
             "sensibleness", "zeros", "color"]
    v = CNNValue(feats + ["zeros", "zeros"], layers=3, filters_per_layer=32)
    assert v.preprocessor.output_dim == 49 and v.model.output_shape == (None, 1)  # value.py:16 input (49, 19, 19)
    gs = go.GameState()
    gs.do_move((3, 3))
    val = v.eval_state(gs)
    assert isinstance(val, float) and -1.0 <= val <= 1.0
    vals = v.batch_eval_state([gs, go.GameState()])
    assert len(vals) == 2 and abs(vals[0] - val) < 1e-6
    assert len(v.model.get_weights()) == 2 * 4 + 4


def test_config0_real_positions_batch16_vs_oracle(golden_dir):
    """BASELINE configs[0]: default 48-plane Preprocess + CNNPolicy with the reference defaults (12 layers, 128 filters)
    on positions of the AlphaGo - Lee Sedol games (tests/test_data SGFs, via the reference's own HDF5 fixture), batch 16:
    batch_eval_state against the CPU oracle (oracle encoder + fp32 / bf16-emulating network restatement)."""
    from test_oracle import load_hdf5_fixture
    _, actions, offsets, _ = load_hdf5_fixture(golden_dir)
    policy = CNNPolicy(["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
                        "ladder_capture", "ladder_escape", "sensibleness", "zeros"], seed=11)
    assert policy.preprocessor.output_dim == 48 and policy.model.filters == [128] * 12
    w = policy.model.get_weights()
    w[-2] = w[-2] * 40  # sharpen the random-init head so the distribution is not flat
    policy.model.set_weights(w)
    st = go.GameState(19)
    states = []
    start, count = offsets[0]
    for i in range(start, start + 160):
        st.do_move((int(actions[i][0]), int(actions[i][1])))
        if i % 10 == 9:
            states.append(st.copy())
    assert len(states) == 16
    got = policy.batch_eval_state(states)
    packed = [s.pack() for s in states]
    planes = og.encode_batch(np.stack([p[0] for p in packed]), np.stack([p[1] for p in packed]), [p[2] for p in packed],
                             [p[3] for p in packed], 19, og.DEFAULT_FEATURES)
    assert np.array_equal(policy.preprocessor.states_to_tensor(states).cpu().numpy(), planes)
    params = {"convs": [(no.keras_to_xcorr(w[2 * i]), w[2 * i + 1]) for i in range(12)], "head_w": w[24].reshape(-1),
              "head_b": w[25], "pos_bias": None}
    em = no.forward(params, planes, emulate_bf16=True)["probs"].numpy()
    fp = no.forward(params, planes)["probs"].numpy()
    for k, s in enumerate(states):
        legal = s.get_legal_moves()
        idx = [x * 19 + y for x, y in legal]
        for ref, tol in ((em, 1e-2), (fp, 5e-2)):
            want = ref[k][idx] / ref[k][idx].sum()
            have = np.asarray([p for _, p in got[k]])
            assert [m for m, _ in got[k]] == legal
            assert np.abs(have - want).max() <= tol * want.max(), (k, tol)
