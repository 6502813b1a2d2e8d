"""GPU test of the leaf-evaluation queue (alphago_b200/leaf_queue.py): batched, shared-encoding results
must be the ones CNNPolicy.eval_state / CNNValue.eval_state give state by state (policy.py:81-91, mcts.py:91-126)."""
import numpy as np
import pytest

from alphago_b200 import go
from alphago_b200.leaf_queue import LeafEvaluator
from alphago_b200.models import CNNPolicy, CNNValue
from alphago_b200.preprocessing import DEFAULT_FEATURES

pytestmark = pytest.mark.gpu


def _states(n, seed):
    rng = np.random.RandomState(seed)
    out = []
    for g in range(n):
        st = go.GameState(19)
        for _ in range(int(rng.randint(5, 120))):
            moves = st.get_legal_moves(include_eyes=False)
            if not moves:
                break
            st.do_move(moves[rng.randint(len(moves))])
        out.append(st)
    return out


def test_leaf_queue_matches_per_state_eval():
    policy = CNNPolicy(DEFAULT_FEATURES, layers=4, filters_per_layer=64, seed=5, max_batch=16)
    value = CNNValue(DEFAULT_FEATURES + ["color"], layers=4, filters_per_layer=64, seed=6, max_batch=16)
    q = LeafEvaluator(policy, value, max_batch=16)
    assert q.shared
    states = _states(37, 1)          # three flush chunks, the last one ragged
    for st in states:
        q.submit(st)
    st_mut = states[3]
    st_mut.do_move(st_mut.get_legal_moves()[0])   # a state may change after submit: it was packed at submit time
    got = q.flush()
    assert len(got) == 37 and len(q) == 0 and q.flush() == []
    ref_states = _states(37, 1)
    for i, st in enumerate(ref_states):
        want_p = policy.eval_state(st)
        want_v = value.eval_state(st)
        priors, v = got[i]
        assert [m for m, _ in priors] == [m for m, _ in want_p]
        assert np.allclose([p for _, p in priors], [p for _, p in want_p], rtol=0, atol=1e-7)
        assert abs(v - want_v) < 1e-6
        assert abs(sum(p for _, p in priors) - 1.0) < 1e-5


def test_flush_arrays_matches_flush():
    policy = CNNPolicy(DEFAULT_FEATURES, layers=3, filters_per_layer=32, seed=8, max_batch=8)
    value = CNNValue(DEFAULT_FEATURES + ["color"], layers=3, filters_per_layer=32, seed=9, max_batch=8)
    q = LeafEvaluator(policy, value, max_batch=8)
    states = _states(19, 3)
    lists = q.evaluate(states)
    for st in states:
        q.submit(st, want_moves=False)
    P, V = q.flush_arrays()
    assert P.shape == (19, 361) and V.shape == (19,)
    for i, (priors, v) in enumerate(lists):
        dense = np.zeros(361, np.float32)
        for (x, y), p in priors:
            dense[x * 19 + y] = p
        assert np.allclose(P[i], dense, rtol=1e-5, atol=1e-7) and abs(V[i] - v) < 1e-6


def test_leaf_queue_policy_only_and_errors():
    policy = CNNPolicy(["board", "ones", "turns_since"], layers=3, filters_per_layer=32, seed=7)
    q = LeafEvaluator(policy, None)
    sts = _states(5, 2)
    res = q.evaluate(sts, moves_lists=[[(0, 0), (3, 3)]] * 5)
    assert all(v is None and [m for m, _ in pr] == [(0, 0), (3, 3)] for pr, v in res)
    q.submit(go.GameState(19))
    q.submit(go.GameState(9))
    with pytest.raises(ValueError):
        q.flush()
