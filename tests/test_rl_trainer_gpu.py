"""GPU tests of the policy players and the REINFORCE trainer (alphago_b200/ai.py, reinforcement_policy_trainer.py), the
drop-ins for AlphaGo/ai.py and AlphaGo/training/reinforcement_policy_trainer.py.  The reference's own versions need Keras
and Python 2 and have no tests; what is checked is the contract read off their source."""
import argparse

import numpy as np
import pytest

from alphago_b200 import go
from alphago_b200.ai import GreedyPolicyPlayer, ProbabilisticPolicyPlayer
from alphago_b200.models import CNNPolicy
from alphago_b200.reinforcement_policy_trainer import make_training_pairs, run, train_batch

pytestmark = pytest.mark.gpu
FEATS = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after", "sensibleness", "zeros"]


def _policy(seed, board=7):
    return CNNPolicy(FEATS, board=board, layers=3, filters_per_layer=32, seed=seed, max_batch=16)


def test_players_choose_sensible_moves():
    pol = _policy(1)
    st = go.GameState(7)
    for mv in [(3, 3), (3, 4), (2, 2)]:
        st.do_move(mv)
    sensible = st.get_legal_moves(include_eyes=False)
    g = GreedyPolicyPlayer(pol).get_move(st)
    probs = dict(pol.eval_state(st, sensible))
    assert g in sensible and probs[g] == max(probs.values())
    np.random.seed(0)
    p = ProbabilisticPolicyPlayer(pol, temperature=0.5)
    assert p.get_move(st) in sensible
    np.random.seed(3)
    batch = p.get_moves([st, go.GameState(7), st])
    assert len(batch) == 3 and all(m in s.get_legal_moves(include_eyes=False) for m, s in zip(batch, [st, go.GameState(7), st]))
    # nothing sensible left -> pass (ai.py:22,49)
    full = go.GameState(7)
    full.get_legal_moves = lambda include_eyes=True: []
    assert GreedyPolicyPlayer(pol).get_move(full) is go.PASS_MOVE and p.get_move(full) is go.PASS_MOVE


def test_make_training_pairs_and_train_batch():
    learner, opponent = ProbabilisticPolicyPlayer(_policy(2)), ProbabilisticPolicyPlayer(_policy(3))
    np.random.seed(11)
    X_list, y_list, winners = make_training_pairs(learner, opponent, FEATS, 4, max_plies=40)
    colour = make_training_pairs.player_color
    assert len(X_list) == len(y_list) == len(winners) == 4 and colour in (go.BLACK, go.WHITE)
    for X, y in zip(X_list, y_list):
        assert X.dtype == np.uint8 and X.shape[1:] == (learner.policy.preprocessor.output_dim, 7, 7)
        assert y.shape == (len(X), 49) and (y.sum(1) == 1).all() and 15 <= len(X) <= 20   # the learner's plies of 40
        # the move was legal where it was recorded: its point is empty in the 'board' planes (plane 2)
        if colour == go.BLACK:   # (for White the reference pairs the move with the position one ply earlier)
            assert all(X[i, 2].reshape(-1)[y[i].argmax()] == 1 for i in range(len(X)))
    # same seed, same games
    np.random.seed(11)
    X2, y2, w2 = make_training_pairs(learner, opponent, FEATS, 4, max_plies=40)
    assert all(np.array_equal(a, b) for a, b in zip(X_list, X2)) and w2 == winners
    # without the reference's stale-position quirk every recorded move is on an empty point
    np.random.seed(12)
    X3, y3, w3 = make_training_pairs(learner, opponent, FEATS, 3, reference_quirks=False, max_plies=30)
    assert all(X[i, 2].reshape(-1)[y[i].argmax()] == 1 for X, y in zip(X3, y3) for i in range(len(X)))
    # one update per game, sign by result
    from alphago_b200.models import SGD
    learner.policy.model.compile(loss="binary_crossentropy", optimizer=SGD(lr=0.05))
    before = [w.copy() for w in learner.policy.model.get_weights()]
    train_batch(learner, X_list, y_list, winners, 0.05)
    after = learner.policy.model.get_weights()
    assert any(not np.array_equal(a, b) for a, b in zip(before, after)) and all(np.isfinite(a).all() for a in after)
    assert abs(learner.policy.model.optimizer.lr.get_value()) == 0.05


def test_fast_move_selection_is_sensible():
    import torch
    pl = ProbabilisticPolicyPlayer(_policy(6), temperature=0.7)
    torch.manual_seed(0)
    np.random.seed(2)
    X, y, w = make_training_pairs(pl, pl, FEATS, 5, reference_quirks=False, max_plies=60, fast=True)
    assert all(len(a) > 0 for a in X)
    assert all(X_[i, 2].reshape(-1)[y_[i].argmax()] == 1 for X_, y_ in zip(X, y) for i in range(len(X_)))   # played on empty points
    st = go.GameState(7)
    for mv in [(3, 3), (3, 4), (2, 2), (4, 4)]:
        st.do_move(mv)
    for m in pl.get_moves_fast([st] * 8):
        assert m in st.get_legal_moves(include_eyes=False)


def test_run_loop_plays_to_the_end_on_a_small_board():
    learner = ProbabilisticPolicyPlayer(_policy(4, board=5))
    args = argparse.Namespace(learning_rate=0.02, iterations=1, save_every=2, game_batch_size=3)
    np.random.seed(5)
    out = run(learner, args, [learner], FEATS)
    assert out == [learner] and len(run.player_wins_per_batch) == 2 and all(0 <= w <= 3 for w in run.player_wins_per_batch)


class FakePolicy(object):
    """the stand-in network of oracle/gen_golden.py `selfplay` (same weights for the same position and move)"""

    def __init__(self, salt, size=19, planes=46):
        self.salt = salt
        self.model = type("M", (), {"input_shape": (None, planes, size, size)})()

    def batch_eval_state(self, states, moves_lists=None):
        return [[(m, 1.0 + ((m[0] * 7 + m[1] * 13 + len(st.history) * 3 + self.salt) % 11)) for m in ml]
                for st, ml in zip(states, moves_lists)]


@pytest.mark.parametrize("tag", ["black", "white"])
def test_make_training_pairs_equals_the_reference_loop(golden_dir, tag):
    """tests/golden/ref_selfplay.npz: the reference's make_training_pairs + ProbabilisticPolicyPlayer.get_moves run from their
    source on the reference GameState / Preprocess with FakePolicy for the networks, two 19x19 games to the end per colour
    of the learner.  The product's loop (host rules engine, CUDA encoder) must yield the same pairs bit for bit: same moves
    drawn from np.random, the same positions -- one ply old for a White learner --, turns_since planes aged by the history
    the copied state shares with the live one (go.py:167), the same winners."""
    import os
    g = np.load(os.path.join(golden_dir, "ref_selfplay.npz"))
    feats = [str(f) for f in g["features"]]
    learner, opponent = ProbabilisticPolicyPlayer(FakePolicy(3), temperature=0.67), ProbabilisticPolicyPlayer(FakePolicy(8))
    np.random.seed(int(g["seed_" + tag]))
    X_list, y_list, winners = make_training_pairs(learner, opponent, feats, 2)
    assert make_training_pairs.player_color == (go.BLACK if tag == "black" else go.WHITE)
    assert [int(w) for w in winners] == [int(w) for w in g["winners_" + tag]]
    for i in range(2):
        want_y = g["y_%s_%d" % (tag, i)]
        want_X = np.unpackbits(g["X_%s_%d" % (tag, i)], axis=1)[:, :46 * 361].reshape(-1, 46, 19, 19)
        assert np.array_equal(y_list[i].argmax(1), want_y), (tag, i)
        assert X_list[i].shape == want_X.shape
        if not np.array_equal(X_list[i], want_X):
            bad = np.argwhere(X_list[i] != want_X)
            raise AssertionError("%s game %d: %d bytes differ, planes %s, first %s" % (tag, i, len(bad), sorted(set(bad[:, 1].tolist())), bad[:3].tolist()))
