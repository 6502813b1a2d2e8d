"""CPU test of the self-play loop's host logic (alphago_b200/reinforcement_policy_trainer.make_training_pairs + ai.
ProbabilisticPolicyPlayer.get_moves) against the reference's own loop (tests/golden/ref_selfplay.npz, made by
oracle/gen_golden.py `selfplay` from reinforcement_policy_trainer.py:16-76 and ai.py:24-68): the moves drawn, the positions
they are paired with, the ages the encoder is given (the reference's copied states share their history list, go.py:167)
and the winners.  The encoder itself is not run here: what would be sent to it is captured and compared with the stone
and turns_since planes of the reference's tensors.  tests/test_rl_trainer_gpu.py does the same with the CUDA encoder."""
import os

import numpy as np
import pytest

from alphago_b200 import go, preprocessing
from alphago_b200.ai import ProbabilisticPolicyPlayer
from alphago_b200.reinforcement_policy_trainer import make_training_pairs


class FakePolicy(object):
    """the stand-in network of oracle/gen_golden.py `selfplay`"""

    def __init__(self, salt, size=19, planes=46):
        self.salt = salt
        self.model = type("M", (), {"input_shape": (None, planes, size, size)})()

    def batch_eval_state(self, states, moves_lists=None):
        return [[(m, 1.0 + ((m[0] * 7 + m[1] * 13 + len(st.history) * 3 + self.salt) % 11)) for m in ml]
                for st, ml in zip(states, moves_lists)]


@pytest.mark.parametrize("tag", ["black", "white"])
def test_selfplay_loop_host_logic_equals_the_reference(golden_dir, tag, monkeypatch):
    g = np.load(os.path.join(golden_dir, "ref_selfplay.npz"))
    feats = [str(f) for f in g["features"]]
    captured = []

    def capture(self, board, age, player, ko, size, device="cuda"):
        captured.append((np.array(board), np.array(age), np.array(player)))
        return np.zeros((len(player), self.output_dim, size, size), np.uint8)
    monkeypatch.setattr(preprocessing.Preprocess, "encode_host", capture)
    learner, opponent = ProbabilisticPolicyPlayer(FakePolicy(3), temperature=0.67), ProbabilisticPolicyPlayer(FakePolicy(8))
    np.random.seed(int(g["seed_" + tag]))
    X_list, y_list, winners = make_training_pairs(learner, opponent, feats, 2)
    assert make_training_pairs.player_color == (go.BLACK if tag == "black" else go.WHITE)
    assert [int(w) for w in winners] == [int(w) for w in g["winners_" + tag]]
    assert len(captured) == 2
    for i in range(2):
        assert np.array_equal(y_list[i].argmax(1), g["y_%s_%d" % (tag, i)]), (tag, i)
        want = np.unpackbits(g["X_%s_%d" % (tag, i)], axis=1)[:, :46 * 361].reshape(-1, 46, 361)
        board, age, player = captured[i]
        assert len(board) == len(want)
        own = (board == player[:, None]).astype(np.uint8)
        opp = (board == -player[:, None]).astype(np.uint8)
        assert np.array_equal(own, want[:, 0]) and np.array_equal(opp, want[:, 1]) and np.array_equal(board == 0, want[:, 2] == 1)
        for t in range(8):   # turns_since planes 4..11 (preprocessing.py:20-43) from the ages handed to the encoder
            assert np.array_equal(((board != 0) & (age == t)).astype(np.uint8), want[:, 4 + t]), (tag, i, t)
