"""GPU tests of device-resident self-play (alphago_b200/selfplay.py, csrc/selfplay.cu): the rules applied on
the device must be the host engine's -- which tests/test_host.py holds to the reference's games move for move -- after every
ply: board, ko, side to move, the turns_since ages the encoder reads, pass counters, end of game, winner."""
import numpy as np
import pytest

from alphago_b200 import go
from alphago_b200.preprocessing import Preprocess

pytestmark = pytest.mark.gpu
FEATS = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after", "sensibleness", "zeros"]


@pytest.mark.parametrize("size,games,plies", [(19, 128, 200), (9, 64, 260), (5, 32, 120)])
def test_device_games_follow_the_host_rules_ply_by_ply(size, games, plies):
    import torch
    from alphago_b200.selfplay import DeviceGames
    torch.manual_seed(size)
    pre = Preprocess(["sensibleness"])
    dg = DeviceGames(games, size)
    host = [go.GameState(size) for _ in range(games)]
    n = size * size
    probs = torch.rand(games, n, device="cuda") + 0.05      # a random "policy": every sensible move has weight
    captures = kos = passes = over = 0
    for t in range(plies):
        planes = dg.encode(pre)
        mv = dg.step(probs, planes[:, 0], beta=1.0).cpu().numpy()
        board, age = dg.board.cpu().numpy(), dg.age.cpu().numpy()
        player, ko, flags = dg.player.cpu().numpy(), dg.ko.cpu().numpy(), dg.flags.cpu().numpy()
        for g, st in enumerate(host):
            if st.is_end_of_game:
                assert mv[g] == -2 and flags[g] & 1, (t, g)
                over += 1
                continue
            stones = int((st.board != 0).sum())
            if mv[g] >= 0:
                move = (int(mv[g]) // size, int(mv[g]) % size)
                assert move in st.get_legal_moves(include_eyes=False), (t, g, move)   # drawn from the sensible moves only
                st.do_move(move)
                captures += int((st.board != 0).sum()) < stones + 1
            else:
                assert mv[g] == -1 and len(st.get_legal_moves(include_eyes=False)) == 0, (t, g)   # pass only when nothing is sensible
                st.do_move(go.PASS_MOVE)
                passes += 1
            b, a, pl, k = st.pack()
            assert np.array_equal(board[g], b), (t, g)
            assert np.array_equal(age[g][b != 0], a[b != 0]), (t, g)
            assert player[g] == pl and ko[g] == k, (t, g, int(ko[g]), k)
            assert bool(flags[g] & 1) == bool(st.is_end_of_game), (t, g)
            kos += k >= 0
        probs = torch.rand(games, n, device="cuda") + 0.05
    want = [st.get_winner() for st in host]
    assert dg.winners() == want
    p = dg.passes.cpu().numpy()
    assert all(p[g, 0] == st.passes_black and p[g, 1] == st.passes_white for g, st in enumerate(host))
    print("size %d: %d captures, %d ko positions, %d passes, %d game-over plies" % (size, captures, kos, passes, over))
    assert captures > 0 and kos > 0
    if size == 5:
        assert passes > 0 and over > 0   # small boards run into passes and finished games


def test_sampling_follows_the_weights():
    """p^beta over the masked points: frequencies against the exact distribution (4-sigma), pass when nothing is allowed"""
    import torch
    from alphago_b200.selfplay import DeviceGames
    games, size = 4096, 5
    n = size * size
    w = torch.tensor([0.0, 1.0, 2.0, 0.5, 4.0] + [0.0] * (n - 5), device="cuda")
    probs = w.repeat(games, 1).contiguous()
    mask = torch.zeros(games, n, dtype=torch.uint8, device="cuda")
    mask[:, :5] = 1
    mask[:, 3] = 0                       # masked out although it has weight
    mask[games - 7:, :] = 0              # nothing allowed -> pass
    dg = DeviceGames(games, size)
    torch.manual_seed(1)
    mv = dg.step(probs, mask, beta=2.0).cpu().numpy()
    assert (mv[games - 7:] == -1).all() and set(np.unique(mv[:games - 7])) <= {1, 2, 4}
    expect = np.array([1.0, 4.0, 16.0]) / 21.0
    m = games - 7
    for point, e in zip((1, 2, 4), expect):
        f = (mv[:m] == point).mean()
        assert abs(f - e) < 4 * np.sqrt(e * (1 - e) / m), (point, f, e)


def test_make_training_pairs_device_and_update():
    import torch
    from alphago_b200.ai import ProbabilisticPolicyPlayer
    from alphago_b200.models import SGD, CNNPolicy
    from alphago_b200.reinforcement_policy_trainer import train_batch
    from alphago_b200.selfplay import make_training_pairs_device
    pol = CNNPolicy(FEATS, board=7, layers=3, filters_per_layer=32, seed=2, max_batch=64)
    pl = ProbabilisticPolicyPlayer(pol, temperature=0.7)
    torch.manual_seed(0)
    np.random.seed(4)
    for quirks in (False, True):
        X, y, winners = make_training_pairs_device(pl, pl, FEATS, 6, reference_quirks=quirks, max_plies=40)
        colour = make_training_pairs_device.player_color
        assert len(X) == len(y) == len(winners) == 6 and all(w in (-1, 0, 1) for w in winners)
        for Xg, yg in zip(X, y):
            assert Xg.is_cuda and Xg.dtype == torch.uint8 and tuple(Xg.shape[1:]) == (pol.preprocessor.output_dim, 7, 7)
            assert len(Xg) == len(yg) and 15 <= len(yg) <= 20
            if not quirks or colour == go.BLACK:   # the move was played on a point that is empty in the paired position
                xe = Xg[:, 2].reshape(len(yg), -1).cpu().numpy()
                assert all(xe[i, yg[i]] == 1 for i in range(len(yg)))
    pol.model.compile(loss="binary_crossentropy", optimizer=SGD(lr=0.05))
    before = [w.copy() for w in pol.model.get_weights()]
    train_batch(pl, X, y, winners, 0.05)
    after = pol.model.get_weights()
    assert any(not np.array_equal(a, b) for a, b in zip(before, after)) and all(np.isfinite(a).all() for a in after)


def test_run_loop_with_device_games():
    """reinforcement_policy_trainer.run with --device-games: the same loop, the mini-batch's games on the device"""
    import argparse
    import torch
    from alphago_b200.ai import ProbabilisticPolicyPlayer
    from alphago_b200.models import CNNPolicy
    from alphago_b200.reinforcement_policy_trainer import run
    learner = ProbabilisticPolicyPlayer(CNNPolicy(FEATS, board=5, layers=3, filters_per_layer=32, seed=4, max_batch=16))
    args = argparse.Namespace(learning_rate=0.02, iterations=1, save_every=2, game_batch_size=3, device_games=True)
    torch.manual_seed(3)
    np.random.seed(5)
    before = [w.copy() for w in learner.policy.model.get_weights()]
    out = run(learner, args, [learner], FEATS)
    assert out == [learner] and len(run.player_wins_per_batch) == 2 and all(0 <= w <= 3 for w in run.player_wins_per_batch)
    after = learner.policy.model.get_weights()
    assert any(not np.array_equal(a, b) for a, b in zip(before, after)) and all(np.isfinite(a).all() for a in after)
