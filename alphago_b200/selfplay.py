"""Device-resident self-play: the games of a REINFORCE mini-batch live in device memory and a ply is
encode -> policy forward -> `ag_selfplay_step`, with no host round trip.

`DeviceGames` holds what `GameState` holds for a batch of games (AlphaGo/go.py:13-45: board, ko, side to move, pass
counters, end-of-game flag, and the turns_since ages the encoder reads instead of a history list).
`make_training_pairs_device` is `reinforcement_policy_trainer.make_training_pairs` (rpt.py:16-76) on top of it: same
arguments, same meaning of the result, but the positions and moves of the learner stay on the device until the games are
over and the planes of ALL recorded positions come from one encoder launch.  Moves are drawn as
`ProbabilisticPolicyPlayer` draws them (ai.py:38-68: the policy over the sensible moves, raised to 1/temperature) from
uniform numbers generated on the device, so the random stream differs from the host players' `np.random.choice`.
Of the reference behaviours `reinforcement_policy_trainer.make_training_pairs` keeps under `reference_quirks=True`, this path
reproduces the pairing of a White learner's moves with positions one ply old; the turns_since ageing through the shared
history list of copied states (go.py:167) is NOT reproduced here: positions carry their own ages.
"""
import ctypes as C

import numpy as np

from . import _lib, go
from .preprocessing import FEATURE_SIZES, Preprocess


def _plane_offset(feature_list, name):
    off = 0
    for f in feature_list:
        if f.lower() == name:
            return off
        off += FEATURE_SIZES[f.lower()]
    raise ValueError("the policy's feature list needs %r (it marks the moves a player may choose, ai.py:15)" % name)


class DeviceGames(object):
    """`games` GameStates of one board size in device memory"""

    def __init__(self, games, size=19, komi=7.5):
        import torch
        self.torch, self.games, self.size, self.komi = torch, int(games), int(size), float(komi)
        n = self.size * self.size
        dev = "cuda"
        self.board = torch.zeros((self.games, n), dtype=torch.int8, device=dev)
        self.age = torch.full((self.games, n), 255, dtype=torch.uint8, device=dev)
        self.player = torch.ones(self.games, dtype=torch.int8, device=dev)       # BLACK moves first (go.py:16)
        self.ko = torch.full((self.games,), -1, dtype=torch.int16, device=dev)
        self.flags = torch.zeros(self.games, dtype=torch.uint8, device=dev)      # bit 0 is_end_of_game, bit 1 last move was a pass
        self.passes = torch.zeros((self.games, 2), dtype=torch.int32, device=dev)
        self.moves = torch.empty(self.games, dtype=torch.int16, device=dev)
        self.plies = 0

    def encode(self, preprocessor):
        """uint8 planes (games, F, size, size) of the current positions"""
        return preprocessor.encode_packed(self.board, self.age, self.player, self.ko, self.size)

    def step(self, probs, mask, beta=1.0, uniform=None):
        """one ply in every unfinished game.  probs float32 (games, size*size) on the device; mask: uint8 view whose row g
        marks the points game g may play (e.g. planes[:, k] of the sensibleness plane: a strided view is fine);
        uniform: float32 (games,) in [0, 1), drawn with torch.rand when not given.  Returns the moves tensor (int16:
        x*size+y, -1 pass, -2 game already over), valid until the next call."""
        torch = self.torch
        n = self.size * self.size
        if uniform is None:
            uniform = torch.rand(self.games, device="cuda", dtype=torch.float32)
        if mask.dtype != torch.uint8 or mask.shape[0] != self.games or mask[0].numel() != n or not mask[0].is_contiguous():
            raise ValueError("mask must be uint8 (games, size*size) or (games, size, size) with contiguous rows")
        stride = mask.stride(0)
        if not probs.is_contiguous() or probs.dtype != torch.float32 or tuple(probs.shape) != (self.games, n):
            raise ValueError("probs must be contiguous float32 (games, size*size)")
        _lib.check(_lib.lib().ag_selfplay_step(
            C.c_void_p(self.board.data_ptr()), C.c_void_p(self.age.data_ptr()), C.c_void_p(self.player.data_ptr()),
            C.c_void_p(self.ko.data_ptr()), C.c_void_p(self.flags.data_ptr()), C.c_void_p(self.passes.data_ptr()),
            C.c_void_p(probs.data_ptr()), C.c_void_p(mask.data_ptr()), int(stride), C.c_void_p(uniform.data_ptr()), float(beta),
            self.games, self.size, C.c_void_p(self.moves.data_ptr()), _lib.current_stream()), "ag_selfplay_step")
        self.plies += 1
        return self.moves

    def all_over(self):
        """synchronises: one byte per game comes back"""
        return bool((self.flags & 1).all().item())

    def winners(self):
        """GameState.get_winner() of every game (go.py:269-293) as a list of +1 / -1 / 0"""
        torch = self.torch
        out = torch.empty(self.games, dtype=torch.int8, device="cuda")
        _lib.check(_lib.lib().ag_selfplay_winner(C.c_void_p(self.board.data_ptr()), C.c_void_p(self.passes.data_ptr()), self.games,
                                                 self.size, self.komi, C.c_void_p(out.data_ptr()), _lib.current_stream()),
                   "ag_selfplay_winner")
        return [int(v) for v in out.cpu().numpy()]


def policy_ply(games, player, planes=None):
    """one ply of `player` (a ProbabilisticPolicyPlayer) in all games: encode, forward, sample + play on the device"""
    pre = player.policy.preprocessor
    if planes is None:
        planes = games.encode(pre)
    probs = player.policy.model.forward_device(planes)
    off = _plane_offset(pre.feature_list, "sensibleness")
    return games.step(probs, planes[:, off], beta=player.beta)


def make_training_pairs_device(player, opp, features, mini_batch_size, reference_quirks=True, max_plies=None, check_every=16):
    """rpt.py:16-76 with the games on the device.  Returns (X_list, y_list, winners): per game a uint8 DEVICE tensor
    (n_moves, F, size, size) of the positions the learner moved in, an int32 numpy array of its moves (x*size+y; the
    reference's one-hot rows are `np.eye(size*size)[y]`), and the winner.  `SequentialModel.fit` takes both forms.
    reference_quirks as in reinforcement_policy_trainer.make_training_pairs (a White learner is paired with positions one ply
    old).  check_every: plies between looks at the end-of-game flags (the only host synchronisation of the loop)."""
    import torch
    bsize = player.policy.model.input_shape[-1]
    n = bsize * bsize
    G = int(mini_batch_size)
    games = DeviceGames(G, bsize)
    player_color = int(np.random.choice([go.BLACK, go.WHITE]))
    make_training_pairs_device.player_color = player_color
    player1, player2 = (player, opp) if player_color == go.BLACK else (opp, player)
    rec = {"board": [], "age": [], "player": [], "ko": [], "move": []}

    def snapshot():
        return (games.board.clone(), games.age.clone(), games.player.clone(), games.ko.clone())

    def record(prev, moves):
        rec["board"].append(prev[0]); rec["age"].append(prev[1]); rec["player"].append(prev[2]); rec["ko"].append(prev[3])
        rec["move"].append(moves.clone())

    plies = 0
    while True:
        prev = snapshot()                                   # "cache states before moves" (rpt.py:56)
        mv = policy_ply(games, player1)
        if player_color == go.BLACK:
            record(prev, mv)
        if not reference_quirks:
            prev = snapshot()
        mv = policy_ply(games, player2)
        if player_color == go.WHITE:
            record(prev, mv)
        plies += 2
        if max_plies is not None and plies >= max_plies:
            break
        if plies % check_every == 0 and games.all_over():
            break
    winners = games.winners()
    T = len(rec["move"])
    moves = torch.stack(rec["move"]).cpu().numpy()          # (T, G): >= 0 where the learner placed a stone
    keep = moves >= 0                                       # passes and finished games leave no pair (rpt.py:46-49)
    pre = Preprocess(features)
    # every recorded position of every game through ONE encoder launch, then cut per game
    board = torch.stack(rec["board"]).reshape(T * G, n); age = torch.stack(rec["age"]).reshape(T * G, n)
    pl = torch.stack(rec["player"]).reshape(T * G); ko = torch.stack(rec["ko"]).reshape(T * G)
    # kept (position, game) pairs ordered by game, then by ply: one gather, one encoder launch, per-game views of its output
    flat_keep = np.flatnonzero(keep.reshape(-1))                      # index = ply * G + game
    order = np.argsort(flat_keep % G, kind="stable")
    sel = torch.from_numpy(flat_keep[order]).cuda()
    planes = pre.encode_packed(board[sel].contiguous(), age[sel].contiguous(), pl[sel].contiguous(), ko[sel].contiguous(), bsize)
    counts = keep.sum(0)
    starts = np.concatenate([[0], np.cumsum(counts)])
    X_list = [planes[int(starts[g]):int(starts[g + 1])] for g in range(G)]
    y_list = [moves[:, g][keep[:, g]].astype(np.int32) for g in range(G)]
    make_training_pairs_device.games = games
    return X_list, y_list, winners
