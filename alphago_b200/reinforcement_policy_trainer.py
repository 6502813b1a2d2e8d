"""REINFORCE self-play trainer, drop-in for AlphaGo/training/reinforcement_policy_trainer.py:16-176.

Same functions and meaning: `make_training_pairs(player, opp, features, mini_batch_size)` plays a mini-batch of games in
lock step (both sides move through `ProbabilisticPolicyPlayer.get_moves`, i.e. one encoder launch + one forward pass per
ply for the whole batch) and returns, per game, the learner's (position planes, one-hot move) pairs and the winner;
`train_batch` makes one SGD step per game on the compiled `binary_crossentropy` loss with the learning rate's sign set by
the game's result (rpt.py:95-102); `run` loops over mini-batches.  What changes is where the work happens: positions are
PACKED when they occur (726 bytes) and each game's planes are produced by one encoder launch at the end of the game
instead of one Python feature pass per move, and the update is the CUDA training step.

Three things the reference does are kept because a drop-in must learn the same thing, and can be switched off with
`reference_quirks=False`: (1) the positions cached before Black's move are also the ones paired with White's reply, so a
learner playing White is trained on positions one ply old (rpt.py:56-65: `states_prev` is not refreshed between the two
`do_move` calls); (2) the update is negative iff `winner == -1`, i.e. iff WHITE won, whatever colour the learner had
(rpt.py:95-100); (3) `GameState.copy()` shares the history list with the original (go.py:167), and the cached position is
encoded only AFTER the move(s) were played on the original (rpt.py:39-41), so its turns_since planes are computed from a
history that already holds those moves: every stone looks one (Black learner) or two (White learner) entries older, and a
stone captured by Black's move shows up with age 0 if White then plays on its point.  `tests/test_rl_trainer_gpu.py` holds
this to the reference's own `make_training_pairs`, run from its source on fake policies (`tests/golden/ref_selfplay.npz`).
With the switch off, White's moves are paired with the positions they were played in, ages are the position's own and the
sign is positive iff the learner won.
"""
import argparse
import os

import numpy as np

from . import _lib, go
from .ai import ProbabilisticPolicyPlayer
from .go import GameState
from .models import SGD, CNNPolicy
from .preprocessing import Preprocess, pack_state
from .util import flatten_idx


def make_training_pairs(player, opp, features, mini_batch_size, reference_quirks=True, max_plies=None, fast=False):
    """Make training pairs for a batch of matches.  Returns (X_list, y_list, winners): per game the uint8 planes
    (n_moves, F, size, size) of the positions the learner moved in, the one-hot moves (n_moves, size*size), and
    `GameState.get_winner()`.  After the call `make_training_pairs.player_color` holds the learner's colour.
    max_plies (not in the reference) ends games early for tests; fast=True lets the players pick their moves with
    `get_moves_fast` (masking and sampling on the device) instead of per-state Python move lists."""
    preprocessor = Preprocess(features)
    bsize = player.policy.model.input_shape[-1]
    bsize_flat = bsize * bsize
    packed = [list() for _ in range(mini_batch_size)]
    moves_made = [list() for _ in range(mini_batch_size)]
    states = [GameState(bsize) for _ in range(mini_batch_size)]
    # Randomly choose who goes first (i.e. color of 'player')
    player_color = np.random.choice([go.BLACK, go.WHITE])
    make_training_pairs.player_color = int(player_color)
    player1, player2 = (player, opp) if player_color == go.BLACK else (opp, player)

    def do_move(prev_packed, moves):
        for i, (st, mv) in enumerate(zip(states, moves)):
            if not st.is_end_of_game:   # only do more moves if not end of game already
                st.do_move(mv)
                if st.current_player != player_color and mv is not go.PASS_MOVE:   # the learner just moved
                    pk = prev_packed[i]
                    if reference_quirks:
                        # the reference encodes its cached copy NOW, and the copy shares the history list that has grown
                        # meanwhile (go.py:167, rpt.py:39-41): the literal turns_since walk over (history now, board then)
                        hist = np.asarray([-1 if m is None else m[0] * bsize + m[1] for m in st.history] or [0], np.int16)
                        age = np.empty(bsize * bsize, np.uint8)
                        _lib.lib().ag_age_from_history(bsize, _lib.np_ptr(pk[0]), _lib.np_ptr(hist), len(st.history), _lib.np_ptr(age))
                        pk = (pk[0], age, pk[2], pk[3])
                    packed[i].append(pk)
                    moves_made[i].append(flatten_idx(mv, bsize))

    plies = 0
    while True:
        prev = [pack_state(st) for st in states]   # cache states before moves (copies: the engine packs into new arrays)
        do_move(prev, player1.get_moves_fast(states) if fast else player1.get_moves(states))
        if not reference_quirks:
            prev = [pack_state(st) for st in states]
        do_move(prev, player2.get_moves_fast(states) if fast else player2.get_moves(states))
        plies += 2
        if all(st.is_end_of_game for st in states) or (max_plies is not None and plies >= max_plies):
            break
    winners = [st.get_winner() for st in states]
    X_list, y_list = [], []
    for i in range(mini_batch_size):
        n = len(packed[i])
        if n == 0:
            X_list.append(np.zeros((0, preprocessor.output_dim, bsize, bsize), np.uint8))
            y_list.append(np.zeros((0, bsize_flat)))
            continue
        board = np.stack([p[0] for p in packed[i]]); age = np.stack([p[1] for p in packed[i]])
        pl = np.asarray([p[2] for p in packed[i]], np.int8); ko = np.asarray([p[3] for p in packed[i]], np.int16)
        X_list.append(preprocessor.encode_host(board, age, pl, ko, bsize))
        y = np.zeros((n, bsize_flat))
        y[np.arange(n), moves_made[i]] = 1
        y_list.append(y)
    return X_list, y_list, winners


def train_batch(player, X_list, y_list, winners, lr, reference_quirks=True, player_color=None):
    """Given the outcomes of a mini-batch of play against a fixed opponent, update the weights with reinforcement
    learning: one fit() per game, learning rate negative for a lost game (rpt.py:82-104)."""
    for X, y, winner in zip(X_list, y_list, winners):
        if len(X) == 0:
            continue
        if reference_quirks:
            lost = winner == -1
        else:
            lost = winner != (player_color if player_color is not None else getattr(make_training_pairs, "player_color", go.BLACK))
        # setting the learning rate negative is the reference's way of making a negative weights update
        player.policy.model.optimizer.lr.set_value(-lr if lost else lr)
        player.policy.model.fit(X, y, nb_epoch=1, batch_size=len(X))
    return player


def run(player, args, opponents, features, reference_quirks=True):
    # Set SGD and compile
    sgd = SGD(lr=args.learning_rate)
    player.policy.model.compile(loss="binary_crossentropy", optimizer=sgd)
    player_wins_per_batch = []
    for i_iter in range(args.iterations):
        for i_batch in range(args.save_every):
            opp = opponents[np.random.randint(len(opponents))]   # randomly choose opponent from pool
            if getattr(args, "device_games", False):
                # the games live on the device (selfplay.py): a ply is encode -> forward -> one step kernel, no host round trip
                from .selfplay import make_training_pairs_device
                X_list, y_list, winners = make_training_pairs_device(player, opp, features, args.game_batch_size, reference_quirks,
                                                                     getattr(args, "max_plies", None))
                make_training_pairs.player_color = make_training_pairs_device.player_color
            else:
                X_list, y_list, winners = make_training_pairs(player, opp, features, args.game_batch_size, reference_quirks,
                                                              getattr(args, "max_plies", None))
            n_wins = int(np.sum(np.array(winners) == 1))
            player_wins_per_batch.append(n_wins)
            print("Number of wins this batch: {}/{}".format(n_wins, args.game_batch_size))
            player = train_batch(player, X_list, y_list, winners, args.learning_rate, reference_quirks)
    run.player_wins_per_batch = player_wins_per_batch
    return opponents


def main(cmd_line_args=None):
    parser = argparse.ArgumentParser(description="Perform reinforcement learning to improve given policy network. Second phase of pipeline.")
    parser.add_argument("initial_weights", help="Path to file with weights to start from.")
    parser.add_argument("initial_json", help="Path to file with initial network params.")
    parser.add_argument("--model_folder", help="Path to folder where the model params will be saved after each epoch. Default: None", default=None)
    parser.add_argument("--learning_rate", help="Keras learning rate (Default: .03)", type=float, default=.03)
    parser.add_argument("--save_every", help="Save policy every n mini-batches (Default: 500)", type=int, default=500)
    parser.add_argument("--game_batch_size", help="Number of games per mini-batch (Default: 20)", type=int, default=20)
    parser.add_argument("--iterations", help="Number of training iterations (i.e. mini-batch) (Default: 20)", type=int, default=20)
    parser.add_argument("--device-games", help="Keep the games of a mini-batch in device memory (added flag; moves are then drawn from torch's generator)", default=False, action="store_true")
    args = parser.parse_args(cmd_line_args)
    # the reference ignores both positional files for now and starts from a fresh default net (rpt.py:155-165); here they are
    # honoured when they exist
    features = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after", "sensibleness", "zeros"]
    if os.path.exists(args.initial_json):
        policy = CNNPolicy.load_model(args.initial_json)
        features = policy.preprocessor.feature_list
        if os.path.exists(args.initial_weights):
            policy.model.load_weights(args.initial_weights)
    else:
        policy = CNNPolicy(features)
    player = ProbabilisticPolicyPlayer(policy)
    opponents = [player]   # no opponent pool yet (rpt.py:167-176): play against self
    return run(player, args, opponents, features)


if __name__ == "__main__":
    main()
