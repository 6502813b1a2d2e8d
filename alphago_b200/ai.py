"""Policy players with the surface of the reference `AlphaGo/ai.py:6-68`: `GreedyPolicyPlayer` and
`ProbabilisticPolicyPlayer` (`get_move`, and the batch form `get_moves` the RL trainer drives, rpt.py:62-67).

A player wraps a policy object with `eval_state(state, moves)` / `batch_eval_state(states, moves_lists)`
(`models.CNNPolicy`); the candidate moves are the legal moves that are not the mover's own eyes (ai.py:15,38,55).  In the
batch form all states go through one encoder launch and one forward pass; sampling uses numpy's global generator like the
reference (`np.random.choice`), so `np.random.seed` makes games reproducible.
"""
import numpy as np

from . import go


def _sensible_moves(state):
    """legal moves that do not fill the mover's own eye: state.get_legal_moves(include_eyes=False) is the same filter
    (go.py:261-267) evaluated inside the rules engine instead of one is_eye call per move from Python"""
    return state.get_legal_moves(include_eyes=False)


class GreedyPolicyPlayer(object):
    """A player that uses a greedy policy (i.e. chooses the highest probability move each turn)"""

    def __init__(self, policy_function):
        self.policy = policy_function

    def get_move(self, state):
        sensible_moves = _sensible_moves(state)
        if len(sensible_moves) > 0:
            move_probs = self.policy.eval_state(state, sensible_moves)
            max_prob = max(move_probs, key=lambda ap: ap[1])
            return max_prob[0]
        return go.PASS_MOVE   # no 'sensible' moves available


class ProbabilisticPolicyPlayer(object):
    """A player that samples a move in proportion to the probability given by the policy.  The 'temperature' pushes play
    towards uniformly random (high) or greedy (low)."""

    def __init__(self, policy_function, temperature=1.0):
        assert (temperature > 0.0)
        self.policy = policy_function
        self.beta = 1.0 / temperature

    def _sample(self, move_probs):
        moves, probabilities = zip(*move_probs)
        probabilities = np.array(probabilities, dtype=np.float64)
        probabilities = probabilities ** self.beta
        probabilities = probabilities / probabilities.sum()
        # numpy reads a list of tuples as 2D: choose an index, then look the move up
        choice_idx = np.random.choice(len(moves), p=probabilities)
        return moves[choice_idx]

    def get_move(self, state):
        sensible_moves = _sensible_moves(state)
        if len(sensible_moves) > 0:
            return self._sample(self.policy.eval_state(state, sensible_moves))
        return go.PASS_MOVE

    def get_moves(self, states):
        """Batch version of get_move. A list of moves is returned (one per state)"""
        sensible_move_lists = [_sensible_moves(st) for st in states]
        all_moves_distributions = self.policy.batch_eval_state(states, sensible_move_lists)
        move_list = [None] * len(states)
        for i, move_probs in enumerate(all_moves_distributions):
            move_list[i] = go.PASS_MOVE if len(move_probs) == 0 else self._sample(move_probs)
        return move_list

    def get_moves_fast(self, states):
        """get_moves without per-move Python objects: the candidate mask is the encoder's own `sensibleness` plane (legal and
        not the mover's eye, preprocessing.py:155-161), the policy is raised to 1/temperature, renormalised and sampled
        on the device (torch.multinomial: a different random stream from get_moves' np.random.choice, same distribution).
        Needs 'sensibleness' among the policy's features; returns a list of (x, y) / PASS_MOVE."""
        import torch
        from .preprocessing import FEATURE_SIZES, pack_state
        pre = self.policy.preprocessor
        off = 0
        for f in pre.feature_list:
            if f.lower() == "sensibleness":
                break
            off += FEATURE_SIZES[f.lower()]
        else:
            raise ValueError("get_moves_fast needs the 'sensibleness' feature plane")
        n = len(states)
        if n == 0:
            return []
        size = states[0].size
        packed = [pack_state(st) for st in states]
        board = torch.from_numpy(np.stack([p[0] for p in packed])).cuda(non_blocking=True)
        age = torch.from_numpy(np.stack([p[1] for p in packed])).cuda(non_blocking=True)
        player = torch.from_numpy(np.asarray([p[2] for p in packed], np.int8)).cuda(non_blocking=True)
        ko = torch.from_numpy(np.asarray([p[3] for p in packed], np.int16)).cuda(non_blocking=True)
        planes = pre.encode_packed(board, age, player, ko, size)
        probs = self.policy.model.forward_device(planes)
        mask = planes[:, off].reshape(n, -1) > 0
        w = (probs.clamp_min(1e-30) ** self.beta) * mask
        any_move = mask.any(1)
        w = torch.where(any_move[:, None], w, torch.ones_like(w))   # rows without a sensible move: sampled, then turned into a pass
        choice = torch.multinomial(w, 1)[:, 0].cpu().numpy()
        has = any_move.cpu().numpy()
        return [divmod(int(c), size) if h else go.PASS_MOVE for c, h in zip(choice, has)]

