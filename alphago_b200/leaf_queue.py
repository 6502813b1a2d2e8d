"""Leaf-evaluation queue for tree search: what `AlphaGo/mcts.py:91-126` does one state at a time --
`self._policy(state)` for the priors of a node and `self._value(state)` for a leaf -- done for many states per launch.

A search (any search: this module does not build the tree) `submit()`s the states it wants evaluated and `flush()`es
when its batch is full or it has nothing else to do.  One flush packs the states on the host, encodes them with ONE
encoder launch (when the value net's feature list is the policy net's plus extra planes, the planes are computed once
and shared), runs the two conv stacks once each and copies back only (N, size*size) probabilities and N values.  Results
have the shape the reference's callables return: a list of ((x, y), probability) over the state's legal moves
(`CNNPolicy.eval_state`, policy.py:81-91) and a float in [-1, 1] (`mcts.py:71`).
"""
import numpy as np

from .preprocessing import pack_state


class LeafEvaluator(object):

    def __init__(self, policy, value=None, max_batch=256):
        self.policy, self.value = policy, value
        self.max_batch = int(max_batch)
        pf = [f.lower() for f in policy.preprocessor.feature_list]
        vf = [f.lower() for f in value.preprocessor.feature_list] if value is not None else None
        # one encoding serves both nets when the value features start with the policy's (the usual 48 planes + colour)
        self.shared = vf is not None and vf[:len(pf)] == pf
        self._states, self._moves = [], []

    def __len__(self):
        return len(self._states)

    def submit(self, state, moves=None, want_moves=True):
        """queue a state (it is packed now: the caller may go on mutating it); returns its index in the next flush.
        want_moves=False skips the host-side legal-move list (for flush_arrays, which masks on the device)"""
        self._states.append((state.size, pack_state(state)))
        self._moves.append(list(moves) if moves is not None else (state.get_legal_moves() if want_moves else None))
        return len(self._states) - 1

    def full(self):
        return len(self._states) >= self.max_batch

    def _encode(self, pre, packed, size):
        import torch
        board = torch.from_numpy(np.stack([p[0] for p in packed])).cuda(non_blocking=True)
        age = torch.from_numpy(np.stack([p[1] for p in packed])).cuda(non_blocking=True)
        player = torch.from_numpy(np.asarray([p[2] for p in packed], np.int8)).cuda(non_blocking=True)
        ko = torch.from_numpy(np.asarray([p[3] for p in packed], np.int16)).cuda(non_blocking=True)
        return pre.encode_packed(board, age, player, ko, size)

    def _legal_plane_range(self):
        """plane range of the policy features whose union is exactly the legal moves (capture_size sets one of its 8 planes
        on every legal move, preprocessing.py:64-88), or None"""
        off = 0
        from .preprocessing import FEATURE_SIZES
        for f in self.policy.preprocessor.feature_list:
            f = f.lower()
            if f == "capture_size":
                return off, off + FEATURE_SIZES[f]
            off += FEATURE_SIZES[f]
        return None

    def _run(self, masked):
        n = len(self._states)
        size = self._states[0][0]
        if not all(s == size for s, _ in self._states):
            raise ValueError("all states must have the same size")   # policy.py:69
        packed = [p for _, p in self._states]
        moves, self._states, self._moves = self._moves, [], []
        out_p = np.empty((n, size * size), np.float32)
        out_v = np.empty(n, np.float32) if self.value is not None else None
        legal = self._legal_plane_range() if masked else None
        if masked and legal is None:
            raise ValueError("flush_arrays needs 'capture_size' among the policy features (its planes mark the legal moves)")
        for s in range(0, n, self.max_batch):
            chunk = packed[s:s + self.max_batch]
            if self.shared:
                planes_v = self._encode(self.value.preprocessor, chunk, size)
                planes_p = planes_v[:, :self.policy.preprocessor.output_dim].contiguous()
            else:
                planes_p = self._encode(self.policy.preprocessor, chunk, size)
                planes_v = self._encode(self.value.preprocessor, chunk, size) if self.value is not None else None
            probs = self.policy.model.forward_device(planes_p)
            vals = self.value.model.forward_device(planes_v) if self.value is not None else None
            if masked:   # renormalise over the legal moves on the device (what _select_moves_and_normalize does per state)
                mask = planes_p[:, legal[0]:legal[1]].sum(1).reshape(len(chunk), -1) > 0
                probs = probs * mask
                probs = probs / probs.sum(1, keepdim=True).clamp_min(1e-30)
            out_p[s:s + len(chunk)] = probs.cpu().numpy()
            if vals is not None:
                out_v[s:s + len(chunk)] = vals.cpu().numpy().reshape(-1)
        return n, size, moves, out_p, out_v

    def flush_arrays(self):
        """evaluate everything queued, results as arrays: (P float32 (N, size*size), V float32 (N,) or None) with P[i]
        the policy renormalised over state i's legal moves and 0 elsewhere (index x*size+y, util.flatten_idx).  No
        per-move Python objects: the fast path for a search that keeps its priors in arrays."""
        if len(self._states) == 0:
            return np.zeros((0, 0), np.float32), (np.zeros(0, np.float32) if self.value is not None else None)
        n, size, moves, out_p, out_v = self._run(masked=True)
        return out_p, out_v

    def flush(self):
        """evaluate everything queued: list of (priors, value) in submit order; value is None without a value net"""
        if len(self._states) == 0:
            return []
        n, size, moves, out_p, out_v = self._run(masked=False)
        results = []
        for i in range(n):
            priors = self.policy._select_moves_and_normalize(out_p[i], moves[i], size)
            results.append((priors, float(out_v[i]) if out_v is not None else None))
        return results

    def evaluate(self, states, moves_lists=None):
        """submit + flush for a list of states"""
        for i, st in enumerate(states):
            self.submit(st, None if moves_lists is None else moves_lists[i])
        return self.flush()
