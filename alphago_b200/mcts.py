"""Monte Carlo tree search over the policy / value networks: drop-in for AlphaGo/mcts.py:4-175, plus the batched search
the reference leaves as an empty class (`ParallelMCTS`, mcts.py:174-175).

`TreeNode` and `MCTS` keep the reference's names, arguments and arithmetic: one simulation at a time, `playout_depth`
policy calls down the tree, a value call and a greedy rollout at the leaf, the backup from the root downward.  Two
behaviours of the reference are kept because they decide which move comes out: the root's own visit count is never
incremented (it is not on the `visited` list), so the exploration bonus of the root's children is c_puct * P * sqrt(0)
= 0 after their first visit (mcts.py:58,123-124); and the chosen move is the most visited child (mcts.py:157).  One is
not: a simulation that runs into the end of the game leaves `None` entries in `visited` and the reference then fails
on `None.update` (mcts.py:105,123-124); here those entries are skipped.

`ParallelMCTS` runs `n_parallel` simulations in lock step so that the networks see batches: at every
ply the states of all simulations go through ONE `batch_eval_state` call (one encoder launch + one forward pass), the
leaves through one value call, the rollouts through one rollout-policy call per ply.  Simulations of a wave are kept
apart by a virtual loss on the nodes already chosen in that wave.
"""
import numpy as np


class TreeNode(object):
    """Tree representation of MCTS that covers selection, expansion, evaluation and backup (aka 'update()')"""

    def __init__(self, parent, prior_p):
        self.parent = parent
        self.nVisits = 0
        self.Q_value = 0
        self.u_value = prior_p
        self.children = {}
        self.P = prior_p
        self.vloss = 0   # simulations of the current wave that went through this node (ParallelMCTS only)

    def expansion(self, actions):
        """Expand subtree: `actions` is the policy function's output, a list of ((x, y), prior probability); moves
        that already have a child keep it (mcts.py:15-27)"""
        for action, prob in actions:
            if action not in self.children:
                self.children[action] = TreeNode(self, prob)

    def selection(self):
        """(action, next_node) with the largest action value Q plus bonus u(P); the first one wins a tie"""
        return max(self.children.items(), key=lambda an: an[1].toValue())

    def isLeaf(self):
        """no nodes below this have been expanded"""
        return self.children == {}

    def update(self, leaf_value, c_puct):
        """Update node values from a leaf evaluation (mcts.py:45-60): running mean of the leaf values, and
        u = c_puct * P * sqrt(parent visits) / (1 + visits) (not normalised to a distribution)"""
        self.nVisits += 1
        mean_V = self.Q_value * (self.nVisits - 1)
        self.Q_value = (mean_V + leaf_value) / self.nVisits
        self.u_value = c_puct * self.P * np.sqrt(self.parent.nVisits) / (1 + self.nVisits)

    def toValue(self):
        """action value Q plus bonus u(P); every simulation in flight through this node counts as a lost visit"""
        if self.vloss:
            n = self.nVisits + self.vloss
            return (self.Q_value * self.nVisits - self.vloss) / n + self.u_value * (1 + self.nVisits) / (1 + n)
        return self.Q_value + self.u_value


class MCTS(object):
    """Monte Carlo tree search, takes a game state, a value network function, a policy network function and a rollout
    policy function; get_move returns an action after the lookahead search is complete.

    The value function takes a state and returns a number in [-1, 1]; the policy and rollout functions take a state
    and return a list of (action, prob) tuples where action is an (x, y) tuple.  0 <= lmbda <= 1 weighs the value
    network against the rollouts in a leaf's value; 0 < c_puct < inf controls how quickly exploration converges to the
    maximum-value policy (mcts.py:68-91)."""

    def __init__(self, state, value_network, policy_network, rollout_policy, lmbda=0.5, c_puct=5, rollout_limit=500,
                 playout_depth=20, n_search=10000):
        self.root = TreeNode(None, 1.0)
        self._value = value_network
        self._policy = policy_network
        self._rollout = rollout_policy
        self._lmbda = lmbda
        self._c_puct = c_puct
        self._rollout_limit = rollout_limit
        self._L = playout_depth
        self._n_search = n_search

    def _DFS(self, nDepth, treenode, state):
        """One simulation: play nDepth moves down the tree with the full policy network, evaluate the leaf, update the
        action values and visit counts of the traversed nodes (mcts.py:93-124).  `state` is modified."""
        visited = [None] * nDepth
        for index in range(nDepth):
            action_probs = self._policy(state)
            if len(action_probs) == 0:   # end of game
                break
            treenode.expansion(action_probs)
            action, treenode = treenode.selection()
            state.do_move(action)
            visited[index] = treenode
        v = self._value(state)
        z = self._evaluate_rollout(state, self._rollout_limit)
        leaf_value = (1 - self._lmbda) * v + self._lmbda * z
        # from the root downward, so that 'parent' visit counts are correct
        for node in visited:
            if node is not None:
                node.update(leaf_value, self._c_puct)

    def _evaluate_rollout(self, state, limit):
        """play the rollout policy's best move until the end of the game; the winner (0 if tie) (mcts.py:126-139)"""
        for i in range(limit):
            action_probs = self._rollout(state)
            if len(action_probs) == 0:
                break
            max_action = max(action_probs, key=lambda ap: ap[1])[0]
            state.do_move(max_action)
        else:
            print("WARNING: rollout reached move limit")
        return state.get_winner()

    def get_move(self, state):
        """run n_search simulations from `state`, then return the most visited child of the root (mcts.py:141-158)"""
        action_probs = self._policy(state)
        self.root.expansion(action_probs)
        for n in range(0, self._n_search):
            state_copy = state.copy()
            self._DFS(self._L, self.root, state_copy)
        return max(self.root.children.items(), key=lambda an: an[1].nVisits)[0]

    def update_with_move(self, last_move):
        """step forward in the tree and discard everything that isn't still reachable (mcts.py:160-169)"""
        if last_move in self.root.children:
            self.root = self.root.children[last_move]
            self.root.parent = None
        else:
            self.root = TreeNode(None, 1.0)


def _batched(fn):
    """states -> list of results: through the owner's batch_eval_state when `fn` is a network (or its eval_state), else
    one call per state"""
    owner = fn if hasattr(fn, "batch_eval_state") else getattr(fn, "__self__", None)
    if owner is not None and hasattr(owner, "batch_eval_state") and getattr(fn, "__name__", "eval_state") == "eval_state":
        return lambda states: list(owner.batch_eval_state(states)) if states else []
    return lambda states: [fn(s) for s in states]


class ParallelMCTS(MCTS):
    """MCTS with `n_parallel` simulations per wave, evaluated in batches.

    Same arguments as MCTS; each network argument may be a CNNPolicy / CNNValue (or its `eval_state`), whose
    `batch_eval_state` then serves a whole wave per call, or any callable state -> result (called state by state).
    With n_parallel=1 a wave is one simulation and the search is MCTS's, move for move.  With lmbda == 0 the rollouts,
    which would not count, are not played."""

    def __init__(self, state, value_network, policy_network, rollout_policy, lmbda=0.5, c_puct=5, rollout_limit=500,
                 playout_depth=20, n_search=10000, n_parallel=16):
        MCTS.__init__(self, state, value_network, policy_network, rollout_policy, lmbda, c_puct, rollout_limit,
                      playout_depth, n_search)
        self._n_parallel = max(1, int(n_parallel))
        self._policy_batch = _batched(policy_network)
        self._value_batch = _batched(value_network)
        self._rollout_batch = _batched(rollout_policy) if rollout_policy is not None else None

    def _wave(self, nDepth, root, states):
        """len(states) simulations from `root` in lock step; `states` are modified"""
        k = len(states)
        nodes = [root] * k
        visited = [[] for _ in range(k)]
        live = list(range(k))
        for index in range(nDepth):
            if not live:
                break
            results = self._policy_batch([states[i] for i in live])
            still = []
            for i, action_probs in zip(live, results):
                if len(action_probs) == 0:   # end of game for this simulation
                    continue
                nodes[i].expansion(action_probs)
                action, nodes[i] = nodes[i].selection()
                nodes[i].vloss += 1              # the next simulation of this wave looks elsewhere
                states[i].do_move(action)
                visited[i].append(nodes[i])
                still.append(i)
            live = still
        values = self._value_batch(states)
        if self._lmbda != 0:
            outcomes = self._rollout_wave(states, self._rollout_limit)
        else:
            outcomes = [0.0] * k
        for i in range(k):
            leaf_value = (1 - self._lmbda) * values[i] + self._lmbda * outcomes[i]
            for node in visited[i]:
                node.vloss -= 1
            for node in visited[i]:
                node.update(leaf_value, self._c_puct)

    def _rollout_wave(self, states, limit):
        """MCTS._evaluate_rollout for all states at once: one rollout-policy batch per ply"""
        live = list(range(len(states)))
        for _ in range(limit):
            if not live:
                break
            results = self._rollout_batch([states[i] for i in live])
            still = []
            for i, action_probs in zip(live, results):
                if len(action_probs) == 0:
                    continue
                states[i].do_move(max(action_probs, key=lambda ap: ap[1])[0])
                still.append(i)
            live = still
        if live:
            print("WARNING: rollout reached move limit")
        return [s.get_winner() for s in states]

    def _DFS(self, nDepth, treenode, state):
        self._wave(nDepth, treenode, [state])

    def get_move(self, state):
        action_probs = self._policy_batch([state])[0]
        self.root.expansion(action_probs)
        done = 0
        while done < self._n_search:
            k = min(self._n_parallel, self._n_search - done)
            self._wave(self._L, self.root, [state.copy() for _ in range(k)])
            done += k
        return max(self.root.children.items(), key=lambda an: an[1].nVisits)[0]
