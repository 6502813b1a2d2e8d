"""CPU tests of alphago_b200/mcts.py after the reference's tests/test_mcts.py:8-50 (same toy networks), plus the batched
search: ParallelMCTS with one simulation per wave must build the tree MCTS builds, and with many per wave it must send
its states through the networks' batch entry points."""
import numpy as np

from alphago_b200.go import GameState
from alphago_b200.mcts import MCTS, ParallelMCTS, TreeNode


def policy_network(state):
    moves = state.get_legal_moves(include_eyes=False)
    # 'random' distribution over positions that is smallest at (0,0) and largest at (18,18)
    probs = np.arange(361, dtype=float)
    probs = probs / probs.sum()
    return list(zip(moves, probs))


def value_network(state):
    return 0.0   # it's not very confident


def rollout_policy(state):
    return policy_network(state)


def test_treenode_selection():
    gs = GameState()
    treenode = TreeNode(None, 1.0)
    treenode.expansion(policy_network(gs))
    action, node = treenode.selection()
    assert action == (18, 18) and node is not None   # according to the policy above
    assert not treenode.isLeaf() and node.isLeaf() and node.parent is treenode


def test_treenode_update_arithmetic():
    root = TreeNode(None, 1.0)
    root.expansion([((0, 0), 0.25)])
    child = root.children[(0, 0)]
    assert child.toValue() == 0.25                      # u starts as the prior (mcts.py:10)
    root.nVisits = 4
    child.update(1.0, 5)
    child.update(0.0, 5)
    assert child.nVisits == 2 and child.Q_value == 0.5
    assert np.isclose(child.u_value, 5 * 0.25 * np.sqrt(4) / 3)   # c_puct * P * sqrt(parent visits) / (1 + visits)


def test_mcts_DFS():
    gs = GameState()
    mcts = MCTS(gs, value_network, policy_network, rollout_policy, n_search=2, rollout_limit=20)
    treenode = TreeNode(None, 1.0)
    mcts._DFS(8, treenode, gs.copy())
    assert treenode.children[(18, 18)].nVisits == 1, "DFS visits incorrect"
    node, depth = treenode, 0
    while not node.isLeaf():   # exactly one path of 8 visited nodes
        visited = [c for c in node.children.values() if c.nVisits]
        if not visited:
            break
        assert len(visited) == 1
        node, depth = visited[0], depth + 1
    assert depth == 8


def test_mcts_get_move_and_update_with_move():
    gs = GameState()
    mcts = MCTS(gs, value_network, policy_network, rollout_policy, n_search=2, rollout_limit=10, playout_depth=4)
    move = mcts.get_move(gs)
    assert move in mcts.root.children and len(gs.history) == 0   # the search worked on copies
    kept = mcts.root.children[move]
    mcts.update_with_move(move)
    assert mcts.root is kept and mcts.root.parent is None
    mcts.update_with_move((0, 0) if move != (0, 0) else (1, 1))   # a move the tree has not seen: fresh root
    assert mcts.root.isLeaf() and mcts.root.parent is None


def test_simulation_that_reaches_the_end_of_the_game_is_backed_up():
    """a policy with no moves ends the descent early; the nodes reached so far are still updated (the reference
    fails on the None entries of its visited list, mcts.py:105,123-124)"""
    calls = {"n": 0}

    def short_policy(state):
        calls["n"] += 1
        return policy_network(state) if calls["n"] <= 2 else []

    gs = GameState()
    mcts = MCTS(gs, value_network, short_policy, lambda s: [], n_search=1, playout_depth=6)
    root = TreeNode(None, 1.0)
    mcts._DFS(6, root, gs.copy())
    assert root.children[(18, 18)].nVisits == 1


def _tree_signature(node, depth=0, out=None, path=()):
    out = [] if out is None else out
    for action, child in node.children.items():
        if child.nVisits:
            out.append((path + (action,), child.nVisits, round(float(child.Q_value), 6), round(float(child.u_value), 6)))
            _tree_signature(child, depth + 1, out, path + (action,))
    return sorted(out)


def test_parallel_mcts_with_one_simulation_per_wave_is_mcts():
    gs = GameState(9)
    for mv in [(2, 2), (6, 6), (2, 6)]:
        gs.do_move(mv)

    def pol(state):
        moves = state.get_legal_moves(include_eyes=False)
        w = np.asarray([1.0 + ((x * 7 + y * 3) % 5) for x, y in moves])
        return list(zip(moves, w / w.sum()))

    def val(state):
        return 0.1 * (len(state.history) % 3 - 1)

    kw = dict(lmbda=0.5, c_puct=5, rollout_limit=200, playout_depth=3, n_search=12)
    a = MCTS(gs, val, pol, pol, **kw)
    b = ParallelMCTS(gs, val, pol, pol, n_parallel=1, **kw)
    assert a.get_move(gs) == b.get_move(gs)
    assert _tree_signature(a.root) == _tree_signature(b.root)
    assert all(n.vloss == 0 for n in b.root.children.values())


class _BatchNet(object):
    """stands in for CNNPolicy / CNNValue: counts what goes through eval_state and through batch_eval_state"""

    def __init__(self, kind):
        self.kind, self.single, self.batches = kind, 0, []

    def _one(self, state):
        if self.kind == "value":
            return 0.05
        moves = state.get_legal_moves(include_eyes=False)
        return [(m, 1.0 / len(moves)) for m in moves]

    def eval_state(self, state, moves=None):
        self.single += 1
        return self._one(state)

    def batch_eval_state(self, states, moves_lists=None):
        self.batches.append(len(states))
        return [self._one(s) for s in states]


def test_parallel_mcts_batches_its_network_calls():
    gs = GameState(9)
    policy, value = _BatchNet("policy"), _BatchNet("value")
    search = ParallelMCTS(gs, value.eval_state, policy.eval_state, None, lmbda=0, playout_depth=3, n_search=20, n_parallel=8)
    move = search.get_move(gs)
    assert move in search.root.children
    assert policy.single == 0 and value.single == 0                 # everything went through the batch entry points
    assert policy.batches == [1] + [8, 8, 8] * 2 + [4, 4, 4]        # root, then one call per ply per wave (8 + 8 + 4 simulations)
    assert value.batches == [8, 8, 4]
    assert sum(c.nVisits for c in search.root.children.values()) == 20
    # virtual loss spreads a wave over several children and is gone afterwards
    assert sum(1 for c in search.root.children.values() if c.nVisits) > 1

    def no_vloss(node):
        return node.vloss == 0 and all(no_vloss(c) for c in node.children.values())
    assert no_vloss(search.root)


def test_parallel_mcts_rollouts_in_lock_step():
    gs = GameState(7)
    policy, value, rollout = _BatchNet("policy"), _BatchNet("value"), _BatchNet("policy")
    search = ParallelMCTS(gs, value, policy, rollout, lmbda=1.0, playout_depth=1, n_search=4, n_parallel=4, rollout_limit=300)
    search.get_move(gs)
    assert len(rollout.batches) > 10 and rollout.batches[0] == 4 and rollout.single == 0
    qs = [c.Q_value for c in search.root.children.values() if c.nVisits]
    assert qs and all(q in (-1, 0, 1) for q in qs)                  # lmbda = 1: the leaf value is the rollout's winner
