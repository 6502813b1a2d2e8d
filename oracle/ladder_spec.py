"""Ladder planes: the build-owned specification, in Python, over any GameState-like object.

The reference has no ladder semantics (preprocessing.py:147-152 raise NotImplementedError;
tests/test_preprocessing.py:218-222 are `pass`), so parity for these two planes is UNPINNED
against the reference.  This file states the spec on top of a GameState API (`copy`,
`do_move`, `is_legal`, `board`, `current_player`) so that, run on the *reference's own*
GameState, it inherits the reference's rules and ko quirks; go_oracle.c restates it in C and
tests/test_oracle_ladder.py checks the two agree.  CPU ORACLE -- test infrastructure only.

Spec, with the bounds shared by the C oracle and the CUDA kernel:
  DEPTH = 80 plies, NODES = 160 predicate entries per top-level query (a clean ladder across a 19x19 board
  takes 90; fewer than 1 query in 10,000 on random playouts needs more -- those are scored as budget-exhausted).
Candidates (prey groups, replies) are visited in ascending flat index x*size+y; prey groups
are identified by the first neighbour of the move (order (x-1,y),(x+1,y),(x,y-1),(x,y+1),
go.py:96) that touches them.
"""
DEPTH = 80
NODES = 160


def _nbrs(p, size):
    x, y = p
    return [q for q in ((x - 1, y), (x + 1, y), (x, y - 1), (x, y + 1))
            if 0 <= q[0] < size and 0 <= q[1] < size]


def _group(board, p, size):
    """flood fill: (stones, liberties) of the group at p"""
    col = board[p[0]][p[1]]
    stones, libs, todo = {p}, set(), [p]
    while todo:
        c = todo.pop()
        for q in _nbrs(c, size):
            v = board[q[0]][q[1]]
            if v == 0:
                libs.add(q)
            elif v == col and q not in stones:
                stones.add(q)
                todo.append(q)
    return stones, libs


def _copy(st):
    """copy() of the reference GameState is not safe to keep playing on: it aliases `history`
    (go.py:167) and gives every stone its OWN copy of its group's liberty/group set
    (go.py:174-177), so a later do_move updates only the adjacent stone's set and can miss a
    capture.  Restore the sharing the rules engine relies on (go.py:133-136)."""
    c = st.copy()
    try:
        c.history = list(st.history)
    except AttributeError:  # oracle / product states own their history already
        pass
    if hasattr(c, "group_sets") and hasattr(c, "liberty_sets"):
        done = set()
        for x in range(c.size):
            for y in range(c.size):
                if c.board[x][y] != 0 and (x, y) not in done:
                    grp, libs = c.group_sets[x][y], c.liberty_sets[x][y]
                    for (gx, gy) in grp:
                        c.group_sets[gx][gy] = grp
                        c.liberty_sets[gx][gy] = libs
                    done |= grp
    return c


def _prey_candidates(st, action, colour, want_libs):
    out, seen = [], []
    for q in _nbrs(action, st.size):
        if st.board[q[0]][q[1]] != colour:
            continue
        stones, libs = _group(st.board, q, st.size)
        if any(q in s for s in seen):
            continue
        seen.append(stones)
        if len(libs) == want_libs:
            out.append(q)
    return out


def is_ladder_capture(st, action, prey=None, depth=DEPTH, budget=None):
    budget = [NODES] if budget is None else budget
    if not st.is_legal(action):
        return False
    if depth <= 0 or budget[0] <= 0:
        return True
    budget[0] -= 1
    hunter = st.current_player
    cands = _prey_candidates(st, action, -hunter, 2) if prey is None else [prey]
    for pr in cands:
        t = _copy(st)
        t.do_move(action)
        if t.board[pr[0]][pr[1]] == 0:
            continue
        stones, libs = _group(t.board, pr, t.size)
        escapes = set(libs)
        for s in stones:
            for q in _nbrs(s, t.size):
                if t.board[q[0]][q[1]] == hunter:
                    _, hl = _group(t.board, q, t.size)
                    if len(hl) == 1:
                        escapes |= hl
        escaped = False
        for e in sorted(escapes):
            if is_ladder_escape(t, e, pr, depth - 1, budget):
                escaped = True
                break
        if not escaped:
            return True
    return False


def is_ladder_escape(st, action, prey=None, depth=DEPTH, budget=None):
    budget = [NODES] if budget is None else budget
    if not st.is_legal(action):
        return False
    if depth <= 0 or budget[0] <= 0:
        return False
    budget[0] -= 1
    me = st.current_player
    cands = _prey_candidates(st, action, me, 1) if prey is None else [prey]
    for pr in cands:
        t = _copy(st)
        t.do_move(action)
        if t.board[pr[0]][pr[1]] == 0:
            continue
        _, libs = _group(t.board, pr, t.size)
        if len(libs) >= 3:
            return True
        if len(libs) <= 1:
            continue
        caught = False
        for c in sorted(libs):
            if is_ladder_capture(t, c, pr, depth - 1, budget):
                caught = True
                break
        if not caught:
            return True
    return False


def ladder_planes(st):
    """(ladder_capture, ladder_escape) uint8 planes [size][size] over legal moves"""
    import numpy as np
    cap = np.zeros((st.size, st.size), np.uint8)
    esc = np.zeros((st.size, st.size), np.uint8)
    for (x, y) in st.get_legal_moves():
        cap[x, y] = is_ladder_capture(st, (x, y))
        esc[x, y] = is_ladder_escape(st, (x, y))
    return cap, esc
