"""Go rules state with the surface of the reference `AlphaGo.go.GameState` (go.py:9-349).

Same attribute and method names, same conventions ((x, y) tuples, BLACK=+1 / WHITE=-1, PASS_MOVE
= None) and the same observable quirks, so code written against the reference (players, MCTS,
the trainers) keeps working.  The bookkeeping is not the reference's per-point Python sets: the
position lives in the C++ engine of libalphago_b200.so (csrc/host_go.hpp) and `liberty_sets`,
`group_sets`, `liberty_counts` are derived views built on demand.
"""
import ctypes as C

import numpy as np

from . import _lib

WHITE = -1
BLACK = +1
EMPTY = 0
PASS_MOVE = None


class IllegalMove(Exception):
    pass


class GameState(object):
    """State of a game of Go (go.py:9-45)."""

    def __init__(self, size=19, komi=7.5, _handle=None):
        L = _lib.lib()
        self._h = _handle if _handle is not None else L.ag_state_new(size)
        if not self._h:
            raise ValueError(_lib.last_error())
        self.size = size
        self.komi = komi

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().ag_state_free(h)
            except Exception:
                pass

    # ------------------------------------------------------------ scalar state
    def _get(self, field):
        return _lib.lib().ag_state_get(self._h, field)

    @property
    def current_player(self):
        return self._get(1)

    @current_player.setter
    def current_player(self, colour):
        _lib.check(_lib.lib().ag_state_set_player(self._h, int(colour)), "set current_player")

    @property
    def ko(self):
        k = self._get(2)
        return None if k < 0 else divmod(k, self.size)

    @ko.setter
    def ko(self, pos):
        _lib.lib().ag_state_set_ko(self._h, -1 if pos is None else pos[0] * self.size + pos[1])

    @property
    def is_end_of_game(self):
        return bool(self._get(4))

    @property
    def turns_played(self):
        return self._get(5)

    @property
    def num_black_prisoners(self):
        return self._get(6)

    @property
    def num_white_prisoners(self):
        return self._get(7)

    @property
    def passes_black(self):
        return self._get(8)

    @property
    def passes_white(self):
        return self._get(9)

    # ------------------------------------------------------------ array views
    @property
    def board(self):
        """numpy (size, size) indexed [x][y] with values in {-1, 0, +1} (go.py:14)"""
        b = np.empty(self.size * self.size, np.int8)
        _lib.lib().ag_state_board(self._h, _lib.np_ptr(b))
        return b.reshape(self.size, self.size)

    @property
    def history(self):
        n = self._get(3)
        h = np.empty(max(n, 1), np.int16)
        _lib.lib().ag_state_history(self._h, _lib.np_ptr(h))
        return [PASS_MOVE if m < 0 else divmod(int(m), self.size) for m in h[:n]]

    def _groups(self):
        n = self.size * self.size
        gid = np.empty(n, np.int16); st = np.empty(n, np.int16); lb = np.empty(n, np.int16)
        _lib.lib().ag_state_groups(self._h, _lib.np_ptr(gid), _lib.np_ptr(st), _lib.np_ptr(lb))
        return gid, st, lb

    @property
    def liberty_counts(self):
        """liberties of the group at each point, -1 on empty points (go.py:40-41)"""
        return self._groups()[2].astype(np.int64).reshape(self.size, self.size)

    @property
    def group_sets(self):
        """group_sets[x][y]: set of (x, y) in the group at that point, empty set on empty points (go.py:45)"""
        gid = self._groups()[0]
        sets = {}
        for p, g in enumerate(gid):
            if g >= 0:
                sets.setdefault(int(g), set()).add(divmod(p, self.size))
        return [[sets[int(gid[x * self.size + y])] if gid[x * self.size + y] >= 0 else set()
                 for y in range(self.size)] for x in range(self.size)]

    @property
    def liberty_sets(self):
        """liberty_sets[x][y]: liberties of the group at (x, y); for an empty point its empty
        neighbours (go.py:34-37)"""
        b = self.board
        gid = self._groups()[0]
        per_group = {}
        out = [[None] * self.size for _ in range(self.size)]
        for x in range(self.size):
            for y in range(self.size):
                empt = set(q for q in self._neighbors((x, y)) if b[q] == EMPTY)
                g = int(gid[x * self.size + y])
                if g < 0:
                    out[x][y] = empt
                else:
                    per_group.setdefault(g, set()).update(empt)
        for x in range(self.size):
            for y in range(self.size):
                g = int(gid[x * self.size + y])
                if g >= 0:
                    out[x][y] = per_group[g]
        return out

    # ------------------------------------------------------------ queries
    def _idx(self, action):
        if action is PASS_MOVE:
            return -1
        (x, y) = action
        if not self._on_board(action):
            return self.size * self.size  # off-board: never legal
        return x * self.size + y

    def _on_board(self, position):
        (x, y) = position
        return x >= 0 and y >= 0 and x < self.size and y < self.size

    def _neighbors(self, position):
        (x, y) = position
        return [p for p in [(x - 1, y), (x + 1, y), (x, y - 1), (x, y + 1)] if self._on_board(p)]

    def _diagonals(self, position):
        (x, y) = position
        return [p for p in [(x - 1, y - 1), (x + 1, y + 1), (x + 1, y - 1), (x - 1, y + 1)] if self._on_board(p)]

    def get_group(self, position):
        return self.group_sets[position[0]][position[1]]

    def get_groups_around(self, position):
        """unique non-empty groups adjacent to position (go.py:63-83)"""
        gs, groups, b = self.group_sets, [], self.board
        for (nx, ny) in self._neighbors(position):
            if b[nx][ny] != EMPTY and not any(gs[nx][ny] is g for g in groups):
                groups.append(gs[nx][ny])
        return groups

    def is_suicide(self, action):
        return self.board[action] == EMPTY and action != self.ko and not self.is_legal(action)

    def is_legal(self, action):
        return bool(_lib.lib().ag_state_is_legal(self._h, self._idx(action)))

    def is_eyeish(self, position, owner):
        return bool(_lib.lib().ag_state_is_eyeish(self._h, self._idx(position), int(owner)))

    def is_eye(self, position, owner, stack=None):
        return bool(_lib.lib().ag_state_is_eye(self._h, self._idx(position), int(owner)))

    def get_legal_moves(self, include_eyes=True):
        buf = np.empty(self.size * self.size, np.int16)
        c = _lib.lib().ag_state_legal_moves(self._h, int(bool(include_eyes)), _lib.np_ptr(buf))
        return [divmod(int(m), self.size) for m in buf[:c]]

    def get_winner(self):
        return _lib.lib().ag_state_winner(self._h, float(self.komi))

    def copy(self):
        """independent copy (the reference aliases `history`, go.py:167; this one does not)"""
        return GameState(self.size, self.komi, _handle=_lib.lib().ag_state_copy(self._h))

    def do_move(self, action, color=None):
        """play at action=(x, y) or PASS_MOVE; raises IllegalMove (go.py:295-349)"""
        rc = _lib.lib().ag_state_do_move(self._h, self._idx(action), int(color or 0))
        if rc != 0:
            raise IllegalMove(str(action))
        return self.is_end_of_game

    # ------------------------------------------------------------ encoder input
    def pack(self):
        """(board int8[n], age uint8[n], player, ko) -- the encoder's view of this state"""
        n = self.size * self.size
        b = np.empty(n, np.int8); a = np.empty(n, np.uint8)
        pl = C.c_int8(); ko = C.c_int16()
        _lib.lib().ag_state_pack(self._h, _lib.np_ptr(b), _lib.np_ptr(a), C.byref(pl), C.byref(ko))
        return b, a, pl.value, ko.value


    def replay_pack(self, moves, players):
        """Play `moves` ((x, y) or PASS_MOVE) for `players` (BLACK / WHITE) on this state, packing the position in front of
        every non-pass move: one library call for a whole game record (ag_state_replay_pack).  Returns (boards int8 (k, n),
        ages uint8 (k, n), players int8 (k,), kos int16 (k,), illegal) with illegal = index into `moves` of the move that
        was refused (the positions up to and including its own are returned), or None."""
        m = len(moves)
        n = self.size * self.size
        pts = np.asarray([self._idx(mv) for mv in moves], np.int16).reshape(m)
        col = np.asarray(players, np.int8).reshape(m)
        b = np.empty((m, n), np.int8); a = np.empty((m, n), np.uint8)
        pl = np.empty(m, np.int8); ko = np.empty(m, np.int16)
        k = C.c_int(0)
        rc = _lib.lib().ag_state_replay_pack(self._h, m, _lib.np_ptr(pts), _lib.np_ptr(col), _lib.np_ptr(b), _lib.np_ptr(a),
                                             _lib.np_ptr(pl), _lib.np_ptr(ko), C.byref(k))
        illegal = None
        if rc != 0:
            msg = _lib.last_error()
            if "illegal move" not in msg:
                raise _lib.AlphaGoB200Error("ag_state_replay_pack failed (%d): %s" % (rc, msg))
            illegal = int(msg.rsplit("entry ", 1)[1].rstrip(")"))
        return b[:k.value], a[:k.value], pl[:k.value], ko[:k.value], illegal


def playout_positions(seed, n, size=19, max_moves=400, per_game=8):
    """Synthetic random-playout positions (uniform over sensible moves, AlphaGo/ai.py:15) as
    packed encoder inputs: board int8 (n, size*size), age uint8, player int8 (n,), ko int16 (n,)."""
    pts = size * size
    board = np.empty((n, pts), np.int8); age = np.empty((n, pts), np.uint8)
    player = np.empty(n, np.int8); ko = np.empty(n, np.int16)
    _lib.check(_lib.lib().ag_playout_positions(int(seed), n, size, max_moves, per_game, _lib.np_ptr(board),
                                               _lib.np_ptr(age), _lib.np_ptr(player), _lib.np_ptr(ko)),
               "ag_playout_positions")
    return board, age, player, ko
