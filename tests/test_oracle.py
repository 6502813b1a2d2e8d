"""Pin the CPU oracle (oracle/go_oracle.c) against the reference.

Sources of truth, in order: the committed golden fixtures generated from the unmodified
reference (oracle/gen_golden.py), the reference's own hand-coded known-answer tests
(tests/test_preprocessing.py, test_gamestate.py, test_liberties.py -- restated here on the
oracle's GameState mirror), and -- when /root/reference is present -- the live reference.
"""
import os

import numpy as np
import pytest

from oracle import go_oracle as og
from oracle import ref_shim, ladder_spec

FEATS46 = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
           "liberties_after", "sensibleness", "zeros"]
BLACK, WHITE = 1, -1


def test_oracle_builds_and_feature_sizes():
    og.build()
    assert og.output_dim(og.DEFAULT_FEATURES) == 48
    assert og.output_dim(FEATS46) == 46
    assert og.output_dim(["board", "ones", "turns_since"]) == 12


# ---------------------------------------------------------------- golden: rules
def test_rules_replay_reference_games(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_games.npz"))
    for gi in range(int(g["n_games"])):
        size = int(g["g%d_size" % gi])
        st = og.OState(size)
        samples = {int(t): i for i, t in enumerate(g["g%d_sample_at" % gi])}
        for t, m in enumerate(g["g%d_moves" % gi]):
            st.do_move(None if m < 0 else divmod(int(m), size))
            assert np.array_equal(st.board, g["g%d_boards" % gi][t]), (gi, t)
            ko = st.ko
            assert (-1 if ko is None else ko[0] * size + ko[1]) == int(g["g%d_ko" % gi][t]), (gi, t)
            assert st.current_player == int(g["g%d_player" % gi][t])
            if t in samples:
                i = samples[t]
                assert np.array_equal(st.liberty_counts, g["g%d_libs" % gi][i]), (gi, t)
                lm = np.zeros((size, size), np.uint8)
                for p in st.get_legal_moves():
                    lm[p] = 1
                sm = np.zeros((size, size), np.uint8)
                for p in st.get_legal_moves(include_eyes=False):
                    sm[p] = 1
                assert np.array_equal(lm, g["g%d_legal" % gi][i]), (gi, t)
                assert np.array_equal(sm, g["g%d_sensible" % gi][i]), (gi, t)


# ---------------------------------------------------------------- golden: encoder
@pytest.mark.parametrize("size", [19, 13, 9, 7])
def test_encoder_matches_reference_planes(golden_dir, size):
    g = np.load(os.path.join(golden_dir, "ref_planes.npz"))
    feats = [str(f) for f in g["features"]]
    out = og.encode_batch(g["board_%d" % size], g["age_%d" % size], g["player_%d" % size], g["ko_%d" % size],
                          size, feats)
    ref = g["planes_%d" % size]
    assert out.shape == ref.shape
    bad = np.argwhere(out != ref)
    assert bad.size == 0, "first mismatches (pos,plane,x,y): %s" % bad[:10]
    # the quirks must actually be exercised by the fixture (SURVEY 7 "hard parts")
    if size == 19:
        assert ref[:, 36:44].sum() > 0 and ref[:, 20:28].sum() > 0


def test_kat_boards_from_reference_tests(golden_dir):
    k = np.load(os.path.join(golden_dir, "ref_kat.npz"))
    st = og.OState(7)
    for m in k["simple_moves"]:
        st.do_move((int(m[0]), int(m[1])))
    assert st.ko == divmod(int(k["simple_ko"]), 7) == (4, 3)
    assert np.array_equal(st.encode(FEATS46)[0], k["simple_planes"])
    st = og.OState(7)
    for m in k["atari_black"]:
        st.do_move((int(m[0]), int(m[1])), BLACK)
    for m in k["atari_white"]:
        st.do_move((int(m[0]), int(m[1])), WHITE)
    assert st.current_player == int(k["atari_player"])
    assert np.array_equal(st.encode(FEATS46)[0], k["atari_planes"])


# ------------------------------------------- the reference's own unit tests, restated
def simple_board():  # tests/test_preprocessing.py:7-41
    gs = og.OState(7)
    for m in [(0, 0), (1, 0), (0, 1), (1, 1), (0, 2), (3, 4), (3, 3), (4, 5), (4, 2), (5, 4), (5, 3), (4, 3), (4, 4)]:
        gs.do_move(m)
    return gs


def self_atari_board():  # tests/test_preprocessing.py:44-75
    gs = og.OState(7)
    for m in [(2, 4), (4, 4), (6, 0)]:
        gs.do_move(m, BLACK)
    for m in [(1, 0), (5, 0), (2, 3), (4, 3), (1, 4), (5, 4), (2, 5), (3, 5), (4, 5)]:
        gs.do_move(m, WHITE)
    return gs


def _feat(gs, names):
    return gs.encode(names)[0].transpose((1, 2, 0))


def test_ref_get_board():  # tests/test_preprocessing.py:87-114
    gs = simple_board()
    f = _feat(gs, ["board"])
    white = np.zeros((7, 7), int); black = np.zeros((7, 7), int)
    for p in [(1, 0), (1, 1), (3, 4), (4, 5), (5, 4)]:
        white[p] = 1
    for p in [(0, 0), (0, 1), (0, 2), (3, 3), (4, 2), (4, 4), (5, 3)]:
        black[p] = 1
    assert np.all(f == np.dstack((white, black, 1 - white - black)))


def test_ref_turns_since():  # tests/test_preprocessing.py:116-132
    gs = simple_board()
    f = _feat(gs, ["turns_since"])
    exp = np.zeros((7, 7, 8))
    rev = gs.history[::-1]
    for x in range(7):
        for y in range(7):
            if gs.board[x, y] != 0:
                exp[x, y, min(rev.index((x, y)), 7)] = 1
    assert np.all(f == exp)


def test_ref_liberties():  # tests/test_preprocessing.py:134-165
    f = _feat(simple_board(), ["liberties"])
    e = np.zeros((7, 7, 8))
    e[4, 4, 0] = 1
    e[0, 0:3, 1] = 1; e[3, 4, 1] = 1; e[5, 4, 1] = 1
    e[1, 0:2, 2] = 1; e[4, 5, 2] = 1; e[3, 3, 2] = 1; e[5, 3, 2] = 1
    e[4, 2, 3] = 1
    assert np.all(f == e)


def test_ref_capture_size_and_sensibleness_and_concat():  # tests/test_preprocessing.py:167-181,224-258
    gs = simple_board()
    f = _feat(gs, ["capture_size"])
    e = np.zeros((7, 7, 8))
    for p in gs.get_legal_moves():
        e[p[0], p[1], 0] = 1
    assert np.all(f == e)
    s = gs.encode(["sensibleness"])[0, 0]
    es = np.zeros((7, 7))
    for p in gs.get_legal_moves():
        if not gs.is_eye(p, WHITE):
            es[p] = 1
    assert np.all(s == es)
    cat = _feat(gs, ["board", "sensibleness", "capture_size"])
    exp = np.zeros((7, 7, 12))
    exp[:, :, 0] = gs.board == WHITE; exp[:, :, 1] = gs.board == BLACK; exp[:, :, 2] = gs.board == 0
    exp[:, :, 3] = es; exp[:, :, 4] = e[:, :, 0]
    assert np.all(cat == exp)


def test_ref_self_atari_size():  # tests/test_preprocessing.py:183-194
    f = _feat(self_atari_board(), ["self_atari_size"])
    e = np.zeros((7, 7, 8)); e[0, 0, 0] = 1; e[3, 4, 2] = 1
    assert np.all(f == e)


def test_ref_liberties_after():  # tests/test_preprocessing.py:196-216
    gs = simple_board()
    f = _feat(gs, ["liberties_after"])
    e = np.zeros((7, 7, 8))
    for (x, y) in gs.get_legal_moves():
        c = gs.copy(); c.do_move((x, y))
        libs = c.liberty_counts[x, y]
        e[x, y, libs - 1 if libs < 7 else 7] = 1
    assert np.all(f == e)


def test_ref_ko():  # tests/test_gamestate.py:9-29
    gs = og.OState(9)
    for m in [(1, 0), (2, 0), (0, 1), (3, 1), (1, 2), (2, 2), (2, 1), (1, 1)]:
        gs.do_move(m)
    assert gs.prisoners == (1, 0)
    assert not gs.is_legal((2, 1))
    gs.do_move((5, 5)); gs.do_move((5, 6))
    assert gs.is_legal((2, 1))


def test_ref_snapback_is_not_ko():  # tests/test_gamestate.py:31-58
    gs = og.OState(5)
    for b, w in zip([(0, 0), (2, 1), (3, 0)], [(0, 1), (1, 1), (2, 0)]):
        gs.do_move(b); gs.do_move(w)
    gs.do_move((1, 0))
    assert gs.ko is None and gs.is_legal((2, 0))
    gs.do_move((2, 0))
    assert gs.prisoners == (2, 1)


def test_ref_eyes():  # tests/test_gamestate.py:63-118
    gs = og.OState(7)
    for m in [(1, 0), (5, 4), (2, 1), (6, 5), (1, 2), (5, 6), (0, 1), (4, 5)]:
        gs.do_move(m)
    assert gs.is_eyeish((1, 1), BLACK) and not gs.is_eyeish((1, 1), WHITE)
    assert gs.is_eyeish((5, 5), WHITE) and not gs.is_eyeish((5, 5), BLACK)
    assert not gs.is_eyeish((1, 0), BLACK) and not gs.is_eyeish((2, 2), WHITE)
    gs = og.OState(7)
    gs.do_move((1, 0), BLACK); gs.do_move((0, 1), BLACK)
    assert gs.is_eyeish((0, 0), BLACK) and not gs.is_eye((0, 0), BLACK)
    for m in [(1, 2), (2, 1), (2, 2), (0, 2)]:
        gs.do_move(m, BLACK)
    assert gs.is_eye((0, 0), BLACK) and gs.is_eye((1, 1), BLACK)
    gs = og.OState(7)  # checkerboard recursion
    for x in range(7):
        for y in range(7):
            if (x + y) % 2 == 1:
                gs.do_move((x, y), BLACK)
    assert gs.is_eye((0, 0), BLACK)


def test_ref_liberties_after_capture():  # tests/test_gamestate.py:120-147
    a, b = og.OState(7), og.OState(7)
    for x in range(2, 5):
        for y in range(2, 5):
            a.do_move((x, y), BLACK)
    for x in range(2, 5):
        for s in (a, b):
            s.do_move((x, 1), WHITE); s.do_move((x, 5), WHITE)
    for s in (a, b):
        s.do_move((1, 1), WHITE)
    for y in range(2, 5):
        for s in (a, b):
            s.do_move((1, y), WHITE); s.do_move((5, y), WHITE)
    assert np.all(a.board == b.board) and np.all(a.liberty_counts == b.liberty_counts)


def test_ref_liberty_counts_and_groups():  # tests/test_liberties.py:18-40
    s = og.OState(19)
    for m in [(4, 5), (5, 5), (5, 6), (10, 10), (4, 6), (10, 11), (6, 6), (9, 10)]:
        s.do_move(m)
    lc = s.liberty_counts
    assert lc[5][5] == 2 and lc[4][5] == 8 and lc[5][6] == 8
    st = og.OState(19)
    for m in [(0, 0), (5, 5), (0, 1), (6, 6), (1, 0), (1, 1)]:
        st.do_move(m)
    gsz = st.group_sizes
    assert gsz[0, 0] == 3 and gsz[4, 4] == 0 and gsz[5, 5] == 1


def test_double_single_capture_leaves_stale_ko():
    """go.py:322-331 sets ko inside the neighbour loop; verified against the live reference
    by test_live_reference_random_playouts when present."""
    st = og.OState(9)
    # white stones at (3,4) and (5,4) each with one liberty at (4,4); black surrounds
    for m in [(2, 4), (3, 3), (3, 5), (6, 4), (5, 3), (5, 5)]:
        st.do_move(m, BLACK)
    st.do_move((3, 4), WHITE); st.do_move((5, 4), WHITE)
    st.do_move((4, 3), BLACK); st.do_move((4, 5), BLACK)
    st.current_player = BLACK
    st.do_move((4, 4))  # captures both single stones
    assert st.board[3, 4] == 0 and st.board[5, 4] == 0
    assert st.ko is None or st.ko in [(3, 4), (5, 4)]


# ---------------------------------------------------------------- ladders (build-owned spec)
def test_ladder_c_matches_python_spec_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "ladder_spec.npz"))
    for i in range(len(g["size"])):
        size = int(g["size"][i]); n = size * size
        out = og.encode_batch(g["board"][i][:n], np.zeros(n, np.uint8), [g["player"][i]], [g["ko"][i]], size,
                              ["ladder_capture", "ladder_escape"])
        assert np.array_equal(out[0, 0].reshape(-1), g["capture"][i][:n]), i
        assert np.array_equal(out[0, 1].reshape(-1), g["escape"][i][:n]), i


def _ladder_board(breaker, to_move, chased=False):
    """White stone at (2,2) with black on (2,1),(1,2),(3,1): black (2,3) starts a ladder that
    runs down the diagonal (2,2),(3,2),(3,3),(4,3)... to the edge.  A white `breaker` stone on
    that diagonal rescues it; one off the path does not."""
    st = og.OState(9)
    for m in [(2, 1), (1, 2), (3, 1)] + ([(2, 3)] if chased else []):
        st.do_move(m, BLACK)
    st.do_move((2, 2), WHITE)
    if breaker:
        st.do_move(breaker, WHITE)
    st.current_player = to_move
    return st


def test_ladder_known_answers():
    for breaker, works in [(None, True), ((6, 6), False), ((7, 6), False), ((0, 8), True)]:
        st = _ladder_board(breaker, BLACK)
        assert st.is_ladder_capture((2, 3)) is works, breaker
        assert ladder_spec.is_ladder_capture(st, (2, 3)) is works, breaker
        # chasing from the wrong side lets the prey out at once
        assert st.is_ladder_capture((3, 2)) is False
        assert ladder_spec.is_ladder_capture(st, (3, 2)) is False
    # escape plane: white to move with (2,2) in atari; extending only helps with a breaker ahead
    for breaker, escapes in [(None, False), ((6, 6), True)]:
        st = _ladder_board(breaker, WHITE, chased=True)
        assert st.is_ladder_escape((3, 2)) is escapes
        assert ladder_spec.is_ladder_escape(st, (3, 2)) is escapes
    # the planes are zero off the legal moves and on moves that touch no prey
    st = _ladder_board(None, BLACK)
    planes = st.encode(["ladder_capture", "ladder_escape"])[0]
    assert planes[0].sum() == 1 and planes[0][2, 3] == 1 and planes[1].sum() == 0


def test_ladder_c_matches_python_spec_random():
    b, a, pl, ko = og.playout_positions(5, 24, size=9, max_moves=70, per_game=4)
    for i in range(len(pl)):
        out = og.encode_batch(b[i], a[i], [pl[i]], [ko[i]], 9, ["ladder_capture", "ladder_escape"])[0]
        gs = _PackedState(b[i].reshape(9, 9), int(pl[i]), int(ko[i]))
        c, e = ladder_spec.ladder_planes(gs)
        assert np.array_equal(out[0], c), i
        assert np.array_equal(out[1], e), i


class _PackedState(object):
    """minimal python GameState over a packed position, rules restated with flood fill;
    only used to drive ladder_spec.py independently of the C code."""

    def __init__(self, board, player, ko):
        self.board = np.array(board, dtype=np.int8); self.size = board.shape[0]
        self.current_player = player
        self.ko = None if ko < 0 else divmod(ko, self.size)
        self.history = []

    def copy(self):
        k = -1 if self.ko is None else self.ko[0] * self.size + self.ko[1]
        return _PackedState(self.board.copy(), self.current_player, k)

    def _grp(self, p):
        return ladder_spec._group(self.board, p, self.size)

    def is_legal(self, a):
        x, y = a
        if self.board[x, y] != 0 or a == self.ko:
            return False
        nb = ladder_spec._nbrs(a, self.size)
        if any(self.board[q] == 0 for q in nb):
            return True
        for q in nb:
            _, libs = self._grp(q)
            other = len(libs - {a}) > 0
            if self.board[q] == self.current_player and other:
                return True
            if self.board[q] == -self.current_player and not other:
                return True
        return False

    def get_legal_moves(self):
        return [(x, y) for x in range(self.size) for y in range(self.size) if self.is_legal((x, y))]

    def do_move(self, a):
        c = self.current_player
        self.ko = None
        self.board[a] = c
        for q in ladder_spec._nbrs(a, self.size):
            if self.board[q] == -c:
                stones, libs = self._grp(q)
                if not libs:
                    for s in stones:
                        self.board[s] = 0
                    if len(stones) == 1:
                        ms, ml = self._grp(a)
                        if len(ml) == 1 and len(ms) == 1:
                            self.ko = q
        self.current_player = -c


# ---------------------------------------------------------------- live reference (build container only)
@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present")
def test_live_reference_random_playouts():
    go, pp = ref_shim.load()
    prep = pp.Preprocess(FEATS46)
    rng = np.random.RandomState(123)
    for size, nm in [(19, 200), (9, 120)]:
        gs = go.GameState(size); st = og.OState(size)
        for t in range(nm):
            mv = gs.get_legal_moves(include_eyes=False)
            assert mv == st.get_legal_moves(include_eyes=False)
            m = mv[rng.randint(len(mv))] if mv else None
            gs.do_move(m); st.do_move(m)
            assert gs.ko == st.ko
            if t % 10 == 0:
                ref = prep.state_to_tensor(gs)[0].astype(np.uint8)
                assert np.array_equal(st.encode(FEATS46)[0], ref), (size, t)
                b, a, pl, ko = og.pack_gamestate(gs)
                assert np.array_equal(og.encode_batch(b, a, [pl], [ko], size, FEATS46)[0], ref)


# ---------------------------------------------------------------- the reference's own HDF5 golden fixture
def load_hdf5_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_hdf5_fixture.npz"))
    shape = tuple(int(v) for v in g["shape"])
    states = np.unpackbits(g["states_bits"], axis=1)[:, :int(np.prod(shape[1:]))].reshape(shape)
    return states, g["actions"], g["offsets"], [str(f) for f in g["features"]]


def test_reference_hdf5_fixture_real_games(golden_dir):
    """tests/test_data/hdf5/alphago-vs-lee-sedol-features.hdf5 (written by the reference's own converter from
    the five AlphaGo - Lee Sedol games): replaying its actions must reproduce its 1033 x 12 planes exactly"""
    states, actions, offsets, feats = load_hdf5_fixture(golden_dir)
    assert states.shape == (1033, 12, 19, 19) and feats == ["board", "ones", "turns_since"]
    for start, count in offsets:
        st = og.OState(19)
        for i in range(start, start + count):
            assert np.array_equal(st.encode(feats)[0], states[i]), i
            st.do_move((int(actions[i][0]), int(actions[i][1])))


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present")
def test_hdf5_lite_reads_reference_file(golden_dir):
    from alphago_b200 import hdf5_lite
    f = hdf5_lite.File(os.path.join(ref_shim.REFERENCE_ROOT, "tests", "test_data", "hdf5", "alphago-vs-lee-sedol-features.hdf5"))
    assert f.keys() == ["actions", "file_offsets", "states"]
    s, a = f["states"], f["actions"]
    assert s.shape == (1033, 12, 19, 19) and s.dtype == np.uint8 and s.chunk == (64, 12, 19, 19) and s.filters == [32000]
    states, actions, offsets, _ = load_hdf5_fixture(golden_dir)
    assert np.array_equal(s[:], states) and np.array_equal(a[:], actions)
    assert np.array_equal(s[100], states[100]) and np.array_equal(s[60:70], states[60:70]) and len(s) == 1033
    assert np.array_equal(s[[5, 900, 64]], states[[5, 900, 64]])
    names = f["file_offsets"].keys()
    assert len(names) == 5 and sorted(f["file_offsets"][n][:].tolist() for n in names) == sorted(offsets.tolist())
