"""CPU tests of the product's host side: the C-ABI library loads and exports every symbol the
header declares, the host rules engine (alphago_b200.go.GameState) replays the reference's games
move for move (golden fixtures), and Preprocess keeps the reference's constructor contract."""
import ctypes
import os

import numpy as np
import pytest

from alphago_b200 import _lib, go
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES, pack_state
from oracle import go_oracle as og

BLACK, WHITE = go.BLACK, go.WHITE


def test_library_exports_every_header_symbol():
    L = ctypes.CDLL(_lib.LIB_PATH)
    names = _lib.header_symbols()
    assert "ag_encode" in names and "ag_state_do_move" in names
    for name in names:
        assert hasattr(L, name), "missing export: " + name
    assert _lib.lib().ag_version() >= 1
    assert [_lib.lib().ag_feature_planes(i) for i in range(12)] == [3, 1, 8, 8, 8, 8, 8, 1, 1, 1, 1, 1]
    assert _lib.lib().ag_feature_planes(99) == -1


def test_gamestate_replays_reference_games(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_games.npz"))
    for gi in range(int(g["n_games"])):
        size = int(g["g%d_size" % gi])
        st = go.GameState(size)
        samples = {int(t): i for i, t in enumerate(g["g%d_sample_at" % gi])}
        for t, m in enumerate(g["g%d_moves" % gi]):
            st.do_move(None if m < 0 else divmod(int(m), size))
            assert np.array_equal(st.board, g["g%d_boards" % gi][t]), (gi, t)
            assert (-1 if st.ko is None else st.ko[0] * size + st.ko[1]) == int(g["g%d_ko" % gi][t])
            assert st.current_player == int(g["g%d_player" % gi][t])
            if t in samples:
                i = samples[t]
                assert np.array_equal(st.liberty_counts, g["g%d_libs" % gi][i])
                lm = np.zeros((size, size), np.uint8)
                for p in st.get_legal_moves():
                    lm[p] = 1
                sm = np.zeros((size, size), np.uint8)
                for p in st.get_legal_moves(include_eyes=False):
                    sm[p] = 1
                assert np.array_equal(lm, g["g%d_legal" % gi][i]), (gi, t)
                assert np.array_equal(sm, g["g%d_sensible" % gi][i]), (gi, t)
        assert len(st.history) == len(g["g%d_moves" % gi])


def test_gamestate_reference_unit_tests():
    # tests/test_gamestate.py:9-58 (ko, snapback), :63-118 (eyes), tests/test_liberties.py:18-40
    gs = go.GameState(size=9)
    for m in [(1, 0), (2, 0), (0, 1), (3, 1), (1, 2), (2, 2), (2, 1), (1, 1)]:
        gs.do_move(m)
    assert gs.num_black_prisoners == 1 and gs.num_white_prisoners == 0
    assert not gs.is_legal((2, 1))
    gs.do_move((5, 5)); gs.do_move((5, 6))
    assert gs.is_legal((2, 1))
    gs = go.GameState(size=5)
    for b, w in zip([(0, 0), (2, 1), (3, 0)], [(0, 1), (1, 1), (2, 0)]):
        gs.do_move(b); gs.do_move(w)
    gs.do_move((1, 0))
    assert gs.ko is None and gs.is_legal((2, 0))
    gs.do_move((2, 0))
    assert gs.num_black_prisoners == 2 and gs.num_white_prisoners == 1
    gs = go.GameState(7)
    gs.do_move((1, 0), BLACK); gs.do_move((0, 1), BLACK)
    assert gs.is_eyeish((0, 0), BLACK) and not gs.is_eye((0, 0), BLACK)
    for m in [(1, 2), (2, 1), (2, 2), (0, 2)]:
        gs.do_move(m, BLACK)
    assert gs.is_eye((0, 0), BLACK) and gs.is_eye((1, 1), BLACK)
    gs = go.GameState(7)
    for x in range(7):
        for y in range(7):
            if (x + y) % 2 == 1:
                gs.do_move((x, y), BLACK)
    assert gs.is_eye((0, 0), BLACK)
    st = go.GameState()
    for m in [(0, 0), (5, 5), (0, 1), (6, 6), (1, 0), (1, 1)]:
        st.do_move(m)
    assert len(st.get_group((0, 0))) == 3 and len(st.get_group((4, 4))) == 0 and len(st.get_group((5, 5))) == 1
    assert st.liberty_counts[5][5] == 4 and st.liberty_counts[4][4] == -1
    with pytest.raises(go.IllegalMove):
        st.do_move((0, 0))
    assert not st.is_legal((-1, 3)) and not st.is_legal((19, 0))


def test_gamestate_views_and_copy():
    st = go.GameState(9)
    for m in [(2, 2), (2, 3), (3, 2), (6, 6), None]:
        st.do_move(m)
    c = st.copy()
    c.do_move((0, 0))
    assert len(c.history) == len(st.history) + 1 and st.board[0, 0] == 0  # history not aliased
    assert st.history[-1] is None and st.passes_black == 1
    ls, gs = st.liberty_sets, st.group_sets
    assert gs[2][2] == {(2, 2), (3, 2)} and gs[2][2] is gs[3][2]
    assert ls[2][2] == {(1, 2), (4, 2), (3, 1), (2, 1), (3, 3)}
    assert ls[0][0] == {(1, 0), (0, 1)} and gs[0][0] == set()
    assert len(st.get_groups_around((2, 4))) == 1 and st.get_winner() in (-1, 0, 1)
    b, a, pl, ko = st.pack()
    assert (b, a, pl, ko)[2:] == (st.current_player, -1)
    b2, a2, pl2, ko2 = og.pack_gamestate(st)  # independent packer
    assert np.array_equal(b, b2) and np.array_equal(a, a2)


def test_pack_state_duck_typed_with_aliased_history():
    class Dummy(object):
        pass
    d = Dummy()
    d.size = 5
    d.board = np.zeros((5, 5)); d.board[1, 1] = 1; d.board[2, 2] = -1
    d.history = [(1, 1), (2, 2), None, (4, 4)]  # (4,4) was never played on this board (go.py:167)
    d.current_player = 1; d.ko = (0, 3)
    b, a, pl, ko = pack_state(d)
    assert ko == 3 and pl == 1
    assert a[6] == 3 and a[12] == 2 and a[24] == 255


def test_preprocess_constructor_contract():
    p = Preprocess()
    assert p.output_dim == 48 and p.feature_list == DEFAULT_FEATURES
    assert Preprocess(["board", "ones", "turns_since"]).output_dim == 12
    assert Preprocess(["Board", "SENSIBLENESS"]).output_dim == 4  # names are lower-cased (preprocessing.py:231)
    with pytest.raises(ValueError):
        Preprocess(["board", "nonsense"])


def test_playout_positions_are_consistent():
    b, a, pl, ko = go.playout_positions(3, 300, size=19, max_moves=400, per_game=8)
    assert b.shape == (300, 361) and set(np.unique(b)) <= {-1, 0, 1} and set(np.unique(pl)) <= {-1, 1}
    assert ((a < 8) | (a == 255)).all() and (a[b == 0] == 255).all() and (a[b != 0] < 8).all()
    assert (ko >= -1).all() and all(b[i, k] == 0 for i, k in enumerate(ko) if k >= 0)
    b2, a2, pl2, ko2 = go.playout_positions(3, 300, size=19, max_moves=400, per_game=8)
    assert np.array_equal(b, b2) and np.array_equal(ko, ko2)  # seeded
    # every group on every board has a liberty (positions are legal)
    planes = og.encode_batch(b, a, pl, ko, 19, ["board", "liberties"])
    stones = planes[:, 0] + planes[:, 1]
    assert np.array_equal(planes[:, 3:].sum(1), stones)


def test_entry_points_refuse_bad_arguments_before_touching_the_device():
    """error behaviour of the C ABI: a negative status (or NULL) and a message, no CUDA call -- checkable without a GPU"""
    L = _lib.lib()
    i32 = np.asarray([0, 1, 2], np.int32)
    p = _lib.np_ptr(i32)
    nul = ctypes.c_void_p(0)
    assert L.ag_encode(nul, nul, nul, nul, 4, 20, p, 3, nul, nul) == -1 and "size" in _lib.last_error()          # AG_EINVAL
    assert L.ag_encode(nul, nul, nul, nul, 4, 19, p, 0, nul, nul) == -1
    bad = np.asarray([0, 99], np.int32)
    assert L.ag_encode(nul, nul, nul, nul, 4, 19, _lib.np_ptr(bad), 2, nul, nul) == -1 and "feature" in _lib.last_error()
    assert L.ag_encode(nul, nul, nul, nul, 4, 19, p, 3, nul, nul) == -1 and "null" in _lib.last_error()
    assert L.ag_encode(nul, nul, nul, nul, 0, 19, p, 3, nul, nul) == 0                                              # empty batch: nothing to do
    f = np.asarray([64, 64], np.int32)
    for widths, what in (([5, 4], "odd"), ([7, 3], "odd")):
        w = np.asarray(widths, np.int32)
        assert not L.ag_net_create(19, 48, 2, _lib.np_ptr(f), _lib.np_ptr(w), 0, 8) and what in _lib.last_error()
    w = np.asarray([5, 3], np.int32)
    g = np.asarray([64, 32], np.int32)
    assert not L.ag_net_create(19, 48, 2, _lib.np_ptr(g), _lib.np_ptr(w), 0, 8) and "same filter count" in _lib.last_error()
    assert L.ag_selfplay_step(nul, nul, nul, nul, nul, nul, nul, nul, 361, nul, 1.0, 4, 25, nul, nul) == -1
    assert L.ag_selfplay_step(nul, nul, nul, nul, nul, nul, nul, nul, 10, nul, 1.0, 4, 19, nul, nul) == -1           # mask rows shorter than a board
    assert L.ag_selfplay_step(nul, nul, nul, nul, nul, nul, nul, nul, 361, nul, 1.0, 4, 19, nul, nul) == -1 and "null" in _lib.last_error()
    assert L.ag_selfplay_winner(nul, nul, 0, 19, 7.5, nul, nul) == 0
    rec = (ctypes.c_ubyte * 144)()
    assert not L.ag_xchg_create(0, 1, nul, 1024, 16, rec) and "world" in _lib.last_error()                          # a lone rank has nothing to exchange
    assert not L.ag_xchg_create(3, 2, nul, 1024, 16, rec)
    assert L.ag_state_new(1) is None and L.ag_state_new(20) is None
    assert L.ag_wgrad_partial_bytes(3, 192, 192) > 9 * 192 * 192 * 4 - 1 and L.ag_wgrad_scratch_bytes(3, 100, 192, 1000) == 0
