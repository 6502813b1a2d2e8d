"""CPU tests of the SGF reader and the converter's host half (alphago_b200/util.py, game_converter.pack_game), after the
reference's tests/test_game_converter.py.  Golden data: the reference's own converter output for the five AlphaGo -
Lee Sedol games (tests/golden/ref_hdf5_fixture.npz, from its alphago-vs-lee-sedol-features.hdf5) and the game records
themselves (tests/golden/sgf, copied by oracle/gen_golden.py sgf).  The planes are produced by the ORACLE here (no GPU):
what is under test is parsing, replay and packing; tests/test_converter_gpu.py runs the whole converter."""
import os

import numpy as np
import pytest

from oracle import go_oracle as og
from test_oracle import load_hdf5_fixture

from alphago_b200 import go, util


def _sgf(golden_dir, name):
    return os.path.join(golden_dir, "sgf", name)


def _fixture_games(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_hdf5_fixture.npz"))
    return [(str(n).split("::")[-1], int(o[0]), int(o[1])) for n, o in zip(g["names"], g["offsets"])]


def test_flatten_idx_roundtrip():
    for size in (7, 19):
        for idx in range(size * size):
            assert util.flatten_idx(util.unflatten_idx(idx, size), size) == idx
    assert util.flatten_idx((3, 4), 19) == 3 * 19 + 4   # util.py:6-8: x * size + y


def test_parse_sgf_move():
    assert util._parse_sgf_move("") is go.PASS_MOVE and util._parse_sgf_move("tt") is go.PASS_MOVE
    assert util._parse_sgf_move("ab") == (0, 1) and util._parse_sgf_move("ss") == (18, 18)


def test_main_line_ignores_variations_and_unescapes():
    text = "(;SZ[9]C[a \\] bracket and \\\\ backslash];B[aa](;W[bb];B[cc])(;W[dd]))(;SZ[19];B[ee])"
    nodes = util.sgf_main_line(text)
    assert [sorted(n) for n in nodes] == [["C", "SZ"], ["B"], ["W"], ["B"]]
    assert nodes[0]["C"] == ["a ] bracket and \\ backslash"]
    assert nodes[2]["W"] == ["bb"] and nodes[3]["B"] == ["cc"]
    with pytest.raises(util.ParseException):
        util.sgf_main_line("no game here")
    with pytest.raises(util.ParseException):
        util.sgf_main_line("(;B[aa")


def test_ab_aw(golden_dir):
    """tests/test_game_converter.py:8-10: set-up stones go through do_move (util.py:38-44), so they are in the history"""
    with open(_sgf(golden_dir, "ab_aw.sgf")) as f:
        gs = util.sgf_to_gamestate(f.read())
    assert gs.size == 19
    assert gs.history[:4] == [(3, 3), (15, 15), (15, 3), (3, 15)]   # AB[dd][pp] AW[pd][dp]
    with open(_sgf(golden_dir, "ab_aw.sgf")) as f:
        first, move, player = next(util.sgf_iter_states(f.read()))
    assert first.board[3][3] == go.BLACK and first.board[15][15] == go.BLACK
    assert first.board[15][3] == go.WHITE and first.board[3][15] == go.WHITE
    assert len(first.history) == 4 and (move, player) == ((5, 2), go.WHITE)   # ;W[fc] is the first move


def test_iter_states_yields_one_object_advanced_in_place(golden_dir):
    name = _fixture_games(golden_dir)[0][0]
    with open(_sgf(golden_dir, name)) as f:
        text = f.read()
    seen, turns = set(), []
    for gs, move, player in util.sgf_iter_states(text):
        seen.add(id(gs))
        turns.append(len(gs.history))
        assert player in (go.BLACK, go.WHITE)
    assert len(seen) == 1 and turns == list(range(len(turns)))


def test_moves_and_planes_match_reference_converter_output(golden_dir):
    """replaying the SGFs gives exactly the reference converter's actions, and the packed positions encode (oracle) to
    exactly its 1033 x 12 planes"""
    from alphago_b200.game_converter import game_converter
    states, actions, offsets, feats = load_hdf5_fixture(golden_dir)
    conv = game_converter.__new__(game_converter)   # host half only: no encoder handle needed
    for name, start, count in _fixture_games(golden_dir):
        packed, acts, error = conv.pack_game(_sgf(golden_dir, name), 19)
        assert error is None and len(packed) == count
        assert np.array_equal(np.asarray(acts, np.uint8), actions[start:start + count])
        board = np.stack([p[0] for p in packed]); age = np.stack([p[1] for p in packed])
        player = np.asarray([p[2] for p in packed], np.int8); ko = np.asarray([p[3] for p in packed], np.int16)
        planes = og.encode_batch(board, age, player, ko, 19, feats)
        assert np.array_equal(planes, states[start:start + count]), name


def test_pack_game_error_handling(golden_dir, tmp_path):
    from alphago_b200.game_converter import SizeMismatchError, game_converter
    conv = game_converter.__new__(game_converter)
    # wrong board size: nothing extracted (game_converter.py:35-36)
    packed, acts, error = conv.pack_game(_sgf(golden_dir, "ab_aw.sgf"), 9)
    assert packed == [] and isinstance(error, SizeMismatchError)
    # an illegal move (playing on a stone) keeps the pairs before it (game_converter.py:111-112)
    bad = tmp_path / "illegal.sgf"
    bad.write_text("(;SZ[19];B[aa];W[bb];B[bb];W[cc])")
    packed, acts, error = conv.pack_game(str(bad), 19)
    assert isinstance(error, go.IllegalMove) and acts == [(0, 0), (1, 1), (1, 1)][:len(acts)] and len(packed) == 3
    # passes are skipped, both spellings
    p = tmp_path / "pass.sgf"
    p.write_text("(;SZ[19];B[aa];W[];B[tt];W[cc])")
    packed, acts, error = conv.pack_game(str(p), 19)
    assert error is None and acts == [(0, 0), (2, 2)]
    # not SGF at all
    q = tmp_path / "junk.sgf"
    q.write_text("hello")
    packed, acts, error = conv.pack_game(str(q), 19)
    assert packed == [] and isinstance(error, util.ParseException)
    # a malformed move value part-way through the record: the pairs in front of it stand, the error is reported (the
    # reference's generator has yielded them before the exception surfaces, game_converter.py:103-123)
    m = tmp_path / "malformed.sgf"
    m.write_text("(;SZ[19];B[aa];W[bb];B[c];W[dd])")
    packed, acts, error = conv.pack_game(str(m), 19)
    assert acts == [(0, 0), (1, 1)] and len(packed) == 2 and isinstance(error, (IndexError, ValueError))
    # a first node that carries no move: nothing to repeat yet (util.py:81 fails the same way)
    n = tmp_path / "nomove.sgf"
    n.write_text("(;SZ[19];C[hello];B[aa])")
    packed, acts, error = conv.pack_game(str(n), 19)
    assert packed == [] and acts == [] and isinstance(error, UnboundLocalError)
