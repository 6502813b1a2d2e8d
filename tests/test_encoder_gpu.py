"""GPU parity tests of the CUDA feature encoder, through the C ABI (ag_encode), against
(i) golden planes from the unmodified reference and (ii) the CPU oracle on seeded random
playouts.  Bit-exact: the planes are 0/1 bytes."""
import os

import numpy as np
import pytest

from alphago_b200 import go
from alphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES
from oracle import go_oracle as og

pytestmark = pytest.mark.gpu

FEATS46 = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
           "liberties_after", "sensibleness", "zeros"]


def gpu_encode(feats, board, age, player, ko, size):
    import torch
    pp = Preprocess(feats)
    d = "cuda"
    out = pp.encode_packed(torch.from_numpy(np.ascontiguousarray(board, np.int8).reshape(len(player), -1)).to(d),
                           torch.from_numpy(np.ascontiguousarray(age, np.uint8).reshape(len(player), -1)).to(d),
                           torch.from_numpy(np.ascontiguousarray(player, np.int8)).to(d),
                           torch.from_numpy(np.ascontiguousarray(ko, np.int16)).to(d), size)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def assert_same(got, want, feats):
    assert got.shape == want.shape
    if not np.array_equal(got, want):
        bad = np.argwhere(got != want)
        names = []
        for f in feats:
            names += [f] * og.output_dim([f])
        per_plane = {}
        for b in bad:
            per_plane[names[b[1]] + "#%d" % b[1]] = per_plane.get(names[b[1]] + "#%d" % b[1], 0) + 1
        raise AssertionError("%d mismatching bytes in %d positions; per plane %s; first %s" % (
            len(bad), len(set(bad[:, 0])), per_plane, bad[:5].tolist()))


@pytest.mark.parametrize("size", [19, 13, 9, 7])
def test_matches_reference_golden_planes(golden_dir, size):
    g = np.load(os.path.join(golden_dir, "ref_planes.npz"))
    feats = [str(f) for f in g["features"]]
    got = gpu_encode(feats, g["board_%d" % size], g["age_%d" % size], g["player_%d" % size], g["ko_%d" % size], size)
    assert_same(got, g["planes_%d" % size], feats)


def test_reference_kat_boards_via_state_to_tensor(golden_dir):
    k = np.load(os.path.join(golden_dir, "ref_kat.npz"))
    st = go.GameState(7)
    for m in k["simple_moves"]:
        st.do_move((int(m[0]), int(m[1])))
    t = Preprocess(FEATS46).state_to_tensor(st)
    assert t.dtype == np.float64 and t.shape == (1, 46, 7, 7)
    assert np.array_equal(t[0], k["simple_planes"])
    st = go.GameState(7)
    for m in k["atari_black"]:
        st.do_move((int(m[0]), int(m[1])), go.BLACK)
    for m in k["atari_white"]:
        st.do_move((int(m[0]), int(m[1])), go.WHITE)
    assert np.array_equal(Preprocess(FEATS46).state_to_tensor(st)[0], k["atari_planes"])


@pytest.mark.parametrize("n,size,seed", [(4096, 19, 11), (1000, 13, 12), (1001, 9, 13), (333, 7, 14), (64, 5, 15)])
def test_matches_oracle_on_random_playouts(n, size, seed):
    b, a, pl, ko = go.playout_positions(seed, n, size=size, max_moves=size * size + 40, per_game=8)
    want = og.encode_batch(b, a, pl, ko, size, FEATS46, threads=os.cpu_count() or 1)
    assert_same(gpu_encode(FEATS46, b, a, pl, ko, size), want, FEATS46)


def test_feature_subsets_orders_and_repeats():
    b, a, pl, ko = go.playout_positions(21, 500, size=19)
    for feats in (["board"], ["sensibleness"], ["zeros", "ones", "color"], ["liberties_after", "board", "liberties_after"],
                  ["board", "ones", "turns_since"], ["capture_size", "self_atari_size"],
                  ["turns_since", "sensibleness", "capture_size", "board", "color"]):
        want = og.encode_batch(b, a, pl, ko, 19, feats, threads=os.cpu_count() or 1)
        assert_same(gpu_encode(feats, b, a, pl, ko, 19), want, feats)


def test_full_size_batch_properties_and_oracle():
    """BASELINE configs[1] upper size: 16384 positions, ALL 48 planes (ladder planes included); one-hot invariants
    plus the oracle on the whole batch (the C oracle does this in seconds with all host cores)."""
    n = 16384
    b, a, pl, ko = go.playout_positions(99, n, size=19, per_game=16)
    got = gpu_encode(DEFAULT_FEATURES, b, a, pl, ko, 19)
    assert got.shape == (n, 48, 19, 19)
    stones = (b != 0).reshape(n, 19, 19)
    assert np.array_equal(got[:, 0:3].sum(1), np.ones((n, 19, 19), np.uint8))          # board is a partition
    assert (got[:, 3] == 1).all() and (got[:, 47] == 0).all()                          # ones / zeros
    assert np.array_equal(got[:, 4:12].sum(1), stones.astype(np.uint8))                # one age per stone
    assert np.array_equal(got[:, 12:20].sum(1), stones.astype(np.uint8))               # one liberty class per stone
    legal = got[:, 20:28].sum(1)
    assert legal.max() <= 1 and np.array_equal(got[:, 36:44].sum(1), legal)            # capture/libs_after on legal moves
    assert (got[:, 28:36].sum(1) <= legal).all() and (got[:, 46] <= legal).all()       # self-atari, sensible subset of legal
    assert (got[:, 44] <= legal).all() and (got[:, 45] <= legal).all()                 # ladder moves are legal moves
    want = og.encode_batch(b, a, pl, ko, 19, DEFAULT_FEATURES, threads=os.cpu_count() or 1)
    assert_same(got, want, DEFAULT_FEATURES)
    assert want[:, 44].sum() > 1000 and want[:, 45].sum() > 1000


def test_edge_cases():
    import torch
    feats = FEATS46
    # empty batch
    z = Preprocess(feats).encode_packed(torch.empty((0, 361), dtype=torch.int8, device="cuda"),
                                        torch.empty((0, 361), dtype=torch.uint8, device="cuda"),
                                        torch.empty((0,), dtype=torch.int8, device="cuda"),
                                        torch.empty((0,), dtype=torch.int16, device="cuda"), 19)
    assert z.shape == (0, 46, 19, 19)
    # empty board, single positions, both colours to move
    for size in (19, 9, 2):
        n = size * size
        for player in (1, -1):
            b = np.zeros((1, n), np.int8); a = np.full((1, n), 255, np.uint8)
            assert_same(gpu_encode(feats, b, a, [player], [-1], size), og.encode_batch(b, a, [player], [-1], size, feats), feats)
    # checkerboards: every empty point is an eye candidate -> deep is_eye recursion (tests/test_gamestate.py:110-118)
    for size in (7,):  # (9x9 already takes the reference-faithful recursion minutes: it is exponential)
        n = size * size
        b = np.zeros((2, n), np.int8)
        for p in range(n):
            if (p // size + p % size) % 2 == 1:
                b[0, p] = 1; b[1, p] = -1
        a = np.where(b != 0, 7, 255).astype(np.uint8)
        for player in (1, -1):
            want = og.encode_batch(b, a, [player, player], [-1, -1], size, feats)
            assert_same(gpu_encode(feats, b, a, [player, player], [-1, -1], size), want, feats)
    # one huge group with > 8 liberties next to small groups; a ko point; stale ages
    st = go.GameState(19)
    for x in range(19):
        st.do_move((x, 9), go.BLACK)
    for m in [(0, 0), (1, 0), (0, 1)]:
        st.do_move(m, go.WHITE)
    st.do_move((5, 5), go.BLACK); st.do_move(None); st.do_move(None, go.BLACK)
    b, a, pl, ko = st.pack()
    for k in (-1, 200):
        assert_same(gpu_encode(feats, b[None], a[None], [pl], [k], 19), og.encode_batch(b[None], a[None], [pl], [k], 19, feats), feats)
    # unsupported requests fail loudly
    with pytest.raises(Exception):
        gpu_encode(feats, np.zeros((1, 400), np.int8), np.zeros((1, 400), np.uint8), [1], [-1], 20)


def test_host_buffer_path_matches_device_path():
    b, a, pl, ko = go.playout_positions(5, 777, size=19)
    pp = Preprocess(FEATS46)
    host = pp.encode_host(b, a, pl, ko, 19)
    assert host.dtype == np.uint8 and host.shape == (777, 46, 19, 19)
    assert_same(host, og.encode_batch(b, a, pl, ko, 19, FEATS46, threads=os.cpu_count() or 1), FEATS46)


# ---------------------------------------------------------------- ladder planes (build-owned spec)
LADDER = ["ladder_capture", "ladder_escape"]


def test_ladder_planes_match_python_spec_fixture(golden_dir):
    """planes produced by oracle/ladder_spec.py on the reference's own GameState (tests/golden/ladder_spec.npz)"""
    g = np.load(os.path.join(golden_dir, "ladder_spec.npz"))
    for size in (19, 9, 7):
        sel = np.flatnonzero(g["size"] == size)
        n = size * size
        b = g["board"][sel][:, :n]
        a = np.where(b != 0, 7, 255).astype(np.uint8)
        got = gpu_encode(LADDER, b, a, g["player"][sel], g["ko"][sel], size)
        assert np.array_equal(got[:, 0].reshape(len(sel), n), g["capture"][sel][:, :n]), size
        assert np.array_equal(got[:, 1].reshape(len(sel), n), g["escape"][sel][:, :n]), size


def test_ladder_known_answers():
    def board(breaker, to_move, chased=False):
        st = go.GameState(9)
        for m in [(2, 1), (1, 2), (3, 1)] + ([(2, 3)] if chased else []):
            st.do_move(m, go.BLACK)
        st.do_move((2, 2), go.WHITE)
        if breaker:
            st.do_move(breaker, go.WHITE)
        st.current_player = to_move
        return st
    pp = Preprocess(LADDER)
    for breaker, works in [(None, 1), ((6, 6), 0), ((7, 6), 0), ((0, 8), 1)]:
        st = board(breaker, go.BLACK)
        t = pp.state_to_tensor(st)[0]
        assert t[0][2, 3] == works and t[0][3, 2] == 0 and t[1].sum() == 0
        if breaker is None:
            assert t[0].sum() == 1
        b, a, pl, ko = st.pack()  # (a lone corner stone is itself ladder-capturable: compare whole planes with the spec)
        assert np.array_equal(t, og.encode_batch(b[None], a[None], [pl], [ko], 9, LADDER)[0])
    for breaker, escapes in [(None, 0), ((6, 6), 1)]:
        t = pp.state_to_tensor(board(breaker, go.WHITE, chased=True))[0]
        assert t[1][3, 2] == escapes


@pytest.mark.parametrize("n,size,seed", [(3000, 19, 31), (1500, 9, 32), (600, 7, 33), (500, 13, 34)])
def test_all_48_planes_match_oracle(n, size, seed):
    b, a, pl, ko = go.playout_positions(seed, n, size=size, max_moves=size * size + 40, per_game=8)
    want = og.encode_batch(b, a, pl, ko, size, DEFAULT_FEATURES, threads=os.cpu_count() or 1)
    got = gpu_encode(DEFAULT_FEATURES, b, a, pl, ko, size)
    assert got.shape == (n, 48, size, size)
    assert_same(got, want, DEFAULT_FEATURES)
    assert want[:, 44].sum() > 0 and want[:, 45].sum() > 0  # the fixture exercises both planes


def test_reference_hdf5_fixture_real_games(golden_dir):
    """1033 real positions of the five AlphaGo - Lee Sedol games, planes written by the reference's own converter
    (tests/golden/ref_hdf5_fixture.npz): product GameState replay + CUDA encoder, bit-exact"""
    from test_oracle import load_hdf5_fixture
    states, actions, offsets, feats = load_hdf5_fixture(golden_dir)
    pp = Preprocess(feats)
    for start, count in offsets:
        st = go.GameState(19)
        batch = []
        for i in range(start, start + count):
            batch.append(st.copy())
            st.do_move((int(actions[i][0]), int(actions[i][1])))
        got = pp.states_to_tensor(batch).cpu().numpy()
        assert_same(got, states[start:start + count], feats)
