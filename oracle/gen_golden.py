"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

    python -m oracle.gen_golden            (needs /root/reference; takes ~1-2 minutes)

CPU ORACLE tooling -- test infrastructure only.  The fixtures are what pins oracle/go_oracle.c
(and, through it, the CUDA encoder) to the reference's own code on machines where the
reference tree is absent (the GPU box).  Files written:

  ref_games.npz      full move sequences of seeded random playouts on the reference GameState
                     (go.py) with the board / ko / player after every move, liberty_counts and
                     the legal + sensible move masks at sampled moves -> pins the rules.
  ref_planes.npz     packed encoder inputs + the reference Preprocess output (46 planes: the
                     default list minus the two ladder features, which raise
                     NotImplementedError in the reference) for random-playout positions on
                     19x19 / 13x13 / 9x9 / 7x7 boards -> pins the encoder.
  ref_kat.npz        the hand-made 7x7 boards of the reference's tests/test_preprocessing.py
                     (simple_board, self_atari_board) as move lists + reference planes.
  ladder_spec.npz    ladder planes from oracle/ladder_spec.py run on the reference GameState
                     (build-owned spec; parity unpinned against the reference).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim, ladder_spec  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
FEATS46 = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
           "liberties_after", "sensibleness", "zeros"]


def pack(gs):
    """literal packing of a reference GameState (history walk of preprocessing.py:20-43)"""
    size = gs.size
    board = gs.board.astype(np.int8).reshape(-1)
    age = np.full(size * size, 255, np.uint8)
    depth = 0
    for m in gs.history[::-1]:
        if m is not None and gs.board[m] != 0 and age[m[0] * size + m[1]] == 255:
            age[m[0] * size + m[1]] = depth
        if depth < 7:
            depth += 1
    ko = -1 if gs.ko is None else gs.ko[0] * size + gs.ko[1]
    return board, age, np.int8(gs.current_player), np.int16(ko)


def playout(go, size, rng, nmoves):
    """uniform over get_legal_moves(include_eyes=False) (AlphaGo/ai.py:15); pass if none"""
    gs = go.GameState(size)
    for _ in range(nmoves):
        mv = gs.get_legal_moves(include_eyes=False)
        gs.do_move(mv[rng.randint(len(mv))] if mv else None)
        yield gs
        if gs.is_end_of_game:
            return


def main():
    go, pp = ref_shim.load()
    os.makedirs(OUT, exist_ok=True)

    # ---- rules: full games
    rng = np.random.RandomState(20160309)
    games = {}
    for gi, (size, nm) in enumerate([(19, 420), (19, 420), (19, 300), (13, 260), (9, 160), (9, 160), (7, 90), (5, 60)]):
        moves, boards, kos, players, libs, legal, sens, sample_at = [], [], [], [], [], [], [], []
        for t, gs in enumerate(playout(go, size, rng, nm)):
            m = gs.history[-1]
            moves.append(-1 if m is None else m[0] * size + m[1])
            boards.append(gs.board.astype(np.int8).copy())
            kos.append(-1 if gs.ko is None else gs.ko[0] * size + gs.ko[1])
            players.append(gs.current_player)
            if t % 7 == 0:
                sample_at.append(t)
                libs.append(gs.liberty_counts.astype(np.int16).copy())
                lm = np.zeros((size, size), np.uint8)
                for p in gs.get_legal_moves():
                    lm[p] = 1
                sm = np.zeros((size, size), np.uint8)
                for p in gs.get_legal_moves(include_eyes=False):
                    sm[p] = 1
                legal.append(lm); sens.append(sm)
        games["g%d_size" % gi] = np.int32(size)
        games["g%d_moves" % gi] = np.asarray(moves, np.int16)
        games["g%d_boards" % gi] = np.asarray(boards)
        games["g%d_ko" % gi] = np.asarray(kos, np.int16)
        games["g%d_player" % gi] = np.asarray(players, np.int8)
        games["g%d_sample_at" % gi] = np.asarray(sample_at, np.int32)
        games["g%d_libs" % gi] = np.asarray(libs)
        games["g%d_legal" % gi] = np.asarray(legal)
        games["g%d_sensible" % gi] = np.asarray(sens)
    games["n_games"] = np.int32(8)
    np.savez_compressed(os.path.join(OUT, "ref_games.npz"), **games)
    print("ref_games.npz written")

    # ---- encoder planes
    prep = pp.Preprocess(FEATS46)
    out = {"features": np.asarray(FEATS46)}
    rng = np.random.RandomState(48)
    for size, ngames, nm, every in [(19, 10, 420, 11), (13, 3, 250, 9), (9, 4, 150, 6), (7, 4, 80, 4)]:
        B, A, P, K, X = [], [], [], [], []
        for _ in range(ngames):
            off = rng.randint(every)
            for t, gs in enumerate(playout(go, size, rng, nm)):
                if (t + off) % every:
                    continue
                b, a, p, k = pack(gs)
                B.append(b); A.append(a); P.append(p); K.append(k)
                X.append(prep.state_to_tensor(gs)[0].astype(np.uint8))
        # positions reached through copy() carry an aliased, longer history (go.py:167)
        gs = go.GameState(size)
        for _ in range(12):
            mv = gs.get_legal_moves(include_eyes=False)
            prev = gs.copy()
            gs.do_move(mv[rng.randint(len(mv))])
            b, a, p, k = pack(prev)
            B.append(b); A.append(a); P.append(p); K.append(k)
            X.append(prep.state_to_tensor(prev)[0].astype(np.uint8))
        out["board_%d" % size] = np.asarray(B); out["age_%d" % size] = np.asarray(A)
        out["player_%d" % size] = np.asarray(P); out["ko_%d" % size] = np.asarray(K)
        out["planes_%d" % size] = np.asarray(X)
        print("size", size, "positions", len(B))
    np.savez_compressed(os.path.join(OUT, "ref_planes.npz"), **out)
    print("ref_planes.npz written")

    # ---- the reference's own hand-made KAT boards (tests/test_preprocessing.py:7-75)
    simple = [(0, 0), (1, 0), (0, 1), (1, 1), (0, 2), (3, 4), (3, 3), (4, 5), (4, 2), (5, 4), (5, 3), (4, 3), (4, 4)]
    atari_b = [(2, 4), (4, 4), (6, 0)]
    atari_w = [(1, 0), (5, 0), (2, 3), (4, 3), (1, 4), (5, 4), (2, 5), (3, 5), (4, 5)]
    kat = {}
    gs = go.GameState(size=7)
    for m in simple:
        gs.do_move(m)
    kat["simple_moves"] = np.asarray(simple, np.int16)
    kat["simple_planes"] = prep.state_to_tensor(gs)[0].astype(np.uint8)
    kat["simple_ko"] = np.int16(gs.ko[0] * 7 + gs.ko[1])
    gs = go.GameState(size=7)
    for m in atari_b:
        gs.do_move(m, go.BLACK)
    for m in atari_w:
        gs.do_move(m, go.WHITE)
    kat["atari_black"] = np.asarray(atari_b, np.int16)
    kat["atari_white"] = np.asarray(atari_w, np.int16)
    kat["atari_planes"] = prep.state_to_tensor(gs)[0].astype(np.uint8)
    kat["atari_player"] = np.int8(gs.current_player)
    np.savez_compressed(os.path.join(OUT, "ref_kat.npz"), **kat)
    print("ref_kat.npz written")

    # ---- ladder planes from the python spec on the reference GameState
    rng = np.random.RandomState(7)
    B, A, P, K, LC, LE = [], [], [], [], [], []
    for size, ngames, nm, every in [(19, 2, 260, 37), (9, 3, 90, 13), (7, 2, 50, 9)]:
        for _ in range(ngames):
            for t, gs in enumerate(playout(go, size, rng, nm)):
                if t % every != every - 1:
                    continue
                b, a, p, k = pack(gs)
                c, e = ladder_spec.ladder_planes(gs)
                pad = np.zeros(361, np.int8); pad[:size * size] = b
                B.append(pad); P.append(p); K.append(k)
                A.append(np.int32(size))
                cc = np.zeros(361, np.uint8); cc[:size * size] = c.reshape(-1)
                ee = np.zeros(361, np.uint8); ee[:size * size] = e.reshape(-1)
                LC.append(cc); LE.append(ee)
                print("ladder", size, t, int(c.sum()), int(e.sum()), flush=True)
    np.savez_compressed(os.path.join(OUT, "ladder_spec.npz"), board=np.asarray(B), size=np.asarray(A),
                        player=np.asarray(P), ko=np.asarray(K), capture=np.asarray(LC), escape=np.asarray(LE))
    print("ladder_spec.npz written")


def export_reference_hdf5_fixture():
    """tests/test_data/hdf5/alphago-vs-lee-sedol-features.hdf5 of the reference (written by its own
    game_converter.py from the five AlphaGo - Lee Sedol SGFs: 1033 (state planes, action) pairs, features
    board/ones/turns_since) re-packed as tests/golden/ref_hdf5_fixture.npz: planes bit-packed, actions, per-game
    [start, count].  Read with alphago_b200/hdf5_lite.py (no h5py in this image)."""
    from alphago_b200 import hdf5_lite
    path = os.path.join(ref_shim.REFERENCE_ROOT, "tests", "test_data", "hdf5", "alphago-vs-lee-sedol-features.hdf5")
    f = hdf5_lite.File(path)
    states, actions = f["states"][:], f["actions"][:]
    assert states.shape == (1033, 12, 19, 19) and states.max() == 1
    names = f["file_offsets"].keys()
    offsets = np.asarray([f["file_offsets"][k][:] for k in names], np.int32)
    np.savez_compressed(os.path.join(OUT, "ref_hdf5_fixture.npz"), states_bits=np.packbits(states.reshape(1033, -1), axis=1),
                        shape=np.asarray(states.shape, np.int32), actions=actions, offsets=offsets,
                        names=np.asarray(names), features=np.asarray(["board", "ones", "turns_since"]))
    print("ref_hdf5_fixture.npz written", offsets.tolist())


def export_reference_sgf_fixtures():
    """Game records the reference keeps for its own converter tests (tests/test_data/sgf: the five AlphaGo - Lee Sedol
    games behind alphago-vs-lee-sedol-features.hdf5, and ab_aw.sgf with set-up stones) copied to tests/golden/sgf as
    test INPUT data for alphago_b200/game_converter.py.  Run: python oracle/gen_golden.py sgf"""
    import shutil
    src = os.path.join(ref_shim.REFERENCE_ROOT, "tests", "test_data", "sgf")
    dst = os.path.join(OUT, "sgf")
    os.makedirs(dst, exist_ok=True)
    for name in sorted(os.listdir(src)):
        if name.endswith(".sgf") and (name == "ab_aw.sgf" or "Sedol" in name):
            shutil.copyfile(os.path.join(src, name), os.path.join(dst, name))
            print("copied", name)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "hdf5":
        export_reference_hdf5_fixture()
    elif len(sys.argv) > 1 and sys.argv[1] == "sgf":
        export_reference_sgf_fixtures()
    else:
        main()
        export_reference_hdf5_fixture()

