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


# ----------------------------------------------------------------------------------------------------------------------
# stage-2 pieces of the reference that ARE runnable here (pure Python around the Keras calls)
def _load_reference_policy_module():
    """import AlphaGo/models/policy.py UNMODIFIED.  Its imports of keras (absent here) and of the third-party `sgf`
    package (pulled in by AlphaGo/util.py:1) are satisfied by empty stand-in modules in sys.modules: none of the code
    exercised below (policy.py:44-91) touches them."""
    import types
    ref_shim.load()
    for name in ["keras", "keras.models", "keras.layers", "keras.layers.core", "keras.layers.convolutional",
                 "keras.backend", "sgf"]:
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    km, kl = sys.modules["keras.models"], sys.modules["keras.layers"]
    km.Sequential = km.model_from_json = None
    kl.convolutional = sys.modules["keras.layers.convolutional"]
    sys.modules["keras.layers.core"].Activation = sys.modules["keras.layers.core"].Flatten = None
    sys.modules["keras"].backend = sys.modules["keras.backend"]
    sys.modules["keras"].models, sys.modules["keras"].layers = km, kl
    import AlphaGo.models.policy as pol
    return pol


def export_policy_glue():
    """tests/golden/ref_policy_glue.npz: what CNNPolicy.eval_state / batch_eval_state / _select_moves_and_normalize
    (policy.py:44-91) return, run UNMODIFIED on reference GameStates, with `forward` replaced by a table lookup (the
    network itself needs Keras).  Recorded per case: the move lists that rebuild the states, the fake network output,
    the moves asked for, and the reference's [(move, prob)] answers."""
    go, pp = ref_shim.load()
    pol = _load_reference_policy_module()
    rng = np.random.RandomState(1444)
    size = 19
    policy = object.__new__(pol.CNNPolicy)
    policy.preprocessor = pp.Preprocess(["board", "ones", "turns_since"])
    # positions: seeded playouts; record move lists so that the product's GameState can replay them
    games, states = [], []
    for g in range(6):
        gs = go.GameState(size)
        nm = [0, 30, 120, 250, 330, 400][g]
        for _ in range(nm):
            mv = gs.get_legal_moves(include_eyes=False)
            gs.do_move(mv[rng.randint(len(mv))] if mv else None)
        games.append(np.asarray([-1 if m is None else m[0] * size + m[1] for m in gs.history], np.int16))
        states.append(gs)
    n = len(states)
    net_out = rng.gamma(0.3, 1.0, size=(n, size * size)).astype(np.float32)
    net_out /= net_out.sum(axis=1, keepdims=True)
    seen = []

    def fake_forward(inpt):
        seen.append(inpt.shape)
        assert inpt.shape[1:] == (12, size, size)
        return net_out[:inpt.shape[0]] if inpt.shape[0] > 1 else net_out[fake_forward.row:fake_forward.row + 1]
    fake_forward.row = 0
    policy.forward = fake_forward
    out = {"n_states": np.int32(n), "size": np.int32(size), "net_out": net_out}
    for i, g in enumerate(games):
        out["moves_%d" % i] = g

    def flat(res):
        res = list(res)
        return (np.asarray([m[0] * size + m[1] for m, _ in res], np.int32), np.asarray([p for _, p in res], np.float64))
    # batch_eval_state, default (legal) moves
    res = policy.batch_eval_state(states)
    for i in range(n):
        out["batch_default_idx_%d" % i], out["batch_default_p_%d" % i] = flat(res[i])
    # batch_eval_state with explicit move lists, incl. an empty one and a repeated move
    asked = []
    for i, gs in enumerate(states):
        legal = gs.get_legal_moves()
        if i == 1:
            asked.append([])
        elif i == 2:
            asked.append([legal[0], legal[0], legal[-1]])
        else:
            asked.append([legal[j] for j in rng.permutation(len(legal))[:1 + 7 * i]])
    res = policy.batch_eval_state(states, asked)
    for i in range(n):
        r = list(res[i]) if not isinstance(res[i], list) else res[i]
        out["asked_%d" % i] = np.asarray([m[0] * size + m[1] for m in asked[i]], np.int32)
        out["batch_asked_idx_%d" % i], out["batch_asked_p_%d" % i] = flat(r)
    # eval_state one by one (default and explicit moves)
    for i, gs in enumerate(states):
        fake_forward.row = i
        out["eval_default_idx_%d" % i], out["eval_default_p_%d" % i] = flat(policy.eval_state(gs))
        out["eval_asked_idx_%d" % i], out["eval_asked_p_%d" % i] = flat(policy.eval_state(gs, asked[i])) \
            if asked[i] else flat(policy.eval_state(gs, asked[i]))
    # contracts
    out["empty_batch_is_empty_list"] = np.int8(policy.batch_eval_state([]) == [])
    try:
        policy.batch_eval_state([go.GameState(19), go.GameState(13)])
        out["mixed_size_error"] = np.asarray("")
    except ValueError as e:
        out["mixed_size_error"] = np.asarray(str(e))
    np.savez_compressed(os.path.join(OUT, "ref_policy_glue.npz"), **out)
    print("ref_policy_glue.npz written:", n, "states, forward calls", len(seen), "error text:", out["mixed_size_error"])


def _exec_reference_lines(path, ranges, namespace):
    """run line ranges (1-based, inclusive) of a reference source file as they stand.  supervised_policy_trainer.py as a
    whole is Python 2 (print statement at :115) and imports keras/h5py; the pieces pinned here are valid Python 3."""
    lines = open(path).read().split("\n")
    for a, b in ranges:
        exec(compile("\n" * (a - 1) + "\n".join(lines[a - 1:b]), path, "exec"), namespace)
    return namespace


def export_generator():
    """tests/golden/ref_generator.npz: shuffled_hdf5_batch_generator + one_hot_action + BOARD_TRANSFORMATIONS
    (supervised_policy_trainer.py:10-40,71-80) and MetadataWriterCallback (spt.py:43-68), executed from the reference
    source by line range.  Records, for seeded np.random: the dataset, the index list (its length is NOT a multiple of
    the batch, so batches straddle the wrap, spt.py:27-40), the data index and transform the generator used for every
    sample, and the (X, Y) batches it yielded; each of the 8 transforms applied to an asymmetric plane; and the metadata
    the callback holds after every epoch of two made-up histories."""
    spt = os.path.join(ref_shim.REFERENCE_ROOT, "AlphaGo", "training", "supervised_policy_trainer.py")
    ns = _exec_reference_lines(spt, [(10, 40), (71, 80)], {"np": np})
    T = ns["BOARD_TRANSFORMATIONS"]
    out = {}
    rng = np.random.RandomState(99)
    # each transform alone, on asymmetric arrays (19x19 and 7x7), and on a one-hot action
    for size in (19, 7):
        plane = rng.randint(0, 255, (size, size)).astype(np.uint8)
        out["plane_%d" % size] = plane
        out["plane_tf_%d" % size] = np.asarray([t(plane) for t in T])
        acts = np.asarray([(0, 0), (0, size - 1), (2, 5), (size - 1, 3), (size // 2, size // 2)], np.uint8)
        out["acts_%d" % size] = acts
        out["acts_tf_%d" % size] = np.asarray([[int(np.argmax(t(ns["one_hot_action"](tuple(a), size)).flatten()))
                                                 for a in acts] for t in T], np.int32)
    # the generator
    n_data, planes, size, batch = 37, 5, 19, 8
    states = (rng.rand(n_data, planes, size, size) < 0.3).astype(np.uint8)
    actions = rng.randint(0, size, (n_data, 2)).astype(np.uint8)
    indices = rng.permutation(n_data)[:21]   # 21 indices, batch 8: the third batch straddles the wrap
    log_idx, log_tf = [], []

    class LoggedStates(object):
        shape = states.shape

        def __getitem__(self, i):
            log_idx.append(int(i))
            return states[i]

    class LoggedActions(object):
        def __getitem__(self, i):
            return tuple(actions[i])   # the HDF5 row indexes the one-hot as (x, y): spt.py:14

    def tagged(k):
        def f(feature):
            if feature.shape == (size, size) and (not log_tf or len(log_tf) < len(log_idx)):
                log_tf.append(k)
            return T[k](feature)
        return f
    np.random.seed(2016)
    gen = ns["shuffled_hdf5_batch_generator"](LoggedStates(), LoggedActions(), indices, batch, [tagged(k) for k in range(8)])
    X, Y = [], []
    for _ in range(7):   # 56 samples = 2 full passes + 14
        xb, yb = next(gen)
        assert xb.dtype == np.float64 and yb.dtype == np.float64 and set(np.unique(xb)) <= {0.0, 1.0}
        X.append(xb.astype(np.uint8).copy()); Y.append(yb.copy())   # the generator reuses its buffers (spt.py:24-25)
    assert len(log_idx) == len(log_tf) == 7 * batch
    out.update(gen_states=states, gen_actions=actions, gen_indices=indices.astype(np.int64), gen_batch=np.int32(batch),
               gen_used_index=np.asarray(log_idx, np.int64), gen_used_transform=np.asarray(log_tf, np.uint8),
               gen_X=np.asarray(X), gen_Y=np.asarray(Y, np.uint8))
    # MetadataWriterCallback: best-epoch bookkeeping with and without val_loss
    import json
    import tempfile
    ns2 = _exec_reference_lines(spt, [(43, 68)], {"np": np, "json": json, "Callback": object})
    hist = {"with_val": [{"loss": 5.0, "acc": .01, "val_loss": 4.0, "val_acc": .02}, {"loss": 4.0, "acc": .02, "val_loss": 4.5, "val_acc": .03},
                         {"loss": 3.0, "acc": .03, "val_loss": 3.5, "val_acc": .04}, {"loss": 2.5, "acc": .04, "val_loss": 3.5, "val_acc": .05}],
            "without_val": [{"loss": 5.0, "acc": .01}, {"loss": 5.5, "acc": .02}, {"loss": 4.0, "acc": .03}]}
    record = {}
    for name, logs in hist.items():
        with tempfile.TemporaryDirectory() as d:
            cb = ns2["MetadataWriterCallback"](os.path.join(d, "metadata.json"))
            after = []
            for e, lg in enumerate(logs):
                cb.on_epoch_end(e, dict(lg))
                after.append(json.load(open(cb.file)))
            record[name] = {"logs": logs, "after": after}
    out["metadata_json"] = np.asarray(json.dumps(record))
    np.savez_compressed(os.path.join(OUT, "ref_generator.npz"), **out)
    print("ref_generator.npz written; used index", log_idx[:24], "transforms", log_tf[:16])


class FakePolicy(object):
    """stands in for CNNPolicy in the self-play pins: a deterministic, position-dependent weight for every move asked for
    (the players normalise), and the `model.input_shape` the trainer reads (rpt.py:51).  tests/test_rl_trainer_gpu.py holds
    the same class."""

    def __init__(self, salt, size=19, planes=46):
        self.salt = salt
        self.model = type("M", (), {"input_shape": (None, planes, size, size)})()

    def batch_eval_state(self, states, moves_lists=None):
        return [[(m, 1.0 + ((m[0] * 7 + m[1] * 13 + len(st.history) * 3 + self.salt) % 11)) for m in ml]
                for st, ml in zip(states, moves_lists)]


def export_selfplay():
    """tests/golden/ref_selfplay.npz: the reference's own self-play loop -- `make_training_pairs`
    (reinforcement_policy_trainer.py:16-76) driving `ProbabilisticPolicyPlayer.get_moves` (ai.py:24-68), both executed from
    the reference source by line range (the files as wholes are Python 2) on the reference GameState and Preprocess --
    with FakePolicy in place of the Keras networks: two mini-batches of two 19x19 games played to the end, one with the
    learner as Black and one as White.  Recorded: the planes and moves of every training pair, the winners.  This pins the
    pairing of positions and moves (incl. positions one ply old for a White learner), the history aliasing of
    GameState.copy() in the turns_since planes, pass and end-of-game handling, the sampling's use of np.random, and
    get_winner."""
    go, pp = ref_shim.load()
    ai = os.path.join(ref_shim.REFERENCE_ROOT, "AlphaGo", "ai.py")
    rpt = os.path.join(ref_shim.REFERENCE_ROOT, "AlphaGo", "training", "reinforcement_policy_trainer.py")
    ns = _exec_reference_lines(ai, [(24, 68)], {"np": np, "go": go})
    ns2 = _exec_reference_lines(rpt, [(16, 76)], {"np": np, "go": go, "GameState": go.GameState, "Preprocess": pp.Preprocess,
                                                  "flatten_idx": lambda pos, size: pos[0] * size + pos[1], "xrange": range})
    Player = ns["ProbabilisticPolicyPlayer"]
    out = {"features": np.asarray(FEATS46)}
    want = {go.BLACK: None, go.WHITE: None}
    seed = 100
    while any(v is None for v in want.values()):
        np.random.seed(seed)
        colour = int(np.random.choice([go.BLACK, go.WHITE]))   # what make_training_pairs draws first (rpt.py:54)
        if want[colour] is None:
            want[colour] = seed
        seed += 1
    for colour, seed in want.items():
        tag = "black" if colour == go.BLACK else "white"
        learner, opponent = Player(FakePolicy(3), temperature=0.67), Player(FakePolicy(8))
        np.random.seed(seed)
        X_list, y_list, winners = ns2["make_training_pairs"](learner, opponent, FEATS46, 2)
        out["seed_" + tag] = np.int32(seed)
        out["winners_" + tag] = np.asarray(winners, np.int8)
        for g, (X, y) in enumerate(zip(X_list, y_list)):
            assert X.shape[1:] == (46, 19, 19) and set(np.unique(X)) <= {0.0, 1.0} and (y.sum(1) == 1).all()
            out["X_%s_%d" % (tag, g)] = np.packbits(X.astype(np.uint8).reshape(len(X), -1), axis=1)
            out["y_%s_%d" % (tag, g)] = y.argmax(1).astype(np.int16)
            print("self-play", tag, "game", g, "pairs", len(X), "winner", winners[g], flush=True)
    np.savez_compressed(os.path.join(OUT, "ref_selfplay.npz"), **out)
    print("ref_selfplay.npz written")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "hdf5":
        export_reference_hdf5_fixture()
    elif len(sys.argv) > 1 and sys.argv[1] == "sgf":
        export_reference_sgf_fixtures()
    elif len(sys.argv) > 1 and sys.argv[1] == "selfplay":
        export_selfplay()
    elif len(sys.argv) > 1 and sys.argv[1] == "stage2":
        export_policy_glue()
        export_generator()
        export_selfplay()
    else:
        main()
        export_reference_hdf5_fixture()
        export_policy_glue()
        export_generator()

