"""GPU tests of the SGF -> dataset converter (alphago_b200/game_converter.py), after the reference's
tests/test_game_converter.py: the whole path (SGF reader, host rules engine, CUDA encoder in bulk, dataset file) must
reproduce the reference converter's own output for the five AlphaGo - Lee Sedol games bit for bit."""
import os

import numpy as np
import pytest

from test_oracle import load_hdf5_fixture

from alphago_b200.game_converter import game_converter, run_game_converter
from alphago_b200.supervised_policy_trainer import open_dataset

pytestmark = pytest.mark.gpu


def _read_offsets(path):
    with open(path, "rb") as f:
        magic = f.read(2)
    if magic == b"PK":
        z = np.load(path)
        return {str(n): [int(o[0]), int(o[1])] for n, o in zip(z["file_offsets_names"], z["file_offsets"])}
    from alphago_b200 import hdf5_lite
    g = hdf5_lite.open_file(path, "r")["file_offsets"]
    return {k: [int(v) for v in g[k][:]] for k in g.keys()}


def test_sgfs_to_hdf5_reproduces_reference_fixture(golden_dir, tmp_path):
    states, actions, offsets, feats = load_hdf5_fixture(golden_dir)
    g = np.load(os.path.join(golden_dir, "ref_hdf5_fixture.npz"))
    files = [os.path.join(golden_dir, "sgf", str(n).split("::")[-1]) for n in g["names"]]   # the fixture's own order
    out = str(tmp_path / "features.hdf5")
    conv = game_converter(feats, batch=300)   # several encoder launches, games straddling them
    conv.sgfs_to_hdf5(files, out, bd_size=19)
    assert os.path.exists(out) and not os.path.exists(str(tmp_path / ".tmp.features.hdf5"))
    s, a = open_dataset(out)
    assert np.array_equal(np.asarray(s[:]), states) and np.array_equal(np.asarray(a[:]), actions)
    got = _read_offsets(out)
    assert [got[f.replace("/", ":")] for f in files] == [[int(o[0]), int(o[1])] for o in offsets]


def test_convert_game_generator_contract(golden_dir):
    """game_converter.py:22-40: (1, F, size, size) planes and the (x, y) move, one pair per non-pass move"""
    states, actions, offsets, feats = load_hdf5_fixture(golden_dir)
    g = np.load(os.path.join(golden_dir, "ref_hdf5_fixture.npz"))
    name = str(g["names"][0]).split("::")[-1]
    conv = game_converter(feats)
    pairs = []
    for x, move in conv.convert_game(os.path.join(golden_dir, "sgf", name), 19):
        pairs.append((x, move))
        if len(pairs) == 12:
            break
    for i, (x, move) in enumerate(pairs):
        assert x.shape == (1, 12, 19, 19) and np.array_equal(x[0].astype(np.uint8), states[i])
        assert tuple(move) == (int(actions[i][0]), int(actions[i][1]))


def test_directory_conversion_and_training_on_it(golden_dir, tmp_path):
    """tests/test_game_converter.py:15-18 (--directory) and then the SL trainer on the file it wrote"""
    import warnings
    from alphago_b200.models import CNNPolicy
    from alphago_b200.supervised_policy_trainer import run_training
    out = str(tmp_path / "all.h5")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        run_game_converter(["--features", "board,ones,turns_since", "--outfile", out, "--directory", os.path.join(golden_dir, "sgf")])
    s, a = open_dataset(out)
    offs = _read_offsets(out)
    assert len(offs) == 6 and len(s) == len(a) == sum(v[1] for v in offs.values()) == 1033 + 290   # + ab_aw.sgf's 290 moves
    model = str(tmp_path / "mini.json")
    CNNPolicy(["board", "ones", "turns_since"], layers=3, filters_per_layer=32, seed=2).save_model(model)
    meta = run_training([model, out, str(tmp_path / "run"), "--epochs", "1", "--minibatch", "32", "--epoch-length", "256",
                         "--learning-rate", "0.05", "--seed", "3"])
    assert len(meta["epochs"]) == 1 and np.isfinite(meta["epochs"][0]["loss"])
