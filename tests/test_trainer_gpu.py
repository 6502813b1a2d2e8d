"""GPU test of the run_training drop-in (AlphaGo/training/supervised_policy_trainer.py:83-213), after the
reference's tests/test_supervised_policy_trainer.py: train the minimodel architecture for an epoch on the five
AlphaGo - Lee Sedol games and check the files it must leave behind; then resume."""
import json
import os

import numpy as np
import pytest

from alphago_b200.models import CNNPolicy
from alphago_b200.supervised_policy_trainer import run_training

pytestmark = pytest.mark.gpu


def test_run_training_outputs_and_resume(tmp_path, golden_dir):
    from test_oracle import load_hdf5_fixture
    states, actions, _, feats = load_hdf5_fixture(golden_dir)
    d = str(tmp_path)
    data = os.path.join(d, "features.npz")
    np.savez(data, states=states, actions=actions)
    model = os.path.join(d, "minimodel.json")
    # the shape of tests/test_data/minimodel.json: 12 planes, 5x5 + 4x 3x3 convs with 16 filters
    CNNPolicy(feats, layers=5, filters_per_layer=16, seed=3).save_model(model)
    out = os.path.join(d, "out")
    meta = run_training([model, data, out, "--epochs", "2", "--minibatch", "32", "--learning-rate", "0.05", "--seed", "1"])
    for name in ("metadata.json", "shuffle.npz", "weights.00000.hdf5", "weights.00001.hdf5"):
        assert os.path.exists(os.path.join(out, name)), name
    on_disk = json.load(open(os.path.join(out, "metadata.json")))
    assert on_disk == meta and len(meta["epochs"]) == 2 and meta["training_data"] == data and meta["model_file"] == model
    for logs in meta["epochs"]:
        assert set(logs) == {"loss", "acc", "val_loss", "val_acc"} and np.isfinite(logs["loss"]) and 0 <= logs["acc"] <= 1
    # a 16-filter net with the reference's tiny uniform init barely moves in two short epochs: it must not diverge
    assert meta["epochs"][0]["loss"] < np.log(361) + 0.2 and meta["epochs"][1]["loss"] < meta["epochs"][0]["loss"] + 0.05, meta["epochs"]
    with open(os.path.join(out, "shuffle.npz"), "rb") as f:
        shuffle = np.load(f)
    assert sorted(shuffle.tolist()) == list(range(1033))
    # resume: weights come from the checkpoint, metadata is appended, the split is reused
    meta2 = run_training([model, data, out, "--epochs", "1", "--minibatch", "32", "--weights", "weights.00001.hdf5",
                          "--epoch-length", "128"])
    assert len(meta2["epochs"]) == 3
    with open(os.path.join(out, "shuffle.npz"), "rb") as f:
        assert np.array_equal(np.load(f), shuffle)
    assert np.isfinite(meta2["epochs"][2]["loss"]) and meta2["epochs"][2]["loss"] < np.log(361) + 0.2
