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


def test_host_feeder_matches_device_resident_training(golden_dir):
    """HostFeeder (host dataset, pinned staging, copy stream, metrics one step late) must train exactly like the
    device-resident path: same minibatches, same symmetries -> bitwise the same weights, the same metrics."""
    import torch
    from test_oracle import load_hdf5_fixture
    from alphago_b200 import nn
    from alphago_b200.training import HostFeeder, SLTrainer
    states, actions, _, feats = load_hdf5_fixture(golden_dir)
    states = np.ascontiguousarray(states[:256], np.uint8)
    actions = np.ascontiguousarray(actions[:256], np.uint8)
    batch, steps = 32, 6
    rng = np.random.RandomState(5)
    plan = [(rng.randint(0, len(states), batch), rng.randint(0, 8, batch)) for _ in range(steps)]

    def make():
        net = nn.ConvNet(19, states.shape[1], [32, 32, 32], [5, 3, 3], max_batch=batch)
        net.init_uniform(seed=9)
        return net, SLTrainer(net, lr=0.05, decay=0.0, momentum=0.0)

    net_a, tr_a = make()
    d_states, d_actions = torch.from_numpy(states).cuda(), torch.from_numpy(actions).cuda()
    want = []
    for idx, tf in plan:
        x, y = tr_a.assemble(d_states, d_actions, idx, tf)
        tr_a.step(x, y)
        want.append(tuple(float(v) for v in tr_a.loss_acc.cpu()))

    net_b, tr_b = make()
    feeder = HostFeeder(tr_b, states, actions, batch)
    got = []
    for idx, tf in plan:
        m = feeder.step(idx, tf)
        if m is not None:
            got.append(m)
    got.append(feeder.flush())
    # the loss sum is accumulated with atomics (order varies run to run): equal to rounding; the counts are exact
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert abs(g[0] - w[0]) <= 1e-5 * abs(w[0]) and g[1] == w[1]
    for l in range(3):
        wa, ba = net_a.get_conv(l)
        wb, bb = net_b.get_conv(l)
        assert np.array_equal(wa, wb) and np.array_equal(ba, bb)
    with pytest.raises(ValueError):
        feeder.step(plan[0][0][:5], plan[0][1][:5])


def test_run_training_streams_a_dataset_larger_than_the_device_budget(tmp_path, golden_dir):
    """A dataset above --device-dataset-mb is never copied to the device as a whole: run_training reads minibatches from
    the mapped HDF5 file on a worker thread (the reference's generator reads sample by sample, spt.py:27-36).  Same seed,
    same minibatches, same symmetries -> bitwise the same weights as the device-resident run."""
    from test_oracle import load_hdf5_fixture
    from alphago_b200 import hdf5_lite
    states, actions, _, feats = load_hdf5_fixture(golden_dir)
    d = str(tmp_path)
    data = os.path.join(d, "features.hdf5")
    with hdf5_lite.File(data, "w") as f:   # the converter's layout: LZF chunks of 64 / 1024 rows (game_converter.py:71-88)
        f.create_dataset("states", data=states, chunks=(64,) + states.shape[1:], compression="lzf")
        f.create_dataset("actions", data=actions, chunks=(1024, 2), compression="lzf")
    model = os.path.join(d, "model.json")
    # with its weights: a model file without them is initialised afresh by every load_model (policy.py:158-169)
    CNNPolicy(feats, layers=3, filters_per_layer=32, seed=5).save_model(model, os.path.join(d, "start.hdf5"))
    common = ["--epochs", "2", "--minibatch", "32", "--epoch-length", "320", "--learning-rate", "0.05", "--seed", "4"]
    meta_a = run_training([model, data, os.path.join(d, "resident")] + common)
    meta_b = run_training([model, data, os.path.join(d, "streamed")] + common + ["--device-dataset-mb", "1", "--io-threads", "3"])
    assert states.nbytes > 2 ** 20
    for ea, eb in zip(meta_a["epochs"], meta_b["epochs"]):
        for k in ("loss", "acc", "val_loss", "val_acc"):
            assert abs(ea[k] - eb[k]) <= 1e-5 * max(1.0, abs(ea[k])), (k, ea, eb)   # loss sums use float atomics
    for name in ("weights.00000.hdf5", "weights.00001.hdf5"):
        a = CNNPolicy.load_model(model); a.model.load_weights(os.path.join(d, "resident", name))
        b = CNNPolicy.load_model(model); b.model.load_weights(os.path.join(d, "streamed", name))
        assert all(np.array_equal(x, y) for x, y in zip(a.model.get_weights(), b.model.get_weights())), name
