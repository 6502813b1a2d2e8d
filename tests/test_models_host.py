"""CPU tests of the host-side model plumbing that needs no device: the Keras-1.0 JSON layer list
CNNPolicy writes/reads (tests/test_data/minimodel.json shape) and the weight-file container."""
import json
import os

import numpy as np

from alphago_b200 import weights_io
from alphago_b200.models import SequentialModel, CNNPolicy


def test_policy_spec_matches_reference_minimodel_shape():
    # minimodel.json: 12 planes, 5x5 conv + 4x 3x3 conv with 16 filters, 1x1 conv, Flatten, softmax
    spec = SequentialModel.build_policy(12, 19, 16, 5, [5, 3, 3, 3, 3])
    assert [l["class_name"] for l in spec] == ["Convolution2D"] * 6 + ["Flatten", "Activation"]
    c0 = spec[0]["config"]
    assert c0["batch_input_shape"] == [None, 12, 19, 19] and c0["nb_row"] == 5 and c0["border_mode"] == "same"
    assert c0["dim_ordering"] == "th" and c0["activation"] == "relu" and c0["init"] == "uniform"
    assert spec[5]["config"]["nb_filter"] == 1 and spec[5]["config"]["nb_row"] == 1 and spec[5]["config"]["activation"] == "linear"
    assert spec[7]["config"]["activation"] == "softmax"
    json.loads(json.dumps({"class_name": "Sequential", "config": spec}))
    params, widths = CNNPolicy._params({"layers": 4, "filter_width_3": 1})
    assert widths == [5, 3, 1, 3] and params["filters_per_layer"] == 128 and params["board"] == 19
    v = SequentialModel.build_value(49, 19, 152, 12, [5] + [3] * 11)
    assert [l["class_name"] for l in v][-3:] == ["Flatten", "Dense", "Dense"] and v[-1]["config"]["activation"] == "tanh"


def test_weight_file_roundtrip(tmp_path):
    spec = SequentialModel.build_policy(12, 19, 16, 2, [5, 3])
    rng = np.random.RandomState(0)
    weights = [rng.randn(16, 12, 5, 5), rng.randn(16), rng.randn(16, 16, 3, 3), rng.randn(16), rng.randn(1, 16, 1, 1), rng.randn(1)]
    weights = [w.astype(np.float32) for w in weights]
    path = os.path.join(str(tmp_path), "weights.00000.hdf5")
    weights_io.save(path, spec, weights)
    assert os.path.exists(path)
    back = weights_io.load(path, spec)
    assert len(back) == 6 and all(np.array_equal(a, b) for a, b in zip(weights, back))


def test_reference_minimodel_json_parses():
    """the reference's own fixture (build container only; the tree is absent on the GPU box)"""
    import pytest
    path = "/root/reference/tests/test_data/minimodel.json"
    if not os.path.exists(path):
        pytest.skip("reference tree not present")
    spec = json.load(open(path))
    assert spec["feature_list"] == ["board", "ones", "turns_since"]
    g = SequentialModel.parse_layers(json.loads(spec["keras_model"])["config"])
    assert g == {"in_planes": 12, "board": 19, "filters": [16] * 5, "widths": [5, 3, 3, 3, 3], "head": 0}
