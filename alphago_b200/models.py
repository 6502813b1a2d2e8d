"""`CNNPolicy` / `CNNValue`: drop-in mirrors of AlphaGo/models/policy.py:12-193 (and the value stack of
AlphaGo/models/value.py:12-28) on top of the device network (alphago_b200.nn.ConvNet).

`policy.model` is a small Keras-1.0-shaped object (`SequentialModel`): the reference's callers use
`.model.input_shape`, `.get_weights()/.set_weights()`, `.save_weights()/.load_weights()`, `.to_json()`
and, in the RL trainer, `.compile/.fit/.optimizer.lr.set_value` (policy.py:185-189,
reinforcement_policy_trainer.py:51,99-109, tests/test_policy.py:42-51).  Model JSON files written by
the reference (`{"keras_model": <Keras Sequential JSON>, "feature_list": [...], "weights_file"?}`,
policy.py:184-193) load unchanged, e.g. tests/test_data/minimodel.json.

Filter convention: Keras 1.0 on Theano computes a true convolution, i.e. it flips the stored filters;
the CUDA kernels cross-correlate.  `get_weights/set_weights` therefore flip both spatial axes
(`SequentialModel.flip_kernels`, default True) so arrays exchanged with Keras files mean the same net.
"""
import json
import os

import numpy as np

from . import nn
from .preprocessing import Preprocess


def flatten_idx(position, size):  # AlphaGo/util.py:6-8
    (x, y) = position
    return x * size + y


def unflatten_idx(idx, size):  # AlphaGo/util.py:11-13
    x, y = divmod(idx, size)
    return (x, y)


class _Scalar(object):
    """stands in for a Keras backend variable (optimizer.lr.set_value, rpt.py:99-101)"""

    def __init__(self, v):
        self.v = float(v)

    def set_value(self, v):
        self.v = float(v)

    def get_value(self):
        return self.v


class _SGD(object):
    """stands in for keras.optimizers.SGD (spt.py:199, rpt.py:108): exported as `SGD`"""

    def __init__(self, lr=0.01, momentum=0.0, decay=0.0):
        self.lr, self.momentum, self.decay = _Scalar(lr), float(momentum), float(decay)
        self.iterations = 0


SGD = _SGD


def _conv_layer_json(name, nb_filter, width, activation, input_shape=None):
    cfg = {"name": name, "trainable": True, "dim_ordering": "th", "nb_filter": nb_filter, "nb_row": width, "nb_col": width,
           "subsample": [1, 1], "init": "uniform", "border_mode": "same", "activation": activation,
           "W_regularizer": None, "b_regularizer": None, "activity_regularizer": None, "W_constraint": None,
           "b_constraint": None}
    if input_shape is not None:
        cfg["batch_input_shape"] = [None] + list(input_shape)
        cfg["input_dtype"] = "float32"
    return {"class_name": "Convolution2D", "config": cfg}


class SequentialModel(object):
    """The conv stack as a Keras-1.0-like Sequential: a layer list (JSON round-trippable) + weights on the GPU."""

    @staticmethod
    def parse_layers(layers):
        """Keras 1.0 Sequential layer list (the `config` of model.to_json()) -> geometry of the device net"""
        convs = [l["config"] for l in layers if l["class_name"] == "Convolution2D"]
        dense = [l["config"] for l in layers if l["class_name"] == "Dense"]
        if len(convs) < 2 or convs[-1]["nb_filter"] != 1 or convs[-1]["nb_row"] != 1:
            raise ValueError("expected KxK convolutions followed by a 1x1 convolution to one plane")
        shape = convs[0]["batch_input_shape"]
        for c in convs:
            if c.get("border_mode", "same") != "same" or c["nb_row"] != c["nb_col"]:
                raise ValueError("only square 'same' convolutions are supported")
            if c.get("dim_ordering", "th") != "th":
                raise ValueError("only dim_ordering 'th' (NCHW) models are supported")
        if dense and [d["output_dim"] for d in dense] != [256, 1]:
            raise ValueError("value head must be Dense(256) + Dense(1)")
        return {"in_planes": int(shape[1]), "board": int(shape[2]), "filters": [int(c["nb_filter"]) for c in convs[:-1]],
                "widths": [int(c["nb_row"]) for c in convs[:-1]], "head": nn.VALUE_HEAD if dense else nn.POLICY_HEAD}

    def __init__(self, layers, max_batch=64, flip_kernels=True):
        self.layers = layers
        self.flip_kernels = flip_kernels
        g = self.parse_layers(layers)
        self.in_planes, self.board, self.filters, self.widths, self.head = g["in_planes"], g["board"], g["filters"], g["widths"], g["head"]
        self.max_batch = int(max_batch)
        self.net = nn.ConvNet(self.board, self.in_planes, self.filters, self.widths, head=self.head, max_batch=self.max_batch)
        self.optimizer = None
        self.loss = None
        self._trainer = None

    # ------------------------------------------------------------------ construction
    @staticmethod
    def build_policy(input_dim, board, filters_per_layer, layers, widths):
        spec = [_conv_layer_json("convolution2d_1", filters_per_layer, widths[0], "relu", (input_dim, board, board))]
        for i in range(2, layers + 1):
            spec.append(_conv_layer_json("convolution2d_%d" % i, filters_per_layer, widths[i - 1], "relu"))
        spec.append(_conv_layer_json("convolution2d_%d" % (layers + 1), 1, 1, "linear"))
        spec.append({"class_name": "Flatten", "config": {"trainable": True, "name": "flatten_1"}})
        spec.append({"class_name": "Activation", "config": {"activation": "softmax", "trainable": True, "name": "activation_1"}})
        return spec

    @staticmethod
    def build_value(input_dim, board, filters_per_layer, layers, widths):
        spec = SequentialModel.build_policy(input_dim, board, filters_per_layer, layers, widths)[:-1]
        spec.append({"class_name": "Dense", "config": {"name": "dense_1", "output_dim": 256, "init": "uniform", "activation": "linear", "trainable": True}})
        spec.append({"class_name": "Dense", "config": {"name": "dense_2", "output_dim": 1, "init": "uniform", "activation": "tanh", "trainable": True}})
        return spec

    def ensure_batch(self, n):
        """the device buffers are sized for max_batch; grow (re-create, keep weights) on demand"""
        if n <= self.max_batch:
            return
        weights = self.get_weights()
        self.net.close()
        self.max_batch = int(max(n, 2 * self.max_batch))
        self.net = nn.ConvNet(self.board, self.in_planes, self.filters, self.widths, head=self.head, max_batch=self.max_batch)
        self.set_weights(weights)
        if self.head == nn.POLICY_HEAD and getattr(self, "loss", None) == "binary_crossentropy":
            self.net.set_loss(1)
        self._trainer = None

    # ------------------------------------------------------------------ Keras surface
    @property
    def input_shape(self):
        return (None, self.in_planes, self.board, self.board)

    @property
    def output_shape(self):
        return (None, self.board * self.board) if self.head == nn.POLICY_HEAD else (None, 1)

    def to_json(self):
        return json.dumps({"class_name": "Sequential", "config": self.layers})

    def _flip(self, w):
        return np.ascontiguousarray(w[:, :, ::-1, ::-1]) if self.flip_kernels else w

    def get_weights(self):
        out = []
        for l in range(self.net.n_conv):
            w, b = self.net.get_conv(l)
            out += [self._flip(w), b]
        hw, hb = self.net.get_head(0)
        out += [hw.reshape(1, -1, 1, 1), hb]
        if self.head == nn.VALUE_HEAD:
            w1, b1 = self.net.get_head(2)
            w2, b2 = self.net.get_head(3)
            out += [np.ascontiguousarray(w1.T), b1, w2.reshape(256, 1), b2]  # Keras Dense stores (input_dim, output_dim)
        return out

    def set_weights(self, weights):
        weights = [np.asarray(w, np.float32) for w in weights]
        need = 2 * (self.net.n_conv + 1) + (4 if self.head == nn.VALUE_HEAD else 0)
        if len(weights) != need:
            raise ValueError("expected %d weight arrays, got %d" % (need, len(weights)))
        for l in range(self.net.n_conv):
            self.net.set_conv(l, self._flip(weights[2 * l]), weights[2 * l + 1])
        k = 2 * self.net.n_conv
        self.net.set_head(0, weights[k].reshape(-1), weights[k + 1].reshape(1))
        if self.head == nn.VALUE_HEAD:
            self.net.set_head(2, np.ascontiguousarray(weights[k + 2].T), weights[k + 3])
            self.net.set_head(3, weights[k + 4].reshape(-1), weights[k + 5].reshape(1))

    def save_weights(self, path, overwrite=True):
        from . import weights_io
        weights_io.save(path, self.layers, self.get_weights())

    def load_weights(self, path):
        from . import weights_io
        self.set_weights(weights_io.load(path, self.layers))

    def predict(self, x, batch_size=None):
        return self.forward(x)

    def forward(self, x):
        """numpy (N, C, S, S) of 0/1 planes (any dtype) -> numpy float32 (N, S*S) / (N, 1)"""
        import torch
        x = np.asarray(x)
        self.ensure_batch(len(x))
        planes = torch.from_numpy(np.ascontiguousarray(x, dtype=np.uint8)).cuda()
        return self.forward_device(planes).cpu().numpy()

    def forward_device(self, planes):
        self.ensure_batch(planes.shape[0])
        out = self.net.forward(planes)
        return out if self.head == nn.POLICY_HEAD else out.reshape(-1, 1)

    def compile(self, loss=None, optimizer=None, metrics=None):
        """Keras' compile as far as the two trainers use it: loss 'categorical_crossentropy' (spt.py:200, the default) or
        'binary_crossentropy' (rpt.py:109), an SGD-like optimizer object with .lr/.momentum/.decay"""
        if loss not in (None, "categorical_crossentropy", "binary_crossentropy"):
            raise ValueError("unsupported loss %r" % (loss,))
        self.loss = loss
        if self.head == nn.POLICY_HEAD:
            self.net.set_loss(1 if loss == "binary_crossentropy" else 0)
        if optimizer is None or isinstance(optimizer, str):
            optimizer = _SGD()
        self.optimizer = optimizer

    def fit(self, X, Y, nb_epoch=1, batch_size=32, verbose=0):
        """SGD on (X planes, Y one-hot moves) -- the call the RL trainer makes (rpt.py:102).  The compiled loss against the
        arg-max of each Y row; a negative optimizer.lr flips the update as rpt.py:95-101 intends."""
        import torch
        from .training import SLTrainer
        if self.optimizer is None:
            self.compile()
        on_device = torch.is_tensor(X) and X.is_cuda   # selfplay.make_training_pairs_device keeps the planes on the device
        X = X.contiguous() if on_device else np.ascontiguousarray(X, dtype=np.uint8)
        Y = np.asarray(Y)
        # one-hot rows (the reference's form) or the move indices themselves
        labels = np.ascontiguousarray(Y if Y.ndim == 1 else Y.reshape(len(X), -1).argmax(1), dtype=np.int32)
        self.ensure_batch(min(batch_size, len(X)))
        if self._trainer is None:
            self._trainer = SLTrainer(self.net, lr=self.optimizer.lr.get_value(), decay=self.optimizer.decay, momentum=self.optimizer.momentum)
        tr = self._trainer
        for _ in range(nb_epoch):
            for s in range(0, len(X), batch_size):
                tr.lr0 = self.optimizer.lr.get_value()
                xb = X[s:s + batch_size] if on_device else torch.from_numpy(X[s:s + batch_size]).cuda()
                tr.step(xb, torch.from_numpy(labels[s:s + batch_size]).cuda())
        loss, acc = tr.pop_metrics(len(X) * nb_epoch)
        return {"loss": loss, "acc": acc}


def model_from_json(text, max_batch=64):
    spec = json.loads(text)
    if spec.get("class_name") != "Sequential":
        raise ValueError("not a Keras Sequential model")
    return SequentialModel(spec["config"], max_batch=max_batch)


class CNNPolicy(object):
    """uses a convolutional neural network to evaluate the state of the game and compute a probability
    distribution over the next action (policy.py:12-24)"""

    def __init__(self, feature_list, **kwargs):
        self.preprocessor = Preprocess(feature_list)
        kwargs["input_dim"] = self.preprocessor.output_dim
        self.model = self.create_network(**kwargs)
        self.forward = self._model_forward()

    def _model_forward(self):
        """batch of inputs (N, F, S, S) -> (N, S*S) move distributions, flattened as util.flatten_idx (policy.py:26-42)"""
        return lambda inpt: self.model.forward(inpt)

    def _select_moves_and_normalize(self, nn_output, moves, size):
        """normalise the distribution over the given moves -> list of (move, prob) (policy.py:44-54)"""
        if len(moves) == 0:
            return []
        move_indices = [flatten_idx(m, size) for m in moves]
        distribution = nn_output[move_indices]
        distribution = distribution / distribution.sum()
        return list(zip(moves, distribution))

    def batch_eval_state(self, states, moves_lists=None):
        """evaluate a list of states at once (policy.py:56-79): planes are encoded and consumed on the
        device; only the (N, S*S) probabilities come back to the host"""
        n_states = len(states)
        if n_states == 0:
            return []
        state_size = states[0].size
        if not all([st.size == state_size for st in states]):
            raise ValueError("all states must have the same size")
        planes = self.preprocessor.states_to_tensor(states)
        network_output = self.model.forward_device(planes).cpu().numpy()
        moves_lists = moves_lists or [st.get_legal_moves() for st in states]
        results = [None] * n_states
        for i in range(n_states):
            results[i] = self._select_moves_and_normalize(network_output[i], moves_lists[i], state_size)
        return results

    def eval_state(self, state, moves=None):
        """list of (action, probability) pairs for one state (policy.py:81-91)"""
        planes = self.preprocessor.states_to_tensor([state])
        network_output = self.model.forward_device(planes).cpu().numpy()
        moves = moves or state.get_legal_moves()
        return self._select_moves_and_normalize(network_output[0], moves, state.size)

    @staticmethod
    def _params(kwargs):
        params = {"board": 19, "filters_per_layer": 128, "layers": 12, "filter_width_1": 5}  # policy.py:106-111
        params.update(kwargs)
        widths = [params["filter_width_1"]] + [params.get("filter_width_%d" % i, 3) for i in range(2, params["layers"] + 1)]
        for w in widths:
            if w % 2 != 1:
                raise ValueError("filter widths must be odd")
        return params, widths

    @staticmethod
    def create_network(**kwargs):
        """input_dim (required), board=19, filters_per_layer=128, layers=12, filter_width_K (policy.py:93-156)"""
        params, widths = CNNPolicy._params(kwargs)
        spec = SequentialModel.build_policy(params["input_dim"], params["board"], params["filters_per_layer"], params["layers"], widths)
        model = SequentialModel(spec, max_batch=params.get("max_batch", 64))
        model.net.init_uniform(seed=params.get("seed", np.random.randint(1 << 30)))
        return model

    @classmethod
    def load_model(cls, json_file):
        """new object from the architecture in json_file (policy.py:158-169)"""
        with open(json_file, "r") as f:
            object_specs = json.load(f)
        new = cls.__new__(cls)
        new.preprocessor = Preprocess(object_specs["feature_list"])
        new.model = model_from_json(object_specs["keras_model"])
        if new.model.in_planes != new.preprocessor.output_dim:
            raise ValueError("model expects %d planes, feature_list gives %d" % (new.model.in_planes, new.preprocessor.output_dim))
        new.model.net.init_uniform(seed=np.random.randint(1 << 30))
        if "weights_file" in object_specs:
            new.model.load_weights(object_specs["weights_file"])
        new.forward = new._model_forward()
        return new

    def save_model(self, json_file, weights_file=None):
        """write the network model and preprocessing features (policy.py:171-193)"""
        object_specs = {"keras_model": self.model.to_json(), "feature_list": self.preprocessor.feature_list}
        if weights_file is not None:
            self.model.save_weights(weights_file)
            object_specs["weights_file"] = weights_file
        with open(json_file, "w") as f:
            json.dump(object_specs, f)


class CNNValue(CNNPolicy):
    """Value network: the conv stack of value.py:14-28 with Flatten + Dense(256) + Dense(1, tanh); evaluates to a
    scalar in [-1, 1] from the side to move's point of view, the contract MCTS expects of its value function
    (AlphaGo/mcts.py:71,118).  The reference only has an untrained stub (value.py:12-39); this class gives it the
    same object surface as CNNPolicy."""

    @staticmethod
    def create_network(**kwargs):
        kwargs.setdefault("filters_per_layer", 152)  # value.py:7 K = 152
        params, widths = CNNPolicy._params(kwargs)
        spec = SequentialModel.build_value(params["input_dim"], params["board"], params["filters_per_layer"], params["layers"], widths)
        model = SequentialModel(spec, max_batch=params.get("max_batch", 64))
        model.net.init_uniform(seed=params.get("seed", np.random.randint(1 << 30)))
        return model

    def batch_eval_state(self, states, moves_lists=None):
        if len(states) == 0:
            return []
        if not all([st.size == states[0].size for st in states]):
            raise ValueError("all states must have the same size")
        planes = self.preprocessor.states_to_tensor(states)
        return [float(v) for v in self.model.forward_device(planes).cpu().numpy().reshape(-1)]

    def eval_state(self, state, moves=None):
        return self.batch_eval_state([state])[0]
