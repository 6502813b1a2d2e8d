"""`run_training`: drop-in for AlphaGo/training/supervised_policy_trainer.py:83-213.

Same command line (positional `model train_data out_directory`; --minibatch/-B, --epochs/-E, --epoch-length/-l,
--learning-rate/-r, --decay/-d, --verbose/-v, --weights, --train-val-test), same files in `out_directory`
(`metadata.json` with {"epochs": [{loss, acc, val_loss, val_acc}...], "best_epoch", "training_data",
"model_file"}, `shuffle.npz`, `weights.NNNNN.hdf5`), same data order semantics: one shuffled index list split into
train/val/test, an endless generator over the train indices that picks one of the 8 board symmetries per sample
(spt.py:18-40).  What changes is where the work happens: minibatch assembly is a gather kernel and the step is the
CUDA network (`SLTrainer`).  A dataset that fits the device budget (--device-dataset-mb, default 16 GiB) is copied to
device memory once; a larger one STREAMS, as the reference's generator does (it reads `state_dataset[data_idx]` sample
by sample and never holds the file, spt.py:27-36): the HDF5 file stays mapped, a worker thread reads and decodes only the
chunks that hold the next minibatch while the GPU runs the current one, and the rows travel through pinned staging
buffers (`training.HostFeeder`).  Both paths see the same minibatches and symmetries and give bitwise the same weights.
Added flags (non-breaking): --momentum, --seed, --device-dataset-mb, --io-threads.
Launched under torchrun, every rank trains on its share of each minibatch and rank 0 writes the files.
"""
import argparse
import json
import os

import numpy as np

from .models import CNNPolicy


class MetadataWriter(object):
    """spt.py:43-68 MetadataWriterCallback"""

    def __init__(self, path):
        self.file = path
        self.metadata = {"epochs": [], "best_epoch": 0}

    def on_epoch_end(self, logs):
        epoch = len(self.metadata["epochs"])
        self.metadata["epochs"].append(logs)
        key = "val_loss" if "val_loss" in logs else "loss"
        best_loss = self.metadata["epochs"][self.metadata["best_epoch"]][key]
        if logs.get(key) < best_loss:
            self.metadata["best_epoch"] = epoch
        with open(self.file, "w") as f:
            json.dump(self.metadata, f)


def open_dataset(path):
    """(states, actions) array-likes from an HDF5 file in the converter's layout (game_converter.py:71-88) or an .npz
    archive with the same two members (what game_converter writes with container="npz", under whatever name it was
    given: the file's first bytes decide, not its suffix)"""
    with open(path, "rb") as f:
        magic = f.read(4)
    if magic[:2] == b"PK":
        z = np.load(path)
        if "states_bits" in z.files:   # the converter's archive: one-hot planes bit-packed
            shape = tuple(int(v) for v in z["states_shape"])
            row = int(np.prod(shape[1:]))
            return np.unpackbits(z["states_bits"], axis=1)[:, :row].reshape(shape), z["actions"]
        return z["states"], z["actions"]
    from . import hdf5_lite
    f = hdf5_lite.open_file(path, "r")
    return f["states"], f["actions"]


def _index_stream(indices, batch):
    """the reference generator's index order: cycle over `indices` forever, `batch` at a time (spt.py:26-40)"""
    pos = 0
    n = len(indices)
    while True:
        take = [indices[(pos + i) % n] for i in range(batch)]
        pos = (pos + batch) % n
        yield np.asarray(take, np.int64)


def run_training(cmd_line_args=None):
    """Run training. command-line args may be passed in as a list"""
    parser = argparse.ArgumentParser(description="Perform supervised training on a policy network.")
    parser.add_argument("model", help="Path to a JSON model file (i.e. from CNNPolicy.save_model())")
    parser.add_argument("train_data", help="A .h5 file of training data")
    parser.add_argument("out_directory", help="directory where metadata and weights will be saved")
    parser.add_argument("--minibatch", "-B", help="Size of training data minibatches. Default: 16", type=int, default=16)
    parser.add_argument("--epochs", "-E", help="Total number of iterations on the data. Default: 10", type=int, default=10)
    parser.add_argument("--epoch-length", "-l", help="Number of training examples considered 'one epoch'. Default: # training data", type=int, default=None)
    parser.add_argument("--learning-rate", "-r", help="Learning rate - how quickly the model learns at first. Default: .03", type=float, default=.03)
    parser.add_argument("--decay", "-d", help="The rate at which learning decreases. Default: .0001", type=float, default=.0001)
    parser.add_argument("--verbose", "-v", help="Turn on verbose mode", default=False, action="store_true")
    parser.add_argument("--weights", help="Name of a .h5 weights file (in the output directory) to load to resume training", default=None)
    parser.add_argument("--train-val-test", help="Fraction of data to use for training/val/test. Must sum to 1. Invalid if restarting training", nargs=3, type=float, default=[0.93, .05, .02])
    parser.add_argument("--momentum", help="SGD momentum. Default: 0 (as the reference)", type=float, default=0.0)
    parser.add_argument("--seed", help="seed for the shuffle and the symmetry choice", type=int, default=None)
    parser.add_argument("--device-dataset-mb", help="Keep the dataset in device memory when it is at most this many MB, else stream it from the file. Default: 16384", type=float, default=16384.0)
    parser.add_argument("--io-threads", help="Threads decoding dataset chunks when streaming. Default: 4", type=int, default=4)
    args = parser.parse_args(cmd_line_args)

    import torch
    import torch.distributed as dist
    from .training import SLTrainer

    world, rank = 1, 0
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ and int(os.environ["WORLD_SIZE"]) > 1:
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        world, rank = dist.get_world_size(), dist.get_rank()
    chief = rank == 0
    say = print if (args.verbose and chief) else (lambda *a, **k: None)
    resume = args.weights is not None
    if resume:
        say("trying to resume from %s with weights %s" % (args.out_directory, os.path.join(args.out_directory, args.weights)))
    elif os.path.exists(args.out_directory):
        say("directory %s exists. any previous data will be overwritten" % args.out_directory)
    else:
        say("starting fresh output directory %s" % args.out_directory)

    model = CNNPolicy.load_model(args.model).model
    if resume:
        model.load_weights(os.path.join(args.out_directory, args.weights))

    states, actions = open_dataset(args.train_data)
    n_total_data = len(states)
    n_train_data = int(args.train_val_test[0] * n_total_data)
    n_val_data = int(args.train_val_test[1] * n_total_data)
    say("datset loaded")
    say("\t%d total samples" % n_total_data)
    say("\t%d training samples" % n_train_data)
    say("\t%d validaion samples" % n_val_data)

    if chief and not os.path.exists(args.out_directory):
        os.makedirs(args.out_directory)
    if world > 1:
        dist.barrier()

    meta_file = os.path.join(args.out_directory, "metadata.json")
    meta_writer = MetadataWriter(meta_file)
    if os.path.exists(meta_file) and resume:
        with open(meta_file, "r") as f:
            meta_writer.metadata = json.load(f)
        say("previous metadata loaded: %d epochs. new epochs will be appended." % len(meta_writer.metadata["epochs"]))
    else:
        say("starting with empty metadata")
    meta_writer.metadata["training_data"] = args.train_data
    meta_writer.metadata["model_file"] = args.model

    rng = np.random.RandomState(args.seed)
    shuffle_file = os.path.join(args.out_directory, "shuffle.npz")
    if os.path.exists(shuffle_file) and resume:
        with open(shuffle_file, "rb") as f:
            shuffle_indices = np.load(f)
        say("loading previous data shuffling indices")
    else:
        shuffle_indices = rng.permutation(n_total_data)
        if world > 1:  # every rank must walk the same order
            t = torch.from_numpy(shuffle_indices).cuda()
            dist.broadcast(t, 0)
            shuffle_indices = t.cpu().numpy()
        if chief:
            with open(shuffle_file, "wb") as f:
                np.save(f, shuffle_indices)
        say("created new data shuffling indices")
    train_indices = shuffle_indices[0:n_train_data]
    val_indices = shuffle_indices[n_train_data:n_train_data + n_val_data]

    shape = tuple(int(v) for v in states.shape)
    if shape[1] != model.in_planes or shape[-1] != model.board:
        raise ValueError("dataset has %d planes on %dx%d boards, the model expects %d on %dx%d" % (
            shape[1], shape[-1], shape[-1], model.in_planes, model.board, model.board))
    if args.minibatch % world != 0:
        raise ValueError("--minibatch %d must be a multiple of the %d ranks" % (args.minibatch, world))
    share = args.minibatch // world  # this rank's part of every minibatch
    model.ensure_batch(share)
    trainer = SLTrainer(model.net, lr=args.learning_rate, decay=args.decay, momentum=args.momentum)
    mine = slice(rank * share, (rank + 1) * share)
    dataset_mb = float(np.prod(shape)) / 2 ** 20
    streaming = dataset_mb > args.device_dataset_mb
    if streaming:
        # the dataset stays where it is (a mapped file or host arrays); rows reach the device minibatch by minibatch
        from concurrent.futures import ThreadPoolExecutor
        from .training import HostFeeder
        say("streaming the dataset (%.0f MB > --device-dataset-mb %.0f)" % (dataset_mb, args.device_dataset_mb))
        feeder = HostFeeder(trainer, states, actions, share, io_threads=max(0, args.io_threads))
        reader = ThreadPoolExecutor(max_workers=1)
    else:
        # dataset resident in device memory (uint8: 17 KB per 48-plane position)
        d_states = torch.from_numpy(np.ascontiguousarray(states[:], dtype=np.uint8)).cuda()
        d_actions = torch.from_numpy(np.ascontiguousarray(actions[:], dtype=np.uint8)).cuda()

    def batches(stream, total):
        """(index, transform) of this rank for minibatch after minibatch until `total` samples have been seen"""
        seen = 0
        while seen < total:
            idx = next(stream)
            tf = rng.randint(0, 8, len(idx))  # np.random.choice(transforms) per sample, spt.py:30
            seen += len(idx)
            yield idx[mine], tf[mine]

    def run(stream, total, train):
        """one pass: returns the number of local samples"""
        local = 0
        if not streaming:
            for idx_r, tf_r in batches(stream, total):
                x, y = trainer.assemble(d_states, d_actions, idx_r, tf_r)
                (trainer.step if train else trainer.evaluate)(x, y)
                local += len(idx_r)
            return local
        it = batches(stream, total)
        nxt = next(it, None)
        pending = reader.submit(feeder.stage, nxt[0]) if nxt is not None else None
        while nxt is not None:
            cur, nxt = nxt, next(it, None)
            pending.result()                       # minibatch `cur` is in its staging buffers
            feeder.step(cur[0], cur[1], train=train)   # queue copy + step ...
            if nxt is not None:                    # ... and read the next one meanwhile
                pending = reader.submit(feeder.stage, nxt[0])
            local += len(cur[0])
        return local

    train_stream = _index_stream(train_indices, args.minibatch)
    val_stream = _index_stream(val_indices, args.minibatch) if n_val_data > 0 else None
    samples_per_epoch = args.epoch_length or n_train_data
    say("STARTING TRAINING")

    for epoch in range(args.epochs):
        # Keras 1.0 fit_generator: batches until samples_per_epoch is reached
        local_seen = run(train_stream, samples_per_epoch, True)
        loss, acc = trainer.pop_metrics(local_seen)
        logs = {"loss": loss, "acc": acc}
        if val_stream is not None:
            vlocal = run(val_stream, n_val_data, False)
            logs["val_loss"], logs["val_acc"] = trainer.pop_metrics(vlocal)
        if chief:
            # ModelCheckpoint numbers files from 0 in every run (spt.py:161-162)
            model.save_weights(os.path.join(args.out_directory, "weights.%05d.hdf5" % epoch))
            meta_writer.on_epoch_end(logs)
        say("epoch %d: %s" % (epoch, json.dumps(logs)))
    if streaming:
        reader.shutdown()
    if world > 1:
        dist.barrier()
    return meta_writer.metadata


if __name__ == "__main__":
    run_training()
