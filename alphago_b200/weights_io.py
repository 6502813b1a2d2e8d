"""Weight files for `SequentialModel.save_weights / load_weights` (policy.py:165-167,188-190;
ModelCheckpoint "weights.{epoch:05d}.hdf5", supervised_policy_trainer.py:161-162).

The file is HDF5 in the layout of Keras 1.0's `save_weights`, so files move between this package and the reference:
root attribute `layer_names` (every layer of the Sequential, in order, the weightless Flatten / Activation included,
because Keras 1.0's loader wants one entry per layer), one group per layer with the attribute `weight_names`, and one
float32 dataset per array (`<layer>_W`, `<layer>_b`).  It is written through h5py where that is importable and through
`hdf5_lite`'s writer otherwise (h5py is not in this image).  `load` goes by names, not by position, so files that hold
only the weight-owning layers load too; so do the `.npz` archives earlier versions of this package wrote under the same
file names (the first bytes decide, not the suffix).
"""
import zipfile

import numpy as np

from . import hdf5_lite

HDF5_MAGIC = hdf5_lite.MAGIC


def _layer_table(layers):
    """[(layer name, [weight names])] for every layer, in Keras order (weightless layers have an empty list)"""
    out = []
    for l in layers:
        n = l["config"]["name"]
        out.append((n, [n + "_W", n + "_b"] if l["class_name"] in ("Convolution2D", "Dense") else []))
    return out


def have_h5py():
    try:
        import h5py  # noqa: F401
        return True
    except ImportError:
        return False


def save(path, layers, weights):
    table = _layer_table(layers)
    assert len(weights) == sum(len(wn) for _, wn in table)
    k = 0
    with hdf5_lite.open_file(path, "w") as f:
        f.attrs["layer_names"] = [n.encode("utf8") for n, _ in table]
        for n, wn in table:
            g = f.create_group(n)
            g.attrs["weight_names"] = [w.encode("utf8") for w in wn]
            for w in wn:
                g.create_dataset(w, data=np.asarray(weights[k], np.float32))
                k += 1


def load(path, layers):
    names = [(n, wn) for n, wn in _layer_table(layers) if wn]
    with open(path, "rb") as f:
        magic = f.read(8)
    out = []
    if magic == HDF5_MAGIC:
        f = hdf5_lite.open_file(path, "r")
        try:
            for n, wn in names:
                for w in wn:
                    out.append(np.asarray(f[n][w][()], np.float32))
        finally:
            f.close()
        return out
    if not zipfile.is_zipfile(path):
        raise ValueError("%s is neither an HDF5 nor an .npz weight file" % path)
    with np.load(path) as z:
        for n, wn in names:
            for w in wn:
                out.append(np.asarray(z[n + "/" + w], np.float32))
    return out
