"""Weight files for `SequentialModel.save_weights / load_weights` (policy.py:165-167,188-190;
ModelCheckpoint "weights.{epoch:05d}.hdf5", supervised_policy_trainer.py:161-162).

With h5py importable the file is Keras 1.0's HDF5 layout (root attr `layer_names`; one group per
layer with attr `weight_names` and one dataset per array), so files move between this package and
the reference.  Without h5py (it is not in this image) the SAME path holds a NumPy .npz archive with
the same names; `load` sniffs the magic bytes, so either kind loads wherever it can be read.
"""
import io
import zipfile

import numpy as np

HDF5_MAGIC = b"\x89HDF\r\n\x1a\n"


def _layer_names(layers):
    """[(layer name, [weight names])] for the layers that own weights, in Keras order"""
    out = []
    for l in layers:
        if l["class_name"] in ("Convolution2D", "Dense"):
            n = l["config"]["name"]
            out.append((n, [n + "_W", n + "_b"]))
    return out


def have_h5py():
    try:
        import h5py  # noqa: F401
        return True
    except ImportError:
        return False


def save(path, layers, weights):
    names = _layer_names(layers)
    assert len(weights) == 2 * len(names)
    if have_h5py():
        import h5py
        with h5py.File(path, "w") as f:
            f.attrs["layer_names"] = [n.encode("utf8") for n, _ in names]
            for i, (n, wn) in enumerate(names):
                g = f.create_group(n)
                g.attrs["weight_names"] = [w.encode("utf8") for w in wn]
                for j, w in enumerate(wn):
                    g.create_dataset(w, data=np.asarray(weights[2 * i + j], np.float32))
        return
    arrays = {}
    for i, (n, wn) in enumerate(names):
        for j, w in enumerate(wn):
            arrays[n + "/" + w] = np.asarray(weights[2 * i + j], np.float32)
    buf = io.BytesIO()
    np.savez(buf, **arrays)
    with open(path, "wb") as f:
        f.write(buf.getvalue())


def load(path, layers):
    names = _layer_names(layers)
    with open(path, "rb") as f:
        magic = f.read(8)
    out = []
    if magic == HDF5_MAGIC:
        if not have_h5py():
            raise ImportError("%s is an HDF5 file and h5py is not installed" % path)
        import h5py
        with h5py.File(path, "r") as f:
            for n, wn in names:
                for w in wn:
                    out.append(np.asarray(f[n][w], np.float32))
        return out
    if not zipfile.is_zipfile(path):
        raise ValueError("%s is neither an HDF5 nor an .npz weight file" % path)
    with np.load(path) as z:
        for n, wn in names:
            for w in wn:
                out.append(np.asarray(z[n + "/" + w], np.float32))
    return out
