"""CPU tests of hdf5_lite's writer (datasets and Keras-1.0 weight files without h5py).

There is no libhdf5 in this image, so the files are checked by hdf5_lite's own READER, which is pinned on the
reference's real h5py-written file (tests/test_oracle.py::test_hdf5_lite_reads_reference_file), and by walking the raw
bytes for the invariants libhdf5 insists on (sizes, alignment, sorted symbol tables, end-of-file address)."""
import os
import struct

import numpy as np
import pytest

from alphago_b200 import hdf5_lite as h5
from alphago_b200 import weights_io


def _planes(n, f=12, seed=0):
    rng = np.random.RandomState(seed)
    return (rng.rand(n, f, 19, 19) < 0.1).astype(np.uint8)


def test_datasets_groups_and_attributes_round_trip(tmp_path):
    path = str(tmp_path / "t.h5")
    rng = np.random.RandomState(1)
    arrays = {"u1": rng.randint(0, 255, (7, 3)).astype(np.uint8), "i2": rng.randint(-9, 9, (5,)).astype(np.int16),
              "i8": rng.randint(-9, 9, (2, 2, 2)).astype(np.int64), "f4": rng.randn(3, 4).astype(np.float32),
              "f8": rng.randn(6), "s": np.asarray([b"ab", b"cdef"]), "empty": np.zeros((0, 2), np.uint8)}
    with h5.File(path, "w") as f:
        for k, v in arrays.items():
            f.create_dataset(k, data=v)
        f["scalar"] = np.float32(2.5)
        f["nested/deeper/list"] = [3, 4]           # python ints -> int64, like h5py
        f.create_group("void")
        f.attrs["layer_names"] = [b"convolution2d_1", b"dense_12"]
        f["f4"].attrs["n"] = 3
        f["nested"].attrs["note"] = "text"
        with pytest.raises(ValueError):
            f.create_dataset("u1", data=[1])
        with pytest.raises(ValueError):
            f.create_group("void")
        assert f.require_group("void") is f["void"] and "nested/deeper" in f and "nope" not in f
    r = h5.File(path)
    assert r.keys() == sorted(list(arrays) + ["scalar", "nested", "void"])
    for k, v in arrays.items():
        assert r[k].dtype == v.dtype and r[k].shape == v.shape and np.array_equal(r[k][()] if v.size else r[k][:], v), k
    assert r["scalar"].shape == () and r["scalar"][()] == np.float32(2.5)
    assert r["nested/deeper/list"][:].tolist() == [3, 4] and r["nested/deeper/list"].dtype == np.int64
    assert r["void"].keys() == []
    assert r.attrs["layer_names"].tolist() == [b"convolution2d_1", b"dense_12"]
    assert r["f4"].attrs["n"] == 3 and r["nested"].attrs["note"] == b"text"


def test_many_links_build_a_multi_level_group_tree(tmp_path):
    """8 symbols per node, 32 nodes per B-tree node: 700 links need two B-tree levels (a KGS conversion has thousands)"""
    path = str(tmp_path / "g.h5")
    names = ["games:2016:%04d-x.sgf" % (i * 7919 % 700) for i in range(700)]
    with h5.File(path, "w") as f:
        g = f.require_group("file_offsets")
        for i, n in enumerate(names):
            g[n] = [i * 3, 3]
    g = h5.File(path)["file_offsets"]
    assert g.keys() == sorted(names)
    assert all(g[n][:].tolist() == [i * 3, 3] for i, n in enumerate(names))


def test_chunked_compressed_and_streamed_like_the_reference_converter(tmp_path):
    """game_converter.py:71-105: require_dataset with one row, resize as rows arrive, assign row by row"""
    path = str(tmp_path / "d.h5")
    states, acts = _planes(300), np.random.RandomState(2).randint(0, 19, (300, 2)).astype(np.uint8)
    with h5.File(path, "w") as f:
        st = f.require_dataset("states", dtype=np.uint8, shape=(1, 12, 19, 19), maxshape=(None, 12, 19, 19), exact=False,
                               chunks=(64, 12, 19, 19), compression="lzf")
        ac = f.require_dataset("actions", dtype=np.uint8, shape=(1, 2), maxshape=(None, 2), exact=False, chunks=(1024, 2), compression="lzf")
        assert f.require_dataset("states", dtype=np.uint8, shape=(1, 12, 19, 19)) is st
        for i in range(100):
            if i >= len(st):
                st.resize((i + 1, 12, 19, 19))
                ac.resize((i + 1, 2))
            st[i] = states[i:i + 1].astype(np.float64)   # state_to_tensor hands over (1, F, s, s) float64
            ac[i] = (int(acts[i][0]), int(acts[i][1]))
        st.resize(300, axis=0)
        ac.resize((300, 2))
        st[100:300] = states[100:]
        ac[100:] = acts[100:]
        with pytest.raises(NotImplementedError):
            st[5] = states[5]                             # append-only
        with pytest.raises(IndexError):
            st[300] = states[0]                           # beyond the current size: resize first
    r = h5.File(path)
    s = r["states"]
    assert s.shape == (300, 12, 19, 19) and s.chunk == (64, 12, 19, 19) and s.filters == [32000] and s.dtype == np.uint8   # real LZF, as the reference
    assert np.array_equal(s[:], states) and np.array_equal(r["actions"][:], acts)
    assert np.array_equal(s[[5, 299, 64]], states[[5, 299, 64]]) and np.array_equal(s[60:70], states[60:70])
    assert os.path.getsize(path) < states.nbytes // 2    # the chunks really are compressed


def test_chunk_tree_with_more_than_one_node_and_nd_chunks(tmp_path):
    path = str(tmp_path / "c.h5")
    a = np.arange(1000 * 6, dtype=np.int32).reshape(1000, 6)
    grid = np.arange(35 * 11, dtype=np.float32).reshape(35, 11)
    with h5.File(path, "w") as f:
        f.create_dataset("rows", data=a, chunks=(4, 6), maxshape=(None, 6))            # 250 chunks: 4 leaves under a root
        f.create_dataset("grid", data=grid, chunks=(8, 4), compression="gzip", compression_opts=6)
    r = h5.File(path)
    assert np.array_equal(r["rows"][:], a) and np.array_equal(r["rows"][997:], a[997:]) and r["rows"].filters == []
    assert np.array_equal(r["grid"][:], grid)


def test_lzf_codec_and_chunks_it_cannot_shrink(tmp_path):
    """the host LZF codec (csrc/lzf.cu) against the pure-Python decoder kept in hdf5_lite as the format's statement, and
    the optional-filter rule: a chunk LZF cannot shrink is stored raw with its filter-mask bit set"""
    rng = np.random.RandomState(3)
    samples = [_planes(8).tobytes(), bytes(70000), b"abcabcabcabcabc", b"the quick brown fox " * 999,
               bytes(rng.randint(0, 4, 50000).astype(np.uint8)), b"ab" * 5000 + rng.bytes(300) + b"ab" * 5000]
    for raw in samples:
        enc = h5.lzf_encode(raw)
        assert enc is not None and len(enc) < len(raw)
        assert h5.lzf_decode(enc, len(raw)) == raw and h5.lzf_decompress(enc, len(raw)) == raw
        with pytest.raises(ValueError):
            h5.lzf_decode(enc[:-1], len(raw))
    assert h5.lzf_encode(rng.bytes(4096)) is None        # noise does not compress
    path = str(tmp_path / "raw.h5")
    noise = np.frombuffer(rng.bytes(3 * 1024), np.uint8).reshape(3, 1024)
    mixed = np.concatenate([noise, np.zeros((3, 1024), np.uint8)])
    with h5.File(path, "w") as f:
        f.create_dataset("mixed", data=mixed, chunks=(1, 1024), compression="lzf")
    d = h5.File(path)["mixed"]
    assert d.filters == [32000] and np.array_equal(d[:], mixed)
    d._chunk((0, 0))
    masks = [d._chunks[(i, 0)][2] for i in range(6)]
    sizes = [d._chunks[(i, 0)][1] for i in range(6)]
    assert masks == [1, 1, 1, 0, 0, 0] and sizes[:3] == [1024] * 3 and max(sizes[3:]) < 64


def _walk_headers(b, addr, seen):
    """every object header reachable from `addr`: checks the invariants libhdf5 asserts when it reads one"""
    ver, _, nmsg, refs, size = struct.unpack_from("<BBHII", b, addr)
    assert ver == 1 and refs == 1 and addr % 8 == 0 and size % 8 == 0
    p, end, n, stab = addr + 16, addr + 16 + size, 0, None
    while p < end:
        mtype, msize, _ = struct.unpack_from("<HHB", b, p)
        assert msize % 8 == 0
        if mtype == 0x0011:
            stab = struct.unpack_from("<QQ", b, p + 8)
        p += 8 + msize
        n += 1
    assert p == end and n == nmsg
    seen.append(addr)
    if stab is None:
        return
    btree, heap = stab
    assert b[heap:heap + 4] == b"HEAP" and b[btree:btree + 4] == b"TREE"
    hsize, hfree, hdata = struct.unpack_from("<QQQ", b, heap + 8)
    assert hfree == 1 and hsize % 8 == 0 and hdata + hsize <= len(b)   # 1 = libhdf5's "no free block"

    def name(off):
        return bytes(b[hdata + off:b.index(b"\0", hdata + off)])

    def tree(node, lo):
        assert b[node:node + 4] == b"TREE" and b[node + 4] == 0
        level, used = b[node + 5], struct.unpack_from("<H", b, node + 6)[0]
        assert used <= 32 and node + 544 <= len(b)
        names, p = [], node + 24
        assert name(struct.unpack_from("<Q", b, p)[0]) == lo
        for _ in range(used):
            child, key = struct.unpack_from("<QQ", b, p + 8)
            if level:
                names += tree(child, lo)
            else:
                assert b[child:child + 4] == b"SNOD" and child + 328 <= len(b)
                cnt = struct.unpack_from("<H", b, child + 6)[0]
                assert 1 <= cnt <= 8
                for i in range(cnt):
                    noff, oaddr, cache = struct.unpack_from("<QQI", b, child + 8 + 40 * i)
                    names.append(name(noff))
                    _walk_headers(b, oaddr, seen)
                    assert cache in (0, 1)
            assert names[-1] == name(key)      # the key right of a child is the largest name below it
            lo = name(key)
            p += 16
        return names

    names = tree(btree, b"")
    assert names == sorted(names) and len(set(names)) == len(names)


def test_file_structure_invariants(tmp_path):
    path = str(tmp_path / "s.h5")
    with h5.File(path, "w") as f:
        f.create_dataset("states", data=_planes(130), maxshape=(None, 12, 19, 19), chunks=(64, 12, 19, 19), compression="lzf")
        g = f.require_group("file_offsets")
        for i in range(300):
            g["n%03d" % i] = [i, 1]
        f.create_group("a/b/c")
    b = open(path, "rb").read()
    assert b[:8] == h5.MAGIC and b[8:16] == bytes([0, 0, 0, 0, 0, 8, 8, 0])
    leaf_k, node_k = struct.unpack_from("<HH", b, 16)
    assert (leaf_k, node_k) == (4, 16)
    base, free, eof, driver = struct.unpack_from("<QQQQ", b, 24)
    assert (base, free, driver) == (0, h5.UNDEF, h5.UNDEF) and eof == len(b)   # libhdf5 refuses a file shorter than this
    _, root, cache, _, btree, heap = struct.unpack_from("<QQIIQQ", b, 56)
    seen = []
    _walk_headers(b, root, seen)
    assert cache == 1 and len(seen) == 1 + 1 + 1 + 300 + 3


def test_keras_weight_file_layout(tmp_path):
    """Keras 1.0 save_weights: `layer_names` names every layer of the Sequential, weightless ones have an empty
    `weight_names`; datasets are float32 named <layer>_W / <layer>_b"""
    from alphago_b200.models import SequentialModel
    spec = SequentialModel.build_policy(12, 19, 16, 2, [5, 3])
    rng = np.random.RandomState(0)
    ws = [rng.randn(16, 12, 5, 5), rng.randn(16), rng.randn(16, 16, 3, 3), rng.randn(16), rng.randn(1, 16, 1, 1), rng.randn(1)]
    path = str(tmp_path / "weights.00003.hdf5")
    weights_io.save(path, spec, ws)
    with open(path, "rb") as f:
        assert f.read(8) == h5.MAGIC
    r = h5.File(path)
    layers = [n.decode() for n in r.attrs["layer_names"]]
    assert layers == [l["config"]["name"] for l in spec] and layers[-2:] == ["flatten_1", "activation_1"]
    assert len(r["flatten_1"].attrs["weight_names"]) == 0 and r["flatten_1"].keys() == []
    assert r["convolution2d_2"].attrs["weight_names"].tolist() == [b"convolution2d_2_W", b"convolution2d_2_b"]
    assert r["convolution2d_2"]["convolution2d_2_W"].dtype == np.float32 and r["convolution2d_2"]["convolution2d_2_W"].shape == (16, 16, 3, 3)
    back = weights_io.load(path, spec)
    assert all(np.array_equal(np.asarray(a, np.float32), b) for a, b in zip(ws, back))
    # archives written by earlier versions under the same kind of name still load
    old = str(tmp_path / "weights.00001.hdf5")
    names = [l["config"]["name"] for l in spec if l["class_name"] == "Convolution2D"]
    with open(old, "wb") as f:
        np.savez(f, **{"%s/%s_%s" % (n, n, s): np.asarray(ws[2 * i + j], np.float32) for i, n in enumerate(names) for j, s in enumerate("Wb")})
    assert all(np.array_equal(a, b) for a, b in zip(back, weights_io.load(old, spec)))


def test_converter_writes_the_reference_layout(golden_dir, tmp_path, monkeypatch):
    """sgfs_to_hdf5 end to end on the host: the five AlphaGo - Lee Sedol games -> a real HDF5 file equal, member by member,
    to the reference converter's own output (tests/golden/ref_hdf5_fixture.npz).  The encoder is replaced by the CPU
    oracle here; tests/test_converter_gpu.py runs the same call with the CUDA encoder."""
    from oracle import go_oracle as og
    from test_oracle import load_hdf5_fixture
    from alphago_b200.game_converter import game_converter
    from alphago_b200.supervised_policy_trainer import open_dataset
    states, actions, offsets, feats = load_hdf5_fixture(golden_dir)
    g = np.load(os.path.join(golden_dir, "ref_hdf5_fixture.npz"))
    files = [os.path.join(golden_dir, "sgf", str(n).split("::")[-1]) for n in g["names"]]

    def oracle_encode(self, packed, bd_size, bits=False):
        board = np.stack([p[0] for p in packed]); age = np.stack([p[1] for p in packed])
        player = np.asarray([p[2] for p in packed], np.int8); ko = np.asarray([p[3] for p in packed], np.int16)
        planes = og.encode_batch(board, age, player, ko, bd_size, feats)
        return np.packbits(planes.reshape(len(packed), -1), axis=1) if bits else planes

    monkeypatch.setattr(game_converter, "encode_packed_positions", oracle_encode)
    conv = game_converter.__new__(game_converter)
    conv.n_features, conv.batch = 12, 300
    for container in ("hdf5", "npz"):
        out = str(tmp_path / ("features." + container))
        conv.sgfs_to_hdf5(files, out, bd_size=19, container=container)
        assert not os.path.exists(str(tmp_path / (".tmp.features." + container)))
        s, a = open_dataset(out)
        assert np.array_equal(np.asarray(s[:]), states) and np.array_equal(np.asarray(a[:]), actions)
    f = h5.File(str(tmp_path / "features.hdf5"))
    assert f.keys() == ["actions", "file_offsets", "states"]            # the reference file's members, no more
    assert f["states"].chunk == (64, 12, 19, 19) and f["actions"].chunk == (1024, 2) and f["states"].dtype == np.uint8
    got = {k: f["file_offsets"][k][:].tolist() for k in f["file_offsets"].keys()}
    assert [got[p.replace("/", ":")] for p in files] == [[int(o[0]), int(o[1])] for o in offsets]


def test_round_trip_property(tmp_path):
    """random shapes, dtypes, chunk shapes and filters: what the writer stores is what the reader returns"""
    from hypothesis import given, settings, strategies as st, HealthCheck

    dtypes = [np.uint8, np.int8, np.int16, np.uint16, np.int32, np.int64, np.float32, np.float64]
    counter = {"n": 0}

    @settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
    @given(st.integers(1, 3).flatmap(lambda nd: st.tuples(st.lists(st.integers(0, 13), min_size=nd, max_size=nd),
                                                         st.lists(st.integers(1, 6), min_size=nd, max_size=nd))),
           st.sampled_from(dtypes), st.sampled_from([None, "lzf", "gzip"]), st.booleans(), st.integers(0, 2 ** 31 - 1))
    def check(shape_chunks, dtype, compression, chunked, seed):
        shape, chunks = tuple(shape_chunks[0]), tuple(shape_chunks[1])
        rng = np.random.RandomState(seed)
        data = (rng.randint(0, 3, shape) * rng.randint(0, 2, shape)).astype(dtype)   # sparse: compressible
        counter["n"] += 1
        path = str(tmp_path / ("p%d.h5" % counter["n"]))
        kw = {}
        if chunked or compression:
            kw = dict(chunks=chunks, compression=compression, maxshape=(None,) + shape[1:])
        with h5.File(path, "w") as f:
            f.create_dataset("grp/d", data=data, **kw)
            f["grp"].attrs["shape"] = list(shape)
        r = h5.File(path)
        d = r["grp/d"]
        assert d.shape == shape and d.dtype == np.dtype(dtype)
        assert np.array_equal(d[:], data)
        if shape[0] > 0:
            k = seed % shape[0]
            assert np.array_equal(d[k], data[k]) and np.array_equal(d[k:], data[k:])
        assert r["grp"].attrs["shape"].tolist() == list(shape)

    check()


def test_lzf_round_trip_property():
    """whatever the compressor emits, both decoders give the input back; streams never exceed the input"""
    from hypothesis import given, settings, strategies as st

    pieces = st.one_of(st.binary(min_size=0, max_size=40),
                       st.tuples(st.binary(min_size=1, max_size=5), st.integers(1, 400)).map(lambda t: t[0] * t[1]))

    @settings(max_examples=150, deadline=None)
    @given(st.lists(pieces, min_size=0, max_size=12).map(b"".join))
    def check(raw):
        enc = h5.lzf_encode(raw)
        if enc is None:
            return                      # not compressible (or shorter than 4 bytes): stored raw
        assert len(enc) < len(raw)
        assert h5.lzf_decode(enc, len(raw)) == raw and h5.lzf_decompress(enc, len(raw)) == raw

    check()


def test_take_rows_reads_only_what_a_shuffled_minibatch_needs(tmp_path):
    """Dataset.take_rows: rows in any order, with repeats, from LZF chunks (the converter's layout), equal to plain
    indexing; with and without a thread pool; on a mapped file (the reader never loads the whole file)"""
    from concurrent.futures import ThreadPoolExecutor
    rng = np.random.RandomState(3)
    data = (rng.rand(300, 5, 7, 7) < 0.3).astype(np.uint8)
    acts = rng.randint(0, 7, (300, 2)).astype(np.uint8)
    path = os.path.join(str(tmp_path), "d.h5")
    with h5.File(path, "w") as f:
        f.create_dataset("states", data=data, chunks=(64, 5, 7, 7), compression="lzf")
        f.create_dataset("actions", data=acts, chunks=(128, 2), compression="lzf")
        f.create_dataset("plain", data=acts)
    f = h5.File(path, "r")
    assert f._map is not None
    idx = rng.randint(0, 300, 97)
    idx[:3] = [299, 0, 299]
    with ThreadPoolExecutor(4) as pool:
        first = np.empty((len(idx),) + data.shape[1:], data.dtype)   # the very first read goes through the pool (chunk index not built yet)
        assert np.array_equal(h5.File(path, "r")["states"].take_rows(idx, first, pool), data[idx])
        for name, ref in (("states", data), ("actions", acts), ("plain", acts)):
            ds = f[name]
            assert np.array_equal(ds.take_rows(idx), ref[idx])
            out = np.empty((len(idx),) + ref.shape[1:], ref.dtype)
            assert ds.take_rows(idx, out, pool) is out and np.array_equal(out, ref[idx])
    assert f["states"].take_rows([]).shape == (0, 5, 7, 7)
    with pytest.raises(IndexError):
        f["states"].take_rows([300])


def _structure(path):
    """what libhdf5 reads before it touches any data: superblock fields, and per object the header messages that define
    it -- datatype, dataspace, layout, filter pipeline, fill value -- byte for byte (addresses and times left out)"""
    f = h5.File(path, "r")
    b = f.b
    sb = {"version": b[8], "free_space_version": b[9], "root_group_version": b[10], "shared_header_version": b[12],
          "size_of_offsets": b[13], "size_of_lengths": b[14], "group_leaf_k": int.from_bytes(b[16:18], "little"),
          "group_internal_k": int.from_bytes(b[18:20], "little")}
    keep = {0x0001: "dataspace", 0x0003: "datatype", 0x0005: "fill_value", 0x000B: "filter_pipeline"}

    def obj(o):
        out = {"header_version": b[o.addr], "messages": sorted(t for t, _, _ in o.msgs if t not in (0x0000, 0x0010, 0x0012))}
        for t, d, s in o.msgs:
            if t in keep:
                out[keep[t]] = bytes(b[d:d + s]).rstrip(b"\0")
            if t == 0x0008:   # layout: version, class, then (chunked) rank and dims -- not the B-tree address
                ver, cls = b[d], b[d + 1]
                out["layout"] = (ver, cls) if cls != 2 else (ver, cls, b[d + 2], bytes(b[d + 11:d + 11 + 4 * b[d + 2]]))
        return out
    tree = {"/": obj(f.obj)}
    for name in f.keys():
        node = f[name]
        tree[name] = obj(node.obj)
        if isinstance(node, h5.Group):
            for sub in node.keys():
                tree[name + "/" + sub] = obj(node[sub].obj)
    return sb, tree


def test_writer_reproduces_the_structure_of_the_references_h5py_file(tmp_path):
    """The reference's own dataset file (tests/test_data/hdf5/alphago-vs-lee-sedol-features.hdf5, written by h5py 2.6 /
    libhdf5 through its converter) re-written by hdf5_lite with the converter's calls (game_converter.py:71-88,122-125):
    superblock fields and, object by object, the datatype / dataspace / layout / filter-pipeline / fill-value messages must
    come out byte for byte as libhdf5 wrote them, the same message types must be present, and the data must read back
    equal.  No libhdf5 exists in this image, so this is the closest stand-in for opening the file with h5py.  Needs the
    reference tree (build container only)."""
    from oracle import ref_shim
    src = os.path.join(ref_shim.REFERENCE_ROOT, "tests", "test_data", "hdf5", "alphago-vs-lee-sedol-features.hdf5")
    if not os.path.exists(src):
        pytest.skip("reference tree not present")
    ref = h5.File(src, "r")
    states, actions = ref["states"][:], ref["actions"][:]
    path = os.path.join(str(tmp_path), "rewritten.hdf5")
    with h5.File(path, "w") as f:   # the converter's calls
        s = f.require_dataset("states", dtype=np.uint8, shape=(1, 12, 19, 19), maxshape=(None, 12, 19, 19), exact=False,
                              chunks=(64, 12, 19, 19), compression="lzf")
        a = f.require_dataset("actions", dtype=np.uint8, shape=(1, 2), maxshape=(None, 2), exact=False, chunks=(1024, 2),
                              compression="lzf")
        offs = f.require_group("file_offsets")
        s.resize((len(states), 12, 19, 19)); a.resize((len(actions), 2))
        for i in range(len(states)):
            s[i] = states[i]; a[i] = actions[i]
        for name in ref["file_offsets"].keys():
            offs[name] = [int(v) for v in ref["file_offsets"][name][:]]
    got = h5.File(path, "r")
    assert np.array_equal(got["states"][:], states) and np.array_equal(got["actions"][:], actions)
    sb_ref, tree_ref = _structure(src)
    sb_got, tree_got = _structure(path)
    assert sb_got == sb_ref
    assert sorted(tree_got) == sorted(tree_ref)
    for name in sorted(tree_ref):
        for key in sorted(tree_ref[name]):
            assert tree_got[name].get(key) == tree_ref[name][key], (name, key, tree_got[name].get(key), tree_ref[name][key])
