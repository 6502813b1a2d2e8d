"""Minimal HDF5 for the reference's dataset and weight files, for images without h5py: a reader and a writer.

Reader: what `game_converter.py:67-88` / Keras 1.0 `save_weights` produce with h5py 2.x defaults (HDF5 1.8 "earliest"
format): superblock v0/v1, symbol-table groups (v1 B-tree + local heap), v1 object headers with continuation
blocks, fixed-point / IEEE-float / fixed-length-string datatypes, contiguous, compact and chunked layouts (v1 chunk
B-tree) with the gzip (1), shuffle (2) and LZF (32000) filters, and v1-v3 attribute messages.  That is the on-disk
contract of the SL trainer (`states` uint8 (N,F,19,19) LZF-chunked, `actions` uint8 (N,2), group `file_offsets`;
supervised_policy_trainer.py:128-129).

Writer (`File(path, "w")`, the subset of h5py's API that game_converter.py:67-88,122-125 and Keras 1.0's save_weights
use: create_dataset / create_group / require_group / group[name] = array / .attrs[name] = value): the same "earliest"
format, laid out the way libhdf5 lays out the reference's own fixture file (superblock v0, v1 object headers, one local
heap + v1 B-tree of symbol nodes per group, contiguous datasets, chunked datasets behind a v1 chunk B-tree).  Chunks are
compressed with the filter asked for: `compression="lzf"` (the reference's converter) gives LZF chunks under h5py's
filter id 32000 and client data, through the library's host codec (csrc/lzf.cu; deflate level 1 where the library is not
built); `compression="gzip"` gives deflate (id 1).  A chunk LZF cannot shrink is stored raw with its filter-mask bit
set, as libhdf5 does for an optional filter.  Data goes to the file as it is
handed over, metadata at close().  No libhdf5 is available in this image to cross-check the files: the writer is
checked by this module's reader (which is itself pinned on the reference's real file) in tests/test_hdf5_lite.py.

`open_file()` prefers h5py when it is importable.
"""
import os
import struct
import zlib

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
MAGIC = b"\x89HDF\r\n\x1a\n"


def lzf_decompress(src, out_len):
    """liblzf stream -> bytes (the h5py LZF filter, registered id 32000)"""
    out = bytearray()
    ip, n = 0, len(src)
    while ip < n:
        ctrl = src[ip]; ip += 1
        if ctrl < 32:
            out += src[ip:ip + ctrl + 1]
            ip += ctrl + 1
        else:
            length = ctrl >> 5
            ref = len(out) - ((ctrl & 0x1F) << 8) - 1
            if length == 7:
                length += src[ip]; ip += 1
            ref -= src[ip]; ip += 1
            length += 2
            if ref < 0:
                raise ValueError("corrupt LZF stream")
            dist = len(out) - ref
            if dist >= length:
                out += out[ref:ref + length]
            else:  # overlapping run: the pattern repeats
                pat = bytes(out[ref:])
                out += (pat * (length // dist + 1))[:length]
    if len(out) != out_len:
        raise ValueError("LZF stream decoded to %d bytes, expected %d" % (len(out), out_len))
    return bytes(out)


def _native():
    """libalphago_b200.so's host LZF codec (csrc/lzf.cu), or None where the library is not built: the reader then decodes in
    Python (slow: seconds per MB) and the writer falls back to deflate chunks"""
    try:
        from . import _lib
        return _lib.lib()
    except Exception:
        return None


def lzf_decode(src, out_len):
    """LZF stream -> bytes, natively when the library is there"""
    L = _native()
    if L is None:
        return lzf_decompress(src, out_len)
    import ctypes
    out = ctypes.create_string_buffer(out_len)
    n = L.ag_lzf_decompress(bytes(src), len(src), ctypes.cast(out, ctypes.c_void_p), out_len)
    if n != out_len:
        raise ValueError("LZF stream decoded to %d bytes, expected %d" % (n, out_len))
    return out.raw


def lzf_encode(raw):
    """bytes -> LZF stream, or None when it would not be shorter than `raw` (the chunk is then stored unfiltered) or the
    library is not built"""
    L = _native()
    if L is None or len(raw) < 4:
        return None
    import ctypes
    cap = len(raw) - 1
    out = ctypes.create_string_buffer(cap)
    src = np.frombuffer(raw, np.uint8)
    n = L.ag_lzf_compress(ctypes.c_void_p(src.ctypes.data), len(raw), ctypes.cast(out, ctypes.c_void_p), cap)
    return out.raw[:n] if n > 0 else None


class _Reader(object):
    def __init__(self, buf):
        self.b = buf

    def u(self, off, size):
        return int.from_bytes(self.b[off:off + size], "little")


def _parse_datatype(b, off):
    cls_ver, bf0, bf1, bf2, size = struct.unpack_from("<BBBBI", b, off)
    cls = cls_ver & 0x0F
    order = ">" if (bf0 & 1) else "<"
    if cls == 0:
        kind = "i" if (bf0 & 8) else "u"
        return np.dtype("%s%s%d" % (order if size > 1 else "|", kind, size))
    if cls == 1:
        return np.dtype("%sf%d" % (order, size))
    if cls == 3:
        return np.dtype("S%d" % size)
    raise NotImplementedError("HDF5 datatype class %d" % cls)


def _parse_dataspace(b, off):
    ver, rank, flags = b[off], b[off + 1], b[off + 2]
    p = off + (8 if ver == 1 else 4)
    dims = [int.from_bytes(b[p + 8 * i:p + 8 * i + 8], "little") for i in range(rank)]
    return tuple(dims)


class _Object(object):
    """an object header: list of (type, data offset, size) messages"""

    def __init__(self, f, addr):
        self.f, self.addr = f, addr
        b = f.b
        ver = b[addr]
        if ver != 1:
            raise NotImplementedError("object header version %d (file written with libver='latest'?)" % ver)
        nmsg = int.from_bytes(b[addr + 2:addr + 4], "little")
        hsize = int.from_bytes(b[addr + 8:addr + 12], "little")
        self.msgs = []
        blocks = [(addr + 16, hsize)]
        while blocks and len(self.msgs) < nmsg:
            p, left = blocks.pop(0)
            end = p + left
            while p + 8 <= end and len(self.msgs) < nmsg:
                mtype, msize, mflags = struct.unpack_from("<HHB", b, p)
                data = p + 8
                if mtype == 0x0010:
                    blocks.append((int.from_bytes(b[data:data + 8], "little"), int.from_bytes(b[data + 8:data + 16], "little")))
                self.msgs.append((mtype, data, msize))
                p = data + msize

    def find(self, mtype):
        return [(d, s) for t, d, s in self.msgs if t == mtype]

    def attrs(self):
        out, b = {}, self.f.b
        for d, s in self.find(0x000C):
            ver = b[d]
            nsize, tsize, ssize = struct.unpack_from("<HHH", b, d + 2)
            p = d + 8 + (1 if ver == 3 else 0)
            pad = (lambda v: (v + 7) & ~7) if ver == 1 else (lambda v: v)
            name = bytes(b[p:p + nsize]).split(b"\0")[0].decode("utf8")
            p += pad(nsize)
            dt = _parse_datatype(b, p)
            p += pad(tsize)
            shape = _parse_dataspace(b, p)
            p += pad(ssize)
            count = int(np.prod(shape)) if shape else 1
            arr = np.frombuffer(b, dtype=dt, count=count, offset=p).reshape(shape)
            out[name] = arr.copy() if shape else arr.reshape(())[()]
        return out


class Dataset(object):
    def __init__(self, f, obj, name):
        self.f, self.obj, self.name = f, obj, name
        b = f.b
        self.dtype = _parse_datatype(b, obj.find(0x0003)[0][0])
        self.shape = _parse_dataspace(b, obj.find(0x0001)[0][0])
        d = obj.find(0x0008)[0][0]
        ver, cls = b[d], b[d + 1]
        if ver != 3:
            raise NotImplementedError("data layout message version %d" % ver)
        self.layout = cls
        self.filters = []
        if cls == 1:
            self.data_addr = int.from_bytes(b[d + 2:d + 10], "little")
        elif cls == 0:
            self.data_addr = d + 4
        elif cls == 2:
            nd = b[d + 2]
            self.btree = int.from_bytes(b[d + 3:d + 11], "little")
            self.chunk = tuple(int.from_bytes(b[d + 11 + 4 * i:d + 15 + 4 * i], "little") for i in range(nd - 1))
            fm = obj.find(0x000B)
            if fm:
                p = fm[0][0]
                fver, nf = b[p], b[p + 1]
                p += 8 if fver == 1 else 2
                for _ in range(nf):
                    fid = int.from_bytes(b[p:p + 2], "little")
                    if fver == 1 or fid >= 256:
                        nlen, _fl, ncd = struct.unpack_from("<HHH", b, p + 2)
                        p += 8
                    else:
                        nlen = 0
                        _fl, ncd = struct.unpack_from("<HH", b, p + 2)
                        p += 6
                    p += ((nlen + 7) & ~7) if fver == 1 else nlen
                    p += 4 * ncd + (4 if (fver == 1 and ncd % 2) else 0)
                    self.filters.append(fid)
            self._chunks = None
        self._cache = {}

    def __len__(self):
        return self.shape[0]

    @property
    def attrs(self):
        return self.obj.attrs()

    # ---- chunk index
    def _walk(self, addr, out):
        b = self.f.b
        if bytes(b[addr:addr + 4]) != b"TREE":
            raise ValueError("bad chunk B-tree node")
        level, used = b[addr + 5], int.from_bytes(b[addr + 6:addr + 8], "little")
        nd = len(self.chunk) + 1
        ksz = 8 + 8 * nd
        p = addr + 24
        for _ in range(used):
            csize, fmask = struct.unpack_from("<II", b, p)
            offs = tuple(int.from_bytes(b[p + 8 + 8 * i:p + 16 + 8 * i], "little") for i in range(nd - 1))
            child = int.from_bytes(b[p + ksz:p + ksz + 8], "little")
            if level == 0:
                out[offs] = (child, csize, fmask)
            else:
                self._walk(child, out)
            p += ksz + 8

    def _index(self):
        if self._chunks is None:   # built aside and published whole: take_rows decodes chunks on several threads
            chunks = {}
            if self.btree != UNDEF:
                self._walk(self.btree, chunks)
            self._chunks = chunks
        return self._chunks

    def _chunk(self, offs):
        """one decoded chunk, through a small cache (row-by-row readers come back to the same chunk)"""
        if offs in self._cache:
            return self._cache[offs]
        arr = self._decode(offs)
        if len(self._cache) > 64:
            self._cache.clear()
        self._cache[offs] = arr
        return arr

    def _decode(self, offs):
        self._index()
        n = int(np.prod(self.chunk))
        if offs not in self._chunks:
            arr = np.zeros(self.chunk, self.dtype)
        else:
            addr, csize, fmask = self._chunks[offs]
            raw = bytes(self.f.b[addr:addr + csize])
            for i, fid in reversed(list(enumerate(self.filters))):
                if fmask & (1 << i):
                    continue
                if fid == 32000:
                    raw = lzf_decode(raw, n * self.dtype.itemsize)
                elif fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    raw = np.frombuffer(raw, np.uint8).reshape(self.dtype.itemsize, -1).T.tobytes()
                else:
                    raise NotImplementedError("HDF5 filter %d" % fid)
            arr = np.frombuffer(raw, self.dtype, count=n).reshape(self.chunk)
        return arr

    def take_rows(self, index, out=None, pool=None):
        """rows `index` (any order, repeats allowed) along the first axis, the way a shuffled minibatch reads a dataset
        (supervised_policy_trainer.py:27-36 reads them one `state_dataset[data_idx]` at a time): every chunk that holds
        one of the rows is decoded once per call, the chunks in parallel when a thread pool is given (the LZF / deflate
        decoders release the GIL).  Nothing but those chunks is read from the file."""
        index = np.asarray(index, np.int64)
        if out is None:
            out = np.empty((len(index),) + self.shape[1:], self.dtype)
        if len(index) and (index.min() < 0 or index.max() >= self.shape[0]):
            raise IndexError("row index out of range")
        if self.layout != 2:
            np.take(self._rows(0, self.shape[0]), index, axis=0, out=out)
            return out
        c0 = self.chunk[0]
        order = np.argsort(index // c0, kind="stable")
        firsts, starts = np.unique(index[order] // c0, return_index=True)
        bounds = list(starts) + [len(index)]

        def one(k):
            first = int(firsts[k]) * c0
            rows = self._rows(first, min(first + c0, self.shape[0]), cached=False)
            sel = order[bounds[k]:bounds[k + 1]]
            out[sel] = rows[index[sel] - first]
        if pool is not None and len(firsts) > 1:
            list(pool.map(one, range(len(firsts))))
        else:
            for k in range(len(firsts)):
                one(k)
        return out

    def _rows(self, lo, hi, cached=True):
        """rows [lo, hi) along the first axis"""
        if self.layout != 2:
            count = int(np.prod(self.shape)) if self.shape else 1
            if count == 0 or self.data_addr == UNDEF:   # nothing allocated: empty, or never written (fill value)
                return np.zeros(self.shape, self.dtype)[lo:hi]
            full = np.frombuffer(self.f.b, self.dtype, count=count, offset=self.data_addr).reshape(self.shape)
            return full[lo:hi]
        out = np.zeros((hi - lo,) + self.shape[1:], self.dtype)
        c0 = self.chunk[0]
        grid = [range(0, s, c) for s, c in zip(self.shape[1:], self.chunk[1:])]
        idxs = [()]
        for g in grid:
            idxs = [i + (o,) for i in idxs for o in g]
        keys = [(first,) + rest for first in range(lo // c0 * c0, hi, c0) for rest in idxs]
        if len(keys) > 8:   # a bulk read (the trainer takes the whole dataset): decode on all cores, past the cache
            from concurrent.futures import ThreadPoolExecutor
            with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as pool:
                decoded = dict(zip(keys, pool.map(self._decode, keys)))
        else:
            decoded = None
        for first in range(lo // c0 * c0, hi, c0):
            for rest in idxs:
                ch = decoded.pop((first,) + rest) if decoded is not None else (self._chunk((first,) + rest) if cached else self._decode((first,) + rest))
                a, z = max(lo, first), min(hi, first + c0, self.shape[0])
                sl = (slice(a - first, z - first),) + tuple(slice(0, min(c, s - o)) for c, s, o in zip(self.chunk[1:], self.shape[1:], rest))
                dst = (slice(a - lo, z - lo),) + tuple(slice(o, min(o + c, s)) for c, s, o in zip(self.chunk[1:], self.shape[1:], rest))
                out[dst] = ch[sl]
        return out

    def __getitem__(self, key):
        if isinstance(key, tuple) and key == ():
            return self._rows(0, self.shape[0]) if self.shape else np.frombuffer(self.f.b, self.dtype, 1, self.data_addr)[0]
        if isinstance(key, (int, np.integer)):
            k = int(key) + (self.shape[0] if key < 0 else 0)
            return self._rows(k, k + 1)[0]
        if isinstance(key, slice):
            lo, hi, st = key.indices(self.shape[0])
            return self._rows(lo, max(lo, hi))[::st]
        if isinstance(key, (list, np.ndarray)):
            return np.stack([self[int(k)] for k in key])
        if isinstance(key, tuple):
            return self[key[0]][key[1:]] if not isinstance(key[0], slice) else self[key[0]][(slice(None),) + key[1:]]
        raise TypeError("unsupported index %r" % (key,))


class Group(object):
    def __init__(self, f, obj, name="/"):
        self.f, self.obj, self.name = f, obj, name
        d = obj.find(0x0011)
        if not d:
            raise NotImplementedError("group without a symbol table (new-style links)")
        b = f.b
        btree = int.from_bytes(b[d[0][0]:d[0][0] + 8], "little")
        heap = int.from_bytes(b[d[0][0] + 8:d[0][0] + 16], "little")
        if bytes(b[heap:heap + 4]) != b"HEAP":
            raise ValueError("bad local heap")
        self._heap = int.from_bytes(b[heap + 24:heap + 32], "little")
        self._links = {}
        self._walk(btree)

    def _walk(self, addr):
        b = self.f.b
        sig = bytes(b[addr:addr + 4])
        if sig == b"TREE":
            level, used = b[addr + 5], int.from_bytes(b[addr + 6:addr + 8], "little")
            p = addr + 24 + 8
            for _ in range(used):
                self._walk(int.from_bytes(b[p:p + 8], "little"))
                p += 16
        elif sig == b"SNOD":
            n = int.from_bytes(b[addr + 6:addr + 8], "little")
            p = addr + 8
            for _ in range(n):
                noff = int.from_bytes(b[p:p + 8], "little")
                oaddr = int.from_bytes(b[p + 8:p + 16], "little")
                s = self._heap + noff
                e = s
                while b[e] != 0:
                    e += 1
                self._links[bytes(b[s:e]).decode("utf8")] = oaddr
                p += 40
        else:
            raise ValueError("bad group B-tree node")

    def keys(self):
        return sorted(self._links)

    def __contains__(self, name):
        return name in self._links

    def __iter__(self):
        return iter(self.keys())

    @property
    def attrs(self):
        return self.obj.attrs()

    def __getitem__(self, name):
        node = self
        for part in [p for p in name.split("/") if p]:
            if part not in node._links:
                raise KeyError(name)
            obj = _Object(self.f, node._links[part])
            node = Dataset(self.f, obj, part) if obj.find(0x0008) else Group(self.f, obj, part)
        return node


class _ReadFile(Group):
    def __init__(self, path):
        # mapped, not read: a KGS-scale dataset is hundreds of GB and only the chunks a reader asks for are touched
        import mmap
        with open(path, "rb") as fh:
            size = os.fstat(fh.fileno()).st_size
            self._map = mmap.mmap(fh.fileno(), 0, access=mmap.ACCESS_READ) if size > 0 else None
        buf = memoryview(self._map) if self._map is not None else memoryview(b"")
        if bytes(buf[:8]) != MAGIC:
            raise ValueError("%s is not an HDF5 file" % path)
        ver = buf[8]
        if ver not in (0, 1) or buf[13] != 8 or buf[14] != 8:
            raise NotImplementedError("HDF5 superblock version %d / offset size %d" % (ver, buf[13]))
        self.b = buf
        root_entry = 24 + 32 + (4 if ver == 1 else 0)
        root_addr = int.from_bytes(buf[root_entry + 8:root_entry + 16], "little")
        Group.__init__(self, self, _Object(self, root_addr), "/")

    def close(self):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


# ------------------------------------------------------------------------------------------------ writer
GROUP_LEAF_K, GROUP_NODE_K, CHUNK_NODE_K = 4, 16, 32   # libhdf5 defaults (the superblock v0 stores the first two)


def _pad8(b):
    return b + b"\0" * (-len(b) % 8)


def _enc_datatype(dt):
    dt = np.dtype(dt)
    if dt.kind in "ui" and dt.itemsize in (1, 2, 4, 8) and dt.byteorder in "<|=":
        return struct.pack("<BBBBIHH", 0x10, 0x08 if dt.kind == "i" else 0, 0, 0, dt.itemsize, 0, 8 * dt.itemsize)
    if dt.kind == "f" and dt.itemsize in (4, 8) and dt.byteorder in "<=":
        if dt.itemsize == 4:
            return struct.pack("<BBBBIHHBBBBI", 0x11, 0x20, 31, 0, 4, 0, 32, 23, 8, 0, 23, 127)
        return struct.pack("<BBBBIHHBBBBI", 0x11, 0x20, 63, 0, 8, 0, 64, 52, 11, 0, 52, 1023)
    if dt.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, max(dt.itemsize, 1))   # null-padded ASCII, as h5py writes numpy S arrays
    raise NotImplementedError("cannot write dtype %r" % dt)


def _enc_dataspace(shape, maxshape=None):
    maxshape = shape if maxshape is None else maxshape
    out = struct.pack("<BBBBI", 1, len(shape), 1 if shape else 0, 0, 0)
    out += b"".join(struct.pack("<Q", int(d)) for d in shape)
    if shape:
        out += b"".join(struct.pack("<Q", UNDEF if m is None else int(m)) for m in maxshape)
    return out


def _as_array(value):
    if isinstance(value, str):
        value = value.encode("utf8")
    arr = np.asarray(value)
    if arr.dtype.kind == "U":
        arr = np.char.encode(arr, "utf8")
    if arr.dtype.kind == "O":
        raise NotImplementedError("variable-length data")
    if arr.dtype == np.bool_:
        arr = arr.astype(np.uint8)
    return arr


class _WAttrs(dict):
    def __setitem__(self, name, value):
        dict.__setitem__(self, name, _as_array(value))

    def messages(self):
        out = []
        for name, arr in self.items():
            nm, dt, sp = name.encode("utf8") + b"\0", _enc_datatype(arr.dtype), _enc_dataspace(arr.shape)
            body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(sp)) + _pad8(nm) + _pad8(dt) + _pad8(sp) + np.ascontiguousarray(arr).tobytes()
            if len(body) > 0xFFF0:
                raise NotImplementedError("attribute %r is too large for an object header message" % name)
            out.append((0x000C, 0, body))
        return out


class _WNode(object):
    def __init__(self, f, name):
        self.f, self.name, self.attrs = f, name, _WAttrs()


class WDataset(_WNode):
    """Contiguous datasets and datasets chunked in several dimensions are written when created (`data` is the whole
    array).  Datasets chunked along the first axis only -- the converter's `states` and `actions` -- can also grow the
    way game_converter.py:100-105 fills them: `resize()` and then `ds[i] = row` / `ds[lo:hi] = rows` in ascending order;
    every completed chunk is deflated and appended to the file at once, so memory holds one chunk, not the dataset."""

    def __init__(self, f, name, shape, dtype, data, maxshape, chunks, compression, compression_opts):
        _WNode.__init__(self, f, name)
        self.shape, self.dtype = tuple(int(v) for v in shape), np.dtype(dtype)
        _enc_datatype(self.dtype)
        self.maxshape = tuple(maxshape) if maxshape is not None else None
        if self.maxshape is not None and len(self.maxshape) != len(self.shape):
            raise ValueError("maxshape must have the rank of the data")
        if chunks is None and (compression or (self.maxshape and any(m is None or m > s for m, s in zip(self.maxshape, self.shape)))):
            chunks = tuple(max(1, s) for s in self.shape)
        self.chunks = tuple(int(c) for c in chunks) if chunks else None
        if self.chunks and (not self.shape or len(self.chunks) != len(self.shape) or min(self.chunks) < 1):
            raise ValueError("bad chunk shape %r for data of shape %r" % (chunks, self.shape))
        if compression not in (None, False, "gzip", "lzf", "deflate"):
            raise NotImplementedError("compression %r" % (compression,))
        # "lzf" (the reference's call, game_converter.py:73-74) gives LZF chunks through the library's host codec, and deflate
        # level 1 where the library is not built
        self.lzf = compression == "lzf" and _native() is not None
        self.level = None if not compression or self.lzf else (1 if compression == "lzf" else int(compression_opts) if compression_opts is not None else 4)
        self.index, self._next, self._pending, self._npending = [], 0, [], 0
        self._rowwise = bool(self.chunks) and self.chunks[1:] == self.shape[1:]
        if self.chunks is None:
            arr = np.zeros(self.shape, self.dtype) if data is None else data
            self.addr = f._append(arr.tobytes()) if arr.size else UNDEF
            self.nbytes = arr.nbytes
        elif data is not None:
            if self._rowwise:
                self[0:self.shape[0]] = data
            else:
                self._write_chunks(data)
        elif not self._rowwise and int(np.prod(self.shape)):
            self._write_chunks(np.zeros(self.shape, self.dtype))

    def __len__(self):
        return self.shape[0]

    def resize(self, size, axis=None):
        if not self._rowwise or self.maxshape is None:
            raise TypeError("only datasets chunked along their first axis, created with maxshape, can be resized")
        size = (int(size),) + self.shape[1:] if axis is not None or np.ndim(size) == 0 else tuple(int(v) for v in size)
        if axis not in (None, 0) or size[1:] != self.shape[1:]:
            raise NotImplementedError("resize along the first axis only")
        if size[0] < self._next or (self.maxshape[0] is not None and size[0] > self.maxshape[0]):
            raise ValueError("cannot resize to %d rows (%d already written)" % (size[0], self._next))
        self.shape = size

    def __setitem__(self, key, value):
        self.f._check_open()
        if not self._rowwise:
            raise NotImplementedError("this dataset was written when it was created")
        if isinstance(key, (int, np.integer)):
            lo, hi = int(key), int(key) + 1
        elif isinstance(key, slice) and key.step in (None, 1):
            lo, hi, _ = key.indices(self.shape[0])
        elif key is Ellipsis:
            lo, hi = 0, self.shape[0]
        else:
            raise NotImplementedError("rows are assigned by index or by contiguous slice")
        if not 0 <= lo <= hi <= self.shape[0]:
            raise IndexError("rows %d:%d are outside the dataset (%d rows; resize first)" % (lo, hi, self.shape[0]))
        if hi == lo:
            return
        if lo != self._next:
            raise NotImplementedError("rows must be written in ascending order without gaps (next row: %d, got %d)" % (self._next, lo))
        rows = np.broadcast_to(np.asarray(value, self.dtype).reshape((-1,) + self.shape[1:]) if np.ndim(value) else np.asarray(value, self.dtype),
                               (hi - lo,) + self.shape[1:])
        # a copy: the rows wait here until a chunk is full, and the caller may reuse its buffer meanwhile (the reference's
        # converter and generator both reuse theirs)
        self._pending.append(np.array(rows, dtype=self.dtype, copy=True))
        self._npending += hi - lo
        self._next = hi
        if self._npending >= self.chunks[0]:
            self._flush(False)

    def _pack(self, raw):
        """chunk bytes -> (stored bytes, filter mask); zlib and the native LZF codec release the interpreter lock"""
        if self.lzf:
            enc = lzf_encode(raw)
            return (enc, 0) if enc is not None else (raw, 1)   # an optional filter that does not help is skipped for that chunk
        return (zlib.compress(raw, self.level), 0) if self.level is not None else (raw, 0)

    def _flush(self, final):
        """deflate and append every complete chunk of the pending rows (and, at the end, the partial one, zero-padded)"""
        if not self._npending:
            return
        from concurrent.futures import ThreadPoolExecutor
        c0 = self.chunks[0]
        rows = self._pending[0] if len(self._pending) == 1 else np.concatenate(self._pending)
        first = self._next - self._npending   # a multiple of c0: complete chunks leave at once
        whole = self._npending // c0
        pieces = [rows[i * c0:(i + 1) * c0] for i in range(whole)]
        rest = rows[whole * c0:]
        if final and len(rest):
            pad = np.zeros((c0,) + self.shape[1:], self.dtype)
            pad[:len(rest)] = rest
            pieces.append(pad)
            rest = rest[:0]
        tail = (0,) * (len(self.shape) - 1)
        with ThreadPoolExecutor(max_workers=max(1, min(16, os.cpu_count() or 1, len(pieces)))) as pool:
            for i, (blob, mask) in enumerate(pool.map(lambda p: self._pack(np.ascontiguousarray(p).tobytes()), pieces)):
                self.index.append(((first + i * c0,) + tail, self.f._append(blob), len(blob), mask))
        self._pending, self._npending = ([np.array(rest)], len(rest)) if len(rest) else ([], 0)

    def _write_chunks(self, arr):
        from concurrent.futures import ThreadPoolExecutor
        grids = [range(0, s, c) for s, c in zip(self.shape, self.chunks)]
        offs = [()]
        for g in grids:
            offs = [o + (v,) for o in offs for v in g]   # row-major = the B-tree's key order
        full = tuple(self.chunks)

        def cut(o):
            sl = tuple(slice(v, min(v + c, s)) for v, c, s in zip(o, full, self.shape))
            piece = arr[sl]
            if piece.shape != full:   # edge chunks are stored whole
                whole = np.zeros(full, arr.dtype)
                whole[tuple(slice(0, n) for n in piece.shape)] = piece
                piece = whole
            return self._pack(np.ascontiguousarray(piece).tobytes())

        with ThreadPoolExecutor(max_workers=min(16, (os.cpu_count() or 1))) as pool:
            for o, (blob, mask) in zip(offs, pool.map(cut, offs)):
                self.index.append((o, self.f._append(blob), len(blob), mask))

    def _chunk_btree(self):
        """v1 B-tree (node type 1) over self.index; returns the root's address (UNDEF without chunks)"""
        if not self.index:
            return UNDEF
        nd = len(self.shape) + 1
        key = lambda size, offs, tail=0, mask=0: struct.pack("<II", size, mask) + b"".join(struct.pack("<Q", v) for v in offs) + struct.pack("<Q", tail)
        # the key behind the last chunk: one chunk further in every dimension (what libhdf5 leaves there)
        end_key = key(0, [o + c for o, c in zip(self.index[-1][0], self.chunks)], self.dtype.itemsize)
        # level 0 entries: (first key, child address); one node holds up to 2K children
        entries = [(key(n, o, 0, m), a) for o, a, n, m in self.index]
        level = 0
        node_size = 24 + (2 * CHUNK_NODE_K) * 8 + (2 * CHUNK_NODE_K + 1) * (8 + 8 * nd)
        while True:
            groups = [entries[i:i + 2 * CHUNK_NODE_K] for i in range(0, len(entries), 2 * CHUNK_NODE_K)]
            addrs = [self.f._reserve(node_size) for _ in groups]
            nxt = []
            for gi, grp in enumerate(groups):
                body = b"TREE" + struct.pack("<BBH", 1, level, len(grp))
                body += struct.pack("<QQ", addrs[gi - 1] if gi > 0 else UNDEF, addrs[gi + 1] if gi + 1 < len(groups) else UNDEF)
                for k, a in grp:
                    body += k + struct.pack("<Q", a)
                body += groups[gi + 1][0][0] if gi + 1 < len(groups) else end_key
                self.f._fill(addrs[gi], body.ljust(node_size, b"\0"))
                nxt.append((grp[0][0], addrs[gi]))
            if len(groups) == 1:
                return addrs[0]
            entries, level = nxt, level + 1

    def _emit(self):
        self._flush(True)
        esize = self.dtype.itemsize
        msgs = [(0x0001, 0, _enc_dataspace(self.shape, self.maxshape)), (0x0003, 1, _enc_datatype(self.dtype))]
        if self.chunks is None:
            msgs.append((0x0005, 1, bytes([2, 2, 2, 1]) + struct.pack("<I", 0)))   # allocate late, fill if set, default fill value
            msgs.append((0x0008, 1, struct.pack("<BBQQ", 3, 1, self.addr, self.nbytes)))
        else:
            msgs.append((0x0005, 1, bytes([2, 3, 0, 1]) + struct.pack("<I", 0)))   # allocate incrementally
            if self.lzf:   # h5py's registration of the filter: id 32000, optional, client data = (filter revision 4, liblzf 1.5, chunk bytes)
                chunk_bytes = int(np.prod(self.chunks)) * esize
                msgs.append((0x000B, 1, struct.pack("<BB6x", 1, 1) + struct.pack("<HHHH", 32000, 8, 1, 3) + b"lzf\0\0\0\0\0" +
                             struct.pack("<IIII", 4, 0x0105, chunk_bytes, 0)))
            elif self.level is not None:
                msgs.append((0x000B, 1, struct.pack("<BB6x", 1, 1) + struct.pack("<HHHH", 1, 8, 1, 1) + b"deflate\0" + struct.pack("<II", self.level, 0)))
            msgs.append((0x0008, 1, struct.pack("<BBBQ", 3, 2, len(self.shape) + 1, self._chunk_btree()) +
                         b"".join(struct.pack("<I", c) for c in self.chunks) + struct.pack("<I", esize)))
        msgs.append((0x0012, 0, struct.pack("<B3xI", 1, self.f._mtime)))
        return self.f._object_header(msgs + self.attrs.messages()), None


class WGroup(_WNode):
    def __init__(self, f, name):
        _WNode.__init__(self, f, name)
        self._links = {}

    def _split(self, name, create):
        parts = [p for p in name.split("/") if p]
        if not parts:
            raise ValueError("empty name")
        node = self
        for p in parts[:-1]:
            if p not in node._links:
                if not create:
                    raise KeyError(name)
                node._links[p] = WGroup(self.f, p)
            node = node._links[p]
            if not isinstance(node, WGroup):
                raise ValueError("%s is not a group" % p)
        return node, parts[-1]

    def create_group(self, name):
        node, leaf = self._split(name, True)
        if leaf in node._links:
            raise ValueError("unable to create group (name already exists): %s" % name)
        node._links[leaf] = WGroup(self.f, leaf)
        return node._links[leaf]

    def require_group(self, name):
        node, leaf = self._split(name, True)
        if leaf not in node._links:
            node._links[leaf] = WGroup(self.f, leaf)
        if not isinstance(node._links[leaf], WGroup):
            raise TypeError("incompatible object (dataset) already exists: %s" % name)
        return node._links[leaf]

    def create_dataset(self, name, shape=None, dtype=None, data=None, maxshape=None, chunks=None, compression=None,
                       compression_opts=None, **_ignored):
        if data is not None:
            data = _as_array(data)
            if dtype is not None:
                data = data.astype(dtype, copy=False)
            if shape is not None:
                data = data.reshape(shape)
            data = np.require(data, requirements="C")   # (ascontiguousarray would turn a scalar into a vector)
            shape, dtype = data.shape, data.dtype
        elif shape is None:
            raise TypeError("one of data or shape must be given")
        else:
            shape, dtype = tuple(np.atleast_1d(shape)) if np.ndim(shape) else (int(shape),), np.dtype(dtype or np.float32)
        node, leaf = self._split(name, True)
        if leaf in node._links:
            raise ValueError("unable to create dataset (name already exists): %s" % name)
        self.f._check_open()
        node._links[leaf] = WDataset(self.f, leaf, shape, dtype, data, maxshape, chunks if chunks is not True else None, compression, compression_opts)
        return node._links[leaf]

    def require_dataset(self, name, shape, dtype, exact=False, **kwds):
        """the dataset of that name, created if it is not there yet (game_converter.py:71-86)"""
        if name in self:
            ds = self[name]
            if not isinstance(ds, WDataset) or ds.shape != tuple(shape):
                raise TypeError("shapes do not match (existing %r vs new %r)" % (getattr(ds, "shape", None), tuple(shape)))
            return ds
        return self.create_dataset(name, shape=shape, dtype=dtype, **kwds)

    def __setitem__(self, name, value):
        self.create_dataset(name, data=value)

    def __getitem__(self, name):
        node, leaf = self._split(name, False)
        return node._links[leaf]

    def __contains__(self, name):
        try:
            self[name]
            return True
        except KeyError:
            return False

    def keys(self):
        return sorted(self._links)

    def __iter__(self):
        return iter(self.keys())

    def _emit(self):
        """children first, then this group's heap, symbol nodes, B-tree and object header; returns (header, (btree, heap))"""
        f = self.f
        names = sorted(self._links, key=lambda n: n.encode("utf8"))   # strcmp order
        heap, offs = bytearray(8), {}
        for n in names:
            offs[n] = len(heap)
            heap += _pad8(n.encode("utf8") + b"\0")
        symbols = []
        for n in names:
            child = self._links[n]
            addr, cache = child._emit()
            scratch = struct.pack("<QQ", *cache) if cache else bytes(16)
            symbols.append((offs[n], struct.pack("<QQII", offs[n], addr, 1 if cache else 0, 0) + scratch))
        heap_addr = f._append(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), 1, f._end() + 32) + bytes(heap))   # 1 = no free block
        per = 2 * GROUP_LEAF_K
        snod_size = 8 + per * 40
        entries = []   # (heap offset of the largest name below, child address)
        for i in range(0, len(symbols), per):
            grp = symbols[i:i + per]
            body = b"SNOD" + struct.pack("<BBH", 1, 0, len(grp)) + b"".join(e for _, e in grp)
            entries.append((grp[-1][0], f._append(body.ljust(snod_size, b"\0"))))
        level, node_size = 0, 24 + (2 * GROUP_NODE_K) * 8 + (2 * GROUP_NODE_K + 1) * 8
        while True:
            groups = [entries[i:i + 2 * GROUP_NODE_K] for i in range(0, len(entries), 2 * GROUP_NODE_K)] or [[]]   # an empty group: a leaf without entries
            addrs = [f._reserve(node_size) for _ in groups]
            nxt = []
            for gi, grp in enumerate(groups):
                body = b"TREE" + struct.pack("<BBH", 0, level, len(grp))
                body += struct.pack("<QQ", addrs[gi - 1] if gi > 0 else UNDEF, addrs[gi + 1] if gi + 1 < len(groups) else UNDEF)
                body += struct.pack("<Q", groups[gi - 1][-1][0] if gi > 0 else 0)   # left key: largest name before this node
                for k, a in grp:
                    body += struct.pack("<QQ", a, k)
                f._fill(addrs[gi], body.ljust(node_size, b"\0"))
                nxt.append((grp[-1][0] if grp else 0, addrs[gi]))
            if len(groups) == 1:
                break
            entries, level = nxt, level + 1
        btree = addrs[0]
        header = f._object_header([(0x0011, 0, struct.pack("<QQ", btree, heap_addr))] + self.attrs.messages())
        return header, (btree, heap_addr)


class _WriteFile(WGroup):
    def __init__(self, path):
        import time
        WGroup.__init__(self, self, "/")
        self._fh = open(path, "wb")
        self._fh.write(bytes(96))   # the superblock goes in at close()
        self._pos = 96
        self._mtime = int(time.time()) & 0xFFFFFFFF
        self._closed = False

    # ---- file space: append-only, 8-byte aligned
    def _check_open(self):
        if self._closed:
            raise ValueError("file is closed")

    def _end(self):
        return self._pos

    def _append(self, blob):
        addr = self._pos
        blob = _pad8(bytes(blob))
        self._fh.write(blob)
        self._pos += len(blob)
        return addr

    def _reserve(self, size):
        return self._append(bytes(size))

    def _fill(self, addr, blob):
        self._fh.seek(addr)
        self._fh.write(blob)
        self._fh.seek(self._pos)

    def _object_header(self, msgs):
        body = b""
        for mtype, flags, data in msgs:
            data = _pad8(data)
            body += struct.pack("<HHB3x", mtype, len(data), flags) + data
        return self._append(struct.pack("<BBHII4x", 1, 0, len(msgs), 1, len(body)) + body)

    def close(self):
        if self._closed:
            return
        root, (btree, heap) = self._emit()
        sb = MAGIC + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", GROUP_LEAF_K, GROUP_NODE_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, self._pos, UNDEF)
        sb += struct.pack("<QQII", 0, root, 1, 0) + struct.pack("<QQ", btree, heap)
        assert len(sb) == 96
        self._fill(0, sb)
        self._fh.close()
        self._closed = True

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
        return False


def File(path, mode="r"):
    """reader (mode "r") or writer (mode "w") for the subset of HDF5 described in the module docstring"""
    if mode == "r":
        return _ReadFile(path)
    if mode == "w":
        return _WriteFile(path)
    raise NotImplementedError("hdf5_lite opens files with mode 'r' or 'w', not %r" % mode)


def open_file(path, mode="r"):
    """h5py.File when h5py is importable, else this module's reader / writer"""
    try:
        import h5py
        return h5py.File(path, mode)
    except ImportError:
        return File(path, mode)
