"""Minimal read-only HDF5 for the reference's dataset files, for images without h5py.

Covers what `game_converter.py:67-88` / Keras 1.0 `save_weights` produce with h5py 2.x defaults (HDF5 1.8 "earliest"
format): superblock v0/v1, symbol-table groups (v1 B-tree + local heap), v1 object headers with continuation
blocks, fixed-point / IEEE-float / fixed-length-string datatypes, contiguous, compact and chunked layouts (v1 chunk
B-tree) with the gzip (1), shuffle (2) and LZF (32000) filters, and v1-v3 attribute messages.  That is the on-disk
contract of the SL trainer (`states` uint8 (N,F,19,19) LZF-chunked, `actions` uint8 (N,2), group `file_offsets`;
supervised_policy_trainer.py:128-129).  `open_file()` prefers h5py when it is importable.
"""
import struct
import zlib

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


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

    def _chunk(self, offs):
        if self._chunks is None:
            self._chunks = {}
            if self.btree != UNDEF:
                self._walk(self.btree, self._chunks)
        if offs in self._cache:
            return self._cache[offs]
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
                    raw = lzf_decompress(raw, n * self.dtype.itemsize)
                elif fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    raw = np.frombuffer(raw, np.uint8).reshape(self.dtype.itemsize, -1).T.tobytes()
                else:
                    raise NotImplementedError("HDF5 filter %d" % fid)
            arr = np.frombuffer(raw, self.dtype, count=n).reshape(self.chunk)
        if len(self._cache) > 64:
            self._cache.clear()
        self._cache[offs] = arr
        return arr

    def _rows(self, lo, hi):
        """rows [lo, hi) along the first axis"""
        if self.layout != 2:
            count = int(np.prod(self.shape)) if self.shape else 1
            full = np.frombuffer(self.f.b, self.dtype, count=count, offset=self.data_addr).reshape(self.shape)
            return full[lo:hi]
        out = np.zeros((hi - lo,) + self.shape[1:], self.dtype)
        c0 = self.chunk[0]
        grid = [range(0, s, c) for s, c in zip(self.shape[1:], self.chunk[1:])]
        for first in range(lo // c0 * c0, hi, c0):
            idxs = [()]
            for g in grid:
                idxs = [i + (o,) for i in idxs for o in g]
            for rest in idxs:
                ch = self._chunk((first,) + rest)
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


class File(Group):
    def __init__(self, path, mode="r"):
        if mode != "r":
            raise NotImplementedError("hdf5_lite is read-only")
        with open(path, "rb") as fh:
            buf = memoryview(fh.read())
        if bytes(buf[:8]) != b"\x89HDF\r\n\x1a\n":
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


def open_file(path, mode="r"):
    """h5py.File when h5py is importable, else the read-only reader above"""
    try:
        import h5py
        return h5py.File(path, mode)
    except ImportError:
        return File(path, mode)
