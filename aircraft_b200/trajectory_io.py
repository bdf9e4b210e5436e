"""Trajectory files in the reference's HDF5 layout, without h5py.

The reference stores solver iterates and simulations as
    iteration_<k>/state   [13, N+1] float64      (column per node)
    iteration_<k>/control [ 7, N(+1)] float64
    iteration_<k>/times   [N+1] float64          ('time' in the SaveMixin variant)
(control/base.py:89-104, main/dynamics/dynamics.py:134-139; read back by
plotting/plotting.py:72-95).  h5py/libhdf5 are not installed on this image, so this module carries
a small pure-Python implementation of the subset of the HDF5 file format those files use:

  reader  superblock v0, object headers v1 (+continuations), old-style groups (symbol table, v1
          B-tree, local heap), little-endian IEEE float / fixed-point datasets with contiguous,
          compact or chunked (+gzip / shuffle) layout.  Checked against the files the reference
          ships (tests/golden/simulation.h5 and friends, written by libhdf5 through h5py).
  writer  the same structures, contiguous float64/float32/int datasets in one level of groups.
          The output is read back by the reader above in the tests; it has NOT been opened with
          libhdf5 on this box (none available), which DESIGN.md states.

This is host-side I/O around the hot path; nothing here runs on the GPU.
"""
from __future__ import annotations

import struct
import zlib
from typing import Dict, Mapping, Optional

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


# ------------------------------------------------------------------------------------------ reader
class _Reader:
    def __init__(self, buf: bytes):
        self.b = buf
        if buf[:8] != _SIG:
            raise ValueError("not an HDF5 file")
        if buf[8] != 0:
            raise ValueError(f"superblock version {buf[8]} is not supported (only 0)")
        self.O, self.L = buf[13], buf[14]
        if (self.O, self.L) != (8, 8):
            raise ValueError("only 8-byte offsets/lengths are supported")
        self.base = self.u(24, 8)
        ste = 24 + 4 * 8                       # base, free-space, eof, driver-info addresses
        self.root = self._ste(ste)

    def u(self, off: int, n: int) -> int:
        return int.from_bytes(self.b[off:off + n], "little")

    def _ste(self, off: int) -> dict:
        return {"name_off": self.u(off, 8), "ohdr": self.u(off + 8, 8), "cache": self.u(off + 16, 4),
                "btree": self.u(off + 24, 8), "heap": self.u(off + 32, 8)}

    # -- object header v1 -> list of (type, bytes)
    def messages(self, addr: int):
        b = self.b
        if b[addr] != 1:
            raise ValueError(f"object header version {b[addr]} is not supported (only 1)")
        nmsg, size = self.u(addr + 2, 2), self.u(addr + 8, 4)
        blocks = [(addr + 16, size)]
        out = []
        while blocks and len(out) < nmsg:
            off, left = blocks.pop(0)
            end = off + left
            while off + 8 <= end and len(out) < nmsg:
                mtype, msize = self.u(off, 2), self.u(off + 2, 2)
                data = b[off + 8:off + 8 + msize]
                off += 8 + msize
                if mtype == 0x10:              # continuation
                    blocks.append((self.u_bytes(data, 0, 8), self.u_bytes(data, 8, 8)))
                out.append((mtype, data))
        return out

    @staticmethod
    def u_bytes(d: bytes, off: int, n: int) -> int:
        return int.from_bytes(d[off:off + n], "little")

    # -- old-style group: {name: object header address}
    def group_entries(self, btree: int, heap: int) -> Dict[str, int]:
        b = self.b
        if b[heap:heap + 4] != b"HEAP":
            raise ValueError("bad local heap")
        hdata = self.u(heap + 24, 8)
        out: Dict[str, int] = {}

        def name_at(o):
            e = b.index(b"\0", hdata + o)
            return b[hdata + o:e].decode()

        def walk(node):
            if b[node:node + 4] == b"SNOD":
                n = self.u(node + 6, 2)
                for i in range(n):
                    s = self._ste(node + 8 + 40 * i)
                    out[name_at(s["name_off"])] = s["ohdr"]
                return
            if b[node:node + 4] != b"TREE" or b[node + 4] != 0:
                raise ValueError("bad group B-tree node")
            n = self.u(node + 6, 2)
            p = node + 24
            for i in range(n):
                walk(self.u(p + 8 + 16 * i, 8))   # key_i (8) child_i (8) ...

        walk(btree)
        return out

    def open_group(self, ohdr: int) -> Optional[Dict[str, int]]:
        for t, d in self.messages(ohdr):
            if t == 0x11:
                return self.group_entries(self.u_bytes(d, 0, 8), self.u_bytes(d, 8, 8))
        return None

    # -- dataset
    def dataset(self, ohdr: int) -> Optional[np.ndarray]:
        shape = dtype = layout = None
        filters = []
        for t, d in self.messages(ohdr):
            if t == 0x01:
                ver, rank = d[0], d[1]
                p = 8 if ver == 1 else 4
                shape = tuple(self.u_bytes(d, p + 8 * i, 8) for i in range(rank))
            elif t == 0x03:
                cls, size = d[0] & 0x0F, self.u_bytes(d, 4, 4)
                if d[1] & 1:
                    raise ValueError("big-endian datasets are not supported")
                if cls == 1:
                    dtype = np.dtype(f"<f{size}")
                elif cls == 0:
                    dtype = np.dtype(("<i" if d[1] & 0x08 else "<u") + str(size))
                else:
                    raise ValueError(f"datatype class {cls} is not supported")
            elif t == 0x08:
                if d[0] != 3:
                    raise ValueError(f"data layout message version {d[0]} is not supported (only 3)")
                layout = d
            elif t == 0x0B:
                nf = d[1]
                p = 8
                for _ in range(nf):
                    fid, nlen, ncv = self.u_bytes(d, p, 2), self.u_bytes(d, p + 2, 2), self.u_bytes(d, p + 6, 2)
                    p += 8 + ((nlen + 7) // 8) * 8
                    cv = [self.u_bytes(d, p + 4 * i, 4) for i in range(ncv)]
                    p += 4 * ncv + (4 if ncv % 2 else 0)
                    filters.append((fid, cv))
        if shape is None or dtype is None or layout is None:
            return None
        count = int(np.prod(shape)) if shape else 1
        cls = layout[1]
        if cls == 0:
            size = self.u_bytes(layout, 2, 2)
            return np.frombuffer(layout[4:4 + size], dtype, count).reshape(shape).copy()
        if cls == 1:
            addr = self.u_bytes(layout, 2, 8)
            if addr == _UNDEF:
                return np.zeros(shape, dtype)
            return np.frombuffer(self.b, dtype, count, self.base + addr).reshape(shape).copy()
        if cls == 2:
            nd = layout[2]
            bt = self.u_bytes(layout, 3, 8)
            cdims = tuple(self.u_bytes(layout, 11 + 4 * i, 4) for i in range(nd - 1))
            out = np.zeros(shape, dtype)
            if bt != _UNDEF:
                self._chunks(bt, nd, cdims, dtype, filters, out)
            return out
        raise ValueError(f"layout class {cls} is not supported")

    def _chunks(self, node, nd, cdims, dtype, filters, out):
        b = self.b
        if b[node:node + 4] != b"TREE" or b[node + 4] != 1:
            raise ValueError("bad chunk B-tree node")
        level, n = b[node + 5], self.u(node + 6, 2)
        ksz = 8 + 8 * nd
        p = node + 24
        for i in range(n):
            k = p + i * (ksz + 8)
            csize, fmask = self.u(k, 4), self.u(k + 4, 4)
            offs = tuple(self.u(k + 8 + 8 * j, 8) for j in range(nd - 1))
            child = self.u(k + ksz, 8)
            if level > 0:
                self._chunks(child, nd, cdims, dtype, filters, out)
                continue
            raw = b[self.base + child:self.base + child + csize]
            for idx in range(len(filters) - 1, -1, -1):   # undo the pipeline back to front
                fid, cv = filters[idx]
                if fmask & (1 << idx):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    es = cv[0] if cv else dtype.itemsize
                    a = np.frombuffer(raw, np.uint8)
                    m = len(a) // es
                    raw = a[:m * es].reshape(es, m).T.tobytes() + a[m * es:].tobytes()
                elif fid == 3:
                    raw = raw[:-4]                         # fletcher32 checksum
                else:
                    raise ValueError(f"filter {fid} is not supported")
            chunk = np.frombuffer(raw, dtype, int(np.prod(cdims))).reshape(cdims)
            sl_out = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, cdims, out.shape))
            sl_in = tuple(slice(0, s.stop - s.start) for s in sl_out)
            out[sl_out] = chunk[sl_in]


def read_trajectories(path) -> Dict[str, Dict[str, np.ndarray]]:
    """{group name: {dataset name: array}} for every group under the root of an HDF5 trajectory file
    (datasets directly under the root are returned under the key '')."""
    with open(path, "rb") as f:
        r = _Reader(f.read())
    root = r.open_group(r.root["ohdr"])
    if root is None:
        raise ValueError("root group has no symbol table")
    out: Dict[str, Dict[str, np.ndarray]] = {}
    for name, ohdr in root.items():
        sub = r.open_group(ohdr)
        if sub is None:
            a = r.dataset(ohdr)
            if a is not None:
                out.setdefault("", {})[name] = a
            continue
        out[name] = {}
        for dn, dh in sub.items():
            a = r.dataset(dh)
            if a is not None:
                out[name][dn] = a
    return out


def load_iteration(path, iteration: int) -> Optional[Dict[str, np.ndarray]]:
    """What TrajectoryData.load (plotting/plotting.py:72-95) extracts: the datasets of
    'iteration_<k>', or None when that group does not exist."""
    return read_trajectories(path).get(f"iteration_{iteration}")


# ------------------------------------------------------------------------------------------ writer
_LEAF_K, _INT_K = 4, 16            # libhdf5 defaults: 8 symbols per SNOD, 32 children per TREE node
_FREE_NULL = 1                     # H5HL_FREE_NULL: "no free block" in a local heap


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def _msg(mtype: int, data: bytes) -> bytes:
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), 0) + data


def _ohdr(msgs) -> bytes:
    body = b"".join(msgs)
    return struct.pack("<BxHII4x", 1, len(msgs), 1, len(body)) + body


def _datatype(dt: np.dtype) -> bytes:
    if dt.kind == "f" and dt.itemsize in (4, 8):
        e, m, bias = (8, 23, 127) if dt.itemsize == 4 else (11, 52, 1023)
        return struct.pack("<BBBBIHHBBBBI", 0x11, 0x20, 8 * dt.itemsize - 1, 0, dt.itemsize, 0,
                           8 * dt.itemsize, m, e, 0, m, bias)
    if dt.kind in "iu":
        return struct.pack("<BBBBIHH", 0x10, 0x08 if dt.kind == "i" else 0, 0, 0, dt.itemsize, 0, 8 * dt.itemsize)
    raise ValueError(f"dtype {dt} is not supported by the HDF5 writer")


class _Writer:
    def __init__(self):
        self.buf = bytearray(96)               # superblock + root symbol-table entry, filled last

    def alloc(self, data: bytes) -> int:
        self.buf += b"\0" * (-len(self.buf) % 8)
        addr = len(self.buf)
        self.buf += data
        return addr

    def dataset(self, a: np.ndarray) -> int:
        a = np.ascontiguousarray(a)
        if a.dtype.byteorder == ">":
            a = a.astype(a.dtype.newbyteorder("<"))
        raw = a.tobytes()
        daddr = self.alloc(raw) if raw else _UNDEF
        space = struct.pack("<BBB5x", 1, a.ndim, 0) + b"".join(struct.pack("<Q", s) for s in a.shape)
        msgs = [_msg(0x01, space), _msg(0x03, _datatype(a.dtype)), _msg(0x05, bytes([2, 2, 2, 1, 0, 0, 0, 0])),
                _msg(0x08, struct.pack("<BBQQ", 3, 1, daddr, len(raw)))]
        return self.alloc(_ohdr(msgs))

    def group(self, entries: Mapping[str, tuple]) -> tuple:
        """entries: name -> (object header address, cache type, btree, heap).  Returns
        (object header address, btree address, heap address) of the new group."""
        names = sorted(entries, key=lambda s: s.encode())
        heap = bytearray(8)                    # offset 0: the empty string
        off = {}
        for nm in names:
            off[nm] = len(heap)
            heap += _pad8(nm.encode() + b"\0")
        hdata = self.alloc(bytes(heap))
        haddr = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), _FREE_NULL, hdata))
        # leaves
        leaves = []                            # (address, last name)
        for i in range(0, max(len(names), 1), 2 * _LEAF_K):
            chunk = names[i:i + 2 * _LEAF_K]
            body = b"SNOD" + struct.pack("<BxH", 1, len(chunk))
            for nm in chunk:
                oh, cache, bt, hp = entries[nm]
                body += struct.pack("<QQI4xQQ", off[nm], oh, cache, bt if cache == 1 else 0, hp if cache == 1 else 0)
            body += b"\0" * (8 + 2 * _LEAF_K * 40 - len(body))
            leaves.append((self.alloc(body), off[chunk[-1]] if chunk else 0))
        level, nodes = 0, leaves
        while True:
            size = 24 + (2 * _INT_K + 1) * 8 + 2 * _INT_K * 8
            groups_ = [nodes[i:i + 2 * _INT_K] for i in range(0, len(nodes), 2 * _INT_K)]
            addrs = [self.alloc(b"\0" * size) for _ in groups_]
            up = []
            prev_key = 0
            for gi, (grp, addr) in enumerate(zip(groups_, addrs)):
                left = addrs[gi - 1] if gi > 0 else _UNDEF
                right = addrs[gi + 1] if gi + 1 < len(addrs) else _UNDEF
                body = b"TREE" + struct.pack("<BBHQQ", 0, level, len(grp), left, right) + struct.pack("<Q", prev_key)
                for child, last in grp:
                    body += struct.pack("<QQ", child, last)
                    prev_key = last
                self.buf[addr:addr + len(body)] = body
                up.append((addr, grp[-1][1]))
            if len(up) == 1:
                btree = up[0][0]
                break
            level, nodes = level + 1, up
        oh = self.alloc(_ohdr([_msg(0x11, struct.pack("<QQ", btree, haddr))]))
        return oh, btree, haddr

    def finish(self, root: tuple) -> bytes:
        oh, bt, hp = root
        sb = _SIG + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", _LEAF_K, _INT_K, 0)
        sb += struct.pack("<QQQQ", 0, _UNDEF, len(self.buf), _UNDEF)
        sb += struct.pack("<QQI4xQQ", 0, oh, 1, bt, hp)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)


def write_trajectories(path, groups: Mapping[str, Mapping[str, np.ndarray]]) -> None:
    """Writes {group: {dataset: array}} as an HDF5 file (superblock v0, old-style groups, contiguous
    little-endian datasets): the structure h5py's File/create_group/create_dataset produce for the
    reference's trajectory files, minus attributes and compression."""
    w = _Writer()
    top = {}
    for gname, dsets in groups.items():
        if gname == "":                                  # datasets stored directly under the root group
            for dn, a in dsets.items():
                top[dn] = (w.dataset(np.asarray(a)), 0, 0, 0)
            continue
        ents = {dn: (w.dataset(np.asarray(a)), 0, 0, 0) for dn, a in dsets.items()}
        oh, bt, hp = w.group(ents)
        top[gname] = (oh, 1, bt, hp)
    data = w.finish(w.group(top))
    with open(path, "wb") as f:
        f.write(data)


def save_iteration(path, state, control, times=None, iteration: int = 0, append: bool = True, **extra) -> None:
    """Stores one iterate as 'iteration_<k>/{state,control,times}' (+ optional lam/mu/nu ...), column per
    node like the reference (state [13, N+1], control [7, N]); with append=True existing groups of the
    file are kept, datasets stored directly under the root included (the reference opens its files in mode 'a',
    main/dynamics/dynamics.py:134-139).  The file is re-read and rewritten as a whole (O(file size) per iterate);
    attributes (the reference's grp.attrs['timestamp']) and gzip compression of the existing content are NOT
    preserved.  EXPERIMENTAL: the writer's output is checked by this module's own reader and structurally against
    two reference files (tests/test_trajectory_io.py); it has never been opened by libhdf5 / h5py (neither is in
    the image)."""
    import os
    groups: Dict[str, Dict[str, np.ndarray]] = {}
    if append and os.path.exists(path):
        groups = dict(read_trajectories(path))
    g = {"state": np.asarray(state, np.float64), "control": np.asarray(control, np.float64)}
    if times is not None:
        g["times"] = np.asarray(times, np.float64)
    for k, v in extra.items():
        g[k] = np.asarray(v)
    groups[f"iteration_{iteration}"] = g
    write_trajectories(path, groups)


def save_rollout(path, X, U, dt, iteration: int = 0, append: bool = True) -> None:
    """A rollout in the product's layout (X[H+1,13], U[H,7]) as a reference trajectory file."""
    X = np.asarray(X.detach().cpu() if hasattr(X, "detach") else X, np.float64)
    U = np.asarray(U.detach().cpu() if hasattr(U, "detach") else U, np.float64)
    if X.ndim != 2 or X.shape[1] != 13 or U.ndim != 2 or U.shape[1] != 7:
        raise RuntimeError(f"save_rollout wants X[H+1,13], U[H,7]; got {X.shape}, {U.shape}")
    save_iteration(path, X.T, U.T, np.arange(X.shape[0]) * float(dt), iteration=iteration, append=append)
