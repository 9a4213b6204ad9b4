"""Minimal HDF5 reader (numpy only) -- TEST INFRASTRUCTURE, not product code.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may
import this module.  The product's own reader is the C++ one behind
`include/highfive/*.hpp` / `nsdf_h5_*`; this file is an *independent* second
implementation used to cross-check it.

What it understands is exactly what the reference's weight files need
(`/root/reference/src/neuralNetwork.cpp:85-151` walks them through HighFive ->
libhdf5, an un-vendored system library):

  * superblock v0, 8-byte offsets/lengths,
  * v1 object headers (+ continuation blocks),
  * "old style" groups: symbol-table message -> v1 B-tree (TREE) -> symbol
    nodes (SNOD) -> local heap (HEAP) for the link names,
  * datasets with a v1 dataspace, an IEEE little-endian float32 datatype and a
    v3 *contiguous* (or compact) data layout.

The HDF5 file-format specification (public, "HDF5 File Format Specification
Version 2.0") is the source for every structure below; nothing here is derived
from libhdf5 or HighFive code.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5FormatError(ValueError):
    pass


@dataclass
class H5Dataset:
    name: str
    dims: Tuple[int, ...]
    data_offset: int  # absolute file offset of the raw data (or -1 if compact)
    data: np.ndarray


@dataclass
class H5Group:
    name: str
    # name-sorted, exactly like HighFive::listObjectNames (H5_INDEX_NAME, native order)
    children: Dict[str, object] = field(default_factory=dict)

    def list_object_names(self) -> List[str]:
        return sorted(self.children.keys(), key=lambda s: s.encode())


class H5File:
    def __init__(self, path: str):
        with open(path, "rb") as f:
            self.buf = f.read()
        b = self.buf
        if b[:8] != _SIG:
            raise H5FormatError("not an HDF5 file (bad signature)")
        ver = b[8]
        if ver != 0:
            raise H5FormatError(f"unsupported superblock version {ver}")
        self.size_of_offsets = b[13]
        self.size_of_lengths = b[14]
        if self.size_of_offsets != 8 or self.size_of_lengths != 8:
            raise H5FormatError("only 8-byte offsets/lengths are supported")
        self.group_leaf_k, self.group_internal_k = struct.unpack_from("<HH", b, 16)
        base, _free, self.eof, _drv = struct.unpack_from("<QQQQ", b, 24)
        self.base = base
        # root group symbol table entry at offset 56
        _name_off, root_hdr, cache_type = struct.unpack_from("<QQI", b, 56)
        self.root_header = root_hdr
        self.root = self._read_group("/", root_hdr)

    # ---- object header -------------------------------------------------
    def _messages(self, addr: int):
        b = self.buf
        ver, _r, nmsg, _refc, hsize = struct.unpack_from("<BBHII", b, addr)
        if ver != 1:
            raise H5FormatError(f"unsupported object header version {ver} @ {addr}")
        blocks = [(addr + 16, hsize)]  # v1 header: 12 bytes + 4 pad to 8-byte boundary
        out = []
        while blocks and len(out) < nmsg:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and len(out) < nmsg:
                mtype, msize, mflags = struct.unpack_from("<HHB", b, pos)
                body = pos + 8
                if mtype == 0x0010:  # continuation
                    coff, clen = struct.unpack_from("<QQ", b, body)
                    blocks.append((coff + self.base, clen))
                out.append((mtype, body, msize, mflags))
                pos = body + msize
        return out

    # ---- groups ----------------------------------------------------------
    def _heap_name(self, heap_addr: int, off: int) -> str:
        b = self.buf
        if b[heap_addr:heap_addr + 4] != b"HEAP":
            raise H5FormatError("bad local heap signature")
        _dsize, _free, daddr = struct.unpack_from("<QQQ", b, heap_addr + 8)
        s = daddr + self.base + off
        e = b.index(b"\x00", s)
        return b[s:e].decode("utf-8")

    def _walk_btree(self, addr: int, heap_addr: int, out: List[Tuple[str, int]]):
        b = self.buf
        sig = b[addr:addr + 4]
        if sig == b"TREE":
            ntype, level, used = struct.unpack_from("<BBH", b, addr + 4)
            if ntype != 0:
                raise H5FormatError("expected a group B-tree node")
            pos = addr + 8 + 16  # skip sibling pointers
            # key0, child0, key1, child1, ..., key_used
            for i in range(used):
                child = struct.unpack_from("<Q", b, pos + 8)[0]
                self._walk_btree(child + self.base, heap_addr, out)
                pos += 16
        elif sig == b"SNOD":
            nsym = struct.unpack_from("<H", b, addr + 6)[0]
            pos = addr + 8
            for _ in range(nsym):
                name_off, hdr, _ctype = struct.unpack_from("<QQI", b, pos)
                out.append((self._heap_name(heap_addr, name_off), hdr + self.base))
                pos += 40
        else:
            raise H5FormatError(f"unexpected node signature {sig!r} @ {addr}")

    def _read_group(self, name: str, hdr: int) -> H5Group:
        g = H5Group(name)
        stab = None
        for mtype, body, _sz, _fl in self._messages(hdr):
            if mtype == 0x0011:
                stab = struct.unpack_from("<QQ", self.buf, body)
        if stab is None:
            raise H5FormatError(f"group {name} has no symbol-table message")
        entries: List[Tuple[str, int]] = []
        self._walk_btree(stab[0] + self.base, stab[1] + self.base, entries)
        for child_name, child_hdr in entries:
            g.children[child_name] = self._read_object(child_name, child_hdr)
        return g

    # ---- objects -----------------------------------------------------------
    def _read_object(self, name: str, hdr: int):
        msgs = self._messages(hdr)
        types = {m[0] for m in msgs}
        if 0x0011 in types:
            return self._read_group(name, hdr)
        if 0x0008 in types:
            return self._read_dataset(name, msgs)
        raise H5FormatError(f"object {name}: neither group nor dataset")

    def _read_dataset(self, name: str, msgs) -> H5Dataset:
        b = self.buf
        dims = None
        dtype_ok = False
        layout = None
        for mtype, body, msize, _fl in msgs:
            if mtype == 0x0001:  # dataspace
                ver, rank, flags = struct.unpack_from("<BBB", b, body)
                if ver == 1:
                    p = body + 8
                elif ver == 2:
                    p = body + 4
                else:
                    raise H5FormatError("unsupported dataspace version")
                dims = struct.unpack_from("<" + "Q" * rank, b, p)
            elif mtype == 0x0003:  # datatype
                cls_ver = b[body]
                cls = cls_ver & 0x0F
                bits0 = b[body + 1]
                size = struct.unpack_from("<I", b, body + 4)[0]
                # class 1 = floating point; bit0 of the class bits = byte order (0 = LE)
                if cls == 1 and size == 4 and (bits0 & 1) == 0:
                    bit_off, prec, epos, esize, mpos, msz, ebias = struct.unpack_from(
                        "<HHBBBBI", b, body + 8)
                    dtype_ok = (prec == 32 and epos == 23 and esize == 8 and mpos == 0
                                and msz == 23 and ebias == 127)
            elif mtype == 0x0008:  # layout
                ver = b[body]
                if ver != 3:
                    raise H5FormatError("unsupported layout message version")
                lclass = b[body + 1]
                if lclass == 1:
                    addr, size = struct.unpack_from("<QQ", b, body + 2)
                    layout = ("contiguous", addr + self.base, size)
                elif lclass == 0:
                    size = struct.unpack_from("<H", b, body + 2)[0]
                    layout = ("compact", body + 4, size)
                else:
                    raise H5FormatError("chunked datasets are not supported")
        if dims is None or layout is None:
            raise H5FormatError(f"dataset {name}: missing dataspace/layout")
        if not dtype_ok:
            raise H5FormatError(f"dataset {name}: only little-endian IEEE float32 is supported")
        n = int(np.prod(dims)) if dims else 1
        kind, off, size = layout
        if size < 4 * n:
            raise H5FormatError(f"dataset {name}: storage smaller than dims")
        data = np.frombuffer(b, dtype="<f4", count=n, offset=off).reshape(dims).copy()
        return H5Dataset(name, tuple(int(d) for d in dims), off if kind == "contiguous" else -1, data)


def load_keras_mlp(path: str):
    """Return [(kernel[in][out], bias[out])] in the reference's layer order.

    Follows `/root/reference/src/neuralNetwork.cpp:85-151`: top-level object
    names in name-sorted order; each must be a group holding exactly one group
    of the same name; inside it, rank-1 datasets are biases and rank-2 datasets
    kernels.
    """
    f = H5File(path)
    layers = []
    for lname in f.root.list_object_names():
        g = f.root.children[lname]
        if not isinstance(g, H5Group) or len(g.children) != 1:
            raise H5FormatError("Unsupported Layer")
        inner = g.children.get(lname)
        if not isinstance(inner, H5Group):
            raise H5FormatError("Unsupported Layer")
        kernel = bias = None
        for dname in inner.list_object_names():
            d = inner.children[dname]
            if not isinstance(d, H5Dataset):
                raise H5FormatError("Unsupported Layer")
            if len(d.dims) == 1:
                bias = d.data
            elif len(d.dims) == 2:
                kernel = d.data
            else:
                raise H5FormatError("Unsupported layer, to many dims!")
        layers.append((kernel, bias))
    return layers


def pack_model(layers):
    """Pack [(kernel[in][out], bias)] into the flat layout the oracle / C-ABI take:
    per layer W[out][in] row-major (the transpose the reference makes in
    `src/layers/denseLayer.cu:217-227`) followed by b[out]; dims = [in0, out0, out1, ...].
    """
    dims = [int(layers[0][0].shape[0])]
    chunks = []
    for k, b in layers:
        dims.append(int(k.shape[1]))
        chunks.append(np.ascontiguousarray(k.T, dtype=np.float32).ravel())
        chunks.append(np.ascontiguousarray(b, dtype=np.float32).ravel())
    return np.concatenate(chunks), np.asarray(dims, dtype=np.int32)
