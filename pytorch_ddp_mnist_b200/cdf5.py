"""CDF-5 ("64BIT_DATA") classic-format netCDF reader/writer, host side, numpy only.

This is the on-disk format the reference's converter produces
(`mnist_to_netcdf.ipynb` cell 2, `to_nc`: `pncpy.File(..., format="64BIT_DATA")`, dims
``Y, X, idx``, variables ``images (idx,Y,X)`` and ``labels (idx,)``, both NC_UBYTE) and that
`MNISTNetCDF` reads one sample at a time (`mnist_pnetcdf_cpu_mp.py:31-46`).  PnetCDF itself is not
in this image, so the subset of the format the path needs is implemented here from the netCDF
classic-format grammar:

    header   = magic numrecs dim_list gatt_list var_list
    magic    = 'C' 'D' 'F' 0x05
    numrecs  = INT64
    dim_list = ABSENT | NC_DIMENSION(0x0A) nelems(INT64) [name dim_length(INT64)] ...
    att_list = ABSENT | NC_ATTRIBUTE(0x0C) nelems(INT64) [name nc_type nelems(INT64) values] ...
    var_list = ABSENT | NC_VARIABLE(0x0B)  nelems(INT64)
               [name ndims(INT64) dimid(INT64)... vatt_list nc_type vsize(INT64) begin(INT64)] ...
    name     = nelems(INT64) bytes padded to 4
    ABSENT   = 0x00000000 INT64(0)

Everything is big-endian.  Only fixed-size (non-record) variables are supported, which is all the
reference writes (``idx`` is a fixed dimension).  Readers MUST use each variable's ``begin`` — the
data section does not necessarily follow the header directly (PnetCDF aligns it).

The same grammar is implemented in C++ in ``csrc/cdf5_reader.cpp`` (the `ddpm_cdf5_*` C-ABI entry
points); this module is the writer and the numpy-facing reader used by the `pncpy` stand-in.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import numpy as np

MAGIC = b"CDF\x05"
NC_DIMENSION = 0x0A
NC_VARIABLE = 0x0B
NC_ATTRIBUTE = 0x0C

# nc_type codes (classic + CDF-5 extended types)
NC_BYTE, NC_CHAR, NC_SHORT, NC_INT, NC_FLOAT, NC_DOUBLE = 1, 2, 3, 4, 5, 6
NC_UBYTE, NC_USHORT, NC_UINT, NC_INT64, NC_UINT64 = 7, 8, 9, 10, 11

_NP_OF_TYPE = {
    NC_BYTE: np.dtype("i1"), NC_CHAR: np.dtype("S1"), NC_SHORT: np.dtype(">i2"),
    NC_INT: np.dtype(">i4"), NC_FLOAT: np.dtype(">f4"), NC_DOUBLE: np.dtype(">f8"),
    NC_UBYTE: np.dtype("u1"), NC_USHORT: np.dtype(">u2"), NC_UINT: np.dtype(">u4"),
    NC_INT64: np.dtype(">i8"), NC_UINT64: np.dtype(">u8"),
}


def _pad4(n: int) -> int:
    return (n + 3) & ~3


@dataclass
class VarInfo:
    name: str
    dimids: Tuple[int, ...]
    nc_type: int
    vsize: int
    begin: int
    shape: Tuple[int, ...] = ()
    atts: Dict[str, np.ndarray] = field(default_factory=dict)

    @property
    def dtype(self) -> np.dtype:
        return _NP_OF_TYPE[self.nc_type]


@dataclass
class Header:
    numrecs: int
    dims: List[Tuple[str, int]]
    gatts: Dict[str, np.ndarray]
    vars: Dict[str, VarInfo]
    header_size: int


# ----------------------------------------------------------------------------- encoding
def _enc_name(name: str) -> bytes:
    raw = name.encode("utf-8")
    return struct.pack(">q", len(raw)) + raw + b"\0" * (_pad4(len(raw)) - len(raw))


def _enc_absent() -> bytes:
    return struct.pack(">iq", 0, 0)


def encode_header(dims: Sequence[Tuple[str, int]],
                  variables: Sequence[Tuple[str, int, Sequence[int]]],
                  data_align: int = 512) -> Tuple[bytes, Dict[str, VarInfo]]:
    """Build a CDF-5 header for fixed-size variables.

    ``variables`` is ``[(name, nc_type, dimids)]`` in definition order.  Returns the header bytes
    (un-padded) and the per-variable layout; the first variable begins at the header size rounded
    up to ``data_align``, later ones follow back to back at 4-byte alignment (the classic-format
    rule: ``vsize`` is the byte size rounded up to 4).
    """
    def build(begins: Sequence[int]) -> bytes:
        out = [MAGIC, struct.pack(">q", 0)]
        if dims:
            out.append(struct.pack(">iq", NC_DIMENSION, len(dims)))
            for name, length in dims:
                out.append(_enc_name(name) + struct.pack(">q", length))
        else:
            out.append(_enc_absent())
        out.append(_enc_absent())                              # no global attributes
        if variables:
            out.append(struct.pack(">iq", NC_VARIABLE, len(variables)))
            for (name, nc_type, dimids), begin in zip(variables, begins):
                nelem = 1
                for d in dimids:
                    nelem *= dims[d][1]
                vsize = _pad4(nelem * _NP_OF_TYPE[nc_type].itemsize)
                out.append(_enc_name(name))
                out.append(struct.pack(">q", len(dimids)))
                out.append(b"".join(struct.pack(">q", d) for d in dimids))
                out.append(_enc_absent())                      # no variable attributes
                out.append(struct.pack(">iqq", nc_type, vsize, begin))
        else:
            out.append(_enc_absent())
        return b"".join(out)

    hsize = len(build([0] * len(variables)))
    begins, cur = [], -(-hsize // data_align) * data_align if data_align > 1 else hsize
    layout: Dict[str, VarInfo] = {}
    for name, nc_type, dimids in variables:
        shape = tuple(dims[d][1] for d in dimids)
        nelem = int(np.prod(shape, dtype=np.int64)) if shape else 1
        vsize = _pad4(nelem * _NP_OF_TYPE[nc_type].itemsize)
        begins.append(cur)
        layout[name] = VarInfo(name, tuple(dimids), nc_type, vsize, cur, shape)
        cur += vsize
    return build(begins), layout


# ----------------------------------------------------------------------------- decoding
class _Cursor:
    def __init__(self, buf: bytes):
        self.buf, self.pos = buf, 0

    def take(self, fmt: str):
        vals = struct.unpack_from(fmt, self.buf, self.pos)
        self.pos += struct.calcsize(fmt)
        return vals if len(vals) > 1 else vals[0]

    def name(self) -> str:
        n = self.take(">q")
        s = self.buf[self.pos:self.pos + n].decode("utf-8")
        self.pos += _pad4(n)
        return s

    def atts(self) -> Dict[str, np.ndarray]:
        tag, n = self.take(">iq")
        if tag == 0:
            if n != 0:
                raise ValueError("CDF-5: ABSENT list with non-zero count")
            return {}
        if tag != NC_ATTRIBUTE:
            raise ValueError(f"CDF-5: expected attribute list tag, got {tag:#x}")
        out = {}
        for _ in range(n):
            nm = self.name()
            nc_type, nelem = self.take(">iq")
            dt = _NP_OF_TYPE[nc_type]
            nbytes = nelem * dt.itemsize
            out[nm] = np.frombuffer(self.buf, dt, nelem, self.pos).copy()
            self.pos += _pad4(nbytes)
        return out


def decode_header(buf: bytes) -> Header:
    """Parse a CDF-5 header from the first bytes of a file (raises ValueError if malformed)."""
    if len(buf) < 12 or buf[:3] != b"CDF":
        raise ValueError("not a netCDF classic file (bad magic)")
    if buf[3] != 5:
        raise ValueError(f"unsupported netCDF classic version {buf[3]} (only CDF-5 / 64BIT_DATA)")
    c = _Cursor(buf)
    c.pos = 4
    numrecs = c.take(">q")
    tag, n = c.take(">iq")
    dims: List[Tuple[str, int]] = []
    if tag == NC_DIMENSION:
        for _ in range(n):
            nm = c.name()
            dims.append((nm, c.take(">q")))
    elif tag != 0 or n != 0:
        raise ValueError(f"CDF-5: bad dim_list tag {tag:#x}")
    gatts = c.atts()
    tag, n = c.take(">iq")
    variables: Dict[str, VarInfo] = {}
    if tag == NC_VARIABLE:
        for _ in range(n):
            nm = c.name()
            nd = c.take(">q")
            dimids = tuple(c.take(">q") for _ in range(nd))
            vatts = c.atts()
            nc_type, vsize, begin = c.take(">iqq")
            if nc_type not in _NP_OF_TYPE:
                raise ValueError(f"CDF-5: unknown nc_type {nc_type} for variable {nm}")
            for d in dimids:
                if not 0 <= d < len(dims):
                    raise ValueError(f"CDF-5: variable {nm} references dimension id {d}")
            shape = tuple(dims[d][1] for d in dimids)
            if any(s == 0 for s in shape[:1]) and numrecs:
                raise ValueError("CDF-5: record variables are not supported")
            variables[nm] = VarInfo(nm, dimids, nc_type, vsize, begin, shape, vatts)
    elif tag != 0 or n != 0:
        raise ValueError(f"CDF-5: bad var_list tag {tag:#x}")
    return Header(numrecs, dims, gatts, variables, c.pos)


def read_header(path: str) -> Header:
    size = os.path.getsize(path)
    want = 8192
    with open(path, "rb") as f:
        while True:
            f.seek(0)
            buf = f.read(min(want, size))
            try:
                return decode_header(buf)
            except struct.error:
                if want >= size:
                    raise ValueError("CDF-5: truncated header")
                want *= 4


# ----------------------------------------------------------------------------- MNIST-shaped files
def write_mnist_cdf5(path: str, images: np.ndarray, labels: np.ndarray, data_align: int = 512,
                     chunk: int = 1 << 16) -> Dict[str, VarInfo]:
    """Write ``images u8[N,28,28]`` and ``labels u8[N]`` with the converter's schema
    (`mnist_to_netcdf.ipynb` c2:91-98: dims Y, X, idx in that order; images dimids (2,0,1);
    labels dimids (2,))."""
    images = np.ascontiguousarray(images, dtype=np.uint8)
    labels = np.ascontiguousarray(labels, dtype=np.uint8)
    n = images.shape[0]
    if images.shape[1:] != (28, 28) or labels.shape != (n,):
        raise ValueError("expected images u8[N,28,28] and labels u8[N]")
    dims = [("Y", 28), ("X", 28), ("idx", n)]
    hdr, layout = encode_header(dims, [("images", NC_UBYTE, (2, 0, 1)), ("labels", NC_UBYTE, (2,))],
                                data_align)
    if os.path.exists(path):
        os.remove(path)
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(b"\0" * (layout["images"].begin - len(hdr)))
        flat = images.reshape(n, 784)
        for s in range(0, n, chunk):
            f.write(flat[s:s + chunk].tobytes())
        f.write(b"\0" * (layout["labels"].begin - f.tell()))
        f.write(labels.tobytes())
        f.write(b"\0" * (layout["labels"].begin + layout["labels"].vsize - f.tell()))
    return layout


def read_mnist_cdf5(path: str) -> Tuple[np.ndarray, np.ndarray]:
    """Whole-file read: returns (images u8[N,28,28], labels u8[N])."""
    h = read_header(path)
    vi, vl = h.vars["images"], h.vars["labels"]
    n = vi.shape[0]
    images = np.fromfile(path, dtype=np.uint8, count=n * 784, offset=vi.begin).reshape(n, 28, 28)
    labels = np.fromfile(path, dtype=np.uint8, count=n, offset=vl.begin)
    return images, labels
