"""`pncpy` (pnetcdf-python) stand-in: the subset of its API that the reference calls.

PnetCDF, pnetcdf-python and MPI are not in this image.  Putting
``pytorch_ddp_mnist_b200/compat`` on ``sys.path`` lets the reference scripts import unchanged:

  reader  `mnist_pnetcdf_cpu_mp.py:31-46`, `mnist_pnetcdf_cpu.py:33,47`
      pncpy.File(path, 'r', comm=comm); .begin_indep(); .variables[name].shape;
      .variables['images'][i] -> u8[28,28]; .variables['labels'].get_var(buf, index=(i,))
      and the collective spelling .get_var_all(buf, index=(i,))
  writer  `mnist_to_netcdf.ipynb` cell 2, lines 89-104
      with pncpy.File(path, comm=, mode="w", format="64BIT_DATA") as f:
          f.def_dim(name, n); f.def_var(name, pncpy.NC_UBYTE, (dims...)); f.enddef()
          var[:] = array; var[i, :, :] = array

Each element access is one positioned read/write on the file, like the independent-mode
`ncmpi_get_vara` calls it stands in for; ``comm`` is accepted and ignored (one process per file
handle).  The file layout is the real CDF-5 grammar (see ``pytorch_ddp_mnist_b200/cdf5.py``).
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Sequence, Tuple

import numpy as np

from pytorch_ddp_mnist_b200 import cdf5 as _cdf5
from pytorch_ddp_mnist_b200.cdf5 import (NC_BYTE, NC_CHAR, NC_SHORT, NC_INT, NC_FLOAT, NC_DOUBLE,  # noqa: F401
                                         NC_UBYTE, NC_USHORT, NC_UINT, NC_INT64, NC_UINT64)

__version__ = "0.0-b200-compat"


class Dimension:
    def __init__(self, name: str, size: int, dimid: int):
        self.name, self.size, self._dimid = name, int(size), dimid

    def __len__(self) -> int:
        return self.size


class Variable:
    def __init__(self, f: "File", info: _cdf5.VarInfo):
        self._f, self._info = f, info

    # ---- metadata
    @property
    def name(self) -> str:
        return self._info.name

    @property
    def shape(self) -> Tuple[int, ...]:
        return self._info.shape

    @property
    def dtype(self) -> np.dtype:
        return self._info.dtype

    @property
    def ndim(self) -> int:
        return len(self._info.shape)

    # ---- hyperslab helpers: (start, count) over leading index + full trailing dims
    def _slab(self, key) -> Tuple[int, Tuple[int, ...]]:
        """Resolve a basic index into (byte offset, result shape); the region must be contiguous."""
        shape = self.shape
        if not isinstance(key, tuple):
            key = (key,)
        if any(k is Ellipsis for k in key):
            i = key.index(Ellipsis)
            key = key[:i] + (slice(None),) * (len(shape) - len(key) + 1) + key[i + 1:]
        key = key + (slice(None),) * (len(shape) - len(key))
        if len(key) != len(shape):
            raise IndexError("too many indices for variable " + self.name)
        start, out_shape, seen_partial = [], [], False
        for k, n in zip(key, shape):
            if isinstance(k, (int, np.integer)):
                k = int(k)
                if k < 0:
                    k += n
                if not 0 <= k < n:
                    raise IndexError(f"index {k} out of range for dimension of size {n}")
                if seen_partial:
                    raise IndexError("non-contiguous hyperslab")
                start.append(k)
            elif isinstance(k, slice):
                lo, hi, st = k.indices(n)
                if st != 1:
                    raise IndexError("strided access is not supported by the pncpy stand-in")
                full = (lo == 0 and hi == n)
                if seen_partial and not full:
                    raise IndexError("non-contiguous hyperslab")
                if not full or out_shape:
                    pass
                start.append(lo)
                out_shape.append(max(hi - lo, 0))
                if not full:
                    seen_partial = True
            else:
                raise IndexError("unsupported index type " + type(k).__name__)
        strides, acc = [], 1
        for n in reversed(shape):
            strides.append(acc)
            acc *= n
        strides.reverse()
        elem_off = sum(s * st for s, st in zip(start, strides))
        return self._info.begin + elem_off * self.dtype.itemsize, tuple(out_shape)

    def __getitem__(self, key) -> np.ndarray:
        off, shp = self._slab(key)
        n = int(np.prod(shp, dtype=np.int64)) if shp else 1
        raw = os.pread(self._f._fd, n * self.dtype.itemsize, off)
        if len(raw) != n * self.dtype.itemsize:
            raise IOError("short read from " + self._f.path)
        arr = np.frombuffer(raw, self.dtype, n)
        return arr.reshape(shp).astype(self.dtype.newbyteorder("="), copy=True)

    def __setitem__(self, key, value) -> None:
        self._f._require_data_mode(write=True)
        off, shp = self._slab(key)
        arr = np.ascontiguousarray(np.broadcast_to(np.asarray(value), shp), dtype=self.dtype)
        os.pwrite(self._f._fd, arr.tobytes(), off)

    def _get(self, buff: np.ndarray, index: Optional[Sequence[int]]) -> None:
        if index is None:
            buff[...] = self[...]
        else:
            buff[...] = self[tuple(int(i) for i in index)]

    def get_var(self, buff: np.ndarray, index: Optional[Sequence[int]] = None, **_kw) -> None:
        """Independent-mode read into ``buff`` (filled in place), `mnist_pnetcdf_cpu_mp.py:46`."""
        self._get(buff, index)

    def get_var_all(self, buff: np.ndarray, index: Optional[Sequence[int]] = None, **_kw) -> None:
        """Collective spelling of :meth:`get_var`, `mnist_pnetcdf_cpu.py:47`."""
        self._get(buff, index)

    def put_var(self, data, index: Optional[Sequence[int]] = None, **_kw) -> None:
        if index is None:
            self[...] = data
        else:
            self[tuple(int(i) for i in index)] = data

    put_var_all = put_var


class File:
    def __init__(self, filename, mode: str = "r", format: Optional[str] = None, comm=None,
                 info=None, **_kw):
        self.path = os.fspath(filename)
        self.comm = comm
        self.dimensions: Dict[str, Dimension] = {}
        self.variables: Dict[str, Variable] = {}
        self._defs = []
        self._closed = False
        self._indep = False
        if mode in ("r", "a", "r+"):
            self._fd = os.open(self.path, os.O_RDONLY if mode == "r" else os.O_RDWR)
            self._writable = mode != "r"
            self._define_mode = False
            hdr = _cdf5.read_header(self.path)
            for i, (nm, n) in enumerate(hdr.dims):
                self.dimensions[nm] = Dimension(nm, n, i)
            for nm, info_ in hdr.vars.items():
                self.variables[nm] = Variable(self, info_)
            self.file_format = "64BIT_DATA"
        elif mode in ("w", "x"):
            if format not in (None, "64BIT_DATA", "NETCDF3_64BIT_DATA"):
                raise ValueError("the pncpy stand-in only writes CDF-5 (format='64BIT_DATA')")
            flags = os.O_RDWR | os.O_CREAT | (os.O_EXCL if mode == "x" else os.O_TRUNC)
            self._fd = os.open(self.path, flags, 0o644)
            self._writable = True
            self._define_mode = True
            self.file_format = "64BIT_DATA"
        else:
            raise ValueError(f"unsupported mode {mode!r}")

    # ---- define mode
    def def_dim(self, dimname: str, size: int = -1) -> Dimension:
        if not self._define_mode:
            raise RuntimeError("def_dim: not in define mode")
        if size is None or size < 0:
            raise ValueError("record (unlimited) dimensions are not supported")
        d = Dimension(dimname, size, len(self.dimensions))
        self.dimensions[dimname] = d
        return d

    createDimension = def_dim

    def def_var(self, varname: str, nc_dtype, dimensions=(), **_kw):
        if not self._define_mode:
            raise RuntimeError("def_var: not in define mode")
        if isinstance(dimensions, (Dimension, str)):
            dimensions = (dimensions,)
        ids = tuple((d if isinstance(d, Dimension) else self.dimensions[d])._dimid
                    for d in dimensions)
        self._defs.append((varname, int(nc_dtype), ids))
        v = Variable(self, _cdf5.VarInfo(varname, ids, int(nc_dtype), 0, 0,
                                         tuple(list(self.dimensions.values())[i].size for i in ids)))
        self.variables[varname] = v
        return v

    createVariable = def_var

    def enddef(self) -> None:
        if not self._define_mode:
            raise RuntimeError("enddef: not in define mode")
        dims = [(d.name, d.size) for d in self.dimensions.values()]
        hdr, layout = _cdf5.encode_header(dims, self._defs)
        os.pwrite(self._fd, hdr, 0)
        end = max((v.begin + v.vsize for v in layout.values()), default=len(hdr))
        os.ftruncate(self._fd, end)           # fixed-size variables: file length is known now
        for nm, info_ in layout.items():
            self.variables[nm]._info = info_
        self._define_mode = False

    def redef(self) -> None:
        raise NotImplementedError("redef is not part of the subset the reference uses")

    # ---- data mode
    def _require_data_mode(self, write: bool = False) -> None:
        if self._closed:
            raise RuntimeError("file is closed")
        if self._define_mode:
            raise RuntimeError("file is in define mode; call enddef() first")
        if write and not self._writable:
            raise PermissionError("file opened read-only")

    def begin_indep(self) -> None:
        self._indep = True

    def end_indep(self) -> None:
        self._indep = False

    def sync(self) -> None:
        if self._writable and not self._closed:
            os.fsync(self._fd)

    def close(self) -> None:
        if not self._closed:
            if self._define_mode:
                self.enddef()
            os.close(self._fd)
            self._closed = True

    def __enter__(self) -> "File":
        return self

    def __exit__(self, *exc) -> None:
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
