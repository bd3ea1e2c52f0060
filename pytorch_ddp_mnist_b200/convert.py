"""IDX -> netCDF (CDF-5) converter: the reference's `mnist_to_netcdf.ipynb` as a script.

`read_images_labels` parses the big-endian IDX files (magic 2051 images / 2049 labels,
nb:c2:24-45); `to_nc` writes the file with the notebook's exact schema through the `pncpy` write
subset (nb:c2:83-104): dims Y=28, X=28, idx=N in that order, `images` NC_UBYTE (idx,Y,X), `labels`
NC_UBYTE (idx,), `enddef`, labels in one put, then the images.  The notebook issues one put per
image; here images go in slabs of 4096 rows (same bytes, ~1000x fewer calls).

    python -m pytorch_ddp_mnist_b200.convert --input_path . --out_dir .
expects train-images-idx3-ubyte/train-images-idx3-ubyte etc. under --input_path (nb:c2:62-65).
"""
from __future__ import annotations

import argparse
import os
import struct
from typing import Tuple

import numpy as np

from .compat import pncpy


def read_images_labels(images_filepath: str, labels_filepath: str) -> Tuple[np.ndarray, np.ndarray]:
    with open(labels_filepath, 'rb') as f:
        magic, size = struct.unpack(">II", f.read(8))
        if magic != 2049:
            raise ValueError('Magic number mismatch, expected 2049, got {}'.format(magic))
        labels = np.frombuffer(f.read(), dtype=np.uint8)
    with open(images_filepath, 'rb') as f:
        hdr = f.read(16)
        magic = struct.unpack(">I", hdr[:4])[0] if len(hdr) >= 4 else -1
        if magic != 2051:
            raise ValueError('Magic number mismatch, expected 2051, got {}'.format(magic))
        if len(hdr) < 16:
            raise ValueError('truncated IDX image header')
        _, n, rows, cols = struct.unpack(">IIII", hdr)
        data = np.frombuffer(f.read(), dtype=np.uint8)
    if rows != 28 or cols != 28:
        raise ValueError(f"expected 28x28 images, got {rows}x{cols}")
    if data.size != n * rows * cols or labels.size != size or size != n:
        raise ValueError("IDX payload length does not match its header")
    return data.reshape(n, rows, cols), labels


def to_nc(samples, labels, comm=None, out_file_path='mnist_train_images.nc', slab: int = 4096) -> None:
    samples = np.asarray(samples, dtype=np.uint8).reshape(len(labels), 28, 28)
    if os.path.exists(out_file_path):
        os.remove(out_file_path)
    with pncpy.File(out_file_path, comm=comm, mode="w", format="64BIT_DATA") as fnc:
        dim_y = fnc.def_dim("Y", 28)
        dim_x = fnc.def_dim("X", 28)
        dim_num = fnc.def_dim("idx", len(samples))
        v = fnc.def_var("images", pncpy.NC_UBYTE, (dim_num, dim_y, dim_x))
        v_label = fnc.def_var("labels", pncpy.NC_UBYTE, (dim_num,))
        fnc.enddef()
        v_label[:] = np.asarray(labels, dtype=np.uint8)
        for s in range(0, len(samples), slab):
            v[s:s + slab, :, :] = samples[s:s + slab]


def main(argv=None) -> None:
    ap = argparse.ArgumentParser(description="convert MNIST IDX files to the reference's netCDF layout")
    ap.add_argument("--input_path", default=".")
    ap.add_argument("--out_dir", default=".")
    a = ap.parse_args(argv)
    j = os.path.join
    sets = (("train", 'train-images-idx3-ubyte/train-images-idx3-ubyte', 'train-labels-idx1-ubyte/train-labels-idx1-ubyte'),
            ("test", 't10k-images-idx3-ubyte/t10k-images-idx3-ubyte', 't10k-labels-idx1-ubyte/t10k-labels-idx1-ubyte'))
    for name, im, lb in sets:
        x, y = read_images_labels(j(a.input_path, im), j(a.input_path, lb))
        print('=> ========= Converting %s Images ========= <=' % name.capitalize())
        to_nc(x, y, None, j(a.out_dir, f"mnist_{name}_images.nc"))


if __name__ == "__main__":
    main()
