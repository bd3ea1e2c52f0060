"""Seeded synthetic MNIST-shaped data (there is no network for the real IDX files).

Generator: chunks of 100,000 images; chunk ``c`` draws from
``numpy.random.default_rng([seed, c])``; ten fixed u8 class templates come from
``default_rng([TEMPLATE_SEED, 2**31])`` (shared by the train and test files); ``image = (template[label] >> 1) + U{0..128}`` so every byte value
0..255 occurs (exercises the whole decode table) and the classes are learnable (the loss moves).
numpy's PCG64 stream is stable across numpy versions, so the bytes are a function of (N, seed)
only; tests pin a checksum.

Files are written with the converter's schema and names (`mnist_to_netcdf.ipynb` c2:83,114).
"""
from __future__ import annotations

import argparse
import os
from typing import Iterator, Tuple

import numpy as np

from . import cdf5

CHUNK = 100_000
TRAIN_N, TRAIN_SEED = 60_000, 1234
TEST_N, TEST_SEED = 10_000, 4321
TEMPLATE_SEED = 1234          # train and test share the class templates, so val loss is meaningful


def templates(seed: int) -> np.ndarray:
    return np.random.default_rng([seed, 2 ** 31]).integers(0, 256, (10, 28, 28), dtype=np.uint8)


def synth_chunk(n: int, seed: int, c: int, tpl_half: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Chunk `c` (samples c*CHUNK ...) of ``synth(n, seed)``: chunks have independent generators, so they can be
    produced in any order and by several processes."""
    m = min(CHUNK, n - c * CHUNK)
    rng = np.random.default_rng([seed, c])
    labels = rng.integers(0, 10, m, dtype=np.uint8)
    noise = rng.integers(0, 129, (m, 28, 28), dtype=np.uint8)
    return tpl_half[labels] + noise, labels


def synth_chunks(n: int, seed: int, template_seed: int = TEMPLATE_SEED
                 ) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
    tpl = templates(template_seed) >> 1
    for c in range(-(-n // CHUNK)):
        yield synth_chunk(n, seed, c, tpl)


def synth(n: int, seed: int) -> Tuple[np.ndarray, np.ndarray]:
    """Return (images u8[n,28,28], labels u8[n])."""
    imgs = np.empty((n, 28, 28), np.uint8)
    labs = np.empty((n,), np.uint8)
    pos = 0
    for im, lb in synth_chunks(n, seed):
        imgs[pos:pos + len(lb)] = im
        labs[pos:pos + len(lb)] = lb
        pos += len(lb)
    return imgs, labs


def write_synth_file_part(path: str, n: int, seed: int, part: int, nparts: int, create: bool) -> None:
    """One of `nparts` cooperating writers of the same file as write_synth_file (chunks part, part+nparts, ...).
    The writer with `create=True` must run first (it writes the header and sizes the file); the others then fill
    their chunks in place.  The result is byte-identical to write_synth_file's."""
    dims = [("Y", 28), ("X", 28), ("idx", n)]
    hdr, layout = cdf5.encode_header(
        dims, [("images", cdf5.NC_UBYTE, (2, 0, 1)), ("labels", cdf5.NC_UBYTE, (2,))])
    if create:
        if os.path.exists(path):
            os.remove(path)
        with open(path, "wb") as f:
            f.write(hdr)
            f.truncate(layout["labels"].begin + layout["labels"].vsize)
        return
    tpl = templates(TEMPLATE_SEED) >> 1
    with open(path, "r+b") as f:
        for c in range(part, -(-n // CHUNK), nparts):
            im, lb = synth_chunk(n, seed, c, tpl)
            f.seek(layout["images"].begin + c * CHUNK * 784)
            f.write(im.tobytes())
            f.seek(layout["labels"].begin + c * CHUNK)
            f.write(lb.tobytes())


def write_synth_file(path: str, n: int, seed: int) -> None:
    """Stream ``synth(n, seed)`` to a CDF-5 file without holding all of it (n may be 6,000,000)."""
    dims = [("Y", 28), ("X", 28), ("idx", n)]
    hdr, layout = cdf5.encode_header(
        dims, [("images", cdf5.NC_UBYTE, (2, 0, 1)), ("labels", cdf5.NC_UBYTE, (2,))])
    if os.path.exists(path):
        os.remove(path)
    with open(path, "wb") as f:
        f.write(hdr)
        f.truncate(layout["labels"].begin + layout["labels"].vsize)
        ipos, lpos = layout["images"].begin, layout["labels"].begin
        for im, lb in synth_chunks(n, seed):
            f.seek(ipos)
            f.write(im.tobytes())
            ipos += im.size
            f.seek(lpos)
            f.write(lb.tobytes())
            lpos += lb.size


def write_default_files(root: str, train_n: int = TRAIN_N, test_n: int = TEST_N,
                        train_seed: int = TRAIN_SEED, test_seed: int = TEST_SEED) -> Tuple[str, str]:
    os.makedirs(root, exist_ok=True)
    tr = os.path.join(root, "mnist_train_images.nc")
    te = os.path.join(root, "mnist_test_images.nc")
    write_synth_file(tr, train_n, train_seed)
    write_synth_file(te, test_n, test_seed)
    return tr, te


def main(argv=None) -> None:
    ap = argparse.ArgumentParser(description="write seeded synthetic MNIST-shaped CDF-5 files")
    ap.add_argument("--root", default=".")
    ap.add_argument("--train_n", type=int, default=TRAIN_N)
    ap.add_argument("--test_n", type=int, default=TEST_N)
    ap.add_argument("--train_seed", type=int, default=TRAIN_SEED)
    ap.add_argument("--test_seed", type=int, default=TEST_SEED)
    a = ap.parse_args(argv)
    for p in write_default_files(a.root, a.train_n, a.test_n, a.train_seed, a.test_seed):
        print("=> wrote", p, os.path.getsize(p), "bytes")


if __name__ == "__main__":
    main()
