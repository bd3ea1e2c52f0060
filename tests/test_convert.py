"""IDX -> CDF-5 converter (the notebook's to_nc) — CPU only."""
import os
import struct

import numpy as np
import pytest

from pytorch_ddp_mnist_b200 import cdf5, convert, synth


def _write_idx(d, im, lb, img_name, lab_name):
    for n in (img_name, lab_name):
        os.makedirs(os.path.join(d, os.path.dirname(n)), exist_ok=True)
    with open(os.path.join(d, img_name), "wb") as f:
        f.write(struct.pack(">IIII", 2051, len(im), 28, 28) + im.tobytes())
    with open(os.path.join(d, lab_name), "wb") as f:
        f.write(struct.pack(">II", 2049, len(lb)) + lb.tobytes())


def test_idx_to_nc_roundtrip(tmp_path):
    d = str(tmp_path)
    tr, te = synth.synth(300, 3), synth.synth(100, 4)
    _write_idx(d, *tr, 'train-images-idx3-ubyte/train-images-idx3-ubyte', 'train-labels-idx1-ubyte/train-labels-idx1-ubyte')
    _write_idx(d, *te, 't10k-images-idx3-ubyte/t10k-images-idx3-ubyte', 't10k-labels-idx1-ubyte/t10k-labels-idx1-ubyte')
    convert.main(["--input_path", d, "--out_dir", d])
    for name, (im, lb) in (("train", tr), ("test", te)):
        a, b = cdf5.read_mnist_cdf5(os.path.join(d, f"mnist_{name}_images.nc"))
        assert np.array_equal(a, im) and np.array_equal(b, lb)
    # byte-identical to the direct writer (same schema, same order of definitions)
    ref = os.path.join(d, "ref.nc")
    cdf5.write_mnist_cdf5(ref, *tr)
    assert open(ref, "rb").read() == open(os.path.join(d, "mnist_train_images.nc"), "rb").read()


def test_idx_magic_checked(tmp_path):
    d = str(tmp_path)
    im, lb = synth.synth(4, 1)
    _write_idx(d, im, lb, "i", "l")
    with pytest.raises(ValueError, match="expected 2049"):
        convert.read_images_labels(os.path.join(d, "i"), os.path.join(d, "i"))
    with pytest.raises(ValueError, match="expected 2051"):
        convert.read_images_labels(os.path.join(d, "l"), os.path.join(d, "l"))
