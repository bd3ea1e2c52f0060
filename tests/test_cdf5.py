"""CDF-5 writer/reader and the pncpy stand-in (host side, CPU only)."""
import os
import struct
import sys

import numpy as np
import pytest

from pytorch_ddp_mnist_b200 import cdf5, synth

COMPAT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                      "pytorch_ddp_mnist_b200", "compat")


@pytest.fixture()
def pncpy():
    sys.path.insert(0, COMPAT)
    try:
        import pncpy as m
        yield m
    finally:
        sys.path.remove(COMPAT)


def test_header_layout_matches_survey(tmp_path):
    im, lb = synth.synth(1000, 7)
    path = str(tmp_path / "a.nc")
    layout = cdf5.write_mnist_cdf5(path, im, lb)
    raw = open(path, "rb").read()
    assert raw[:4] == b"CDF\x05"
    h = cdf5.read_header(path)
    assert h.header_size == 252                       # 252 bytes for these names
    assert h.dims == [("Y", 28), ("X", 28), ("idx", 1000)]
    assert h.vars["images"].dimids == (2, 0, 1) and h.vars["labels"].dimids == (2,)
    assert h.vars["images"].nc_type == cdf5.NC_UBYTE == 7
    assert h.vars["images"].vsize == 784 * 1000 and h.vars["labels"].vsize == 1000
    assert h.vars["images"].begin == layout["images"].begin == 512
    assert h.vars["labels"].begin == 512 + 784000
    # row i of images is the 784 bytes at begin + 784*i
    i = 123
    assert raw[512 + 784 * i: 512 + 784 * (i + 1)] == im[i].tobytes()
    assert raw[h.vars["labels"].begin + i] == lb[i]
    assert len(raw) == h.vars["labels"].begin + 1000


def test_vsize_padding_and_begin_honoured(tmp_path):
    # N not a multiple of 4: labels vsize rounds up; a reader must use `begin`, not header end
    im, lb = synth.synth(1001, 9)
    path = str(tmp_path / "b.nc")
    cdf5.write_mnist_cdf5(path, im, lb, data_align=4096)
    h = cdf5.read_header(path)
    assert h.vars["images"].begin == 4096
    assert h.vars["labels"].vsize == 1004
    a, b = cdf5.read_mnist_cdf5(path)
    assert np.array_equal(a, im) and np.array_equal(b, lb)


def test_same_grammar_as_scipy_cdf2(tmp_path):
    """Cross-check the header grammar against an independent writer: scipy writes CDF-1/2, whose
    header is identical except that the fields CDF-5 widens to 64 bits are 32-bit there."""
    from scipy.io import netcdf_file
    p = str(tmp_path / "s.nc")
    f = netcdf_file(p, "w", version=2)
    f.createDimension("Y", 28); f.createDimension("X", 28); f.createDimension("idx", 5)
    v = f.createVariable("images", "b", ("idx", "Y", "X")); v[:] = 0
    w = f.createVariable("labels", "b", ("idx",)); w[:] = 0
    f.close()
    raw = open(p, "rb").read()
    assert raw[:4] == b"CDF\x02"
    # CDF-2: numrecs(4) tag(4) count(4) then name_len(4) name dimlen(4)
    numrecs, tag, cnt = struct.unpack_from(">iii", raw, 4)
    assert (tag, cnt) == (cdf5.NC_DIMENSION, 3)
    mine, _ = cdf5.encode_header([("Y", 28), ("X", 28), ("idx", 5)],
                                 [("images", cdf5.NC_BYTE, (2, 0, 1)), ("labels", cdf5.NC_BYTE, (2,))])
    tag5, cnt5 = struct.unpack_from(">iq", mine, 12)
    assert (tag5, cnt5) == (tag, cnt)
    # same token sequence: names appear in the same order in both headers
    for name in (b"Y", b"X", b"idx", b"images", b"labels"):
        assert raw.find(name) > 0 and mine.find(name) > 0
    order_s = [raw.find(n) for n in (b"idx", b"images", b"labels")]
    order_m = [mine.find(n) for n in (b"idx", b"images", b"labels")]
    assert order_s == sorted(order_s) and order_m == sorted(order_m)


def test_rejects_other_formats(tmp_path):
    p = tmp_path / "x.nc"
    p.write_bytes(b"CDF\x01" + b"\0" * 64)
    with pytest.raises(ValueError, match="CDF-5"):
        cdf5.read_header(str(p))
    p.write_bytes(b"\x89HDF\r\n\x1a\n" + b"\0" * 64)
    with pytest.raises(ValueError, match="magic"):
        cdf5.read_header(str(p))


def test_pncpy_reader_subset(tmp_path, pncpy):
    im, lb = synth.synth(300, 11)
    path = str(tmp_path / "mnist_train_images.nc")
    cdf5.write_mnist_cdf5(path, im, lb)
    nc = pncpy.File(path, "r", comm=None)
    nc.begin_indep()
    assert nc.variables["images"].shape == (300, 28, 28)          # mp:37
    for i in (0, 17, 299):
        img = np.array(nc.variables["images"][i])                 # mp:43
        assert img.dtype == np.uint8 and img.shape == (28, 28) and np.array_equal(img, im[i])
        buff = np.empty((), np.uint8)
        nc.variables["labels"].get_var(buff, index=(i,))          # mp:46
        assert buff == lb[i] and buff.shape == ()
        nc.variables["labels"].get_var_all(buff, index=(i,))      # pc:47
        assert buff == lb[i]
    with pytest.raises(IndexError):
        nc.variables["images"][300]
    nc.close()


def test_pncpy_writer_subset_matches_direct_writer(tmp_path, pncpy):
    """The notebook's `to_nc` call sequence (nb:c2:89-104) yields the same bytes as write_mnist_cdf5."""
    im, lb = synth.synth(50, 13)
    p1, p2 = str(tmp_path / "w.nc"), str(tmp_path / "d.nc")
    with pncpy.File(p1, comm=None, mode="w", format="64BIT_DATA") as fnc:
        dim_y = fnc.def_dim("Y", 28)
        dim_x = fnc.def_dim("X", 28)
        dim_num = fnc.def_dim("idx", 50)
        v = fnc.def_var("images", pncpy.NC_UBYTE, (dim_num, dim_y, dim_x))
        v_label = fnc.def_var("labels", pncpy.NC_UBYTE, (dim_num,))
        fnc.enddef()
        v_label[:] = np.array(list(lb), dtype=np.uint8)
        for idx, img in enumerate(im):
            v[idx, :, :] = img
    cdf5.write_mnist_cdf5(p2, im, lb)
    assert open(p1, "rb").read() == open(p2, "rb").read()


def test_streaming_synth_writer_equals_in_memory(tmp_path):
    p = str(tmp_path / "s.nc")
    synth.write_synth_file(p, 2500, 21)
    im, lb = synth.synth(2500, 21)
    a, b = cdf5.read_mnist_cdf5(p)
    assert np.array_equal(a, im) and np.array_equal(b, lb)
    assert np.bincount(im.ravel(), minlength=256).min() > 0      # every byte value occurs


def test_parallel_synth_writer_is_byte_identical(tmp_path):
    """bench.py --config scaled6m lets every rank (and a few worker processes per rank) write its share of the
    100,000-sample chunks of the synthetic CDF-5 file: the result must be the very file the serial writer produces."""
    import hashlib
    from pytorch_ddp_mnist_b200 import synth
    n, seed = 230_017, 99                                    # three chunks, the last one ragged
    a, b = str(tmp_path / "serial.nc"), str(tmp_path / "parallel.nc")
    synth.write_synth_file(a, n, seed)
    synth.write_synth_file_part(b, n, seed, 0, 1, True)      # lay the file out
    for part in (2, 0, 1):                                   # any order, any number of writers
        synth.write_synth_file_part(b, n, seed, part, 3, False)
    sha = lambda p: hashlib.sha256(open(p, "rb").read()).hexdigest()
    assert sha(a) == sha(b)
    import bench_scaled
    c = str(tmp_path / "procs.nc")
    synth.write_synth_file_part(c, n, seed, 0, 1, True)
    bench_scaled._make_file(c, n, seed, 0, 2, procs=2)       # rank 0 of 2, two worker processes
    bench_scaled._make_file(c, n, seed, 1, 2, procs=2)
    assert sha(a) == sha(c)
