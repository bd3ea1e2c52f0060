"""The C-ABI library loads without a GPU and exports every symbol the header declares; the host
side entry points (decode table, CDF-5 reader, argument validation) work on CPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle import mnist_oracle as O
from pytorch_ddp_mnist_b200 import cdf5, synth

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
    so = os.path.join(REPO, "pytorch_ddp_mnist_b200", "libddpmnist_b200.so")
    if not os.path.exists(so):
        import __graft_entry__ as g
        g.build()
    from pytorch_ddp_mnist_b200 import _lib
    return _lib


def test_header_symbols_all_exported_and_bound(L):
    hdr = open(os.path.join(REPO, "include", "ddpmnist_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ddpm_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(L.SIGNATURES), declared ^ set(L.SIGNATURES)
    for name in declared:
        assert hasattr(L.lib, name), name
    assert L.lib.ddpm_abi_version() == L.ABI_VERSION


def test_layout_constants_match_header(L):
    hdr = open(os.path.join(REPO, "include", "ddpmnist_b200.h")).read()
    consts = dict(re.findall(r"#define\s+(DDPM_[A-Z0-9_]+)\s+(\d+)", hdr))
    assert int(consts["DDPM_N_PARAMS"]) == L.N_PARAMS == O.N_PARAMS
    offs, pos = [], 0
    for shp in O.PARAM_SHAPES:
        offs.append(pos)
        pos += int(np.prod(shp))
    assert [int(consts[k]) for k in ("DDPM_OFF_W1", "DDPM_OFF_B1", "DDPM_OFF_W2", "DDPM_OFF_B2", "DDPM_OFF_W3")] == offs
    assert [L.OFF_W1, L.OFF_B1, L.OFF_W2, L.OFF_B2, L.OFF_W3] == offs


def test_decode_lut_host_bit_exact(L):
    buf = np.empty(256, np.float32)
    assert L.lib.ddpm_decode_lut_host(buf.ctypes.data) == 0
    assert np.array_equal(buf.view(np.uint32), O.decode_lut().view(np.uint32))
    assert L.lib.ddpm_decode_lut_host(None) != 0 and "null" in L.last_error()


def test_workspace_bytes_monotonic(L):
    sizes = [L.lib.ddpm_workspace_bytes(b) for b in (1, 12, 64, 65, 148, 600, 1024, 4096)]
    assert sizes == sorted(sizes) and sizes[0] > 0
    assert L.lib.ddpm_workspace_bytes(0) == 0
    assert all(s % 256 == 0 for s in sizes)


def test_validation_errors_need_no_gpu(L):
    # argument checks run before any CUDA call
    assert L.lib.ddpm_gather_decode(None, None, 0, None, 4, None, None, None, None) != 0
    assert "n_samples" in L.last_error()
    assert L.lib.ddpm_step_forward(None, None, None, 0, None, None, None, None, 0, 0, 0, 0, None, None, None, None) != 0
    assert "batch must be positive" in L.last_error()
    assert L.lib.ddpm_sgd_step(None, None, -1, 0.1, 1.0, None) != 0


def _open(L, path):
    h = C.c_void_p()
    rc = L.lib.ddpm_cdf5_open(path.encode(), C.byref(h))
    return rc, h


def test_cdf5_c_reader_matches_python(L, tmp_path):
    im, lb = synth.synth(1001, 5)
    p = str(tmp_path / "a.nc")
    cdf5.write_mnist_cdf5(p, im, lb, data_align=4096)
    rc, h = _open(L, p)
    assert rc == 0, L.last_error()
    nd, dims, t, es, beg = C.c_int(), (C.c_int64 * 8)(), C.c_int(), C.c_int(), C.c_int64()
    assert L.lib.ddpm_cdf5_inq_var(h, b"images", C.byref(nd), dims, C.byref(t), C.byref(es), C.byref(beg)) == 0
    assert (nd.value, list(dims[:3]), t.value, es.value, beg.value) == (3, [1001, 28, 28], 7, 1, 4096)
    assert L.lib.ddpm_cdf5_inq_var(h, b"labels", C.byref(nd), dims, C.byref(t), C.byref(es), C.byref(beg)) == 0
    assert (nd.value, dims[0], beg.value) == (1, 1001, 4096 + 1001 * 784)
    out = np.zeros((100, 28, 28), np.uint8)
    assert L.lib.ddpm_cdf5_read_rows(h, b"images", 901, 100, out.ctypes.data) == 0
    assert np.array_equal(out, im[901:])
    lab = np.zeros(1001, np.uint8)
    assert L.lib.ddpm_cdf5_read_rows(h, b"labels", 0, 1001, lab.ctypes.data) == 0
    assert np.array_equal(lab, lb)
    assert L.lib.ddpm_cdf5_read_rows(h, b"images", 1000, 2, out.ctypes.data) != 0 and "out of range" in L.last_error()
    assert L.lib.ddpm_cdf5_read_rows(h, b"images", 0, 0, None) == 0                      # empty read
    assert L.lib.ddpm_cdf5_inq_var(h, b"nope", None, None, None, None, None) != 0
    L.lib.ddpm_cdf5_close(h)


def test_cdf5_c_reader_rejects_bad_files(L, tmp_path):
    p = tmp_path / "x.nc"
    p.write_bytes(b"CDF\x02" + b"\0" * 100)
    rc, _ = _open(L, str(p))
    assert rc != 0 and "CDF-5" in L.last_error()
    p.write_bytes(b"garbage!")
    rc, _ = _open(L, str(p))
    assert rc != 0 and "magic" in L.last_error()
    rc, _ = _open(L, str(tmp_path / "missing.nc"))
    assert rc != 0 and "cannot open" in L.last_error()
    # header claims more data than the file holds
    im, lb = synth.synth(10, 1)
    q = str(tmp_path / "t.nc")
    cdf5.write_mnist_cdf5(q, im, lb)
    raw = open(q, "rb").read()
    open(q, "wb").write(raw[:2000])
    rc, _ = _open(L, q)
    assert rc != 0 and "outside the file" in L.last_error()


def test_philox_host_matches_spec():
    """Philox4x32-10 known-answer (Random123 kat_vectors): counter=0,key=0 and all-ones."""
    def philox(c, k):
        M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
        c, k = list(c), list(k)
        for _ in range(10):
            p0, p1 = M0 * c[0], M1 * c[2]
            c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xffffffff, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xffffffff]
            k = [(k[0] + W0) & 0xffffffff, (k[1] + W1) & 0xffffffff]
        return c
    assert philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
