"""Pins oracle/mnist_oracle.py against fixtures produced by the reference's own code
(oracle/gen_golden.py).  CPU only."""
import hashlib

import numpy as np
import pytest

from oracle import mnist_oracle as O

TOL_ACT = 2e-5      # fp32 restatement vs ATen sgemm: summation order only
# 100 free-running steps.  Not rounding noise alone: at step 8 of this trajectory one fc2
# pre-activation is -7.5e-9 in numpy's fp32 summation order and +4.5e-8 in float64 / ATen's order,
# so its ReLU gate flips and that step's W2/b2 gradients differ by 3e-2 relative.  Any fp32
# implementation with a different summation order can hit such a flip; SURVEY §8(c) budgets 5e-3.
TOL_TRAJ = 5e-3


def _relerr(a, b):
    return float(np.abs(a.astype(np.float64) - b).max() / max(np.abs(b).max(), 1e-30))


def _fro(a, b):
    """relative Frobenius error — the right norm for end-of-trajectory parameters, where a few
    elements carry the gate-flip differences described above"""
    b = b.astype(np.float64)
    return float(np.linalg.norm(a.astype(np.float64) - b) / max(np.linalg.norm(b), 1e-30))


def test_decode_lut_bit_exact(golden):
    k = golden["kats"]
    lut = O.decode_lut()
    assert lut.dtype == np.float32
    assert np.array_equal(lut.view(np.uint32), k["lut"].view(np.uint32))
    assert lut[0] == np.float32(-0.4242129623889923) and lut[255] == np.float32(2.821486711502075)
    # the reciprocal-multiply form is NOT the reference's arithmetic
    v = np.arange(256, dtype=np.float32)
    recip = (v * np.float32(1 / 255.0) - O.MEAN) * np.float32(1 / 0.3081)
    assert (recip != lut).sum() > 100


@pytest.mark.parametrize("n", [60000, 4000, 10])
@pytest.mark.parametrize("ws", [1, 2, 4, 8])
def test_sampler_shards_bit_exact(golden, n, ws):
    k = golden["kats"]
    for epoch in (0, 1):
        for rank in sorted({0, ws - 1}):
            sh = O.shard_indices(n, ws, rank, epoch)
            key = f"shard.n{n}.ws{ws}.e{epoch}.r{rank}"
            assert len(sh) == int(k[key + ".len"]) == -(-n // ws)
            assert np.array_equal(sh[:8], k[key + ".head"])
            assert np.array_equal(sh[-8:], k[key + ".tail"])
            assert np.array_equal(np.frombuffer(hashlib.sha256(sh.tobytes()).digest(), np.uint8),
                                  k[key + ".sha"])


def test_sampler_survey_kats():
    # SURVEY.md §8(a) R1 known answers (torch 2.11 CPU)
    assert O.shard_indices(60000, 1, 0)[:5].tolist() == [37542, 43087, 2992, 25810, 39634]
    assert O.shard_indices(60000, 2, 0)[:5].tolist() == [37542, 2992, 39634, 55768, 44310]
    assert O.shard_indices(60000, 8, 0)[:5].tolist() == [37542, 44310, 57971, 24611, 7787]
    # step counts of R4
    assert [len(O.batches(O.shard_indices(60000, ws, 0), 64)) for ws in (1, 2, 4, 8)] == [938, 469, 235, 118]
    assert [len(O.batches(O.shard_indices(60000, ws, 0), 64)[-1]) for ws in (1, 2, 4, 8)] == [32, 48, 24, 12]


def test_synth_checksums(golden, synth_train, synth_test):
    k = golden["kats"]
    for name, (im, lb) in (("train", synth_train), ("test", synth_test)):
        assert np.array_equal(np.frombuffer(hashlib.sha256(im.tobytes()).digest(), np.uint8),
                              k[f"synth.{name}.images_sha"])
        assert np.array_equal(np.frombuffer(hashlib.sha256(lb.tobytes()).digest(), np.uint8),
                              k[f"synth.{name}.labels_sha"])


def test_cross_entropy_u8_targets(golden):
    k = golden["kats"]
    loss, dz = O.cross_entropy(k["ce.z"], k["ce.y"])
    assert k["ce.y"].dtype == np.uint8
    assert abs(float(loss) - float(k["ce.loss_u8"])) < 1e-6
    assert float(k["ce.loss_u8"]) == float(k["ce.loss_i64"])
    assert abs(dz.sum()) < 1e-6


def _params(g, prefix):
    return {k: g[f"{prefix}.{k}"] for k in O.PARAM_KEYS}


def test_step0_against_reference(golden, synth_train):
    g = golden["traj_ws1"]
    im, lb = synth_train
    idx0 = g["idx"][0]
    assert np.array_equal(idx0, O.shard_indices(60000, 1, 0)[:64])
    x, y = O.gather_decode(im, lb, idx0)
    assert np.array_equal(x.view(np.uint32), g["x0"].view(np.uint32))          # bit-exact input
    assert np.array_equal(y, g["y0"]) and str(g["y0_dtype"]) == "torch.uint8"
    p0 = _params(g, "params0")
    mask = O.unpack_masks(g["masks"])[0]
    p1, loss, grads, a = O.train_step(p0, x, y, mask)
    for name, ref in (("z1", "acts0.0"), ("h1", "acts0.1"), ("h1d", "acts0.2"), ("z2", "acts0.3"),
                      ("h2", "acts0.4"), ("z3", "acts0.5")):
        assert _relerr(a[name], g[ref]) < TOL_ACT, name
    assert abs(float(loss) - float(g["losses"][0])) < 1e-6
    for k in O.PARAM_KEYS:
        assert _relerr(grads[k], g[f"grads0.{k}"]) < TOL_ACT, k
    # SGD is one FMA per element: bit-exact from the reference's own gradients
    p1_ref = O.sgd_step(p0, _params(g, "grads0"))
    for k in O.PARAM_KEYS:
        assert np.array_equal(p1_ref[k].view(np.uint32), g[f"params1.{k}"].view(np.uint32)), k


def test_trajectory_ws1(golden, synth_train):
    g = golden["traj_ws1"]
    im, lb = synth_train
    p = _params(g, "params0")
    masks = O.unpack_masks(g["masks"])
    shard = O.shard_indices(60000, 1, 0)
    losses = []
    for t in range(100):
        idx = shard[t * 64:(t + 1) * 64]
        assert np.array_equal(idx, g["idx"][t])
        x, y = O.gather_decode(im, lb, idx)
        p, loss, _, _ = O.train_step(p, x, y, masks[t])
        losses.append(loss)
    assert np.abs(np.array(losses) - g["losses"]).max() < TOL_TRAJ * np.abs(g["losses"]).max()
    for k in O.PARAM_KEYS:
        assert _fro(p[k], g[f"paramsN.{k}"]) < TOL_TRAJ, k


@pytest.mark.parametrize("ws", [2, 4, 8])
def test_trajectory_ddp(golden, synth_train, ws):
    """The oracle's data-parallel step against the reference's own gloo runs at 2, 4 and 8 ranks (mp:439)."""
    g = golden[f"traj_ws{ws}"]
    im, lb = synth_train
    p = _params(g, "params0")
    masks = [O.unpack_masks(g[f"rank{r}.masks"]) for r in range(ws)]
    shards = [O.shard_indices(60000, ws, r) for r in range(ws)]
    assert int(g["rank0.shard_len"]) == 60000 // ws
    for t in range(20):
        batch = []
        for r in range(ws):
            idx = shards[r][t * 64:(t + 1) * 64]
            assert np.array_equal(idx, g[f"rank{r}.idx"][t])
            x, y = O.gather_decode(im, lb, idx)
            batch.append((x, y, masks[r][t]))
        p_new, loss0, grads, _ = O.train_step(p, *batch[0], peers=batch[1:])
        assert abs(float(loss0) - float(g["rank0.losses"][t])) < TOL_TRAJ
        for r in range(1, ws):
            lr_ = O.cross_entropy(O.forward(p, batch[r][0], batch[r][2])["z3"], batch[r][1])[0]
            assert abs(float(lr_) - float(g[f"rank{r}.losses"][t])) < TOL_TRAJ
        if t == 0:
            for k in O.PARAM_KEYS:
                assert _relerr(grads[k], g[f"grads0.{k}"]) < TOL_ACT, k
        p = p_new
    for k in O.PARAM_KEYS:
        assert _fro(p[k], g[f"paramsN.{k}"]) < TOL_TRAJ, k


def test_flatten_roundtrip(golden):
    g = golden["traj_ws1"]
    p = _params(g, "params0")
    flat = O.flatten_params(p)
    assert flat.size == O.N_PARAMS == 118272 and flat.nbytes == 473088
    q = O.unflatten_params(flat)
    assert all(np.array_equal(p[k], q[k]) for k in O.PARAM_KEYS)
