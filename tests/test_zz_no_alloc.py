"""ddpm_step_forward_static_scratch: the forward/eval entry point for callers that cannot allocate.

Some hosts give a plug-in input and output buffers and nothing else (a closed memory pool, a
"no cudaMalloc after start-up" rule because cudaMalloc synchronises the device, a fixed memory
budget).  They have no workspace to pass to ddpm_step_forward, so this variant keeps its scratch
in a fixed device array inside the library.  The test calls it through the C ABI and checks logits
and loss against the workspace-taking entry point, twice in a row with different inputs — the
second call runs on scratch that still holds the first call's contents.

Collected last on purpose (file name), so a run that stops at the first failure has already run
every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

TOL_F32 = 2e-5          # same bound the other GPU parity tests use


def test_static_scratch_covers_its_stated_limit():
    """CPU-side: the built-in array (3,490,816 B) is large enough for the stated 128-row limit."""
    from pytorch_ddp_mnist_b200 import _lib
    assert _lib.lib.ddpm_workspace_forward_bytes(_lib.STATIC_SCRATCH_MAX_BATCH) <= 3490816


@pytest.mark.gpu
def test_static_scratch_forward_matches_workspace_forward():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from pytorch_ddp_mnist_b200 import _lib, ops, synth
    from pytorch_ddp_mnist_b200._lib import lib, ptr
    dev = torch.device("cuda:0")
    im, lb = synth.synth(4096, 77)
    images = torch.from_numpy(im.reshape(-1, 784)).to(dev)
    labels = torch.from_numpy(lb).to(dev)
    g = torch.Generator().manual_seed(3)
    params = ((torch.rand(_lib.N_PARAMS, generator=g) - 0.5) * 0.1).to(dev)
    stream = torch.cuda.current_stream().cuda_stream

    def both(batch, train, seed):
        rng = np.random.default_rng(seed)
        idx = torch.from_numpy(rng.integers(0, 4096, batch).astype(np.int64)).to(dev)
        mask = torch.from_numpy((rng.random((batch, 128)) < 0.8).astype(np.uint8)).to(dev) if train else None
        # existing entry point, caller-owned workspace
        ws = ops.Workspace(batch, dev)
        ref_logits = torch.empty(batch, 10, device=dev)
        ref_loss = torch.zeros(2, device=dev)
        ops.step_forward(params, ws, batch, images=images, labels=labels, idx=idx, mask=mask,
                         logits=ref_logits, loss=ref_loss)
        # new entry point: same inputs and outputs, no workspace argument
        logits = torch.empty(batch, 10, device=dev)
        loss = torch.zeros(2, device=dev)
        rc = lib.ddpm_step_forward_static_scratch(ptr(params), ptr(images), ptr(labels), 4096, ptr(idx), None,
                                                  None, ptr(mask), 0, 0, batch, 0, ptr(logits), ptr(loss), stream)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        scale = max(float(ref_logits.abs().max()), 1e-30)
        assert float((logits - ref_logits).abs().max()) / scale < TOL_F32
        assert abs(float(loss[0]) - float(ref_loss[0])) < TOL_F32 * max(1.0, abs(float(ref_loss[0])))
        assert abs(float(loss[1]) - float(ref_loss[1])) < TOL_F32

    both(64, train=True, seed=1)        # first call on this device also resolves the array's address
    both(33, train=False, seed=2)       # second call: ragged batch, eval mode, reused scratch
    both(128, train=True, seed=3)       # the stated maximum

    # beyond the maximum, or asking it to prepare a backward pass, fails with a clear message
    idx = torch.zeros(129, dtype=torch.int64, device=dev)
    rc = lib.ddpm_step_forward_static_scratch(ptr(params), ptr(images), ptr(labels), 4096, ptr(idx), None, None,
                                              None, 0, 0, 129, 0, None, None, stream)
    assert rc != 0 and "exceeds the built-in scratch limit of 128" in _lib.last_error()
    rc = lib.ddpm_step_forward_static_scratch(ptr(params), ptr(images), ptr(labels), 4096, ptr(idx), None, None,
                                              None, 0, 0, 64, _lib.FWD_PREP_BACKWARD, None, None, stream)
    assert rc != 0 and "forward/eval only" in _lib.last_error()
