"""Drop-in surface on the GPU: loaders, create_model/loss/optimizer with the reference's loop body,
and `main()` end to end against what the UNMODIFIED reference `main()` produced (tests/golden/main_e2e.npz)."""
import io
import contextlib
import os
import re

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import mnist_oracle as O           # noqa: E402  (checker only)
from oracle.determinism import mask_for_step   # noqa: E402

TOL_TRAJ = 5e-3


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def _fro(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


@pytest.fixture(scope="module")
def nc_dir(tmp_path_factory):
    from pytorch_ddp_mnist_b200 import synth
    d = tmp_path_factory.mktemp("nc_e2e")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "main_e2e.npz"))
    synth.write_default_files(str(d), int(g["train_n"]), int(g["test_n"]))
    return str(d)


def test_resident_dataset_from_file_and_loader(dev, nc_dir):
    from pytorch_ddp_mnist_b200 import synth
    from pytorch_ddp_mnist_b200.dataset import DeviceBatchLoader, MNISTNetCDF, ResidentDataset
    im, lb = synth.synth(4000, synth.TRAIN_SEED)
    r = ResidentDataset.from_file(os.path.join(nc_dir, "mnist_train_images.nc"), dev, chunk_rows=1024)
    assert np.array_equal(r.images.cpu().numpy().reshape(-1, 28, 28), im)
    assert np.array_equal(r.labels.cpu().numpy(), lb)
    with contextlib.redirect_stdout(io.StringIO()):
        ds = MNISTNetCDF(nc_dir, is_train=True, transforms=True, device=dev)
    assert len(ds) == 4000
    x, y = ds[17]                                         # mp:39-49 contract
    assert x.shape == (1, 28, 28) and x.dtype == torch.float32 and y.dtype == np.uint8 and y.shape == ()
    assert np.array_equal(x.cpu().numpy().reshape(-1).view(np.uint32), O.decode_lut()[im[17].reshape(-1)].view(np.uint32))
    assert y == lb[17]
    from torch.utils.data import DistributedSampler
    from pytorch_ddp_mnist_b200.dataset import _IndexOnly
    s = DistributedSampler(_IndexOnly(4000), num_replicas=2, rank=1, shuffle=True, seed=42)
    loader = DeviceBatchLoader(ds, batch_size=64, sampler=s)
    s.set_epoch(0)
    shard = O.shard_indices(4000, 2, 1, 0)
    batches = list(loader)
    assert len(batches) == len(loader) == 32 and batches[-1][0].shape[0] == 2000 - 31 * 64
    xb, yb = batches[3]
    xr, yr = O.gather_decode(im, lb, shard[3 * 64:4 * 64])
    assert xb.shape == (64, 1, 28, 28) and yb.dtype == torch.uint8
    assert np.array_equal(xb.cpu().numpy().reshape(64, 784).view(np.uint32), xr.view(np.uint32))
    assert np.array_equal(yb.cpu().numpy(), yr)


def test_backward_after_a_second_forward_raises(dev):
    """The drop-in model keeps its activations in ONE workspace: a backward that follows a later forward of the
    same model must raise instead of returning the gradients of the wrong batch."""
    from pytorch_ddp_mnist_b200.model import CrossEntropyLoss, create_model
    torch.manual_seed(0)
    model = create_model().to(dev)
    model.train()
    loss = CrossEntropyLoss()
    x1, x2 = torch.randn(8, 784, device=dev), torch.randn(8, 784, device=dev)
    y = torch.randint(0, 10, (8,), device=dev, dtype=torch.uint8)
    l1 = loss(model(x1), y)
    with torch.no_grad():
        model(x2)                                            # e.g. a validation forward inside the step
    with pytest.raises(RuntimeError, match="overwritten by a later forward"):
        l1.backward()
    l3 = loss(model(x1), y)                                  # the normal order still works
    l3.backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())


def test_reference_loop_body_with_dropin_objects(dev, golden, synth_train):
    """mp:453-466 written out with this package's create_model / CrossEntropyLoss / SGD."""
    from pytorch_ddp_mnist_b200.dataset import DeviceBatchLoader, ResidentDataset
    from pytorch_ddp_mnist_b200.model import SGD, CrossEntropyLoss, create_model
    g = golden["traj_ws1"]
    im, lb = synth_train
    res = ResidentDataset.from_arrays(im, lb, dev)
    torch.manual_seed(0)
    model = create_model()
    assert list(model.state_dict().keys()) == list(O.PARAM_KEYS)
    for k in O.PARAM_KEYS:                                   # same init as the reference under the same seed
        assert np.array_equal(model.state_dict()[k].numpy(), g[f"params0.{k}"]), k
    device = dev
    model = model.to(device)
    model.set_mask_provider(lambda t, b: mask_for_step(t, b))
    optimizer = SGD(model.parameters(), lr=0.01)
    loss = CrossEntropyLoss()
    from torch.utils.data import DistributedSampler
    from pytorch_ddp_mnist_b200.dataset import _IndexOnly
    sampler = DistributedSampler(_IndexOnly(60000), num_replicas=1, rank=0, shuffle=True, seed=42)
    train_loader = DeviceBatchLoader(res, batch_size=64, sampler=sampler)
    model.train()
    train_loader.sampler.set_epoch(0)
    epoch_loss = 0
    losses = []
    for n, (x, y) in enumerate(train_loader):
        if n == 30:
            break
        x = x.to(device, non_blocking=True)
        y = y.to(device, non_blocking=True)
        x = x.view(x.shape[0], -1)
        optimizer.zero_grad()
        y_hat = model(x)
        batch_loss = loss(y_hat, y)
        batch_loss.backward()
        optimizer.step()
        batch_loss_scalar = batch_loss.item()
        epoch_loss += batch_loss_scalar / x.shape[0]
        losses.append(batch_loss_scalar)
    assert np.abs(np.array(losses[:5]) - g["losses"][:5]).max() < 1e-5
    assert np.abs(np.array(losses) - g["losses"][:30]).max() < TOL_TRAJ * 2.4
    # eval mode + no_grad (mp:469-477)
    with torch.no_grad():
        model.eval()
        x, y = next(iter(train_loader))
        v = loss(model(x.view(x.shape[0], -1)), y).item()
    sd = {k: t.detach().cpu().numpy() for k, t in model.state_dict().items()}
    xr, yr = O.gather_decode(im, lb, O.shard_indices(60000, 1, 0)[:64])
    assert abs(v - float(O.eval_loss(sd, xr, yr))) < 2e-5
    with pytest.raises(RuntimeError, match="no CPU path"):
        model(torch.zeros(2, 784))


def test_main_end_to_end_matches_unmodified_reference(dev, golden, nc_dir):
    """`main()` on the engine vs the reference's own `main()` (2 epochs, 4000/1000 samples,
    batch 64, short last batches, shuffled validation loader): same printed lines to 4 decimals
    within tolerance, same final state_dict."""
    from pytorch_ddp_mnist_b200 import train
    g = golden["main_e2e"]
    config = train.configure(["--batch_size", str(int(g["batch"])), "--n_epochs", str(int(g["epochs"]))])
    config["trainer"].update(rank=0, world_size=1, device=dev)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        train_loader, test_loader = train.create_data_loaders(config, None, dataset_loc=nc_dir, device=dev)
        torch.manual_seed(0)
        model = train.main(train.create_model(), train_loader, test_loader, config,
                           mask_fn=lambda t, b: mask_for_step(t, b))
    lines = re.findall(r"Epoch=(\d+), train_loss=([0-9.]+), val_loss=([0-9.]+)", buf.getvalue())
    got = np.array([[float(a), float(b), float(c)] for a, b, c in lines])
    want = g["epoch_lines"]
    assert got.shape == want.shape == (2, 3)
    assert np.abs(got[:, 1:] - want[:, 1:]).max() <= TOL_TRAJ * np.abs(want[:, 1:]).max() + 1e-4, (got, want)
    sd = model.state_dict()
    assert list(sd.keys()) == list(O.PARAM_KEYS)
    for k in O.PARAM_KEYS:
        assert _fro(sd[k].cpu().numpy(), g[f"state.{k}"]) < TOL_TRAJ, k


def test_engine_graph_replay_equals_eager(dev, golden, synth_train):
    from pytorch_ddp_mnist_b200 import engine as E
    from pytorch_ddp_mnist_b200.dataset import ResidentDataset
    g = golden["traj_ws1"]
    im, lb = synth_train
    res = ResidentDataset.from_arrays(im, lb, dev)
    sd = {k: torch.from_numpy(g[f"params0.{k}"]) for k in O.PARAM_KEYS}
    idx = torch.from_numpy(g["idx"][:20].reshape(-1)).to(dev)
    outs = []
    for graphed in (False, True):
        eng = E.TrainEngine(dev, 64)
        eng.load_state_dict(sd)
        batches = list(torch.split(idx.clone(), 64))
        if graphed:
            gr = eng.capture_steps(res, batches)
            eng.load_state_dict(sd)                 # capture ran nothing; make sure
            eng.zero_epoch_loss()
            gr.replay()
        else:
            for t, b in enumerate(batches):
                eng.train_step(res, b, philox_step=t)
        torch.cuda.synchronize()
        outs.append((eng.params.clone(), float(eng.loss[1])))
    assert torch.equal(outs[0][0], outs[1][0]) and abs(outs[0][1] - outs[1][1]) < 1e-6


def test_checkpoint_roundtrip(dev, tmp_path, golden):
    from pytorch_ddp_mnist_b200 import engine as E
    g = golden["traj_ws1"]
    eng = E.TrainEngine(dev, 8)
    eng.load_state_dict({k: torch.from_numpy(g[f"paramsN.{k}"]) for k in O.PARAM_KEYS})
    p = str(tmp_path / "model.pt")
    torch.save(eng.state_dict(), p)                          # mp:510-511
    sd = torch.load(p)
    assert list(sd.keys()) == list(O.PARAM_KEYS)
    ref = torch.nn.Sequential(torch.nn.Linear(784, 128), torch.nn.ReLU(), torch.nn.Dropout(0.2),
                              torch.nn.Linear(128, 128), torch.nn.ReLU(), torch.nn.Linear(128, 10, bias=False))
    ref.load_state_dict(sd)                                   # loads into the reference's architecture
    eng2 = E.TrainEngine(dev, 8)
    eng2.load_state_dict(sd)
    assert torch.equal(eng.params, eng2.params)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_ddp_matches_reference(dev, world):
    """2 / 4 / 8 ranks vs the reference's gloo run at the same world size (tests/golden/traj_ws{2,4,8}.npz):
    NCCL between kernels, the fused in-kernel exchange (default and plain bucket exchange), the host-fed fused
    loop, the drop-in model under DistributedDataParallel, the slab + all-gather loader (tests/multigpu_check.py)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                          "--master-addr", "127.0.0.1", "--master-port", str(29731 + world),
                          os.path.join(here, "multigpu_check.py")], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "MULTIGPU_OK" in out.stdout


def _run_script(nc_dir, nproc, extra=()):
    import subprocess
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = os.path.join(repo, "mnist_pnetcdf_b200.py")
    if nproc == 1:
        cmd = [sys.executable, script]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
               "--master-addr", "127.0.0.1", "--master-port", "29751", script, "--parallel", "--wireup_method", "nccl-mpich"]
    cmd += ["--batch_size", "64", "--n_epochs", "2", "--seed", "0", *extra]
    out = subprocess.run(cmd, cwd=nc_dir, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-3000:]
    return out.stdout


def test_train_script_single_gpu(dev, nc_dir):
    """`python mnist_pnetcdf_b200.py --batch_size 64 --n_epochs 2` from a directory holding the .nc files:
    prints the reference's epoch lines, learns, and writes a model.pt the reference architecture loads."""
    out = _run_script(nc_dir, 1)
    lines = re.findall(r"Epoch=(\d+), train_loss=([0-9.]+), val_loss=([0-9.]+)", out)
    assert [l[0] for l in lines] == ["0", "1"]
    assert float(lines[1][1]) < float(lines[0][1]) and float(lines[1][2]) < float(lines[0][2])
    sd = torch.load(os.path.join(nc_dir, "model.pt"))
    assert list(sd.keys()) == list(O.PARAM_KEYS) and all(torch.isfinite(v).all() for v in sd.values())


def test_train_script_convnet(dev, nc_dir):
    """`python mnist_pnetcdf_b200.py --model json_convnet`: BASELINE.json's ConvNet recipe (Adadelta, StepLR, test
    accuracy) through the same script, loaders and sampler - learns the synthetic classes and writes convnet.pt."""
    out = _run_script(nc_dir, 1, extra=("--model", "json_convnet"))
    lines = re.findall(r"Epoch=(\d+), train_loss=([0-9.]+), val_loss=([0-9.]+), val_acc=([0-9.]+)", out)
    assert [l[0] for l in lines] == ["0", "1"], out[-2000:]
    assert float(lines[1][2]) < float(lines[0][2]) or float(lines[1][3]) >= 0.99
    assert float(lines[1][3]) > 0.9, lines
    sd = torch.load(os.path.join(nc_dir, "convnet.pt"))
    assert "conv1.weight" in sd and sd["fc1.weight"].shape == (128, 9216)
    assert all(torch.isfinite(v).all() for v in sd.values())


def test_train_script_two_gpus(dev, nc_dir):
    """Same script under torchrun on 2 GPUs (fused data-parallel step): every rank prints its epoch lines."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = _run_script(nc_dir, 2)
    lines = re.findall(r"Epoch=(\d+), train_loss=([0-9.]+), val_loss=([0-9.]+)", out)
    assert len(lines) == 4                                   # 2 epochs x 2 ranks (mp:484: every rank prints)
    vals = sorted(float(l[2]) for l in lines if l[0] == "1")
    assert abs(vals[0] - vals[1]) < 1e-4                     # identical replicas -> identical validation loss
