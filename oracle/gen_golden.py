#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by running the REAL reference code.

TEST INFRASTRUCTURE — runs only in the build container (needs /root/reference, which does not
exist on the GPU box).  The reference ships no tests and no recorded numbers, so
every pin is produced here by importing `/root/reference/mnist_pnetcdf_cpu_mp.py` UNCHANGED, with
two stand-ins on sys.path for packages that are absent from this image (`pncpy` -> the repo's real
CDF-5 reader, `mpi4py` -> singleton world; see pytorch_ddp_mnist_b200/compat), on the seeded
synthetic files of SURVEY.md §8(d).  The reference seeds nothing but the sampler, so the harness
adds determinism without touching reference code:

  * `torch.manual_seed(S)` before the reference's `create_model()`;
  * `nn.Dropout.forward` is replaced by a mask-injecting form: step `t` uses
    `bernoulli(0.8)` from `torch.Generator().manual_seed(MASK_SEED0 + t)`; forward value and
    gradient are `x * mask * 1.25`, identical to ATen's native_dropout for a given mask;
  * one OpenMP thread count, `torch.use_deterministic_algorithms(True)`.

Fixtures (all small, .npz, committed):
  traj_ws1.npz   ws=1, B=64, 100 steps of the reference loop body (`mp:453-466`) driven by the
                 reference's own `create_data_loaders()` + `create_model()` + DDP + SGD + CE:
                 per-step batch indices, losses, step-0 activations/logits/grads, params at
                 step 0 / 1 / 100, packed dropout masks.
  traj_ws2.npz   same loop on 2 gloo ranks (20 steps): per-rank indices and losses, step-0
                 averaged grads, final params.
  traj_ws4.npz, traj_ws8.npz   the same on 4 and 8 gloo ranks (1 thread each).
  main_e2e.npz   the reference's `main()` itself, unmodified, `--parallel --wireup_method gloo
                 --batch_size 64 --n_epochs 2` on a 4000/1000-sample dataset: the printed
                 `Epoch=…` lines and the final state_dict.
  kats.npz       sampler heads, decode table, dataset checksums.

Usage:  python oracle/gen_golden.py [--out tests/golden]
"""
from __future__ import annotations

import argparse
import contextlib
import hashlib
import io
import os
import re
import sys
import tempfile

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = "/root/reference"


def _setup_path():
    for p in (os.path.join(REPO, "pytorch_ddp_mnist_b200", "compat"), REPO, REF_DIR):
        if p not in sys.path:
            sys.path.insert(0, p)


_setup_path()
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch import nn  # noqa: E402

from pytorch_ddp_mnist_b200 import synth  # noqa: E402


from oracle.determinism import MASK_SEED0, MODEL_SEED, MaskFeed  # noqa: E402,F401


def _import_reference():
    import importlib
    return importlib.import_module("mnist_pnetcdf_cpu_mp")


def _state(model):
    return {k: v.detach().clone().numpy() for k, v in model.state_dict().items()}


def _init_pg(rank, ws, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(ws))
    dist.init_process_group("gloo", init_method="env://", rank=rank, world_size=ws)


# ----------------------------------------------------------------------------- trajectory runs
def run_traj(rank, ws, port, data_dir, steps, batch, out_q=None):
    """The reference's loop body (`mp:453-466`) with the reference's own objects, instrumented."""
    torch.set_num_threads(4 if ws <= 2 else 1)
    torch.use_deterministic_algorithms(True)
    os.chdir(data_dir)
    ref = _import_reference()
    _init_pg(rank, ws, port)
    feed = MaskFeed()
    feed.install()
    config = {"trainer": {"batch_size": batch, "rank": rank, "world_size": ws}}
    with contextlib.redirect_stdout(io.StringIO()):
        train_loader, _ = ref.create_data_loaders(config, ref.MPI.COMM_WORLD)
    torch.manual_seed(MODEL_SEED + rank)            # ranks differ BEFORE the DDP wrap (mp:439)
    model = ref.create_model()
    model = ref.DistributedDataParallel(model)      # broadcasts rank-0 parameters
    optimizer = ref.optim.SGD(model.parameters(), lr=0.01)
    loss = ref.nn.CrossEntropyLoss()
    model.train()
    train_loader.sampler.set_epoch(0)
    out = {"params0": _state(model.module)}
    acts = {}
    hooks = []
    for i, m in enumerate(model.module):
        hooks.append(m.register_forward_hook(
            lambda mod, inp, o, i=i: acts.__setitem__(i, o.detach().clone().numpy())))
    shard = np.array(list(iter(train_loader.sampler)), dtype=np.int64)
    losses, idxs = [], []
    it = iter(train_loader)
    for t in range(steps):
        x, y = next(it)
        b = x.shape[0]
        idxs.append(shard[t * batch:t * batch + b])
        x = x.view(x.shape[0], -1)
        optimizer.zero_grad()
        y_hat = model(x)
        batch_loss = loss(y_hat, y)
        batch_loss.backward()
        if t == 0:
            out["x0"] = x.numpy().copy()
            out["y0"] = y.numpy().copy()
            out["y0_dtype"] = str(y.dtype)
            out["acts0"] = {k: v.copy() for k, v in acts.items()}
            out["grads0"] = {k: p.grad.detach().clone().numpy()
                             for k, p in model.module.named_parameters()}
            for h in hooks:
                h.remove()
        optimizer.step()
        losses.append(batch_loss.item())
        if t == 0:
            out["params1"] = _state(model.module)
    out["paramsN"] = _state(model.module)
    out["losses"] = np.array(losses, np.float32)
    out["idx"] = np.stack(idxs)
    out["shard_head"] = shard[:16]
    out["shard_len"] = len(shard)
    out["masks"] = np.packbits(np.stack(feed.masks).astype(bool), axis=-1)
    feed.remove()
    dist.destroy_process_group()
    if out_q is not None:
        out_q.put((rank, out))
    return out


def _flatten(prefix, d, dst):
    for k, v in d.items():
        if isinstance(v, dict):
            _flatten(f"{prefix}{k}.", v, dst)
        else:
            dst[f"{prefix}{k}"] = np.asarray(v)


def gen_traj_ws1(out_dir, data_dir):
    o = run_traj(0, 1, 29611, data_dir, steps=100, batch=64)
    flat = {}
    _flatten("", o, flat)
    np.savez_compressed(os.path.join(out_dir, "traj_ws1.npz"), **flat)
    print("traj_ws1: loss[0]=%.6f loss[99]=%.6f" % (o["losses"][0], o["losses"][-1]))


def gen_traj_ws(out_dir, data_dir, ws, port):
    """`ws` gloo ranks of the reference loop, 20 steps: per-rank indices, losses and masks; step-0 averaged
    gradients; initial and final parameters (asserted identical on every rank, as DDP guarantees)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=run_traj, args=(r, ws, port, data_dir, 20, 64, q)) for r in range(ws)]
    for p in ps:
        p.start()
    res = dict(q.get() for _ in ps)
    for p in ps:
        p.join()
    flat = {}
    for r in range(ws):
        o = res[r]
        keep = {"losses": o["losses"], "idx": o["idx"], "masks": o["masks"],
                "shard_head": o["shard_head"], "shard_len": o["shard_len"]}
        _flatten(f"rank{r}.", keep, flat)
    for r in range(1, ws):
        for k in res[0]["paramsN"]:                 # DDP keeps the replicas identical
            assert np.array_equal(res[0]["paramsN"][k], res[r]["paramsN"][k]), k
            assert np.array_equal(res[0]["params0"][k], res[r]["params0"][k]), k
            assert np.array_equal(res[0]["grads0"][k], res[r]["grads0"][k]), k
    _flatten("", {"grads0": res[0]["grads0"], "params0": res[0]["params0"],
                  "paramsN": res[0]["paramsN"]}, flat)
    np.savez_compressed(os.path.join(out_dir, f"traj_ws{ws}.npz"), **flat)
    print(f"traj_ws{ws}: " + " ".join("rank%d loss[0]=%.6f" % (r, res[r]["losses"][0]) for r in range(ws)))


def gen_traj_ws2(out_dir, data_dir):
    gen_traj_ws(out_dir, data_dir, 2, 29612)


# ----------------------------------------------------------------------------- unmodified main()
E2E_TRAIN_N, E2E_TEST_N, E2E_EPOCHS, E2E_BATCH = 4000, 1000, 2, 64


def gen_main_e2e(out_dir):
    """Run the reference `__main__` flow (`mp:489-511`) unmodified on a small seeded dataset."""
    torch.set_num_threads(4)
    torch.use_deterministic_algorithms(True)
    ref = _import_reference()
    with tempfile.TemporaryDirectory() as d:
        synth.write_default_files(d, E2E_TRAIN_N, E2E_TEST_N)
        cwd = os.getcwd()
        os.chdir(d)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29613", RANK="0", WORLD_SIZE="1")
        argv = sys.argv
        sys.argv = ["mnist_pnetcdf_cpu_mp.py", "--parallel", "--wireup_method", "gloo",
                    "--batch_size", str(E2E_BATCH), "--n_epochs", str(E2E_EPOCHS)]
        feed = MaskFeed()
        feed.install()
        buf = io.StringIO()
        try:
            with contextlib.redirect_stdout(buf), contextlib.redirect_stderr(io.StringIO()):
                config = ref.configure()
                comm = ref.init_parallel(config)          # noqa: F841
                train_loader, test_loader = ref.create_data_loaders(config, ref.MPI.COMM_WORLD)
                torch.manual_seed(MODEL_SEED)
                model = ref.main(model=ref.create_model(), train_loader=train_loader,
                                 test_loader=test_loader, config=config)
        finally:
            sys.argv = argv
            feed.remove()
            os.chdir(cwd)
            if dist.is_initialized():
                dist.destroy_process_group()
    lines = re.findall(r"Epoch=(\d+), train_loss=([0-9.]+), val_loss=([0-9.]+)", buf.getvalue())
    assert len(lines) == E2E_EPOCHS, buf.getvalue()
    flat = {"epoch_lines": np.array([[float(a), float(b), float(c)] for a, b, c in lines]),
            "train_n": E2E_TRAIN_N, "test_n": E2E_TEST_N, "epochs": E2E_EPOCHS, "batch": E2E_BATCH,
            "n_train_forwards": feed.step}
    _flatten("state.", _state(model), flat)
    np.savez_compressed(os.path.join(out_dir, "main_e2e.npz"), **flat)
    print("main_e2e:", lines)


# ----------------------------------------------------------------------------- known-answer tests
def gen_kats(out_dir):
    from torch.utils.data import DistributedSampler
    from torchvision import transforms
    flat = {}

    class _Len:
        def __init__(self, n):
            self.n = n

        def __len__(self):
            return self.n
    for n in (60000, 4000, 10):
        for ws in (1, 2, 4, 8):
            for epoch in (0, 1):
                for rank in sorted({0, ws - 1}):
                    s = DistributedSampler(_Len(n), num_replicas=ws, rank=rank, shuffle=True, seed=42)
                    s.set_epoch(epoch)
                    sh = np.array(list(iter(s)), np.int64)
                    key = f"shard.n{n}.ws{ws}.e{epoch}.r{rank}"
                    flat[key + ".head"] = sh[:8]
                    flat[key + ".tail"] = sh[-8:]
                    flat[key + ".len"] = len(sh)
                    flat[key + ".sum"] = sh.sum()
                    flat[key + ".sha"] = np.frombuffer(hashlib.sha256(sh.tobytes()).digest(), np.uint8)
    # decode table through the reference's transform stack (mp:374-377)
    tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
    ramp = np.arange(256, dtype=np.uint8).reshape(16, 16)
    flat["lut"] = tf(ramp).reshape(256).numpy()
    for name, n, seed in (("train", synth.TRAIN_N, synth.TRAIN_SEED), ("test", synth.TEST_N, synth.TEST_SEED)):
        im, lb = synth.synth(n, seed)
        flat[f"synth.{name}.images_sha"] = np.frombuffer(hashlib.sha256(im.tobytes()).digest(), np.uint8)
        flat[f"synth.{name}.labels_sha"] = np.frombuffer(hashlib.sha256(lb.tobytes()).digest(), np.uint8)
        flat[f"synth.{name}.head"] = im[0, 0, :8]
    # CrossEntropyLoss with uint8 targets equals int64 targets (mp:461 passes u8)
    g = torch.Generator().manual_seed(7)
    z = torch.randn(64, 10, generator=g)
    y = torch.randint(0, 10, (64,), generator=g)
    flat["ce.z"] = z.numpy()
    flat["ce.y"] = y.numpy().astype(np.uint8)
    flat["ce.loss_u8"] = nn.CrossEntropyLoss()(z, y.to(torch.uint8)).numpy()
    flat["ce.loss_i64"] = nn.CrossEntropyLoss()(z, y).numpy()
    np.savez_compressed(os.path.join(out_dir, "kats.npz"), **flat)
    print("kats: lut[0]=%.16f lut[255]=%.16f" % (flat["lut"][0], flat["lut"][255]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(REPO, "tests", "golden"))
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    if not os.path.isdir(REF_DIR):
        raise SystemExit("gen_golden.py needs /root/reference (build container only)")
    todo = a.only.split(",") if a.only else ["kats", "traj_ws1", "traj_ws2", "traj_ws4", "traj_ws8", "main_e2e"]
    with tempfile.TemporaryDirectory() as data_dir:
        if {"traj_ws1", "traj_ws2", "traj_ws4", "traj_ws8"} & set(todo):
            synth.write_default_files(data_dir)
        if "kats" in todo:
            gen_kats(a.out)
        if "traj_ws1" in todo:
            gen_traj_ws1(a.out, data_dir)
        if "traj_ws2" in todo:
            gen_traj_ws2(a.out, data_dir)
        if "traj_ws4" in todo:
            gen_traj_ws(a.out, data_dir, 4, 29614)
        if "traj_ws8" in todo:
            gen_traj_ws(a.out, data_dir, 8, 29618)
        if "main_e2e" in todo:
            gen_main_e2e(a.out)


if __name__ == "__main__":
    main()
