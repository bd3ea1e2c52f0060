"""Run under torchrun with 2, 4 or 8 ranks (tests/test_gpu_train.py::test_multi_gpu_ddp_matches_reference and
`bench.py`'s parity gate use the same code): every data-parallel step path against the reference's own gloo
run at the same world size (tests/golden/traj_ws{2,4,8}.npz), plus what the gate does not cover — a second fused
launch on the same communicator, the drop-in `create_model()` under `DistributedDataParallel`, the slab +
all-gather dataset loader, and the BASELINE.json ConvNet's bucket all-reduce."""
import os
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

from oracle import mnist_oracle as O                      # noqa: E402
from pytorch_ddp_mnist_b200 import engine as E, synth     # noqa: E402
from pytorch_ddp_mnist_b200.dataset import ResidentDataset  # noqa: E402
import parity_gate                                        # noqa: E402


def check_ddp_wrapped_model(rank, world, local, dev, g):
    """tg:72 / mp:439: `DistributedDataParallel(create_model().to(dev), device_ids=[rank])` over the drop-in
    model, driven by the reference's loop body (mp:455-465) textually, against the reference's gloo run."""
    from torch.nn.parallel import DistributedDataParallel
    from pytorch_ddp_mnist_b200 import model as M, ops
    im, lb = synth.synth(synth.TRAIN_N, synth.TRAIN_SEED)
    images = torch.from_numpy(im.reshape(-1, 784)).to(dev)
    labels = torch.from_numpy(lb).to(dev)
    torch.manual_seed(100 + rank)                          # ranks start different, the wrap broadcasts rank 0's
    model = M.create_model()
    if rank == 0:
        model.load_state_dict({k: torch.from_numpy(g[f"params0.{k}"]) for k in O.PARAM_KEYS})
    model = model.to(dev)
    masks = torch.from_numpy(O.unpack_masks(g[f"rank{rank}.masks"])).to(dev)
    model.set_mask_provider(lambda t, b: masks[t])
    model = DistributedDataParallel(model, device_ids=[local], output_device=local)
    for k, v in model.module.state_dict().items():
        assert np.array_equal(v.cpu().numpy(), g[f"params0.{k}"]), ("DDP wrap did not adopt rank 0's parameters", k)
    optimizer = M.SGD(model.parameters(), lr=0.01)
    loss = M.CrossEntropyLoss()
    model.train()
    idx = torch.from_numpy(g[f"rank{rank}.idx"]).to(dev)
    losses = []
    for t in range(idx.shape[0]):
        x, y = ops.gather_decode(images, labels, idx[t])
        x = x.view(x.shape[0], -1)
        optimizer.zero_grad()
        y_hat = model(x)
        batch_loss = loss(y_hat, y)
        batch_loss.backward()
        optimizer.step()
        losses.append(batch_loss.item())
    ref = g[f"rank{rank}.losses"]
    err = np.abs(np.array(losses) - ref).max()
    assert err < 5e-3 * np.abs(ref).max(), ("DDP-wrapped model: loss trajectory", rank, err)
    assert np.abs(np.array(losses[:3]) - ref[:3]).max() < 1e-5
    sd = model.module.state_dict()
    for k in O.PARAM_KEYS:
        r = g[f"paramsN.{k}"].astype(np.float64)
        fro = np.linalg.norm(sd[k].cpu().numpy() - r) / np.linalg.norm(r)
        assert fro < 5e-3, ("DDP-wrapped model: final parameters", k, fro)
    mine = torch.cat([sd[k].reshape(-1) for k in O.PARAM_KEYS])
    other = mine.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(mine, other), "DDP-wrapped model: replicas diverged"
    return float(err)


def check_exchange_only(rank, world, dev):
    """The fused step's gradient exchange alone, both algorithms (one-shot: everyone pushes to everyone; two-stage:
    thread slices have a home rank that sums and pushes back), over three launches on one communicator: every
    element of the bucket must come out as the exact sum of the ranks' contributions, identically on all ranks."""
    from pytorch_ddp_mnist_b200 import ops
    e = np.arange(O.N_PARAMS)
    unit = (1 + (e & 1023)).astype(np.float64) / 1024.0
    for algo in ("one_shot", "two_stage"):
        os.environ["DDPM_EXCHANGE"] = algo
        try:
            comm = ops.FusedComm(rank, world)
        finally:
            os.environ.pop("DDPM_EXCHANGE", None)
        out = torch.zeros(O.N_PARAMS, device=dev)
        for nsteps in (1, 6, 3):
            out.fill_(-1.0)
            ops.exchange_only(comm, nsteps, out)
            torch.cuda.synchronize()
            assert comm.error() == 0, (algo, "exchange timed out")
            want = unit * (world * (world + 1) / 2) + world * ((nsteps - 1) & 255)
            got = out.cpu().numpy().astype(np.float64)
            assert np.array_equal(got, want), (algo, nsteps, int((got != want).sum()), got[:4], want[:4])
        dist.barrier()
        comm.close()


def check_slab_loader(rank, world, dev):
    """`ResidentDataset.from_file` at ws > 1: every rank reads one contiguous slab of the CDF-5 file and an
    all-gather replicates the u8 dataset - the result must be the file's bytes on every rank."""
    n = 10_007                                              # not divisible by the world size: ragged last slab
    im, lb = synth.synth(n, 77)
    d = [None]
    if rank == 0:
        d[0] = tempfile.mkdtemp(prefix="slab_")
        from pytorch_ddp_mnist_b200 import cdf5
        cdf5.write_mnist_cdf5(os.path.join(d[0], "mnist_train_images.nc"), im, lb)
    dist.broadcast_object_list(d, src=0)
    res = ResidentDataset.from_file(os.path.join(d[0], "mnist_train_images.nc"), dev, rank=rank, world_size=world)
    assert res.images.shape[0] == n and res.labels.shape[0] == n
    assert torch.equal(res.images.cpu(), torch.from_numpy(im.reshape(-1, 784))), "slab loader: pixels differ"
    assert torch.equal(res.labels.cpu(), torch.from_numpy(lb)), "slab loader: labels differ"
    dist.barrier()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", init_method="env://", device_id=dev)

    # ---- every step path vs the reference's gloo run at this world size
    rec = parity_gate.run_gate(dev, rank, world, dist)
    assert rec["ok"], rec
    if world in parity_gate.GOLDEN_WS:
        assert rec["golden"] and rec["max_loss_err"] is not None

    if world == 2:
        # two ranks default to the one-shot exchange: run the gate once more through the two-stage algorithm (the
        # default beyond 2 ranks), so that a 2-GPU box checks it against the reference as well
        os.environ["DDPM_EXCHANGE"] = "two_stage"
        try:
            rec2 = parity_gate.run_gate(dev, rank, world, dist, include_host_fed=False)
        finally:
            os.environ.pop("DDPM_EXCHANGE", None)
        assert rec2["ok"], rec2
    check_exchange_only(rank, world, dev)

    im, lb = synth.synth(synth.TRAIN_N, synth.TRAIN_SEED)
    res = ResidentDataset.from_arrays(im, lb, dev)
    ddp_err = None
    if world in parity_gate.GOLDEN_WS:
        g = np.load(os.path.join(REPO, "tests", "golden", f"traj_ws{world}.npz"))
        assert np.array_equal(g[f"rank{rank}.idx"][0], O.shard_indices(60000, world, rank)[:64])
        # ---- a second launch continues the flag sequence (communicator step counter) and stays in lock-step
        eng2 = E.TrainEngine(dev, 64, world_size=world, rank=rank)
        eng2.load_state_dict({k: torch.from_numpy(g[f"params0.{k}"]) for k in O.PARAM_KEYS})
        eng2.enable_fused_ddp()
        idx = torch.from_numpy(g[f"rank{rank}.idx"]).to(dev)
        idx_all = torch.from_numpy(np.stack([g[f"rank{r}.idx"].reshape(-1) for r in range(world)])).to(dev)
        for n0, n1 in ((0, 7), (7, 20), (0, 4)):
            eng2.train_steps(res, idx[n0:n1].reshape(-1).contiguous(), 64, n1 - n0,
                             idx_all=idx_all[:, n0 * 64:n1 * 64].contiguous())
            torch.cuda.synchronize()
            eng2.check_comm()
            mine = eng2.params.clone()
            other = mine.clone()
            dist.broadcast(other, src=0)
            assert torch.equal(mine, other), "fused replicas diverged after a follow-up launch"
        dist.barrier()
        ddp_err = check_ddp_wrapped_model(rank, world, local, dev, g)
    check_slab_loader(rank, world, dev)

    # ---- large batch (BASELINE.json configs[3]): the bucket leaves in two pieces, the small one all-reduced while fc1's
    #      weight gradient is still being computed; the result must equal the single all-reduce after backward
    big_b = 1024
    brng = np.random.default_rng(99)
    bidx = torch.from_numpy(brng.integers(0, 60000, (world, 3, big_b)).astype(np.int64)[rank]).to(dev)
    finals = []
    for overlap in (True, False):
        e = E.TrainEngine(dev, big_b, world_size=world, rank=rank)
        e.overlap_allreduce = overlap
        torch.manual_seed(3)
        e.load_state_dict(E.init_params_like_reference())
        e.broadcast_params(0)
        for t in range(3):
            e.train_step(res, bidx[t], philox_step=t)
        torch.cuda.synchronize()
        finals.append(e.params.clone())
        other = e.params.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(e.params, other), "large-batch replicas diverged"
    # (NCCL may pick another algorithm, hence another summation order, for the smaller messages: tolerance, not bits)
    rel = float((finals[0] - finals[1]).abs().max() / finals[1].abs().max())
    assert rel < 1e-5, ("overlapped two-piece all-reduce differs from the single all-reduce", rel)
    dist.barrier()

    # ---- BASELINE.json's ConvNet (not in the reference; oracle = restatement): bucket all-reduce over NCCL.
    # Each rank's gradient of its own 16-sample batch, pre-divided by ws and summed, must equal the gradient of the
    # mean loss over the global batch (the restatement evaluated on all ws x 16 samples), and replicas must agree.
    from oracle import convnet_oracle as CO
    from pytorch_ddp_mnist_b200 import convnet as cn
    cb = 16
    crng = np.random.default_rng(4242)
    cidx = crng.integers(0, 60000, (world, cb)).astype(np.int64)
    cm1 = (crng.random((world, cb, 9216)) < 0.75).astype(np.uint8)
    cm2 = (crng.random((world, cb, 128)) < 0.5).astype(np.uint8)
    cp = CO.init_params(5)
    ceng = cn.ConvNetEngine(dev, cb, world_size=world, rank=rank)
    ceng.load_state_dict(cp)
    ceng.train_step(res, torch.from_numpy(cidx[rank]).to(dev), torch.from_numpy(cm1[rank]).to(dev),
                    torch.from_numpy(cm2[rank]).to(dev))
    torch.cuda.synchronize()
    xs, ys = O.gather_decode(im, lb, cidx.reshape(-1))
    _, _, cg, _ = CO.forward_backward(cp, xs.reshape(world * cb, 784), ys, cm1.reshape(world * cb, 9216),
                                      cm2.reshape(world * cb, 128))
    gref = CO.flatten_params(cg).astype(np.float64)
    cerr = float(np.linalg.norm(ceng.grads.cpu().numpy() - gref) / np.linalg.norm(gref))
    assert cerr < 1e-4, ("convnet all-reduced gradient vs global-batch restatement", cerr)
    mine = ceng.params.clone()
    other = mine.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(mine, other), "convnet replicas diverged"
    dist.barrier()
    # ---- a peer that never shows up: the waiting kernel gives up after ~3 s, every later wait within microseconds
    #      (the error word is sticky), and the host sees the error - no rank holds its GPU for minutes
    import time
    lone = E.TrainEngine(dev, 64, world_size=world, rank=rank)
    torch.manual_seed(5)
    lone.load_state_dict(E.init_params_like_reference())
    lone.enable_fused_ddp()
    dist.barrier()
    if rank == 0:
        lidx = torch.from_numpy(O.shard_indices(60000, world, 0)[:40 * 64]).to(dev)
        t0 = time.perf_counter()
        lone.train_steps(res, lidx, 64, 40)                   # 40 steps, nobody else joins
        torch.cuda.synchronize()
        waited = time.perf_counter() - t0
        assert lone.comm.error() != 0, "the exchange time-out did not raise the communicator's error word"
        assert waited < 12.0, f"a missing peer held the kernel for {waited:.1f} s (the bound is ~3 s, sticky afterwards)"
        try:
            lone.check_comm()
            raise AssertionError("check_comm() did not raise")
        except RuntimeError:
            pass
    dist.barrier()
    del lone

    if rank == 0:
        import json
        print("PARITY_GATE " + json.dumps(rec))
        print("MULTIGPU_OK world %d max loss err vs reference %s ddp-wrapped model err %s" %
              (world, rec["max_loss_err"], ddp_err))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
