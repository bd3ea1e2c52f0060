"""world_size-2 gloo tests on CPU for the host-side N>1 logic: env wire-up, sampler sharding, and
the flat-bucket protocol (pre-divide by ws, allreduce SUM, identical SGD on every rank) with the
oracle doing the arithmetic — compared with the reference's own 2-rank gloo run."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mnist_oracle as O

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    try:
        for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "LOCAL_RANK"):
            os.environ.pop(k, None)
        # launcher-style discovery: PMI_* variables only, the gloo branch fills the rest (mp:213-254)
        os.environ.update(PMI_RANK=str(rank), PMI_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        from pytorch_ddp_mnist_b200 import synth, wireup
        comm = wireup.distributed("gloo")
        assert (comm.get_rank(), comm.get_size()) == (rank, world)
        assert os.environ["RANK"] == str(rank) and os.environ["WORLD_SIZE"] == str(world)
        mx = comm.reduceMAX(np.array([float(rank), 5.0 - rank]), 0)
        if rank == 0:
            assert mx.tolist() == [1.0, 5.0]
        comm.barrier()
        g = np.load(os.path.join(REPO, "tests", "golden", "traj_ws2.npz"))
        im, lb = synth.synth(synth.TRAIN_N, synth.TRAIN_SEED)
        # DDP wrap: rank 0's parameters win
        flat = torch.from_numpy(O.flatten_params({k: g[f"params0.{k}"] for k in O.PARAM_KEYS})) if rank == 0 \
            else torch.randn(O.N_PARAMS)
        dist.broadcast(flat, src=0)
        p = O.unflatten_params(flat.numpy())
        shard = O.shard_indices(60000, world, rank)
        masks = O.unpack_masks(g[f"rank{rank}.masks"])
        losses = []
        for t in range(8):
            idx = shard[t * 64:(t + 1) * 64]
            assert np.array_equal(idx, g[f"rank{rank}.idx"][t])
            x, y = O.gather_decode(im, lb, idx)
            a = O.forward(p, x, masks[t])
            loss, dz3 = O.cross_entropy(a["z3"], y)
            gr = O.backward(p, x, a, masks[t], dz3)
            bucket = torch.from_numpy(O.flatten_params(gr) * np.float32(1.0 / world))   # pre-divided
            dist.all_reduce(bucket, op=dist.ReduceOp.SUM)                               # ONE collective per step
            p = O.sgd_step(p, O.unflatten_params(bucket.numpy()))
            losses.append(float(loss))
            if t == 0:
                gu = O.unflatten_params(bucket.numpy())
                for k in O.PARAM_KEYS:
                    ref = g[f"grads0.{k}"]
                    assert np.abs(gu[k] - ref).max() <= 2e-5 * np.abs(ref).max(), k
        assert np.abs(np.array(losses) - g[f"rank{rank}.losses"][:8]).max() < 1e-4
        # replicas identical after 8 steps
        mine = torch.from_numpy(O.flatten_params(p))
        other = mine.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(mine, other)
        comm.finalize()
        q.put((rank, "ok"))
    except Exception as e:                                   # surface the failure in the parent
        import traceback
        q.put((rank, traceback.format_exc()))


def test_gloo_ws2_flat_bucket_protocol():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, 29741, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = dict(q.get(timeout=300) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    assert res == {0: "ok", 1: "ok"}, res


def test_discover_env_precedence(monkeypatch):
    from pytorch_ddp_mnist_b200 import wireup
    for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "PMI_RANK", "PMI_SIZE",
              "OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "PMIX_SERVER_URI2", "SLURM_NTASKS",
              "SLURM_PROCID", "SLURM_LAUNCH_NODE_IPADDR", "SLURM_JOB_NUM_NODES", "SLURM_NTASKS_PER_NODE"):
        monkeypatch.delenv(k, raising=False)
    wireup.discover_env("gloo")                              # nothing set: singleton on localhost:29500
    assert (os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"], os.environ["WORLD_SIZE"], os.environ["RANK"]) == \
        ("localhost", "29500", "1", "0")
    for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("PMIX_SERVER_URI2", "pmix-server.123;tcp4://10.1.2.3:4567")
    monkeypatch.setenv("OMPI_COMM_WORLD_SIZE", "4")
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "3")
    wireup.discover_env("gloo")
    assert (os.environ["MASTER_ADDR"], os.environ["WORLD_SIZE"], os.environ["RANK"]) == ("10.1.2.3", "4", "3")
    # nccl-slurm: the reference multiplies two strings (mp:118); here it is an integer product
    for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        monkeypatch.delenv(k, raising=False)
    with pytest.raises(Exception, match="SLURM_LAUNCH_NODE_IPADDR"):
        wireup.discover_env("nccl-slurm")
    monkeypatch.setenv("SLURM_LAUNCH_NODE_IPADDR", "10.0.0.9")
    monkeypatch.setenv("SLURM_JOB_NUM_NODES", "2")
    monkeypatch.setenv("SLURM_NTASKS_PER_NODE", "8")
    monkeypatch.setenv("SLURM_PROCID", "5")
    wireup.discover_env("nccl-slurm")
    assert (os.environ["MASTER_ADDR"], os.environ["WORLD_SIZE"], os.environ["RANK"]) == ("10.0.0.9", "16", "5")
    with pytest.raises(NotImplementedError):
        wireup.discover_env("carrier-pigeon")


def test_configure_flags_match_reference():
    from pytorch_ddp_mnist_b200 import train
    c = train.configure([])
    assert c["trainer"]["batch_size"] == 128 and c["trainer"]["n_epochs"] == 1            # mp:294-297
    assert c["trainer"]["wireup_method"] == "nccl-slurm" and c["trainer"]["parallel"] is False
    c = train.configure(["--wireup_method", "gloo", "--batch_size", "64", "--n_epochs", "3", "--parallel",
                         "--data_limit", "10", "--num_workers", "2", "--hdf5", "--local-rank", "1"])
    assert (c["trainer"]["batch_size"], c["trainer"]["n_epochs"], c["trainer"]["parallel"]) == (64, 3, True)
    assert c["data"]["limit"] == 10 and c["data"]["hdf5"] is True and c["trainer"]["num_workers"] == 2
    train.configure(["--local_rank", "2"])                                               # both spellings (tg:123)
