"""Profiling aid: per-phase timeline of the fused data-parallel persistent kernel (run under torchrun)."""
import ctypes as C
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from pytorch_ddp_mnist_b200 import ops, synth, engine as E, _lib
from pytorch_ddp_mnist_b200.dataset import ResidentDataset

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", init_method="env://", device_id=dev)
B = 64
im, lb = synth.synth(60000, 1234)
res = ResidentDataset.from_arrays(im, lb, dev)
torch.manual_seed(0)
eng = E.TrainEngine(dev, B, world_size=world, rank=rank)
eng.load_state_dict(E.init_params_like_reference())
eng.enable_fused_ddp()
idx = torch.randperm(60000, device=dev)[:40 * B].to(torch.int64)
G = 148
dbg = torch.zeros(G * 8 * 16, dtype=torch.int64, device=dev)
ia = torch.stack([idx.roll(r * 7) for r in range(world)])  # every rank passes the same table; row `rank` is its own
idx_r = ia[rank].contiguous()
eng.train_steps(res, idx_r[:20 * B], B, 20, idx_all=ia[:, :20 * B].contiguous())
torch.cuda.synchronize(); dist.barrier()
_lib.lib.ddpm_debug_set_persist_timing.argtypes = [C.c_void_p]
_lib.lib.ddpm_debug_set_persist_timing(dbg.data_ptr())
eng.train_steps(res, idx_r[20 * B:], B, 20, idx_all=ia[:, 20 * B:].contiguous())
torch.cuda.synchronize()
_lib.lib.ddpm_debug_set_persist_timing(None)
eng.check_comm()
d = dbg.cpu().numpy().reshape(G, 8, 16)
if rank == 0:
    names = ["phaseA", "wait1", "phaseB", "wait2", "phaseC"]
    for cta in (0, 63, 64, 111, 112, 146):
        seg = (d[cta, 2:8, 1:6] - d[cta, 2:8, 0:5]).mean(axis=0)
        tot = (d[cta, 3:8, 0] - d[cta, 2:7, 0]).mean()
        print(f"cta {cta:3d}: " + " ".join(f"{n}={int(v):6d}" for n, v in zip(names, seg)) + f"  step={int(tot)}")
    for cta in (0, 100):
        x = d[cta, 2:8]
        print(f"cta {cta} phase B: partials+wait={int((x[:,6]-x[:,2]).mean())} z2={int((x[:,7]-x[:,6]).mean())} logits+softmax={int((x[:,8]-x[:,7]).mean())} dz2+dz1={int((x[:,3]-x[:,8]).mean())}")
dist.barrier()
os._exit(0)
