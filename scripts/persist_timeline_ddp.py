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
    def seg(cta, a_, b_):
        return int((d[cta, 2:8, b_] - d[cta, 2:8, a_]).mean())
    print(f"world {world}: cycles per segment, averaged over steps 2..7 of a 20-step launch (rank 0)")
    for cta in (0, 63, 64, 111, 112, 146):
        tot = int((d[cta, 3:8, 0] - d[cta, 2:7, 0]).mean())
        if cta < B:
            print(f"cta {cta:3d}: A={seg(cta,0,1):6d} wait_partials={seg(cta,1,2):6d} B={seg(cta,2,3):6d} "
                  f"wait_rows={seg(cta,3,4):6d} C={seg(cta,4,5):6d}  step={tot}")
        else:
            print(f"cta {cta:3d}: A={seg(cta,0,1):6d} (idle)={seg(cta,1,3):6d} wait_rows={seg(cta,3,4):6d} "
                  f"C={seg(cta,4,5):6d}  step={tot}")
    for cta in (0, 100):
        x = d[cta, 2:8]
        if (x[:, 12] > 0).all():
            print(f"cta {cta} phase C: gradient tile ready after {int((x[:,11]-x[:,4]).mean())}, "
                  f"exchange (push + wait + sum) {int((x[:,12]-x[:,11]).mean())}, update {int((x[:,5]-x[:,12]).mean())}")
dist.barrier()
os._exit(0)
