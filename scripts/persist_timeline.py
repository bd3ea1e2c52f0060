"""Profiling aid: per-phase clock64 timeline of the persistent kernel (first 8 steps of a launch)."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pytorch_ddp_mnist_b200 import ops, synth, engine as E, _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
NROWS = int(sys.argv[2]) if len(sys.argv) > 2 else 240000
dev = torch.device("cuda:0")
im, lb = synth.synth(240000, 1234)
images = torch.from_numpy(im.reshape(-1, 784)).to(dev); labels = torch.from_numpy(lb).to(dev)
torch.manual_seed(0)
eng = E.TrainEngine(dev, B)
eng.load_state_dict(E.init_params_like_reference())
idx = (torch.randperm(240000, device=dev)[:40 * B] % NROWS).to(torch.int64)
G = 148
dbg = torch.zeros(G * 8 * 16, dtype=torch.int64, device=dev)
ops.train_steps_persistent(eng.params, eng.ws, B, 20, images=images, labels=labels, idx=idx[:20 * B], philox=(1, 0), lr=0.01)
torch.cuda.synchronize()
_lib.lib.ddpm_debug_set_persist_timing.argtypes = [C.c_void_p]
_lib.lib.ddpm_debug_set_persist_timing(dbg.data_ptr())
ops.train_steps_persistent(eng.params, eng.ws, B, 20, images=images, labels=labels, idx=idx[20 * B:], philox=(1, 20), lr=0.01)
torch.cuda.synchronize()
_lib.lib.ddpm_debug_set_persist_timing(None)
d = dbg.cpu().numpy().reshape(G, 8, 16)
names = ["phaseA", "wait1", "phaseB", "wait2", "phaseC"]
print("cycles per segment, averaged over steps 2..7 (clock64, ~1.9 GHz)")
for cta in (0, 1, 63, 64, 100, 111, 112, 130, 143, 144, 145, 146, 147):
    seg = (d[cta, 2:8, 1:6] - d[cta, 2:8, 0:5]).mean(axis=0)
    tot = (d[cta, 3:8, 0] - d[cta, 2:7, 0]).mean()
    print(f"cta {cta:3d}: " + " ".join(f"{n}={int(v):6d}" for n, v in zip(names, seg)) + f"  step={int(tot)}")
seg_all = (d[:, 2:8, 1:6] - d[:, 2:8, 0:5]).mean(axis=1)
print("max over CTAs:", dict(zip(names, seg_all.max(axis=0).astype(int))))
print("mean over CTAs:", dict(zip(names, seg_all.mean(axis=0).astype(int))))

# finer stamps: phase B of a row CTA (2 -> 6 staged+partials, 6 -> 7 h1d+z2, 7 -> 8 logits+softmax, 8 -> 3 dz2+dz1)
for cta in (0, 63):
    x = d[cta, 2:8]
    print(f"cta {cta} phase B: partials+wait={int((x[:,6]-x[:,2]).mean())} z2={int((x[:,7]-x[:,6]).mean())} "
          f"logits+softmax={int((x[:,8]-x[:,7]).mean())} dz2+dz1={int((x[:,3]-x[:,8]).mean())}")
for cta in (0, 100):
    x = d[cta, 2:8]
    print(f"cta {cta} phase C: load+sync={int((x[:,9]-x[:,4]).mean())} loop={int((x[:,10]-x[:,9]).mean())} update={int((x[:,5]-x[:,10]).mean())}")
