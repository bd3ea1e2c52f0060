"""Profiling aid: phase stamps of the persistent kernel on ONE clock (%globaltimer, build with
DDPM_DEFS=-DDDPM_DBG_GLOBALTIMER into another .so and point DDPM_LIB at it): every CTA's events of one step in
nanoseconds relative to the earliest step start, so that producers and consumers can be lined up."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pytorch_ddp_mnist_b200 import ops, synth, engine as E, _lib

B = 64
dev = torch.device("cuda:0")
im, lb = synth.synth(240000, 1234)
images = torch.from_numpy(im.reshape(-1, 784)).to(dev); labels = torch.from_numpy(lb).to(dev)
torch.manual_seed(0)
eng = E.TrainEngine(dev, B)
eng.load_state_dict(E.init_params_like_reference())
idx = torch.randperm(240000, device=dev)[:40 * B].to(torch.int64)
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
names = {0: "start", 1: "A_done", 13: "param_loads", 14: "past_ready", 15: "partials", 12: "b1", 2: "B_first_bar", 7: "h2", 8: "dz2",
         3: "B_done", 4: "C_operands", 10: "C_loop_done", 5: "C_done"}
for step in (4, 5):
    t0 = d[:, step, 0].min()
    print(f"--- step {step}: ns after the earliest CTA's step start (globaltimer); spread of step starts: "
          f"{int(d[:, step, 0].max() - t0)} ns")
    for cta in (0, 1, 14, 63, 64, 100, 111, 112, 127, 143, 144, 146, 147):
        ev = [(int(d[cta, step, k] - t0), n) for k, n in names.items() if d[cta, step, k] > 0 and d[cta, step, k] >= d[cta, step, 0]]
        ev.sort()
        print(f"cta {cta:3d}: " + "  ".join(f"{n}={t}" for t, n in ev))
