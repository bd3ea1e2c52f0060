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
# stamps: 0 step start | 1 phase A done (partials stored) | 2 first row's partials arrived (row CTAs) | 3 phase B done
#         4 first operands of phase C arrived | 5 phase C done.  Non-row CTAs have no stamp 2 (3 follows 1 directly).
print("cycles per segment, averaged over steps 2..7 (clock64, ~1.9 GHz)")
def seg(cta, a, b):
    return int((d[cta, 2:8, b] - d[cta, 2:8, a]).mean())
for cta in (0, 1, 63, 64, 100, 111, 112, 130, 143, 144, 145, 146, 147):
    tot = int((d[cta, 3:8, 0] - d[cta, 2:7, 0]).mean())
    row = cta < B
    if row:
        print(f"cta {cta:3d}: A={seg(cta,0,1):6d} wait_partials={seg(cta,1,2):6d} B={seg(cta,2,3):6d} "
              f"wait_rows={seg(cta,3,4):6d} C={seg(cta,4,5):6d}  step={tot}")
    else:
        print(f"cta {cta:3d}: A={seg(cta,0,1):6d} {'':21s} (idle)={seg(cta,1,3):6d} "
              f"wait_rows={seg(cta,3,4):6d} C={seg(cta,4,5):6d}  step={tot}")
for cta in (0, 63):
    x = d[cta, 2:8]
    print(f"cta {cta} phase B: z1+h1d={int((x[:,7]-x[:,6]).mean())} (incl. z2) logits+softmax+dz2={int((x[:,8]-x[:,7]).mean())} dz1={int((x[:,3]-x[:,8]).mean())}")
for cta in (0, 1, 63):
    x = d[cta, 3:8]
    f = lambda k: int((x[:, k] - x[:, 0]).mean())
    print(f"cta {cta} (cycles into the step): A done {f(1)}, W2/W3 loads issued {f(13)}, past the ready check {f(14)}, "
          f"partials polled {f(15)}, b1 polled {f(12)}, first barrier of phase B {f(2)}")
for cta in (0, 100):
    x = d[cta, 2:8]
    print(f"cta {cta} phase C: loop={int((x[:,10]-x[:,9]).mean())} update={int((x[:,5]-x[:,10]).mean())}")
