"""Workload for the ncu captures of train_persist_kernel: after two warm-up launches, one launch each of 1, 20 and 100
steps at batch 64 on fresh rows of a 240,000-sample resident dataset (the traffic per step is the slope between them)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_ddp_mnist_b200 import ops, synth, engine as E

B = 64
dev = torch.device("cuda:0")
im, lb = synth.synth(240000, 1234)
images = torch.from_numpy(im.reshape(-1, 784)).to(dev); labels = torch.from_numpy(lb).to(dev)
torch.manual_seed(0)
eng = E.TrainEngine(dev, B)
eng.load_state_dict(E.init_params_like_reference())
idx = torch.randperm(240000, device=dev)[:200 * B].to(torch.int64)
pos = 0
for n in (20, 20, 1, 20, 100):
    ops.train_steps_persistent(eng.params, eng.ws, B, n, images=images, labels=labels, idx=idx[pos * B:(pos + n) * B],
                               philox=(1, pos), lr=0.01, loss=eng.loss)
    pos += n
    torch.cuda.synchronize()
assert eng.ws.error() == 0
print("final loss", float(eng.loss[0]))
