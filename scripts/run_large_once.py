"""Profiling aid: a few large-batch three-kernel steps (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_ddp_mnist_b200 import ops, synth, engine as E
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device("cuda:0")
im, lb = synth.synth(60000, 1234)
images = torch.from_numpy(im.reshape(-1, 784)).to(dev); labels = torch.from_numpy(lb).to(dev)
torch.manual_seed(0)
eng = E.TrainEngine(dev, B); eng.load_state_dict(E.init_params_like_reference())
idx = torch.randperm(60000, device=dev).to(torch.int64)
for i in range(4):
    ops.step_forward(eng.params, eng.ws, B, images=images, labels=labels, idx=idx[i * B:(i + 1) * B], philox=(1, i), prep_backward=True, loss=eng.loss, tf32=True)
    ops.step_backward(eng.params, eng.ws, B, images=images, idx=idx[i * B:(i + 1) * B], lr=0.01, tf32=True)
torch.cuda.synchronize()
print("loss", float(eng.loss[0]))
