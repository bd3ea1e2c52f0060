"""Timing aid: the three-kernel step at large batch (fp32 FFMA fc1 vs tcgen05 TF32 fc1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_ddp_mnist_b200 import ops, synth, engine as E
dev = torch.device("cuda:0")
im, lb = synth.synth(240000, 1234)
images = torch.from_numpy(im.reshape(-1, 784)).to(dev); labels = torch.from_numpy(lb).to(dev)
torch.manual_seed(0)
for B in (1024, 4096):
    eng = E.TrainEngine(dev, B); eng.load_state_dict(E.init_params_like_reference())
    idx = torch.randperm(240000, device=dev).to(torch.int64)
    for tf32 in (False, True):
        def fwd(i): ops.step_forward(eng.params, eng.ws, B, images=images, labels=labels, idx=idx[(i % 50) * B:(i % 50 + 1) * B], philox=(1, i), prep_backward=True, loss=eng.loss, tf32=tf32)
        def bwd(i): ops.step_backward(eng.params, eng.ws, B, images=images, idx=idx[(i % 50) * B:(i % 50 + 1) * B], lr=0.01, tf32=tf32)
        for i in range(5): fwd(i); bwd(i)
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        for i in range(100): fwd(i)
        e[1].record()
        for i in range(100): fwd(i); bwd(i)
        e[2].record(); torch.cuda.synchronize()
        tf = e[0].elapsed_time(e[1]) / 100; ts = e[1].elapsed_time(e[2]) / 100
        print(f"B={B} tf32={tf32}: forward {tf*1000:.1f} us, fwd+bwd step {ts*1000:.1f} us -> {B/ts*1000:.3e} img/s, loss {float(eng.loss[0]):.4f}")
