"""Per-kernel and whole-step timing of the ConvNet path (CUDA events)."""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
from pytorch_ddp_mnist_b200 import convnet as cn, synth
from pytorch_ddp_mnist_b200.dataset import ResidentDataset
from oracle import convnet_oracle as CO
dev = torch.device('cuda:0')
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
TF32 = len(sys.argv) > 2 and sys.argv[2] == 'tf32'
im, lb = synth.synth(60000, 1234)
res = ResidentDataset.from_arrays(im, lb, dev)
eng = cn.ConvNetEngine(dev, B, tf32=TF32)
eng.load_state_dict(CO.init_params(0))
idx = [torch.randint(0, 60000, (B,), device=dev) for _ in range(50)]
def timed(fn, n=50):
    for i in range(5): fn(idx[i])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(idx[i % 50])
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
fwd = lambda ix: cn.step_forward(eng.params, eng.ws, images=res.images, labels=res.labels, idx=ix, philox=(1, 0), prep_backward=True, loss=eng.loss, tf32=TF32)
bwd = lambda ix: cn.step_backward(eng.params, eng.ws, eng.grads, images=res.images, idx=ix, tf32=TF32, forward_was_tf32=TF32)
opt = lambda ix: cn.adadelta_step(eng.params, eng.grads, eng.square_avg, eng.acc_delta)
step = lambda ix: eng.train_step(res, ix)
t = {k: timed(f) for k, f in (("forward", fwd), ("backward", bwd), ("adadelta", opt), ("step", step))}
print(f"B={B} tf32={TF32}: " + ", ".join(f"{k} {v:.1f} us" for k, v in t.items()) + f" -> {B / t['step'] * 1e6:.3e} img/s, loss {eng.last_batch_loss():.4f}")
