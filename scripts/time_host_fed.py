"""Where does the host-fed loop spend its time?  Host enqueue time of the call vs device time per step."""
import sys, time
import numpy as np, torch
sys.path.insert(0, '.')
from pytorch_ddp_mnist_b200 import ops, synth, engine as E
dev = torch.device('cuda:0')
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
rng = np.random.default_rng(0)
hx = torch.from_numpy(rng.integers(0, 256, (n, B, 784), dtype=np.uint8)).pin_memory()
hy = torch.from_numpy(rng.integers(0, 10, (n, B), dtype=np.uint8)).pin_memory()
hl = torch.zeros(n).pin_memory()
for slots in (64, 256, 1024):
    eng = E.TrainEngine(dev, B); eng.load_state_dict(E.init_params_like_reference())
    eng.train_steps_host(hx[:20], hy[:20], B, 20, host_loss_steps=hl[:20], slots=slots)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    eng.train_steps_host(hx, hy, B, n, host_loss_steps=hl, slots=slots)
    t1 = time.perf_counter(); e1.record()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    eng.check_feed()
    print(f"B={B} slots={slots}: host enqueue {1e6*(t1-t0)/n:.2f} us/step, wall {1e6*(t2-t0)/n:.2f} us/step, device {1e3*e0.elapsed_time(e1)/n:.2f} us/step -> {n*B/(t2-t0):.3e} img/s")
# resident baseline on the same bytes: the batches already in device memory, identity indices, 512 steps per launch
xd = hx.view(-1, 784).to(dev); yd = hy.view(-1).to(dev)
ident = torch.arange(n * B, dtype=torch.int64, device=dev)
eng = E.TrainEngine(dev, B); eng.load_state_dict(E.init_params_like_reference())
class R: images, labels = xd, yd
for per in (128, 512):
    eng.train_steps(R, ident[:per * B], B, per)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s0 in range(0, n - per + 1, per):
        eng.train_steps(R, ident[s0 * B:(s0 + per) * B], B, per)
    e1.record(); torch.cuda.synchronize()
    print(f"B={B} resident, {per} steps/launch: device {1e3*e0.elapsed_time(e1)/(n // per * per):.2f} us/step")
