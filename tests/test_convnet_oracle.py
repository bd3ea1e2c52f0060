"""The ConvNet oracle (oracle/convnet_oracle.py) against torch itself on the CPU.

The model is not in the reference - parity unpinned; what can be pinned is that the
restatement agrees with the standard torch modules / optimizer it restates."""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import convnet_oracle as CO


class _Net(nn.Module):                      # the layer list BASELINE.json names (J1-J3), eval-mode forward
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        self.fc1 = nn.Linear(9216, 128)
        self.fc2 = nn.Linear(128, 10)

    def forward(self, x):
        x = F.relu(self.conv1(x))
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = torch.flatten(x, 1)
        x = F.relu(self.fc1(x))
        return F.log_softmax(self.fc2(x), dim=1)


def test_param_layout():
    net = _Net()
    assert tuple(net.state_dict().keys()) == CO.PARAM_KEYS
    assert tuple(tuple(v.shape) for v in net.state_dict().values()) == CO.PARAM_SHAPES
    assert sum(v.numel() for v in net.parameters()) == CO.N_PARAMS == 1_199_882
    assert CO.param_offsets() == (0, 288, 320, 18752, 18816, 1198464, 1198592, 1199872)
    p = CO.init_params(1)
    assert np.array_equal(CO.flatten_params(CO.unflatten_params(CO.flatten_params(p))), CO.flatten_params(p))


def test_forward_backward_matches_module():
    p = CO.init_params(3)
    net = _Net()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()})
    rng = np.random.default_rng(0)
    x = rng.standard_normal((5, 784)).astype(np.float32)
    y = rng.integers(0, 10, 5).astype(np.uint8)
    lp = net(torch.from_numpy(x).reshape(-1, 1, 28, 28))
    loss = F.nll_loss(lp, torch.from_numpy(y).long())
    loss.backward()
    l2, lp2, g2, _ = CO.forward_backward(p, x, y)
    assert abs(float(loss) - l2) < 1e-6
    assert np.abs(lp.detach().numpy() - lp2).max() < 1e-5
    for k, v in net.named_parameters():
        assert np.abs(v.grad.numpy() - g2[k]).max() < 1e-6, k


def test_dropout_masks_scale_and_gate():
    p = CO.init_params(4)
    rng = np.random.default_rng(1)
    x = rng.standard_normal((3, 784)).astype(np.float32)
    y = rng.integers(0, 10, 3).astype(np.uint8)
    ones1, ones2 = np.ones((3, 9216), np.uint8), np.ones((3, 128), np.uint8)
    l_eval, _, _, _ = CO.forward_backward(p, x, y, want_grads=False)
    l_ones, _, _, _ = CO.forward_backward(p, x, y, ones1, ones2, want_grads=False)
    assert l_eval != l_ones                                   # all-kept masks still scale by 1/0.75 and 2
    _, _, g, _ = CO.forward_backward(p, x, y, ones1, np.zeros((3, 128), np.uint8))
    assert np.all(g["fc1.weight"] == 0) and np.all(g["conv2.weight"] == 0) and np.any(g["fc2.bias"] != 0)


def test_adadelta_matches_torch_optimizer():
    rng = np.random.default_rng(2)
    p0 = rng.standard_normal(4097).astype(np.float32)
    t = torch.tensor(p0.copy(), requires_grad=True)
    opt = torch.optim.Adadelta([t], lr=1.0, rho=0.9, eps=1e-6)
    p, sq, acc = p0.copy(), np.zeros_like(p0), np.zeros_like(p0)
    for step in range(5):
        g = (rng.standard_normal(4097) * 10.0 ** rng.integers(-6, 2)).astype(np.float32)
        t.grad = torch.from_numpy(g.copy())
        opt.step()
        p, sq, acc = CO.adadelta_step(p, g, sq, acc)
        assert np.abs(t.detach().numpy() - p).max() <= 2e-7 * max(1.0, np.abs(p).max()), step
    st = opt.state[t]
    assert np.allclose(st["square_avg"].numpy(), sq, rtol=1e-6, atol=0)
    assert np.allclose(st["acc_delta"].numpy(), acc, rtol=1e-5, atol=1e-12)
