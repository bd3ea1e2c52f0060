"""The CPU port that bench.py times (oracle/ref_port.py) reproduces what the real reference
produced — so the baseline being timed is the reference's arithmetic.  CPU only."""
import numpy as np
import pytest
import torch

from oracle import ref_port
from oracle.determinism import MaskFeed
from pytorch_ddp_mnist_b200 import synth


@pytest.fixture(scope="module")
def data_dir(tmp_path_factory):
    d = tmp_path_factory.mktemp("nc")
    synth.write_default_files(str(d))
    return str(d)


def test_port_reproduces_reference_losses(golden, data_dir):
    g = golden["traj_ws1"]
    torch.set_num_threads(4)
    feed = MaskFeed().install()
    try:
        r = ref_port.port_train(data_dir, 64, steps=20, warmup=0, model_seed=0, record_losses=True)
    finally:
        feed.remove()
    assert np.abs(np.array(r["losses"], np.float32) - g["losses"][:20]).max() < 1e-5
    assert r["images"] == 20 * 64 and r["seconds"] > 0


def test_port_gloo_ws2_runs(data_dir):
    r = ref_port.time_port_gloo(data_dir, 64, steps=6, warmup=2, world_size=2, port=29671)
    assert r["images"] == 6 * 64 * 2 and r["images_per_s"] > 0 and r["world_size"] == 2
