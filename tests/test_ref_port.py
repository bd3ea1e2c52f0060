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
    r = ref_port.time_port_gloo(data_dir, 64, steps=6, warmup=2, world_size=2, timeout_s=240.0)
    assert r["images"] == 6 * 64 * 2 and r["images_per_s"] > 0 and r["world_size"] == 2


def test_port_survives_a_launcher_environment(data_dir, monkeypatch):
    """Called from a torchrun worker the port's children used to inherit TORCHELASTIC_USE_AGENT_STORE (rank 0 then
    becomes a client of a store nobody serves) and hang: they must drop the launcher's environment."""
    monkeypatch.setenv("TORCHELASTIC_USE_AGENT_STORE", "True")
    monkeypatch.setenv("TORCHELASTIC_RUN_ID", "x")
    monkeypatch.setenv("RANK", "0")
    monkeypatch.setenv("LOCAL_RANK", "0")
    monkeypatch.setenv("WORLD_SIZE", "2")
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", "29999")               # nobody listens here
    monkeypatch.setenv("OMP_NUM_THREADS", "1")
    r = ref_port.time_port_gloo(data_dir, 64, steps=4, warmup=1, world_size=2, timeout_s=240.0)
    assert r["images"] == 4 * 64 * 2 and r["threads_per_rank"] >= 1
    import os
    assert os.environ["TORCHELASTIC_USE_AGENT_STORE"] == "True"       # the caller's environment is left alone


def test_port_reports_a_dead_worker_instead_of_hanging(tmp_path):
    with pytest.raises(RuntimeError, match="time_port_gloo"):
        ref_port.time_port_gloo(str(tmp_path / "no_such_dir"), 64, steps=2, warmup=1, world_size=2, timeout_s=120.0)


def test_reference_arm_under_torchrun(tmp_path):
    """`bench.py --impl reference` launched the way the driver launches it at N = 2 (python -m torch.distributed.run):
    rank 0 prints the JSON line, the other rank exits 0, nothing hangs.  Needs no GPU."""
    import json
    import os
    import subprocess
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29877", os.path.join(repo, "bench.py"), "--impl", "reference", "--gpus", "2",
           "--steps", "6", "--warmup", "2"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=420, cwd=str(tmp_path))
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-3000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, out.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["value"] > 0 and d["n_gpus"] == 2
    assert d["config"]["reference_world_size"] == 2 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["value"] == d["value"]
