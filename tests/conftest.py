import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    class _G:
        def __getitem__(self, name):
            return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return _G()


@pytest.fixture(scope="session")
def synth_train():
    from pytorch_ddp_mnist_b200 import synth
    return synth.synth(synth.TRAIN_N, synth.TRAIN_SEED)


@pytest.fixture(scope="session")
def synth_test():
    from pytorch_ddp_mnist_b200 import synth
    return synth.synth(synth.TEST_N, synth.TEST_SEED)
