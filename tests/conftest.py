import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "slow: multi-second CPU oracle runs")


def pkg():
    """The product package (its directory name carries a hyphen, so import by string)."""
    return importlib.import_module("hunyuanworld-mirror_b200")


@pytest.fixture(scope="session")
def wmpkg():
    return pkg()


@pytest.fixture(scope="session")
def state_dict_full():
    w = importlib.import_module("hunyuanworld-mirror_b200.weights")
    return w.synthetic_state_dict(seed=0)


@pytest.fixture(scope="session")
def c1_tokens(state_dict_full):
    """token_list of the oracle aggregator on config C1 (2 views 252 x 252, no priors); shared by the head oracles' tests."""
    import torch
    from oracle import aggregator_oracle as O
    w = importlib.import_module("hunyuanworld-mirror_b200.weights")
    with torch.no_grad():
        outs, _ = O.aggregator_forward(w.synthetic_images(2, 252, 252, seed=1), state_dict_full)
    return outs
