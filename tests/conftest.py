import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(autouse=True)
def _inference_tests_run_without_autograd(request):
    """Everything except the training tests exercises the inference path: run it under torch.no_grad() (with grad mode on,
    WeCLIP.forward takes the trainable-decoder path and returns tensors that carry an autograd graph)."""
    import torch
    needs_grad = request.module.__name__.rsplit(".", 1)[-1] in ("test_train_gpu", "test_training_gpu", "test_training_cpu")
    with torch.set_grad_enabled(needs_grad):
        yield
