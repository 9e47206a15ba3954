import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    import torch

    cache = {}

    def load(name):
        if name not in cache:
            cache[name] = torch.load(os.path.join(GOLDEN, name), weights_only=False)
        return cache[name]

    return load


def rel_err(a, b):
    """||a - b|| / ||b|| in fp64 — the per-op parity metric."""
    import torch

    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    denom = torch.linalg.vector_norm(b)
    if denom == 0:
        return float(torch.linalg.vector_norm(a - b))
    return float(torch.linalg.vector_norm(a - b) / denom)


def close(a, b, rtol, atol=1e-7):
    """Norm-wise closeness with an absolute floor (per element) for mathematically-zero tensors,
    e.g. the key-bias gradient of softmax attention."""
    import torch

    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    diff = float(torch.linalg.vector_norm(a - b))
    return diff <= rtol * float(torch.linalg.vector_norm(b)) + atol * (b.numel() ** 0.5)
