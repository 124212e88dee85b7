import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return load


def speech_like(gen, B, L, sr):
    """same generator as oracle/make_golden.py: 3 random sines x slow AM envelope + small noise."""
    import torch
    t = torch.arange(L, dtype=torch.float64) / sr
    out = torch.zeros(B, L, dtype=torch.float64)
    for b in range(B):
        f = 80.0 + 3000.0 * torch.rand(3, generator=gen, dtype=torch.float64)
        ph = 6.283185307179586 * torch.rand(3, generator=gen, dtype=torch.float64)
        env = 0.5 * (1.0 + torch.sin(6.283185307179586 * (0.5 + 2.5 * torch.rand(1, generator=gen, dtype=torch.float64)) * t))
        sig = sum(torch.sin(6.283185307179586 * f[i] * t + ph[i]) for i in range(3)) / 3.0
        out[b] = 0.8 * env * sig
    out += 0.01 * torch.randn(B, L, generator=gen, dtype=torch.float64)
    return out.clamp(-1, 1).to(torch.float32)
