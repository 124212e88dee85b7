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


def kaldi_check(out_raw, x, what=""):
    """Parity criterion for the kaldi fbank (raw = before CMN) against the fp64 oracle.

    float32 FFTs (the reference's included) carry an error ~1e-7 of the FRAME's energy, so the log of a
    nearly empty mel bin (e.g. 20-50 Hz after pre-emphasis, 1e-8 of the frame maximum) is off by a few
    1e-4 in the reference itself (tests/golden/kaldi_rand.npz: torchaudio fp32 vs fp64 = 4.5e-4).  Hence
      (a) linear mel energies: |M - M64| <= 1e-5 x max_m M64[frame]            (north_star: power, 1e-5)
      (b) log features with M64 >= 1e-2 x frame max: |f - f64| <= 1e-4          (north_star: 1e-4)
      (c) every log feature: |f - f64| <= oracle.kaldi_log_tolerance = max(1e-4, KALDI_TOL_FACTOR (4) x
          max|f32_reference - f64|) on the same input -- the one place the factor is written is oracle/__init__.py.
    out_raw: (B, T, 80) numpy; x: (B, L) torch CPU tensor."""
    import numpy as np
    import torch
    import oracle
    for b in range(x.shape[0]):
        f64 = oracle.kaldi_fbank(x[b], subtract_mean=False, dtype=torch.float64).numpy()
        f32 = oracle.kaldi_fbank(x[b], subtract_mean=False).numpy()
        got = out_raw[b, :f64.shape[0]].astype(np.float64)
        M64, M = np.exp(f64), np.exp(got)
        fmax = M64.max(axis=1, keepdims=True)
        assert (np.abs(M - M64) / fmax).max() <= 1e-5, f"{what} clip {b}: linear mel energy"
        strong = M64 >= 1e-2 * fmax
        assert np.abs(got - f64)[strong].max() <= 1e-4, f"{what} clip {b}: log features of the strong bins"
        bound = oracle.kaldi_log_tolerance(f32, f64)
        assert np.abs(got - f64).max() <= bound, f"{what} clip {b}: {np.abs(got - f64).max()} > {bound}"
