"""Tensor-Memory kaldi fbank kernel (fe_kaldi3.cuh) on the GPU: fbank with and without CMN,
the round-1 shared-memory kernel as a cross-check, ragged clips, rows the box copy cannot deliver, and the fall-back for
waveforms a tensor map cannot describe.  Parity criterion: tests/conftest.py::kaldi_check (fp64 oracle)."""
import os

import numpy as np
import pytest
import torch

import oracle
from tests.conftest import kaldi_check, speech_like

pytestmark = pytest.mark.gpu


@pytest.fixture
def fe():
    import expressive_speech_translation_b200 as fe
    return fe


def _with_env(**kw):
    class _Ctx:
        def __enter__(self):
            self.old = {k: os.environ.get(k) for k in kw}
            for k, v in kw.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
        def __exit__(self, *a):
            for k, v in self.old.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
    return _Ctx()


@pytest.mark.parametrize("B,L", [(5, 80000), (3, 16000), (2, 5444), (7, 33333), (2, 110000)])
def test_kaldi3_raw_and_cmn(fe, B, L):
    x = (0.1 * torch.randn(B, L, generator=torch.Generator().manual_seed(B * 1000 + L))).clamp(-1, 1)
    raw = fe.kaldi_fbank_batch(x.cuda(), subtract_mean=False).cpu().numpy()
    cmn = fe.kaldi_fbank_batch(x.cuda(), subtract_mean=True).cpu().numpy()      # 110000 samples: the CMN's two-pass path
    kaldi_check(raw, x, "kaldi3")
    want = raw - raw.mean(axis=1, keepdims=True, dtype=np.float64)
    assert np.abs(cmn - want).max() <= 2e-5                      # the CMN itself: mean in fp32, fixed order
    for b in range(B):
        ref = oracle.kaldi_fbank(x[b], subtract_mean=True, dtype=torch.float64).numpy()
        r64 = oracle.kaldi_fbank(x[b], subtract_mean=False, dtype=torch.float64).numpy()
        r32 = oracle.kaldi_fbank(x[b], subtract_mean=False).numpy()
        assert np.abs(cmn[b] - ref).max() <= oracle.kaldi_log_tolerance(r32, r64)


def test_kaldi3_ragged_and_speech(fe):
    gen = torch.Generator().manual_seed(4)
    x = speech_like(gen, 4, 48000, 16000)
    lens = [48000, 12345, 400, 30001]
    got = fe.kaldi_fbank_batch(x.cuda(), lengths=lens, subtract_mean=True).cpu().numpy()
    raw = fe.kaldi_fbank_batch(x.cuda(), lengths=lens, subtract_mean=False).cpu().numpy()
    for b, n in enumerate(lens):
        T = 1 + (n - 400) // 160
        ref = oracle.kaldi_fbank(x[b, :n], subtract_mean=True, dtype=torch.float64).numpy()
        r64 = oracle.kaldi_fbank(x[b, :n], subtract_mean=False, dtype=torch.float64).numpy()
        r32 = oracle.kaldi_fbank(x[b, :n], subtract_mean=False).numpy()
        tol = oracle.kaldi_log_tolerance(r32, r64)
        assert np.abs(raw[b, :T] - r64).max() <= tol
        assert np.abs(got[b, :T] - ref).max() <= tol + 2e-5
        assert not got[b, T:].any() and not raw[b, T:].any()      # rows past a ragged clip's own frames are zero


def test_kaldi3_matches_round1_kernel_and_falls_back(fe):
    x = (0.1 * torch.randn(3, 24000, generator=torch.Generator().manual_seed(9))).clamp(-1, 1)
    new = fe.kaldi_fbank_batch(x.cuda(), subtract_mean=False).cpu().numpy()
    kaldi_check(new, x, "kaldi3")
    # a waveform view whose rows are not 16-byte aligned cannot be described by a tensor map: the library must fall
    # back to the shared-memory kernel and still be right
    buf = torch.zeros(3, 24004).cuda()
    view = buf[:, 1:24001]
    view.copy_(x)
    assert view.data_ptr() % 16 != 0
    old = fe.kaldi_fbank_batch(view, subtract_mean=False).cpu().numpy()
    kaldi_check(old, x, "kaldi round-1 kernel (fall-back)")
    for b in range(3):
        r64 = oracle.kaldi_fbank(x[b], subtract_mean=False, dtype=torch.float64).numpy()
        r32 = oracle.kaldi_fbank(x[b], subtract_mean=False).numpy()
        assert np.abs(new[b] - old[b]).max() <= 2 * oracle.kaldi_log_tolerance(r32, r64)


def test_kaldi3_result_does_not_depend_on_the_batch(fe):
    """the reference featurizes one clip per call: a clip inside a 296-clip batch must equal the clip alone, bit for bit"""
    B = 296
    g = torch.Generator(device="cuda").manual_seed(3456)
    x = (0.1 * torch.randn(B, 80000, generator=g, device="cuda")).clamp_(-1, 1)
    n0 = fe.launch_count()
    out = fe.kaldi_fbank_batch(x, subtract_mean=True)
    torch.cuda.synchronize()
    assert fe.launch_count() - n0 == 2                             # featurizer + CMN
    assert out.shape == (B, 498, 80) and float(out.mean(dim=1).abs().max()) <= 2e-5
    for b in (0, 147, 295):
        assert torch.equal(fe.kaldi_fbank_batch(x[b:b + 1], subtract_mean=True)[0], out[b])
    with _with_env(B200FEAT_KALDI_KERNEL="1"):                     # round-1 kernel: same features within the tolerance
        old = fe.kaldi_fbank_batch(x[:4], subtract_mean=True)
    assert float((old - out[:4]).abs().max()) <= 1e-3
