"""Spectral-gate denoiser (/root/reference/Backend/services/audio_processor.py:321-358).
The oracle is the reference's own torch lines (oracle/gate.py).  CPU: the kernels' phase functions under g++ against it;
GPU: the kernels through the C-ABI against it."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import gate
from tests.conftest import speech_like
from tests.emu import build as emu_build

TOL = 2e-5        # on a waveform normalised to a peak of 0.95; measured 3e-7 (the reference's own float32 is 2.5e-7 from float64)
vp = ctypes.c_void_p


def _clips(B, L, seed):
    g = torch.Generator().manual_seed(seed)
    x = speech_like(g, B, L, 16000).float()
    x += 0.03 * torch.randn(B, L, generator=g)            # a noise floor for the gate to find
    return x.clamp(-1, 1).contiguous()


def test_oracle_is_the_reference_lines_and_keeps_their_frame_range_boost():
    """audio_processor.py:338-340 takes `mag.shape[1]` of a (513, T) tensor: the 1.2 x boost covers frames
    [int(T/16), int(T/4)) for every bin.  Pin that reading: the output changes if the boost is moved to the bins."""
    x = _clips(1, 16000, 3)[0]
    ref = gate.reference_spectral_gate(x)
    assert ref.shape == (1, 256 * (16000 // 256)) and abs(float(ref.abs().max()) - 0.95) < 1e-6
    w = torch.hann_window(1024)
    spec = torch.stft(x, 1024, 256, 1024, window=w, return_complex=True)
    mag = spec.abs()
    T = mag.shape[1]
    nf = torch.quantile(mag, 0.1, dim=1, keepdim=True)
    gm = torch.clamp((mag - 1.5 * nf) / (mag + 1e-8), 0, 1)
    by_frames, by_bins = gm.clone(), gm.clone()
    by_frames[:, int(T * 0.0625):int(T * 0.25)] *= 1.2
    by_bins[int(513 * 0.0625):int(513 * 0.25)] *= 1.2
    def fin(m):
        y = torch.istft(spec * m, 1024, 256, 1024, window=w)
        return y / y.abs().max() * 0.95
    assert (fin(by_frames) - ref[0]).abs().max() < 1e-5
    assert (fin(by_bins) - ref[0]).abs().max() > 1e-3


@pytest.fixture(scope="module")
def emu():
    return emu_build.load()


def emu_gate(emu, x, lengths=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    out = np.full((B, 256 * (L // 256)), 7.0, dtype=np.float32)
    la = None if lengths is None else np.asarray(lengths, dtype=np.int64)
    assert emu.emu_gate(x.ctypes.data_as(vp), None if la is None else la.ctypes.data_as(vp), B, L, L, out.ctypes.data_as(vp)) == 0
    return out


def test_emulated_kernels_against_the_reference_lines(emu):
    x = _clips(3, 16000 + 77, 5)
    x[2] *= 0.02
    got = emu_gate(emu, x.numpy())
    for b in range(3):
        ref = gate.reference_spectral_gate(x[b])[0].numpy()
        assert got[b].shape == ref.shape and np.abs(got[b] - ref).max() <= TOL, b
    # ragged: a clip's own 256 * (len // 256) samples, zeros behind; <= 512 samples and silence give zeros
    lens = [16077, 9000, 400]
    got = emu_gate(emu, x.numpy(), lens)
    for b, n in enumerate(lens[:2]):
        ref = gate.reference_spectral_gate(x[b, :n])[0].numpy()
        assert np.abs(got[b, :ref.shape[0]] - ref).max() <= TOL
        assert not got[b, ref.shape[0]:].any()
    assert not got[2].any()
    assert not emu_gate(emu, np.zeros((1, 4000), np.float32)).any()      # max_amp == 0: left as it is (all zeros)


@pytest.fixture(scope="module")
def fe():
    import expressive_speech_translation_b200 as fe_
    return fe_


@pytest.mark.gpu
def test_gpu_against_the_reference_lines(fe):
    x = _clips(5, 16000 * 3 + 123, 7)
    x[3] *= 0.01
    lens = [x.shape[1], 20000, 513, 48000, 512]
    out = fe.spectral_gate_batch(x.cuda(), lengths=lens)
    assert out.shape == (5, 256 * (x.shape[1] // 256))
    assert torch.equal(out, fe.spectral_gate_batch(x.cuda(), lengths=lens))              # deterministic: no atomics on samples
    out = out.cpu()
    for b, n in enumerate(lens):
        if n <= 512:
            assert not out[b].any()
            continue
        ref = gate.reference_spectral_gate(x[b, :n])[0]
        assert (out[b, :ref.shape[0]] - ref).abs().max() <= TOL, (b, float((out[b, :ref.shape[0]] - ref).abs().max()))
        assert not out[b, ref.shape[0]:].any()
        assert abs(float(out[b].abs().max()) - 0.95) < 1e-6
    # a clip's result does not depend on its batch; unaligned rows; silence; the refusal of rows that cannot be reflect-padded
    assert torch.equal(fe.spectral_gate_batch(x[1:2, :20000].contiguous().cuda()).cpu()[0], out[1, :256 * (20000 // 256)])
    y = torch.zeros(5, x.shape[1] + 1)
    y[:, 1:] = x
    assert torch.equal(fe.spectral_gate_batch(y.cuda()[:, 1:], lengths=lens).cpu(), out)
    assert not fe.spectral_gate_batch(torch.zeros(2, 5000).cuda()).any()
    with pytest.raises(ValueError):
        fe.spectral_gate_batch(torch.zeros(1, 512).cuda())


@pytest.mark.gpu
def test_gpu_long_clip_more_frames_than_the_quantile_cache(fe):
    """40 s = 2501 frames: the quantile kernel keeps rows of up to 2048 magnitudes in shared memory and re-reads longer ones"""
    x = _clips(2, 16000 * 40, 11)
    out = fe.spectral_gate_batch(x.cuda(), lengths=[16000 * 40, 16000 * 25]).cpu()
    for b, n in enumerate([16000 * 40, 16000 * 25]):
        ref = gate.reference_spectral_gate(x[b, :n])[0]
        assert (out[b, :ref.shape[0]] - ref).abs().max() <= TOL, b


@pytest.mark.gpu
def test_gpu_full_size_properties(fe):
    """256 x 30 s: finite, peak 0.95 per clip, deterministic, one clip against the reference lines"""
    g = torch.Generator(device="cuda").manual_seed(2)
    x = (0.2 * torch.randn(256, 480000, generator=g, device="cuda")).clamp_(-1, 1)
    t = torch.arange(480000, device="cuda") / 16000.0
    x += 0.5 * torch.sin(2 * torch.pi * 440.0 * t) * (torch.sin(2 * torch.pi * 0.5 * t) > 0)
    out = fe.spectral_gate_batch(x)
    assert out.shape == (256, 480000) and bool(torch.isfinite(out).all())
    assert bool(((out.abs().amax(dim=1) - 0.95).abs() < 1e-6).all())
    assert torch.equal(out, fe.spectral_gate_batch(x))
    ref = gate.reference_spectral_gate(x[101].cpu())[0]
    assert (out[101].cpu() - ref).abs().max() <= TOL
