"""Wav2Lip / Diff2Lip mel (/root/reference/Backend/diff2lip/audio/audio.py:45-51).
CPU: the oracle against its committed goldens and two independent implementations; the baked filter bank; the
emulated kernel against the goldens.  GPU: the kernel through the C-ABI against the oracle."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import wav2lip as w2l
from tests.conftest import speech_like
from tests.emu import build as emu_build

NAMES = ("noise", "speech", "quiet", "harvard")
TOL = 1e-4      # in the normalised [-4, 4] scale (0.08 per dB): 1.25e-3 dB


# ------------------------------------------------------------------------------------------------ oracle (CPU)
@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_its_golden_and_both_twins(golden, name):
    g = golden("w2l_rand.npz")
    x = g[name + "_wave"]
    ref = w2l.wav2lip_mel(x)
    assert ref.dtype == np.float64 and ref.shape == (80, 1 + len(x) // 200)
    assert np.abs(ref - g[name + "_oracle_f64"]).max() <= 1e-9
    assert np.abs(w2l.wav2lip_mel(x, pad_mode="constant") - g[name + "_oracle_zeros_f64"]).max() <= 1e-9
    assert np.abs(ref - g[name + "_twin_hf"]).max() <= 2e-6          # transformers.audio_utils.spectrogram
    assert np.abs(ref - g[name + "_twin_torch"]).max() <= 3e-5       # torch.stft + torchaudio melscale_fbanks (float32 bank)
    assert ref.min() >= -4.0 and ref.max() <= 4.0


def test_oracle_against_the_twins_live():
    """the same two independent implementations, run here on a fresh seed (both libraries are on both machines)"""
    pytest.importorskip("torchaudio")
    pytest.importorskip("transformers")
    from oracle.make_golden_w2l import twin_hf, twin_torch
    x = speech_like(torch.Generator().manual_seed(77), 1, 12345, 16000)[0].numpy().astype(np.float32)
    ref = w2l.wav2lip_mel(x)
    assert np.abs(twin_hf(x) - ref).max() <= 2e-6
    assert np.abs(twin_torch(x) - ref).max() <= 3e-5


def test_oracle_closed_forms_and_steps():
    # audio.py:19-22: y[n] = x[n] - 0.97 x[n-1], y[0] = x[0]
    x = np.arange(1, 6, dtype=np.float32)
    assert np.allclose(w2l.preemphasis(x), [1.0, 2 - 0.97, 3 - 1.94, 4 - 2.91, 5 - 3.88])
    assert w2l.preemphasis(x).dtype == np.float64                     # the reference computes this mel in float64
    # silence: 20 log10(1e-5) - 20 = -120 dB -> 8 * (-20 / 100) - 4 = -5.6 -> clipped to -4
    assert np.all(w2l.wav2lip_mel(np.zeros(2000, np.float32)) == -4.0)
    assert w2l.wav2lip_num_frames(16000) == 81 and w2l.wav2lip_num_frames(199) == 1
    # a full-scale sine in the pass band: one bright row near its frequency, level from first principles
    t = np.arange(16000) / 16000.0
    s = np.sin(2 * np.pi * 1000.0 * t).astype(np.float32)
    m = w2l.wav2lip_mel(s)
    fb = w2l.mel_basis()
    assert fb.shape == (80, 401) and fb.dtype == np.float32 and not fb[:, :2].any() and not fb[:, 381:].any()
    row = int(np.argmax(m[:, 40]))
    assert fb[row, 50] > 0                                             # 1 kHz = bin 50 of 800 at 16 kHz
    gain = abs(1 - 0.97 * np.exp(-2j * np.pi * 1000 / 16000))          # pre-emphasis at 1 kHz
    peak = 0.5 * 400 * gain                                            # |X[50]| of a unit sine under a periodic Hann (sum w / 2)
    # the window's main lobe spreads over bins 49..51 with weights (1/2, 1, 1/2) of the peak
    mel_val = peak * (fb[row, 50] + 0.5 * fb[row, 49] + 0.5 * fb[row, 51])
    want = np.clip(8 * ((20 * np.log10(mel_val) - 20 + 100) / 100) - 4, -4, 4)
    assert abs(m[row, 40] - want) < 2e-3
    with pytest.raises(ValueError):
        w2l.wav2lip_mel(np.zeros(400, np.float32))                     # reflect padding needs more than 400 samples


def test_baked_filter_bank_is_the_oracles(golden):
    assert np.array_equal(golden("baked_filters.npz")["w2l80"], w2l.mel_basis())
    tfa = pytest.importorskip("transformers.audio_utils")
    hf = tfa.mel_filter_bank(num_frequency_bins=401, num_mel_filters=80, min_frequency=55.0, max_frequency=7600.0,
                             sampling_rate=16000, norm="slaney", mel_scale="slaney")
    assert np.abs(hf.T - w2l.mel_basis()).max() <= 1e-9


# ------------------------------------------------------------------------------------------------ emulated kernel (CPU)
@pytest.fixture(scope="module")
def emu():
    return emu_build.load()


def test_emulated_kernel_against_the_goldens(emu, golden):
    g = golden("w2l_rand.npz")
    vp = ctypes.c_void_p
    for name in NAMES:
        x = np.ascontiguousarray(g[name + "_wave"], dtype=np.float32)[None]
        T = 1 + x.shape[1] // 200
        for pad_mode, key in ((0, "_oracle_f64"), (1, "_oracle_zeros_f64")):
            out = np.full((1, 80, T), 7.0, dtype=np.float32)
            assert emu.emu_wav2lip(x.ctypes.data_as(vp), None, 1, x.shape[1], x.shape[1], out.ctypes.data_as(vp), T, 0, pad_mode) == 0
            assert np.abs(out[0] - g[name + key]).max() <= TOL, (name, pad_mode)


# ------------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def fe():
    import expressive_speech_translation_b200 as fe_
    return fe_


@pytest.mark.gpu
def test_gpu_goldens_layouts_ragged(fe, golden):
    g = golden("w2l_rand.npz")
    for name in NAMES:
        x = torch.from_numpy(g[name + "_wave"])[None]
        got = fe.wav2lip_mel_batch(x.cuda()).cpu().numpy()
        assert got.shape == (1, 80, 1 + x.shape[1] // 200)
        assert np.abs(got[0] - g[name + "_oracle_f64"]).max() <= TOL, name
        got = fe.wav2lip_mel_batch(x.cuda(), pad_mode="constant").cpu().numpy()
        assert np.abs(got[0] - g[name + "_oracle_zeros_f64"]).max() <= TOL, name
    gen = torch.Generator().manual_seed(5)
    x = torch.cat([(0.1 * torch.randn(3, 20011, generator=gen)).clamp(-1, 1), speech_like(gen, 4, 20011, 16000).float()])
    lens = [20011, 401, 777, 20010, 16000, 4001, 12345]
    out = fe.wav2lip_mel_batch(x.cuda(), lengths=lens)
    fm = fe.wav2lip_mel_batch(x.cuda(), lengths=lens, frame_major=True)
    assert torch.equal(fm, out.transpose(1, 2).contiguous())           # generate.py:53 takes melspectrogram(wav).T
    out = out.cpu().numpy()
    for b, n in enumerate(lens):
        ref = w2l.wav2lip_mel(x[b, :n].numpy())
        assert np.abs(out[b, :, :ref.shape[1]] - ref).max() <= TOL, b
        assert not out[b, :, ref.shape[1]:].any()
    # unaligned rows take the element-wise loader: same bits
    y = torch.zeros(7, 20012)
    y[:, 1:] = x
    assert torch.equal(fe.wav2lip_mel_batch(y.cuda()[:, 1:], lengths=lens).cpu(), torch.from_numpy(out))
    # silence and the refusals
    assert bool((fe.wav2lip_mel_batch(torch.zeros(2, 5000).cuda()) == -4.0).all())
    with pytest.raises(ValueError):
        fe.wav2lip_mel_batch(torch.zeros(1, 400).cuda())               # reflect padding needs more than 400 samples
    assert fe.wav2lip_mel_batch(torch.zeros(1, 400).cuda(), pad_mode="constant").shape == (1, 80, 3)


@pytest.mark.gpu
def test_gpu_drop_in_melspectrogram_and_install(fe):
    """`audio.melspectrogram(wav)` as /root/reference/Backend/diff2lip/generate.py:53 calls it: numpy in, numpy (80, T) out;
    install() patches a module named audio.audio (the reference does `from audio import audio`)."""
    import sys
    import types
    x = speech_like(torch.Generator().manual_seed(3), 1, 16000 * 3, 16000)[0].numpy().astype(np.float32)
    m = fe.melspectrogram(x)
    assert isinstance(m, np.ndarray) and m.shape == (80, 241) and m.dtype == np.float32
    assert np.abs(m - w2l.wav2lip_mel(x)).max() <= TOL
    # the crop the caller makes: 16 mel steps per window at 80 steps per second (generate.py:35-42)
    assert m.T[10:26].shape == (16, 80)
    pkg, mod = types.ModuleType("audio"), types.ModuleType("audio.audio")
    mod.melspectrogram = lambda wav: None
    pkg.audio = mod
    sys.modules["audio"], sys.modules["audio.audio"] = pkg, mod
    try:
        assert "audio.audio.melspectrogram" in fe.install()
        from audio import audio
        assert np.array_equal(audio.melspectrogram(x), m)
    finally:
        fe.uninstall()
        del sys.modules["audio"], sys.modules["audio.audio"]
    assert mod.melspectrogram(x) is None


@pytest.mark.gpu
def test_gpu_full_size_properties(fe):
    """256 x 10 s: determinism, range, and batch independence (a clip's mel does not depend on its neighbours)."""
    g = torch.Generator(device="cuda").manual_seed(9)
    x = (0.1 * torch.randn(256, 160000, generator=g, device="cuda")).clamp_(-1, 1)
    out = fe.wav2lip_mel_batch(x)
    assert out.shape == (256, 80, 801) and bool(torch.isfinite(out).all())
    assert float(out.min()) >= -4.0 and float(out.max()) <= 4.0
    assert torch.equal(out, fe.wav2lip_mel_batch(x))
    assert torch.equal(out[37:38], fe.wav2lip_mel_batch(x[37:38]))
    ref = w2l.wav2lip_mel(x[200].cpu().numpy())
    assert np.abs(out[200].cpu().numpy() - ref).max() <= TOL
