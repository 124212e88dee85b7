"""CPU suite: the oracle against the golden fixtures, the real third-party implementations (when they are
importable) and closed-form known answers."""
import numpy as np
import pytest
import torch

import oracle
from oracle import numpy_ref


def test_slaney_filters_match_hf_and_baked(golden):
    baked = golden("baked_filters.npz")
    for nm, key in ((80, "whisper80"), (128, "whisper128")):
        f = oracle.slaney_mel_filters(16000, 400, nm, 0.0, 8000.0)
        assert f.shape == (nm, 201)
        assert np.array_equal(f, baked[key])
        assert not f[:, 200].any()          # bin 200 carries no weight
        assert (np.count_nonzero(f, axis=0) <= 2).all()
    cz = oracle.slaney_mel_filters(24000, 1920, 80, 0.0, 8000.0)
    assert np.array_equal(cz, baked["cosy80"])
    assert not cz[:, 640:].any()
    tfa = pytest.importorskip("transformers.audio_utils")
    hf = tfa.mel_filter_bank(num_frequency_bins=201, num_mel_filters=80, min_frequency=0.0, max_frequency=8000.0,
                             sampling_rate=16000, norm="slaney", mel_scale="slaney")
    assert np.array_equal(hf.T.astype(np.float32), oracle.slaney_mel_filters(16000, 400, 80, 0.0, 8000.0))


def test_kaldi_banks_match_torchaudio_and_baked(golden):
    kb = oracle.kaldi_mel_banks()
    assert kb.shape == (80, 257) and not kb[:, 256].any() and not kb[:, 0].any()
    baked = golden("baked_filters.npz")["kaldi80"]
    # torch's float32 log may differ by an ulp between CPU types; the effect on a weight is <= ~2e-5
    assert np.abs(kb - baked).max() < 5e-5
    ta = pytest.importorskip("torchaudio")
    tb, _ = ta.compliance.kaldi.get_mel_banks(80, 512, 16000.0, 20.0, 0.0, 100.0, -500.0, 1.0)
    assert np.abs(kb[:, :256] - tb.numpy()).max() < 5e-5


def test_whisper_harvard_fixture(golden):
    g = golden("harvard_w80.npz")
    wav = g["wave16k"]
    assert wav.shape == (293700,)
    o32 = oracle.whisper_log_mel(wav, 80).numpy()
    assert o32.shape == (80, 1835)
    # the oracle's fp32 path is HF's torch path op for op; allow ulp-level FFT backend differences
    assert np.abs(o32 - g["hf_torch"]).max() < 5e-5
    o64 = oracle.whisper_log_mel(wav, 80, dtype=torch.float64).numpy()
    assert np.abs(o64 - g["oracle_fp64"]).max() < 1e-6
    assert np.abs(o32 - o64).max() < 1e-4


@pytest.mark.parametrize("name", ["noise", "speech"])
@pytest.mark.parametrize("nm", [80, 128])
def test_whisper_random_fixture(golden, name, nm):
    g = golden("whisper_rand.npz")
    x = g[f"{name}_wave"]
    o64 = oracle.whisper_log_mel(x, nm, dtype=torch.float64, norm="per_clip").numpy()
    assert np.abs(o64 - g[f"{name}_fp64_{nm}"]).max() < 1e-6
    assert np.abs(o64 - g[f"{name}_hf_torch_{nm}"]).max() < 1e-4
    twin = np.stack([numpy_ref.whisper_log_mel_np(x[b], nm) for b in range(x.shape[0])])
    assert np.abs(twin - o64).max() < 1e-9


def test_whisper_against_hf_live():
    tf = pytest.importorskip("transformers")
    x = (0.1 * torch.randn(2, 16000, generator=torch.Generator().manual_seed(7))).clamp(-1, 1)
    fe = tf.WhisperFeatureExtractor(feature_size=80)
    hf = fe._torch_extract_fbank_features(x.numpy())
    mine = oracle.whisper_log_mel(x, 80, norm="per_clip").numpy()
    assert np.abs(hf - mine).max() < 1e-6


def test_whisper_closed_forms():
    z = oracle.whisper_log_mel(np.zeros(16000, np.float32), 80).numpy()
    assert z.shape == (80, 100) and np.all(z == -1.5)
    t = np.arange(48000) / 16000.0
    s = oracle.whisper_log_mel((0.5 * np.sin(2 * np.pi * 440 * t)).astype(np.float32), 80).numpy()
    assert s.shape == (80, 300)
    assert abs((s.max() - s.min()) - 2.0) < 1e-5          # the 8-decade clamp is active
    assert int(np.argmax(s.mean(axis=1))) == 11           # 440 Hz lands in mel bin 11
    # padding / frame count / per-call vs per-clip
    x = (0.1 * torch.randn(2, 12345, generator=torch.Generator().manual_seed(3)))
    a = oracle.whisper_log_mel(x, 128, padding=777)
    assert a.shape == (2, 128, (12345 + 777) // 160)
    pc = oracle.whisper_log_mel(x, 80, norm="per_clip")
    for b in range(2):
        assert torch.equal(pc[b], oracle.whisper_log_mel(x[b], 80))
    assert oracle.whisper_num_frames(480000) == 3000 and oracle.whisper_num_frames(960000, 480000) == 9000


@pytest.mark.parametrize("name", ["noise", "speech"])
def test_kaldi_fixture(golden, name):
    g = golden("kaldi_rand.npz")
    x = g[f"{name}_wave"]
    for b in range(x.shape[0]):
        o32 = oracle.kaldi_fbank(x[b], subtract_mean=False).numpy()
        assert np.abs(o32 - g[f"{name}_torchaudio"][b]).max() < 2e-4
        o64 = oracle.kaldi_fbank(x[b], subtract_mean=True, dtype=torch.float64).numpy()
        assert np.abs(o64 - g[f"{name}_fp64_cmn"][b]).max() < 1e-5
        assert np.abs(numpy_ref.kaldi_fbank_np(x[b]) - o64).max() < 1e-9


def test_kaldi_against_torchaudio_live_and_closed_forms():
    ta = pytest.importorskip("torchaudio")
    x = (0.1 * torch.randn(1, 20000, generator=torch.Generator().manual_seed(11))).clamp(-1, 1)
    ref = ta.compliance.kaldi.fbank(x, num_mel_bins=80, dither=0, sample_frequency=16000)
    mine = oracle.kaldi_fbank(x, subtract_mean=False)
    assert mine.shape == ref.shape == (123, 80)
    assert (ref - mine).abs().max() < 1e-4
    z = oracle.kaldi_fbank(torch.zeros(1, 4000), subtract_mean=False)
    assert torch.allclose(z, torch.full_like(z, float(np.log(1.1920928955078125e-07))))
    assert oracle.kaldi_fbank(torch.zeros(1, 4000), subtract_mean=True).abs().max() < 1e-5
    assert oracle.kaldi_num_frames(80000) == 498 and oracle.kaldi_num_frames(399) == 0
    with pytest.raises(AssertionError):
        oracle.kaldi_fbank(torch.zeros(1, 399))


@pytest.mark.parametrize("name", ["noise", "speech"])
def test_cosy_fixture(golden, name):
    g = golden("cosy_rand.npz")
    x = g[f"{name}_wave"]
    o64 = oracle.cosyvoice_mel(x, dtype=torch.float64).numpy()
    assert o64.shape == (2, 80, 50)
    assert np.abs(o64 - g[f"{name}_fp64"]).max() < 1e-5
    assert np.abs(oracle.cosyvoice_mel(x).numpy() - o64).max() < 1e-4
    for b in range(2):
        assert np.abs(numpy_ref.cosyvoice_mel_np(x[b]) - o64[b]).max() < 1e-9


def _cosy_independent(x):
    """Two implementations of the CosyVoice / Matcha mel that share no code with oracle.cosyvoice_mel:
    torchaudio.transforms.MelSpectrogram and transformers.audio_utils.spectrogram, both fed the reflect-padded
    signal ((n_fft - hop) / 2 = 720 each side) with the parameter block of /root/reference/greek_sft.yaml:55-63
    (n_fft 1920, num_mels 80, sampling_rate 24000, hop_size 480, win_size 1920, fmin 0, fmax 8000, center False).
    Neither has matcha's `+ 1e-9` under the square root -- bounded analytically by the caller."""
    ta = pytest.importorskip("torchaudio")
    au = pytest.importorskip("transformers.audio_utils")
    x = torch.as_tensor(x, dtype=torch.float64)
    xp = torch.nn.functional.pad(x.unsqueeze(1), (720, 720), mode="reflect").squeeze(1)
    ms = ta.transforms.MelSpectrogram(sample_rate=24000, n_fft=1920, win_length=1920, hop_length=480, f_min=0.0,
                                      f_max=8000.0, n_mels=80, power=1.0, center=False, norm="slaney",
                                      mel_scale="slaney").to(torch.float64)
    mel_ta = ms(xp).numpy()
    mf = au.mel_filter_bank(961, 80, 0.0, 8000.0, 24000, norm="slaney", mel_scale="slaney")
    win = au.window_function(1920, "hann", periodic=True)
    mel_hf = np.stack([au.spectrogram(xp[b].numpy(), win, frame_length=1920, hop_length=480, fft_length=1920, power=1.0,
                                      center=False, mel_filters=mf, mel_floor=1e-30) for b in range(x.shape[0])])
    return mel_ta, mel_hf


@pytest.mark.parametrize("name", ["noise", "speech"])
def test_cosy_pinned_by_two_independent_implementations(golden, name):
    """§8(c) a2: the Matcha mel is not importable here, so the restatement is pinned against torchaudio's
    MelSpectrogram and transformers' spectrogram (linear mel energies, then ln(max(., 1e-5)) as matcha does).

    The one constant neither twin carries is the 1e-9 under the square root: sqrt(m^2 + 1e-9) - m <= 5e-10 / m, so
    on a mel energy whose contributing magnitudes are >= 1e-2 the relative effect is <= 5e-6 (and <= 5e-8 for
    magnitudes >= 1e-1); it is bounded here as a relative tolerance on energies >= 1e-2 instead of being trusted."""
    g = golden("cosy_rand.npz")
    x = g[f"{name}_wave"]
    o64 = oracle.cosyvoice_mel(x, dtype=torch.float64).numpy()
    mel_ta, mel_hf = _cosy_independent(x)
    assert mel_ta.shape == mel_hf.shape == o64.shape
    M = np.exp(o64)
    strong = M >= 1e-2
    assert strong.mean() > 0.5            # the check must cover most of the fixture
    # transformers' filter bank is bit-equal to the librosa formula: only the 1e-9 term separates them
    assert (np.abs(mel_hf - M)[strong] / M[strong]).max() <= 5e-6
    # torchaudio builds its slaney filter bank in float32 (weights off by ~1e-6 absolute: 6e-6 relative on noise,
    # 3e-5 on tones that sit on a filter's skirt)
    assert (np.abs(mel_ta - M)[strong] / M[strong]).max() <= 1e-4
    assert np.abs(np.log(np.maximum(mel_hf, 1e-5)) - o64)[strong].max() <= 1e-5
    assert np.abs(np.log(np.maximum(mel_ta, 1e-5)) - o64)[strong].max() <= 1e-4
    assert np.abs(np.log(np.maximum(mel_hf, 1e-5)) - o64).max() <= 1e-4       # weak bins: the 1e-9 term + floor
    # and the stored arrays of the generating run (oracle/make_golden.py) say the same
    assert np.abs(g[f"{name}_torchaudio_mel"] - mel_ta).max() <= 1e-6 * mel_ta.max()
    assert np.abs(g[f"{name}_hf_mel"] - mel_hf).max() <= 1e-6 * mel_hf.max()


def test_cosy_closed_forms():
    z = oracle.cosyvoice_mel(torch.zeros(1, 24000))
    assert z.shape == (1, 80, 50) and torch.allclose(z, torch.full_like(z, float(np.log(1e-5))))
    assert oracle.cosyvoice_num_frames(240000) == 500 and oracle.cosyvoice_num_frames(479) == 0
