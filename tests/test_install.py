"""install() / uninstall(): the switch §8(b) names.  whisper and matcha are not importable in this image, so stand-in
modules in sys.modules exercise the patch logic for every target install() lists; the -m gpu test then calls the way
whisper.transcribe does (`log_mel_spectrogram(audio, n_mels, padding=N_SAMPLES)` resolved in its own namespace)."""
import sys
import types

import numpy as np
import pytest
import torch

import expressive_speech_translation_b200 as fe
from expressive_speech_translation_b200 import api


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


@pytest.fixture
def stub_modules(monkeypatch):
    def upstream_log_mel(audio, n_mels=80, padding=0, device=None):
        raise AssertionError("the upstream function must not run once install() has patched it")

    def upstream_mel(y, n_fft, num_mels, sampling_rate, hop_size, win_size, fmin, fmax, center=False):
        raise AssertionError("the upstream function must not run once install() has patched it")

    whisper = _stub("whisper", log_mel_spectrogram=upstream_log_mel)
    whisper.audio = _stub("whisper.audio", log_mel_spectrogram=upstream_log_mel, N_SAMPLES=480000, N_FRAMES=3000)
    # whisper/transcribe.py does `from .audio import log_mel_spectrogram`: its own binding must be patched too
    whisper.transcribe = _stub("whisper.transcribe", log_mel_spectrogram=upstream_log_mel)
    matcha = _stub("matcha")
    matcha.utils = _stub("matcha.utils")
    matcha.utils.audio = _stub("matcha.utils.audio", mel_spectrogram=upstream_mel)
    for m in (whisper, whisper.audio, whisper.transcribe, matcha, matcha.utils, matcha.utils.audio):
        monkeypatch.setitem(sys.modules, m.__name__, m)
    return dict(upstream_log_mel=upstream_log_mel, upstream_mel=upstream_mel, whisper=whisper, matcha=matcha)


def test_install_patches_every_listed_target_and_uninstall_restores(stub_modules):
    import torchaudio
    before = (torchaudio.compliance.kaldi.fbank, torchaudio.functional.resample, torchaudio.transforms.Resample)
    done = fe.install()
    try:
        assert set(done) == {"whisper.audio.log_mel_spectrogram", "whisper.transcribe.log_mel_spectrogram",
                             "whisper.log_mel_spectrogram", "matcha.utils.audio.mel_spectrogram",
                             "torchaudio.compliance.kaldi.fbank", "torchaudio.functional.resample",
                             "torchaudio.transforms.Resample"}
        w = stub_modules["whisper"]
        assert w.audio.log_mel_spectrogram is api.log_mel_spectrogram
        assert w.transcribe.log_mel_spectrogram is api.log_mel_spectrogram
        assert w.log_mel_spectrogram is api.log_mel_spectrogram
        assert stub_modules["matcha"].utils.audio.mel_spectrogram is api.mel_spectrogram
        assert torchaudio.compliance.kaldi.fbank is api.fbank
        assert torchaudio.transforms.Resample is api.Resample
    finally:
        fe.uninstall()
    w = stub_modules["whisper"]
    assert w.audio.log_mel_spectrogram is stub_modules["upstream_log_mel"]
    assert w.transcribe.log_mel_spectrogram is stub_modules["upstream_log_mel"]
    assert stub_modules["matcha"].utils.audio.mel_spectrogram is stub_modules["upstream_mel"]
    assert (torchaudio.compliance.kaldi.fbank, torchaudio.functional.resample, torchaudio.transforms.Resample) == before


def test_install_twice_then_uninstall_is_clean(stub_modules):
    fe.install()
    fe.install()
    fe.uninstall()
    assert stub_modules["whisper"].audio.log_mel_spectrogram is stub_modules["upstream_log_mel"]


def test_resample_is_an_nn_module_like_the_class_it_replaces():
    """install() swaps torchaudio.transforms.Resample process-wide: hosts that do more than call it must keep working"""
    r = fe.Resample(orig_freq=44100, new_freq=16000, lowpass_filter_width=128, rolloff=0.95,
                    resampling_method="kaiser_window", beta=14.769656459379492)     # audio_processor.py:296-303
    assert isinstance(r, torch.nn.Module)
    assert r.state_dict() == {} and r.to("cpu") is r and r.eval() is r
    seq = torch.nn.Sequential(r)
    assert list(seq.children()) == [r]
    assert "orig_freq=44100" in repr(r)
    x = torch.zeros(2, 100)
    assert fe.Resample(16000, 16000)(x) is x                                         # equal rates: the input itself


def test_hyperpyyaml_style_partial_of_mel_spectrogram():
    """greek_sft.yaml:55-63 binds the keyword set with `!name:`, i.e. functools.partial; feat_extractor(y) must be
    callable with the waveform alone and reject other parameter sets loudly"""
    import functools
    feat_extractor = functools.partial(fe.mel_spectrogram, n_fft=1920, num_mels=80, sampling_rate=24000, hop_size=480,
                                       win_size=1920, fmin=0, fmax=8000, center=False)
    with pytest.raises(ValueError):
        feat_extractor(torch.zeros(24000))                      # (L,) instead of (B, L): rejected before any device work
    bad = functools.partial(fe.mel_spectrogram, n_fft=1024, num_mels=80, sampling_rate=22050, hop_size=256,
                            win_size=1024, fmin=0, fmax=8000, center=False)
    with pytest.raises(NotImplementedError):
        bad(torch.zeros(1, 24000))


@pytest.mark.gpu
def test_gpu_patched_transcribe_namespace_call(stub_modules):
    """whisper.transcribe's call: mel = log_mel_spectrogram(audio, model.dims.n_mels, padding=N_SAMPLES), then 3000-frame
    segments (UPSTREAM whisper/transcribe.py; reached from /root/reference/Backend/services/cascaded_backend.py:158)"""
    import oracle
    fe.install()
    try:
        tr = stub_modules["whisper"].transcribe
        audio = (0.1 * torch.randn(48000, generator=torch.Generator().manual_seed(5))).clamp(-1, 1)
        for n_mels in (80, 128):
            mel = tr.log_mel_spectrogram(audio.numpy(), n_mels, padding=480000)     # numpy in -> host tensor out
            assert mel.shape == (n_mels, 3300) and not mel.is_cuda
            ref = oracle.whisper_log_mel(audio, n_mels, padding=480000, dtype=torch.float64)
            assert (mel.double() - ref).abs().max() <= 1e-4
            mel_d = tr.log_mel_spectrogram(audio.cuda(), n_mels, padding=480000)    # device in -> device out
            assert mel_d.is_cuda and torch.equal(mel_d.cpu(), mel)
        y = (0.1 * torch.randn(2, 24000, generator=torch.Generator().manual_seed(6))).clamp(-1, 1)
        m = stub_modules["matcha"].utils.audio.mel_spectrogram(y.cuda(), 1920, 80, 24000, 480, 1920, 0, 8000, center=False)
        assert (m.cpu().double() - oracle.cosyvoice_mel(y, dtype=torch.float64)).abs().max() <= 1e-4
    finally:
        fe.uninstall()


@pytest.mark.gpu
def test_gpu_cosy_prompt_feat_trim():
    import oracle
    y = (0.1 * torch.randn(1, 24000 * 3 + 100, generator=torch.Generator().manual_seed(8))).clamp(-1, 1)
    feat, tl = fe.cosy_prompt_feat(y.cuda(), token_len=70)
    ref = oracle.cosyvoice_mel(y, dtype=torch.float64)[0].T                  # (T, 80)
    T = ref.shape[0]
    assert tl == min(T // 2, 70) and feat.shape == (1, 2 * tl, 80)
    assert (feat[0].cpu().double() - ref[:2 * tl]).abs().max() <= 1e-4
    feat_all, tl_all = fe.cosy_prompt_feat(y.cuda())
    assert feat_all.shape == (1, T, 80) and tl_all == T // 2
