"""Second, independent restatement in plain numpy float64 (oracle; test infrastructure only).

It does NOT call torch.stft: frames are cut by hand with the explicit padding rules
of SURVEY.md appendix A.4 and transformed with np.fft.rfft.  Its only purpose is to
catch a wrong edge rule (reflect without the edge sample, dropped last frame,
snip_edges frame count, replicate-left pre-emphasis ...) in `oracle/featurizers.py`
or in the CUDA kernels, whose framing is written the same explicit way.
"""
import numpy as np

from .filters import slaney_mel_filters, kaldi_mel_banks


def _reflect_index(i, n):
    """index into a length-n signal reflect-padded without repeating the edge sample."""
    i = np.where(i < 0, -i, i)
    i = np.where(i >= n, 2 * (n - 1) - i, i)
    return i


def whisper_log_mel_np(audio, n_mels=80, padding=0):
    x = np.asarray(audio, dtype=np.float64)
    assert x.ndim == 1
    x = np.concatenate([x, np.zeros(padding)])
    n = x.shape[0]
    T = n // 160
    idx = (np.arange(T)[:, None] * 160 - 200) + np.arange(400)[None, :]
    frames = x[_reflect_index(idx, n)]
    win = 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(400) / 400.0)
    spec = np.fft.rfft(frames * win[None], axis=1)
    power = spec.real ** 2 + spec.imag ** 2  # (T, 201)
    filt = slaney_mel_filters(16000, 400, n_mels, 0.0, 8000.0, dtype=np.float32).astype(np.float64)
    mel = power @ filt.T
    logs = np.log10(np.maximum(mel, 1e-10))
    logs = np.maximum(logs, logs.max() - 8.0)
    return ((logs + 4.0) / 4.0).T  # (n_mels, T)


def cosyvoice_mel_np(y):
    x = np.asarray(y, dtype=np.float64)
    assert x.ndim == 1
    n = x.shape[0]
    T = 1 + (n + 1440 - 1920) // 480
    idx = (np.arange(T)[:, None] * 480 - 720) + np.arange(1920)[None, :]
    frames = x[_reflect_index(idx, n)]
    win = 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(1920) / 1920.0)
    spec = np.fft.rfft(frames * win[None], axis=1)
    mag = np.sqrt(spec.real ** 2 + spec.imag ** 2 + 1e-9)
    filt = slaney_mel_filters(24000, 1920, 80, 0.0, 8000.0, dtype=np.float32).astype(np.float64)
    mel = mag @ filt.T
    return np.log(np.maximum(mel, 1e-5)).T  # (80, T)


def kaldi_fbank_np(wave, subtract_mean=True):
    x = np.asarray(wave, dtype=np.float64)
    assert x.ndim == 1
    n = x.shape[0]
    m = 1 + (n - 400) // 160
    idx = np.arange(m)[:, None] * 160 + np.arange(400)[None, :]
    fr = x[idx]
    fr = fr - fr.mean(axis=1, keepdims=True)
    prev = np.concatenate([fr[:, :1], fr[:, :-1]], axis=1)
    fr = fr - 0.97 * prev
    win = (0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(400) / 399.0)) ** 0.85
    fr = fr * win[None]
    spec = np.fft.rfft(fr, n=512, axis=1)
    power = spec.real ** 2 + spec.imag ** 2
    banks = kaldi_mel_banks(80, 512, 16000.0, 20.0, 0.0, dtype=np.float32).astype(np.float64)
    feat = np.log(np.maximum(power @ banks.T, 1.1920928955078125e-07))
    if subtract_mean:
        feat = feat - feat.mean(axis=0, keepdims=True)
    return feat  # (m, 80)
