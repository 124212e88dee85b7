"""Mel filterbanks of the three featurizers (oracle; test infrastructure only).

Slaney filters: restates librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax,
htk=False, norm="slaney") -- the generator of openai-whisper's mel_filters.npz and
the `librosa_mel_fn` call in matcha/utils/audio.py::mel_spectrogram (bound at
/root/reference/greek_sft.yaml:55-63).  Same arithmetic as
transformers.audio_utils.mel_filter_bank(norm="slaney", mel_scale="slaney")
(SURVEY.md appendix A.1).

Kaldi banks: restates torchaudio.compliance.kaldi.get_mel_banks
(site-packages/torchaudio/compliance/kaldi.py:436-511, mel scale :318-331) for
vtln_warp == 1.0, which is the only branch the reference's CosyVoice frontend
reaches (Docker/debug_trigger_test_cosyvoice.py:102,258 -> campplus.onnx input).
"""
import math
import os

import numpy as np


def _hz_to_mel_slaney(f):
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3.0
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    lin = f / f_sp
    with np.errstate(divide="ignore", invalid="ignore"):
        log = min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep
    return np.where(f >= min_log_hz, log, lin)


def _mel_to_hz_slaney(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3.0
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    lin = f_sp * m
    log = min_log_hz * np.exp(logstep * (m - min_log_mel))
    return np.where(m >= min_log_mel, log, lin)


def slaney_mel_filters(sr, n_fft, n_mels, fmin=0.0, fmax=None, dtype=np.float32):
    """(n_mels, n_fft//2+1) Slaney-scale, Slaney-area-normalised triangles.

    Computed in float64 and cast to `dtype` last (librosa computes the ramps in
    float64 and stores float32; whisper ships that float32 array)."""
    if fmax is None:
        fmax = sr / 2.0
    n_bins = n_fft // 2 + 1
    fftfreqs = np.arange(n_bins, dtype=np.float64) * (float(sr) / n_fft)
    mels = np.linspace(_hz_to_mel_slaney(fmin), _hz_to_mel_slaney(fmax), n_mels + 2)
    mel_f = _mel_to_hz_slaney(mels)
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    w = np.zeros((n_mels, n_bins), dtype=np.float64)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0.0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2 : n_mels + 2] - mel_f[:n_mels])
    w *= enorm[:, None]
    return w.astype(dtype)


_KALDI_TABLE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kaldi_banks_80x257_f32.npy")


def kaldi_mel_banks(num_bins=80, padded_window=512, sample_freq=16000.0, low_freq=20.0,
                    high_freq=0.0, dtype=np.float32, recompute=False):
    """(num_bins, padded_window//2 + 1) kaldi triangles in the mel domain; the last
    (Nyquist) column is the zero column torchaudio pads on (kaldi.py:627).

    The reference's filter IS a float32 table: torchaudio builds it with float32 torch ops, whose `log`
    may differ by an ulp between CPU types, and one ulp of a mel value moves a weight by ~7e-6 (1e-4
    relative on the narrow low filters).  So for the reference configuration dtype=float32 returns the
    table torchaudio.compliance.kaldi.get_mel_banks(80, 512, 16000, 20, 0, 100, -500, 1.0) produced in
    the authoring container (committed as oracle/kaldi_banks_80x257_f32.npy by oracle/make_golden.py;
    tests compare it with a live torchaudio when that is importable).  recompute=True (or any other
    configuration) rebuilds it with numpy float32 arithmetic, step by step as torchaudio does;
    dtype=float64 is the analytic higher-precision twin."""
    if (dtype == np.float32 and not recompute and (num_bins, padded_window, sample_freq, low_freq, high_freq)
            == (80, 512, 16000.0, 20.0, 0.0) and os.path.exists(_KALDI_TABLE)):
        return np.load(_KALDI_TABLE)
    ft = np.float32 if dtype == np.float32 else np.float64
    nyquist = 0.5 * sample_freq
    if high_freq <= 0.0:
        high_freq += nyquist
    num_fft_bins = padded_window // 2
    fft_bin_width = sample_freq / padded_window
    mel_low = 1127.0 * math.log(1.0 + low_freq / 700.0)
    mel_high = 1127.0 * math.log(1.0 + high_freq / 700.0)
    delta = (mel_high - mel_low) / (num_bins + 1)
    b = np.arange(num_bins, dtype=ft)[:, None]
    # torch: python-float scalars combine with float32 tensors in float32
    left = (ft(mel_low) + b * ft(delta)).astype(ft)
    center = (ft(mel_low) + (b + ft(1.0)) * ft(delta)).astype(ft)
    right = (ft(mel_low) + (b + ft(2.0)) * ft(delta)).astype(ft)
    freqs = (ft(fft_bin_width) * np.arange(num_fft_bins, dtype=ft)).astype(ft)
    mel = (ft(1127.0) * np.log((ft(1.0) + freqs / ft(700.0)).astype(ft)).astype(ft)).astype(ft)[None, :]
    up = (mel - left) / (center - left)
    down = (right - mel) / (right - center)
    bins = np.maximum(ft(0.0), np.minimum(up, down)).astype(ft)
    out = np.zeros((num_bins, num_fft_bins + 1), dtype=ft)
    out[:, :num_fft_bins] = bins
    return out.astype(dtype)
