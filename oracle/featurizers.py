"""CPU restatement of the three featurizers on the hot path (oracle; test infrastructure only).

Every function computes with the SAME torch CPU ops the reference's third-party
code calls (torch.stft / torch.fft.rfft / matmul / clamp / log), in float32 (what
the reference computes) or float64 (ground truth for tolerance checks).

a1  whisper_log_mel  : openai-whisper whisper/audio.py::log_mel_spectrogram
                       (UPSTREAM, not vendored; reached from
                       /root/reference/Backend/services/cascaded_backend.py:158,
                       /root/reference/Backend/speech_translation_pipeline.py:140,
                       /root/reference/Backend/gtts_pipeline_fixed.py:64).  Line-for-line
                       twin installed here:
                       transformers/models/whisper/feature_extraction_whisper.py:135-164.
a2  cosyvoice_mel    : matcha/utils/audio.py::mel_spectrogram (UPSTREAM), parameters
                       bound at /root/reference/greek_sft.yaml:55-63, reached from
                       /root/reference/Docker/cosyvoice_api.py:128.
a3  kaldi_fbank      : torchaudio/compliance/kaldi.py:514-645 (fbank), :154-217
                       (_get_window), :44-83 (_get_strided) + the CosyVoice frontend's
                       `feat - feat.mean(dim=0, keepdim=True)` (UPSTREAM call site;
                       model at /root/reference/Docker/debug_trigger_test_cosyvoice.py:102).
"""
import numpy as np
import torch

from .filters import slaney_mel_filters, kaldi_mel_banks

WHISPER_SR = 16000
WHISPER_N_FFT = 400
WHISPER_HOP = 160

COSY_SR = 24000
COSY_N_FFT = 1920
COSY_HOP = 480
COSY_N_MELS = 80
COSY_FMIN = 0.0
COSY_FMAX = 8000.0

KALDI_SR = 16000
KALDI_WIN = 400
KALDI_SHIFT = 160
KALDI_PADDED = 512
KALDI_N_MELS = 80
KALDI_PREEMPH = 0.97
KALDI_EPS = 1.1920928955078125e-07  # torch.finfo(float32).eps, kaldi.py:33,633


def whisper_num_frames(n_samples, padding=0):
    """stft(center=True) gives 1 + L//160 frames, the last one is dropped."""
    return (int(n_samples) + int(padding)) // WHISPER_HOP


def cosyvoice_num_frames(n_samples):
    """reflect pad 720 each side, stft(center=False): 1 + (L + 1440 - 1920)//480."""
    L = int(n_samples)
    if L + 2 * ((COSY_N_FFT - COSY_HOP) // 2) < COSY_N_FFT:
        return 0
    return 1 + (L + 2 * ((COSY_N_FFT - COSY_HOP) // 2) - COSY_N_FFT) // COSY_HOP


def kaldi_num_frames(n_samples):
    """snip_edges=True (kaldi.py:67-71)."""
    L = int(n_samples)
    if L < KALDI_WIN:
        return 0
    return 1 + (L - KALDI_WIN) // KALDI_SHIFT


def _as_tensor(x, dtype):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    return x.detach().to("cpu").to(dtype)


def whisper_log_mel(audio, n_mels=80, padding=0, dtype=torch.float32, norm="per_call",
                    return_power=False):
    """audio: (L,) or (B, L) in [-1, 1].  Returns (n_mels, T) or (B, n_mels, T).

    norm="per_call": max over the whole tensor of the call (whisper/audio.py semantics);
    norm="per_clip": max per batch item (HF feature_extraction_whisper.py:156-158).
    """
    assert n_mels in (80, 128), f"Unsupported n_mels: {n_mels}"
    x = _as_tensor(audio, dtype)
    if padding > 0:
        x = torch.nn.functional.pad(x, (0, padding))
    window = torch.hann_window(WHISPER_N_FFT, dtype=dtype)
    stft = torch.stft(x, WHISPER_N_FFT, WHISPER_HOP, window=window, return_complex=True)
    magnitudes = stft[..., :-1].abs() ** 2
    if return_power:
        return magnitudes
    filt64 = slaney_mel_filters(WHISPER_SR, WHISPER_N_FFT, n_mels, 0.0, 8000.0, dtype=np.float32)
    filters = torch.from_numpy(filt64).to(dtype)  # the float32 table is THE filter, also in fp64 mode
    mel_spec = filters @ magnitudes
    log_spec = torch.clamp(mel_spec, min=1e-10).log10()
    if norm == "per_clip" and log_spec.dim() == 3:
        mx = log_spec.amax(dim=(1, 2), keepdim=True)
    else:
        mx = log_spec.max()
    log_spec = torch.maximum(log_spec, mx - 8.0)
    log_spec = (log_spec + 4.0) / 4.0
    return log_spec


def cosyvoice_mel(y, n_fft=COSY_N_FFT, num_mels=COSY_N_MELS, sampling_rate=COSY_SR,
                  hop_size=COSY_HOP, win_size=COSY_N_FFT, fmin=COSY_FMIN, fmax=COSY_FMAX,
                  center=False, dtype=torch.float32, return_magnitude=False):
    """y: (B, L) at 24 kHz.  Returns (B, num_mels, T)."""
    x = _as_tensor(y, dtype)
    squeeze = x.dim() == 1
    if squeeze:
        x = x[None]
    filt = slaney_mel_filters(sampling_rate, n_fft, num_mels, fmin, fmax, dtype=np.float32)
    mel_basis = torch.from_numpy(filt).to(dtype)
    window = torch.hann_window(win_size, dtype=dtype)
    pad = int((n_fft - hop_size) / 2)
    x = torch.nn.functional.pad(x.unsqueeze(1), (pad, pad), mode="reflect").squeeze(1)
    spec = torch.view_as_real(
        torch.stft(x, n_fft, hop_length=hop_size, win_length=win_size, window=window,
                   center=center, pad_mode="reflect", normalized=False, onesided=True,
                   return_complex=True))
    spec = torch.sqrt(spec.pow(2).sum(-1) + 1e-9)
    if return_magnitude:
        return spec[0] if squeeze else spec
    spec = torch.matmul(mel_basis, spec)
    spec = torch.log(torch.clamp(spec, min=1e-5) * 1)
    return spec[0] if squeeze else spec


def kaldi_fbank(waveform, subtract_mean=True, dtype=torch.float32, return_power=False):
    """waveform: (L,) or (1, L) at 16 kHz in [-1, 1].  Returns (m, 80), frame-major.

    Restates torchaudio.compliance.kaldi.fbank(num_mel_bins=80, dither=0,
    sample_frequency=16000) step by step (so it can also run in float64) followed by
    the CosyVoice frontend's CMN when subtract_mean=True."""
    x = _as_tensor(waveform, dtype)
    if x.dim() == 2:
        x = x[0]  # channel 0, kaldi.py:135-137
    L = x.numel()
    assert 2 <= KALDI_WIN <= L, f"choose a window size {KALDI_WIN} that is [2, {L}]"  # kaldi.py:142
    m = kaldi_num_frames(L)
    frames = x.as_strided((m, KALDI_WIN), (KALDI_SHIFT, 1))  # kaldi.py:82-83
    frames = frames - frames.mean(dim=1, keepdim=True)  # remove_dc_offset, kaldi.py:183-186
    off = torch.nn.functional.pad(frames.unsqueeze(0), (1, 0), mode="replicate").squeeze(0)
    frames = frames - KALDI_PREEMPH * off[:, :-1]  # kaldi.py:193-198
    window = torch.hann_window(KALDI_WIN, periodic=False, dtype=dtype).pow(0.85)  # povey, :100
    frames = frames * window[None]
    frames = torch.nn.functional.pad(frames, (0, KALDI_PADDED - KALDI_WIN))  # :207-211
    spectrum = torch.fft.rfft(frames).abs().pow(2.0)  # :616-618
    if return_power:
        return spectrum
    banks = kaldi_mel_banks(KALDI_N_MELS, KALDI_PADDED, float(KALDI_SR), 20.0, 0.0,
                            dtype=np.float32)
    banks = torch.from_numpy(banks).to(dtype)
    mel = torch.mm(spectrum, banks.T)  # :630
    eps = torch.tensor(KALDI_EPS, dtype=dtype)
    feat = torch.max(mel, eps).log()  # :633
    if subtract_mean:
        feat = feat - feat.mean(dim=0, keepdim=True)
    return feat
