"""CPU restatement of the Wav2Lip / Diff2Lip mel (oracle; TEST INFRASTRUCTURE ONLY -- nothing under
expressive_speech_translation_b200/ may import this).

SURVEY.md section 8(f) item 3.  This is the one mel on the path whose source is held by the reference itself:

    /root/reference/Backend/diff2lip/audio/audio.py:45-51    melspectrogram(wav)
    /root/reference/Backend/diff2lip/audio/audio.py:19-22    preemphasis      (scipy.signal.lfilter([1, -k], [1], wav))
    /root/reference/Backend/diff2lip/audio/audio.py:57-61    _stft            (librosa.stft(y, n_fft, hop, win_length))
    /root/reference/Backend/diff2lip/audio/audio.py:92-103   _linear_to_mel / _build_mel_basis (librosa.filters.mel)
    /root/reference/Backend/diff2lip/audio/audio.py:105-107  _amp_to_db
    /root/reference/Backend/diff2lip/audio/audio.py:112-118  _normalize       (symmetric, clipped)
    /root/reference/Backend/diff2lip/audio/hparams.py:22-65  the parameters below
    call sites: /root/reference/Backend/diff2lip/generate.py:53, generate_dist.py:49  (`audio.melspectrogram(wav).T`)

Third-party dependency absent from this image: librosa, pinned `librosa==0.9.2` at
/root/reference/Backend/requirements.txt:14.  What the reference takes from it, restated from its published
algorithm:
  librosa.stft (0.9.2: center=True, pad_mode="reflect", window="hann" = scipy.signal.get_window("hann", win_length,
      fftbins=True) i.e. the PERIODIC Hann, dtype None -> complex of the input's precision): frames
      y_padded[t * hop : t * hop + n_fft] * window, t = 0 .. len(y) // hop, rfft.  (librosa >= 0.10 changed the default
      pad_mode to "constant"; `pad_mode` below selects either -- the pin says "reflect".)
  librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax) (htk=False, norm="slaney"): oracle.filters.slaney_mel_filters,
      the same restatement that is pinned bit for bit to openai-whisper's mel_filters.npz (made by this very librosa
      function) through transformers' copy of it (tests/test_oracle.py).
scipy.signal.lfilter IS importable here and is called as the reference calls it.

Precision: lfilter with Python-float coefficients returns float64 for a float32 wav, librosa.stft keeps that precision
(complex128), the float32 mel basis is promoted by np.dot: the reference computes this mel in float64 and returns
float64.  `dtype=np.float64` (default) is therefore the reference's own arithmetic, not a ground-truth stand-in.

Pinned (tests/test_oracle.py::test_wav2lip_*, oracle/make_golden.py): against two independent implementations of
everything but the lfilter call -- torch.stft + torchaudio.functional.melscale_fbanks, and
transformers.audio_utils.spectrogram + mel_filter_bank -- on seeded inputs, arrays stored in tests/golden/w2l_*.npz.
"""
import numpy as np

from .filters import slaney_mel_filters

# /root/reference/Backend/diff2lip/audio/hparams.py:22-65
NUM_MELS = 80            # :23
N_FFT = 800              # :32
HOP_SIZE = 200           # :33
WIN_SIZE = 800           # :34
SAMPLE_RATE = 16000      # :35
MAX_ABS_VALUE = 4.0      # :46   (signal_normalization, allow_clipping_in_normalization, symmetric_mels: all True, :40-43)
PREEMPHASIS = 0.97       # :55   (preemphasize=True, :54)
MIN_LEVEL_DB = -100      # :58
REF_LEVEL_DB = 20        # :59
FMIN = 55                # :60
FMAX = 7600              # :63


def wav2lip_num_frames(n_samples):
    """librosa.stft(center=True): 1 + len(y) // hop_length frames."""
    return 1 + int(n_samples) // HOP_SIZE


def preemphasis(wav, k=PREEMPHASIS):
    """audio.py:19-22 -- the call the reference makes, scipy's own lfilter (float64 out for a float32 wav)."""
    from scipy import signal
    return signal.lfilter([1, -k], [1], wav)


def _reflect_pad(y, pad, mode):
    if mode == "reflect":
        return np.pad(y, (pad, pad), mode="reflect")
    if mode == "constant":
        return np.pad(y, (pad, pad), mode="constant")
    raise ValueError(f"pad_mode {mode!r}")


def stft_magnitude(y, pad_mode="reflect", dtype=np.float64):
    """|librosa.stft(y, n_fft=800, hop_length=200, win_length=800)| (audio.py:57-61, :47 `np.abs(D)`): (401, 1 + len // 200)."""
    y = np.asarray(y, dtype=dtype)
    if pad_mode == "reflect" and y.shape[0] <= N_FFT // 2:
        raise ValueError("reflect padding needs more than n_fft // 2 samples")
    yp = _reflect_pad(y, N_FFT // 2, pad_mode)
    n = np.arange(WIN_SIZE, dtype=np.float64)
    window = (0.5 - 0.5 * np.cos(2.0 * np.pi * n / WIN_SIZE)).astype(dtype)      # periodic Hann (fftbins=True)
    T = wav2lip_num_frames(y.shape[0])
    idx = HOP_SIZE * np.arange(T)[:, None] + np.arange(N_FFT)[None, :]
    frames = yp[idx] * window[None, :]
    return np.abs(np.fft.rfft(frames, axis=1)).T.astype(dtype)


def mel_basis():
    """audio.py:99-103: librosa.filters.mel(sr=16000, n_fft=800, n_mels=80, fmin=55, fmax=7600), float32 (80, 401)."""
    assert FMAX <= SAMPLE_RATE // 2        # audio.py:100
    return slaney_mel_filters(SAMPLE_RATE, N_FFT, NUM_MELS, float(FMIN), float(FMAX), dtype=np.float32)


def amp_to_db(x):
    """audio.py:105-107"""
    min_level = np.exp(MIN_LEVEL_DB / 20 * np.log(10))
    return 20 * np.log10(np.maximum(min_level, x))


def normalize(S):
    """audio.py:112-118, the branch hparams select (allow_clipping_in_normalization and symmetric_mels)."""
    return np.clip((2 * MAX_ABS_VALUE) * ((S - MIN_LEVEL_DB) / (-MIN_LEVEL_DB)) - MAX_ABS_VALUE, -MAX_ABS_VALUE, MAX_ABS_VALUE)


def wav2lip_mel(wav, pad_mode="reflect", dtype=np.float64, return_db=False):
    """audio.py:45-51 melspectrogram(wav): (80, 1 + len(wav) // 200).  wav: 1-D, 16 kHz, float.
    dtype=np.float32 runs the transform in float32 (what a float32 implementation can be held to)."""
    wav = np.asarray(wav)
    if wav.ndim != 1:
        raise ValueError("wav must be 1-D")
    y = preemphasis(wav.astype(np.float32) if dtype == np.float32 else wav)       # :46
    D = stft_magnitude(y, pad_mode=pad_mode, dtype=dtype)                            # :46-47
    S = amp_to_db(np.dot(mel_basis().astype(dtype), D)) - REF_LEVEL_DB               # :47
    if return_db:
        return S
    return normalize(S)                                                              # :49-50
