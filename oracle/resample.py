"""CPU oracle of the waveform conditioning in front of the featurizers (TEST INFRASTRUCTURE ONLY, see
oracle/__init__.py): int16 / float PCM -> mono mix -> band-limited resampling.

The reference does this with torchaudio:
  /root/reference/Backend/services/audio_processor.py:295-304  torchaudio.transforms.Resample(orig, 16000,
        lowpass_filter_width=128, rolloff=0.95, resampling_method='kaiser_window', beta=14.769656459379492)
  /root/reference/Backend/services/audio_processor.py:306-319  stereo -> mono: mean of the channels when they correlate
        (> 0.5), else mid + 0.3 * side = 0.65 L + 0.35 R
  SURVEY.md section 8(d) cfg1: harvard.wav -> mean(2 ch) -> torchaudio.functional.resample(44100 -> 16000) (defaults)
torchaudio IS importable in the authoring container, so `reference_resample` runs the real code
(site-packages/torchaudio/functional/functional.py:1435 `resample`, :1305 `_get_sinc_resample_kernel`, :1405
`_apply_sinc_resample_kernel`); `resample_restated` is the plain restatement of the same arithmetic (dense
polyphase sum, any dtype) that the tests pin against it and then use in float64 as ground truth.
"""
import math

import numpy as np
import torch


def sinc_kernel(orig_freq, new_freq, lowpass_filter_width=6, rolloff=0.99, resampling_method="sinc_interp_hann",
                beta=None, dtype=torch.float64):
    """(new, K) taps and `width`, functional.py:1305-1402 with dtype=None semantics (float64 arithmetic; this is what
    torchaudio.transforms.Resample stores -- rounded to float32 -- and what the reference therefore applies)."""
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    if lowpass_filter_width <= 0:
        raise ValueError("Low pass filter width should be positive.")
    base = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base)
    idx = torch.arange(-width, width + orig, dtype=torch.float64)[None] / orig
    # functional.py:1377 with dtype=None: `torch.arange(0, -new_freq, -1, dtype=None) / new_freq` is an int64 tensor
    # divided by an int, i.e. a FLOAT32 phase offset -p/new, which is then added to the float64 grid.  The deployed
    # taps carry that rounding (1e-6-level), so the oracle and the kernel's table reproduce it.
    t = (torch.arange(0, -new, -1)[:, None] / new).to(torch.float64) + idx
    t = (t * base).clamp(-lowpass_filter_width, lowpass_filter_width)
    if resampling_method in ("sinc_interp_hann", "sinc_interpolation"):
        window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    elif resampling_method in ("sinc_interp_kaiser", "kaiser_window"):
        if beta is None:
            beta = 14.769656459379492
        b = torch.tensor(float(beta))      # float32, as in the reference
        window = torch.i0(b * torch.sqrt(1 - (t / lowpass_filter_width) ** 2)) / torch.i0(b)
    else:
        raise ValueError(f"Invalid resampling method: {resampling_method}")
    t = t * math.pi
    k = torch.where(t == 0, torch.ones_like(t), t.sin() / t) * window * (base / orig)
    return k.to(dtype), width, orig, new


def resample_restated(x, orig_freq, new_freq, dtype=torch.float64, **kw):
    """(..., L) -> (..., ceil(new * L / orig)); zero padding `width` left, `width + orig` right, stride-`orig`
    correlation with every phase's taps, phases interleaved, cut to the target length (functional.py:1405-1432)."""
    if orig_freq == new_freq:
        return x.to(dtype)
    k, width, orig, new = sinc_kernel(orig_freq, new_freq, dtype=torch.float32 if dtype == torch.float32 else dtype, **kw)
    k = k.to(dtype)
    shape = x.shape
    x = x.reshape(-1, shape[-1]).to(dtype)
    L = x.shape[-1]
    xp = torch.nn.functional.pad(x, (width, width + orig))
    y = torch.nn.functional.conv1d(xp[:, None], k[:, None], stride=orig)       # (N, new, blocks)
    y = y.transpose(1, 2).reshape(x.shape[0], -1)
    target = -(-new * L // orig)
    return y[:, :target].reshape(shape[:-1] + (target,))


def reference_resample(x, orig_freq, new_freq, **kw):
    """the real torchaudio call (float32 in -> float32 out), exactly as SURVEY.md's cfg1 recipe makes it"""
    import torchaudio
    return torchaudio.functional.resample(x, orig_freq, new_freq, **kw)


def reference_resample_transform(x, orig_freq, new_freq, **kw):
    """torchaudio.transforms.Resample, the class /root/reference/Backend/services/audio_processor.py:296 builds
    (taps computed in float64 and stored as float32)"""
    import torchaudio
    return torchaudio.transforms.Resample(orig_freq=orig_freq, new_freq=new_freq, **kw)(x)


def mono_mix(pcm, weights=None, dtype=torch.float64):
    """(B, L, C) int16 or float PCM -> (B, L): int16 is scaled by 1/32768 (whisper.audio.load_audio), channels are
    averaged (audio_processor.py:314) or weighted (0.65 / 0.35 = mid + 0.3 side, :316-318)."""
    x = pcm.to(dtype)
    if pcm.dtype == torch.int16:
        x = x / 32768.0
    if x.dim() == 2:
        x = x[..., None]
    C = x.shape[-1]
    w = torch.full((C,), 1.0 / C, dtype=dtype) if weights is None else torch.as_tensor(weights, dtype=dtype)
    return (x * w).sum(-1)


def condition(pcm, orig_freq, new_freq, weights=None, dtype=torch.float64, **kw):
    """mono mix + resample: the whole front row, ground truth in float64"""
    return resample_restated(mono_mix(pcm, weights, dtype), orig_freq, new_freq, dtype=dtype, **kw)


def reference_stereo_to_mono(audio):
    """/root/reference/Backend/services/audio_processor.py:306-319, the reference's own torch operations in its order,
    on a (2, L) tensor of RESAMPLED channels.  Returns ((1, L) mono, correlation, took_the_mean_branch)."""
    correlation = torch.sum(audio[0] * audio[1]) / torch.sqrt(                       # :309-311
        torch.sum(audio[0] ** 2) * torch.sum(audio[1] ** 2))
    if correlation > 0.5:                                                            # :313
        return torch.mean(audio, dim=0, keepdim=True), float(correlation), True     # :314
    mid = (audio[0] + audio[1]) / 2                                                  # :316
    side = (audio[0] - audio[1]) / 2                                                 # :317
    return (mid + 0.3 * side).unsqueeze(0), float(correlation), False               # :318


def reference_condition_stereo(audio, orig_freq, new_freq=16000):
    """audio_processor.py:294-319 on one (2, L) float32 clip: torchaudio.transforms.Resample with the reference's kaiser
    filter (the real class, float32), then the stereo -> mono rule above."""
    if orig_freq != new_freq:
        audio = reference_resample_transform(audio, orig_freq, new_freq, lowpass_filter_width=128, rolloff=0.95,
                                             resampling_method="sinc_interp_kaiser", beta=14.769656459379492)
    return reference_stereo_to_mono(audio)
