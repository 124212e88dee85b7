#!/usr/bin/env python
"""Generates tests/golden/w2l_rand.npz (Wav2Lip / Diff2Lip mel) in the AUTHORING
container.  Run from the repo root:  python -m oracle.make_golden_w2l

The reference's own module (/root/reference/Backend/diff2lip/audio/audio.py) cannot be imported here: it imports
librosa at module level and librosa (pinned 0.9.2, Backend/requirements.txt:14) is not in this image.  What is stored
instead, per input clip:
  oracle_f64  oracle.wav2lip.wav2lip_mel -- audio.py:45-51 restated line by line, in the reference's float64 arithmetic,
              calling the same scipy.signal.lfilter the reference calls
  twin_torch  scipy lfilter -> torch.stft(800, 200, 800, periodic Hann, center=True, pad_mode="reflect") -> abs ->
              torchaudio.functional.melscale_fbanks(401, 55, 7600, 80, 16000, norm="slaney", mel_scale="slaney") -> dB, clip
  twin_hf     scipy lfilter -> transformers.audio_utils.spectrogram(power=1, center=True, pad_mode="reflect",
              mel_filters=mel_filter_bank(401, 80, 55, 7600, 16000, "slaney", "slaney"), log_mel="dB", min_value=1e-5)
two implementations of the STFT, the mel filter bank and the dB conversion that share no code with the oracle.
Inputs: seeded noise, two speech-like clips (one at 1 % level: bins at the 1e-5 floor), 2 s of
/root/reference/Backend/harvard.wav at 16 kHz (taken from tests/golden/harvard_w80.npz)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import wav2lip as w2l  # noqa: E402
from oracle.make_golden import speech_like  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def twin_torch(x):
    import torchaudio.functional as taf
    y = torch.from_numpy(w2l.preemphasis(x))
    D = torch.stft(y, 800, 200, 800, window=torch.hann_window(800, periodic=True, dtype=torch.float64), center=True,
                   pad_mode="reflect", return_complex=True).abs()
    fb = taf.melscale_fbanks(401, 55.0, 7600.0, 80, 16000, norm="slaney", mel_scale="slaney").double()
    S = 20 * torch.log10(torch.clamp(fb.t() @ D, min=1e-5)) - 20
    return torch.clamp(8 * ((S + 100) / 100) - 4, -4, 4).numpy()


def twin_hf(x):
    from transformers.audio_utils import mel_filter_bank, spectrogram, window_function
    mf = mel_filter_bank(401, 80, 55.0, 7600.0, 16000, norm="slaney", mel_scale="slaney")
    S = spectrogram(w2l.preemphasis(x), window_function(800, "hann", periodic=True), frame_length=800, hop_length=200,
                    fft_length=800, power=1.0, center=True, pad_mode="reflect", onesided=True, mel_filters=mf,
                    mel_floor=1e-5, log_mel="dB", reference=1.0, min_value=1e-5)
    return np.clip(8 * ((S - 20 + 100) / 100) - 4, -4, 4)


def main():
    gen = torch.Generator().manual_seed(20260)
    clips = {"noise": (0.1 * torch.randn(9000, generator=gen)).clamp(-1, 1).numpy().astype(np.float32)}
    sp = speech_like(gen, 2, 9000, 16000).numpy().astype(np.float32)
    clips["speech"] = sp[0]
    clips["quiet"] = (0.01 * sp[1]).astype(np.float32)
    clips["harvard"] = np.load(os.path.join(GOLD, "harvard_w80.npz"))["wave16k"][16000:48000].astype(np.float32)
    d = {}
    for name, x in clips.items():
        ref = w2l.wav2lip_mel(x)
        t1, t2 = twin_torch(x), twin_hf(x)
        assert ref.shape == t1.shape == t2.shape == (80, 1 + len(x) // 200)
        assert np.abs(t1 - ref).max() <= 3e-5, (name, np.abs(t1 - ref).max())      # torchaudio builds its bank in float32
        assert np.abs(t2 - ref).max() <= 2e-6, (name, np.abs(t2 - ref).max())
        d[name + "_wave"] = x
        d[name + "_oracle_f64"] = ref
        d[name + "_twin_torch"] = t1
        d[name + "_twin_hf"] = t2
        d[name + "_oracle_zeros_f64"] = w2l.wav2lip_mel(x, pad_mode="constant")
        print(name, x.shape, "twin_torch", np.abs(t1 - ref).max(), "twin_hf", np.abs(t2 - ref).max(),
              "f32 pipeline", np.abs(w2l.wav2lip_mel(x, dtype=np.float32) - ref).max())
    np.savez_compressed(os.path.join(GOLD, "w2l_rand.npz"), **d)


if __name__ == "__main__":
    main()
