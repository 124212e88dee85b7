#!/usr/bin/env python
"""Small driver for ncu captures: a few launches of one featurizer on a short batch.
    python tools/prof_run.py whisper128|whisper80|kaldi|cosy|resample|resample_kaiser [clips]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import expressive_speech_translation_b200 as fe  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "whisper128"
g = torch.Generator(device="cuda").manual_seed(1)
if kind.startswith("whisper"):
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    x = (0.1 * torch.randn(n, 480000, generator=g, device="cuda")).clamp_(-1, 1)
    f = lambda: fe.whisper_log_mel_batch(x, n_mels=int(kind[7:]))
elif kind.startswith("resample"):           # resample | resample_kaiser : 30-s clips of 44.1 kHz stereo int16 -> 16 kHz mono
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    x = torch.randint(-20000, 20000, (n, 1323000, 2), generator=g, dtype=torch.int16, device="cuda")
    kw = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="sinc_interp_kaiser") if kind.endswith("kaiser") else {}
    f = lambda: fe.condition_pcm(x, 44100, 16000, **kw)
elif kind == "kaldi":
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    x = (0.1 * torch.randn(n, 80000, generator=g, device="cuda")).clamp_(-1, 1)
    f = lambda: fe.kaldi_fbank_batch(x)
else:
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    x = (0.1 * torch.randn(n, 240000, generator=g, device="cuda")).clamp_(-1, 1)
    f = lambda: fe.cosy_mel_batch(x)
for _ in range(3):
    out = f()
torch.cuda.synchronize()
print(kind, tuple(out.shape), float(out.float().mean()))
