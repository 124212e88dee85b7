#!/usr/bin/env python
"""Device-resident time of the four dtype variants of the Whisper kernel on the cfg2 shape."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import expressive_speech_translation_b200 as fe  # noqa: E402

g = torch.Generator(device="cuda").manual_seed(1)
x = (0.1 * torch.randn(1024, 480000, generator=g, device="cuda")).clamp_(-1, 1)
x16 = (x * 32767.0).round().to(torch.int16)
for name, w, dt in (("f32->f32", x, torch.float32), ("f32->f16", x, torch.float16), ("s16->f32", x16, torch.float32), ("s16->f16", x16, torch.float16)):
    out = torch.empty((1024, 128, 3000), dtype=dt, device="cuda")
    for _ in range(3):
        fe.whisper_log_mel_batch(w, n_mels=128, out=out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fe.whisper_log_mel_batch(w, n_mels=128, out=out)
    e1.record()
    torch.cuda.synchronize()
    print(name, round(e0.elapsed_time(e1) / 20, 4), "ms")
