#!/usr/bin/env python
"""BASELINE configs[0] (one real clip, harvard.wav after 44.1 k -> 16 k resampling: 293 700 samples, 80 bins) as a
latency measurement of the drop-in call: what a caller of whisper.audio.log_mel_spectrogram sees per clip."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import expressive_speech_translation_b200 as fe  # noqa: E402
import oracle  # noqa: E402  (CPU leg of a measurement script, like bench.py's cpu_baseline)

g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "harvard_w80.npz"))
x = np.ascontiguousarray(g[[k for k in g.files if g[k].ndim == 1][0]], dtype=np.float32)
xd = torch.from_numpy(x).cuda()


def wall(f, n):
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


dev = wall(lambda: fe.log_mel_spectrogram(xd, n_mels=80), 200)
host = wall(lambda: fe.log_mel_spectrogram(x, n_mels=80, device="cuda").cpu(), 200)
xt = torch.from_numpy(x)
torch.set_num_threads(len(os.sched_getaffinity(0)))
cpu = wall(lambda: oracle.whisper_log_mel(xt, 80), 20)
torch.set_num_threads(1)
cpu1 = wall(lambda: oracle.whisper_log_mel(xt, 80), 20)
print(f"clip {len(x)} samples ({len(x) / 16000:.2f} s): drop-in call, device-resident input {dev:.0f} us; numpy in -> host tensor out {host:.0f} us; "
      f"CPU reference path {cpu:.0f} us ({torch.get_num_threads()} thread: {cpu1:.0f} us, all {len(os.sched_getaffinity(0))} threads: {cpu:.0f} us)")
