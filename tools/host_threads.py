"""Concurrency of the host-buffer drop-in calls: N threads x K calls of log_mel_spectrogram(np.ndarray) against one thread,
for a single 10-s clip per call (launch-overhead-bound) and for 8 x 30 s per call (pageable-copy-bound).
B200FEAT_LIB may point at another build (A/B against the whole-call arena lock of round 1)."""
import sys
import threading
import time

import numpy as np

sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe

rng = np.random.default_rng(3)
for name, shape, K in (("1 x 10 s", (160000,), 40), ("8 x 30 s", (8, 480000), 10)):
    clips = [np.clip(0.1 * rng.standard_normal(shape), -1, 1).astype(np.float32) for _ in range(4)]

    def loop(i):
        for _ in range(K):
            fe.log_mel_spectrogram(clips[i], n_mels=80)

    loop(0)
    t0 = time.perf_counter(); loop(0); t1 = time.perf_counter() - t0
    for rnd in range(2):      # round 0 warms up: every thread's first call leases (allocates) its own arena and streams
        th = [threading.Thread(target=loop, args=(i,)) for i in range(4)]
        t0 = time.perf_counter()
        [t.start() for t in th]; [t.join() for t in th]
        t4 = time.perf_counter() - t0
    print(f"{name}: one thread {t1 / K * 1e3:.3f} ms per call; four threads {t4 / K * 1e3:.3f} ms per round of 4 calls = {t4 / t1:.2f} x one thread")
