#!/usr/bin/env python
"""Device time of the conditioning kernel (b200feat_resample) on the shapes that sit in front of the featurizers,
CUDA events on the launching stream, against the algorithmic bytes (PCM in + float32 waveform out).
    python tools/resample_timing.py [clips]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import expressive_speech_translation_b200 as fe  # noqa: E402

PEAK = 6449.4
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def timeit(f, iters=10):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        f()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    g = torch.Generator(device="cuda").manual_seed(1)
    kaiser = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="sinc_interp_kaiser")
    cases = []
    pcm = torch.randint(-20000, 20000, (n, 1323000, 2), generator=g, dtype=torch.int16, device="cuda")
    cases.append(("s16 stereo 44.1k -> 16k hann (cfg2 in front)", pcm, 44100, 16000, {}, n * 30.0))
    cases.append(("s16 stereo 44.1k -> 16k kaiser-128 (audio_processor.py)", pcm[: max(1, n // 8)], 44100, 16000, kaiser, max(1, n // 8) * 30.0))
    for name, x, of, nf, kw, secs in cases:
        out = fe.condition_pcm(x, of, nf, **kw)
        ms = timeit(lambda: fe.condition_pcm(x, of, nf, out=out, **kw))
        nbytes = x.numel() * x.element_size() + out.numel() * 4
        print(f"{name}: {ms:.3f} ms  {nbytes / ms / 1e6:.0f} GB/s = {nbytes / ms / 1e6 / PEAK:.3f} of {PEAK} GB/s  "
              f"{secs / ms * 1e3 / 1e6:.2f} M audio-s/s", flush=True)
    del pcm, out
    x48 = torch.randint(-20000, 20000, (n, 1440000, 2), generator=g, dtype=torch.int16, device="cuda")
    out = fe.condition_pcm(x48, 48000, 16000)
    ms = timeit(lambda: fe.condition_pcm(x48, 48000, 16000, out=out))
    nbytes = x48.numel() * 2 + out.numel() * 4
    print(f"s16 stereo 48k -> 16k hann: {ms:.3f} ms  {nbytes / ms / 1e6:.0f} GB/s = {nbytes / ms / 1e6 / PEAK:.3f}", flush=True)
    del x48, out
    w = (0.1 * torch.randn(256, 160000, generator=g, device="cuda")).clamp_(-1, 1)
    out = fe.condition_pcm(w, 16000, 24000)
    ms = timeit(lambda: fe.condition_pcm(w, 16000, 24000, out=out), 50)
    nbytes = w.numel() * 4 + out.numel() * 4
    print(f"f32 mono 16k -> 24k hann (cfg3 in front, 256 x 10 s): {ms:.4f} ms  {nbytes / ms / 1e6:.0f} GB/s = {nbytes / ms / 1e6 / PEAK:.3f}", flush=True)
    # chain: PCM -> 16 kHz -> Whisper 128-bin log-mel, everything device-resident
    pcm = torch.randint(-20000, 20000, (n, 1323000, 2), generator=g, dtype=torch.int16, device="cuda")
    wav = fe.condition_pcm(pcm, 44100, 16000)
    mel = fe.whisper_log_mel_batch(wav, n_mels=128)

    def chain():
        fe.condition_pcm(pcm, 44100, 16000, out=wav)
        fe.whisper_log_mel_batch(wav, n_mels=128, out=mel)
    ms = timeit(chain)
    print(f"chain s16 stereo 44.1k -> log-mel 128 ({n} x 30 s): {ms:.3f} ms  {n * 30.0 / ms * 1e3 / 1e6:.2f} M audio-s/s", flush=True)


if __name__ == "__main__":
    main()
