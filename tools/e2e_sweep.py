#!/usr/bin/env python
"""End-to-end (pinned host in -> pinned host out) time of the Whisper host API for one chunk size.
    B200FEAT_HOST_CHUNK_MB=32 python tools/e2e_sweep.py [clips]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import expressive_speech_translation_b200 as fe  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
L, M = 480000, 128
xh = torch.empty((B, L), dtype=torch.float32, pin_memory=True).normal_(0, 0.1).clamp_(-1, 1)
oh = torch.empty((B, M, L // 160), dtype=torch.float32, pin_memory=True)
dev = torch.device("cuda", 0)
fe.whisper_log_mel_host(xh, n_mels=M, norm="per_clip", out=oh, device=dev)
torch.cuda.synchronize()
ts = []
for _ in range(4):
    t0 = time.perf_counter()
    fe.whisper_log_mel_host(xh, n_mels=M, norm="per_clip", out=oh, device=dev)
    ts.append(time.perf_counter() - t0)
gb = (xh.numel() + oh.numel()) * 4 / 1e9
print(f"chunk {os.environ.get('B200FEAT_HOST_CHUNK_MB', '16')} MB: best {min(ts)*1e3:.2f} ms  ({gb/min(ts):.1f} GB/s both ways, "
      f"{B*30/min(ts)/1e3:.0f} k audio-s/s)")
