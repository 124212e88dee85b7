"""A/B timing of the Whisper kernels on one B200: B200FEAT_WHISPER_KERNEL=2 (shared-memory frame-pair kernel) against
the default (Tensor-Memory kernel), 1024 x 30 s, 128 mel bins, device-resident, CUDA events.  One process per variant
(the choice is read once per process).  Usage: python tools/whisper_ab.py [n_mels] [B]"""
import os
import subprocess
import sys

CHILD = r'''
import sys, torch, numpy as np
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe
n_mels, B = int(sys.argv[1]), int(sys.argv[2])
g = torch.Generator(device="cuda").manual_seed(1234)
x = (0.1 * torch.randn(B, 480000, generator=g, device="cuda")).clamp_(-1, 1)
out = torch.empty(B, n_mels, 3000, device="cuda")
for _ in range(3):
    fe.whisper_log_mel_batch(x, n_mels=n_mels, out=out)
torch.cuda.synchronize()
fe.profile_begin()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20):
    fe.whisper_log_mel_batch(x, n_mels=n_mels, out=out)
b.record(); torch.cuda.synchronize()
pr = fe.profile_end()
ms = a.elapsed_time(b) / 20
print(f"step {ms:.4f} ms  kernel {pr['feature_ms'] / pr['feature_launches']:.4f} ms  fixup {pr['epilogue_ms'] / pr['epilogue_launches']:.4f} ms  "
      f"checksum {float(out.double().sum()):.6f}  roofline {B * 3456000 * n_mels / 128 / 1e6 / (pr['feature_ms'] / pr['feature_launches']) / 6542.1 if n_mels == 128 else 0:.3f}")
'''
n_mels = sys.argv[1] if len(sys.argv) > 1 else "128"
B = sys.argv[2] if len(sys.argv) > 2 else "1024"
if os.environ.get("AB_LIBS"):       # A/B of two builds of the library instead: AB_LIBS=path_a,path_b -> a b a b on the same box
    for lib in os.environ["AB_LIBS"].split(",") * 2:
        env = dict(os.environ, B200FEAT_LIB=os.path.abspath(lib))
        r = subprocess.run([sys.executable, "-c", CHILD, n_mels, B], env=env, capture_output=True, text=True, timeout=600)
        print(f"{os.path.basename(lib)}: {r.stdout.strip()} {r.stderr.strip()[-400:]}")
    sys.exit(0)
for k in ("2", "3"):
    env = dict(os.environ, B200FEAT_WHISPER_KERNEL=k)
    r = subprocess.run([sys.executable, "-c", CHILD, n_mels, B], env=env, capture_output=True, text=True, timeout=600)
    print(f"kernel {k}: {r.stdout.strip()} {r.stderr.strip()[-400:]}")
