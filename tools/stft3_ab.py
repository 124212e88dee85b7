"""A/B timing of the three-stage kernels between two builds of the library (B200FEAT_LIB), one process per build, a b a b.
Usage: AB_LIBS=a.so,b.so python tools/stft3_ab.py"""
import os
import subprocess
import sys

CHILD = r'''
import sys, torch
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe
g = torch.Generator(device="cuda").manual_seed(1)
x24 = (0.1 * torch.randn(256, 240000, generator=g, device="cuda")).clamp_(-1, 1)
x16 = (0.1 * torch.randn(256, 160000, generator=g, device="cuda")).clamp_(-1, 1)
res = []
for name, fn, x in (("cosy", fe.cosy_mel_batch, x24), ("w2l", fe.wav2lip_mel_batch, x16)):
    for _ in range(3):
        o = fn(x)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        o = fn(x)
    b.record(); torch.cuda.synchronize()
    res.append(f"{name} {a.elapsed_time(b) / 20:.4f} ms checksum {float(o.double().sum()):.4f}")
print("  ".join(res))
'''
for lib in os.environ["AB_LIBS"].split(",") * 2:
    env = dict(os.environ, B200FEAT_LIB=os.path.abspath(lib))
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=600)
    print(f"{os.path.basename(lib)}: {r.stdout.strip()} {r.stderr.strip()[-300:]}")
