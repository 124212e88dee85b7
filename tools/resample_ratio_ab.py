"""torchaudio's default filter (hann, width 6) on the tensor-core path against the CUDA-core kernel (B200FEAT_RESAMPLE_TC=0), per ratio.
One process per variant (the choice is made when a filter's plan is first built)."""
import os
import subprocess
import sys

CHILD = r'''
import sys, torch
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe
gen = torch.Generator(device="cuda").manual_seed(1)
res = []
for of, nf, B, secs in ((44100, 16000, 256, 30), (48000, 16000, 256, 30), (16000, 24000, 256, 10), (32000, 16000, 256, 30), (22050, 16000, 256, 30)):
    x = (0.1 * torch.randn(B, of * secs, generator=gen, device="cuda")).clamp_(-1, 1)
    out = torch.empty(B, fe.resample_out_len(of * secs, of, nf), device="cuda")
    for _ in range(3):
        fe.condition_pcm(x, of, nf, out=out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        fe.condition_pcm(x, of, nf, out=out)
    b.record(); torch.cuda.synchronize()
    res.append(f"{of}->{nf} {B}x{secs}s {a.elapsed_time(b) / 10:.3f} ms")
    del x, out
print("  ".join(res))
'''
for tc in ("3", "0"):
    r = subprocess.run([sys.executable, "-c", CHILD], env=dict(os.environ, B200FEAT_RESAMPLE_TC=tc), capture_output=True, text=True, timeout=900)
    print(f"TC={tc}: {r.stdout.strip()} {r.stderr.strip()[-300:]}")
