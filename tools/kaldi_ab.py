"""A/B timing of the kaldi fbank kernels on one B200: B200FEAT_KALDI_KERNEL=1 (shared-memory kernel of round 1) against
the default (Tensor-Memory kernel), cfg4 = 4096 x 5 s, fbank + CMN, device-resident, CUDA events.  One process per variant."""
import os
import subprocess
import sys

CHILD = r'''
import sys, torch
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe
B = int(sys.argv[1])
g = torch.Generator(device="cuda").manual_seed(3456)
x = (0.1 * torch.randn(B, 80000, generator=g, device="cuda")).clamp_(-1, 1)
out = torch.empty(B, 498, 80, device="cuda")
for _ in range(3):
    fe.kaldi_fbank_batch(x, out=out)
torch.cuda.synchronize()
fe.profile_begin()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20):
    fe.kaldi_fbank_batch(x, out=out)
b.record(); torch.cuda.synchronize()
pr = fe.profile_end()
ms = a.elapsed_time(b) / 20
fk, ek = pr['feature_ms'] / 20, pr['epilogue_ms'] / 20
print(f"step {ms:.4f} ms  kernel {fk:.4f} ms ({pr['feature_launches'] // 20} launches)  cmn {ek:.4f} ms  checksum {float(out.double().sum()):.4f} "
      f"abs {float(out.double().abs().sum()):.2f}  roofline(step) {B * 479360 / 1e6 / ms / 6542.1:.3f}")
'''
B = sys.argv[1] if len(sys.argv) > 1 else "4096"
for k in ("1", "3"):
    env = dict(os.environ, B200FEAT_KALDI_KERNEL=k)
    r = subprocess.run([sys.executable, "-c", CHILD, B], env=env, capture_output=True, text=True, timeout=600)
    print(f"kernel {k}: {r.stdout.strip()} {r.stderr.strip()[-600:]}")
