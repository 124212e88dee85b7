"""Tensor-core resampler (fe_resample_tc.cuh) against the float64 restatement of torchaudio's resampler and against the
CUDA-core kernel (B200FEAT_RESAMPLE_TC=0, read when a filter's plan is first built: one process per variant), with timing.
Usage: python tools/resample_tc_check.py [B]"""
import os
import subprocess
import sys

CHILD = r'''
import sys, torch, numpy as np
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe
from oracle import resample as orc
B = int(sys.argv[1])
kw = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="sinc_interp_kaiser", beta=14.769656459379492)
g = torch.Generator().manual_seed(3)
x = (0.3 * torch.randn(3, 44100 + 777, generator=g)).clamp(-1, 1)
lens = [44100 + 777, 20000, 441 * 128 + 5]
got = fe.condition_pcm(x.cuda(), 44100, 16000, lengths=lens, **kw).cpu()
err = 0.0
for b, n in enumerate(lens):
    ref = orc.resample_restated(x[b, :n], 44100, 16000, dtype=torch.float64, **kw)
    m = ref.shape[0]
    err = max(err, float((got[b, :m].double() - ref).abs().max()))
    assert not got[b, m:].any(), "zeros behind a ragged clip"
st = torch.randint(-20000, 20000, (2, 50000, 2), generator=g, dtype=torch.int16)
gs = fe.condition_pcm(st.cuda(), 44100, 16000, mix=(0.65, 0.35), **kw).cpu()
rs = orc.condition(st, 44100, 16000, weights=(0.65, 0.35), **kw)
err2 = float((gs.double() - rs).abs().max())
for of, nf in ((48000, 16000), (24000, 16000), (16000, 24000), (32000, 16000)):
    y = (0.3 * torch.randn(2, 30011, generator=g)).clamp(-1, 1)
    got_r = fe.condition_pcm(y.cuda(), of, nf, lengths=[30011, 9999], **kw).cpu()
    e = 0.0
    for b, n in enumerate([30011, 9999]):
        ref = orc.resample_restated(y[b, :n], of, nf, dtype=torch.float64, **kw)
        e = max(e, float((got_r[b, :ref.shape[0]].double() - ref).abs().max()))
        assert not got_r[b, ref.shape[0]:].any()
    print(f"[{of}->{nf}: {e:.2e}]", end=" ")
gen = torch.Generator(device="cuda").manual_seed(1)
b48 = (0.1 * torch.randn(B, 1440000, generator=gen, device="cuda")).clamp_(-1, 1)
o48 = torch.empty(B, 480000, device="cuda")
for _ in range(2):
    fe.condition_pcm(b48, 48000, 16000, out=o48, **kw)
torch.cuda.synchronize()
a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    fe.condition_pcm(b48, 48000, 16000, out=o48, **kw)
b_.record(); torch.cuda.synchronize()
print(f"[48k->16k {B} x 30 s: {a.elapsed_time(b_) / 5:.3f} ms]", end=" ")
del b48, o48
big = (0.1 * torch.randn(B, 1323000, generator=gen, device="cuda")).clamp_(-1, 1)
out = torch.empty(B, 480000, device="cuda")
for _ in range(2):
    fe.condition_pcm(big, 44100, 16000, out=out, **kw)
torch.cuda.synchronize()
a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    fe.condition_pcm(big, 44100, 16000, out=out, **kw)
b_.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b_) / 5
print(f"max |err| vs fp64: mono ragged {err:.3e}, int16 stereo mix {err2:.3e};  {B} x 30 s: {ms:.3f} ms = {B * 30 / ms / 1e3:.3f} M audio-s/s  checksum {float(out.double().sum()):.6f}")
'''
B = sys.argv[1] if len(sys.argv) > 1 else "128"
for tc in ("1", "0"):
    env = dict(os.environ, B200FEAT_RESAMPLE_TC=tc)
    r = subprocess.run([sys.executable, "-c", CHILD, B], env=env, capture_output=True, text=True, timeout=900)
    print(f"TC={tc}: {r.stdout.strip()} {r.stderr.strip()[-1500:]}")
