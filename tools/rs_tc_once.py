"""two launches of the tensor-core resampler (reference's kaiser filter, 128 x 30 s float32 mono at 44.1 kHz) for ncu"""
import sys
import torch
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe
g = torch.Generator(device="cuda").manual_seed(1)
x = (0.1 * torch.randn(128, 1323000, generator=g, device="cuda")).clamp_(-1, 1)
out = torch.empty(128, 480000, device="cuda")
for _ in range(3):
    fe.condition_pcm(x, 44100, 16000, out=out, **fe.REFERENCE_RESAMPLER)
torch.cuda.synchronize()
print(float(out.double().sum()))
