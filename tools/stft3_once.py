"""one launch each of the two three-stage kernels (CosyVoice mel 256 x 10 s @ 24 kHz, Wav2Lip mel 256 x 10 s @ 16 kHz) for ncu"""
import sys
import torch
sys.path.insert(0, ".")
import expressive_speech_translation_b200 as fe
g = torch.Generator(device="cuda").manual_seed(1)
x24 = (0.1 * torch.randn(256, 240000, generator=g, device="cuda")).clamp_(-1, 1)
x16 = (0.1 * torch.randn(256, 160000, generator=g, device="cuda")).clamp_(-1, 1)
for _ in range(2):
    a = fe.cosy_mel_batch(x24)
    b = fe.wav2lip_mel_batch(x16)
torch.cuda.synchronize()
print(float(a.sum()), float(b.sum()))
