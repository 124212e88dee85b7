"""CPU oracle of the reference's spectral-gate denoiser (TEST INFRASTRUCTURE ONLY -- nothing under
expressive_speech_translation_b200/ may import this).

/root/reference/Backend/services/audio_processor.py:321-358, the reference's own torch lines in its order, for one mono clip.
torch IS importable here, so this is the reference's arithmetic itself (float32), not a restatement; `dtype=torch.float64`
runs the same lines in float64 as ground truth.

Note on :339-341: `freq_bins = mag.shape[1]` is taken from a (513, T) tensor (the clip is squeezed to 1-D before torch.stft),
so it is the number of FRAMES, and `gain_mask[:, int(freq_bins * 0.0625):int(freq_bins * 0.25)] *= 1.2` boosts a range of
frames for every bin.  The oracle keeps the lines as they are; so does the kernel.
"""
import torch


def reference_spectral_gate(audio, dtype=torch.float32):
    """audio: (1, L) or (L,) -> (1, 256 * (L // 256)) ; audio_processor.py:321-358"""
    audio = torch.as_tensor(audio, dtype=dtype)
    if audio.dim() == 1:
        audio = audio[None]
    window = torch.hann_window(1024, dtype=dtype)
    spec = torch.stft(                                                               # :322-329
        audio.squeeze(), n_fft=1024, hop_length=256, win_length=1024, window=window, return_complex=True)
    mag = torch.abs(spec)                                                            # :331
    phase = torch.angle(spec)                                                        # :332
    noise_floor = torch.quantile(mag, 0.1, dim=1, keepdim=True)                      # :334
    gain_mask = torch.clamp((mag - 1.5 * noise_floor) / (mag + 1e-8), 0, 1)          # :335
    freq_bins = mag.shape[1]                                                         # :338
    speech_range = (int(freq_bins * 0.0625), int(freq_bins * 0.25))                  # :339
    gain_mask[:, speech_range[0]:speech_range[1]] *= 1.2                             # :340
    enhanced_spec = mag * gain_mask * torch.exp(1j * phase)                          # :342
    audio = torch.istft(enhanced_spec, n_fft=1024, hop_length=256, win_length=1024, window=window).unsqueeze(0)   # :343-349
    max_amp = audio.abs().max()                                                      # :352
    if max_amp > 0:
        audio = audio / max_amp * 0.95                                               # :354
    return audio
