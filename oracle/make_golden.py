#!/usr/bin/env python
"""Generates tests/golden/*.npz in the AUTHORING container (needs /root/reference, transformers,
torchaudio).  The GPU box has neither /root/reference nor a reason to re-run this: the fixtures are
committed.  Run from the repo root:  python -m oracle.make_golden

What pins what
  harvard_w80.npz   input  : /root/reference/Backend/harvard.wav (44.1 kHz stereo s16, sha256 971b4163...)
                             -> mean of channels -> torchaudio.functional.resample(44100 -> 16000), the
                             downmix + resample the reference applies before Whisper
                             (/root/reference/Backend/app.py:36-45, Backend/speech_translation_pipeline.py:133)
                    expect : transformers.WhisperFeatureExtractor._torch_extract_fbank_features (an
                             independent implementation of whisper.audio.log_mel_spectrogram) and the fp64 oracle
  whisper_rand.npz  seeded noise / speech-like clips, 80 and 128 bins: HF torch path + HF numpy path + fp64 oracle
  kaldi_rand.npz    torchaudio.compliance.kaldi.fbank(num_mel_bins=80, dither=0, sample_frequency=16000)
                    itself (+ CMN) and the fp64 oracle
  cosy_rand.npz     fp64 oracle, its numpy twin, and TWO independent implementations of the same mel fed the
                    reflect-padded signal: torchaudio.transforms.MelSpectrogram(24000, 1920, 1920, 480, 0, 8000,
                    n_mels=80, power=1, center=False, norm="slaney", mel_scale="slaney") and
                    transformers.audio_utils.spectrogram(power=1, mel_filters=mel_filter_bank(961, 80, 0, 8000,
                    24000, "slaney", "slaney")); parameters = /root/reference/greek_sft.yaml:55-63.  They differ from
                    matcha only by its `+ 1e-9` under the sqrt (bounded in tests/test_oracle.py)
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from oracle import numpy_ref  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def speech_like(gen, B, L, sr):
    """sum of 3 random sines x slow AM envelope + small noise: wide dynamic range (exercises max-8)."""
    t = torch.arange(L, dtype=torch.float64) / sr
    out = torch.zeros(B, L, dtype=torch.float64)
    for b in range(B):
        f = 80.0 + 3000.0 * torch.rand(3, generator=gen, dtype=torch.float64)
        ph = 6.283185307179586 * torch.rand(3, generator=gen, dtype=torch.float64)
        env = 0.5 * (1.0 + torch.sin(6.283185307179586 * (0.5 + 2.5 * torch.rand(1, generator=gen, dtype=torch.float64)) * t))
        sig = sum(torch.sin(6.283185307179586 * f[i] * t + ph[i]) for i in range(3)) / 3.0
        out[b] = 0.8 * env * sig
    out += 0.01 * torch.randn(B, L, generator=gen, dtype=torch.float64)
    return out.clamp(-1, 1).to(torch.float32)


def main():
    os.makedirs(GOLD, exist_ok=True)
    from transformers import WhisperFeatureExtractor
    import torchaudio
    from scipy.io import wavfile

    # ---------------------------------------------------------------- the kaldi filter table (see oracle/filters.py)
    tb, _ = torchaudio.compliance.kaldi.get_mel_banks(80, 512, 16000.0, 20.0, 0.0, 100.0, -500.0, 1.0)
    table = np.zeros((80, 257), dtype=np.float32)
    table[:, :256] = tb.numpy()
    np.save(os.path.join(ROOT, "oracle", "kaldi_banks_80x257_f32.npy"), table)
    baked = np.load(os.path.join(GOLD, "baked_filters.npz"))["kaldi80"]
    assert np.array_equal(table, baked), "kernel-baked kaldi table != torchaudio's: re-run tools/gen_kernels.py"

    # ---------------------------------------------------------------- config 1: harvard.wav
    path = "/root/reference/Backend/harvard.wav"
    sha = hashlib.sha256(open(path, "rb").read()).hexdigest()
    sr, pcm = wavfile.read(path)
    assert sr == 44100 and pcm.shape == (809508, 2) and pcm.dtype == np.int16, (sr, pcm.shape, pcm.dtype)
    mono = torch.from_numpy(pcm.astype(np.float32) / 32768.0).mean(dim=1)
    wav16 = torchaudio.functional.resample(mono, 44100, 16000).contiguous()
    assert wav16.shape == (293700,), wav16.shape
    fe80 = WhisperFeatureExtractor(feature_size=80)
    hf = fe80._torch_extract_fbank_features(wav16.numpy())
    o32 = oracle.whisper_log_mel(wav16, 80).numpy()
    o64 = oracle.whisper_log_mel(wav16, 80, dtype=torch.float64).numpy()
    assert hf.shape == (80, 1835)
    assert np.array_equal(hf, o32), "oracle fp32 path must be bit-identical to HF's torch path"
    print("harvard: |hf - fp64| max", np.abs(hf - o64).max())
    np.savez_compressed(os.path.join(GOLD, "harvard_w80.npz"), wave16k=wav16.numpy(), hf_torch=hf,
                        oracle_fp64=o64.astype(np.float32), source_sha256=np.array(sha),
                        versions=np.array(f"torch {torch.__version__} torchaudio {torchaudio.__version__}"))

    # ---------------------------------------------------------------- whisper random / speech-like
    gen = torch.Generator().manual_seed(1234)
    noise = (0.1 * torch.randn(3, 24000, generator=gen)).clamp(-1, 1)
    speech = speech_like(gen, 3, 24000, 16000)
    d = {}
    for name, x in (("noise", noise), ("speech", speech)):
        d[f"{name}_wave"] = x.numpy()
        for nm in (80, 128):
            fe = WhisperFeatureExtractor(feature_size=nm)
            hf_t = fe._torch_extract_fbank_features(x.numpy())  # per-item max (HF semantics)
            hf_n = fe._np_extract_fbank_features(x.numpy().astype(np.float64), "cpu")
            o64 = oracle.whisper_log_mel(x, nm, dtype=torch.float64, norm="per_clip").numpy()
            o32 = oracle.whisper_log_mel(x, nm, norm="per_clip").numpy()
            assert np.array_equal(hf_t, o32)
            assert np.abs(hf_n - o64).max() < 2e-5, np.abs(hf_n - o64).max()
            d[f"{name}_hf_torch_{nm}"] = hf_t
            d[f"{name}_fp64_{nm}"] = o64.astype(np.float32)
            print(f"whisper {name} {nm}: |hf_torch - fp64| {np.abs(hf_t - o64).max():.2e}  |hf_numpy - fp64| {np.abs(hf_n - o64).max():.2e}")
    np.savez_compressed(os.path.join(GOLD, "whisper_rand.npz"), **d)

    # ---------------------------------------------------------------- kaldi fbank
    gen = torch.Generator().manual_seed(3456)
    xs = {"noise": (0.1 * torch.randn(2, 16000, generator=gen)).clamp(-1, 1), "speech": speech_like(gen, 2, 16000, 16000)}
    d = {}
    for name, x in xs.items():
        d[f"{name}_wave"] = x.numpy()
        ta = torch.stack([torchaudio.compliance.kaldi.fbank(x[b:b + 1], num_mel_bins=80, dither=0, sample_frequency=16000)
                          for b in range(x.shape[0])])
        ta_cmn = ta - ta.mean(dim=1, keepdim=True)
        o32 = torch.stack([oracle.kaldi_fbank(x[b], subtract_mean=False) for b in range(x.shape[0])])
        o64 = torch.stack([oracle.kaldi_fbank(x[b], subtract_mean=False, dtype=torch.float64) for b in range(x.shape[0])])
        o64c = torch.stack([oracle.kaldi_fbank(x[b], subtract_mean=True, dtype=torch.float64) for b in range(x.shape[0])])
        print(f"kaldi {name}: |torchaudio - oracle32| {(ta - o32).abs().max():.2e}  |torchaudio - fp64| {(ta.double() - o64).abs().max():.2e}")
        assert (ta - o32).abs().max() < 1e-4
        d[f"{name}_torchaudio"] = ta.numpy()
        d[f"{name}_torchaudio_cmn"] = ta_cmn.numpy()
        d[f"{name}_fp64"] = o64.numpy().astype(np.float32)
        d[f"{name}_fp64_cmn"] = o64c.numpy().astype(np.float32)
    np.savez_compressed(os.path.join(GOLD, "kaldi_rand.npz"), **d)

    # ---------------------------------------------------------------- cosyvoice mel
    gen = torch.Generator().manual_seed(2345)
    xs = {"noise": (0.1 * torch.randn(2, 24000, generator=gen)).clamp(-1, 1), "speech": speech_like(gen, 2, 24000, 24000)}
    d = {}
    for name, x in xs.items():
        d[f"{name}_wave"] = x.numpy()
        o64 = oracle.cosyvoice_mel(x, dtype=torch.float64).numpy()
        o32 = oracle.cosyvoice_mel(x).numpy()
        twin = np.stack([numpy_ref.cosyvoice_mel_np(x[b].numpy()) for b in range(x.shape[0])])
        assert np.abs(twin - o64).max() < 1e-9
        print(f"cosy {name}: |fp32 - fp64| {np.abs(o32 - o64).max():.2e}")
        d[f"{name}_fp32"] = o32
        d[f"{name}_fp64"] = o64.astype(np.float32)
        # independent implementations (linear mel energies, float64 in, stored float32)
        from transformers import audio_utils as au
        xp = torch.nn.functional.pad(x.double().unsqueeze(1), (720, 720), mode="reflect").squeeze(1)
        ms = torchaudio.transforms.MelSpectrogram(sample_rate=24000, n_fft=1920, win_length=1920, hop_length=480,
                                                  f_min=0.0, f_max=8000.0, n_mels=80, power=1.0, center=False,
                                                  norm="slaney", mel_scale="slaney").to(torch.float64)
        mel_ta = ms(xp).numpy()
        mf = au.mel_filter_bank(961, 80, 0.0, 8000.0, 24000, norm="slaney", mel_scale="slaney")
        win = au.window_function(1920, "hann", periodic=True)
        mel_hf = np.stack([au.spectrogram(xp[b].numpy(), win, frame_length=1920, hop_length=480, fft_length=1920,
                                          power=1.0, center=False, mel_filters=mf, mel_floor=1e-30)
                           for b in range(x.shape[0])])
        M = np.exp(o64)
        strong = M >= 1e-2
        print(f"cosy {name}: torchaudio rel {(np.abs(mel_ta - M)[strong] / M[strong]).max():.2e}  "
              f"transformers rel {(np.abs(mel_hf - M)[strong] / M[strong]).max():.2e}")
        assert (np.abs(mel_ta - M)[strong] / M[strong]).max() <= 1e-4   # torchaudio builds its filter bank in float32
        assert (np.abs(mel_hf - M)[strong] / M[strong]).max() <= 5e-6
        d[f"{name}_torchaudio_mel"] = mel_ta.astype(np.float32)
        d[f"{name}_hf_mel"] = mel_hf.astype(np.float32)
    np.savez_compressed(os.path.join(GOLD, "cosy_rand.npz"), **d)
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
