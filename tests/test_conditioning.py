"""SURVEY.md section 8(f) item 1, the remainder: the reference's stereo -> mono rule on the device, and PCM in ->
log-mel out as one library call (device and host-buffer entry points)."""
import numpy as np
import pytest
import torch

import oracle
from oracle import resample as orc


def _stereo_batch():
    """clips that take each branch of audio_processor.py:313: correlated channels, independent channels, anti-phase,
    one silent channel (correlation NaN -> mid/side branch)"""
    g = torch.Generator().manual_seed(17)
    L = 22050
    a = 0.3 * torch.randn(L, generator=g)
    b = 0.3 * torch.randn(L, generator=g)
    clips = [torch.stack([a, 0.9 * a + 0.1 * b]),            # corr ~ 0.99 -> mean
             torch.stack([a, b]),                              # corr ~ 0    -> mid + 0.3 side
             torch.stack([a, -a]),                             # corr = -1   -> mid + 0.3 side
             torch.stack([a, 0.6 * a + 0.8 * b]),              # corr ~ 0.6  -> mean
             torch.stack([a, torch.zeros(L)])]                 # 0 / 0       -> NaN > 0.5 is False -> mid/side
    return torch.stack(clips).clamp(-1, 1)                      # (B, 2, L)


def test_oracle_stereo_rule_branches():
    x = _stereo_batch()
    want = [True, False, False, True, False]
    for b in range(x.shape[0]):
        mono, corr, mean = orc.reference_stereo_to_mono(x[b])
        assert mean is want[b], (b, corr)
        if mean:
            assert torch.equal(mono[0], (x[b, 0] + x[b, 1]) / 2)
        else:
            # mid + 0.3 side = 0.65 L + 0.35 R up to float32 rounding
            assert (mono[0] - (0.65 * x[b, 0] + 0.35 * x[b, 1])).abs().max() < 1e-6


@pytest.mark.gpu
def test_gpu_stereo_auto_matches_the_reference_code():
    """b200feat_condition_stereo_auto against the reference's own torch code (torchaudio.transforms.Resample with its
    kaiser filter + audio_processor.py:306-319), float32 and int16 PCM: same branch per clip, correlation within 1e-5,
    samples within 4e-6 (the resampler's tolerance against the real torchaudio, tests/test_resample.py)."""
    import expressive_speech_translation_b200 as fe
    x = _stereo_batch()                                          # (B, 2, L) at 44.1 kHz
    pcm = x.transpose(1, 2).contiguous()                         # (B, L, 2) interleaved
    out, (dec, cor) = fe.condition_stereo_auto(pcm.cuda(), 44100, 16000, return_decisions=True)
    assert out.shape == (5, fe.resample_out_len(x.shape[2], 44100, 16000))
    for b in range(5):
        mono, corr, mean = orc.reference_condition_stereo(x[b], 44100)
        assert bool(dec[b].item()) is mean, b
        if not np.isnan(corr):
            assert abs(float(cor[b]) - corr) < 1e-5
        assert (out[b].cpu() - mono[0]).abs().max() <= 4e-6, b
    # the mix itself is the reference's float32 arithmetic on the resampled channels: bit-exact given r0, r1
    r0 = fe.condition_pcm(pcm.cuda(), 44100, 16000, mix=(1.0, 0.0), **fe.REFERENCE_RESAMPLER)
    r1 = fe.condition_pcm(pcm.cuda(), 44100, 16000, mix=(0.0, 1.0), **fe.REFERENCE_RESAMPLER)
    for b in range(5):
        mono, _, _ = orc.reference_stereo_to_mono(torch.stack([r0[b], r1[b]]).cpu())
        assert torch.equal(out[b].cpu(), mono[0]), b
    # int16 PCM and ragged lengths
    p16 = (pcm * 32767).round().to(torch.int16)
    lens = [22050, 9000, 22050, 15001, 22050]
    o16, (d16, _) = fe.condition_stereo_auto(p16.cuda(), 44100, 16000, lengths=lens, return_decisions=True)
    for b, n in enumerate(lens):
        xb = (p16[b, :n].to(torch.float32) / 32768.0).t().contiguous()
        mono, corr, mean = orc.reference_condition_stereo(xb, 44100)
        m = mono.shape[1]
        assert bool(d16[b].item()) is mean
        assert (o16[b, :m].cpu() - mono[0]).abs().max() <= 4e-6
        assert not o16[b, m:].any()
    with pytest.raises(ValueError):
        fe.condition_stereo_auto(pcm[:, :, :1].cuda(), 44100, 16000)


@pytest.mark.gpu
def test_gpu_pcm_to_log_mel_in_one_call_device_and_host():
    """b200feat_pcm_whisper_logmel == b200feat_resample followed by b200feat_whisper_logmel_ex, bit for bit, for int16 stereo
    and float32 mono PCM, ragged, padded, float16 out, with the reference's kaiser filter; the host-buffer entry point gives
    the same bits; and the result is within the featurizer's tolerance of the float64 oracle chain."""
    import expressive_speech_translation_b200 as fe
    g = torch.Generator().manual_seed(23)
    B, L = 6, 44100 * 2            # 32000 samples at 16 kHz: rows of a multiple of 4 floats, so that the two-call chain below
                                   # runs the same (Tensor-Memory) log-mel kernel as the one-call path, whose waveform rows
                                   # are always 16-byte aligned
    pcm = torch.randint(-20000, 20000, (B, L, 2), generator=g, dtype=torch.int16)
    lens = [L, 30000, L - 1, 44100, 5000, L]
    for kw in (dict(), dict(lengths=lens, padding=4800, n_mels=128), dict(out_dtype=torch.float16, norm="per_call"),
               dict(mix=(0.65, 0.35), lengths=lens, **fe.REFERENCE_RESAMPLER)):
        rs = {k: kw[k] for k in ("lowpass_filter_width", "rolloff", "resampling_method", "beta") if k in kw}
        wav = fe.condition_pcm(pcm.cuda(), 44100, 16000, lengths=kw.get("lengths"), mix=kw.get("mix"), **rs)
        wl = None if "lengths" not in kw else [fe.resample_out_len(n, 44100, 16000) for n in lens]
        want = fe.whisper_log_mel_batch(wav, lengths=wl, n_mels=kw.get("n_mels", 80), padding=kw.get("padding", 0),
                                        norm=kw.get("norm", "per_clip"), out_dtype=kw.get("out_dtype", torch.float32))
        got = fe.pcm_whisper_log_mel_batch(pcm.cuda(), 44100, **kw)
        assert got.dtype == want.dtype and torch.equal(got, want), kw
        host = fe.pcm_whisper_log_mel_host(pcm.pin_memory(), 44100, **kw)
        assert torch.equal(host, want.cpu()), kw
        host = fe.pcm_whisper_log_mel_host(pcm.numpy(), 44100, pin_output=False, **kw)        # pageable numpy in
        assert torch.equal(host, want.cpu()), kw
    # an output length that is not a multiple of 4: the two-call chain then takes the frame-pair kernel (unaligned rows),
    # the one-call path the Tensor-Memory kernel on its aligned workspace rows -- equal within the two kernels' 2e-6
    odd = torch.randint(-20000, 20000, (2, 44100 + 321, 2), generator=g, dtype=torch.int16)
    two = fe.whisper_log_mel_batch(fe.condition_pcm(odd.cuda(), 44100, 16000), n_mels=80)
    one = fe.pcm_whisper_log_mel_batch(odd.cuda(), 44100, n_mels=80)
    assert one.shape == two.shape and (one - two).abs().max() <= 2e-6
    assert torch.equal(fe.pcm_whisper_log_mel_host(odd, 44100, n_mels=80), one.cpu())
    mono = (0.1 * torch.randn(3, 48000, generator=g)).clamp(-1, 1)
    got = fe.pcm_whisper_log_mel_batch(mono.cuda(), 48000, n_mels=80)
    ref = oracle.whisper_log_mel(orc.condition(mono[..., None], 48000, 16000), 80, dtype=torch.float64)
    assert (got.cpu().double() - ref).abs().max() <= 1e-4
    # a large chunked host call: more clips than one arena slot holds at B200FEAT_HOST_CHUNK_MB (64 MB)
    big = torch.randint(-20000, 20000, (96, 44100 * 10, 2), generator=g, dtype=torch.int16)
    h = fe.pcm_whisper_log_mel_host(big.pin_memory(), 44100, n_mels=80)
    d = fe.pcm_whisper_log_mel_batch(big.cuda(), 44100, n_mels=80)
    assert torch.equal(h, d.cpu())


@pytest.mark.gpu
def test_gpu_resample_host_twin_equals_device_path():
    """b200feat_resample_host: pinned and pageable PCM, both filters (CUDA-core and tensor-core kernels), ragged, chunked"""
    import expressive_speech_translation_b200 as fe
    g = torch.Generator().manual_seed(4)
    pcm = torch.randint(-20000, 20000, (40, 44100 * 6, 2), generator=g, dtype=torch.int16)      # 42 MB of PCM + 15 MB out: two chunks
    lens = [pcm.shape[1]] * 38 + [50000, 3]
    for kw in (dict(), dict(fe.REFERENCE_RESAMPLER)):
        dev = fe.condition_pcm(pcm.cuda(), 44100, 16000, lengths=lens, mix=(0.65, 0.35), **kw).cpu()
        assert torch.equal(fe.condition_pcm_host(pcm.pin_memory(), 44100, 16000, lengths=lens, mix=(0.65, 0.35), **kw), dev)
        assert torch.equal(fe.condition_pcm_host(pcm.numpy(), 44100, 16000, lengths=lens, mix=(0.65, 0.35), pin_output=False, **kw), dev)
    mono = (0.2 * torch.randn(3, 24000, generator=g)).clamp(-1, 1)
    assert torch.equal(fe.condition_pcm_host(mono, 16000, 24000), fe.condition_pcm(mono.cuda(), 16000, 24000).cpu())
