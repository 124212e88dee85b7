"""Waveform conditioning in front of the featurizers: int16 / float PCM -> mono mix ->
torchaudio-compatible band-limited resampling (b200feat_resample).

CPU part: the oracle (oracle/resample.py) is pinned against the REAL torchaudio code (functional.resample and
transforms.Resample, both importable here) and against the committed golden made from the reference's own fixture
Backend/harvard.wav; the kernel's phase functions + the host-side plan (tap table) are run through the g++ emulation
against the float64 oracle; the C ABI's argument checks are exercised without a device.
GPU part (-m gpu): the CUDA kernel through the Python mirror, same cases, plus the chain PCM -> 16 kHz -> log-mel and
the full cfg2-sized batch.

Tolerance: waveforms live in [-1, 1]; max abs error <= 2e-6 against the float64 oracle (hann, 35 taps) and <= 4e-6 for
the reference's 128-zero-crossing kaiser filter (743 taps, float32 accumulation).  The reference's own float32 path is
2.5e-6 .. 1.8e-5 away from float64 on the same inputs (printed by oracle/make_golden_resample.py).
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import resample as orc
from tests.emu import build as emu_build

vp = ctypes.c_void_p
TOL_HANN, TOL_KAISER = 2e-6, 4e-6
KAISER = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="sinc_interp_kaiser", beta=14.769656459379492)
RATIOS = [(44100, 16000, 50000), (48000, 16000, 30011), (16000, 24000, 20000), (22050, 16000, 12345),
          (8000, 16000, 999), (16000, 48000, 100), (44100, 48000, 9000), (16000, 8000, 5), (32000, 16000, 4097)]


def rnd(B, L, seed, C=None):
    g = torch.Generator().manual_seed(seed)
    shape = (B, L) if C is None else (B, L, C)
    return (0.3 * torch.randn(*shape, generator=g)).clamp(-1, 1)


def rnd_pcm16(B, L, C, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(-32768, 32768, (B, L, C), generator=g, dtype=torch.int32).to(torch.int16)


# ------------------------------------------------------------------------------------------ oracle pinned
@pytest.mark.parametrize("of,nf,L", RATIOS[:5])
def test_oracle_matches_real_torchaudio(of, nf, L):
    x = rnd(2, L, 11)
    ref_fn = orc.reference_resample(x, of, nf)                       # taps built in float32
    ref_tr = orc.reference_resample_transform(x, of, nf)             # taps built in float64, stored float32
    mine32 = orc.resample_restated(x, of, nf, dtype=torch.float32)
    mine64 = orc.resample_restated(x, of, nf, dtype=torch.float64)
    assert mine32.shape == ref_fn.shape == ref_tr.shape
    assert (mine32 - ref_tr).abs().max() <= 1e-6                     # same taps, same conv
    assert (mine64 - ref_tr.double()).abs().max() <= 1e-6
    assert (mine64 - ref_fn.double()).abs().max() <= 3e-5            # functional's float32 taps are its own error


def test_oracle_kaiser_matches_real_torchaudio():
    x = rnd(1, 30000, 12)
    ref = orc.reference_resample_transform(x, 44100, 16000, **KAISER)
    mine = orc.resample_restated(x, 44100, 16000, dtype=torch.float64, **KAISER)
    assert (mine - ref.double()).abs().max() <= 1e-5


def test_oracle_against_harvard_golden(golden):
    g = golden("harvard_pcm_2s.npz")
    pcm = torch.from_numpy(g["pcm"])[None]
    assert np.abs(orc.condition(pcm, 44100, 16000)[0].numpy() - g["ta_hann_f32"]).max() <= 5e-6
    assert np.abs(orc.condition(pcm, 44100, 16000, **KAISER)[0].numpy() - g["ta_kaiser_f32"]).max() <= 5e-6
    whole = golden("harvard_w80.npz")["wave16k"]                     # the full clip, made by the same real call
    assert np.array_equal(whole[:31900], g["ta_hann_f32"][:31900])


# ------------------------------------------------------------------------------------------ emulation
@pytest.fixture(scope="module")
def emu():
    return emu_build.load()


def emu_condition(emu, x, of, nf, mix=None, lengths=None, lpfw=6, rolloff=0.99, method=0, beta=14.769656459379492,
                  misalign=False):
    x = np.ascontiguousarray(x)
    B, L = x.shape[:2]
    C = 1 if x.ndim == 2 else x.shape[2]
    if misalign:                                                    # rows that are not aligned to the frame size
        buf = np.zeros(x.size + 1, dtype=x.dtype)
        buf[1:] = x.ravel()
        src = buf[1:]
    else:
        src = x.ravel()
    g = int(np.gcd(of, nf))
    Lo = -(-(nf // g) * L // (of // g))
    out = np.full((B, Lo), 7.0, np.float32)
    info = np.zeros(8, np.int32)
    mixa = None if mix is None else np.asarray(mix, np.float32)
    la = None if lengths is None else np.asarray(lengths, np.int64)
    rc = emu.emu_resample(src.ctypes.data_as(vp), int(x.dtype == np.int16), C,
                          None if mixa is None else mixa.ctypes.data_as(vp),
                          None if la is None else la.ctypes.data_as(vp), B, L, of, nf, lpfw, rolloff, method, beta,
                          out.ctypes.data_as(vp), Lo, info.ctypes.data_as(vp))
    assert rc == 0
    return out, info


@pytest.mark.parametrize("of,nf,L", RATIOS)
def test_emulation_ratios(emu, of, nf, L):
    x = rnd(2, L, 21)
    out, info = emu_condition(emu, x.numpy(), of, nf)
    ref = orc.resample_restated(x, of, nf).numpy()
    assert out.shape == ref.shape
    assert np.abs(out - ref).max() <= TOL_HANN


def test_emulation_kaiser_reference_filter(emu):
    x = rnd(1, 20000, 22)
    out, info = emu_condition(emu, x.numpy(), 44100, 16000, lpfw=128, rolloff=0.95, method=1)
    assert info[3] == 372 and info[7] == 0                          # width; table too large for shared memory
    ref = orc.resample_restated(x, 44100, 16000, **KAISER).numpy()
    assert np.abs(out - ref).max() <= TOL_KAISER


@pytest.mark.parametrize("misalign", [False, True])
@pytest.mark.parametrize("mix", [None, (0.65, 0.35)])
def test_emulation_pcm16_stereo(emu, mix, misalign):
    pcm = rnd_pcm16(2, 30000, 2, 23)
    out, _ = emu_condition(emu, pcm.numpy(), 44100, 16000, mix=mix, misalign=misalign)
    ref = orc.condition(pcm, 44100, 16000, weights=mix).numpy()
    assert np.abs(out - ref).max() <= TOL_HANN
    xf = (pcm.to(torch.float32) / 32768.0).numpy()
    outf, _ = emu_condition(emu, xf, 44100, 16000, mix=mix, misalign=misalign)
    if mix is None:
        assert np.array_equal(out, outf)                            # 0.5 / 32768 is a power of two: same bits
    else:
        assert np.abs(out - outf).max() <= 2e-7


def test_emulation_ragged_and_mono16(emu):
    pcm = rnd_pcm16(3, 20000, 1, 24)
    lens = [20000, 7777, 0]
    out, _ = emu_condition(emu, pcm.numpy(), 48000, 16000, lengths=lens)
    for b, n in enumerate(lens):
        m = -(-n // 3)
        if n:
            ref = orc.condition(pcm[b:b + 1, :n], 48000, 16000)[0].numpy()
            assert np.abs(out[b, :m] - ref).max() <= TOL_HANN
        assert not out[b, m:].any()


def test_emulation_harvard_golden(emu, golden):
    g = golden("harvard_pcm_2s.npz")
    out, _ = emu_condition(emu, g["pcm"][None], 44100, 16000)
    assert np.abs(out[0] - g["fp64_hann"]).max() <= TOL_HANN
    assert np.abs(out[0] - g["ta_hann_f32"]).max() <= 5e-6          # the real torchaudio output
    out, _ = emu_condition(emu, g["pcm"][None], 44100, 16000, lpfw=128, rolloff=0.95, method=1)
    assert np.abs(out[0] - g["fp64_kaiser"]).max() <= TOL_KAISER
    assert np.abs(out[0] - g["ta_kaiser_f32"]).max() <= 6e-6


def test_emulation_closed_forms(emu):
    z, _ = emu_condition(emu, np.zeros((1, 5000), np.float32), 44100, 16000)
    assert not z.any()
    one, _ = emu_condition(emu, np.ones((1, 20000), np.float32), 44100, 16000)
    assert np.abs(one[0, 100:-100] - 1.0).max() <= 2e-3             # DC gain of the windowed sinc, away from the edges
    t = np.arange(44100, dtype=np.float64) / 44100.0
    s, _ = emu_condition(emu, np.sin(2 * np.pi * 440.0 * t)[None].astype(np.float32), 44100, 16000)
    want = np.sin(2 * np.pi * 440.0 * np.arange(16000) / 16000.0)
    assert np.abs(s[0, 100:-100] - want[100:-100]).max() <= 2e-3    # a 440 Hz tone stays a 440 Hz tone


# ------------------------------------------------------------------------------------------ C ABI without a device
def test_abi_resample_arguments():
    from expressive_speech_translation_b200 import _lib
    lib = _lib.lib
    assert lib.b200feat_resample_out_len(809508, 44100, 16000) == 293700     # harvard.wav
    assert lib.b200feat_resample_out_len(480000, 16000, 24000) == 720000
    assert lib.b200feat_resample_out_len(5, 16000, 8000) == 3
    assert lib.b200feat_resample_out_len(0, 44100, 16000) == 0

    def call(dtype=0, ch=1, B=1, L=1000, stride=None, of=44100, nf=16000, lpfw=6, rolloff=0.99, method=0, Lo=None,
             ostride=None, pcm=8, out=8):
        full = lib.b200feat_resample_out_len(L, of, nf)
        Lo = full if Lo is None else Lo
        return lib.b200feat_resample(pcm, dtype, ch, None, None, B, L, L * ch if stride is None else stride, of, nf,
                                     lpfw, rolloff, method, 0.0, out, Lo if ostride is None else ostride, Lo, None)

    assert call(dtype=1) == -1 and b"pcm_dtype" in lib.b200feat_last_error()
    assert call(ch=3) == -1 and b"channels" in lib.b200feat_last_error()
    assert call(of=0) == -1 and b"positive" in lib.b200feat_last_error()
    assert call(B=-1) == -1
    assert call(stride=10) == -1 and b"pcm_stride" in lib.b200feat_last_error()
    assert call(Lo=10 ** 6) == -1 and b"L_out" in lib.b200feat_last_error()
    assert call(ostride=1) == -1 and b"out_stride" in lib.b200feat_last_error()
    assert call(pcm=None) == -1 and b"null" in lib.b200feat_last_error()
    assert call(B=0) == 0 and call(L=0) == 0                                  # nothing to do: no device needed


# ------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def fe():
    assert torch.cuda.is_available(), "the -m gpu suite needs a CUDA device"
    import expressive_speech_translation_b200 as m
    return m


@pytest.mark.gpu
@pytest.mark.parametrize("of,nf,L", RATIOS)
def test_gpu_ratios(fe, of, nf, L):
    x = rnd(3, L, 31)
    out = fe.resample(x.cuda(), of, nf)
    ref = orc.resample_restated(x, of, nf)
    assert out.is_cuda and out.shape == ref.shape and out.dtype == torch.float32
    assert (out.cpu().double() - ref).abs().max() <= TOL_HANN
    host = fe.resample(x, of, nf)                                   # CPU tensor in -> CPU tensor out, like torchaudio
    assert not host.is_cuda and torch.equal(host, out.cpu())
    assert fe.resample(x, of, of) is x


@pytest.mark.gpu
def test_gpu_leading_dims_and_transform_class(fe):
    x = rnd(6, 9000, 32).reshape(2, 3, 9000)
    out = fe.resample(x.cuda(), 44100, 16000)
    assert tuple(out.shape) == (2, 3, 3266)
    ref = orc.resample_restated(x, 44100, 16000)
    assert (out.cpu().double() - ref).abs().max() <= TOL_HANN
    y = fe.Resample(orig_freq=44100, new_freq=16000, **KAISER)(x[0].cuda())       # audio_processor.py:296-304
    refk = orc.resample_restated(x[0], 44100, 16000, **KAISER)
    assert (y.cpu().double() - refk).abs().max() <= TOL_KAISER
    with pytest.raises(TypeError):
        fe.resample(torch.zeros(4, 100, dtype=torch.int32), 44100, 16000)
    with pytest.raises(ValueError):
        fe.resample(x, 44100, 16000, resampling_method="nearest")
    with pytest.raises(ValueError):
        fe.resample(x, 0, 16000)


@pytest.mark.gpu
@pytest.mark.parametrize("mix", [None, (0.65, 0.35)])
def test_gpu_pcm16_stereo(fe, mix):
    pcm = rnd_pcm16(4, 66150, 2, 33)
    lens = [66150, 44100, 12345, 0]
    out = fe.condition_pcm(pcm.cuda(), 44100, 16000, lengths=lens, mix=mix).cpu()
    for b, n in enumerate(lens):
        m = fe.resample_out_len(n, 44100, 16000)
        if n:
            ref = orc.condition(pcm[b:b + 1, :n], 44100, 16000, weights=mix)[0]
            assert (out[b, :m].double() - ref).abs().max() <= TOL_HANN
        assert not out[b, m:].any()
    xf = pcm.to(torch.float32) / 32768.0
    outf = fe.condition_pcm(xf.cuda(), 44100, 16000, lengths=lens, mix=mix).cpu()
    if mix is None:
        assert torch.equal(out, outf)
    else:
        assert (out - outf).abs().max() <= 2e-7
    odd = fe.condition_pcm(pcm.cuda().reshape(-1)[1:-1].reshape(1, -1, 2)[:, :1000], 44100, 16000)   # contiguous copy
    assert odd.shape == (1, 363)


@pytest.mark.gpu
def test_gpu_harvard_golden_and_chain(fe, golden):
    import oracle
    g = golden("harvard_pcm_2s.npz")
    pcm = torch.from_numpy(g["pcm"])[None]
    w = fe.condition_pcm(pcm.cuda(), 44100, 16000)
    assert np.abs(w[0].cpu().numpy() - g["fp64_hann"]).max() <= TOL_HANN
    assert np.abs(w[0].cpu().numpy() - g["ta_hann_f32"]).max() <= 5e-6
    k = fe.condition_pcm(pcm.cuda(), 44100, 16000, **KAISER)
    assert np.abs(k[0].cpu().numpy() - g["fp64_kaiser"]).max() <= TOL_KAISER
    # int16 stereo 44.1 kHz in -> log-mel out, device-resident all the way (the real cfg-1 pipeline)
    mel = fe.whisper_log_mel_batch(w, n_mels=80)
    ref = oracle.whisper_log_mel(orc.condition(pcm, 44100, 16000)[0], 80, dtype=torch.float64)
    assert (mel[0].cpu().double() - ref).abs().max() <= 1e-4
    # CosyVoice's 16 -> 24 kHz step in front of its mel
    w24 = fe.resample(w, 16000, 24000)
    ref24 = orc.resample_restated(orc.condition(pcm, 44100, 16000), 16000, 24000)
    assert (w24.cpu().double() - ref24).abs().max() <= 2 * TOL_HANN


@pytest.mark.gpu
def test_gpu_full_size_batch(fe):
    """cfg2's batch as it really arrives: 1024 x 30 s of 44.1 kHz stereo int16 (5.4 GB) -> (1024, 480000) float32.
    Tile independence (rows of the big launch == the same clips run alone, bit for bit), spot clips against the
    float64 oracle, and linearity on the float path."""
    B, L = 1024, 1323000
    g = torch.Generator(device="cuda").manual_seed(34)
    pcm = torch.randint(-20000, 20000, (B, L, 2), generator=g, dtype=torch.int16, device="cuda")
    out = fe.condition_pcm(pcm, 44100, 16000)
    assert tuple(out.shape) == (B, 480000)
    for b in (0, 517, 1023):
        alone = fe.condition_pcm(pcm[b:b + 1], 44100, 16000)
        assert torch.equal(alone[0], out[b])
        ref = orc.condition(pcm[b:b + 1, :132300].cpu(), 44100, 16000)[0]
        assert (out[b, :47000].cpu().double() - ref[:47000]).abs().max() <= TOL_HANN
    assert torch.isfinite(out).all() and float(out.abs().max()) < 1.0
    del pcm, out
    a = rnd(2, 200000, 35, C=2).cuda()
    b2 = rnd(2, 200000, 36, C=2).cuda()
    lhs = fe.condition_pcm(a + b2, 44100, 16000)
    rhs = fe.condition_pcm(a, 44100, 16000) + fe.condition_pcm(b2, 44100, 16000)
    assert (lhs - rhs).abs().max() <= 1e-6


@pytest.mark.gpu
def test_gpu_install_patches_resample(fe):
    import torchaudio
    orig = torchaudio.functional.resample
    done = fe.install()
    try:
        assert "torchaudio.functional.resample" in done
        x = rnd(1, 4410, 37)
        y = torchaudio.functional.resample(x, 44100, 16000)
        assert tuple(y.shape) == (1, 1600)
    finally:
        fe.uninstall()
    assert torchaudio.functional.resample is orig


@pytest.mark.gpu
def test_gpu_reference_kaiser_filter_on_tensor_cores_full_size():
    """The filter the reference deploys (audio_processor.py:296-303: kaiser, width 128, rolloff 0.95; 44.1 -> 16 kHz = 160
    phases x 1185 taps) runs as a 3 x TF32 GEMM on the tensor cores (fe_resample_tc.cuh).  At BASELINE configs[1]'s size
    (1024 x 30 s): two clips against the float64 restatement within 4e-6, determinism, batch independence, ragged ends,
    int16 stereo input through the mono pre-pass; and the launch really is the tensor-core kernel."""
    import expressive_speech_translation_b200 as fe
    kw = dict(fe.REFERENCE_RESAMPLER)
    g = torch.Generator(device="cuda").manual_seed(11)
    B, L = 1024, 1323000
    x = torch.empty(B, L, device="cuda")
    for lo in range(0, B, 128):
        x[lo:lo + 128].normal_(generator=g).mul_(0.25).clamp_(-1, 1)
    n0 = fe.launch_count()
    out = fe.condition_pcm(x, 44100, 16000, **kw)
    assert fe.launch_count() - n0 == 1                       # float32 mono, aligned rows: the GEMM kernel alone
    assert out.shape == (B, 480000) and bool(torch.isfinite(out).all())
    assert torch.equal(out, fe.condition_pcm(x, 44100, 16000, **kw))
    assert torch.equal(out[517:518], fe.condition_pcm(x[517:518], 44100, 16000, **kw))
    for b in (0, 1023):
        ref = orc.resample_restated(x[b].cpu(), 44100, 16000, dtype=torch.float64, **kw)
        assert (out[b].cpu().double() - ref).abs().max() <= 4e-6, b
    del out
    lens = torch.full((B,), L, dtype=torch.int64)
    lens[3], lens[900] = 441 * 128 * 7 + 13, 5000
    rag = fe.condition_pcm(x, 44100, 16000, lengths=lens, **kw)
    for b in (3, 900):
        n = int(lens[b])
        ref = orc.resample_restated(x[b, :n].cpu(), 44100, 16000, dtype=torch.float64, **kw)
        m = ref.shape[0]
        assert (rag[b, :m].cpu().double() - ref).abs().max() <= 4e-6
        assert not rag[b, m:].any()
    del rag, x
    pcm = torch.randint(-20000, 20000, (64, 441 * 300 + 7, 2), generator=torch.Generator().manual_seed(2), dtype=torch.int16)
    got = fe.condition_pcm(pcm.cuda(), 44100, 16000, mix=(0.65, 0.35), **kw).cpu()
    ref = orc.condition(pcm[:2], 44100, 16000, weights=(0.65, 0.35), **kw)
    assert (got[:2].double() - ref).abs().max() <= 4e-6
    # rows of a multiple of 4 frames: the kernel reads the int16 stereo frames itself (no mono pre-pass: ONE launch), ragged too
    pcm4 = pcm[:, :441 * 300 + 4].contiguous()
    n0 = fe.launch_count()
    got4 = fe.condition_pcm(pcm4.cuda(), 44100, 16000, lengths=[pcm4.shape[1]] * 63 + [7001], **kw).cpu()
    assert fe.launch_count() - n0 == 1
    ref = orc.condition(pcm4[:1], 44100, 16000, **kw)
    assert (got4[:1].double() - ref).abs().max() <= 4e-6
    ref = orc.condition(pcm4[63:, :7001], 44100, 16000, **kw)
    assert (got4[63:, :ref.shape[1]].double() - ref).abs().max() <= 4e-6 and not got4[63, ref.shape[1]:].any()


@pytest.mark.gpu
@pytest.mark.parametrize("of,nf", [(48000, 16000), (24000, 16000), (16000, 24000), (32000, 16000), (44100, 16000)])
def test_gpu_reference_kaiser_filter_other_ratios(of, nf):
    """The tensor-core path regroups m blocks so that new * m outputs form the GEMM's N (48 -> 16 kHz: 480 -> 160): every
    ratio whose regrouped phase count is a multiple of 16 takes it.  Ragged clips, float64 restatement, 4e-6."""
    import expressive_speech_translation_b200 as fe
    kw = dict(fe.REFERENCE_RESAMPLER)
    g = torch.Generator().manual_seed(of + nf)
    y = (0.3 * torch.randn(3, 30011, generator=g)).clamp(-1, 1)
    lens = [30011, 9999, 1]
    got = fe.condition_pcm(y.cuda(), of, nf, lengths=lens, **kw).cpu()
    for b, n in enumerate(lens):
        ref = orc.resample_restated(y[b, :n], of, nf, dtype=torch.float64, **kw)
        assert (got[b, :ref.shape[0]].double() - ref).abs().max() <= 4e-6, (b, n)
        assert not got[b, ref.shape[0]:].any()
    st = torch.randint(-20000, 20000, (2, 8000, 2), generator=g, dtype=torch.int16)
    got = fe.condition_pcm(st.cuda(), of, nf, **kw).cpu()
    assert (got.double() - orc.condition(st, of, nf, **kw)).abs().max() <= 4e-6
