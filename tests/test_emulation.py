"""CPU suite: the kernels' phase functions (the very source the CUDA kernels call), compiled with g++ and run
thread after thread, against the fp64 oracle.  This checks index maps, codelets, twiddle/window/mel tables
and the ragged / padding logic on the GPU-less authoring box; the `-m gpu` tests check the real kernels."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from tests.emu import build as emu_build
from tests.conftest import speech_like, kaldi_check

vp, i64 = ctypes.c_void_p, ctypes.c_int64


@pytest.fixture(scope="module")
def emu():
    return emu_build.load()


def _lens(lengths):
    if lengths is None:
        return None, None
    la = np.ascontiguousarray(np.asarray(lengths, dtype=np.int64))
    return la, la.ctypes.data_as(vp)


def emu_whisper(emu, x, n_mels=80, padding=0, norm=0, lengths=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    T = (L + padding) // 160
    out = np.full((B, n_mels, T), 7.0, dtype=np.float32)
    keep, lp = _lens(lengths)
    assert emu.emu_whisper(x.ctypes.data_as(vp), lp, B, L, L, n_mels, padding, norm, out.ctypes.data_as(vp), T) == 0
    return out


def emu_whisper_f16(emu, x, n_mels=80, padding=0, norm=0, lengths=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    T = (L + padding) // 160
    out = np.full((B, n_mels, T), 7.0, dtype=np.float16)
    keep, lp = _lens(lengths)
    assert emu.emu_whisper_f16(x.ctypes.data_as(vp), lp, B, L, L, n_mels, padding, norm, out.ctypes.data_as(vp), T) == 0
    return out


def emu_whisper3(emu, x, n_mels=80, padding=0, norm=0, lengths=None, f16=False, misalign=False):
    """the Tensor-Memory kernel's phases (fe_whisper3.cuh); misalign=True shifts the waveform off 16-byte alignment,
    which sends every unit down the element-wise loader path instead of the bulk-copy path"""
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    if misalign:
        buf = np.zeros(B * L + 1, dtype=np.float32)
        buf[1:] = x.reshape(-1)
        x = buf[1:].reshape(B, L)
        assert x.ctypes.data % 16 != 0
    T = (L + padding) // 160
    out = np.full((B, n_mels, T), 7.0, dtype=np.float16 if f16 else np.float32)
    keep, lp = _lens(lengths)
    assert emu.emu_whisper3(x.ctypes.data_as(vp), lp, B, L, L, n_mels, padding, norm, out.ctypes.data_as(vp), int(f16), T) == 0
    return out


def emu_kaldi(emu, x, cmn=1, lengths=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    T = 1 + (L - 400) // 160
    out = np.full((B, T, 80), 7.0, dtype=np.float32)
    keep, lp = _lens(lengths)
    assert emu.emu_kaldi(x.ctypes.data_as(vp), lp, B, L, L, cmn, out.ctypes.data_as(vp), T) == 0
    return out


def emu_cosy(emu, x, lengths=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    T = 1 + (L + 1440 - 1920) // 480
    out = np.full((B, 80, T), 7.0, dtype=np.float32)
    keep, lp = _lens(lengths)
    assert emu.emu_cosy(x.ctypes.data_as(vp), lp, B, L, L, out.ctypes.data_as(vp), T) == 0
    return out


@pytest.mark.parametrize("B,L,nm,pad", [(2, 16000, 80, 0), (1, 12345, 128, 777), (3, 4000, 128, 0), (1, 8000, 80, 48000)])
def test_whisper_noise(emu, B, L, nm, pad):
    x = (0.1 * torch.randn(B, L, generator=torch.Generator().manual_seed(L))).clamp(-1, 1)
    ref = oracle.whisper_log_mel(x, nm, padding=pad, dtype=torch.float64, norm="per_clip").numpy()
    got = emu_whisper(emu, x.numpy(), nm, pad)
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 1e-4


def test_whisper_speech_sine_zeros_percall_ragged(emu):
    gen = torch.Generator().manual_seed(5)
    x = speech_like(gen, 2, 16000, 16000)
    ref = oracle.whisper_log_mel(x, 128, dtype=torch.float64, norm="per_clip").numpy()
    assert np.abs(emu_whisper(emu, x.numpy(), 128) - ref).max() <= 1e-4
    ref = oracle.whisper_log_mel(x, 80, dtype=torch.float64, norm="per_call").numpy()
    assert np.abs(emu_whisper(emu, x.numpy(), 80, norm=1) - ref).max() <= 1e-4
    t = np.arange(48000) / 16000.0
    s = (0.5 * np.sin(2 * np.pi * 440 * t)).astype(np.float32)[None]
    ref = oracle.whisper_log_mel(s, 80, dtype=torch.float64).numpy()
    assert np.abs(emu_whisper(emu, s, 80) - ref).max() <= 1e-4
    z = emu_whisper(emu, np.zeros((1, 16000), np.float32), 80)
    assert np.abs(z + 1.5).max() < 1e-6
    got = emu_whisper(emu, x.numpy(), 80, lengths=[16000, 5003])
    ref = oracle.whisper_log_mel(x[1, :5003], 80, dtype=torch.float64).numpy()
    assert np.abs(got[1, :, :ref.shape[1]] - ref).max() <= 1e-4
    assert not got[1, :, ref.shape[1]:].any()


def test_float_to_half_converter(emu):
    """fe_f2h_rn's host form (the device form is cvt.rn.f16.f32) against torch's float32 -> float16 cast."""
    g = torch.Generator().manual_seed(3)
    v = torch.cat([torch.randn(200000, generator=g) * 2.0, torch.randn(50000, generator=g) * 1e-5,
                   torch.randn(50000, generator=g) * 7e4,
                   torch.tensor([0.0, -0.0, 1.0, -1.5, 65504.0, 65519.9, 65520.0, 1e9, float("inf"), -float("inf"),
                                 2.0 ** -24, 2.0 ** -25, 2.0 ** -25 * 1.0001, 6.1e-5, 6.097e-5, 1.00048828125, 1.0009765625])])
    # every half value and the midpoints between neighbours (ties to even)
    h = torch.arange(0, 0x7c00, dtype=torch.int16).view(torch.float16).float()
    v = torch.cat([v, h, -h, (h[:-1] + h[1:]) / 2]).contiguous()
    out = np.zeros(v.numel(), dtype=np.uint16)
    emu.emu_f2h(v.numpy().ctypes.data_as(vp), out.ctypes.data_as(vp), v.numel())
    ref = v.to(torch.float16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(out, ref)


def test_whisper_f16_output_is_the_rounded_f32_output(emu):
    """float16 store path of phase M + the fix-up in the half domain == float32 result cast to half, bit for bit:
    noise (no clamp), speech-like (clamp active), per-call norm, ragged with zero fill."""
    gen = torch.Generator().manual_seed(11)
    cases = [((0.1 * torch.randn(2, 12000, generator=gen)).clamp(-1, 1), dict(n_mels=128)),
             (speech_like(gen, 2, 16000, 16000), dict(n_mels=80)),
             (speech_like(gen, 2, 16000, 16000), dict(n_mels=128, norm=1)),
             (speech_like(gen, 2, 16000, 16000), dict(n_mels=80, lengths=[16000, 5003], padding=320))]
    for x, kw in cases:
        f32 = emu_whisper(emu, x.numpy(), **kw)
        f16 = emu_whisper_f16(emu, x.numpy(), **kw)
        assert f16.dtype == np.float16 and f16.shape == f32.shape
        assert np.array_equal(f16.view(np.uint16), torch.from_numpy(f32).to(torch.float16).numpy().view(np.uint16))


def test_whisper_int16_pcm_input(emu):
    """int16 PCM in (scaled by 1/32768 in the loader, interior and edge units) == the float32 path on pcm / 32768,
    bit for bit, and within tolerance of the oracle on the same converted samples."""
    gen = torch.Generator().manual_seed(13)
    x = speech_like(gen, 2, 24000, 16000)
    pcm = (x * 32767.0).round().clamp(-32768, 32767).to(torch.int16)
    xf = pcm.to(torch.float32) / 32768.0
    for kw in (dict(n_mels=80), dict(n_mels=128, padding=1600, lengths=[24000, 7001])):
        n_mels, padding, lengths = kw.get("n_mels"), kw.get("padding", 0), kw.get("lengths")
        B, L = pcm.shape
        T = (L + padding) // 160
        out = np.full((B, n_mels, T), 7.0, dtype=np.float32)
        keep, lp = _lens(lengths)
        a = np.ascontiguousarray(pcm.numpy())
        assert emu.emu_whisper_s16(a.ctypes.data_as(vp), lp, B, L, L, n_mels, padding, 0, out.ctypes.data_as(vp), T) == 0
        assert np.array_equal(out, emu_whisper(emu, xf.numpy(), n_mels, padding, 0, lengths))
    ref = oracle.whisper_log_mel(xf, 80, dtype=torch.float64, norm="per_clip").numpy()
    assert np.abs(emu_whisper(emu, xf.numpy(), 80) - ref).max() <= 1e-4


@pytest.mark.parametrize("B,L", [(2, 16000), (1, 80000), (3, 999), (1, 400)])
def test_kaldi(emu, B, L):
    x = (0.1 * torch.randn(B, L, generator=torch.Generator().manual_seed(L))).clamp(-1, 1)
    raw = emu_kaldi(emu, x.numpy(), 0)
    assert raw.shape == (B, oracle.kaldi_num_frames(L), 80)
    kaldi_check(raw, x)
    cmn = emu_kaldi(emu, x.numpy(), 1)
    assert np.abs(raw - raw.mean(axis=1, keepdims=True) - cmn).max() < 2e-5


def test_kaldi_speech_and_ragged(emu):
    x = speech_like(torch.Generator().manual_seed(9), 2, 16000, 16000)
    kaldi_check(emu_kaldi(emu, x.numpy(), 0), x)
    raw = emu_kaldi(emu, x.numpy(), 0, lengths=[16000, 5000])
    cmn = emu_kaldi(emu, x.numpy(), 1, lengths=[16000, 5000])
    T1 = oracle.kaldi_num_frames(5000)
    kaldi_check(raw[1:, :T1], x[1:, :5000])
    assert not raw[1, T1:].any() and not cmn[1, T1:].any()
    assert np.abs(cmn[1, :T1] - (raw[1, :T1] - raw[1, :T1].mean(axis=0, keepdims=True))).max() < 2e-5


@pytest.mark.parametrize("B,L", [(2, 24000), (1, 48000), (2, 5003), (1, 1000)])
def test_cosy(emu, B, L):
    y = (0.1 * torch.randn(B, L, generator=torch.Generator().manual_seed(L))).clamp(-1, 1)
    ref = oracle.cosyvoice_mel(y, dtype=torch.float64).numpy()
    got = emu_cosy(emu, y.numpy())
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 1e-4


def test_cosy_speech_zeros_ragged(emu):
    y = speech_like(torch.Generator().manual_seed(4), 2, 24000, 24000)
    ref = oracle.cosyvoice_mel(y, dtype=torch.float64).numpy()
    assert np.abs(emu_cosy(emu, y.numpy()) - ref).max() <= 1e-4
    z = emu_cosy(emu, np.zeros((1, 24000), np.float32))
    assert np.abs(z - np.log(1e-5)).max() < 1e-5
    got = emu_cosy(emu, y.numpy(), lengths=[24000, 7777])
    ref = oracle.cosyvoice_mel(y[1, :7777], dtype=torch.float64).numpy()
    assert np.abs(got[1, :, :ref.shape[1]] - ref).max() <= 1e-4
    assert not got[1, :, ref.shape[1]:].any()


# ------------------------------------------------------------------------------------------------ Wav2Lip / Diff2Lip mel
def emu_wav2lip(emu, x, lengths=None, frame_major=False, pad_mode=0):
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    T = 1 + L // 200
    out = np.full((B, T, 80) if frame_major else (B, 80, T), 7.0, dtype=np.float32)
    keep, lp = _lens(lengths)
    assert emu.emu_wav2lip(x.ctypes.data_as(vp), lp, B, L, L, out.ctypes.data_as(vp), T, int(frame_major), pad_mode) == 0
    return out


def test_wav2lip_mel_against_the_reference_restatement(emu):
    """the kernel's phases (N = 800 = 16 x 5 x 10, pre-emphasis and reflect padding in the loader) against
    oracle.wav2lip (the reference's audio.py:45-51 in its own float64 arithmetic): <= 1e-4 in the normalised [-4, 4]
    scale; noise, speech-like, a sine, silence, ragged clips, zero padding mode, frame-major layout."""
    from oracle import wav2lip as w2l
    gen = torch.Generator().manual_seed(8)
    x = np.concatenate([(0.1 * torch.randn(1, 9000, generator=gen)).clamp(-1, 1).numpy(), speech_like(gen, 2, 9000, 16000).numpy()])
    x[2] *= 0.01                                   # a quiet clip: bins near the 1e-5 floor
    got = emu_wav2lip(emu, x)
    for b in range(3):
        ref = w2l.wav2lip_mel(x[b])
        assert got[b].shape == ref.shape and np.abs(got[b] - ref).max() <= 1e-4, b
    assert np.array_equal(emu_wav2lip(emu, x, frame_major=True), got.transpose(0, 2, 1))
    for b in range(3):
        assert np.abs(emu_wav2lip(emu, x, pad_mode=1)[b] - w2l.wav2lip_mel(x[b], pad_mode="constant")).max() <= 1e-4
    t = np.arange(8000) / 16000.0
    s = (0.5 * np.sin(2 * np.pi * 440 * t)).astype(np.float32)[None]
    assert np.abs(emu_wav2lip(emu, s)[0] - w2l.wav2lip_mel(s[0])).max() <= 1e-4
    z = emu_wav2lip(emu, np.zeros((1, 3000), np.float32))
    assert np.all(z == -4.0)                        # 20 log10(1e-5) - 20 = -120 dB -> clipped at -max_abs_value
    got = emu_wav2lip(emu, x, lengths=[9000, 4567, 401])
    for b, n in enumerate([9000, 4567, 401]):
        ref = w2l.wav2lip_mel(x[b, :n])
        assert np.abs(got[b, :, :ref.shape[1]] - ref).max() <= 1e-4
        assert not got[b, :, ref.shape[1]:].any()


# ------------------------------------------------------------------------------------------------ Tensor-Memory kernel
@pytest.mark.parametrize("B,L,nm,pad", [(2, 48000, 80, 0), (1, 12345, 128, 777), (3, 4000, 128, 0), (1, 8000, 80, 48000),
                                        (1, 160 * 300 + 57, 128, 0)])
def test_whisper3_noise(emu, B, L, nm, pad):
    x = (0.1 * torch.randn(B, L, generator=torch.Generator().manual_seed(L))).clamp(-1, 1)
    ref = oracle.whisper_log_mel(x, nm, padding=pad, dtype=torch.float64, norm="per_clip").numpy()
    got = emu_whisper3(emu, x.numpy(), nm, pad)
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 1e-4
    # the element-wise loader path (misaligned rows) must give the bulk-copy path's result bit for bit
    assert np.array_equal(emu_whisper3(emu, x.numpy(), nm, pad, misalign=True), got)


def test_whisper3_equals_frame_pair_kernel_closely(emu):
    """both formulations compute the same 20 x 20 FFT with the same codelets and tables; only the mel summation
    order differs (two chains per filter in the frame-pair kernel, one here)"""
    x = (0.1 * torch.randn(2, 40000, generator=torch.Generator().manual_seed(1))).clamp(-1, 1).numpy()
    for nm in (80, 128):
        assert np.abs(emu_whisper3(emu, x, nm) - emu_whisper(emu, x, nm)).max() <= 2e-6


def test_whisper3_speech_sine_zeros_percall_ragged(emu):
    gen = torch.Generator().manual_seed(5)
    x = speech_like(gen, 2, 48000, 16000)
    ref = oracle.whisper_log_mel(x, 128, dtype=torch.float64, norm="per_clip").numpy()
    assert np.abs(emu_whisper3(emu, x.numpy(), 128) - ref).max() <= 1e-4
    ref = oracle.whisper_log_mel(x, 80, dtype=torch.float64, norm="per_call").numpy()
    assert np.abs(emu_whisper3(emu, x.numpy(), 80, norm=1) - ref).max() <= 1e-4
    t = np.arange(48000) / 16000.0
    s = (0.5 * np.sin(2 * np.pi * 440 * t)).astype(np.float32)[None]
    ref = oracle.whisper_log_mel(s, 80, dtype=torch.float64).numpy()
    assert np.abs(emu_whisper3(emu, s, 80) - ref).max() <= 1e-4
    z = emu_whisper3(emu, np.zeros((1, 16000), np.float32), 80)
    assert np.abs(z + 1.5).max() < 1e-6
    got = emu_whisper3(emu, x.numpy(), 80, lengths=[48000, 25003])
    ref = oracle.whisper_log_mel(x[1, :25003], 80, dtype=torch.float64).numpy()
    assert np.abs(got[1, :, :ref.shape[1]] - ref).max() <= 1e-4
    assert not got[1, :, ref.shape[1]:].any()


def test_whisper3_units_of_pure_padding_are_filled_not_transformed(emu):
    """whisper.transcribe's call (padding = N_SAMPLES) on ragged clips: units that lie wholly in the zero padding are
    skipped by the kernel and filled with the clip's clamp threshold by the fix-up (fe_w3_silent).  Lengths sit on and
    next to the boundary  128 * k * 160 - 200  where a unit turns silent; one clip is all zeros."""
    gen = torch.Generator().manual_seed(12)
    L, pad = 2 * 128 * 160 + 900, 128 * 160 * 3
    x = speech_like(gen, 5, L, 16000)
    x[4] = 0.0
    lengths = [L, 128 * 160 - 200, 128 * 160 - 201, 2 * 128 * 160 - 199, L]
    for nm, norm in ((80, 0), (128, 1)):
        got = emu_whisper3(emu, x.numpy(), nm, pad, norm=norm, lengths=lengths)
        refs = [oracle.whisper_log_mel(x[b, :n], nm, padding=pad, dtype=torch.float64).numpy() for b, n in enumerate(lengths)]
        for b, r in enumerate(refs):
            T = r.shape[1]
            if norm == 0:
                assert np.abs(got[b, :, :T] - r).max() <= 1e-4, (nm, b)
                # the padded tail is one constant: max(-10, max - 8) in the output scale
                tail = got[b, :, (lengths[b] + 200) // 160 + 2:T]
                assert tail.size and np.all(tail == tail.flat[0])
                assert abs(float(tail.flat[0]) - max(-1.5, float(got[b, :, :T].max()) - 2.0)) <= 1e-6
            assert not got[b, :, T:].any()
        if norm == 1:      # per-call: one maximum, hence one threshold, for all clips
            thr = max(-1.5, float(got.max()) - 2.0)
            for b in range(5):
                T = (lengths[b] + pad) // 160
                tail = got[b, :, (lengths[b] + 200) // 160 + 2:T]
                assert np.all(np.abs(tail - thr) <= 1e-6)
        # float16 output: the rounded float32 output, silent units included
        h = emu_whisper3(emu, x.numpy(), nm, pad, norm=norm, lengths=lengths, f16=True)
        assert np.array_equal(h, got.astype(np.float16))
        # element-wise loader path agrees bit for bit
        assert np.array_equal(emu_whisper3(emu, x.numpy(), nm, pad, norm=norm, lengths=lengths, misalign=True), got)


def test_whisper3_windows_are_slices_of_the_padded_mel(emu):
    """b200feat_whisper_logmel_windows: (B, n_windows, n_mels, W) == the W-frame slices of the padded
    log-mel, bit for bit against the plain layout of the same kernel, and within 1e-4 of the oracle's slices.  Small W so
    that the emulation stays fast: W = 300 (two 128-frame units + one of 44), padding = W * 160 like transcribe's."""
    W, pad = 300, 300 * 160
    gen = torch.Generator().manual_seed(44)
    L = 2 * W * 160 + 160 * 31 + 76           # content 631 frames -> 3 windows; a multiple of 4: 16-byte aligned rows
    x = np.ascontiguousarray(speech_like(gen, 3, L, 16000).numpy())
    lengths = [L, W * 160, 12345]             # W * 160: content exactly one window, signal reaches into the frames behind it
    la = np.asarray(lengths, dtype=np.int64)
    nw = -(-(L // 160) // W)
    assert nw == 3
    for nm, f16 in ((80, False), (128, True)):
        out = np.full((3, nw, nm, W), 7.0, dtype=np.float16 if f16 else np.float32)
        assert emu.emu_whisper3_windows(x.ctypes.data_as(vp), la.ctypes.data_as(vp), 3, L, L, nm, pad, 0, W, nw,
                                        out.ctypes.data_as(vp), int(f16)) == 0
        plain = emu_whisper3(emu, x, nm, pad, lengths=lengths, f16=f16)          # (3, nm, (L + pad) // 160)
        for w in range(nw):
            assert np.array_equal(out[:, w], plain[:, :, w * W:(w + 1) * W]), (nm, w)
        if not f16:
            for b, n in enumerate(lengths):
                ref = oracle.whisper_log_mel(torch.from_numpy(x[b, :n]), nm, padding=pad, dtype=torch.float64).numpy()
                for w in range(nw):
                    r = ref[:, w * W:(w + 1) * W]
                    if r.shape[1]:
                        assert np.abs(out[b, w, :, :r.shape[1]] - r).max() <= 1e-4
                    assert not out[b, w, :, r.shape[1]:].any()


def test_whisper3_f16_output_is_the_rounded_f32_output(emu):
    x = speech_like(torch.Generator().manual_seed(9), 2, 30000, 16000).numpy()
    for nm in (80, 128):
        f32 = emu_whisper3(emu, x, nm, lengths=[30000, 17001])
        f16 = emu_whisper3(emu, x, nm, lengths=[30000, 17001], f16=True)
        assert np.array_equal(f16, f32.astype(np.float16))


# ------------------------------------------------------------------------------------------------ Tensor-Memory kaldi kernel
def emu_kaldi3(emu, x, cmn=1, lengths=None, T_out=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, L = x.shape
    T = 1 + (L - 400) // 160 if T_out is None else T_out
    out = np.full((B, T, 80), 7.0, dtype=np.float32)
    keep, lp = _lens(lengths)
    assert emu.emu_kaldi3(x.ctypes.data_as(vp), lp, B, L, L, cmn, out.ctypes.data_as(vp), T) == 0
    return out


@pytest.mark.parametrize("B,L", [(2, 16000), (1, 80000), (3, 5444), (2, 23456)])
def test_kaldi3(emu, B, L):
    x = (0.1 * torch.randn(B, L, generator=torch.Generator().manual_seed(L))).clamp(-1, 1)
    kaldi_check(emu_kaldi3(emu, x.numpy(), cmn=0), x, f"kaldi3 {B}x{L}")
    got = emu_kaldi3(emu, x.numpy(), cmn=1)
    for b in range(B):
        ref = oracle.kaldi_fbank(x[b], subtract_mean=True, dtype=torch.float64).numpy()
        raw = oracle.kaldi_fbank(x[b], subtract_mean=False, dtype=torch.float64).numpy()
        raw32 = oracle.kaldi_fbank(x[b], subtract_mean=False).numpy()
        assert np.abs(got[b] - ref).max() <= oracle.kaldi_log_tolerance(raw32, raw)
    # the two formulations (same codelets for N = 512 = 32 x 16, other summation orders) differ by at most the sum of
    # their distances to fp64 -- near-empty bins carry a few 1e-4 each (oracle.kaldi_log_tolerance)
    a, c = emu_kaldi3(emu, x.numpy(), cmn=0), emu_kaldi(emu, x.numpy(), cmn=0)
    for b in range(B):
        raw = oracle.kaldi_fbank(x[b], subtract_mean=False, dtype=torch.float64).numpy()
        raw32 = oracle.kaldi_fbank(x[b], subtract_mean=False).numpy()
        assert np.abs(a[b] - c[b]).max() <= 2 * oracle.kaldi_log_tolerance(raw32, raw)


def test_kaldi3_speech_zeros_ragged(emu):
    gen = torch.Generator().manual_seed(21)
    x = speech_like(gen, 2, 40000, 16000)
    kaldi_check(emu_kaldi3(emu, x.numpy(), cmn=0), x, "kaldi3 speech")
    z = emu_kaldi3(emu, np.zeros((1, 8000), np.float32), cmn=0)
    assert np.allclose(z, np.log(1.1920928955078125e-07))
    assert np.abs(emu_kaldi3(emu, np.zeros((1, 8000), np.float32), cmn=1)).max() < 1e-5
    lens = [40000, 12345]
    got = emu_kaldi3(emu, x.numpy(), cmn=0, lengths=lens)
    T1 = 1 + (12345 - 400) // 160
    ref = oracle.kaldi_fbank(x[1, :12345], subtract_mean=False, dtype=torch.float64).numpy()
    assert np.abs(got[1, :T1] - ref).max() <= 5e-4 and not got[1, T1:].any()
    kaldi_check(got[:1], x[:1], "kaldi3 ragged clip 0")
    # T_out larger than any clip's frame count (rows of a longer batch buffer): whole dead units are zero-filled
    got = emu_kaldi3(emu, x.numpy()[:, :6000], cmn=1, T_out=300)
    assert not got[:, 1 + (6000 - 400) // 160:].any()
