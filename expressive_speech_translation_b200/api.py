"""Host-side mirror of the reference's featurizer call signatures, on top of libb200feat.so.

The three drop-in functions keep the name, argument order, defaults and error behaviour of the calls
the reference makes:

  log_mel_spectrogram(audio, n_mels=80, padding=0, device=None)
      = whisper.audio.log_mel_spectrogram, called inside transcribe() at
        /root/reference/Backend/services/cascaded_backend.py:158,
        /root/reference/Backend/speech_translation_pipeline.py:140,
        /root/reference/Backend/gtts_pipeline_fixed.py:64
  mel_spectrogram(y, n_fft, num_mels, sampling_rate, hop_size, win_size, fmin, fmax, center=False)
      = matcha.utils.audio.mel_spectrogram as bound at /root/reference/greek_sft.yaml:55-63
  fbank(waveform, ...kaldi kwargs...)
      = torchaudio.compliance.kaldi.fbank as the CosyVoice frontend calls it
        (num_mel_bins=80, dither=0, sample_frequency=16000)

plus the batched entry points the manifest runner uses (whisper_log_mel_batch, cosy_mel_batch,
kaldi_fbank_batch) and their host-buffer (end-to-end) twins (*_host).  PyTorch is used for device
memory and streams only; all arithmetic happens in the CUDA library.  There is no CPU fallback.
"""
import ctypes
import wave as _wave

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, KIND_WHISPER, KIND_COSY, KIND_KALDI, NORM_PER_CLIP, NORM_PER_CALL

WHISPER_SR, WHISPER_N_FFT, WHISPER_HOP = 16000, 400, 160
COSY_PARAMS = dict(n_fft=1920, num_mels=80, sampling_rate=24000, hop_size=480, win_size=1920, fmin=0, fmax=8000)
_NORMS = {"per_clip": NORM_PER_CLIP, "per_call": NORM_PER_CALL}
_DTYPES = {torch.float32: _lib.DT_F32, torch.float16: _lib.DT_F16, torch.int16: _lib.DT_S16}


def num_frames(kind, n_samples, padding=0):
    k = {"whisper": KIND_WHISPER, "cosy": KIND_COSY, "kaldi": KIND_KALDI}[kind]
    return int(lib.b200feat_num_frames(k, int(n_samples), int(padding)))


def launch_count():
    """kernel launches issued by the library in this process (bench.py's gpu_launches)."""
    return int(lib.b200feat_launch_count())


def profile_begin():
    """bracket every subsequent kernel launch with CUDA events on its launching stream (roofline leg)."""
    check(lib.b200feat_profile_begin(), "b200feat_profile_begin")


def profile_end():
    """-> dict(feature_ms, feature_launches, epilogue_ms, epilogue_launches) summed since profile_begin."""
    fm, em = ctypes.c_double(0.0), ctypes.c_double(0.0)
    fn, en = ctypes.c_int64(0), ctypes.c_int64(0)
    check(lib.b200feat_profile_end(ctypes.byref(fm), ctypes.byref(fn), ctypes.byref(em), ctypes.byref(en)),
          "b200feat_profile_end")
    return dict(feature_ms=fm.value, feature_launches=fn.value, epilogue_ms=em.value, epilogue_launches=en.value)


# ----------------------------------------------------------------------------------------- helpers
def _cuda_device(device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("expressive_speech_translation_b200 needs a CUDA device (sm_100a); there is no CPU path")
    if device is None:
        return torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    if device.type != "cuda":
        raise ValueError(f"device must be a CUDA device, got {device}")
    if device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    return device


def _as_float_tensor(x, what, allow_pcm16=False):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    if not torch.is_tensor(x):
        raise TypeError(f"{what} must be a torch.Tensor or numpy.ndarray, got {type(x).__name__}")
    if allow_pcm16 and x.dtype == torch.int16:
        return x
    if not x.is_floating_point():
        raise TypeError(f"{what} must be floating point (the reference's torch.stft rejects {x.dtype})")
    return x


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _lengths_dev(lengths, B, device):
    if lengths is None:
        return None
    if isinstance(lengths, (list, tuple, np.ndarray)):
        lengths = torch.as_tensor(np.asarray(lengths, dtype=np.int64))
    lengths = lengths.to(device=device, dtype=torch.int64).contiguous()
    if lengths.numel() != B:
        raise ValueError(f"lengths has {lengths.numel()} entries for a batch of {B}")
    return lengths


def _prep_wave(wave, device):
    """(B, L) float32 (int16 PCM stays int16) on `device` with a usable row stride; returns (tensor, stride)."""
    w = wave.to(device=device, dtype=torch.int16 if wave.dtype == torch.int16 else torch.float32)
    if w.dim() != 2:
        raise ValueError(f"expected a (B, L) waveform batch, got shape {tuple(w.shape)}")
    if w.stride(1) != 1 or (w.shape[0] > 1 and w.stride(0) < w.shape[1]):
        w = w.contiguous()
    stride = w.stride(0) if w.shape[0] > 1 else max(w.shape[1], 1)
    return w, stride


def _workspace(kind, B, L, device):
    n = int(lib.b200feat_workspace_bytes(kind, B, L))
    return torch.empty(n, dtype=torch.uint8, device=device), n


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


# ----------------------------------------------------------------------------------------- batched device API
def _out_dtype(out_dtype):
    if out_dtype not in (torch.float32, torch.float16):
        raise ValueError(f"out_dtype must be torch.float32 or torch.float16, got {out_dtype}")
    return out_dtype


def whisper_log_mel_batch(wave, lengths=None, n_mels=80, padding=0, norm="per_clip", out=None, out_dtype=torch.float32):
    """(B, L) waveform batch on a CUDA device -> (B, n_mels, (L + padding) // 160) log-mel, async on the
    current stream.  `lengths` (B,) marks ragged clips (frames past a clip's own count are 0).
    out_dtype=torch.float16 writes the encoder-ready half tensor directly: bit-identical to `.to(torch.float16)` of
    the float32 result (the cast whisper.transcribe(fp16=True) applies, whisper/transcribe.py `mel_segment.to(dtype)`).
    An int16 `wave` is taken as PCM and scaled by 1/32768 inside the loader (whisper.audio.load_audio's conversion)."""
    assert n_mels in {80, 128}, f"Unsupported n_mels: {n_mels}"
    out_dtype = _out_dtype(out.dtype if out is not None else out_dtype)
    wave = _as_float_tensor(wave, "wave", allow_pcm16=True)
    device = _cuda_device(wave.device if wave.is_cuda else None)
    w, stride = _prep_wave(wave, device)
    B, L = w.shape
    T = (L + padding) // WHISPER_HOP
    if out is None:
        out = torch.empty((B, n_mels, T), dtype=out_dtype, device=device)
    elif out.shape != (B, n_mels, T) or not out.is_contiguous() or out.device != device:
        raise ValueError("out must be a contiguous float32 / float16 (B, n_mels, T) tensor on the wave's device")
    if B == 0 or T == 0:
        return out
    ln = _lengths_dev(lengths, B, device)
    ws, nws = _workspace(KIND_WHISPER, B, L + int(padding), device)
    with torch.cuda.device(device):
        if w.dtype == torch.float32 and out_dtype == torch.float32:
            rc = lib.b200feat_whisper_logmel(_ptr(w), _ptr(ln), B, L, stride, n_mels, int(padding), _NORMS[norm],
                                             _ptr(out), n_mels * T, T, _ptr(ws), nws, _stream(device))
        else:
            rc = lib.b200feat_whisper_logmel_ex(_ptr(w), _DTYPES[w.dtype], _ptr(ln), B, L, stride, n_mels, int(padding),
                                                _NORMS[norm], _ptr(out), _DTYPES[out_dtype], n_mels * T, T, _ptr(ws), nws,
                                                _stream(device))
    check(rc, "b200feat_whisper_logmel")
    return out


N_SAMPLES = 480000      # whisper.audio.N_SAMPLES: 30 s at 16 kHz, the padding whisper.transcribe asks for
N_FRAMES = 3000         # whisper.audio.N_FRAMES: frames of one encoder window


def whisper_num_windows(n_samples, win_frames=N_FRAMES):
    """windows whisper.transcribe decodes for a clip of n_samples: seek = 0, win_frames, ... < content_frames = n_samples // 160"""
    return max(1, -(-(int(n_samples) // WHISPER_HOP) // int(win_frames)))


def whisper_log_mel_windows(wave, lengths=None, n_mels=80, padding=N_SAMPLES, norm="per_clip", win_frames=N_FRAMES,
                            out=None, out_dtype=torch.float16):
    """(B, L) float32 waveform batch on a CUDA device -> (B, n_windows, n_mels, win_frames) encoder-ready windows, the
    `pad_or_trim(mel[:, seek:seek + N_FRAMES], N_FRAMES).to(dtype)` slices whisper.transcribe feeds the encoder
    (UPSTREAM transcribe.py, reached from /root/reference/Backend/services/cascaded_backend.py:158), for seek = 0, 3000,
    ...: written by the kernel's epilogue in that layout and dtype (default float16, transcribe's fp16=True), no
    intermediate (n_mels, T) tensor.  Identical, bit for bit, to slicing whisper_log_mel_batch(wave, padding=padding).
    n_windows = whisper_num_windows(L); a ragged clip's own count is whisper_num_windows(lengths[b])."""
    assert n_mels in {80, 128}, f"Unsupported n_mels: {n_mels}"
    out_dtype = _out_dtype(out.dtype if out is not None else out_dtype)
    wave = _as_float_tensor(wave, "wave")
    device = _cuda_device(wave.device if wave.is_cuda else None)
    w, stride = _prep_wave(wave, device)
    if w.data_ptr() % 16 or stride % 4:             # the Tensor-Memory kernel's bulk copies need 16-byte aligned rows
        Lp = (w.shape[1] + 3) // 4 * 4
        buf = torch.zeros((w.shape[0], Lp), dtype=torch.float32, device=device)
        buf[:, :w.shape[1]] = w
        w, stride = buf[:, :w.shape[1]], Lp
    B, L = w.shape
    nw = whisper_num_windows(L, win_frames)
    if out is None:
        out = torch.empty((B, nw, n_mels, win_frames), dtype=out_dtype, device=device)
    elif out.shape != (B, nw, n_mels, win_frames) or not out.is_contiguous() or out.device != device:
        raise ValueError("out must be a contiguous float32 / float16 (B, n_windows, n_mels, win_frames) tensor on the wave's device")
    if B == 0:
        return out
    ln = _lengths_dev(lengths, B, device)
    ws, nws = _workspace(KIND_WHISPER, B, L + int(padding) + WHISPER_HOP * 128 * (nw + 1), device)
    with torch.cuda.device(device):
        rc = lib.b200feat_whisper_logmel_windows(_ptr(w), _ptr(ln), B, L, stride, n_mels, int(padding), _NORMS[norm],
                                                 int(win_frames), nw, _ptr(out), _DTYPES[out_dtype], nw * n_mels * win_frames,
                                                 _ptr(ws), nws, _stream(device))
    check(rc, "b200feat_whisper_logmel_windows")
    return out


def cosy_mel_batch(wave, lengths=None, out=None, frame_major=False):
    """(B, L) 24 kHz batch on a CUDA device -> (B, 80, T) ln-mel (n_fft 1920, hop 480); frame_major=True writes
    (B, T, 80), the layout the CosyVoice frontend's `_extract_speech_feat` produces with `.transpose(0, 1)`."""
    wave = _as_float_tensor(wave, "y")
    device = _cuda_device(wave.device if wave.is_cuda else None)
    w, stride = _prep_wave(wave, device)
    B, L = w.shape
    T = num_frames("cosy", L)
    shape = (B, T, 80) if frame_major else (B, 80, T)
    if out is None:
        out = torch.empty(shape, dtype=torch.float32, device=device)
    elif out.shape != shape or out.dtype != torch.float32 or not out.is_contiguous() or out.device != device:
        raise ValueError(f"out must be a contiguous float32 {shape} tensor on the wave's device")
    if B == 0 or T == 0:
        return out
    ln = _lengths_dev(lengths, B, device)
    ws, nws = _workspace(KIND_COSY, B, L, device)
    with torch.cuda.device(device):
        fn = lib.b200feat_cosy_mel_frames if frame_major else lib.b200feat_cosy_mel
        rc = fn(_ptr(w), _ptr(ln), B, L, stride, _ptr(out), 80 * T, T, _ptr(ws), nws, _stream(device))
    check(rc, "b200feat_cosy_mel")
    return out


PAD_REFLECT, PAD_ZEROS = 0, 1
_PAD_MODES = {"reflect": PAD_REFLECT, "constant": PAD_ZEROS}


def wav2lip_num_frames(n_samples):
    return 1 + int(n_samples) // 200


def wav2lip_mel_batch(wave, lengths=None, out=None, frame_major=False, pad_mode="reflect"):
    """(B, L) 16 kHz batch on a CUDA device -> (B, 80, 1 + L // 200) Wav2Lip / Diff2Lip mel in [-4, 4]
    (/root/reference/Backend/diff2lip/audio/audio.py:45-51 with hparams.py:22-65); frame_major=True writes (B, T, 80), the
    `.T` its callers take (generate.py:53).  pad_mode: what librosa.stft pads with -- "reflect" (librosa 0.9.2, the
    reference's pin) or "constant" (librosa >= 0.10)."""
    wave = _as_float_tensor(wave, "wav")
    device = _cuda_device(wave.device if wave.is_cuda else None)
    w, stride = _prep_wave(wave, device)
    B, L = w.shape
    T = wav2lip_num_frames(L)
    shape = (B, T, 80) if frame_major else (B, 80, T)
    if out is None:
        out = torch.empty(shape, dtype=torch.float32, device=device)
    elif out.shape != shape or out.dtype != torch.float32 or not out.is_contiguous() or out.device != device:
        raise ValueError(f"out must be a contiguous float32 {shape} tensor on the wave's device")
    if B == 0:
        return out
    ln = _lengths_dev(lengths, B, device)
    with torch.cuda.device(device):
        rc = lib.b200feat_wav2lip_mel(_ptr(w), _ptr(ln), B, L, stride, _PAD_MODES[pad_mode], int(bool(frame_major)),
                                      _ptr(out), 80 * T, T, _stream(device))
    check(rc, "b200feat_wav2lip_mel")
    return out


def melspectrogram(wav):
    """Drop-in for /root/reference/Backend/diff2lip/audio/audio.py:45 `melspectrogram(wav)`: 1-D waveform at 16 kHz
    (numpy, as librosa.load returns it, or a torch tensor) -> (80, 1 + len // 200) in [-4, 4].  numpy in -> numpy out.
    float32 (the reference's lfilter promotes to float64; the values agree within 1e-4)."""
    is_np = not torch.is_tensor(wav)
    w = _as_float_tensor(wav, "wav")
    if w.dim() != 1:
        raise ValueError(f"melspectrogram expects a 1-D wav, got shape {tuple(w.shape)}")
    out = wav2lip_mel_batch(w[None])[0]
    if is_np:
        return out.cpu().numpy()
    return out if w.is_cuda else out.cpu()


def spectral_gate_batch(wave, lengths=None, out=None):
    """(B, L) mono float32 batch on a CUDA device -> (B, 256 * (L // 256)) denoised and peak-normalised waveform: the
    STFT spectral gate + iSTFT + `/ max * 0.95` of /root/reference/Backend/services/audio_processor.py:322-358, per clip."""
    wave = _as_float_tensor(wave, "audio")
    device = _cuda_device(wave.device if wave.is_cuda else None)
    w, stride = _prep_wave(wave, device)
    B, L = w.shape
    L_out = int(lib.b200feat_spectral_gate_out_len(L))
    if out is None:
        out = torch.empty((B, L_out), dtype=torch.float32, device=device)
    elif out.shape != (B, L_out) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != device:
        raise ValueError(f"out must be a contiguous float32 ({B}, {L_out}) tensor on the wave's device")
    if B == 0 or L_out == 0:
        return out
    ln = _lengths_dev(lengths, B, device)
    nws = int(lib.b200feat_spectral_gate_workspace_bytes(B, L))
    ws = torch.empty(nws, dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        check(lib.b200feat_spectral_gate(_ptr(w), _ptr(ln), B, L, stride, _ptr(out), L_out, _ptr(ws), nws, _stream(device)),
              "b200feat_spectral_gate")
    return out


def kaldi_fbank_batch(wave, lengths=None, subtract_mean=True, out=None):
    """(B, L) 16 kHz batch on a CUDA device -> (B, T, 80) kaldi fbank, optionally mean-normalised per clip."""
    wave = _as_float_tensor(wave, "waveform")
    device = _cuda_device(wave.device if wave.is_cuda else None)
    w, stride = _prep_wave(wave, device)
    B, L = w.shape
    T = num_frames("kaldi", L)
    if out is None:
        out = torch.empty((B, T, 80), dtype=torch.float32, device=device)
    elif out.shape != (B, T, 80) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != device:
        raise ValueError("out must be a contiguous float32 (B, T, 80) tensor on the wave's device")
    if B == 0 or T == 0:
        return out
    ln = _lengths_dev(lengths, B, device)
    ws, nws = _workspace(KIND_KALDI, B, L, device)
    with torch.cuda.device(device):
        rc = lib.b200feat_kaldi_fbank(_ptr(w), _ptr(ln), B, L, stride, 1 if subtract_mean else 0, _ptr(out),
                                      T * 80, T, _ptr(ws), nws, _stream(device))
    check(rc, "b200feat_kaldi_fbank")
    return out


# ----------------------------------------------------------------------------------------- host-buffer (end-to-end) API
def _host_2d(x, what, allow_pcm16=False):
    x = _as_float_tensor(x, what, allow_pcm16)
    if x.is_cuda:
        raise ValueError(f"{what}: the *_host entry points take host memory")
    if x.dtype != torch.int16:
        x = x.to(torch.float32)
    if x.dim() != 2:
        raise ValueError(f"expected a (B, L) waveform batch, got shape {tuple(x.shape)}")
    return x.contiguous()


def _host_lengths(lengths, B):
    if lengths is None:
        return None
    ln = torch.as_tensor(np.asarray(lengths, dtype=np.int64)).contiguous()
    if ln.numel() != B:
        raise ValueError(f"lengths has {ln.numel()} entries for a batch of {B}")
    return ln


def _host_out(shape, out, pin, dtype=torch.float32):
    if out is not None:
        if tuple(out.shape) != tuple(shape) or out.dtype != dtype or out.is_cuda or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous {dtype} host tensor of shape {tuple(shape)}")
        return out
    return torch.empty(shape, dtype=dtype, pin_memory=pin)


def whisper_log_mel_host(wave, lengths=None, n_mels=80, padding=0, norm="per_clip", out=None, device=None,
                         pin_output=True, out_dtype=torch.float32):
    """Host (ideally pinned) (B, L) waveform -> host (B, n_mels, T) log-mel through the chunked
    H2D / kernel / D2H pipeline inside the library.  Blocks until the result is complete.
    out_dtype=torch.float16: half features (see whisper_log_mel_batch), half the D2H bytes."""
    assert n_mels in {80, 128}, f"Unsupported n_mels: {n_mels}"
    out_dtype = _out_dtype(out.dtype if out is not None else out_dtype)
    w = _host_2d(wave, "wave", allow_pcm16=True)
    B, L = w.shape
    T = (L + padding) // WHISPER_HOP
    dev = _cuda_device(device)
    out = _host_out((B, n_mels, T), out, pin_output, out_dtype)
    ln = _host_lengths(lengths, B)
    if B and T:
        if w.dtype == torch.float32 and out_dtype == torch.float32:
            rc = lib.b200feat_whisper_logmel_host(_ptr(w), _ptr(ln), B, L, n_mels, int(padding), _NORMS[norm], _ptr(out), dev.index)
        else:
            rc = lib.b200feat_whisper_logmel_ex_host(_ptr(w), _DTYPES[w.dtype], _ptr(ln), B, L, n_mels, int(padding),
                                                     _NORMS[norm], _ptr(out), _DTYPES[out_dtype], dev.index)
        check(rc, "b200feat_whisper_logmel_host")
    return out


def cosy_mel_host(wave, lengths=None, out=None, device=None, pin_output=True):
    w = _host_2d(wave, "y")
    B, L = w.shape
    T = num_frames("cosy", L)
    dev = _cuda_device(device)
    out = _host_out((B, 80, T), out, pin_output)
    ln = _host_lengths(lengths, B)
    if B and T:
        check(lib.b200feat_cosy_mel_host(_ptr(w), _ptr(ln), B, L, _ptr(out), dev.index), "b200feat_cosy_mel_host")
    return out


def kaldi_fbank_host(wave, lengths=None, subtract_mean=True, out=None, device=None, pin_output=True):
    w = _host_2d(wave, "waveform")
    B, L = w.shape
    T = num_frames("kaldi", L)
    dev = _cuda_device(device)
    out = _host_out((B, T, 80), out, pin_output)
    ln = _host_lengths(lengths, B)
    if B and T:
        check(lib.b200feat_kaldi_fbank_host(_ptr(w), _ptr(ln), B, L, 1 if subtract_mean else 0, _ptr(out),
                                            dev.index), "b200feat_kaldi_fbank_host")
    return out


# ----------------------------------------------------------------------------------------- drop-in signatures
def _load_wav_16k(path):
    """whisper.load_audio shells out to ffmpeg (absent here); 16 kHz mono PCM16 .wav files are read
    directly, anything else must be decoded by the caller."""
    with _wave.open(path, "rb") as f:
        if f.getframerate() != WHISPER_SR or f.getnchannels() != 1 or f.getsampwidth() != 2:
            raise ValueError(f"{path}: only 16 kHz mono 16-bit PCM .wav can be loaded without ffmpeg; "
                             "decode and resample first (e.g. Backend/app.py:36-57 does this with ffmpeg)")
        pcm = np.frombuffer(f.readframes(f.getnframes()), dtype=np.int16)
    return pcm.astype(np.float32) / 32768.0


def log_mel_spectrogram(audio, n_mels: int = 80, padding: int = 0, device=None):
    """Drop-in for whisper.audio.log_mel_spectrogram.

    audio: path, np.ndarray or Tensor of shape (L,) / (*, L) at 16 kHz.  Returns (n_mels, T) /
    (*, n_mels, T) float32 with T = (L + padding) // 160; the dynamic-range clamp uses the max over the
    WHOLE call, as upstream.  The result lives on `device` if given, else on the input's device (host
    inputs run through the pipelined host path and come back as host tensors)."""
    if isinstance(audio, str):
        audio = _load_wav_16k(audio)
    audio = _as_float_tensor(audio, "audio")
    assert n_mels in {80, 128}, f"Unsupported n_mels: {n_mels}"
    lead = audio.shape[:-1]
    L = audio.shape[-1] if audio.dim() > 0 else 0
    if audio.dim() == 0:
        raise ValueError("audio must have at least one dimension")
    if L + padding <= WHISPER_N_FFT // 2:
        # torch.stft's reflect pad raises for inputs this short; keep the failure loud and typed
        raise RuntimeError(f"Argument #4: Padding size should be less than the corresponding input dimension, "
                           f"but got: padding (200, 200) at dimension 1 of input of length {L + padding}")
    x = audio.reshape(-1, L)
    if device is not None:
        x = x.to(_cuda_device(device))
    if x.is_cuda:
        res = whisper_log_mel_batch(x, n_mels=n_mels, padding=padding, norm="per_call")
    else:
        res = whisper_log_mel_host(x, n_mels=n_mels, padding=padding, norm="per_call", pin_output=False)
    return res.reshape(*lead, n_mels, res.shape[-1])


def mel_spectrogram(y, n_fft, num_mels, sampling_rate, hop_size, win_size, fmin, fmax, center=False):
    """Drop-in for matcha.utils.audio.mel_spectrogram with the parameter set the reference binds
    (greek_sft.yaml:55-63).  y: (B, L) at 24 kHz in [-1, 1] -> (B, 80, T).  Any other parameter set is
    rejected (this library bakes the 1920/480/80/0-8000 Hz tables)."""
    got = dict(n_fft=n_fft, num_mels=num_mels, sampling_rate=sampling_rate, hop_size=hop_size,
               win_size=win_size, fmin=fmin, fmax=fmax)
    for k, v in COSY_PARAMS.items():
        if float(got[k]) != float(v):
            raise NotImplementedError(f"mel_spectrogram: only the CosyVoice parameter set {COSY_PARAMS} is built; got {got}")
    if center:
        raise NotImplementedError("mel_spectrogram: center=True is not used by the reference and not built")
    y = _as_float_tensor(y, "y")
    if y.dim() != 2:
        raise ValueError(f"mel_spectrogram expects y of shape (B, L), got {tuple(y.shape)}")
    if y.shape[1] <= 720:
        raise RuntimeError(f"Argument #4: Padding size should be less than the corresponding input dimension, "
                           f"but got: padding (720, 720) at dimension 2 of input of length {y.shape[1]}")
    if y.is_cuda:
        return cosy_mel_batch(y)
    return cosy_mel_host(y, pin_output=False)


def fbank(waveform, blackman_coeff: float = 0.42, channel: int = -1, dither: float = 0.0,
          energy_floor: float = 1.0, frame_length: float = 25.0, frame_shift: float = 10.0,
          high_freq: float = 0.0, htk_compat: bool = False, low_freq: float = 20.0, min_duration: float = 0.0,
          num_mel_bins: int = 23, preemphasis_coefficient: float = 0.97, raw_energy: bool = True,
          remove_dc_offset: bool = True, round_to_power_of_two: bool = True, sample_frequency: float = 16000.0,
          snip_edges: bool = True, subtract_mean: bool = False, use_energy: bool = False,
          use_log_fbank: bool = True, use_power: bool = True, vtln_high: float = -500.0, vtln_low: float = 100.0,
          vtln_warp: float = 1.0, window_type: str = "povey"):
    """Drop-in for torchaudio.compliance.kaldi.fbank for the configuration the reference uses
    (num_mel_bins=80, dither=0, sample_frequency=16000, every other option at its default; subtract_mean
    is honoured).  waveform: (c, n) -> (m, 80).  Any other configuration is rejected loudly."""
    fixed = dict(blackman_coeff=(blackman_coeff, 0.42), dither=(dither, 0.0), energy_floor=(energy_floor, 1.0),
                 frame_length=(frame_length, 25.0), frame_shift=(frame_shift, 10.0), high_freq=(high_freq, 0.0),
                 htk_compat=(htk_compat, False), low_freq=(low_freq, 20.0), num_mel_bins=(num_mel_bins, 80),
                 preemphasis_coefficient=(preemphasis_coefficient, 0.97), raw_energy=(raw_energy, True),
                 remove_dc_offset=(remove_dc_offset, True), round_to_power_of_two=(round_to_power_of_two, True),
                 sample_frequency=(sample_frequency, 16000.0), snip_edges=(snip_edges, True),
                 use_energy=(use_energy, False), use_log_fbank=(use_log_fbank, True), use_power=(use_power, True),
                 vtln_warp=(vtln_warp, 1.0), window_type=(window_type, "povey"))
    bad = {k: v[0] for k, v in fixed.items() if v[0] != v[1]}
    if bad:
        raise NotImplementedError(f"fbank: only the CosyVoice speaker-embedding configuration is built "
                                  f"(num_mel_bins=80, dither=0, sample_frequency=16000, defaults otherwise); got {bad}")
    waveform = _as_float_tensor(waveform, "waveform")
    if waveform.dim() != 2:
        raise ValueError(f"fbank expects waveform of shape (c, n), got {tuple(waveform.shape)}")
    ch = max(channel, 0)
    assert ch < waveform.size(0), "Invalid channel {} for size {}".format(ch, waveform.size(0))
    x = waveform[ch:ch + 1, :]
    n = x.shape[1]
    assert 2 <= 400 <= n, "choose a window size {} that is [2, {}]".format(400, n)  # kaldi.py:142
    if n < min_duration * sample_frequency:
        return torch.empty(0, device=waveform.device, dtype=waveform.dtype)  # kaldi.py:595-597
    if x.is_cuda:
        return kaldi_fbank_batch(x, subtract_mean=subtract_mean)[0]
    return kaldi_fbank_host(x, subtract_mean=subtract_mean, pin_output=False)[0]


# ----------------------------------------------------------------------------------------- waveform conditioning
_RS_METHODS = {"sinc_interp_hann": _lib.RESAMPLE_HANN, "sinc_interpolation": _lib.RESAMPLE_HANN,
               "sinc_interp_kaiser": _lib.RESAMPLE_KAISER, "kaiser_window": _lib.RESAMPLE_KAISER}


def resample_out_len(n_frames, orig_freq, new_freq):
    """ceil(new * n / orig) with the ratio reduced by its gcd (torchaudio's target_length)."""
    return int(lib.b200feat_resample_out_len(int(n_frames), int(orig_freq), int(new_freq)))


def condition_pcm(pcm, orig_freq, new_freq, lengths=None, mix=None, lowpass_filter_width=6, rolloff=0.99,
                  resampling_method="sinc_interp_hann", beta=None, out=None):
    """Interleaved PCM -> mono -> resampled float32 waveform batch in ONE kernel (b200feat_resample), the row in
    front of the featurizers: `pcm` is (B, L) mono or (B, L, C) with C in {1, 2}, int16 (scaled by 1/32768 like
    whisper.audio.load_audio) or float32, on a CUDA device (a host tensor is copied over).  `mix` = per-channel
    weights (None: mean of the channels, audio_processor.py:314; (0.65, 0.35) is its mid + 0.3 * side branch, :316-318).
    Filter arguments are torchaudio.functional.resample's.  Returns (B, ceil(new * L / orig)) float32; with
    `lengths` (input frames per clip) every row is zero past its own ceil(new * len / orig).  Async on the current stream."""
    if isinstance(pcm, np.ndarray):
        pcm = torch.from_numpy(pcm)
    if not torch.is_tensor(pcm):
        raise TypeError(f"pcm must be a torch.Tensor or numpy.ndarray, got {type(pcm).__name__}")
    if pcm.dtype != torch.int16:
        if not pcm.is_floating_point():
            raise TypeError(f"Expected floating point type for waveform tensor, but received {pcm.dtype}.")
        pcm = pcm if pcm.dtype == torch.float32 else pcm.to(torch.float32)
    if not (int(orig_freq) == orig_freq and int(new_freq) == new_freq):
        raise Exception("Frequencies must be of integer type to ensure quality resampling computation.")
    if orig_freq <= 0.0 or new_freq <= 0.0:
        raise ValueError("Original frequency and desired frequecy should be positive")
    if resampling_method not in _RS_METHODS:
        raise ValueError("Invalid resampling method: {}".format(resampling_method))
    if lowpass_filter_width <= 0:
        raise ValueError("Low pass filter width should be positive.")
    if pcm.dim() == 2:
        pcm = pcm.unsqueeze(-1)
    if pcm.dim() != 3 or pcm.shape[2] not in (1, 2):
        raise ValueError(f"expected (B, L) or (B, L, C) PCM with C in (1, 2), got shape {tuple(pcm.shape)}")
    device = _cuda_device(pcm.device if pcm.is_cuda else None)
    x = pcm.to(device).contiguous()
    B, L, C = x.shape
    L_out = resample_out_len(L, orig_freq, new_freq)
    if out is None:
        out = torch.empty((B, L_out), dtype=torch.float32, device=device)
    elif out.shape != (B, L_out) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != device:
        raise ValueError(f"out must be a contiguous float32 ({B}, {L_out}) tensor on {device}")
    mix_arr = None
    if mix is not None:
        if len(mix) != C:
            raise ValueError(f"mix has {len(mix)} weights for {C} channel(s)")
        mix_arr = (ctypes.c_float * C)(*[float(v) for v in mix])
    ld = _lengths_dev(lengths, B, device)
    with torch.cuda.device(device):
        check(lib.b200feat_resample(_ptr(x), _DTYPES[x.dtype], C, mix_arr, _ptr(ld), B, L, L * C,
                                    int(orig_freq), int(new_freq), int(lowpass_filter_width), float(rolloff),
                                    _RS_METHODS[resampling_method], 14.769656459379492 if beta is None else float(beta),
                                    _ptr(out), L_out, L_out, _stream(device)), "b200feat_resample")
    return out


# /root/reference/Backend/services/audio_processor.py:296-303: the filter the reference's AudioProcessor deploys
REFERENCE_RESAMPLER = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="sinc_interp_kaiser", beta=14.769656459379492)


def _pcm_tensor(pcm):
    if isinstance(pcm, np.ndarray):
        pcm = torch.from_numpy(pcm)
    if not torch.is_tensor(pcm):
        raise TypeError(f"pcm must be a torch.Tensor or numpy.ndarray, got {type(pcm).__name__}")
    if pcm.dtype != torch.int16:
        if not pcm.is_floating_point():
            raise TypeError(f"Expected floating point type for waveform tensor, but received {pcm.dtype}.")
        pcm = pcm if pcm.dtype == torch.float32 else pcm.to(torch.float32)
    if pcm.dim() == 2:
        pcm = pcm.unsqueeze(-1)
    if pcm.dim() != 3 or pcm.shape[2] not in (1, 2):
        raise ValueError(f"expected (B, L) or (B, L, C) PCM with C in (1, 2), got shape {tuple(pcm.shape)}")
    return pcm


def _rs_args(orig_freq, new_freq, lowpass_filter_width, rolloff, resampling_method, beta):
    if not (int(orig_freq) == orig_freq and int(new_freq) == new_freq):
        raise Exception("Frequencies must be of integer type to ensure quality resampling computation.")
    if orig_freq <= 0.0 or new_freq <= 0.0:
        raise ValueError("Original frequency and desired frequecy should be positive")
    if resampling_method not in _RS_METHODS:
        raise ValueError("Invalid resampling method: {}".format(resampling_method))
    if lowpass_filter_width <= 0:
        raise ValueError("Low pass filter width should be positive.")
    return (int(orig_freq), int(new_freq), int(lowpass_filter_width), float(rolloff), _RS_METHODS[resampling_method],
            14.769656459379492 if beta is None else float(beta))


def condition_stereo_auto(pcm, orig_freq, new_freq, lengths=None, lowpass_filter_width=128, rolloff=0.95,
                          resampling_method="sinc_interp_kaiser", beta=14.769656459379492, return_decisions=False):
    """The reference's conditioning of a stereo clip, batched, on the device
    (/root/reference/Backend/services/audio_processor.py:294-319): resample BOTH channels (defaults = the reference's
    kaiser filter), take the inter-channel correlation of the resampled channels, and mix per clip:
    mean of the channels if correlation > 0.5, else mid + 0.3 * side.  pcm: (B, L, 2) int16 or float32 ->
    (B, ceil(new * L / orig)) float32 [, (decisions int32 (B,), correlations float32 (B,))]."""
    x = _pcm_tensor(pcm)
    if x.shape[2] != 2:
        raise ValueError("condition_stereo_auto expects (B, L, 2) stereo PCM")
    of, nf, width, ro, meth, bt = _rs_args(orig_freq, new_freq, lowpass_filter_width, rolloff, resampling_method, beta)
    device = _cuda_device(x.device if x.is_cuda else None)
    x = x.to(device).contiguous()
    B, L, C = x.shape
    L_out = resample_out_len(L, of, nf)
    out = torch.empty((B, L_out), dtype=torch.float32, device=device)
    scratch = torch.empty((B, L_out), dtype=torch.float32, device=device)
    dec = torch.empty((B,), dtype=torch.int32, device=device)
    cor = torch.empty((B,), dtype=torch.float32, device=device)
    ld = _lengths_dev(lengths, B, device)
    with torch.cuda.device(device):
        check(lib.b200feat_condition_stereo_auto(_ptr(x), _DTYPES[x.dtype], _ptr(ld), B, L, L * C, of, nf, width, ro, meth, bt,
                                                 _ptr(out), L_out, L_out, _ptr(scratch), scratch.numel(), _ptr(dec), _ptr(cor),
                                                 _stream(device)), "b200feat_condition_stereo_auto")
    return (out, (dec, cor)) if return_decisions else out


def pcm_whisper_log_mel_batch(pcm, orig_freq, new_freq=WHISPER_SR, lengths=None, mix=None, n_mels=80, padding=0, norm="per_clip",
                              out=None, out_dtype=torch.float32, lowpass_filter_width=6, rolloff=0.99,
                              resampling_method="sinc_interp_hann", beta=None):
    """PCM in, Whisper log-mel out, ONE library call on the device: (B, L) / (B, L, C) int16 or float32 PCM at orig_freq
    -> mono mix + resample to 16 kHz -> (B, n_mels, (ceil(16000 * L / orig) + padding) // 160).  The waveform in between lives
    in the call's workspace.  Same result as whisper_log_mel_batch(condition_pcm(pcm, ...), ...), bit for bit.
    `lengths` = input frames per clip."""
    assert n_mels in {80, 128}, f"Unsupported n_mels: {n_mels}"
    out_dtype = _out_dtype(out.dtype if out is not None else out_dtype)
    x = _pcm_tensor(pcm)
    of, nf, width, ro, meth, bt = _rs_args(orig_freq, new_freq, lowpass_filter_width, rolloff, resampling_method, beta)
    device = _cuda_device(x.device if x.is_cuda else None)
    x = x.to(device).contiguous()
    B, L, C = x.shape
    Lw = resample_out_len(L, of, nf)
    T = (Lw + int(padding)) // WHISPER_HOP
    if out is None:
        out = torch.empty((B, n_mels, T), dtype=out_dtype, device=device)
    elif out.shape != (B, n_mels, T) or not out.is_contiguous() or out.device != device:
        raise ValueError("out must be a contiguous float32 / float16 (B, n_mels, T) tensor on the PCM's device")
    if B == 0 or T == 0:
        return out
    mix_arr = None
    if mix is not None:
        if len(mix) != C:
            raise ValueError(f"mix has {len(mix)} weights for {C} channel(s)")
        mix_arr = (ctypes.c_float * C)(*[float(v) for v in mix])
    ld = _lengths_dev(lengths, B, device)
    nws = int(lib.b200feat_pcm_whisper_workspace_bytes(B, L, of, nf, int(padding)))
    ws = torch.empty(nws, dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        check(lib.b200feat_pcm_whisper_logmel(_ptr(x), _DTYPES[x.dtype], C, mix_arr, _ptr(ld), B, L, L * C, of, nf, width, ro, meth, bt,
                                              n_mels, int(padding), _NORMS[norm], _ptr(out), _DTYPES[out_dtype], n_mels * T, T,
                                              _ptr(ws), nws, _stream(device)), "b200feat_pcm_whisper_logmel")
    return out


def condition_pcm_host(pcm, orig_freq, new_freq, lengths=None, mix=None, out=None, device=None, pin_output=True,
                       lowpass_filter_width=6, rolloff=0.99, resampling_method="sinc_interp_hann", beta=None):
    """condition_pcm for host buffers: host (ideally pinned) int16 / float32 PCM -> host float32 waveform
    (B, ceil(new * L / orig)) through the library's chunked H2D / resampler / D2H pipeline.  Blocks."""
    x = _pcm_tensor(pcm)
    if x.is_cuda:
        raise ValueError("pcm: the *_host entry points take host memory")
    x = x.contiguous()
    of, nf, width, ro, meth, bt = _rs_args(orig_freq, new_freq, lowpass_filter_width, rolloff, resampling_method, beta)
    B, L, C = x.shape
    L_out = resample_out_len(L, of, nf)
    dev = _cuda_device(device)
    out = _host_out((B, L_out), out, pin_output)
    mix_arr = None
    if mix is not None:
        if len(mix) != C:
            raise ValueError(f"mix has {len(mix)} weights for {C} channel(s)")
        mix_arr = (ctypes.c_float * C)(*[float(v) for v in mix])
    ln = _host_lengths(lengths, B)
    if B and L_out:
        check(lib.b200feat_resample_host(_ptr(x), _DTYPES[x.dtype], C, mix_arr, _ptr(ln), B, L, of, nf, width, ro, meth, bt,
                                         _ptr(out), dev.index), "b200feat_resample_host")
    return out


def pcm_whisper_log_mel_host(pcm, orig_freq, new_freq=WHISPER_SR, lengths=None, mix=None, n_mels=80, padding=0, norm="per_clip",
                             out=None, out_dtype=torch.float32, device=None, pin_output=True, lowpass_filter_width=6,
                             rolloff=0.99, resampling_method="sinc_interp_hann", beta=None):
    """The host-buffer twin: host (ideally pinned) int16 / float32 PCM -> host log-mel through the library's chunked
    H2D / resample / log-mel / D2H pipeline.  PCM crosses PCIe, the float32 waveform never does.  Blocks."""
    assert n_mels in {80, 128}, f"Unsupported n_mels: {n_mels}"
    out_dtype = _out_dtype(out.dtype if out is not None else out_dtype)
    x = _pcm_tensor(pcm)
    if x.is_cuda:
        raise ValueError("pcm: the *_host entry points take host memory")
    x = x.contiguous()
    of, nf, width, ro, meth, bt = _rs_args(orig_freq, new_freq, lowpass_filter_width, rolloff, resampling_method, beta)
    B, L, C = x.shape
    Lw = resample_out_len(L, of, nf)
    T = (Lw + int(padding)) // WHISPER_HOP
    dev = _cuda_device(device)
    out = _host_out((B, n_mels, T), out, pin_output, out_dtype)
    mix_arr = None
    if mix is not None:
        if len(mix) != C:
            raise ValueError(f"mix has {len(mix)} weights for {C} channel(s)")
        mix_arr = (ctypes.c_float * C)(*[float(v) for v in mix])
    ln = _host_lengths(lengths, B)
    if B and T:
        check(lib.b200feat_pcm_whisper_logmel_host(_ptr(x), _DTYPES[x.dtype], C, mix_arr, _ptr(ln), B, L, of, nf, width, ro, meth, bt,
                                                   n_mels, int(padding), _NORMS[norm], _ptr(out), _DTYPES[out_dtype], dev.index),
              "b200feat_pcm_whisper_logmel_host")
    return out


def resample(waveform, orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99,
             resampling_method: str = "sinc_interp_hann", beta=None):
    """torchaudio.functional.resample (functional.py:1435): `(..., time)` -> `(..., ceil(new * time / orig))`, same
    arguments, defaults and errors; returns the input itself when the rates are equal.  The result lives on the
    input's device (a CPU tensor makes the round trip to the current CUDA device: there is no CPU path)."""
    if orig_freq <= 0.0 or new_freq <= 0.0:
        raise ValueError("Original frequency and desired frequecy should be positive")
    if orig_freq == new_freq:
        return waveform
    if not torch.is_tensor(waveform):
        raise TypeError(f"waveform must be a torch.Tensor, got {type(waveform).__name__}")
    if not waveform.is_floating_point():
        raise TypeError(f"Expected floating point type for waveform tensor, but received {waveform.dtype}.")
    shape = waveform.shape
    y = condition_pcm(waveform.reshape(-1, shape[-1]), orig_freq, new_freq, None, None, lowpass_filter_width, rolloff,
                      resampling_method, beta)
    return y.reshape(shape[:-1] + y.shape[-1:]).to(device=waveform.device, dtype=waveform.dtype)


class Resample(torch.nn.Module):
    """torchaudio.transforms.Resample as /root/reference/Backend/services/audio_processor.py:296-303 builds it
    (orig_freq, new_freq, resampling_method, lowpass_filter_width, rolloff, beta); calling it resamples `(..., time)`.
    A torch.nn.Module like the class it replaces (install() swaps it in process-wide), so `.to(device)`, `.eval()`,
    nn.Sequential membership and `state_dict()` keep working; it owns no parameters or buffers -- the tap table is a
    library-owned constant built once per (device, ratio, filter)."""

    def __init__(self, orig_freq: int = 16000, new_freq: int = 16000, resampling_method: str = "sinc_interp_hann",
                 lowpass_filter_width: int = 6, rolloff: float = 0.99, beta=None, *, dtype=None):
        super().__init__()
        self.orig_freq, self.new_freq, self.resampling_method = orig_freq, new_freq, resampling_method
        self.lowpass_filter_width, self.rolloff, self.beta = lowpass_filter_width, rolloff, beta
        self.gcd = int(np.gcd(int(orig_freq), int(new_freq)))

    def forward(self, waveform):
        return resample(waveform, self.orig_freq, self.new_freq, self.lowpass_filter_width, self.rolloff,
                        self.resampling_method, self.beta)

    def extra_repr(self):
        return f"orig_freq={self.orig_freq}, new_freq={self.new_freq}, resampling_method={self.resampling_method!r}"


def cosy_prompt_feat(wave, token_len=None, lengths=None):
    """CosyVoice frontend `_extract_speech_feat` + the 24 kHz trim of `frontend_zero_shot` (UPSTREAM cosyvoice/cli/
    frontend.py, reached from /root/reference/Docker/cosyvoice_api.py:128): (B, L) 24 kHz prompt batch -> frame-major
    (B, T, 80) ln-mel written in that layout by the kernel (no transpose pass); with `token_len` (speech tokens of
    the prompt, 25 Hz against 50 Hz mel frames) T is cut to 2 * min(T // 2, token_len) like
    `speech_feat[:, :2 * token_len]`.  Returns (features view, token_len actually used)."""
    feat = cosy_mel_batch(wave, lengths=lengths, frame_major=True)
    if token_len is None:
        return feat, feat.shape[1] // 2
    tl = min(feat.shape[1] // 2, int(token_len))
    return feat[:, :2 * tl], tl


# ----------------------------------------------------------------------------------------- install / uninstall
_PATCHED = []


def install(verbose=False):
    """Point the reference's call sites at this library by patching the third-party module attributes
    they resolve at call time.  Modules that are not importable are skipped.  Returns the patched names."""
    import importlib
    targets = [("whisper.audio", "log_mel_spectrogram", log_mel_spectrogram),
               ("whisper.transcribe", "log_mel_spectrogram", log_mel_spectrogram),
               ("whisper", "log_mel_spectrogram", log_mel_spectrogram),
               ("matcha.utils.audio", "mel_spectrogram", mel_spectrogram),
               ("audio.audio", "melspectrogram", melspectrogram),        # Backend/diff2lip: `from audio import audio`
               ("torchaudio.compliance.kaldi", "fbank", fbank),
               ("torchaudio.functional", "resample", resample),
               ("torchaudio.transforms", "Resample", Resample)]
    done = []
    for mod_name, attr, fn in targets:
        try:
            mod = importlib.import_module(mod_name)
        except Exception:
            continue
        if not hasattr(mod, attr):
            continue
        _PATCHED.append((mod, attr, getattr(mod, attr)))
        setattr(mod, attr, fn)
        done.append(f"{mod_name}.{attr}")
        if verbose:
            print(f"[b200feat] patched {mod_name}.{attr}")
    return done


def uninstall():
    while _PATCHED:
        mod, attr, orig = _PATCHED.pop()
        setattr(mod, attr, orig)
