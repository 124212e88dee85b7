"""CPU suite: libb200feat.so loads, exports every symbol include/b200feat.h declares, and its argument
validation / pure-host helpers behave (no kernel is launched here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from expressive_speech_translation_b200 import build, _lib
    build.build()
    return _lib.lib


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "b200feat.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200feat_[a-z_0-9]+)\s*\(", txt)))


def test_header_symbols_are_exported_and_bound(lib):
    from expressive_speech_translation_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200feat.h but not exported"
    assert sorted(_lib.SYMBOLS) == names, "the ctypes binding must cover exactly the declared ABI"


def test_no_torch_or_cxx_types_in_header():
    txt = open(os.path.join(ROOT, "include", "b200feat.h")).read()
    code = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    for bad in ("std::", "at::", "torch", "Tensor", "template", "class "):
        assert bad not in code
    assert 'extern "C"' in code


def test_version_frames_workspace(lib):
    assert lib.b200feat_version() == 101
    assert lib.b200feat_num_frames(0, 480000, 0) == 3000
    assert lib.b200feat_num_frames(0, 960000, 480000) == 9000
    assert lib.b200feat_num_frames(0, 293700, 0) == 1835
    assert lib.b200feat_num_frames(1, 240000, 0) == 500
    assert lib.b200feat_num_frames(1, 479, 0) == 0
    assert lib.b200feat_num_frames(2, 80000, 0) == 498
    assert lib.b200feat_num_frames(2, 399, 0) == 0
    assert lib.b200feat_num_frames(9, 1, 0) == -1
    # Whisper: 16 + keys (4 B per clip, 16-aligned) + (min, max) per 64-frame unit; L includes the padding
    assert lib.b200feat_workspace_bytes(0, 1024, 480000) == 16 + 4096 + 1024 * 47 * 8
    assert lib.b200feat_workspace_bytes(1, 256, 240000) == 16
    assert lib.b200feat_workspace_bytes(2, 4096, 80000) == 16 + 4096 * 16 * 80 * 4
    assert lib.b200feat_launch_count() >= 0


def test_argument_validation_without_gpu(lib):
    one = ctypes.c_void_p(16)  # never dereferenced: validation fails first
    rc = lib.b200feat_whisper_logmel(one, None, 1, 16000, 16000, 64, 0, 0, one, 64 * 100, 100, one, 64, None)
    assert rc == -1 and b"Unsupported n_mels" in lib.b200feat_last_error()
    rc = lib.b200feat_whisper_logmel(one, None, 1, 16000, 100, 80, 0, 0, one, 8000, 100, one, 64, None)
    assert rc == -1 and b"wave_stride" in lib.b200feat_last_error()
    rc = lib.b200feat_whisper_logmel(one, None, 1, 16000, 16000, 80, 0, 0, one, 8000, 99, one, 64, None)
    assert rc == -1 and b"T_out" in lib.b200feat_last_error()
    rc = lib.b200feat_whisper_logmel(one, None, 1, 100, 100, 80, 0, 0, one, 80, 1, one, 64, None)
    assert rc == -1 and b"reflect" in lib.b200feat_last_error()
    rc = lib.b200feat_whisper_logmel(one, None, 4, 16000, 16000, 80, 0, 0, one, 8000, 100, one, 16, None)
    assert rc == -3
    rc = lib.b200feat_kaldi_fbank(one, None, 1, 16000, 16000, 1, one, 98 * 80, 97, one, 1 << 20, None)
    assert rc == -1
    rc = lib.b200feat_cosy_mel(None, None, 1, 24000, 24000, None, 4000, 50, None, 0, None)
    assert rc == -1 and b"null" in lib.b200feat_last_error()
    # empty batches are a no-op success
    assert lib.b200feat_whisper_logmel(None, None, 0, 16000, 16000, 80, 0, 0, None, 8000, 100, one, 64, None) == 0
    assert lib.b200feat_whisper_logmel_host(None, None, 0, 16000, 80, 0, 0, None, 0) == 0


def test_python_signatures_mirror_reference():
    import inspect
    import expressive_speech_translation_b200 as m
    s = inspect.signature(m.log_mel_spectrogram)
    assert list(s.parameters) == ["audio", "n_mels", "padding", "device"]
    assert s.parameters["n_mels"].default == 80 and s.parameters["padding"].default == 0 and s.parameters["device"].default is None
    s = inspect.signature(m.mel_spectrogram)
    assert list(s.parameters) == ["y", "n_fft", "num_mels", "sampling_rate", "hop_size", "win_size", "fmin", "fmax", "center"]
    assert s.parameters["center"].default is False
    ta = pytest.importorskip("torchaudio")
    ref = inspect.signature(ta.compliance.kaldi.fbank)
    mine = inspect.signature(m.fbank)
    assert list(ref.parameters) == list(mine.parameters)
    for k in ref.parameters:
        assert ref.parameters[k].default == mine.parameters[k].default, k


def test_python_rejects_unsupported_configs_loudly():
    import torch
    import expressive_speech_translation_b200 as m
    with pytest.raises(AssertionError, match="Unsupported n_mels"):
        m.log_mel_spectrogram(torch.zeros(16000), n_mels=64)
    with pytest.raises(TypeError):
        m.log_mel_spectrogram(torch.zeros(16000, dtype=torch.int16))
    with pytest.raises(NotImplementedError):
        m.mel_spectrogram(torch.zeros(1, 24000), 1024, 80, 22050, 256, 1024, 0, 8000)
    with pytest.raises(NotImplementedError):
        m.fbank(torch.zeros(1, 16000), num_mel_bins=23)
    with pytest.raises(AssertionError, match="choose a window size"):
        m.fbank(torch.zeros(1, 399), num_mel_bins=80)
    with pytest.raises(RuntimeError, match="Padding size"):
        m.log_mel_spectrogram(torch.zeros(150))
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU path"):
            m.log_mel_spectrogram(torch.zeros(16000))


def test_hf_feature_extractor_mirror_argument_handling(tmp_path):
    """the transformers.WhisperFeatureExtractor mirror: constructor / from_pretrained / the errors raised before any
    device work (the numerics are a -m gpu test against the installed transformers class)."""
    import numpy as np
    import torch
    import expressive_speech_translation_b200 as m
    fx = m.WhisperFeatureExtractor()
    assert (fx.feature_size, fx.sampling_rate, fx.hop_length, fx.n_fft, fx.n_samples, fx.nb_max_frames) == (80, 16000, 160, 400, 480000, 3000)
    x = np.zeros(16000, dtype=np.float32)
    with pytest.raises(ValueError, match="sampling rate of 16000"):
        fx(x, sampling_rate=44100)
    for kw in (dict(do_normalize=True), dict(pad_to_multiple_of=8), dict(padding="do_not_pad"), dict(return_tensors="tf")):
        with pytest.raises(NotImplementedError):
            fx(x, sampling_rate=16000, **kw)
    with pytest.raises(ValueError, match="mono-channel"):
        fx(np.zeros((2, 2, 100), dtype=np.float32), sampling_rate=16000)
    with pytest.raises(NotImplementedError):
        m.WhisperFeatureExtractor(feature_size=64)
    with pytest.raises(NotImplementedError):
        m.WhisperFeatureExtractor(sampling_rate=22050)
    (tmp_path / "preprocessor_config.json").write_text('{"feature_size": 128, "chunk_length": 30, "hop_length": 160, "n_fft": 400, '
                                                       '"sampling_rate": 16000, "processor_class": "WhisperProcessor"}')
    fx = m.WhisperFeatureExtractor.from_pretrained(str(tmp_path))
    assert fx.feature_size == 128
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU path"):
            fx(x, sampling_rate=16000)
