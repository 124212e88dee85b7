"""ctypes binding of libb200feat.so (the C ABI in include/b200feat.h).

There is no CPU fallback and no other backend: if the shared library is missing or fails to load the
import raises.  Build it with `python -m expressive_speech_translation_b200.build` (or
`__graft_entry__.build()`).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200FEAT_LIB: another build of the same library (A/B timing of kernel variants); there is still no fallback
LIB_PATH = os.environ.get("B200FEAT_LIB") or os.path.join(_HERE, "lib", "libb200feat.so")

KIND_WHISPER, KIND_COSY, KIND_KALDI = 0, 1, 2
NORM_PER_CLIP, NORM_PER_CALL = 0, 1
DT_F32, DT_F16, DT_S16 = 0, 1, 2
RESAMPLE_HANN, RESAMPLE_KAISER = 0, 1

# every symbol include/b200feat.h declares: (restype, argtypes)
_c = ctypes
SYMBOLS = {
    "b200feat_version": (_c.c_int, []),
    "b200feat_last_error": (_c.c_char_p, []),
    "b200feat_num_frames": (_c.c_int64, [_c.c_int32, _c.c_int64, _c.c_int32]),
    "b200feat_workspace_bytes": (_c.c_int64, [_c.c_int32, _c.c_int32, _c.c_int64]),
    "b200feat_whisper_logmel": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64,
                                           _c.c_int32, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_int64,
                                           _c.c_int64, _c.c_void_p, _c.c_int64, _c.c_void_p]),
    "b200feat_whisper_logmel_ex": (_c.c_int, [_c.c_void_p, _c.c_int32, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64,
                                              _c.c_int32, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_int32, _c.c_int64,
                                              _c.c_int64, _c.c_void_p, _c.c_int64, _c.c_void_p]),
    "b200feat_noalloc_workspace_max": (_c.c_int64, []),
    "b200feat_whisper_logmel_noalloc": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64,
                                                   _c.c_int32, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_int64,
                                                   _c.c_int64, _c.c_void_p]),
    "b200feat_cosy_mel": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64,
                                     _c.c_void_p, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_int64,
                                     _c.c_void_p]),
    "b200feat_cosy_mel_frames": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64,
                                            _c.c_void_p, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_int64,
                                            _c.c_void_p]),
    "b200feat_kaldi_fbank": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64,
                                        _c.c_int32, _c.c_void_p, _c.c_int64, _c.c_int64, _c.c_void_p,
                                        _c.c_int64, _c.c_void_p]),
    "b200feat_whisper_logmel_host": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int32,
                                                _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_int32]),
    "b200feat_whisper_logmel_ex_host": (_c.c_int, [_c.c_void_p, _c.c_int32, _c.c_void_p, _c.c_int32, _c.c_int64,
                                                   _c.c_int32, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_int32,
                                                   _c.c_int32]),
    "b200feat_cosy_mel_host": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_void_p,
                                          _c.c_int32]),
    "b200feat_kaldi_fbank_host": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int32,
                                             _c.c_void_p, _c.c_int32]),
    "b200feat_resample_out_len": (_c.c_int64, [_c.c_int64, _c.c_int32, _c.c_int32]),
    "b200feat_resample": (_c.c_int, [_c.c_void_p, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_void_p, _c.c_int32,
                                     _c.c_int64, _c.c_int64, _c.c_int32, _c.c_int32, _c.c_int32, _c.c_double,
                                     _c.c_int32, _c.c_double, _c.c_void_p, _c.c_int64, _c.c_int64, _c.c_void_p]),
    "b200feat_release": (None, []),
    "b200feat_launch_count": (_c.c_int64, []),
    "b200feat_whisper_logmel_windows": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64, _c.c_int32,
                                                   _c.c_int32, _c.c_int32, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_int32,
                                                   _c.c_int64, _c.c_void_p, _c.c_int64, _c.c_void_p]),
    "b200feat_wav2lip_num_frames": (_c.c_int64, [_c.c_int64]),
    "b200feat_wav2lip_mel": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64, _c.c_int32, _c.c_int32,
                                        _c.c_void_p, _c.c_int64, _c.c_int64, _c.c_void_p]),
    "b200feat_condition_stereo_auto": (_c.c_int, [_c.c_void_p, _c.c_int32, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64,
                                                  _c.c_int32, _c.c_int32, _c.c_int32, _c.c_double, _c.c_int32, _c.c_double,
                                                  _c.c_void_p, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_int64, _c.c_void_p,
                                                  _c.c_void_p, _c.c_void_p]),
    "b200feat_pcm_whisper_workspace_bytes": (_c.c_int64, [_c.c_int32, _c.c_int64, _c.c_int32, _c.c_int32, _c.c_int32]),
    "b200feat_pcm_whisper_logmel": (_c.c_int, [_c.c_void_p, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_void_p, _c.c_int32,
                                               _c.c_int64, _c.c_int64, _c.c_int32, _c.c_int32, _c.c_int32, _c.c_double,
                                               _c.c_int32, _c.c_double, _c.c_int32, _c.c_int32, _c.c_int32, _c.c_void_p,
                                               _c.c_int32, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_int64, _c.c_void_p]),
    "b200feat_pcm_whisper_logmel_host": (_c.c_int, [_c.c_void_p, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_void_p, _c.c_int32,
                                                    _c.c_int64, _c.c_int32, _c.c_int32, _c.c_int32, _c.c_double, _c.c_int32,
                                                    _c.c_double, _c.c_int32, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_int32,
                                                    _c.c_int32]),
    "b200feat_resample_host": (_c.c_int, [_c.c_void_p, _c.c_int32, _c.c_int32, _c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64,
                                          _c.c_int32, _c.c_int32, _c.c_int32, _c.c_double, _c.c_int32, _c.c_double, _c.c_void_p,
                                          _c.c_int32]),
    "b200feat_spectral_gate_out_len": (_c.c_int64, [_c.c_int64]),
    "b200feat_spectral_gate_workspace_bytes": (_c.c_int64, [_c.c_int32, _c.c_int64]),
    "b200feat_spectral_gate": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_int64,
                                          _c.c_void_p, _c.c_int64, _c.c_void_p]),
    "b200feat_profile_begin": (_c.c_int, []),
    "b200feat_profile_end": (_c.c_int, [_c.POINTER(_c.c_double), _c.POINTER(_c.c_int64), _c.POINTER(_c.c_double),
                                        _c.POINTER(_c.c_int64)]),
}


class B200FeatError(RuntimeError):
    pass


def load(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: the CUDA extension is the only implementation of this package "
            "(no CPU fallback). Build it with `python -m expressive_speech_translation_b200.build`.")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()


def check(rc, what):
    if rc != 0:
        msg = lib.b200feat_last_error().decode("utf-8", "replace")
        if rc == -1:
            raise ValueError(f"{what}: {msg}")
        raise B200FeatError(f"{what} failed (code {rc}): {msg}")
