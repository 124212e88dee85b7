"""B200-native (sm_100a) waveform -> log-mel / mel / fbank front-end.

Drop-in behind the Python call signatures the reference pipeline uses (Whisper log_mel_spectrogram,
CosyVoice/Matcha mel_spectrogram, torchaudio kaldi fbank); the arithmetic is hand-written CUDA reached
through the C ABI in include/b200feat.h.  Importing this package loads
expressive_speech_translation_b200/lib/libb200feat.so and raises if it is missing: there is no CPU
fallback and no alternative backend.
"""
from ._lib import LIB_PATH, B200FeatError  # noqa: F401
from .api import (  # noqa: F401
    log_mel_spectrogram,
    mel_spectrogram,
    fbank,
    whisper_log_mel_batch,
    whisper_log_mel_windows,
    wav2lip_mel_batch,
    wav2lip_num_frames,
    melspectrogram,
    spectral_gate_batch,
    whisper_num_windows,
    cosy_mel_batch,
    cosy_prompt_feat,
    kaldi_fbank_batch,
    whisper_log_mel_host,
    cosy_mel_host,
    kaldi_fbank_host,
    condition_pcm,
    condition_stereo_auto,
    condition_pcm_host,
    pcm_whisper_log_mel_batch,
    pcm_whisper_log_mel_host,
    REFERENCE_RESAMPLER,
    resample,
    resample_out_len,
    Resample,
    num_frames,
    launch_count,
    profile_begin,
    profile_end,
    install,
    uninstall,
)
from .hf import WhisperFeatureExtractor  # noqa: F401

__version__ = "0.1.0"
