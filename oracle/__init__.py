"""CPU oracle for the waveform -> log-mel / mel / fbank hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker / the timed CPU
baseline.  The product (``expressive_speech_translation_b200``) never imports it
and has no CPU fallback.

Parity status: the reference (RobMills28/Expressive-Speech-Translation) holds NO
golden vectors or numeric tests for this path, and the
arithmetic lives in third-party packages (openai-whisper [unpinned], Matcha-TTS
inside the missing CosyVoice submodule fca1df3e, torchaudio.compliance.kaldi).
The oracle is pinned instead against the implementations that ARE importable in
the authoring container (see ``oracle/make_golden.py``):
  * Whisper log-mel  : transformers.WhisperFeatureExtractor torch + numpy paths
                       (independent restatements of whisper/audio.py)
  * kaldi fbank      : torchaudio.compliance.kaldi.fbank itself (the real code
                       the reference's CosyVoice frontend calls)
  * CosyVoice mel    : restated from matcha/utils/audio.py::mel_spectrogram with
                       the parameters bound at /root/reference/greek_sft.yaml:55-63 and
                       pinned against TWO independent implementations of the same mel
                       (torchaudio.transforms.MelSpectrogram, transformers.audio_utils.
                       spectrogram; tests/test_oracle.py, arrays stored in
                       tests/golden/cosy_rand.npz).  Matcha's `+ 1e-9` under the square
                       root is the one constant neither twin has; its effect is bounded
                       analytically in the test (<= 5e-6 relative on energies >= 1e-2).
                       openai-whisper and matcha themselves are not importable here, so
                       a1 / a2 are pinned to independent twins, not to the upstream files.
  * resampler        : oracle/resample.py runs the real torchaudio.functional.resample /
                       transforms.Resample and pins a float64 restatement to them; the stereo
                       rule is the reference's own torch lines (audio_processor.py:306-319)
  * Wav2Lip mel      : oracle/wav2lip.py restates the reference-held
                       Backend/diff2lip/audio/audio.py:45-51 line by line (librosa 0.9.2 is absent:
                       its stft / filters.mel restated), pinned by torch.stft + torchaudio
                       melscale_fbanks and by transformers.audio_utils.spectrogram
                       (oracle/make_golden_w2l.py, tests/golden/w2l_rand.npz)
  * spectral gate    : oracle/gate.py IS the reference's torch code
                       (Backend/services/audio_processor.py:321-358), runnable here
"""
from .filters import slaney_mel_filters, kaldi_mel_banks  # noqa: F401
from .featurizers import (  # noqa: F401
    whisper_log_mel,
    cosyvoice_mel,
    kaldi_fbank,
    whisper_num_frames,
    cosyvoice_num_frames,
    kaldi_num_frames,
)


# The ONE statement of the kaldi log-feature tolerance (tests/conftest.py::kaldi_check, __graft_entry__.smoke):
# |f - f64| <= max(1e-4, KALDI_TOL_FACTOR x max|f32_reference - f64|) on the same input.  Why a factor at all: a
# float32 FFT carries an error ~1e-7 of the FRAME's energy, so the log of a nearly empty mel bin is off by a few 1e-4
# in the reference's own float32 path (tests/golden/kaldi_rand.npz: torchaudio fp32 vs fp64 = 4.5e-4 .. 5.3e-4).
KALDI_TOL_FACTOR = 4.0


def kaldi_log_tolerance(f32_reference, f64):
    import numpy as _np
    return max(1e-4, KALDI_TOL_FACTOR * float(_np.abs(_np.asarray(f32_reference, dtype=_np.float64) - _np.asarray(f64)).max()))
