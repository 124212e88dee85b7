"""Mirror of transformers' WhisperFeatureExtractor call for the log-mel it computes.

The reference reaches it through MuseTalk's AudioProcessor (`AudioProcessor(feature_extractor_path=...)` at
/root/reference/Docker/api_inference_logic.py:62-65, `get_audio_feature` at :99-102), which loads the clip at
16 kHz, cuts it into 30-s segments and calls `feature_extractor(segment, return_tensors="pt",
sampling_rate=16000).input_features` on each.  The math is the one installed at
transformers/models/whisper/feature_extraction_whisper.py (zero-pad / truncate to 30 s, periodic-Hann STFT 400/160,
Slaney mel, log10, drop the last frame, clamp to (max - 8) PER CLIP, (x + 4) / 4) -- exactly
`whisper_log_mel_host(..., norm="per_clip")` on the zero-padded batch, so this class only does the padding and the
bookkeeping around one library call.  Same constructor / call arguments and error behaviour for the subset the
reference uses; anything else raises NotImplementedError instead of silently doing something different.
"""
import json
import os

import numpy as np
import torch

from . import api


class Features(dict):
    """dict with attribute access, like transformers.BatchFeature (`out.input_features` / `out["input_features"]`)."""

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name) from None


class WhisperFeatureExtractor:
    model_input_names = ["input_features"]

    def __init__(self, feature_size=80, sampling_rate=16000, hop_length=160, chunk_length=30, n_fft=400,
                 padding_value=0.0, dither=0.0, return_attention_mask=False, **kwargs):
        if feature_size not in (80, 128):
            raise NotImplementedError(f"feature_size {feature_size}: the library carries the 80- and 128-bin Whisper filters")
        if (sampling_rate, hop_length, n_fft) != (16000, 160, 400):
            raise NotImplementedError("only Whisper's front-end parameters (16 kHz, n_fft 400, hop 160) are built")
        if padding_value != 0.0 or dither != 0.0:
            raise NotImplementedError("padding_value / dither other than 0 are not supported")
        self.feature_size, self.sampling_rate, self.hop_length = feature_size, sampling_rate, hop_length
        self.chunk_length, self.n_fft, self.padding_value, self.dither = chunk_length, n_fft, padding_value, dither
        self.n_samples = chunk_length * sampling_rate
        self.nb_max_frames = self.n_samples // hop_length
        self.return_attention_mask = return_attention_mask

    @classmethod
    def from_pretrained(cls, path, **kwargs):
        """reads <path>/preprocessor_config.json like transformers does for a local directory (no hub access here)."""
        cfg_path = os.path.join(path, "preprocessor_config.json") if os.path.isdir(path) else path
        with open(cfg_path) as f:
            cfg = json.load(f)
        keys = ("feature_size", "sampling_rate", "hop_length", "chunk_length", "n_fft", "padding_value", "dither",
                "return_attention_mask")
        args = {k: cfg[k] for k in keys if k in cfg}
        args.update(kwargs)
        return cls(**args)

    def __call__(self, raw_speech, truncation=True, pad_to_multiple_of=None, return_tensors=None,
                 return_attention_mask=None, padding="max_length", max_length=None, sampling_rate=None,
                 do_normalize=None, device="cpu", **kwargs):
        if sampling_rate is not None and sampling_rate != self.sampling_rate:
            raise ValueError(
                f"The model corresponding to this feature extractor: {self.__class__.__name__} was trained using a"
                f" sampling rate of {self.sampling_rate}. Please make sure that the provided `raw_speech` input"
                f" was sampled with {self.sampling_rate} and not {sampling_rate}.")
        if pad_to_multiple_of is not None or do_normalize:
            raise NotImplementedError("pad_to_multiple_of / do_normalize are not supported")
        if padding not in ("max_length", "longest", True):
            raise NotImplementedError(f"padding={padding!r}: only 'max_length' and 'longest' are supported")
        if return_tensors not in (None, "pt", "np"):
            raise NotImplementedError(f"return_tensors={return_tensors!r}: only 'pt', 'np' or None")

        if isinstance(raw_speech, torch.Tensor):
            raw_speech = raw_speech.detach().cpu().numpy()
        if isinstance(raw_speech, np.ndarray) and raw_speech.ndim > 2:
            raise ValueError(f"Only mono-channel audio is supported for input to {self}")
        batched = (isinstance(raw_speech, np.ndarray) and raw_speech.ndim == 2) or (
            isinstance(raw_speech, (list, tuple)) and len(raw_speech) > 0
            and isinstance(raw_speech[0], (np.ndarray, list, tuple, torch.Tensor)))
        clips = [np.asarray(c, dtype=np.float32).reshape(-1) for c in (raw_speech if batched else [raw_speech])]

        lens = [len(c) for c in clips]
        if padding == "max_length":
            L = int(max_length) if max_length else self.n_samples
            if not truncation and max(lens) > L:
                raise NotImplementedError("clips longer than max_length with truncation=False")
        else:
            L = max(lens)
            if truncation and max_length:
                L = min(L, int(max_length))
        lens = [min(n, L) for n in lens]
        batch = torch.zeros((len(clips), L), dtype=torch.float32, pin_memory=torch.cuda.is_available())
        for b, (c, n) in enumerate(zip(clips, lens)):
            batch[b, :n] = torch.from_numpy(c[:n])

        # zero-padded clips, maximum taken per clip: transformers' _torch_extract_fbank_features / _np_extract_fbank_features
        feats = api.whisper_log_mel_host(batch, n_mels=self.feature_size, norm="per_clip",
                                         device=None if device in (None, "cpu") else device)
        out = Features(input_features=feats)
        if return_attention_mask if return_attention_mask is not None else self.return_attention_mask:
            mask = torch.zeros((len(clips), L), dtype=torch.int32)
            for b, n in enumerate(lens):
                mask[b, :n] = 1
            m = mask[:, :: self.hop_length]
            if L % self.hop_length != 0:
                m = m[:, :-1]
            out["attention_mask"] = m
        if return_tensors != "pt":
            for k in list(out):
                out[k] = out[k].numpy()
        return out
