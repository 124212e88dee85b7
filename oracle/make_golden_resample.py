"""Golden fixture of the waveform-conditioning row (TEST INFRASTRUCTURE; run in the authoring container only -- it is
the one script besides oracle/make_golden.py that reads /root/reference).

  tests/golden/harvard_pcm_2s.npz
    pcm            int16 (88200, 2): the first 2 s of /root/reference/Backend/harvard.wav (44.1 kHz stereo s16)
    ta_hann_f32    torchaudio.functional.resample(mean(pcm / 32768), 44100, 16000) -- the REAL reference code, float32
                   (SURVEY.md section 8(d) cfg1 recipe; the first 2 s of harvard_w80.npz's `wave16k` up to the edge)
    ta_kaiser_f32  torchaudio.transforms.Resample(44100, 16000, lowpass_filter_width=128, rolloff=0.95,
                   resampling_method='sinc_interp_kaiser', beta=14.769656459379492) on both channels, then the mean --
                   what /root/reference/Backend/services/audio_processor.py:295-314 computes, float32
    fp64_hann / fp64_kaiser   oracle.resample.condition in float64 (stored rounded to float32)
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import resample as orc  # noqa: E402

KAISER = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="sinc_interp_kaiser", beta=14.769656459379492)


def main():
    import torchaudio
    from scipy.io import wavfile
    path = "/root/reference/Backend/harvard.wav"
    sha = hashlib.sha256(open(path, "rb").read()).hexdigest()
    sr, pcm = wavfile.read(path)
    assert sr == 44100 and pcm.dtype == np.int16 and pcm.shape[1] == 2
    pcm = np.ascontiguousarray(pcm[:88200])
    x = torch.from_numpy(pcm.astype(np.float32) / 32768.0)                 # (L, 2)
    ta_hann = torchaudio.functional.resample(x.mean(dim=1), 44100, 16000).numpy()
    ta_kaiser = torchaudio.transforms.Resample(orig_freq=44100, new_freq=16000, **KAISER)(x.T.contiguous()).mean(dim=0).numpy()
    t = torch.from_numpy(pcm)[None]
    f_hann = orc.condition(t, 44100, 16000)[0].numpy()
    f_kaiser = orc.condition(t, 44100, 16000, **KAISER)[0].numpy()
    print("real torchaudio fp32 vs fp64 oracle: hann %.3e  kaiser %.3e" %
          (np.abs(ta_hann - f_hann).max(), np.abs(ta_kaiser - f_kaiser).max()))
    whole = np.load(os.path.join(ROOT, "tests", "golden", "harvard_w80.npz"))["wave16k"]
    print("vs first 2 s of harvard_w80.npz wave16k (away from the cut): %.3e" % np.abs(whole[:31900] - ta_hann[:31900]).max())
    out = os.path.join(ROOT, "tests", "golden", "harvard_pcm_2s.npz")
    np.savez_compressed(out, pcm=pcm, ta_hann_f32=ta_hann, ta_kaiser_f32=ta_kaiser, fp64_hann=f_hann.astype(np.float32),
                        fp64_kaiser=f_kaiser.astype(np.float32), source_sha256=np.array(sha),
                        versions=np.array(f"torch {torch.__version__} torchaudio {torchaudio.__version__}"))
    print(out, os.path.getsize(out))


if __name__ == "__main__":
    main()
