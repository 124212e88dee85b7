#!/usr/bin/env python
"""Benchmark of the hot path: batched waveform -> Whisper large-v3 128-bin log-mel (BASELINE.json configs[1]:
1024 synthetic 30-s 16 kHz chunks per GPU).  One step = one pass over the whole batch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload cfg2|cfg3|cfg4]

--workload (default cfg2, the headline the driver runs) measures the other two featurizers of SURVEY.md section 8(a)
with the same contract: cfg3 = CosyVoice 24 kHz mel on 256 x 10 s, cfg4 = kaldi fbank + CMN on 4096 x 5 s.

N > 1 is launched by torchrun (one rank per GPU); every rank featurizes its own shard of 1024 clips (weak
scaling, clips are independent: no data-path collective), the timed region is bracketed by barrier +
synchronize, the step time is the MAX over ranks and `value` is the whole-job audio-seconds per second.

  value      device-resident inputs, CUDA events on the launching stream
  e2e        same metric through the host-buffer API (pinned host in, pinned host out; H2D + D2H inside)
  roofline   the featurizer kernel's own launches timed with CUDA events (library profiling hooks),
             algorithmic bytes = 4*(samples in) + 4*(features out) per launch
  cpu_baseline / --impl reference
             the reference's own CPU ops for this path (oracle fp32 = torch.stft + dense mel matmul + clamp/
             log10/max-8/(x+4)/4, exactly whisper.audio.log_mel_spectrogram) on a bounded sample, all host threads
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

def _load_manifest_helpers():
    """expressive_speech_translation_b200/manifest.py loaded by path: pure shard arithmetic, usable by the reference arm
    without importing the package (and with it the CUDA library)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_b200_manifest", os.path.join(ROOT, "expressive_speech_translation_b200", "manifest.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


manifest = _load_manifest_helpers()

METRIC = "audio_seconds_per_second_log_mel"
UNIT = "audio-s/s"


class Workload:
    """one BASELINE.json config: shapes, the device / host / CPU-reference step functions and the roofline bytes.
    The default (and the only one the driver runs) is cfg2; cfg3 / cfg4 are the other two featurizers of SURVEY.md
    section 8(a), measured with the same contract through --workload."""

    def __init__(self, key, rank=0, world=1):
        self.key = key
        self.world = world
        if key == "cfg2":
            self.B, self.L, self.sr, self.n_mels = 1024, 480000, 16000, 128
            self.T = self.L // 160
            self.out_shape = (self.B, self.n_mels, self.T)
            self.kernel = "fe_whisper3_kernel<128, 0, 0>"
            self.cpu_clips = 256
            self.desc = "BASELINE configs[1]: Whisper large-v3 128-bin log-mel, 1024 x 30 s @ 16 kHz per GPU"
            self.featurizer = "whisper_log_mel"
            self.traffic_key = "fe_whisper3_kernel_128_dram_bytes_per_1024_clips"
            self.cpu_desc = "oracle fp32 = torch.stft + dense mel matmul + log10/max-8 (the ops whisper.audio runs)"
        elif key == "cfg3":
            self.B, self.L, self.sr, self.n_mels = 256, 240000, 24000, 80
            self.T = self.L // 480
            self.out_shape = (self.B, self.n_mels, self.T)
            self.kernel = "fe_cosy_kernel"
            self.cpu_clips = 64
            self.desc = "BASELINE configs[2]: CosyVoice 24 kHz 80-bin mel_spectrogram (n_fft 1920, hop 480), 256 x 10 s per GPU"
            self.featurizer = "cosy_mel"
            self.traffic_key = None
            self.cpu_desc = "oracle fp32 = reflect pad + torch.stft(1920, 480) + sqrt + dense mel matmul + ln (matcha mel_spectrogram's ops)"
        elif key == "cfg4":
            self.B, self.L, self.sr, self.n_mels = 4096, 80000, 16000, 80
            self.T = 1 + (self.L - 400) // 160
            self.out_shape = (self.B, self.T, self.n_mels)
            self.kernel = "fe_kaldi3_kernel"
            self.cpu_clips = 256
            self.desc = "BASELINE configs[3]: kaldi 80-dim fbank (25 ms / 10 ms, dither 0) + CMN, 4096 x 5 s @ 16 kHz per GPU"
            self.featurizer = "kaldi_fbank"
            self.traffic_key = None
            self.cpu_desc = "torchaudio.compliance.kaldi.fbank per clip + mean subtraction, as the reference calls it"
        elif key == "w2l":
            # SURVEY 8(f)3 (not a BASELINE config): measured inside the default run's other_workloads block only
            self.B, self.L, self.sr, self.n_mels = 256, 160000, 16000, 80
            self.T = 1 + self.L // 200
            self.out_shape = (self.B, self.n_mels, self.T)
            self.kernel = "fe_wav2lip_kernel"
            self.cpu_clips = 64
            self.desc = ("SURVEY 8(f)3: Wav2Lip / Diff2Lip melspectrogram (Backend/diff2lip/audio/audio.py:45-51; pre-emphasis, "
                         "n_fft 800, hop 200, 80 bins 55-7600 Hz, dB, [-4, 4]), 256 x 10 s @ 16 kHz per GPU")
            self.featurizer = "wav2lip_mel"
            self.traffic_key = None
            self.cpu_desc = "oracle.wav2lip (the reference's numpy/scipy arithmetic, float64)"
        elif key == "gate":
            # SURVEY 8(f)3, second half (not a BASELINE config): other_workloads only
            self.B, self.L, self.sr, self.n_mels = 256, 480000, 16000, 0
            self.T = 1 + self.L // 256
            self.out_shape = (self.B, 256 * (self.L // 256))
            self.kernel = "fe_gate_stft_kernel + fe_gate_quantile_kernel + fe_gate_inverse_kernel + fe_gate_scale_kernel"
            self.cpu_clips = 16
            self.desc = ("SURVEY 8(f)3: AudioProcessor's spectral gate (Backend/services/audio_processor.py:321-358; STFT 1024/256, 10 % "
                         "quantile noise floor per bin, gain mask, iSTFT, peak normalisation), 256 x 30 s @ 16 kHz per GPU")
            self.featurizer = "spectral_gate"
            self.traffic_key = None
            self.cpu_desc = "oracle.gate (the reference's own torch lines)"
        elif key == "cfg5":
            # the manifest sweep: a FIXED total of 10 000 clips, contiguous shards per rank (strong scaling), each shard
            # featurized in sub-batches of <= 512 clips like the manifest runner does
            self.total_clips, self.sub = 10000, 512
            lo, hi = manifest.shard_rows(self.total_clips, rank, world)
            self.B, self.L, self.sr, self.n_mels = hi - lo, 960000, 16000, 80
            self.T = self.L // 160
            self.out_shape = (self.B, self.n_mels, self.T)
            self.kernel = "fe_whisper3_kernel<80, 0, 0>"
            self.cpu_clips = 32
            self.desc = ("BASELINE configs[4]: Batch-Processing manifest sweep, 10 000 x 60 s @ 16 kHz, Whisper 80-bin log-mel, "
                         "contiguous clip shards per GPU, sub-batches of 512 clips")
            self.featurizer = "whisper_log_mel"
            self.traffic_key = None
            self.cpu_desc = "oracle fp32 = torch.stft + dense mel matmul + log10/max-8 (the ops whisper.audio runs)"
        elif key in ("cfg2pcm", "cfg2pcmk"):
            # cfg2 as the audio really arrives -- 44.1 kHz stereo int16 PCM -> mono mix +
            # torchaudio-compatible resampling to 16 kHz (b200feat_resample) -> Whisper 128-bin log-mel, device-resident
            self.B, self.L, self.sr, self.n_mels = 1024, 480000, 16000, 128
            self.pcm_sr, self.pcm_L, self.pcm_ch = 44100, 1323000, 2
            self.T = self.L // 160
            self.out_shape = (self.B, self.n_mels, self.T)
            self.kernel = "fe_resample_kernel<1, 2, 1>"
            self.cpu_clips = 32
            self.desc = ("SURVEY 8(f)1 in front of BASELINE configs[1]: 1024 x 30 s of 44.1 kHz stereo int16 PCM -> mono + resample to "
                         "16 kHz -> Whisper large-v3 128-bin log-mel per GPU")
            self.featurizer = "condition_pcm + whisper_log_mel"
            self.traffic_key = "fe_resample_kernel_s16x2_dram_bytes_per_1024_clips"
            self.cpu_desc = ("mean of the channels + torchaudio.functional.resample(44100 -> 16000) + oracle fp32 log-mel "
                             "(torch.stft + dense mel matmul + log10/max-8): the reference's own ops")
            self.rs_kw = {}
            if key == "cfg2pcmk":
                # the filter the reference's AudioProcessor builds (Backend/services/audio_processor.py:296-303): kaiser window,
                # lowpass_filter_width 128, rolloff 0.95 -- 160 phases x 1185 taps, run as a 3 x TF32 GEMM on the tensor cores
                self.rs_kw = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="sinc_interp_kaiser", beta=14.769656459379492)
                self.kernel = "fe_resample_tc_kernel"
                self.desc = self.desc.replace("mono + resample", "mono + resample (the reference's kaiser / width-128 filter, tensor cores)")
                self.featurizer = "condition_pcm(kaiser 128) + whisper_log_mel"
                self.traffic_key = None
        else:
            raise ValueError(key)
        # algorithmic flops per audio-second (sparse mel, 2.5 N log2 N real-FFT model)
        self.flops_per_audio_s = {"cfg2": 1.05e6, "cfg2pcm": None, "cfg2pcmk": None, "cfg3": 2.7e6, "cfg4": 1.4e6, "cfg5": 1.04e6, "w2l": 1.6e6, "gate": None}[key]
        self.audio_s_per_step = self.B * self.L / self.sr              # this rank's share
        self.scaling = "strong" if key == "cfg5" else "weak"
        # whole-job audio per step: fixed total for the sweep, per-GPU work x N otherwise
        self.global_audio_s = (self.total_clips if key == "cfg5" else self.B * world) * self.L / self.sr
        self.e2e_clips = min(self.B, 1024)                              # clips of the shard the host-buffer leg streams per step
        if key in ("cfg2pcm", "cfg2pcmk") and world > 1:
            self.e2e_clips = 256                                        # 5.3 KB of pinned PCM per audio-ms: keep N ranks' pinned buffers small
        self.out_per_clip = self.out_shape[1] * (self.out_shape[2] if len(self.out_shape) > 2 else 1)
        self.bytes_per_clip = 4 * self.L + 4 * self.out_per_clip     # algorithmic: 4*(in) + 4*(out)
        if key in ("cfg2pcm", "cfg2pcmk"):                            # the roofline kernel here is the resampler: PCM in + waveform out
            self.bytes_per_clip = self.pcm_L * self.pcm_ch * 2 + 4 * self.L

    def config(self, n_gpus):
        tot_in, tot_out = self.B * self.L * 4 / 1e9, self.B * self.out_per_clip * 4 / 1e9
        return {"workload": self.desc, "featurizer": self.featurizer, "n_mels": self.n_mels, "clips_per_gpu": self.B,
                "samples_per_clip": self.L, "pcm": (f"{self.pcm_sr} Hz x {self.pcm_ch} ch int16, {self.pcm_L} frames per clip"
                                                    if self.key in ("cfg2pcm", "cfg2pcmk") else None), "global_clips": self.total_clips if self.key == "cfg5" else self.B * n_gpus,
                "norm": "per_clip" if self.key in ("cfg2", "cfg5") else None,
                "parallelism": f"clip-sharded x{n_gpus}, no collective",
                "l2_policy": f"inputs ({tot_in:.2f} GB) and outputs ({tot_out:.2f} GB) per step exceed the 126 MB L2; no flush needed"}

    def device_step(self, fe, x, out):
        if self.key in ("cfg2pcm", "cfg2pcmk"):
            # ONE library call: mono mix + resampler, then the log-mel kernels; the waveform lives in the call's workspace
            return lambda: fe.pcm_whisper_log_mel_batch(x, self.pcm_sr, self.sr, n_mels=self.n_mels, norm="per_clip", out=out, **self.rs_kw)
        if self.key == "cfg2":
            return lambda: fe.whisper_log_mel_batch(x, n_mels=self.n_mels, norm="per_clip", out=out)
        if self.key == "cfg5":
            spans = manifest.sub_batches(0, self.B, self.sub)

            def sweep():
                for lo, hi in spans:
                    fe.whisper_log_mel_batch(x[lo:hi], n_mels=self.n_mels, norm="per_clip", out=out[lo:hi])
            return sweep
        if self.key == "cfg3":
            return lambda: fe.cosy_mel_batch(x, out=out)
        if self.key == "w2l":
            return lambda: fe.wav2lip_mel_batch(x, out=out)
        if self.key == "gate":
            return lambda: fe.spectral_gate_batch(x, out=out)
        return lambda: fe.kaldi_fbank_batch(x, subtract_mean=True, out=out)

    def host_step(self, fe, xh, oh, dev):
        if self.key in ("cfg2pcm", "cfg2pcmk"):
            # pinned int16 PCM in -> pinned float32 features out through the library's chunked pipeline (b200feat_pcm_whisper_logmel_host)
            return lambda: fe.pcm_whisper_log_mel_host(xh, self.pcm_sr, self.sr, n_mels=self.n_mels, norm="per_clip", out=oh, device=dev, **self.rs_kw)
        if self.key in ("cfg2", "cfg5"):
            return lambda: fe.whisper_log_mel_host(xh, n_mels=self.n_mels, norm="per_clip", out=oh, device=dev)
        if self.key == "cfg3":
            return lambda: fe.cosy_mel_host(xh, out=oh, device=dev)
        return lambda: fe.kaldi_fbank_host(xh, subtract_mean=True, out=oh, device=dev)

    def cpu_step(self):
        import oracle
        g = torch.Generator().manual_seed(1234)
        if self.key == "cfg2pcm":
            from oracle import resample as orc
            pcm = torch.randint(-20000, 20000, (self.cpu_clips, self.pcm_L, self.pcm_ch), generator=g, dtype=torch.int16)

            def ref_chain():
                mono = (pcm.to(torch.float32) / 32768.0).mean(dim=2)
                return oracle.whisper_log_mel(orc.reference_resample(mono, self.pcm_sr, self.sr), self.n_mels, norm="per_clip")
            return ref_chain, self.cpu_clips * self.L / self.sr
        x = (0.1 * torch.randn(self.cpu_clips, self.L, generator=g)).clamp_(-1, 1)
        if self.key in ("cfg2", "cfg5"):
            return (lambda: oracle.whisper_log_mel(x, self.n_mels, norm="per_clip")), self.cpu_clips * self.L / self.sr
        if self.key == "cfg3":
            return (lambda: oracle.cosyvoice_mel(x)), self.cpu_clips * self.L / self.sr
        return (lambda: [oracle.kaldi_fbank(x[b]) for b in range(self.cpu_clips)]), self.cpu_clips * self.L / self.sr


WL = Workload("cfg2")


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.rows = []
        self.proc = None
        self.th = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.th = threading.Thread(target=self._read, daemon=True)
        self.th.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- reference arm
def reference_step_fn():
    """the reference's own CPU path for this featurizer on a bounded sample of the workload (WL.cpu_clips clips),
    all host threads."""
    return WL.cpu_step()


def time_cpu(step, reps, warm):
    for _ in range(warm):
        step()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        step()
        ts.append(time.perf_counter() - t0)
    return ts


def run_reference(args, rank, world):
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 for its workers; the reference arm is one process on rank 0 and may use
    # every host core
    try:
        torch.set_num_threads(len(os.sched_getaffinity(0)))
    except Exception:
        torch.set_num_threads(os.cpu_count() or 1)
    step, audio_s = reference_step_fn()
    ts = time_cpu(step, args.steps, args.warmup)
    total = sum(ts)
    value = audio_s * len(ts) / total
    cores = torch.get_num_threads()
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(ts), "higher_is_better": True, "scaling": WL.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": WL.config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{WL.cpu_clips} of the {WL.B} clips per step ({WL.L / WL.sr:.0f} s each), {WL.cpu_desc}, "
                                       f"{cores} torch threads of {os.cpu_count()} host cpus"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- helpers of our arm
FP32_PEAK_TFLOPS = 72.9      # measured on this pool's B200 with tools/ubench/f32x2.cu (FFMA and FFMA2 alike), round 1


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (driver-measured copy bandwidth)"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def speech_like_device(gen, B, L, sr, dev):
    """the generator of tests/conftest.py::speech_like on the device: 3 random sines x slow AM envelope + 1 % noise"""
    t = torch.arange(L, dtype=torch.float32, device=dev) / sr
    out = torch.empty(B, L, dtype=torch.float32, device=dev)
    for lo in range(0, B, 64):
        hi = min(lo + 64, B)
        n = hi - lo
        f = 80.0 + 3000.0 * torch.rand(n, 3, 1, generator=gen, device=dev)
        ph = 6.283185307179586 * torch.rand(n, 3, 1, generator=gen, device=dev)
        fe_ = 0.5 + 2.5 * torch.rand(n, 1, generator=gen, device=dev)
        env = 0.5 * (1.0 + torch.sin(6.283185307179586 * fe_ * t))
        sig = torch.sin(6.283185307179586 * f * t + ph).sum(dim=1) / 3.0
        out[lo:hi] = (0.8 * env * sig + 0.01 * torch.randn(n, L, generator=gen, device=dev)).clamp_(-1, 1)
    return out


def measure_light(fe, wl, dev, rank, world, steps, barrier, max_over_ranks):
    """one of the other workloads inside the default run: device-resident step time + the featurizer kernel's share,
    same contract as the headline (CUDA events, >= 3 warm-ups, barrier + synchronize, MAX over ranks)."""
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    if wl.key == "cfg5":
        x = torch.empty(wl.B, wl.L, dtype=torch.float32, device=dev)
        for lo, hi in manifest.sub_batches(0, wl.B, 256):
            x[lo:hi].normal_(generator=gen).mul_(0.1).clamp_(-1, 1)
    elif wl.key in ("cfg2pcm", "cfg2pcmk"):
        x = torch.randint(-20000, 20000, (wl.B, wl.pcm_L, wl.pcm_ch), generator=gen, dtype=torch.int16, device=dev)
    else:
        x = (0.1 * torch.randn(wl.B, wl.L, generator=gen, device=dev)).clamp_(-1, 1)
    out = torch.empty(wl.out_shape, dtype=torch.float32, device=dev)
    step = wl.device_step(fe, x, out)
    for _ in range(3):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / steps
    fe.profile_begin()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    prof = fe.profile_end()
    del x, out
    torch.cuda.empty_cache()
    peak, _ = hbm_peak()
    k_ms_step = prof["feature_ms"] / steps                       # all featurizer launches of one step
    bytes_step = wl.B * (4 * wl.L + 4 * wl.out_per_clip)        # this rank's algorithmic bytes (whole step: featurizer rows only)
    res = {"workload": wl.desc, "scaling": wl.scaling, "ms_per_step": ms_step, "value": wl.global_audio_s / (ms_step * 1e-3), "unit": UNIT,
           "steps": steps, "kernel": wl.kernel, "kernel_ms_per_step": k_ms_step, "epilogue_ms_per_step": prof["epilogue_ms"] / steps,
           "launches_per_step": (prof["feature_launches"] + prof["epilogue_launches"]) // steps}
    if wl.key == "gate":
        res["roofline"] = {"bound": "hbm", "note": "four kernels; algorithmic bytes = waveform in + waveform out; the complex spectrum (513 x T x 8 "
                                                   "bytes per clip, written once and read three times) is intermediate traffic",
                           "frac_whole_step": bytes_step / 1e9 / (ms_step * 1e-3) / peak}
    elif wl.key in ("cfg2pcm", "cfg2pcmk"):
        chain_bytes = wl.B * (wl.pcm_L * wl.pcm_ch * 2 + 4 * wl.out_per_clip)
        res["roofline"] = {"bound": "hbm", "note": "whole chain (mono mix + resampler, then log-mel) against the bytes a fully fused chain would "
                                                   "move: int16 PCM in + float32 features out; the float32 waveform in between is extra traffic",
                           "frac_whole_step": chain_bytes / 1e9 / (ms_step * 1e-3) / peak}
    else:
        res["roofline"] = {"bound": "hbm" if wl.key not in ("cfg3", "w2l") else "fp32", "frac": bytes_step / 1e9 / (k_ms_step * 1e-3) / peak,
                           "frac_whole_step": bytes_step / 1e9 / (ms_step * 1e-3) / peak if world == 1 or wl.scaling == "weak" else None,
                           "fp32_frac": wl.flops_per_audio_s * wl.audio_s_per_step / (k_ms_step * 1e-3) / 1e12 / FP32_PEAK_TFLOPS}
    return res


def library_baseline(wl, x, dev, steps=3):
    """the eager torch-CUDA lowering of the same featurizer on the same box (cuFFT + cuBLAS + ATen
    elementwise kernels), CUDA events, on a slice of the batch; the "library baseline", never a product path."""
    nb = min(wl.B, 256)
    xs = x[:nb]
    import torchaudio.functional as taf
    filt = taf.melscale_fbanks(201, 0.0, 8000.0, wl.n_mels, 16000, norm="slaney", mel_scale="slaney").t().contiguous().to(dev)
    win = torch.hann_window(400, device=dev)

    def step():
        st = torch.stft(xs, 400, 160, window=win, return_complex=True)
        mag = st[..., :-1].abs() ** 2
        ls = torch.clamp(filt @ mag, min=1e-10).log10()
        ls = torch.maximum(ls, ls.amax(dim=(1, 2), keepdim=True) - 8.0)
        return (ls + 4.0) / 4.0
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    torch.cuda.empty_cache()
    return {"value": nb * wl.L / wl.sr / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "clips": nb,
            "what": "eager PyTorch on the same GPU: torch.stft (cuFFT) + abs^2 + dense mel matmul (cuBLAS) + clamp/log10/max-8/(x+4)/4 "
                    "(ATen elementwise), per-clip max; the library lowering of whisper.audio.log_mel_spectrogram"}


# --------------------------------------------------------------------------------------------- our arm
def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    import expressive_speech_translation_b200 as fe

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    if WL.key == "cfg5":
        x = torch.empty(WL.B, WL.L, dtype=torch.float32, device=dev)    # 38.4 GB at N = 1: filled in place, 256 clips at a time
        for lo, hi in manifest.sub_batches(0, WL.B, 256):
            x[lo:hi].normal_(generator=gen).mul_(0.1).clamp_(-1, 1)
    elif WL.key == "cfg2pcm":
        x = torch.randint(-20000, 20000, (WL.B, WL.pcm_L, WL.pcm_ch), generator=gen, dtype=torch.int16, device=dev)
    else:
        x = (0.1 * torch.randn(WL.B, WL.L, generator=gen, device=dev)).clamp_(-1, 1)
    out = torch.empty(WL.out_shape, dtype=torch.float32, device=dev)
    step = WL.device_step(fe, x, out)

    # ---- device-resident timing
    for _ in range(args.warmup):
        step()
    barrier()
    clocks = ClockSampler(local_rank)
    clocks.start()
    n0 = fe.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = fe.launch_count() - n0
    clk = clocks.stop()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    ms_step = ms_total / args.steps
    value = WL.global_audio_s / (ms_step * 1e-3)

    # ---- the featurizer kernel alone (per-launch CUDA events inside the library), rank 0 reports
    fe.profile_begin()
    if WL.key == "cfg2pcm":
        wav_r = torch.empty((WL.B, WL.L), dtype=torch.float32, device=dev)
        for _ in range(args.steps):
            fe.condition_pcm(x, WL.pcm_sr, WL.sr, out=wav_r)          # the row's own kernel; the log-mel kernel is cfg2's roofline
        del wav_r
    else:
        for _ in range(args.steps):
            step()
    torch.cuda.synchronize()
    prof = fe.profile_end()
    k_ms = prof["feature_ms"] / max(prof["feature_launches"], 1)
    clips_per_launch = WL.B * args.steps / max(prof["feature_launches"], 1)
    achieved = clips_per_launch * WL.bytes_per_clip / (k_ms * 1e-3) / 1e9
    peak, peak_src = hbm_peak()
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            traffic = json.load(f).get(WL.traffic_key) if WL.traffic_key else None
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": WL.kernel, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms_per_launch": k_ms, "clips_per_launch": clips_per_launch,
                "algorithmic_bytes_per_launch": clips_per_launch * WL.bytes_per_clip,
                "kernel_share_of_step": prof["feature_ms"] / max(prof["feature_ms"] + prof["epilogue_ms"], 1e-9),
                "epilogue_ms_per_step": prof["epilogue_ms"] / args.steps,
                "step_frac": (WL.B * WL.bytes_per_clip / (ms_step / 1.0 * 1e-3) / 1e9) / peak if world == 1 else None}
    feat_epi_ms = (prof["feature_ms"] + prof["epilogue_ms"]) / args.steps
    # the same bytes over the featurizer kernel PLUS its second pass (Whisper: the clamp fix-up kernel, which rewrites only the
    # 128-frame units the per-clip floor max-8 reaches -- none on white noise, see realistic_leg; kaldi: the CMN kernel)
    roofline["frac_with_epilogue"] = WL.B * WL.bytes_per_clip / 1e9 / (feat_epi_ms * 1e-3) / peak
    roofline["epilogue_share_of_kernels"] = prof["epilogue_ms"] / max(prof["feature_ms"] + prof["epilogue_ms"], 1e-9)
    if WL.flops_per_audio_s:
        tf = WL.flops_per_audio_s * clips_per_launch * WL.L / WL.sr / (k_ms * 1e-3) / 1e12
        roofline["fp32_frac"] = tf / FP32_PEAK_TFLOPS
        roofline["fp32"] = {"achieved_tflops": tf, "peak_tflops": FP32_PEAK_TFLOPS, "peak_source": "tools/ubench/f32x2.cu on this pool's B200 (round 1)",
                            "flops_model": f"{WL.flops_per_audio_s:.3g} flop per audio-second (SURVEY.md 8(d): real FFT 2.5 N log2 N + sparse mel)"}

    if WL.key == "cfg2pcm":
        roofline["kernel_share_of_step"] = k_ms / ms_step
        roofline["note"] = ("the row's own kernel (mono mix + resampler) timed alone; the rest of the step is cfg2's log-mel kernel, "
                            "whose roofline is the default workload's line")
        roofline["step_frac"] = None

    # ---- device-resident step with the narrow dtypes (int16 PCM in, float16 out), Whisper workloads only
    narrow_dev_ms = None
    if WL.featurizer == "whisper_log_mel":
        nb = min(WL.B, 1024)
        x16 = (x[:nb] * 32767.0).round_().to(torch.int16)
        o16 = torch.empty((nb,) + tuple(WL.out_shape[1:]), dtype=torch.float16, device=dev)
        for _ in range(3):
            fe.whisper_log_mel_batch(x16, n_mels=WL.n_mels, norm="per_clip", out=o16)
        e0.record()
        for _ in range(args.steps):
            fe.whisper_log_mel_batch(x16, n_mels=WL.n_mels, norm="per_clip", out=o16)
        e1.record()
        torch.cuda.synchronize()
        narrow_dev_ms = e0.elapsed_time(e1) / args.steps
        narrow_dev_clips = nb
        del x16, o16

    # ---- realistic leg (N = 1, headline workload): speech-like, ragged, zero-padded clips the way whisper.transcribe calls the
    # path (padding = 480000).  Here the per-clip floor max-8 does reach frames (the zero tail, pauses), so the fix-up kernel
    # has real work; its share is reported, nothing is hidden in the white-noise headline.
    realistic = None
    library = None
    if WL.key == "cfg2" and world == 1 and not args.no_extras:
        nb = 512
        g2 = torch.Generator(device=dev).manual_seed(99)
        xs_ = speech_like_device(g2, nb, WL.L, WL.sr, dev)
        lens = torch.randint(5 * WL.sr, WL.L + 1, (nb,), generator=g2, device=dev, dtype=torch.int64)
        Tpad = (WL.L + 480000) // 160
        o2 = torch.empty((nb, WL.n_mels, Tpad), dtype=torch.float32, device=dev)
        rstep = lambda: fe.whisper_log_mel_batch(xs_, lengths=lens, n_mels=WL.n_mels, padding=480000, norm="per_clip", out=o2)
        for _ in range(3):
            rstep()
        e0.record()
        for _ in range(args.steps):
            rstep()
        e1.record()
        torch.cuda.synchronize()
        r_ms = e0.elapsed_time(e1) / args.steps
        fe.profile_begin()
        for _ in range(args.steps):
            rstep()
        torch.cuda.synchronize()
        rp = fe.profile_end()
        valid = int(lens.sum().item())
        r_bytes = 4 * valid + 4 * o2.numel()
        realistic = {"what": f"{nb} speech-like clips (3 sines x AM envelope + 1 % noise), ragged 5-30 s, padding=480000 "
                             f"(whisper.transcribe's call): output (B, {WL.n_mels}, {Tpad}) incl. the zero-padded tail",
                     "ms_per_step": r_ms, "value": valid / WL.sr / (r_ms * 1e-3), "unit": UNIT,
                     "feature_ms_per_step": rp["feature_ms"] / args.steps, "fixup_ms_per_step": rp["epilogue_ms"] / args.steps,
                     "fixup_share": rp["epilogue_ms"] / max(rp["feature_ms"] + rp["epilogue_ms"], 1e-9),
                     "algorithmic_bytes_per_step": r_bytes,
                     "frac_feature_plus_fixup": r_bytes / 1e9 / ((rp["feature_ms"] + rp["epilogue_ms"]) / args.steps * 1e-3) / peak}
        del xs_, o2, lens
        torch.cuda.empty_cache()
        library = library_baseline(WL, x, dev)

    # ---- end to end: pinned host in -> pinned host out through the host-buffer API
    # (one process per GPU: keep this rank's host buffers on the NUMA node next to its GPU)
    bound = None if os.environ.get("B200FEAT_NO_NUMA_BIND") else manifest.bind_host_to_gpu(local_rank)
    del out
    ne = WL.e2e_clips
    if WL.key == "cfg2pcm":
        xh = torch.empty((ne, WL.pcm_L, WL.pcm_ch), dtype=torch.int16, pin_memory=True)
    else:
        xh = torch.empty((ne, WL.L), dtype=torch.float32, pin_memory=True)
    xh.copy_(x[:ne])
    del x
    oh = torch.empty((ne,) + tuple(WL.out_shape[1:]), dtype=torch.float32, pin_memory=True)
    host_step = WL.host_step(fe, xh, oh, dev)

    def e2e_step():
        host_step()
        return float(oh[0, 0, 0])                                  # read the step's result on the host

    e2e_steps = max(2, min(args.steps, 5))
    e2e_step()
    barrier()
    alone_ms = None
    if world > 1:
        # rank 0 streams alone while the others idle: the single-GPU e2e figure of THIS box, the denominator of e2e_efficiency
        if rank == 0:
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                e2e_step()
            alone_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / e2e_steps
    e2e = {"value": world * ne * WL.L / WL.sr / (e2e_ms * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": xh.numel() * xh.element_size(), "d2h_bytes_per_step": ne * WL.out_per_clip * 4,
           "ms_per_step": e2e_ms, "steps": e2e_steps, "clips_per_step_per_gpu": ne, "host_cpus_bound": len(bound) if bound else None,
           "api": (f"{WL.featurizer}_host -> b200feat_*_host (pinned host buffers, chunked H2D/kernel/D2H pipeline, blocking)"
                   if WL.key != "cfg2pcm" else
                   "pcm_whisper_log_mel_host -> b200feat_pcm_whisper_logmel_host (pinned int16 PCM in, chunked H2D / resample / log-mel / D2H, blocking)")}

    if alone_ms is not None:
        alone = ne * WL.L / WL.sr / (alone_ms * 1e-3)
        e2e["rank0_alone"] = {"value": alone, "unit": UNIT, "ms_per_step": alone_ms}
        e2e["e2e_efficiency"] = e2e["value"] / (world * alone)
        e2e["limiter"] = ("host side: all ranks' pinned buffers stream through the same host DRAM and PCIe root complexes; the device-"
                          "resident `value` scales with N, the host-buffer figure scales with host memory bandwidth")

    # ---- the same leg with the narrow dtypes either side of the path: int16 PCM in, float16
    # features out -- half the PCIe bytes.  Reported next to e2e, never instead of it (the reference API is float32).
    if WL.featurizer == "whisper_log_mel":
        xs = torch.empty((ne, WL.L), dtype=torch.int16, pin_memory=True)
        for lo, hi in manifest.sub_batches(0, ne, 64):
            xs[lo:hi].copy_((xh[lo:hi] * 32767.0).round_().to(torch.int16))
        oh16 = torch.empty((ne,) + tuple(WL.out_shape[1:]), dtype=torch.float16, pin_memory=True)

        def narrow_step():
            fe.whisper_log_mel_host(xs, n_mels=WL.n_mels, norm="per_clip", out=oh16, device=dev)
            return float(oh16[0, 0, 0])

        narrow_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            narrow_step()
        barrier()
        n_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / e2e_steps
        e2e["pcm16_in_f16_out"] = {"value": world * ne * WL.L / WL.sr / (n_ms * 1e-3), "unit": UNIT, "ms_per_step": n_ms,
                                   "h2d_bytes_per_step": ne * WL.L * 2, "d2h_bytes_per_step": ne * WL.out_per_clip * 2,
                                   "api": "whisper_log_mel_host(int16 wave, out float16) -> b200feat_whisper_logmel_ex_host",
                                   "device_resident_ms_per_step": narrow_dev_ms, "device_resident_clips": narrow_dev_clips}
        del xs, oh16

    # ---- CPU baseline beside it (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1:
        stepc, audio_s = reference_step_fn()
        ts = time_cpu(stepc, 3, 1)
        cores = torch.get_num_threads()
        cpu = {"value": audio_s / min(ts), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{WL.cpu_clips} of the {WL.B} clips ({WL.L / WL.sr:.0f} s each), best of 3 after 1 warm-up, {WL.cpu_desc}, "
                         f"{cores} torch threads of {os.cpu_count()} host cpus"}

    cpu1 = None
    if rank == 0 and world == 1 and not args.no_extras:
        import oracle
        nt = torch.get_num_threads()
        torch.set_num_threads(1)
        try:
            g1 = torch.Generator().manual_seed(1234)
            n1 = 8
            x1 = (0.1 * torch.randn(n1, WL.L, generator=g1)).clamp_(-1, 1)
            if WL.key in ("cfg2", "cfg5", "cfg2pcm"):
                f1 = lambda: oracle.whisper_log_mel(x1, WL.n_mels, norm="per_clip")
            elif WL.key == "cfg3":
                f1 = lambda: oracle.cosyvoice_mel(x1)
            else:
                f1 = lambda: [oracle.kaldi_fbank(x1[b]) for b in range(n1)]
            ts1 = time_cpu(f1, 2, 1)
            cpu1 = {"value": n1 * WL.L / WL.sr / min(ts1), "unit": UNIT, "cores": 1, "kind": "port",
                    "sample": f"{n1} clips ({WL.L / WL.sr:.0f} s each), best of 2 after 1 warm-up, torch.set_num_threads(1)"}
        finally:
            torch.set_num_threads(nt)

    # ---- the other workloads of SURVEY.md section 8(a)/(f), same run, same clock: device-resident, reduced steps
    others = None
    if WL.key == "cfg2" and not args.no_extras:
        del xh, oh
        torch.cuda.empty_cache()
        others = {}
        for key in ("cfg3", "cfg4", "cfg2pcm", "cfg2pcmk", "w2l", "gate", "cfg5"):
            wl = Workload(key, rank, world)
            others[key] = measure_light(fe, wl, dev, rank, world, 3 if key == "cfg5" else min(args.steps, 5), barrier, max_over_ranks)

    if world > 1:
        counts = [None] * world
        dist.all_gather_object(counts, {"rank": rank, "clips": WL.B})
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": WL.scaling, "vs_baseline": None, "dtype": "f32",
                "data": ("synthetic (0.1*N(0,1) clipped to [-1,1], seed 1234+rank, generated on device)" if WL.key != "cfg2pcm" else
                         "synthetic (uniform int16 in [-20000, 20000), stereo, seed 1234+rank, generated on device)"),
                "config": WL.config(world), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": launches, "clocks": clk}
        if realistic is not None:
            line["realistic_leg"] = realistic
        if library is not None:
            line["library_baseline"] = library
        if cpu1 is not None:
            line["cpu_baseline_1thread"] = cpu1
        if others is not None:
            line["other_workloads"] = others
        if world > 1:
            line["config"]["clips_per_rank"] = [c["clips"] for c in sorted(counts, key=lambda c: c["rank"])]
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg4", "cfg5", "cfg2pcm"],
                    help="BASELINE.json config: cfg2 = Whisper 128-bin (the headline; default), cfg3 = CosyVoice mel, cfg4 = kaldi fbank, "
                         "cfg5 = manifest sweep (10 000 x 60 s, Whisper 80-bin, strong scaling), cfg2pcm = cfg2 fed with 44.1 kHz stereo "
                         "int16 PCM through the fused mono-mix + resampler (SURVEY 8(f) item 1)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the legs beside the contract's line (realistic leg, library baseline, 1-thread CPU, other workloads)")
    args = ap.parse_args()
    global WL
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    WL = Workload(args.workload, rank if world == args.gpus else 0, world if world == args.gpus else 1)
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"),
                   os.path.abspath(__file__), "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup),
                   "--workload", args.workload] + (["--no-extras"] if args.no_extras else [])
            sys.exit(subprocess.call(cmd))
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
