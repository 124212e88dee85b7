#!/usr/bin/env python
"""Benchmark of the hot path: batched waveform -> Whisper large-v3 128-bin log-mel (BASELINE.json configs[1]:
1024 synthetic 30-s 16 kHz chunks per GPU).  One step = one pass over the whole batch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

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

METRIC = "audio_seconds_per_second_log_mel"
UNIT = "audio-s/s"
B_CLIPS, L_SAMPLES, N_MELS, SR = 1024, 480000, 128, 16000
T_FRAMES = L_SAMPLES // 160
AUDIO_S_PER_STEP = B_CLIPS * L_SAMPLES / SR                       # 30720 audio-seconds per GPU per step
BYTES_PER_CLIP = 4 * L_SAMPLES + 4 * N_MELS * T_FRAMES            # algorithmic: 4*(in) + 4*(out) = 3 456 000 B
CPU_SAMPLE_CLIPS = 64


def config(n_gpus):
    return {"workload": "BASELINE configs[1]: Whisper large-v3 128-bin log-mel, 1024 x 30 s @ 16 kHz per GPU",
            "featurizer": "whisper_log_mel", "n_mels": N_MELS, "clips_per_gpu": B_CLIPS, "samples_per_clip": L_SAMPLES,
            "global_clips": B_CLIPS * n_gpus, "norm": "per_clip", "parallelism": f"clip-sharded x{n_gpus}, no collective",
            "l2_policy": "inputs (1.97 GB) and outputs (1.57 GB) per step exceed the 126 MB L2; no flush needed"}


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
    """the reference's own CPU path for this featurizer (whisper.audio.log_mel_spectrogram's ops) on a bounded
    sample of the workload: CPU_SAMPLE_CLIPS clips of 30 s, all host threads."""
    import oracle
    g = torch.Generator().manual_seed(1234)
    x = (0.1 * torch.randn(CPU_SAMPLE_CLIPS, L_SAMPLES, generator=g)).clamp_(-1, 1)

    def step():
        return oracle.whisper_log_mel(x, N_MELS, norm="per_clip")
    return step, CPU_SAMPLE_CLIPS * L_SAMPLES / SR


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
    step, audio_s = reference_step_fn()
    ts = time_cpu(step, args.steps, args.warmup)
    total = sum(ts)
    value = audio_s * len(ts) / total
    cores = torch.get_num_threads()
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(ts), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{CPU_SAMPLE_CLIPS} of the 1024 clips per step (30 s each), oracle fp32 = "
                                       "torch.stft + dense mel matmul + log10/max-8 (the ops whisper.audio runs), "
                                       f"{cores} torch threads of {os.cpu_count()} host cpus"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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
    x = (0.1 * torch.randn(B_CLIPS, L_SAMPLES, generator=gen, device=dev)).clamp_(-1, 1)
    out = torch.empty((B_CLIPS, N_MELS, T_FRAMES), dtype=torch.float32, device=dev)

    def step():
        fe.whisper_log_mel_batch(x, n_mels=N_MELS, norm="per_clip", out=out)

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
    value = world * AUDIO_S_PER_STEP / (ms_step * 1e-3)

    # ---- the featurizer kernel alone (per-launch CUDA events inside the library), rank 0 reports
    fe.profile_begin()
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize()
    prof = fe.profile_end()
    k_ms = prof["feature_ms"] / max(prof["feature_launches"], 1)
    clips_per_launch = B_CLIPS * args.steps / max(prof["feature_launches"], 1)
    achieved = clips_per_launch * BYTES_PER_CLIP / (k_ms * 1e-3) / 1e9
    peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["hbm_gbs"])
            peak_src = "MEASURED_PEAKS.json hbm_gbs (driver-measured copy bandwidth)"
    except Exception:
        pass
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            traffic = json.load(f).get("fe_whisper2_kernel_128_dram_bytes_per_1024_clips")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "fe_whisper2_kernel<128>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms_per_launch": k_ms, "clips_per_launch": clips_per_launch,
                "algorithmic_bytes_per_launch": clips_per_launch * BYTES_PER_CLIP,
                "kernel_share_of_step": prof["feature_ms"] / max(prof["feature_ms"] + prof["epilogue_ms"], 1e-9),
                "epilogue_ms_per_step": prof["epilogue_ms"] / args.steps,
                "step_frac": (B_CLIPS * BYTES_PER_CLIP / (ms_step / 1.0 * 1e-3) / 1e9) / peak if world == 1 else None}

    # ---- end to end: pinned host in -> pinned host out through the host-buffer API
    del out
    xh = torch.empty((B_CLIPS, L_SAMPLES), dtype=torch.float32, pin_memory=True)
    xh.copy_(x)
    del x
    oh = torch.empty((B_CLIPS, N_MELS, T_FRAMES), dtype=torch.float32, pin_memory=True)

    def e2e_step():
        fe.whisper_log_mel_host(xh, n_mels=N_MELS, norm="per_clip", out=oh, device=dev)
        return float(oh[0, 0, 0])                                  # read the step's result on the host

    e2e_steps = max(2, min(args.steps, 5))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / e2e_steps
    e2e = {"value": world * AUDIO_S_PER_STEP / (e2e_ms * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": B_CLIPS * L_SAMPLES * 4, "d2h_bytes_per_step": B_CLIPS * N_MELS * T_FRAMES * 4,
           "ms_per_step": e2e_ms, "steps": e2e_steps, "api": "whisper_log_mel_host -> b200feat_whisper_logmel_host "
           "(pinned host buffers, chunked H2D/kernel/D2H pipeline, blocking)"}

    # ---- CPU baseline beside it (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1:
        stepc, audio_s = reference_step_fn()
        ts = time_cpu(stepc, 3, 1)
        cores = torch.get_num_threads()
        cpu = {"value": audio_s / min(ts), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{CPU_SAMPLE_CLIPS} of the 1024 clips (30 s each), best of 3 after 1 warm-up, oracle fp32 "
                         f"(torch.stft + dense mel matmul + log10/max-8), {cores} torch threads of {os.cpu_count()} host cpus"}

    if world > 1:
        counts = [None] * world
        dist.all_gather_object(counts, {"rank": rank, "clips": B_CLIPS * args.steps})
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic (0.1*N(0,1) clipped to [-1,1], seed 1234+rank, generated on device)",
                "config": config(world), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": launches, "clocks": clk}
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"),
                   os.path.abspath(__file__), "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup)]
            sys.exit(subprocess.call(cmd))
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
