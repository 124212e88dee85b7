"""Manifest sweep: the Batch-Processing runner's featurization stage, one process per GPU.

    python -m expressive_speech_translation_b200.sweep <manifest.csv> <out_dir> [--featurizer whisper80|whisper128|cosy|kaldi]
    torchrun --nproc-per-node N -m expressive_speech_translation_b200.sweep <manifest.csv> <out_dir> ...

What it mirrors in the reference:
  * the manifest is the CSV /root/reference/Batch-Processing/create_batch_manifest.py:52-56 writes
    (header `job_id,dataset,source_video_path`, one row per job);
  * the reference runs ONE row per SLURM array task (/root/reference/Batch-Processing/submit_slurm_job.sh:41,53);
    here every rank takes a CONTIGUOUS block of rows (manifest.shard_rows) -- clips are independent, there is no
    data-path collective, only one all-gather of (clips, frames, samples) for the report;
  * per-job outputs are named after the job id, and a failed job leaves `<job_id>.error` holding
    "Error processing <path>:\n\n<message>" (/root/reference/Batch-Processing/run_batch_job.py:81-85);
  * re-running a sweep skips every job whose output already exists (idempotent resume) and retries the failed ones.

Inputs: the image has no ffmpeg / torchcodec, so sources must be RIFF/WAVE PCM16 (mono or stereo, any rate; Python's
`wave`).  Clips whose rate is not the featurizer's are conditioned on the GPU with the reference's own filter
(torchaudio Resample kaiser, width 128, rolloff 0.95: /root/reference/Backend/services/audio_processor.py:296-303) and
stereo is mixed down to mono there as well.  Outputs: `<out_dir>/<job_id>.npy` (or `.safetensors`, key "features"):
the feature cache the ASR / TTS stages stream from.
"""
import argparse
import json
import os
import time
import wave as _wave

import numpy as np
import torch

from . import manifest as _manifest

FEATURIZERS = {
    # name: (sample rate, kind, n_mels)
    "whisper80": (16000, "whisper", 80),
    "whisper128": (16000, "whisper", 128),
    "cosy": (24000, "cosy", 80),
    "kaldi": (16000, "kaldi", 80),
}
REF_RESAMPLE = dict(lowpass_filter_width=128, rolloff=0.95, resampling_method="kaiser_window", beta=14.769656459379492)


def write_manifest(path, rows):
    """rows: iterable of (job_id, dataset, source_path) -> CSV with the reference's header
    (/root/reference/Batch-Processing/create_batch_manifest.py:52-56)."""
    import csv
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["job_id", "dataset", "source_video_path"])
        for r in rows:
            w.writerow(list(r))


def probe_wav(path):
    """-> (rate, channels, frames) from the header; anything but 16-bit PCM mono / stereo raises ValueError."""
    try:
        with _wave.open(path, "rb") as f:
            ch, width, rate, n = f.getnchannels(), f.getsampwidth(), f.getframerate(), f.getnframes()
            if width != 2 or ch not in (1, 2) or f.getcomptype() != "NONE":
                raise ValueError(f"only 16-bit PCM mono / stereo .wav is readable here (channels {ch}, sample width {width})")
    except (_wave.Error, EOFError) as e:
        raise ValueError(f"not a readable RIFF/WAVE file: {e}") from None
    if n <= 0:
        raise ValueError("empty sample data")
    return rate, ch, n


def read_wav_pcm16(path):
    """-> (rate, int16 array (frames, channels)); a truncated data chunk raises ValueError."""
    rate, ch, n = probe_wav(path)
    with _wave.open(path, "rb") as f:
        data = f.readframes(n)
    pcm = np.frombuffer(data, dtype="<i2")
    if pcm.size != n * ch:
        raise ValueError(f"truncated sample data: header says {n} frames, file holds {pcm.size // ch}")
    return rate, pcm.reshape(-1, ch)


def output_path(out_dir, job_id, fmt):
    return os.path.join(out_dir, f"{job_id}.{fmt}")


def save_features(path, feat, fmt):
    """atomic: a partially written cache entry must never look like a finished job to the next run"""
    tmp = path + ".tmp"
    if fmt == "npy":
        with open(tmp, "wb") as f:
            np.save(f, feat)
    elif fmt == "safetensors":
        from safetensors.numpy import save_file
        save_file({"features": np.ascontiguousarray(feat)}, tmp)
    else:
        raise ValueError(f"unknown feature format {fmt!r}")
    os.replace(tmp, path)


def load_features(path):
    if path.endswith(".npy"):
        return np.load(path)
    from safetensors.numpy import load_file
    return load_file(path)["features"]


class GpuFeaturizer:
    """batches of PCM16 clips -> per-clip feature arrays through the CUDA library (no CPU path)"""

    def __init__(self, name, device=None, out_dtype=torch.float32, padding=0):
        from . import api
        self.api = api
        self.name = name
        self.rate, self.kind, self.n_mels = FEATURIZERS[name]
        self.device = api._cuda_device(device)
        self.out_dtype = out_dtype
        self.padding = int(padding)

    def __call__(self, clips):
        """clips: list of (rate, int16 (frames, channels)) that share rate and channel count.
        -> list of numpy feature arrays, one per clip (whisper / cosy: (n_mels, T_i); kaldi: (T_i, 80))."""
        api = self.api
        rate, ch = clips[0][0], clips[0][1].shape[1]
        n = [c[1].shape[0] for c in clips]
        L = max(n)
        host = torch.zeros((len(clips), L, ch), dtype=torch.int16, pin_memory=True)
        for i, (_, pcm) in enumerate(clips):
            host[i, :n[i]] = torch.from_numpy(pcm)
        dev = host.to(self.device, non_blocking=True)
        lengths = torch.tensor(n, dtype=torch.int64)
        if rate != self.rate:
            # conditioning row: int16 interleaved PCM -> mono float32 at the featurizer's rate, ONE kernel, with the
            # reference's filter (audio_processor.py:296-303)
            wave = api.condition_pcm(dev, rate, self.rate, lengths=lengths, **REF_RESAMPLE)
            lens = [api.resample_out_len(v, rate, self.rate) for v in n]
        elif ch == 1 and self.kind == "whisper":
            wave, lens = dev[:, :, 0], n      # int16 mono at the right rate: the Whisper loader converts in-kernel
        else:
            # right rate already: only the int16 -> float32 scale (load_audio's / 32768) and the channel mean are left
            wave, lens = dev.to(torch.float32).mul_(1.0 / 32768.0).mean(dim=2), n
        ln = torch.tensor(lens, dtype=torch.int64)
        if self.kind == "whisper":
            out = api.whisper_log_mel_batch(wave, lengths=ln, n_mels=self.n_mels, padding=self.padding, norm="per_clip",
                                            out_dtype=self.out_dtype)
            T = [(v + self.padding) // 160 for v in lens]
            res = out.cpu().numpy()
            return [res[i, :, :T[i]] for i in range(len(clips))]
        if self.kind == "cosy":
            out = api.cosy_mel_batch(wave, lengths=ln).cpu().numpy()
            return [out[i, :, :api.num_frames("cosy", lens[i])] for i in range(len(clips))]
        out = api.kaldi_fbank_batch(wave, lengths=ln, subtract_mean=True).cpu().numpy()
        return [out[i, :api.num_frames("kaldi", lens[i])] for i in range(len(clips))]


def _min_frames(kind):
    return {"whisper": 201, "cosy": 721, "kaldi": 400}[kind]


def run_sweep(manifest_path, out_dir, featurize, rate, kind, rank=0, world=1, fmt="npy", max_clips=512,
              max_batch_samples=64 << 20, log=None):
    """Featurizes this rank's rows of the manifest.  `featurize(clips) -> [feature arrays]` (GpuFeaturizer, or a
    stand-in for the CPU tests).  Returns dict(rows, done, skipped, failed, frames, samples, seconds)."""
    os.makedirs(out_dir, exist_ok=True)
    rows = _manifest.read_manifest(manifest_path)
    lo, hi = _manifest.shard_rows(len(rows), rank, world)
    t_start = time.time()
    stats = dict(rows=hi - lo, done=0, skipped=0, failed=0, frames=0, samples=0)
    pending = []                                     # (job_id, source, rate, channels, frames): headers only

    def fail(job_id, src, err):
        stats["failed"] += 1
        with open(os.path.join(out_dir, f"{job_id}.error"), "w") as f:
            f.write(f"Error processing {src}:\n\n{err}")
        if log:
            log(f"[FAILURE] job {job_id}: {err}")

    def flush(group):
        clips, live = [], []
        for g in group:                              # the samples are read only now: one batch resident at a time
            try:
                r, pcm = read_wav_pcm16(g[1])
                clips.append((r, pcm))
                live.append(g)
            except Exception as e:
                fail(g[0], g[1], e)
        if not live:
            return
        try:
            feats = featurize(clips)
        except Exception as e:                       # a batch-level failure is every member's failure
            for g in live:
                fail(g[0], g[1], e)
            return
        for g, clip, feat in zip(live, clips, feats):
            save_features(output_path(out_dir, g[0], fmt), feat, fmt)
            err = os.path.join(out_dir, f"{g[0]}.error")
            if os.path.exists(err):
                os.remove(err)                       # a retried job that now succeeded
            stats["done"] += 1
            stats["frames"] += int(feat.shape[0] if kind == "kaldi" else feat.shape[-1])
            stats["samples"] += int(clip[1].shape[0])

    for row in rows[lo:hi]:
        if len(row) < 3:
            fail(row[0] if row else "unknown", "", "manifest row needs job_id,dataset,source_video_path")
            continue
        job_id, _dataset, src = row[0], row[1], row[2]
        if os.path.exists(output_path(out_dir, job_id, fmt)):
            stats["skipped"] += 1
            continue
        try:
            r, ch, n = probe_wav(src)
            need = -(-_min_frames(kind) * r // rate)
            if n < need:
                raise ValueError(f"clip too short for the {kind} featurizer: {n} frames at {r} Hz")
        except Exception as e:
            fail(job_id, src, e)
            continue
        pending.append((job_id, src, r, ch, n))
    # batch by (rate, channels), longest first: padding waste stays small and a batch is one launch
    pending.sort(key=lambda g: (g[2], g[3], -g[4]))
    group, key = [], None
    for g in pending:
        k = (g[2], g[3])
        longest = group[0][4] if group else g[4]
        if group and (k != key or len(group) >= max_clips or (len(group) + 1) * longest * g[3] > max_batch_samples):
            flush(group)
            group = []
        group.append(g)
        key = k
    flush(group)
    stats["seconds"] = time.time() - t_start
    return stats


def main(argv=None):
    ap = argparse.ArgumentParser(description="Featurize every clip of a Batch-Processing manifest (one rank per GPU).")
    ap.add_argument("manifest")
    ap.add_argument("out_dir")
    ap.add_argument("--featurizer", default="whisper80", choices=sorted(FEATURIZERS))
    ap.add_argument("--format", default="npy", choices=["npy", "safetensors"])
    ap.add_argument("--max-clips", type=int, default=512)
    ap.add_argument("--padding", type=int, default=0, help="whisper: zeros appended before framing (transcribe uses 480000)")
    ap.add_argument("--fp16", action="store_true", help="whisper: write encoder-ready float16 features")
    args = ap.parse_args(argv)

    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rate, kind, _ = FEATURIZERS[args.featurizer]
    fz = GpuFeaturizer(args.featurizer, device=torch.device("cuda", local),
                       out_dtype=torch.float16 if args.fp16 else torch.float32, padding=args.padding)
    st = run_sweep(args.manifest, args.out_dir, fz, rate, kind, rank, world, args.format, args.max_clips,
                   log=lambda m: print(f"[rank {rank}] {m}", flush=True))
    table = _manifest.gather_counts(st["done"], st["frames"], st["samples"], device=torch.device("cuda", local))
    if world > 1:
        dist.barrier()
    if rank == 0:
        report = dict(featurizer=args.featurizer, world=world, per_rank=table.tolist(),
                      clips=int(table[:, 0].sum()), frames=int(table[:, 1].sum()), samples=int(table[:, 2].sum()),
                      rank0=st)
        with open(os.path.join(args.out_dir, "sweep_report.json"), "w") as f:
            json.dump(report, f, indent=1)
        print(json.dumps(report))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
