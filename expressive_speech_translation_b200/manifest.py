"""Clip sharding for the batch sweep.

The reference's batch driver runs one manifest row = one clip = one process
(/root/reference/Batch-Processing/submit_slurm_job.sh:10,41,53: SLURM array task i reads CSV row i + 1).  Here one
process per GPU takes a CONTIGUOUS block of rows; clips are independent, so there is no collective on the data
path.  The only exchange is an all-gather of three int64 per rank (clips, frames, samples) for the sweep report.
"""
import csv

import torch


def shard_rows(n_rows, rank, world):
    """rows [lo, hi) of rank `rank` out of `world`: lo = rank*n // world (contiguous, sizes differ by <= 1)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    if n_rows < 0:
        raise ValueError("n_rows < 0")
    return rank * n_rows // world, (rank + 1) * n_rows // world


def read_manifest(path):
    """rows of a Batch-Processing manifest (job_id, dataset, source_video_path; header row skipped), as written by
    /root/reference/Batch-Processing/create_batch_manifest.py:52-56."""
    with open(path, newline="") as f:
        rows = list(csv.reader(f))
    if rows and rows[0] and rows[0][0].strip().lower() == "job_id":
        rows = rows[1:]
    return [r for r in rows if r]


def sub_batches(lo, hi, max_clips):
    """[lo, hi) cut into consecutive sub-batches of at most max_clips clips (bounds device residency)."""
    if max_clips < 1:
        raise ValueError("max_clips < 1")
    return [(a, min(a + max_clips, hi)) for a in range(lo, hi, max_clips)]


def gather_counts(clips, frames, samples, device=None, group=None):
    """-> (world, 3) int64 tensor of every rank's (clips, frames, samples).  Works on gloo (CPU tensors) and nccl
    (pass the rank's cuda device).  Without an initialised process group it returns the local row."""
    import torch.distributed as dist
    mine = torch.tensor([int(clips), int(frames), int(samples)], dtype=torch.int64, device=device)
    if not (dist.is_available() and dist.is_initialized()):
        return mine.unsqueeze(0)
    world = dist.get_world_size(group)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return torch.stack(out)


def bind_host_to_gpu(device_index):
    """Pins the calling process to the CPUs NVML lists as local to CUDA device `device_index`, so that host buffers
    it allocates afterwards (first touch) live on the NUMA node next to that GPU's PCIe root.  With one process per
    GPU (the sweep's launch model) this keeps every rank's H2D/D2H traffic off the socket interconnect.
    -> the cpu list it bound to, or None when NVML / the affinity call is unavailable (nothing is changed then)."""
    import os
    try:
        import pynvml
        props = torch.cuda.get_device_properties(device_index)
        bus_id = f"{props.pci_domain_id:08x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus_id)
        n_words = (os.cpu_count() + 63) // 64
        words = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None
