"""N > 1 host logic on CPU: contiguous clip shards and the counts all-gather, world_size 2 on gloo."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from expressive_speech_translation_b200 import manifest


def test_shard_rows_partition():
    for n in (0, 1, 7, 400, 10000):
        for world in (1, 2, 4, 8):
            spans = [manifest.shard_rows(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        manifest.shard_rows(10, 2, 2)


def test_sub_batches_and_manifest(tmp_path):
    assert manifest.sub_batches(3, 10, 4) == [(3, 7), (7, 10)]
    assert manifest.sub_batches(5, 5, 4) == []
    p = tmp_path / "batch_manifest.csv"
    p.write_text("job_id,dataset,source_video_path\n1,a,/x/1.mp4\n2,b,/x/2.mp4\n")
    assert manifest.read_manifest(str(p)) == [["1", "a", "/x/1.mp4"], ["2", "b", "/x/2.mp4"]]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_rows, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = manifest.shard_rows(n_rows, rank, world)
    clips = hi - lo
    got = manifest.gather_counts(clips, clips * 6000, clips * 960000)
    ret[rank] = (lo, hi, got.tolist())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_counts():
    world, n_rows = 2, 10001
    port = _free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        procs = [ctx.Process(target=_worker, args=(r, world, port, n_rows, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        res = dict(ret)
    assert res[0][2] == res[1][2]                       # every rank sees the same table
    table = res[0][2]
    assert sum(row[0] for row in table) == n_rows       # shards cover the manifest exactly once
    assert [row[1] for row in table] == [row[0] * 6000 for row in table]
    assert res[0][1] == res[1][0]                       # contiguous
