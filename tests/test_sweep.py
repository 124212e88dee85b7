"""§8(f)4: the manifest sweep runner -- manifest I/O, per-job outputs, `.error` files, idempotent resume, rank shards.

CPU tests drive run_sweep with a stand-in featurizer (the runner's host logic is what is under test; the oracle plays
the featurizer so the cached arrays are meaningful); the `-m gpu` tests run the real CUDA path and compare the cache
with whisper_log_mel_batch / kaldi_fbank_batch bit for bit."""
import os
import socket
import wave

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from expressive_speech_translation_b200 import manifest, sweep


def _write_wav(path, pcm, rate):
    with wave.open(str(path), "wb") as f:
        f.setnchannels(pcm.shape[1])
        f.setsampwidth(2)
        f.setframerate(rate)
        f.writeframes(np.ascontiguousarray(pcm).astype("<i2").tobytes())


def _make_jobs(tmp_path, n=6, rate=16000, seed=7, corrupt=(3,), stereo=()):
    rng = np.random.default_rng(seed)
    rows = []
    for i in range(n):
        p = tmp_path / f"clip{i}.wav"
        frames = 4000 + 1600 * i
        ch = 2 if i in stereo else 1
        pcm = (rng.standard_normal((frames, ch)) * 3000).astype(np.int16)
        _write_wav(p, pcm, rate)
        if i in corrupt:
            p.write_bytes(b"RIFFxxxxnot a wave file")
        rows.append((f"job-{i:02d}", "TED", str(p)))
    man = tmp_path / "batch_manifest.csv"
    sweep.write_manifest(str(man), rows)
    return man, rows


def _oracle_featurizer(clips):
    out = []
    for rate, pcm in clips:
        assert rate == 16000
        x = torch.from_numpy(pcm.astype(np.float32) / 32768.0).mean(dim=1)
        out.append(oracle.whisper_log_mel(x, 80).numpy())
    return out


def test_manifest_roundtrip_matches_reference_header(tmp_path):
    man, rows = _make_jobs(tmp_path, n=2, corrupt=())
    text = man.read_text().splitlines()
    assert text[0] == "job_id,dataset,source_video_path"          # create_batch_manifest.py:52-56
    assert manifest.read_manifest(str(man)) == [list(r) for r in rows]


def test_sweep_outputs_errors_and_resume(tmp_path):
    man, rows = _make_jobs(tmp_path)
    out = tmp_path / "out"
    calls = []

    def fz(clips):
        calls.append(len(clips))
        return _oracle_featurizer(clips)

    st = sweep.run_sweep(str(man), str(out), fz, 16000, "whisper", max_clips=4)
    assert (st["rows"], st["done"], st["failed"], st["skipped"]) == (6, 5, 1, 0)
    assert sum(calls) == 5 and max(calls) <= 4
    err = (out / "job-03.error").read_text()
    assert err.startswith(f"Error processing {rows[3][2]}:\n\n")     # run_batch_job.py:84-85
    for i in (0, 1, 2, 4, 5):
        f = sweep.load_features(str(out / f"job-{i:02d}.npy"))
        assert f.shape == (80, (4000 + 1600 * i) // 160)
    assert not list(out.glob("*.tmp"))
    assert st["frames"] == sum((4000 + 1600 * i) // 160 for i in (0, 1, 2, 4, 5))
    # resume: finished jobs are skipped, the failed one is retried; once its source is repaired it completes and
    # its .error file goes away
    calls.clear()
    st2 = sweep.run_sweep(str(man), str(out), fz, 16000, "whisper")
    assert (st2["done"], st2["failed"], st2["skipped"]) == (0, 1, 5) and calls == []
    _write_wav(rows[3][2], np.zeros((8000, 1), np.int16) + 5, 16000)
    st3 = sweep.run_sweep(str(man), str(out), fz, 16000, "whisper")
    assert (st3["done"], st3["failed"], st3["skipped"]) == (1, 0, 5)
    assert not (out / "job-03.error").exists() and (out / "job-03.npy").exists()


def test_sweep_rejects_short_and_non_pcm16(tmp_path):
    short = tmp_path / "short.wav"
    _write_wav(short, np.zeros((100, 1), np.int16), 16000)
    wide = tmp_path / "wide.wav"
    with wave.open(str(wide), "wb") as f:
        f.setnchannels(1); f.setsampwidth(4); f.setframerate(16000); f.writeframes(b"\0" * 64000)
    man = tmp_path / "m.csv"
    sweep.write_manifest(str(man), [("a", "d", str(short)), ("b", "d", str(wide)), ("c", "d", str(tmp_path / "missing.wav"))])
    st = sweep.run_sweep(str(man), str(tmp_path / "o"), _oracle_featurizer, 16000, "whisper")
    assert (st["done"], st["failed"]) == (0, 3)
    assert "too short" in (tmp_path / "o" / "a.error").read_text()
    assert "16-bit PCM" in (tmp_path / "o" / "b.error").read_text()


def test_safetensors_cache(tmp_path):
    pytest.importorskip("safetensors")
    man, _ = _make_jobs(tmp_path, n=2, corrupt=())
    sweep.run_sweep(str(man), str(tmp_path / "o"), _oracle_featurizer, 16000, "whisper", fmt="safetensors")
    a = sweep.load_features(str(tmp_path / "o" / "job-00.safetensors"))
    assert a.shape == (80, 25) and a.dtype == np.float32


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _rank_main(rank, world, port, man, out, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    st = sweep.run_sweep(man, out, _oracle_featurizer, 16000, "whisper", rank=rank, world=world)
    table = manifest.gather_counts(st["done"], st["frames"], st["samples"])
    ret[rank] = (st["rows"], st["done"], st["skipped"], st["failed"], table.tolist())
    dist.barrier()
    dist.destroy_process_group()


def _run_two_ranks(man, out):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        procs = [ctx.Process(target=_rank_main, args=(r, world, port, str(man), str(out), ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
            assert p.exitcode == 0
        return dict(ret)


def test_two_rank_gloo_sweep_and_resume(tmp_path):
    man, rows = _make_jobs(tmp_path, n=7, corrupt=(5,))
    out = tmp_path / "out"
    res = _run_two_ranks(man, out)
    assert res[0][0] + res[1][0] == 7 and abs(res[0][0] - res[1][0]) <= 1    # contiguous shards of rows
    assert res[0][4] == res[1][4]                                            # the all-gathered table is shared
    assert sum(r[0] for r in res[0][4]) == 6                                 # clips done, summed over ranks
    assert (out / "job-05.error").exists() and len(list(out.glob("*.npy"))) == 6
    res2 = _run_two_ranks(man, out)                                          # resume: nothing left but the failure
    assert sum(res2[r][1] for r in (0, 1)) == 0 and sum(res2[r][2] for r in (0, 1)) == 6
    assert sum(res2[r][3] for r in (0, 1)) == 1


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_gpu_sweep_cache_equals_batched_call_bit_for_bit(tmp_path):
    import expressive_speech_translation_b200 as fe
    man, rows = _make_jobs(tmp_path, n=6, corrupt=(2,))
    out = tmp_path / "out"
    fz = sweep.GpuFeaturizer("whisper80")
    st = sweep.run_sweep(str(man), str(out), fz, 16000, "whisper", max_clips=4)
    assert (st["done"], st["failed"]) == (5, 1)
    for i in (0, 1, 3, 4, 5):
        _, pcm = sweep.read_wav_pcm16(rows[i][2])
        x = torch.from_numpy(pcm[:, 0].copy()).cuda()[None]
        want = fe.whisper_log_mel_batch(x, n_mels=80, norm="per_clip")[0].cpu().numpy()
        got = sweep.load_features(str(out / f"job-{i:02d}.npy"))
        assert got.shape == want.shape and np.array_equal(got, want)
        ref = oracle.whisper_log_mel(torch.from_numpy(pcm[:, 0].astype(np.float32) / 32768.0), 80, dtype=torch.float64).numpy()
        assert np.abs(got - ref).max() <= 1e-4


@pytest.mark.gpu
def test_gpu_sweep_conditions_44k_stereo_and_kaldi(tmp_path):
    """44.1 kHz stereo sources go through the conditioning kernel (reference filter) before the featurizer"""
    import expressive_speech_translation_b200 as fe
    man, rows = _make_jobs(tmp_path, n=3, rate=44100, corrupt=(), stereo=(0, 1, 2))
    out = tmp_path / "out"
    st = sweep.run_sweep(str(man), str(out), sweep.GpuFeaturizer("kaldi"), 16000, "kaldi")
    assert (st["done"], st["failed"]) == (3, 0)
    for i in range(3):
        _, pcm = sweep.read_wav_pcm16(rows[i][2])
        w = fe.condition_pcm(torch.from_numpy(pcm.copy()).cuda()[None], 44100, 16000, **sweep.REF_RESAMPLE)
        want = fe.kaldi_fbank_batch(w, subtract_mean=True)[0].cpu().numpy()
        got = sweep.load_features(str(out / f"job-{i:02d}.npy"))
        assert got.shape == want.shape
        assert np.abs(got - want).max() <= 2e-5      # batched with ragged neighbours vs alone: same kernel, same taps
