"""b200feat_whisper_logmel_noalloc: the Whisper log-mel entry point that allocates nothing (scratch = a static device array
of 1 MiB inside the library).  No counterpart in the reference.

Collected LAST on purpose: the file name sorts after every other test file.
The new entry point is called through the C ABI (ctypes) and compared with the existing workspace entry point
b200feat_whisper_logmel on the same inputs, on two calls in a row that share the static array (the second call is
smaller and ragged, so stale contents of the first would show)."""
import ctypes

import numpy as np
import pytest
import torch

TOL = 1e-4      # the max-abs tolerance of the Whisper parity tests (tests/test_gpu_parity.py)


def _call_pair(lib, x, lengths, n_mels, padding, norm):
    """-> (out of the workspace entry point, out of the no-alloc entry point), both (B, n_mels, T) float32 on the GPU"""
    B, L = x.shape
    T = (L + padding) // 160
    vp = ctypes.c_void_p
    ln = None if lengths is None else torch.tensor(lengths, dtype=torch.int64, device=x.device)
    lnp = None if ln is None else vp(ln.data_ptr())
    stream = vp(torch.cuda.current_stream().cuda_stream)
    # reference: caller-owned workspace
    nws = int(lib.b200feat_workspace_bytes(0, B, L + padding))
    ws = torch.empty(nws, dtype=torch.uint8, device=x.device)
    ref = torch.full((B, n_mels, T), 7.0, dtype=torch.float32, device=x.device)
    rc = lib.b200feat_whisper_logmel(vp(x.data_ptr()), lnp, B, L, L, n_mels, padding, norm, vp(ref.data_ptr()), n_mels * T, T,
                                     vp(ws.data_ptr()), nws, stream)
    assert rc == 0, lib.b200feat_last_error()
    # under test: no workspace argument at all; inputs and outputs were allocated above, before the call
    got = torch.full((B, n_mels, T), 7.0, dtype=torch.float32, device=x.device)
    torch.cuda.synchronize()
    rc = lib.b200feat_whisper_logmel_noalloc(vp(x.data_ptr()), lnp, B, L, L, n_mels, padding, norm, vp(got.data_ptr()),
                                             n_mels * T, T, stream)
    assert rc == 0, lib.b200feat_last_error()
    torch.cuda.synchronize()
    return ref, got


@pytest.mark.gpu
def test_noalloc_entry_point_matches_workspace_entry_point_twice_in_a_row():
    from expressive_speech_translation_b200 import _lib
    import oracle
    lib = _lib.lib
    assert lib.b200feat_noalloc_workspace_max() == 1 << 20
    g = torch.Generator().manual_seed(77)
    # call 1: 6 clips of 3 s, 128 bins, per-clip normalisation
    x1 = (0.1 * torch.randn(6, 48000, generator=g)).clamp(-1, 1)
    ref1, got1 = _call_pair(lib, x1.cuda(), None, 128, 0, 0)
    assert float((ref1 - got1).abs().max()) <= TOL
    o1 = oracle.whisper_log_mel(x1, 128, dtype=torch.float64, norm="per_clip")
    assert float((got1.cpu().double() - o1).abs().max()) <= TOL
    # call 2, straight after, on the SAME static array: fewer, shorter, ragged clips, 80 bins, per-call normalisation,
    # zero padding behind the audio -- anything left over from call 1 (larger keys, more units) would change the clamp
    x2 = 0.01 * (0.1 * torch.randn(3, 20000, generator=g)).clamp(-1, 1)
    ref2, got2 = _call_pair(lib, x2.cuda(), [20000, 9001, 15555], 80, 3200, 1)
    assert float((ref2 - got2).abs().max()) <= TOL
    assert torch.equal(ref2, got2)          # same kernels, same scratch layout: in fact bit-identical
    assert torch.equal(ref1, got1)
    # and call 1 again reproduces itself after call 2 used the array
    ref3, got3 = _call_pair(lib, x1.cuda(), None, 128, 0, 0)
    assert torch.equal(got3, got1)


@pytest.mark.gpu
def test_noalloc_rejects_calls_beyond_the_static_array():
    from expressive_speech_translation_b200 import _lib
    lib = _lib.lib
    vp = ctypes.c_void_p
    B, L = 4096, 480000                     # needs 4096 * 47 * 8 bytes = 1.5 MiB > 1 MiB: rejected before any launch
    T = L // 160
    dummy = torch.zeros(16, device="cuda")  # never dereferenced
    rc = lib.b200feat_whisper_logmel_noalloc(vp(dummy.data_ptr()), None, B, L, L, 128, 0, 0, vp(dummy.data_ptr()), 128 * T, T, None)
    assert rc == -3                         # B200FEAT_EWORKSPACE
    msg = lib.b200feat_last_error().decode()
    assert "built-in array" in msg and "1048576" in msg
    # an empty call is a no-op like the workspace entry point's
    assert lib.b200feat_whisper_logmel_noalloc(None, None, 0, 16000, 16000, 80, 0, 0, None, 8000, 100, None) == 0


def test_noalloc_symbol_is_declared_and_bound():
    """CPU: the header declares it, the library exports it, the ctypes table binds it (no compute call here)"""
    import os
    from expressive_speech_translation_b200 import _lib
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "b200feat.h")).read()
    assert "b200feat_whisper_logmel_noalloc(" in hdr and "no counterpart in the reference" in hdr.lower()
    assert hasattr(_lib.lib, "b200feat_whisper_logmel_noalloc") and "b200feat_whisper_logmel_noalloc" in _lib.SYMBOLS
    assert _lib.lib.b200feat_noalloc_workspace_max() == 1 << 20
