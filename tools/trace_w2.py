#!/usr/bin/env python
"""Phase timeline of the Whisper frame-pair kernel (debug build with -DFE_W2_TRACE; never the product library).
    python tools/trace_w2.py build      # here: nvcc -> lib/libb200feat_trace.so
    python tools/trace_w2.py run [out]  # on the GPU box: runs 64 clips, prints per-phase cycles of CTA 0
"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "expressive_speech_translation_b200")
LIB = os.environ.get("TR_LIB", os.path.join(PKG, "lib", "libb200feat_trace.so"))
UNITS, WARPS, EV = 12, 20, 8  # 20 = trace slots per iteration (>= warps)

if sys.argv[1] == "build":
    cmd = ["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-DFE_W2_TRACE",
           "-Xcompiler", "-fPIC", "-shared", "-o", LIB, os.path.join(PKG, "csrc", "b200feat.cu")] + sys.argv[2:]
    subprocess.check_call(cmd)
    print(LIB)
    sys.exit(0)

import numpy as np  # noqa: E402
import torch  # noqa: E402

lib = ctypes.CDLL(LIB)
n_mels = 128
B, L = int(os.environ.get("TR_CLIPS", "64")), 480000
g = torch.Generator(device="cuda").manual_seed(1)
x = (0.1 * torch.randn(B, L, generator=g, device="cuda")).clamp_(-1, 1)
T = L // 160
out = torch.empty(B, n_mels, T, device="cuda")
lib.b200feat_workspace_bytes.restype = ctypes.c_int64
lib.b200feat_workspace_bytes.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int64]
wsb = lib.b200feat_workspace_bytes(0, B, L)
ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
f = lib.b200feat_whisper_logmel
f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_int64, ctypes.c_int32,
              ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p,
              ctypes.c_int64, ctypes.c_void_p]
for _ in range(3):
    rc = f(x.data_ptr(), None, B, L, L, n_mels, 0, 0, out.data_ptr(), n_mels * T, T, ws.data_ptr(), wsb, None)
    assert rc == 0
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    f(x.data_ptr(), None, B, L, L, n_mels, 0, 0, out.data_ptr(), n_mels * T, T, ws.data_ptr(), wsb, None)
e1.record()
torch.cuda.synchronize()
print(f"{B} clips: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per call (both kernels) -> {e0.elapsed_time(e1) / 10 * 1024 / B:.3f} ms per 1024 clips")
buf = np.zeros(UNITS * WARPS * EV, dtype=np.int64)
lib.b200feat_debug_trace.argtypes = [ctypes.c_void_p]
assert lib.b200feat_debug_trace(buf.ctypes.data) == 0
tr = buf.reshape(UNITS, WARPS, EV)
t0 = tr[0, :12, 0].min()
lines = []


def ops(w):
    return ["B", "A1", "A2+M" if not (w & 4) else "M", "-" if not (w & 4) else "A2", "keys", "bar"]


for it in range(2, 8):
    lines.append(f"--- iteration {it}  (unit time {int(tr[it+1, 0, 0] - tr[it, 0, 0])} cycles)")
    for w in range(12):
        e = tr[it, w]
        d = [int(e[i + 1] - e[i]) for i in range(6)]
        lines.append(f" cw{w:2d} start {int(e[0]-t0):8d} " + " ".join(f"{n:>4s}={v:5d}" for n, v in zip(ops(w), d)))
    for w in range(12, 16):
        e = tr[it, w]
        lines.append(f" lw{w:2d} start {int(e[0]-t0):8d} waitFree={int(e[1]-e[0]):6d} store={int(e[2]-e[1]):5d} issue={int(e[3]-e[2]):5d}")
txt = "\n".join(lines)
print(txt)
if len(sys.argv) > 2:
    os.makedirs(os.path.dirname(sys.argv[2]) or ".", exist_ok=True)
    open(sys.argv[2], "w").write(txt + "\n")
