#!/usr/bin/env python
"""Phase timeline of the Tensor-Memory Whisper kernel (debug build with -DFE_W3_TRACE; never the product library).
    python tools/trace_w3.py build      # here: nvcc -> lib/libb200feat_trace3.so
    python tools/trace_w3.py run [out]  # on the GPU box: runs TR_CLIPS clips, prints per-phase cycles of CTA 0's 16 warps
"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "expressive_speech_translation_b200")
LIB = os.environ.get("TR_LIB", os.path.join(PKG, "lib", "libb200feat_trace3.so"))
UNITS, WARPS, EV = 16, 16, 10

if sys.argv[1] == "build":
    cmd = ["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-DFE_W3_TRACE",
           "-Xcompiler", "-fPIC", "-shared", "-o", LIB, os.path.join(PKG, "csrc", "b200feat.cu")] + sys.argv[2:]
    subprocess.check_call(cmd)
    print(LIB)
    sys.exit(0)

import numpy as np  # noqa: E402
import torch  # noqa: E402

lib = ctypes.CDLL(LIB)
n_mels = 128
B, L = int(os.environ.get("TR_CLIPS", "148")), 480000
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
buf = np.zeros(UNITS * WARPS * EV, dtype=np.int64)
lib.b200feat_debug_trace3.argtypes = [ctypes.c_void_p]
assert lib.b200feat_debug_trace3(buf.ctypes.data) == 0
tr = buf.reshape(UNITS, WARPS, EV)
lines = []
names = ["waitW", "A", "waitPF", "M", "waitA", "B1", "B2"]
for it in range(3, 9):
    base = tr[it, :, 0].min()
    lines.append(f"--- iteration {it}  (period {int(tr[it + 1, 0, 0] - tr[it, 0, 0])} cycles)")
    for w in range(16):
        e = tr[it, w]
        d = [int(e[i + 1] - e[i]) if e[i + 1] and e[i] else 0 for i in range(7)]
        lines.append(f" team {w % 4} member {w // 4} start {int(e[0] - base):6d} " + " ".join(f"{n}={v:5d}" for n, v in zip(names, d))
                     + f"  next {int(tr[it + 1, w, 0] - e[0]):6d}")
txt = "\n".join(lines)
print(txt)
if len(sys.argv) > 2:
    os.makedirs(os.path.dirname(sys.argv[2]) or ".", exist_ok=True)
    open(sys.argv[2], "w").write(txt + "\n")
