"""Builds the host emulation of the kernel phases (tests/emu/emu.cpp) with g++ into tests/emu/_build/."""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "expressive_speech_translation_b200", "csrc")
OUT = os.path.join(HERE, "_build", "libfe_emu.so")


def build():
    src = os.path.join(HERE, "emu.cpp")
    deps = [src] + [os.path.join(dp, f) for dp, _, fs in os.walk(CSRC) for f in fs if f.endswith((".h", ".cuh"))]
    if os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=fast", "-I", CSRC, src, "-o", OUT]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stderr)
    return OUT


def load():
    lib = ctypes.CDLL(build())
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    lib.emu_whisper.argtypes = [vp, vp, i32, i64, i64, i32, i32, i32, vp, i64]
    lib.emu_whisper_f16.argtypes = [vp, vp, i32, i64, i64, i32, i32, i32, vp, i64]
    lib.emu_whisper3.argtypes = [vp, vp, i32, i64, i64, i32, i32, i32, vp, i32, i64]
    lib.emu_wav2lip.argtypes = [vp, vp, i32, i64, i64, vp, i64, i32, i32]
    lib.emu_gate.argtypes = [vp, vp, i32, i64, i64, vp]
    lib.emu_whisper3_windows.argtypes = [vp, vp, i32, i64, i64, i32, i32, i32, i32, i32, vp, i32]
    lib.emu_whisper_s16.argtypes = [vp, vp, i32, i64, i64, i32, i32, i32, vp, i64]
    lib.emu_f2h.argtypes = [vp, vp, i64]
    lib.emu_f2h.restype = None
    lib.emu_kaldi.argtypes = [vp, vp, i32, i64, i64, i32, vp, i64]
    lib.emu_kaldi3.argtypes = [vp, vp, i32, i64, i64, i32, vp, i64]
    lib.emu_cosy.argtypes = [vp, vp, i32, i64, i64, vp, i64]
    f64 = ctypes.c_double
    lib.emu_resample.argtypes = [vp, i32, i32, vp, vp, i32, i64, i32, i32, i32, f64, i32, f64, vp, i64, vp]
    return lib
