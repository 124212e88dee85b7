"""Builds expressive_speech_translation_b200/lib/libb200feat.so with nvcc for sm_100a (in-tree)."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "b200feat.cu")
LIB = os.path.join(HERE, "lib", "libb200feat.so")
DEPS = [
    SRC,
    os.path.join(HERE, "csrc", "b200feat_host.inl"),
    os.path.join(HERE, "csrc", "fe_common.h"),
    os.path.join(HERE, "csrc", "fe_whisper2.cuh"),
    os.path.join(HERE, "csrc", "fe_whisper3.cuh"),
    os.path.join(HERE, "csrc", "fe_kaldi.cuh"),
    os.path.join(HERE, "csrc", "fe_kaldi3.cuh"),
    os.path.join(HERE, "csrc", "fe_cosy.cuh"),
    os.path.join(HERE, "csrc", "fe_resample.cuh"),
    os.path.join(HERE, "csrc", "fe_resample_plan.h"),
    os.path.join(HERE, "csrc", "generated", "fe_codelets.h"),
    os.path.join(HERE, "csrc", "generated", "fe_tables.h"),
    os.path.join(os.path.dirname(HERE), "include", "b200feat.h"),
]


def nvcc_path():
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def up_to_date():
    if not os.path.exists(LIB):
        return False
    t = os.path.getmtime(LIB)
    return all(os.path.getmtime(d) <= t for d in DEPS)


def build(force=False, verbose=False):
    if not force and up_to_date():
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    cmd = [nvcc_path(), "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
           "-Xcompiler", "-fPIC", "-shared", "-o", LIB, SRC]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
