"""CPU suite: the rolled mel-projection tables the kernels bake (csrc/generated/fe_tables.h) reproduce the filter
matrices of tests/golden/baked_filters.npz exactly -- every filter once, every tap with its weight, padded taps zero
and inside the row -- and the chunk/class ranges partition the descriptors."""
import os
import re

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
TABLES = os.path.join(os.path.dirname(HERE), "expressive_speech_translation_b200", "csrc", "generated", "fe_tables.h")


def _array(text, name, dtype):
    m = re.search(r"FE_CONSTANT \w+ " + re.escape(name) + r"\[(\d+)\] = \{(.*?)\};", text, re.S)
    assert m, name
    vals = [v.strip().rstrip("f") for v in m.group(2).replace("\n", " ").split(",") if v.strip()]
    a = np.array([float(v) for v in vals], dtype=np.float64).astype(dtype)
    assert a.size == int(m.group(1))
    return a


def _define(text, name):
    m = re.search(r"#define " + re.escape(name) + r" (.*)", text)
    assert m, name
    return m.group(1).strip()


@pytest.fixture(scope="module")
def tables():
    with open(TABLES) as f:
        return f.read()


@pytest.mark.parametrize("nm,key", [(80, "whisper80"), (128, "whisper128")])
def test_whisper_melloop_tables(tables, golden, nm, key):
    filt = golden("baked_filters.npz")[key]
    pre = f"fe_melloop_w{nm}"
    caps = [int(c) for c in _define(tables, pre.upper() + "_CAPS").split(",")]
    ncls = int(_define(tables, pre.upper() + "_NCLS"))
    assert ncls == len(caps) and caps == sorted(caps)
    desc = _array(tables, pre + "_desc", np.int64).reshape(-1, 2)
    w = _array(tables, pre + "_w", np.float32)
    assert w.size == int(_define(tables, pre.upper() + "_NW"))
    ranges = _array(tables, pre + "_ranges", np.int64).reshape(-1, ncls + 1)
    assert desc.shape[0] == nm
    # ranges: chunk after chunk, class after class, a partition of the descriptor list
    flat = ranges.reshape(-1)
    assert flat[0] == 0 and flat[-1] == nm and (np.diff(flat) >= 0).all()
    assert (ranges[1:, 0] == ranges[:-1, -1]).all()
    rebuilt = np.zeros_like(filt)
    seen = set()
    for c in range(ranges.shape[0]):
        for k in range(ncls):
            for i in range(ranges[c, k], ranges[c, k + 1]):
                k0, m, off = (desc[i, 0] & 0xffff) // 8, desc[i, 0] >> 16, desc[i, 1] // 4
                cap = caps[k]
                assert m not in seen and 0 <= k0 and k0 + cap <= 200        # padded taps stay inside the 200 bins
                assert off % (1 if cap == 1 else 2 if cap == 2 else 4) == 0   # alignment of the weight loads
                seen.add(int(m))
                rebuilt[m, k0:k0 + cap] = w[off:off + cap]
    assert seen == set(range(nm))
    assert np.array_equal(rebuilt, filt)


def test_cosy_melroll_tables(tables, golden):
    filt = golden("baked_filters.npz")["cosy80"]
    desc = _array(tables, "fe_melroll_cosy_desc", np.int64).reshape(-1, 3)
    w = _array(tables, "fe_melroll_cosy_w", np.float32)
    assert desc.shape[0] == 80
    rebuilt = np.zeros_like(filt)
    for m, (k0, n4, off) in enumerate(desc):
        assert n4 % 4 == 0 and n4 >= 4 and 0 <= k0 and k0 + n4 <= 640
        rebuilt[m, k0:k0 + n4] = w[off:off + n4]
    assert np.array_equal(rebuilt, filt)
