#!/usr/bin/env python
"""Generates the straight-line parts of the featurizer kernels.

  expressive_speech_translation_b200/csrc/generated/fe_codelets.h   DFT codelets (tools/fftgen.py)
  expressive_speech_translation_b200/csrc/generated/fe_tables.h     windows, inter-stage twiddles,
                                                                    unrolled sparse mel projections

Run from the repo root:  python tools/gen_kernels.py
The output is committed; the build does not run this script.

Filter tables baked here:
  * Slaney mel (Whisper 80/128 @16k/400, CosyVoice 80 @24k/1920): librosa.filters.mel formula
    (SURVEY.md appendix A.1), float64 -> float32, identical to whisper's mel_filters.npz /
    transformers.audio_utils.mel_filter_bank(norm="slaney", mel_scale="slaney").
  * kaldi banks: torchaudio.compliance.kaldi.get_mel_banks(80, 512, 16000, 20, 0, 100, -500, 1.0)
    itself (site-packages/torchaudio/compliance/kaldi.py:436-511), float32.
"""
import math
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import fftgen as g  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "expressive_speech_translation_b200", "csrc", "generated")


# ------------------------------------------------------------------ filters
def hz_to_mel(f):
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, f / f_sp)


def mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def slaney(sr, n_fft, n_mels, fmin, fmax):
    nb = n_fft // 2 + 1
    ff = np.arange(nb, dtype=np.float64) * (float(sr) / n_fft)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2))
    fd = np.diff(mel_f)
    ramps = mel_f[:, None] - ff[None, :]
    w = np.zeros((n_mels, nb))
    for i in range(n_mels):
        w[i] = np.maximum(0.0, np.minimum(-ramps[i] / fd[i], ramps[i + 2] / fd[i + 1]))
    w *= (2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels]))[:, None]
    return w.astype(np.float32)


def kaldi_banks():
    import torchaudio
    b, _ = torchaudio.compliance.kaldi.get_mel_banks(80, 512, 16000.0, 20.0, 0.0, 100.0, -500.0, 1.0)
    out = np.zeros((80, 257), dtype=np.float32)
    out[:, :256] = b.numpy()
    return out


# ------------------------------------------------------------------ helpers
def flit(v):
    v = float(np.float32(v))
    s = repr(v)
    if "e" not in s and "." not in s and "inf" not in s and "nan" not in s:
        s += ".0"
    return s + "f"


def emit_array(name, arr, per_line=8, ctype="float"):
    if ctype == "int":
        flat = [str(int(v)) for v in np.asarray(arr).reshape(-1)]
    else:
        flat = [flit(v) for v in np.asarray(arr).reshape(-1)]
    lines = [f"FE_CONSTANT {ctype} {name}[{len(flat)}] = {{"]
    for i in range(0, len(flat), per_line):
        lines.append("    " + ", ".join(flat[i:i + per_line]) + ",")
    lines.append("};")
    return "\n".join(lines)


def gen_codelet(name, n, real_in=False, nz=None, outs=None, real_out=()):
    """void name(const float* xr, [const float* xi,] float* yr, float* yi)
    inputs x[0..nz) (rest zero); outputs listed in `outs` written at yr[j], yi[j] (j = position
    in `outs`); outputs whose index is in `real_out` have no imaginary part emitted."""
    d = g.Dag()
    nz = n if nz is None else nz
    xs = []
    for i in range(n):
        if i < nz:
            xs.append(g.C(d.inp(("r", i)), None if real_in else d.inp(("i", i))))
        else:
            xs.append(g.C(None, None))
    ys = g.dft(d, xs)
    outs = list(range(n)) if outs is None else list(outs)
    vals, slots = [], []
    for j, k in enumerate(outs):
        vals.append(ys[k].re)
        slots.append(("r", j))
        if k not in real_out:
            vals.append(ys[k].im)
            slots.append(("i", j))
    ops = g.count_ops(d, vals)

    def in_expr(nm):
        return f"x{nm[0]}[{nm[1]}]"

    def out_stmt(idx, e):
        s = slots[idx]
        return f"y{s[0]}[{s[1]}] = {e};"

    body = g.emit_body(d, vals, in_expr, out_stmt)
    n_fma = body.count("fe_fmac(") + body.count("fe_fmsc(") + body.count("fe_fnmac(")
    n_ops = body.count("= fe_")
    args = "const V* __restrict__ xr, " + ("" if real_in else "const V* __restrict__ xi, ")
    args += "V* __restrict__ yr, V* __restrict__ yi"
    hdr = (f"// DFT-{n} {'real' if real_in else 'complex'} input ({nz} non-zero), outputs k={outs[0]}..{outs[-1]}"
           f"  [{ops['add'] + ops['sub']} add/sub + {ops['mul']} mul -> {n_ops} instructions, {n_fma} of them fused]\n")
    return hdr + f"template <class V>\nFE_HD FE_INLINE void {name}({args}) {{\n{body}\n}}\n"


# ------------------------------------------------------------------ mel projection
def partition(costs, nch):
    """contiguous partition of `costs` into nch chunks minimising the max chunk cost."""
    n = len(costs)
    lo, hi = max(costs), sum(costs)

    def fits(cap):
        cnt, cur = 1, 0
        for c in costs:
            if cur + c > cap:
                cnt += 1
                cur = c
            else:
                cur += c
        return cnt <= nch

    while lo < hi:
        mid = (lo + hi) // 2
        if fits(mid):
            hi = mid
        else:
            lo = mid + 1
    cap = lo
    bounds, cur, start = [], 0, 0
    for i, c in enumerate(costs):
        if cur + c > cap:
            bounds.append((start, i))
            start, cur = i, c
        else:
            cur += c
    bounds.append((start, n))
    while len(bounds) < nch:  # split the largest multi-filter chunk until we have nch
        j = max(range(len(bounds)), key=lambda q: (bounds[q][1] - bounds[q][0] > 1) * sum(costs[bounds[q][0]:bounds[q][1]]))
        a, b = bounds[j]
        if b - a <= 1:
            break
        mid = (a + b) // 2
        bounds[j:j + 1] = [(a, mid), (mid, b)]
    return bounds


def gen_melproj(name, filt, nch, max_bin, epi_cost=6):
    """template<class V, class Load, class Emit> void name(int chunk, Load&& ld, Emit&& emit)
    ld(k) returns this lane's power (V = float, or fe2 for a frame pair) at FFT bin k.  emit(m, value)
    is called once per filter of the chunk."""
    n_mels = filt.shape[0]
    assert not np.any(filt[:, max_bin + 1:]), "filter weight beyond max_bin"
    costs = []
    for m in range(n_mels):
        nnz = int(np.count_nonzero(filt[m]))
        costs.append(2 * nnz + epi_cost)
    bounds = partition(costs, nch)
    out = [f"// sparse mel projection, {n_mels} filters, {int(np.count_nonzero(filt))} non-zeros, {len(bounds)} chunks",
           f"#define {name.upper()}_NCHUNK {len(bounds)}",
           "template <class V, class Load, class Emit>",
           f"FE_HD FE_INLINE void {name}(int chunk, Load&& ld, Emit&& emit) {{",
           "    switch (chunk) {"]
    for c, (a, b) in enumerate(bounds):
        out.append(f"    case {c}: {{  // filters {a}..{b - 1}")
        bins = sorted(set(int(k) for m in range(a, b) for k in np.nonzero(filt[m])[0]))
        body = []
        decl = set()
        for k in bins:
            body.append(f"        {{ const V p = ld({k});")
            for m in range(a, b):
                w = filt[m, k]
                if w != 0:
                    if m in decl:
                        body.append(f"          a{m} = fe_fmac({flit(w)}, p, a{m});")
                    else:
                        body.append(f"          a{m} = fe_mulc({flit(w)}, p);")
                        decl.add(m)
            body.append("        }")
        out.append("        V " + ", ".join(f"a{m} = fe_zero<V>()" for m in range(a, b)) + ";")
        out.extend(body)
        for m in range(a, b):
            out.append(f"        emit({m}, a{m});")
        out.append("    } break;")
    out.append("    default: break;")
    out.append("    }")
    out.append("}")
    return "\n".join(out) + "\n"


def gen_melloop(name, filt, nch, caps, n_bins=200, emit_cost=17, chunk_weight=None, k0_scale=8, by_class=False):
    """tables for the rolled mel projection of the Whisper frame-pair kernel.  (The fully unrolled form does
    not fit the instruction cache next to the FFT codelets.)  Filters are sorted into classes by tap count
    (`caps`, ascending); a filter is zero-padded to the cap of its class, so each class is a branch-free loop:
      name_desc[2*i]   = 8 * k0 | m << 16   (byte offset of the first tap in a row of power pairs, mel row)
      name_desc[2*i+1] = 4 * off            (byte offset of the first weight in name_w; 4- / 8-byte aligned for
                                             the 1- / 2-tap classes, 16-byte aligned otherwise)
      name_w           = weights, cap per filter
      name_ranges[(NCLS+1)*c + k] = first descriptor of class k in chunk c (k = NCLS: end of the chunk); descriptors
                             are ordered chunk by chunk, class by class; a chunk's share of the filters
                             follows chunk_weight (cost model 2 * cap + emit_cost).  The kernel copies desc
                             and w to shared memory."""
    n_mels = filt.shape[0]
    nmax = max(int(np.count_nonzero(r)) for r in filt)
    ncls = len(caps)
    assert caps[-1] >= nmax and list(caps) == sorted(caps)
    info = []
    for m in range(n_mels):
        nz = np.nonzero(filt[m])[0]
        if len(nz) == 0:
            k0, n = 0, 0
        else:
            k0, n = int(nz[0]), int(nz[-1] - nz[0] + 1)
            assert np.all(filt[m, k0:k0 + n] != 0), "taps of a filter must be contiguous"
        cls = next(i for i, c in enumerate(caps) if n <= c)
        cap = caps[cls]
        k0 = min(k0, n_bins - cap)           # padded taps must stay inside the row
        info.append((m, cls, cap, k0, [float(v) for v in filt[m, k0:k0 + cap]]))
        assert np.count_nonzero(filt[m, :k0]) == 0 and np.count_nonzero(filt[m, k0 + cap:]) == 0
    # weighted longest-processing-time assignment of filters to chunks (rows are independent, so a chunk need
    # not be a contiguous range); chunk_weight[c] = share of the mel work chunk c should carry
    cw = chunk_weight if chunk_weight is not None else [1.0] * nch
    assert len(cw) == nch
    load = [0.0] * nch
    chunks = [[] for _ in range(nch)]
    if by_class:
        # a chunk = a contiguous run of the class-sorted filter list (equal cost per chunk): a warp then sees one or
        # two classes with several filters each, so its class loops can keep several independent filters in flight
        order = sorted(info, key=lambda t: (t[1], t[0]))
        for c, (a, b) in enumerate(partition([3 * t[2] + emit_cost for t in order], nch)):
            chunks[c] = order[a:b]
    else:
        for it in sorted(info, key=lambda t: -(2 * t[2] + emit_cost)):
            c = min(range(nch), key=lambda q: (load[q] + 2 * it[2] + emit_cost) / cw[q])
            chunks[c].append(it)
            load[c] += 2 * it[2] + emit_cost
    desc, wts, ranges = [], [], []
    for c in range(nch):
        chunks[c].sort(key=lambda t: t[0])
        for cls in range(ncls):
            ranges.append(len(desc) // 2)
            for m, c_, cap, k0, w in chunks[c]:
                if c_ != cls:
                    continue
                walign = 1 if cap == 1 else (2 if cap == 2 else 4)     # 32- / 64- / 128-bit weight loads
                while len(wts) % walign:
                    wts.append(0.0)
                desc += [(k0_scale * k0 | (m << 16)) if k0_scale != 512 else (512 * k0 | m), 4 * len(wts)]
                wts += w + [0.0] * ((-cap) % walign)
        ranges.append(len(desc) // 2)
    while len(wts) % 4:
        wts.append(0.0)
    out = [f"// rolled mel projection tables, {n_mels} filters, {int(np.count_nonzero(filt))} taps ({sum(i[2] for i in info)} padded), "
           f"class caps {caps}",
           f"#define {name.upper()}_NCLS {ncls}",
           f"#define {name.upper()}_CAPS {', '.join(str(c) for c in caps)}",
           f"#define {name.upper()}_NW {len(wts)}",
           emit_array(name + "_desc", np.array(desc, dtype=np.int64), per_line=8, ctype="int"),
           emit_array(name + "_w", np.array(wts, dtype=np.float64)),
           emit_array(name + "_ranges", np.array(ranges, dtype=np.int64), per_line=ncls + 1, ctype="int")]
    return "\n".join(out) + "\n"


def gen_mel_segments(name, filt, nwarps, group=4, extra_load=None):
    """tables for the segment form of a triangular filter bank in which every bin feeds at most two ADJACENT filters
    (kaldi's mel banks).  Segment j (j = -1 .. n_mels - 1) = the bins whose lowest non-zero filter is j; in it filter j
    falls (weight wf) and filter j + 1 rises (weight wr).  A lane accumulates `group` consecutive filters at a time:
    group g = filters group * g .. + group - 1 scans segments group * g - 1 .. group * g + group - 1.
      name_seg[j + 1]  = first bin of segment j (j = -1 .. n_mels; entry n_mels + 1 = end)
      name_w[2 * k]    = (wf, wr) of bin k
      name_groups[2 * w], [2 * w + 1] = the (at most two) groups warp w handles, -1 = none; longest-processing-time
                         assignment on the cost model bins * 6 + 44 per group (+ extra_load[w])."""
    n_mels, n_bins = filt.shape
    seg_of = np.full(n_bins, -2, dtype=np.int64)
    w2 = np.zeros((n_bins, 2), dtype=np.float64)
    for k in range(n_bins):
        nz = np.nonzero(filt[:, k])[0]
        assert len(nz) <= 2, "a bin feeds more than two filters"
        if len(nz) == 0:
            continue
        j = int(nz[0])
        if len(nz) == 2:
            assert nz[1] == j + 1, "the two filters of a bin must be adjacent"
            w2[k] = (filt[j, k], filt[j + 1, k])
        else:
            w2[k] = (filt[j, k], 0.0)
        seg_of[k] = j
    live = np.nonzero(seg_of > -2)[0]
    # bins of a segment are contiguous and segments ascend; bins before the first live one / dead gaps carry zero weights
    first = int(live[0])
    for k in range(first, n_bins):
        if seg_of[k] == -2:
            seg_of[k] = seg_of[k - 1]
    assert np.all(np.diff(seg_of[first:]) >= 0)
    seg = np.zeros(n_mels + 2, dtype=np.int64)            # seg[j + 1] = first bin of segment j
    for j in range(-1, n_mels + 1):
        ks = np.nonzero(seg_of[first:] >= j)[0]
        seg[j + 1] = first + int(ks[0]) if len(ks) else n_bins
    # rising part of filter 0 (segment -1) is empty when its left edge lies below the first live bin: fine
    # check: rebuilding the matrix from the tables gives the matrix
    chk = np.zeros_like(filt, dtype=np.float64)
    for j in range(-1, n_mels):
        for k in range(seg[j + 1], seg[j + 2]):
            if j >= 0:
                chk[j, k] += w2[k, 0]
            if j + 1 < n_mels:
                chk[j + 1, k] += w2[k, 1] if seg_of[k] == j else 0.0
    assert np.array_equal(chk.astype(np.float32), filt.astype(np.float32)), "segment tables do not reproduce the filter bank"
    ngroups = n_mels // group
    assert ngroups * group == n_mels
    cost = [6 * int(seg[group * gi + group + 1] - seg[group * gi]) + 44 for gi in range(ngroups)]
    load = list(extra_load) if extra_load is not None else [0.0] * nwarps
    slots = [[] for _ in range(nwarps)]
    for gi in sorted(range(ngroups), key=lambda q: -cost[q]):
        cand = [w for w in range(nwarps) if len(slots[w]) < 2]
        w = min(cand, key=lambda q: load[q])
        slots[w].append(gi)
        load[w] += cost[gi]
    groups = []
    for w in range(nwarps):
        groups += sorted(slots[w]) + [-1] * (2 - len(slots[w]))
    out = [f"// segment tables of a {n_mels}-filter bank over {n_bins} bins ({int(np.count_nonzero(filt))} non-zeros); groups of {group} filters; "
           f"per-warp cost {min(load):.0f} .. {max(load):.0f}",
           f"#define {name.upper()}_GROUP {group}",
           emit_array(name + "_seg", seg, per_line=16, ctype="int"),
           emit_array(name + "_w", w2).replace("FE_CONSTANT float", "FE_CONSTANT __attribute__((aligned(8))) float"),
           emit_array(name + "_groups", np.array(groups, dtype=np.int64), per_line=16, ctype="int")]
    return "\n".join(out) + "\n"


def gen_melrolled(name, filt, n_bins):
    """tables for the rolled mel projection of the one-frame-per-lane kernels (kaldi, CosyVoice): per filter
    name_desc[3*m] = first tap k0, [3*m+1] = taps padded to a multiple of 4, [3*m+2] = offset of its weights in
    name_w (zero-padded; the padded taps stay inside [0, n_bins))."""
    desc, wts = [], []
    for m in range(filt.shape[0]):
        nz = np.nonzero(filt[m])[0]
        if len(nz) == 0:
            k0, n = 0, 0
        else:
            k0, n = int(nz[0]), int(nz[-1] - nz[0] + 1)
            assert np.all(filt[m, k0:k0 + n] != 0), "taps of a filter must be contiguous"
        n4 = max(4, (n + 3) // 4 * 4)
        k0 = min(k0, n_bins - n4)
        assert k0 >= 0 and np.count_nonzero(filt[m, :k0]) == 0 and np.count_nonzero(filt[m, k0 + n4:]) == 0
        desc += [k0, n4, len(wts)]
        wts += [float(v) for v in filt[m, k0:k0 + n4]]
    out = [f"// rolled mel projection tables, {filt.shape[0]} filters, {int(np.count_nonzero(filt))} taps ({len(wts)} padded)",
           emit_array(name + "_desc", np.array(desc, dtype=np.int64), per_line=9, ctype="int"),
           emit_array(name + "_w", np.array(wts, dtype=np.float64))]
    return "\n".join(out) + "\n"


def twiddle_table(N, rows, cols):
    """float2 table T[row][col] = exp(-2*pi*i*row*col/N), rows x cols."""
    t = np.zeros((rows, cols, 2), dtype=np.float64)
    for r in range(rows):
        for c in range(cols):
            wr, wi = g.twiddle(r * c, N)
            t[r, c] = (wr, wi)
    return t


def main():
    os.makedirs(OUT, exist_ok=True)
    # ---------------- codelets
    cl = ["// GENERATED by tools/gen_kernels.py -- do not edit.", "#pragma once", ""]
    # Whisper: N = 400 = 20 x 20
    cl.append(gen_codelet("fe_rdft20_h11", 20, real_in=True, outs=range(11), real_out=(0, 10)))
    cl.append(gen_codelet("fe_cdft20", 20))
    cl.append(gen_codelet("fe_rdft20_h10", 20, real_in=True, outs=range(10), real_out=(0,)))
    cl.append(gen_codelet("fe_cdft20_h10", 20, outs=range(10)))
    # kaldi: N = 512 = 32 x 16, 400 samples -> n1 < 25 non-zero
    cl.append(gen_codelet("fe_rdft32z25_h17", 32, real_in=True, nz=25, outs=range(17), real_out=(0, 16)))
    cl.append(gen_codelet("fe_cdft16", 16))
    cl.append(gen_codelet("fe_rdft16_h8", 16, real_in=True, outs=range(8), real_out=(0,)))
    cl.append(gen_codelet("fe_cdft16_h8", 16, outs=range(8)))
    # CosyVoice: N = 1920 = 16 x 12 x 10
    cl.append(gen_codelet("fe_rdft16_h9", 16, real_in=True, outs=range(9), real_out=(0, 8)))
    cl.append(gen_codelet("fe_cdft12", 12))
    cl.append(gen_codelet("fe_cdft10", 10))
    # Wav2Lip / Diff2Lip: N = 800 = 16 x 5 x 10 (shares fe_rdft16_h9 and fe_cdft10 with CosyVoice)
    cl.append(gen_codelet("fe_cdft5", 5))
    # spectral gate (fe_gate.cuh): N = 1024 = 16 x 8 x 8 (forward: fe_rdft16_h9 first; inverse: complex fe_cdft16 first)
    cl.append(gen_codelet("fe_cdft8", 8))
    with open(os.path.join(OUT, "fe_codelets.h"), "w") as f:
        f.write("\n".join(cl))

    # ---------------- tables
    tb = ["// GENERATED by tools/gen_kernels.py -- do not edit.", "#pragma once", ""]
    n = np.arange(400, dtype=np.float64)
    tb.append("// periodic Hann, N = 400 (torch.hann_window default)")
    tb.append(emit_array("fe_win_whisper", 0.5 - 0.5 * np.cos(2 * np.pi * n / 400.0)))
    tb.append("// povey = symmetric Hann ^ 0.85, N = 400 (torchaudio kaldi.py:100)")
    tb.append(emit_array("fe_win_kaldi", (0.5 - 0.5 * np.cos(2 * np.pi * n / 399.0)) ** 0.85))
    n = np.arange(1920, dtype=np.float64)
    tb.append("// periodic Hann, N = 1920")
    tb.append(emit_array("fe_win_cosy", 0.5 - 0.5 * np.cos(2 * np.pi * n / 1920.0)))
    tb.append("// Whisper stage-A twiddles W_400^(n2*k1), [n2 = 0..19][k1 = 0..10] (re, im)")
    tb.append(emit_array("fe_tw_whisper", twiddle_table(400, 20, 11)))
    # packed Whisper kernel: per n2 one row of 40 floats = window w[20*n1 + n2] (n1 = 0..19), then the
    # stage-A twiddles W_400^(n2*k1) as (re, im) for k1 = 1..10
    w400 = 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(400, dtype=np.float64) / 400.0)
    tw = twiddle_table(400, 20, 11)
    rows = np.zeros((20, 40), dtype=np.float64)
    for n2 in range(20):
        rows[n2, :20] = w400[20 * np.arange(20) + n2]
        rows[n2, 20:] = tw[n2, 1:11].reshape(-1)
    tb.append("// Whisper (frame-pair kernel) stage-A constants [n2][window(n1 = 0..19) | W_400^(n2*k1) (re, im), k1 = 1..10]")
    tb.append(emit_array("fe_w2_tab", rows))
    tb.append("// kaldi stage-A twiddles W_512^(n2*k1), [n2 = 0..15][k1 = 0..16]")
    tb.append(emit_array("fe_tw_kaldi", twiddle_table(512, 16, 17)))
    tb.append("// CosyVoice stage-A twiddles W_1920^(m*k1), [m = 0..119][k1 = 0..8]")
    tb.append(emit_array("fe_tw_cosy_a", twiddle_table(1920, 120, 9)))
    tb.append("// CosyVoice stage-B twiddles W_120^(n3*k2), [n3 = 0..9][k2 = 0..11]")
    tb.append(emit_array("fe_tw_cosy_b", twiddle_table(120, 10, 12)))
    w80 = slaney(16000, 400, 80, 0.0, 8000.0)
    w128 = slaney(16000, 400, 128, 0.0, 8000.0)
    cz = slaney(24000, 1920, 80, 0.0, 8000.0)
    kb = kaldi_banks()
    # compute warps 5, 8, 9, 10 run one phase-A task per unit, the others two (fe_w2_a_second): they take
    # a larger share of the mel projection so that all warps reach the unit's barrier together
    mel_share = [1.0] * 12
    for w_ in (5, 8, 9, 10):
        mel_share[w_] = 1.8
    tb.append(gen_melloop("fe_melloop_w80", w80, 12, (2, 3, 4, 6, 8, 11, 14), chunk_weight=mel_share))
    tb.append(gen_melloop("fe_melloop_w128", w128, 12, (1, 2, 3, 4, 6, 7, 9), chunk_weight=mel_share))
    # Tensor-Memory Whisper kernel (fe_whisper3.cuh): per phase-A task i (sub-sequences n2 = 4 i + c, c = 0..3) 160 floats:
    # window w[20 n1 + 4 i + c] as [n1][c], then W_400^(n2 k1), k1 = 1..10, as [k1][wr_c0 wr_c1 wi_c0 wi_c1 | wr_c2 wr_c3 wi_c2 wi_c3]
    rows3 = np.zeros((5, 160), dtype=np.float64)
    for i in range(5):
        for n1 in range(20):
            for c in range(4):
                rows3[i, 4 * n1 + c] = w400[20 * n1 + 4 * i + c]
        for k1 in range(1, 11):
            for c in range(4):
                wr, wi = tw[4 * i + c, k1]
                base = 80 + 8 * (k1 - 1) + 4 * (c // 2)
                rows3[i, base + (c % 2)] = wr
                rows3[i, base + 2 + (c % 2)] = wi
    tb.append("// Whisper (Tensor-Memory kernel) phase-A constants, see tools/gen_kernels.py")
    tb.append(emit_array("fe_w3_tab", rows3).replace("FE_CONSTANT float fe_w3_tab", "FE_CONSTANT __attribute__((aligned(16))) float fe_w3_tab"))
    # 16 mel chunks (chunk = warp = 4 * member + team), equal shares: the phase-A / phase-B task map (fe_w3_a_hu,
    # fe_w3_b_task) already levels the members' totals
    share3 = [1.0] * 16
    tb.append(gen_melloop("fe_melloop3_w80", w80, 16, (2, 3, 4, 6, 8, 11, 14), chunk_weight=share3, k0_scale=512, emit_cost=22, by_class=True))
    tb.append(gen_melloop("fe_melloop3_w128", w128, 16, (1, 2, 3, 4, 6, 7, 9), chunk_weight=share3, k0_scale=512, emit_cost=22, by_class=True))
    # ---- Tensor-Memory kaldi kernel (fe_kaldi3.cuh): N = 512 = 32 (n1) x 16 (n2), half-unit hu = sub-sequences 2 hu, 2 hu + 1
    wk = (0.5 - 0.5 * np.cos(2 * np.pi * np.arange(400, dtype=np.float64) / 399.0)) ** 0.85
    k3win = np.zeros((8, 25, 2))
    k3tw = np.zeros((8, 15, 4))
    for hu in range(8):
        for n1 in range(25):
            for c in range(2):
                k3win[hu, n1, c] = wk[16 * n1 + 2 * hu + c]
        for k1 in range(1, 16):
            for c in range(2):
                wr, wi = g.twiddle((2 * hu + c) * k1, 512)
                k3tw[hu, k1 - 1, c] = wr
                k3tw[hu, k1 - 1, 2 + c] = wi
    tb.append("// kaldi (Tensor-Memory kernel) phase-A constants: povey window pairs [hu][n1][c], twiddles W_512^((2 hu + c) k1) as "
              "[hu][k1 - 1][wr_c0, wr_c1, wi_c0, wi_c1], and W_512^(16 n2) as [n2][re, im] (applied in phase B)")
    tb.append(emit_array("fe_k3_win", k3win).replace("FE_CONSTANT float", "FE_CONSTANT __attribute__((aligned(16))) float"))
    tb.append(emit_array("fe_k3_tw", k3tw).replace("FE_CONSTANT float", "FE_CONSTANT __attribute__((aligned(16))) float"))
    tb.append(emit_array("fe_k3_tw16", np.array([g.twiddle(16 * n2, 512) for n2 in range(16)])))
    tb.append(gen_mel_segments("fe_k3_mel", kb[:, :256], 16))
    tb.append(gen_melrolled("fe_melroll_cosy", cz, 640))
    tb.append(gen_melproj("fe_melproj_kaldi", kb, 16, 255))
    # ---- Wav2Lip / Diff2Lip mel (fe_cosy.cuh, FeLCfg): N = 800 = 16 x 5 x 10.  These tables live in global memory
    # (FE_TABLE): the 64 KB constant bank is full.
    def table(txt):
        return txt.replace("FE_CONSTANT", "FE_TABLE")
    n = np.arange(800, dtype=np.float64)
    tb.append("// periodic Hann, N = 800 (librosa.stft window='hann', fftbins=True)")
    tb.append(table(emit_array("fe_win_w2l", 0.5 - 0.5 * np.cos(2 * np.pi * n / 800.0))))
    tb.append("// Wav2Lip stage-A twiddles W_800^(m*k1), [m = 0..49][k1 = 0..8]")
    tb.append(table(emit_array("fe_tw_w2l_a", twiddle_table(800, 50, 9))))
    tb.append("// Wav2Lip stage-B twiddles W_50^(n3*k2), [n3 = 0..9][k2 = 0..4]")
    tb.append(table(emit_array("fe_tw_w2l_b", twiddle_table(50, 10, 5))))
    n = np.arange(1024, dtype=np.float64)
    tb.append("// spectral gate: periodic Hann N = 1024; W_1024^(m*k1) [m = 0..63][k1 = 0..8] (forward) and [k1 = 0..15] (inverse); W_64^(n3*k2)")
    tb.append(table(emit_array("fe_win_gate", 0.5 - 0.5 * np.cos(2 * np.pi * n / 1024.0))))
    tb.append(table(emit_array("fe_tw_gate_a", twiddle_table(1024, 64, 9))))
    tb.append(table(emit_array("fe_tw_gate_a16", twiddle_table(1024, 64, 16))))
    tb.append(table(emit_array("fe_tw_gate_b", twiddle_table(64, 8, 8))))
    wl = slaney(16000, 800, 80, 55.0, 7600.0)
    tb.append(table(gen_melrolled("fe_melroll_w2l", wl, 401)))
    with open(os.path.join(OUT, "fe_tables.h"), "w") as f:
        f.write("\n".join(tb))
    # the float32 tables the kernels bake, for the parity tests (tests/ compares them to the oracle's)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "baked_filters.npz"),
                        whisper80=w80, whisper128=w128, cosy80=cz, kaldi80=kb, w2l80=wl)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
