// Whisper log-mel featurizer, frame-pair kernel: phases of the persistent fused kernel.
//
// Replaces whisper/audio.py::log_mel_spectrogram (UPSTREAM openai-whisper; reached from
// /root/reference/Backend/services/cascaded_backend.py:158, Backend/speech_translation_pipeline.py:140,
// Backend/gtts_pipeline_fixed.py:64; same arithmetic as
// transformers/models/whisper/feature_extraction_whisper.py:135-164).
//
// Work unit = 64 consecutive frames of one clip.  Lane l of every warp owns the frame PAIR
// (t0 + l, t0 + 32 + l); the two frames of a pair travel through the whole pipeline side by side in one
// 64-bit register pair, so every floating-point instruction is a packed FADD2 / FMUL2 / FFMA2 (half the
// issue slots of the scalar formulation, constants broadcast by the instruction) and every shared-memory
// access moves both frames at once.  All per-task constants (window, twiddles, mel weights) are uniform
// across the warp: constant bank or immediates.
//
//   loader warps : waveform samples of the unit -> W[i] = (x[g0 + i], x[g0 + 5120 + i]), i < 5360, one pad
//                  slot per 160 samples (lane stride 161 pairs: conflict-free 64-bit accesses); reflect /
//                  zero padding resolved here.  Runs one unit ahead of the compute warps.
//   phase A      : 20 tasks (n2); real radix-20 DFT over n1 (n = 20*n1 + n2) with the Hann window fused into
//                  the loads, Hermitian half k1 = 0..10, times W_400^(n2*k1)  -> Ar/Ai[lane][k1][n2] (pairs)
//   phase B      : 11 tasks; k1 = 1..9: complex radix-20 DFT over n2, |X|^2 of bins k1 + 20*j and (mirrored)
//                  (20-k1) + 20*j written IN PLACE over the task's own row of the real plane; k1 = 0 (real
//                  input) and k1 = 10 (self-mirrored) are two lighter tasks.
//   phase M      : unrolled sparse mel projection (immediates), log10, (x + 4) / 4, coalesced stores, running
//                  min / max of the unit.
// The 1e-10 floor and the dynamic-range clamp max(x, max - 8) need the per-clip (or per-call) maximum: the
// kernel publishes the maximum (atomicMax) and each unit's minimum; a second, tiny kernel re-touches only the
// units whose minimum lies below the threshold (none for noise-like audio, the silent stretches for speech)
// and zero-fills frames past a ragged clip's end.  In the common case every output makes exactly one trip
// to HBM.
#pragma once
#include "fe_common.h"

struct FeWhisper2Params {
    const float* wave;         // (B, wave_stride) fp32 -- or int16 PCM when in_s16 (then a `const short*` in disguise)
    const int64_t* lengths;    // nullable, device: valid samples per clip
    int64_t wave_stride;
    int64_t L;                 // samples per clip when lengths == nullptr (also the clamp for lengths)
    int32_t B;
    int32_t padding;           // zeros appended to every clip before framing
    int32_t n_mels;            // 80 | 128
    int32_t norm_mode;         // FE_NORM_*
    void* out;                 // (B, n_mels, T_out), float32 or float16 (the kernel's OUT16 template argument)
    int64_t out_batch_stride;  // elements between clips in `out`
    int64_t T_out;             // frames per output row (row stride)
    int32_t units_per_clip;    // ceil(T_out / 64)
    int32_t total_units;       // B * units_per_clip
    int32_t vec_ok;            // wave base 16B-aligned and wave_stride % 4 == 0
    int32_t in_s16;            // 1: wave holds int16 PCM, the loader scales by 1/32768 (whisper.audio.load_audio's cast)
    int32_t* max_keys;         // (B) or (1): order-preserving keys of the running max of (x + 4) / 4
    float2* unit_minmax;       // (total_units): (min, max) of the unit's live outputs
    // windowed output (Tensor-Memory kernel only; 0 = off): out is (B, n_windows, n_mels, win_frames), the encoder-ready
    // segments whisper.transcribe cuts with mel[:, seek : seek + N_FRAMES] + pad_or_trim
    int32_t win_frames;
    int32_t n_windows;
    int32_t use_tmap;          // Tensor-Memory kernel: the launch carries a tensor map of the waveform (box copies)
    int32_t tmap_rows;         // rows per clip in that map
    int32_t step_b, step_w;    // Tensor-Memory kernel: gridDim.x units = step_w unit positions + step_b clips
};

struct FeW2Cfg {
    static constexpr int N = 400, HOP = 160, PADL = 200, R = 20;
    static constexpr int LANES = 32, FT = 64;            // frame pairs / frames per unit
    static constexpr int SEG = LANES * HOP;              // sample distance between the frames of a pair (5120)
    static constexpr int NCW = 12, NLW = 4;              // compute / loader warps.  (Measured: 12 + 5 and 15 + 5
                                                         // warps cap the kernel at 96 registers and are slower.)
    static constexpr int NCT = NCW * 32, NLT = NLW * 32, NT = NCT + NLT;
    static constexpr int RT = (LANES - 1) * HOP + N;     // samples per segment of a unit (5360)
    static constexpr int LS = HOP + 1;                   // lane stride in W (pairs)
    static constexpr int SW = RT + RT / HOP + 3;         // padded W (pairs)
    static constexpr int FS = 221;                       // (re pair, im pair) elements per lane in A (11*20 + 1 -> odd)
    static constexpr int PSZ = 201;                      // pairs per lane in P (200 bins + 1 -> odd)
    static constexpr int OFF_AR = SW * 2;                // in floats: A[lane][20*k1 + n2] = (re pair, im pair)
    static constexpr int OFF_P = OFF_AR + LANES * FS * 4;    // power spectrum P[lane][bin] (pairs); own buffer so
                                                             // that phase A of the next unit can overlap B and M
    static constexpr int OFF_MISC = OFF_P + LANES * PSZ * 2; // ints: [0] max key [1] min key [4..7], [8..11] unit info
    static constexpr int OFF_MBAR = OFF_MISC + 12;           // two 8-byte mbarriers: [0] "B(u) has read its rows of
                                                             // the A planes", [1] "P(u) is complete"
    // mel tables, copied once per CTA (the mel loop reads them with per-filter offsets):
    static constexpr int OFF_MDESC = OFF_MISC + 16;          // fe_melloop_*_desc (2 ints per filter, <= 128 filters)
    static constexpr int OFF_MW = OFF_MDESC + 256;           // fe_melloop_*_w (padded taps)
    static constexpr int MW_MAX = 576;
    static constexpr int SMEM_FLOATS = OFF_MW + MW_MAX;
    static_assert(OFF_MBAR % 2 == 0, "mbarriers need 8-byte alignment");
    static_assert(OFF_AR % 4 == 0, "A is accessed with 128-bit loads and stores");
    static_assert(OFF_MDESC % 4 == 0 && OFF_MW % 4 == 0, "tables are read with 128-bit loads");
    static_assert(SMEM_FLOATS * 4 <= 232448, "more than 227 KB of shared memory");
    static constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
    static constexpr float LOG_SCALE = 0.07525749891599530f;  // log10(2) / 4
    static constexpr float Y_FLOOR = -1.5f;                   // (log10(1e-10) + 4) / 4
};

struct FeW2Unit {
    int b;          // clip
    int t0;         // first frame of the unit
    int T;          // frames of this clip (<= T_out)
    int64_t len;    // valid samples
    int64_t lp;     // len + padding
};

FE_HD FE_INLINE FeW2Unit fe_w2_unit(const FeWhisper2Params& p, int u) {
    FeW2Unit t;
    t.b = u / p.units_per_clip;
    t.t0 = (u - t.b * p.units_per_clip) * FeW2Cfg::FT;
    t.len = fe_clip_len(p.lengths, t.b, p.L);
    t.lp = t.len + p.padding;
    int64_t T = t.lp / FeW2Cfg::HOP;
    t.T = (int)(T < p.T_out ? T : p.T_out);
    return t;
}

// ---------------------------------------------------------------- loader
// ltid in [0, NLT).  One sample of both segments per step: a warp reads two coalesced 128-byte rows and writes
// 32 consecutive pairs (conflict-free 64-bit stores; 128-bit loads would make the interleaving stores 4-way
// conflicted).
FE_HD FE_INLINE bool fe_w2_interior(const FeWhisper2Params& p, const FeW2Unit& t) {
    typedef FeW2Cfg C;
    const int64_t g0 = (int64_t)t.t0 * C::HOP - C::PADL;
    return g0 >= 0 && g0 + C::SEG + C::RT <= t.len;
}

// Interior units (fully inside the clip) are register-staged: fe_w2_load_issue puts every load of the unit in flight (it runs BEFORE the loader
// waits for W to be released); fe_w2_load_store writes the pairs once W is free (conflict-free 64-bit stores)
struct FeW2Stage {
    static constexpr int U = (FeW2Cfg::RT + FeW2Cfg::NLT - 1) / FeW2Cfg::NLT;  // samples per loader thread and segment
    float a[U], b[U];
};

// bit moves between int and float registers (host: memcpy)
FE_HD FE_INLINE float fe_int_bits_as_float(int v) {
#if defined(__CUDA_ARCH__)
    return __int_as_float(v);
#else
    float f; memcpy(&f, &v, 4); return f;
#endif
}
FE_HD FE_INLINE int fe_float_bits_as_int(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    int v; memcpy(&v, &f, 4); return v;
#endif
}
// int16 PCM -> float exactly as whisper.audio.load_audio does it: astype(float32) / 32768.0 (a power of two: exact)
// Without I2F (quarter-rate conversion pipe, on the loader's critical path: measured +16 % kernel time): the bits
// 0x4B400000 + v are the float 1.5 * 2^23 + v, and one FMA maps that to v / 32768 exactly (12582912 / 32768 = 384).
FE_HD FE_INLINE float fe_pcm16_to_float(int v) {
    return fmaf(fe_int_bits_as_float(0x4B400000 + v), 1.0f / 32768.0f, -384.0f);
}

// int16 input (IN16, a template argument of the kernel so that the float32 kernel carries none of this code): the
// staged registers carry the raw sample as integer bits, the conversion happens at the store
template <bool IN16 = false>
FE_HD FE_INLINE void fe_w2_load_issue(const FeWhisper2Params& p, const FeW2Unit& t, int ltid, FeW2Stage& r) {
    typedef FeW2Cfg C;
    const int64_t off = (int64_t)t.b * p.wave_stride + ((int64_t)t.t0 * C::HOP - C::PADL) + ltid;
    if (IN16) {
        const short* __restrict__ xs = reinterpret_cast<const short*>(p.wave) + off;
#pragma unroll
        for (int j = 0; j < FeW2Stage::U; ++j) {
            if (j + 1 < FeW2Stage::U || ltid + j * C::NLT < C::RT) {
                r.a[j] = fe_int_bits_as_float((int)xs[j * C::NLT]);
                r.b[j] = fe_int_bits_as_float((int)xs[j * C::NLT + C::SEG]);
            }
        }
        return;
    }
    const float* __restrict__ xa = p.wave + off;
#pragma unroll
    for (int j = 0; j < FeW2Stage::U; ++j) {
        if (j + 1 < FeW2Stage::U || ltid + j * C::NLT < C::RT) { r.a[j] = xa[j * C::NLT]; r.b[j] = xa[j * C::NLT + C::SEG]; }
    }
}

template <bool IN16 = false>
FE_HD FE_INLINE void fe_w2_load_store(float* sm, int ltid, const FeW2Stage& r) {
    typedef FeW2Cfg C;
    // sample i = ltid + NLT * j sits in slot i + i / HOP.  NLT * 5 = 4 * HOP, so for j = 5 * m + c the slot is
    // base_c + (4 * HOP + 4) * m with base_c = i_c + i_c / HOP, i_c = ltid + NLT * c: five bases per thread, every
    // other offset is an immediate
    static_assert(C::NLT * 5 == 4 * C::HOP, "slot arithmetic of the loader");
    fe2* __restrict__ w = reinterpret_cast<fe2*>(sm);
    fe2* base[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) {
        const unsigned ic = (unsigned)(ltid + C::NLT * c);
        base[c] = w + ic + ic / (unsigned)C::HOP;
    }
    if (IN16) {
#pragma unroll
        for (int j = 0; j < FeW2Stage::U; ++j) {
            if (j + 1 < FeW2Stage::U || ltid + j * C::NLT < C::RT)
                base[j % 5][(4 * C::HOP + 4) * (j / 5)] =
                    fe_mk2(fe_pcm16_to_float(fe_float_bits_as_int(r.a[j])), fe_pcm16_to_float(fe_float_bits_as_int(r.b[j])));
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < FeW2Stage::U; ++j) {
        if (j + 1 < FeW2Stage::U || ltid + j * C::NLT < C::RT)
            base[j % 5][(4 * C::HOP + 4) * (j / 5)] = fe_mk2(r.a[j], r.b[j]);
    }
}

// edge units (reflect / zero padding, ragged ends): element-wise, after W is free.  One sample of each segment
// per step like the fast path (coalesced reads, conflict-free pair slots), issued as 4-byte cp.async so that the
// copies of the whole unit are in flight at once (a rolled loop of dependent loads would pay the HBM latency
// 42 times); samples beyond the clip are zero-filled by a zero source size.
template <class I>
FE_HD FE_INLINE I fe_w2_reflect(I g, I lp) {
    if (g < 0) g = -g;
    if (g >= lp) g = 2 * (lp - 1) - g;
    if (g < 0) g = 0;               // degenerate clips shorter than the pad: stay in bounds
    if (g >= lp) g = lp - 1;
    return g;
}

template <class I>
FE_HD FE_INLINE void fe_w2_load_edge_t(const float* __restrict__ x, I g0, I len, I lp, float* sm, int ltid) {
    typedef FeW2Cfg C;
    fe2* __restrict__ w = reinterpret_cast<fe2*>(sm);
#if defined(__CUDA_ARCH__)
    const unsigned w0 = (unsigned)__cvta_generic_to_shared(w);
#pragma unroll 1
    for (int i = ltid; i < C::RT; i += C::NLT) {
        const unsigned d = w0 + 8u * (unsigned)(i + i / C::HOP);
        const I ga = fe_w2_reflect<I>(g0 + i, lp), gb = fe_w2_reflect<I>(g0 + C::SEG + i, lp);
        const unsigned na = ga < len ? 4u : 0u, nb = gb < len ? 4u : 0u;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(x + (na ? ga : 0)), "r"(na) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d + 4u), "l"(x + (nb ? gb : 0)), "r"(nb) : "memory");
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
#else
    for (int i = ltid; i < C::RT; i += C::NLT) {
        const I ga = fe_w2_reflect<I>(g0 + i, lp), gb = fe_w2_reflect<I>(g0 + C::SEG + i, lp);
        w[i + i / C::HOP] = fe_mk2(ga < len ? x[ga] : 0.0f, gb < len ? x[gb] : 0.0f);
    }
#endif
}

// int16 edge units: cp.async cannot convert, so plain loads in batches of four independent pairs per thread
template <class I>
FE_HD FE_INLINE void fe_w2_load_edge_s16_t(const short* __restrict__ x, I g0, I len, I lp, float* sm, int ltid) {
    typedef FeW2Cfg C;
    fe2* __restrict__ w = reinterpret_cast<fe2*>(sm);
    constexpr int NB = 4;
#pragma unroll 1
    for (int i0 = ltid; i0 < C::RT; i0 += NB * C::NLT) {
        int va[NB], vb[NB];
#pragma unroll
        for (int k = 0; k < NB; ++k) {
            const int i = i0 + k * C::NLT;
            va[k] = 0; vb[k] = 0;
            if (i < C::RT) {
                const I ga = fe_w2_reflect<I>(g0 + i, lp), gb = fe_w2_reflect<I>(g0 + C::SEG + i, lp);
                if (ga < len) va[k] = x[ga];
                if (gb < len) vb[k] = x[gb];
            }
        }
#pragma unroll
        for (int k = 0; k < NB; ++k) {
            const int i = i0 + k * C::NLT;
            if (i < C::RT) w[i + i / C::HOP] = fe_mk2(fe_pcm16_to_float(va[k]), fe_pcm16_to_float(vb[k]));
        }
    }
}

// Edge units, register-staged like interior ones: every (reflected, bounds-checked) load of the unit goes in flight
// before the loader waits for W; fe_w2_load_store then writes them.  (The cp.async form below starts its copies only
// once W is free: 1.51 -> 1.47 ms for 1024 x 30 s, two edge units per 47-unit clip.)  32-bit index arithmetic: the
// caller checks that the clip is shorter than 2^30 samples (longer clips take fe_w2_load_edge).
FE_HD FE_INLINE bool fe_w2_edge_fits_int(const FeW2Unit& t) {
    return t.lp < ((int64_t)1 << 30) && (int64_t)t.t0 * FeW2Cfg::HOP < ((int64_t)1 << 30);
}
template <bool IN16>
FE_HD FE_INLINE void fe_w2_load_issue_edge(const FeWhisper2Params& p, const FeW2Unit& t, int ltid, FeW2Stage& r) {
    typedef FeW2Cfg C;
    const short* __restrict__ xs = reinterpret_cast<const short*>(p.wave) + (int64_t)t.b * p.wave_stride;
    const float* __restrict__ xf = p.wave + (int64_t)t.b * p.wave_stride;
    const int g0 = t.t0 * C::HOP - C::PADL + ltid, len = (int)t.len, lp = (int)t.lp;
#pragma unroll
    for (int j = 0; j < FeW2Stage::U; ++j) {
        if (j + 1 < FeW2Stage::U || ltid + j * C::NLT < C::RT) {
            const int ga = fe_w2_reflect<int>(g0 + j * C::NLT, lp), gb = fe_w2_reflect<int>(g0 + j * C::NLT + C::SEG, lp);
            if (IN16) {
                r.a[j] = fe_int_bits_as_float(ga < len ? (int)xs[ga] : 0);
                r.b[j] = fe_int_bits_as_float(gb < len ? (int)xs[gb] : 0);
            } else {
                r.a[j] = ga < len ? xf[ga] : 0.0f;
                r.b[j] = gb < len ? xf[gb] : 0.0f;
            }
        }
    }
}

// how the loader fills W for a unit: 0 nothing (no live frame), 1 interior + register-staged, 2 edge + register-staged,
// 3 edge of a clip of 2^30 samples or more, element-wise after W is free
template <bool IN16>
FE_HD FE_INLINE int fe_w2_load_mode(const FeWhisper2Params& p, const FeW2Unit& t) {
    if (!(t.t0 < t.T)) return 0;
    if (fe_w2_interior(p, t)) return 1;
    return fe_w2_edge_fits_int(t) ? 2 : 3;
}
template <bool IN16>
FE_HD FE_INLINE void fe_w2_load_ahead(const FeWhisper2Params& p, const FeW2Unit& t, int ltid, int mode, FeW2Stage& r) {
    if (mode == 1) fe_w2_load_issue<IN16>(p, t, ltid, r);
    else if (mode == 2) fe_w2_load_issue_edge<IN16>(p, t, ltid, r);
}
template <bool IN16>
FE_HD FE_INLINE void fe_w2_load_finish(const FeWhisper2Params& p, const FeW2Unit& t, float* sm, int ltid, int mode,
                                       const FeW2Stage& r);

template <bool IN16 = false>
FE_HD FE_INLINE void fe_w2_load_edge(const FeWhisper2Params& p, const FeW2Unit& t, float* sm, int ltid) {
    typedef FeW2Cfg C;
    const int64_t g0 = (int64_t)t.t0 * C::HOP - C::PADL;
    if (IN16) {
        const short* __restrict__ xs = reinterpret_cast<const short*>(p.wave) + (int64_t)t.b * p.wave_stride;
        if (t.lp < ((int64_t)1 << 30) && g0 < ((int64_t)1 << 30))
            fe_w2_load_edge_s16_t<int>(xs, (int)g0, (int)t.len, (int)t.lp, sm, ltid);
        else
            fe_w2_load_edge_s16_t<int64_t>(xs, g0, t.len, t.lp, sm, ltid);
        return;
    }
    const float* __restrict__ x = p.wave + (int64_t)t.b * p.wave_stride;
    if (t.lp < ((int64_t)1 << 30) && g0 < ((int64_t)1 << 30))
        fe_w2_load_edge_t<int>(x, (int)g0, (int)t.len, (int)t.lp, sm, ltid);
    else
        fe_w2_load_edge_t<int64_t>(x, g0, t.len, t.lp, sm, ltid);
}

template <bool IN16>
FE_HD FE_INLINE void fe_w2_load_finish(const FeWhisper2Params& p, const FeW2Unit& t, float* sm, int ltid, int mode,
                                       const FeW2Stage& r) {
    if (mode == 1 || mode == 2) fe_w2_load_store<IN16>(sm, ltid, r);
    else if (mode == 3) fe_w2_load_edge<IN16>(p, t, sm, ltid);
}

// copies the constant tables into shared memory; every thread of the CTA calls it, a barrier must follow
template <int NMELS>
FE_HD FE_INLINE void fe_w2_copy_tables(float* sm, int tid, int nthreads) {
    typedef FeW2Cfg C;
    static_assert(FE_MELLOOP_W80_NW <= C::MW_MAX && FE_MELLOOP_W128_NW <= C::MW_MAX, "MW_MAX too small");
    const int* __restrict__ desc = NMELS == 80 ? fe_melloop_w80_desc : fe_melloop_w128_desc;
    const float* __restrict__ w = NMELS == 80 ? fe_melloop_w80_w : fe_melloop_w128_w;
    const int nw = NMELS == 80 ? FE_MELLOOP_W80_NW : FE_MELLOOP_W128_NW;
    int* d = reinterpret_cast<int*>(sm + C::OFF_MDESC);
    for (int i = tid; i < 2 * NMELS; i += nthreads) d[i] = desc[i];
    for (int i = tid; i < nw; i += nthreads) sm[C::OFF_MW + i] = w[i];
}

struct FeNoHook { FE_HD FE_INLINE void operator()() const {} };

// Task map of the compute warps (one CTA-wide barrier per unit, see fe_whisper2_kernel):
//   phase B: warp 0 -> k1 = 0, warps 1..9 -> k1 = warp, warp 10 -> k1 = 10, warp 11 none
//   phase A: first task n2 = warp; a second task for warps 0-4, 6, 7, 11 (n2 = 12..19), chosen so that the four
//            SM sub-partitions (warp % 4) carry about the same FP32 work over A + B
//   phase M: chunk = warp (generated/fe_tables.h: fe_melloop_*_ranges)
FE_HD FE_INLINE int fe_w2_a_second(int warp) {
    return warp < 5 ? warp + 12 : (warp == 6 ? 17 : (warp == 7 ? 18 : (warp == 11 ? 19 : -1)));
}

// A buffer, interleaved: A[lane][20*k1 + n2] = (re pair, im pair), 16 bytes; lane stride FS elements (odd multiple
// of 16 bytes: conflict-free 128-bit accesses).  One STS.128 / LDS.128 moves what two 64-bit accesses moved with
// separate planes: same wavefronts, half the shared-memory instructions in the queue.
FE_HD FE_INLINE float4* fe_w2_a_ptr(float* sm, int lane) {
    return reinterpret_cast<float4*>(sm + FeW2Cfg::OFF_AR) + FeW2Cfg::FS * lane;
}
FE_HD FE_INLINE void fe_w2_a_store(float4* q, fe2 re, fe2 im) { *q = make_float4(re.x, re.y, im.x, im.y); }
FE_HD FE_INLINE void fe_w2_a_load(const float4* q, fe2& re, fe2& im) {
    const float4 v = *q;
    re = fe_mk2(v.x, v.y); im = fe_mk2(v.z, v.w);
}

// ---------------------------------------------------------------- phase A
template <class PreStore>
FE_HD FE_INLINE void fe_w2_stage_a(float* sm, int lane, int n2, PreStore pre_store) {
    typedef FeW2Cfg C;
    const fe2* __restrict__ s = reinterpret_cast<const fe2*>(sm) + C::LS * lane + n2;
    // window and twiddles of this task: constant bank (register-indexed LDC.64; measured faster here than a
    // shared-memory copy, which costs ~200 wavefronts of the already busy shared-memory pipe per unit)
    const float* __restrict__ tab = fe_w2_tab + 40 * n2;
    fe2 x[C::R];
#pragma unroll
    for (int n1 = 0; n1 < C::R; ++n1) {
        // sample n = 20*n1 + n2 of this lane's frames; (20*n1 + n2) / 160 == n1 / 8 for every n2 < 20
        x[n1] = fe_mulc(tab[n1], s[C::R * n1 + n1 / (C::HOP / C::R)]);
    }
    fe2 yr[11], yi[11];
    fe_rdft20_h11(x, yr, yi);
    float4* __restrict__ a = fe_w2_a_ptr(sm, lane) + n2;
    pre_store();   // the A buffer may still be read by phase B of the previous unit
    *reinterpret_cast<fe2*>(a) = yr[0];                              // k1 = 0 is real: only the first half is read
#pragma unroll
    for (int k1 = 1; k1 < 10; ++k1) {
        const float wr = tab[18 + 2 * k1], wi = tab[19 + 2 * k1];
        fe_w2_a_store(a + C::R * k1, fe_fnmac(wi, yi[k1], fe_mulc(wr, yr[k1])),    // wr*yr - wi*yi
                      fe_fmac(wi, yr[k1], fe_mulc(wr, yi[k1])));                   // wr*yi + wi*yr
    }
    fe_w2_a_store(a + C::R * 10, fe_mulc(tab[38], yr[10]), fe_mulc(tab[39], yr[10]));
}

// all phase-A tasks of one warp, in order (prologue of the kernel; host emulation)
FE_HD FE_INLINE void fe_w2_phase_a(float* sm, int lane, int warp) {
    fe_w2_stage_a(sm, lane, warp, FeNoHook());
    const int n2b = fe_w2_a_second(warp);
    if (n2b >= 0) fe_w2_stage_a(sm, lane, n2b, FeNoHook());
}

// ---------------------------------------------------------------- phase B
// Power of bin k (k < 200) of this lane's pair goes to P[lane][k] (lane stride PSZ pairs, odd).
template <class PostLoad>
FE_HD FE_INLINE void fe_w2_stage_b_heavy(float* sm, int lane, int k1, PostLoad post_load) {
    typedef FeW2Cfg C;
    const float4* __restrict__ a = fe_w2_a_ptr(sm, lane) + C::R * k1;
    fe2 xr[20], xi[20], yr[20], yi[20];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) fe_w2_a_load(a + n2, xr[n2], xi[n2]);
    post_load();
    fe_cdft20(xr, xi, yr, yi);
    fe2* __restrict__ pa = reinterpret_cast<fe2*>(sm + C::OFF_P) + C::PSZ * lane + k1;          // bins k1 + 20*j
    fe2* __restrict__ pb = reinterpret_cast<fe2*>(sm + C::OFF_P) + C::PSZ * lane + (C::R - k1);  // bins (20-k1) + 20*j
#pragma unroll
    for (int j = 0; j < 10; ++j) {
        pa[C::R * j] = fe_fma(yr[j], yr[j], fe_mul(yi[j], yi[j]));
        pb[C::R * j] = fe_fma(yr[19 - j], yr[19 - j], fe_mul(yi[19 - j], yi[19 - j]));
    }
}

template <class PostLoad>
FE_HD FE_INLINE void fe_w2_stage_b_k0(float* sm, int lane, PostLoad post_load) {   // k1 = 0: real input, bins 20*j, j = 0..9
    typedef FeW2Cfg C;
    const float4* __restrict__ a = fe_w2_a_ptr(sm, lane);
    fe2 xr[20], yr[10], yi[10];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) xr[n2] = *reinterpret_cast<const fe2*>(a + n2);
    post_load();
    yi[0] = fe_zero<fe2>();
    fe_rdft20_h10(xr, yr, yi);
    fe2* __restrict__ pa = reinterpret_cast<fe2*>(sm + C::OFF_P) + C::PSZ * lane;
#pragma unroll
    for (int j = 0; j < 10; ++j) pa[C::R * j] = fe_fma(yr[j], yr[j], fe_mul(yi[j], yi[j]));
}

template <class PostLoad>
FE_HD FE_INLINE void fe_w2_stage_b_k10(float* sm, int lane, PostLoad post_load) {  // k1 = 10: bins 10 + 20*j, j = 0..9
    typedef FeW2Cfg C;
    const float4* __restrict__ a = fe_w2_a_ptr(sm, lane) + C::R * 10;
    fe2 xr[20], xi[20], yr[10], yi[10];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) fe_w2_a_load(a + n2, xr[n2], xi[n2]);
    post_load();
    fe_cdft20_h10(xr, xi, yr, yi);
    fe2* __restrict__ pa = reinterpret_cast<fe2*>(sm + C::OFF_P) + C::PSZ * lane + 10;
#pragma unroll
    for (int j = 0; j < 10; ++j) pa[C::R * j] = fe_fma(yr[j], yr[j], fe_mul(yi[j], yi[j]));
}

// 11 tasks: warp 0 -> k1 = 0, warps 1..9 -> k1 = warp, warp 10 -> k1 = 10, warp 11 idle
// post_load runs once per warp, after the task's rows of the A planes are in registers (also for warp 11)
template <class PostLoad>
FE_HD FE_INLINE void fe_w2_phase_b(float* sm, int lane, int warp, PostLoad post_load) {
    if (warp == 0) fe_w2_stage_b_k0(sm, lane, post_load);
    else if (warp < 10) fe_w2_stage_b_heavy(sm, lane, warp, post_load);
    else if (warp == 10) fe_w2_stage_b_k10(sm, lane, post_load);
    else post_load();
}
FE_HD FE_INLINE void fe_w2_phase_b(float* sm, int lane, int warp) { fe_w2_phase_b(sm, lane, warp, FeNoHook()); }

// ---------------------------------------------------------------- phase M
FE_HD FE_INLINE float fe_max3(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
#else
    return fmaxf(fmaxf(a, b), c);
#endif
}
FE_HD FE_INLINE float fe_min3(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
#else
    return fminf(fminf(a, b), c);
#endif
}
// base + m * row_bytes in ONE instruction (IMAD.WIDE.U32 with the immediate m); row_bytes < 2^26
FE_HD FE_INLINE char* fe_row_ptr(char* base, unsigned row_bytes, unsigned m) {
#if defined(__CUDA_ARCH__)
    char* r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(row_bytes), "r"(m), "l"(base));
    return r;
#else
    return base + (size_t)row_bytes * m;
#endif
}
// global stores through a pointer built by fe_row_ptr: STG instead of a generic ST, and no "memory" clobber --
// the output is write-only for this kernel, so shared-memory loads of the next filter may move above the store
template <int BYTE_OFF>
FE_HD FE_INLINE void fe_stg(char* q, float v) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.f32 [%0+%1], %2;" ::"l"(q), "n"(BYTE_OFF), "f"(v));
#else
    *reinterpret_cast<float*>(q + BYTE_OFF) = v;
#endif
}
// float32 -> IEEE binary16 bits, round to nearest even (what torch's .to(float16) does); the host form is the plain
// bit manipulation so that the g++ emulation needs no CUDA header
FE_HD FE_INLINE unsigned short fe_f2h_rn(float v) {
#if defined(__CUDA_ARCH__)
    unsigned short h;
    asm("cvt.rn.f16.f32 %0, %1;" : "=h"(h) : "f"(v));
    return h;
#else
    union { float f; unsigned u; } c; c.f = v;
    const unsigned sign = (c.u >> 16) & 0x8000u, a = c.u & 0x7fffffffu;
    if (a > 0x7f800000u) return (unsigned short)(sign | 0x7e00u);            // NaN
    if (a >= 0x477ff000u) return (unsigned short)(sign | 0x7c00u);           // >= 65520 (and inf) -> inf
    if (a < 0x33000001u) return (unsigned short)sign;                        // <= 2^-25 -> 0
    if (a < 0x38800000u) {                                                   // subnormal half
        const int shift = 126 - (int)(a >> 23);                             // 14..24
        const unsigned m = (a & 0x7fffffu) | 0x800000u;
        unsigned r = m >> shift;
        const unsigned rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (r & 1u))) ++r;
        return (unsigned short)(sign | r);
    }
    unsigned r = ((a - 0x38000000u) >> 13);
    const unsigned rem = a & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (r & 1u))) ++r;
    return (unsigned short)(sign | r);
#endif
}
template <int BYTE_OFF>
FE_HD FE_INLINE void fe_stg_h(char* q, float v) {
    const unsigned short h = fe_f2h_rn(v);
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.b16 [%0+%1], %2;" ::"l"(q), "n"(BYTE_OFF), "h"(h));
#else
    *reinterpret_cast<unsigned short*>(q + BYTE_OFF) = h;
#endif
}
// log2 without the 1e-10 floor: 0 (and flushed subnormals) -> -inf; the floor is applied with the clamp
FE_HD FE_INLINE float fe_w2_lg2(float v) {
#if defined(__CUDA_ARCH__)
    return fe_lg2_ftz(v);
#else
    return v < 1.1754944e-38f ? -INFINITY : log2f(v);
#endif
}

// Writes y = (log10(mel) + 4) / 4 for the live frames of the pair (NO floor, NO clamp: both are applied by
// the fix-up pass, which only touches units whose minimum needs them) and returns this thread's (min, max)
// over them; dead frames carry NaN through min/max (max.f32 / min.f32 return the non-NaN operand).
// Rolled form: the warp's filters are sorted into classes by tap count (zero-padded to the class cap), each
// class is a branch-free loop over descriptors in shared memory (generated/fe_tables.h: fe_melloop_*).
// The fully unrolled projection was ~2100 instructions -- with the FFT codelets far beyond the 32 KB
// instruction cache, and every warp ran a different part of it: the whole kernel starved on instruction fetch.
struct FeW2MelCtx {
    const char* Pl;      // this lane's row of power pairs
    const char* wtab;    // padded weights (shared memory)
    const int2* desc;    // descriptors (shared memory)
    char* obase;         // out + clip offset + first frame of the pair
    unsigned rs4;        // output row pitch in bytes
    fe2 one;             // 1 for live frames, NaN for dead ones
    bool la, lb;
    float mn, mx;
};

template <int CAP, bool OUT16>
FE_HD FE_INLINE void fe_w2_mel_class(FeW2MelCtx& c, int i0, int i1) {
    typedef FeW2Cfg C;
#pragma unroll(CAP <= 2 ? 4 : 2)
    for (int i = i0; i < i1; ++i) {
        const int2 d = c.desc[i];
        const fe2* __restrict__ P = reinterpret_cast<const fe2*>(c.Pl + (d.x & 0xffff));
        float w[(CAP + 3) / 4 * 4];
        if (CAP == 1) {
            w[0] = *reinterpret_cast<const float*>(c.wtab + d.y);
        } else if (CAP == 2) {
            const float2 v = *reinterpret_cast<const float2*>(c.wtab + d.y);
            w[0] = v.x; w[1] = v.y;
        } else {
            const float4* __restrict__ w4 = reinterpret_cast<const float4*>(c.wtab + d.y);
#pragma unroll
            for (int j = 0; j < (CAP + 3) / 4; ++j) {
                const float4 v = w4[j];
                w[4 * j] = v.x; w[4 * j + 1] = v.y; w[4 * j + 2] = v.z; w[4 * j + 3] = v.w;
            }
        }
        fe2 acc = fe_mulc(w[0], P[0]);
        if (CAP > 1) {
            fe2 acc2 = fe_mulc(w[1], P[1]);              // two chains: half the dependent latency
#pragma unroll
            for (int j = 2; j < CAP; ++j) {
                if (j & 1) acc2 = fe_fmac(w[j], P[j], acc2);
                else acc = fe_fmac(w[j], P[j], acc);
            }
            acc = fe_add(acc, acc2);
        }
        const fe2 y = fe_fmac(C::LOG_SCALE, fe_mk2(fe_w2_lg2(acc.x), fe_w2_lg2(acc.y)), c.one);   // (log10(v) + 4) / 4
        char* o = fe_row_ptr(c.obase, c.rs4, (unsigned)d.x >> 16);
        if (OUT16) {   // min/max below stay on the unrounded values: the fix-up decides in float32 like the reference
            if (c.la) fe_stg_h<0>(o, y.x);
            if (c.lb) fe_stg_h<2 * C::LANES>(o, y.y);
        } else {
            if (c.la) fe_stg<0>(o, y.x);
            if (c.lb) fe_stg<4 * C::LANES>(o, y.y);
        }
        c.mx = fe_max3(c.mx, y.x, y.y);
        c.mn = fe_min3(c.mn, y.x, y.y);
    }
}

// class caps of the generated tables, and the compile-time walk over the classes
template <int NMELS> struct FeW2MelCaps;
template <> struct FeW2MelCaps<80> {
    static constexpr int N = FE_MELLOOP_W80_NCLS;
    static FE_HD constexpr int cap(int k) { constexpr int c[N] = {FE_MELLOOP_W80_CAPS}; return c[k]; }
};
template <> struct FeW2MelCaps<128> {
    static constexpr int N = FE_MELLOOP_W128_NCLS;
    static FE_HD constexpr int cap(int k) { constexpr int c[N] = {FE_MELLOOP_W128_CAPS}; return c[k]; }
};
template <int NMELS, bool OUT16, int K, bool END = (K == FeW2MelCaps<NMELS>::N)>
struct FeW2MelRun {
    static FE_HD FE_INLINE void run(FeW2MelCtx& c, const int* r) {
        fe_w2_mel_class<FeW2MelCaps<NMELS>::cap(K), OUT16>(c, r[K], r[K + 1]);
        FeW2MelRun<NMELS, OUT16, K + 1>::run(c, r);
    }
};
template <int NMELS, bool OUT16, int K>
struct FeW2MelRun<NMELS, OUT16, K, true> {
    static FE_HD FE_INLINE void run(FeW2MelCtx&, const int*) {}
};

template <int NMELS, bool OUT16 = false>
FE_HD FE_INLINE fe2 fe_w2_phase_mel(const FeWhisper2Params& p, const FeW2Unit& t, float* sm, int lane, int warp) {
    typedef FeW2Cfg C;
    constexpr int NCLS = FeW2MelCaps<NMELS>::N;
    const int* __restrict__ ranges = (NMELS == 80 ? fe_melloop_w80_ranges : fe_melloop_w128_ranges) + (NCLS + 1) * warp;
    const int fa = t.t0 + lane, fb = fa + C::LANES;
    FeW2MelCtx c;
    c.la = fa < t.T; c.lb = fb < t.T;
    c.Pl = reinterpret_cast<const char*>(reinterpret_cast<const fe2*>(sm + C::OFF_P) + C::PSZ * lane);
    c.wtab = reinterpret_cast<const char*>(sm + C::OFF_MW);
    c.desc = reinterpret_cast<const int2*>(sm + C::OFF_MDESC);
    constexpr int ESZ = OUT16 ? 2 : 4;
    c.obase = reinterpret_cast<char*>(p.out) + ((int64_t)t.b * p.out_batch_stride + fa) * ESZ;
    c.rs4 = (unsigned)p.T_out * (unsigned)ESZ;     // row pitch in bytes; the host guarantees T_out < 2^24
    union { unsigned u; float f; } qnan;
    qnan.u = 0x7fc00000u;
    c.one = fe_mk2(c.la ? 1.0f : qnan.f, c.lb ? 1.0f : qnan.f);
    c.mn = 3.0e38f; c.mx = -3.0e38f;
    int r[NCLS + 1];
#pragma unroll
    for (int k = 0; k <= NCLS; ++k) r[k] = ranges[k];
    FeW2MelRun<NMELS, OUT16, 0>::run(c, r);
    return fe_mk2(c.mn, c.mx);
}

// ---------------------------------------------------------------- fix-up (second kernel)
// Floor and clamp in the y = (x + 4) / 4 domain, with x = log10(mel) un-floored:
//   (max(max(x, -10), xmax' - 8) + 4) / 4 == max(y, thr),  thr = max(-1.5, max(ymax, -1.5) - 2)
// where xmax' is the maximum of the floored values, ymax the published maximum of the un-floored ones.
FE_HD FE_INLINE float fe_w2_threshold(float ymax) {
    const float ym = fmaxf(ymax, FeW2Cfg::Y_FLOOR);
    return fmaxf(FeW2Cfg::Y_FLOOR, ym - 2.0f);
}
