// Waveform conditioning in front of the featurizers: interleaved int16 / float32
// PCM with 1 or 2 channels -> weighted mono mix -> band-limited polyphase resampling, one kernel, one trip
// through HBM.  Arithmetic = torchaudio.functional.resample (site-packages/torchaudio/functional/functional.py
// `_get_sinc_resample_kernel` + `_apply_sinc_resample_kernel`), which is what the reference runs at
// /root/reference/Backend/services/audio_processor.py:295-304 (kaiser, width 128) and what SURVEY.md's cfg1
// recipe uses (defaults: hann, width 6) -- but only the taps inside the window's support are touched
// (35 of 475 for 44.1 -> 16 kHz; the reference's conv1d multiplies all 475).
//
//   out[q*new + p] = sum_k h[p][k] * xpad[q*orig + k],   xpad[n] = x[n - width] (0 outside the clip)
//
// lane <-> block q (32 consecutive blocks per CTA tile), warp <-> group of 4 consecutive phases p: the four
// phases share one window of J input samples, so per tap step the warp issues one broadcast 128-bit load of
// four taps, one conflict-free scalar load of the sample (lane stride = block pitch, kept odd) and 4 FMAs.
// Small ratios (48 -> 16 kHz is orig 3 / new 1) are run as `m` blocks at once (orig_e = m*orig, new_e = m*new)
// with the tap table replicated, so a tile is never tiny.  The output tile is transposed through shared memory
// (odd pitch) and written with full 128 B rows.
#pragma once
#include "fe_common.h"

struct FeRsCfg {
    static constexpr int NT = 640;       // 20 warps, two CTAs per SM
    static constexpr int NW = NT / 32;
    static constexpr int QB = 32;        // blocks per tile = lanes
    static constexpr int R = 4;          // phases per group
};

struct FeResampleParams {
    const void* in;            // (B, L_in, CH) interleaved, float32 or int16
    const int64_t* lengths;    // device int64[B] input frames per clip, or null
    float* out;                // (B, out_stride)
    const float4* taps;        // [G][J] x 4 phases
    const int32_t* kstart;     // [G] first tap's index k into xpad, relative to the block start
    int64_t in_stride;         // elements of the input dtype between clips
    int64_t L_in;              // frames per clip
    int64_t out_stride;
    int64_t L_out;             // samples written per clip (zeros past a ragged clip's own length)
    int32_t orig, newf;        // reduced ratio (per-clip output length)
    int32_t orig_e, new_e;     // block sizes actually run (m * orig, m * new)
    int32_t width;
    int32_t P, OP;             // odd pitches of the input rows / output rows in shared memory
    int32_t G, J;
    int32_t nx;                // samples of one input tile = QB * orig_e + 2 * width
    int32_t xs_floats;         // floats reserved for the input tile
    int32_t tiles_per_clip;
    int32_t clip0;
    float w0, w1;              // mono mix weights (already carry 1/32768 for int16 input)
    int32_t vec_ok;            // stereo rows are aligned to the frame size: one load per frame
};

FE_HD FE_INLINE int64_t fe_rs_len(const FeResampleParams& p, int clip) {
    int64_t len = p.L_in;
    if (p.lengths) {
        int64_t v = p.lengths[clip];
        len = v < 0 ? 0 : (v > p.L_in ? p.L_in : v);
    }
    return len;
}

// One frame as it is fetched (raw bits, so that eight loads are in flight before the first conversion) and its
// mono value.  int16 goes through the 1.5 * 2^23 bit trick of fe_pcm16_to_float (fe_whisper2.cuh; no I2F): the
// subtraction is exact, the weights carry the 1/32768.  Stereo frames are one 32-bit (int16) / 64-bit (float) load
// when the rows are aligned to the frame size (p.vec_ok), two scalar loads otherwise.
template <bool IN16, int CH> struct FeRsRaw;
template <> struct FeRsRaw<true, 1> {
    int v;
    static constexpr int FRAME_BYTES = 2;
    FE_HD FE_INLINE void zero() { v = 0; }
    FE_HD FE_INLINE void load(const FeResampleParams&, const char* src, int n) { v = ((const short*)src)[n]; }
    FE_HD FE_INLINE float value(const FeResampleParams& p) const {
        return p.w0 * (fe_int_bits_as_float(0x4B400000 + v) - 12582912.0f);
    }
};
template <> struct FeRsRaw<true, 2> {
    int v;   // low half = channel 0, high half = channel 1
    static constexpr int FRAME_BYTES = 4;
    FE_HD FE_INLINE void zero() { v = 0; }
    FE_HD FE_INLINE void load(const FeResampleParams& p, const char* src, int n) {
        if (p.vec_ok) v = ((const int*)src)[n];
        else {
            const unsigned short* s = (const unsigned short*)src;
            v = (int)((unsigned)s[2 * n] | ((unsigned)s[2 * n + 1] << 16));
        }
    }
    FE_HD FE_INLINE float value(const FeResampleParams& p) const {
        const float a = fe_int_bits_as_float(0x4B400000 + (int)(short)(v & 0xffff)) - 12582912.0f;
        const float b = fe_int_bits_as_float(0x4B400000 + (v >> 16)) - 12582912.0f;
        return fmaf(p.w1, b, p.w0 * a);
    }
};
template <> struct FeRsRaw<false, 1> {
    float v;
    static constexpr int FRAME_BYTES = 4;
    FE_HD FE_INLINE void zero() { v = 0.0f; }
    FE_HD FE_INLINE void load(const FeResampleParams&, const char* src, int n) { v = ((const float*)src)[n]; }
    FE_HD FE_INLINE float value(const FeResampleParams& p) const { return p.w0 * v; }
};
template <> struct FeRsRaw<false, 2> {
    float2 v;
    static constexpr int FRAME_BYTES = 8;
    FE_HD FE_INLINE void zero() { v.x = 0.0f; v.y = 0.0f; }
    FE_HD FE_INLINE void load(const FeResampleParams& p, const char* src, int n) {
        if (p.vec_ok) v = ((const float2*)src)[n];
        else { v.x = ((const float*)src)[2 * n]; v.y = ((const float*)src)[2 * n + 1]; }
    }
    FE_HD FE_INLINE float value(const FeResampleParams& p) const { return fmaf(p.w1, v.y, p.w0 * v.x); }
};

// phase 1: mono-mixed input tile -> shared memory rows of orig_e samples, pitch P.  Tile sample n is clip frame
// g0 + n; frames [nlo, nhi) of the tile exist, the rest is zero padding.  Batches of U = 12 frames per thread that lie
// wholly inside take the unpredicated path (32-bit indices, every later address an immediate off the first).  Measured
// on the cfg2-sized 44.1 kHz batch: U = 8 3.34 ms, 12 3.25 ms, 16 3.73 ms, 24 (one round per tile) 4.0 ms -- past 12 the
// loads queue on the memory-instruction pipe (mio_throttle) instead of overlapping HBM latency.
template <bool IN16, int CH>
FE_HD FE_INLINE void fe_rs_load(const FeResampleParams& p, float* xs, int tid, int cta) {
    typedef FeRsRaw<IN16, CH> Raw;
    constexpr int U = 12, NT = FeRsCfg::NT;
    const int clip = p.clip0 + cta / p.tiles_per_clip;
    const int tile = cta % p.tiles_per_clip;
    const int64_t len = fe_rs_len(p, clip);
    const int64_t g0 = (int64_t)tile * FeRsCfg::QB * p.orig_e - p.width;
    const char* src = (const char*)p.in + ((int64_t)clip * p.in_stride * (IN16 ? 2 : 4) + g0 * Raw::FRAME_BYTES);
    const int nlo = g0 < 0 ? (int)(-g0) : 0;
    const int64_t rest = len - g0;
    const int nhi = rest <= 0 ? 0 : (rest < p.nx ? (int)rest : p.nx);
    const bool skew = p.P != p.orig_e;
    for (int n0 = tid; n0 < p.nx; n0 += U * NT) {
        Raw raw[U];
        if (n0 >= nlo && n0 + (U - 1) * NT < nhi) {
#if defined(__CUDACC__)
#pragma unroll
#endif
            for (int u = 0; u < U; ++u) raw[u].load(p, src, n0 + u * NT);
            if (!skew) {
#if defined(__CUDACC__)
#pragma unroll
#endif
                for (int u = 0; u < U; ++u) xs[n0 + u * NT] = raw[u].value(p);
                continue;
            }
        } else {
#if defined(__CUDACC__)
#pragma unroll
#endif
            for (int u = 0; u < U; ++u) {
                const int n = n0 + u * NT;
                raw[u].zero();
                if (n >= nlo && n < nhi) raw[u].load(p, src, n);
            }
        }
#if defined(__CUDACC__)
#pragma unroll
#endif
        for (int u = 0; u < U; ++u) {
            const int n = n0 + u * NT;
            if (n < p.nx) {
                const int a = skew ? n + (int)((unsigned)n / (unsigned)p.orig_e) : n;
                xs[a] = raw[u].value(p);
            }
        }
    }
}

// phase 2: warp <-> phase group, lane <-> block; results to the transposed output tile.  The four accumulators of a
// group are two packed pairs: one FFMA2 (sample broadcast by the instruction) per two phases.
FE_HD FE_INLINE void fe_rs_compute(const FeResampleParams& p, const float* xs, float* ost, const float4* taps,
                                   int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const float* xl = xs + lane * p.P;
    float* ol = ost + lane * p.OP;
    const int wrap = p.P - p.orig_e;
    if (wrap == 0) {
        // odd orig_e (no pitch padding): the window is one contiguous run.  A warp's groups are taken two at a time so
        // that two independent tap / sample streams are in flight (same per-phase summation order as the loop below):
        // 3.65 -> 3.34 ms on the cfg2-sized 44.1 kHz batch.  Folding the padded-pitch case into the same two-stream
        // loop (segments between the streams' row crossings) was measured slower for every ratio, so it keeps its own loop.
        for (int g = warp; g < p.G; g += 2 * FeRsCfg::NW) {
            const int g2 = g + FeRsCfg::NW;
            const bool two = g2 < p.G;
            const float* xa = xl + p.kstart[g];
            const float* xb = xl + p.kstart[two ? g2 : g];
            const float4* ta = taps + (int64_t)g * p.J;
            const float4* tb = taps + (int64_t)(two ? g2 : g) * p.J;
            fe2 a01 = fe_mk2(0.0f, 0.0f), a23 = fe_mk2(0.0f, 0.0f), b01 = fe_mk2(0.0f, 0.0f), b23 = fe_mk2(0.0f, 0.0f);
#if defined(__CUDACC__)
#pragma unroll 4
#endif
            for (int j = 0; j < p.J; ++j) {
                const float4 t = ta[j], u = tb[j];
                const float x = xa[j], y = xb[j];
                a01 = fe_fmac(x, fe_mk2(t.x, t.y), a01);
                a23 = fe_fmac(x, fe_mk2(t.z, t.w), a23);
                b01 = fe_fmac(y, fe_mk2(u.x, u.y), b01);
                b23 = fe_fmac(y, fe_mk2(u.z, u.w), b23);
            }
            int pe = g * FeRsCfg::R;
            ol[pe] = a01.x;
            if (pe + 1 < p.new_e) ol[pe + 1] = a01.y;
            if (pe + 2 < p.new_e) ol[pe + 2] = a23.x;
            if (pe + 3 < p.new_e) ol[pe + 3] = a23.y;
            if (two) {
                pe = g2 * FeRsCfg::R;
                ol[pe] = b01.x;
                if (pe + 1 < p.new_e) ol[pe + 1] = b01.y;
                if (pe + 2 < p.new_e) ol[pe + 2] = b23.x;
                if (pe + 3 < p.new_e) ol[pe + 3] = b23.y;
            }
        }
        return;
    }
    // even orig_e: rows are padded to an odd pitch, a window may cross into the next row
    for (int g = warp; g < p.G; g += FeRsCfg::NW) {
        const int k0 = p.kstart[g];
        int col = k0 % p.orig_e;
        const int off = (k0 / p.orig_e) * p.P + col;
        const float4* tg = taps + (int64_t)g * p.J;
        fe2 a01 = fe_mk2(0.0f, 0.0f), a23 = fe_mk2(0.0f, 0.0f);
        const float* xp = xl + off;
        for (int j = 0; j < p.J;) {
            const int left = p.J - j, row_left = p.orig_e - col;
            const int run = left < row_left ? left : row_left;
#if defined(__CUDACC__)
#pragma unroll 4
#endif
            for (int i = 0; i < run; ++i) {
                const float4 t = tg[i];
                const float x = xp[i];
                a01 = fe_fmac(x, fe_mk2(t.x, t.y), a01);
                a23 = fe_fmac(x, fe_mk2(t.z, t.w), a23);
            }
            j += run; tg += run; xp += run + wrap; col = 0;
        }
        const int pe = g * FeRsCfg::R;
        ol[pe] = a01.x;
        if (pe + 1 < p.new_e) ol[pe + 1] = a01.y;
        if (pe + 2 < p.new_e) ol[pe + 2] = a23.x;
        if (pe + 3 < p.new_e) ol[pe + 3] = a23.y;
    }
}

// phase 3: output tile -> global, one block row (new_e consecutive samples) per warp step; tile-relative 32-bit
// indices: [0, nvalid) carries samples, [nvalid, nwrite) the zeros past a ragged clip's end
FE_HD FE_INLINE void fe_rs_store(const FeResampleParams& p, const float* ost, int tid, int cta) {
    const int clip = p.clip0 + cta / p.tiles_per_clip;
    const int tile = cta % p.tiles_per_clip;
    const int64_t len = fe_rs_len(p, clip);
    const int64_t len_out = (len * p.newf + p.orig - 1) / p.orig;
    const int64_t i0 = (int64_t)tile * FeRsCfg::QB * p.new_e;
    const int tile_out = FeRsCfg::QB * p.new_e;
    float* orow = p.out + ((int64_t)clip * p.out_stride + i0);
    const int64_t v64 = len_out - i0, w64 = p.L_out - i0;
    const int nvalid = v64 <= 0 ? 0 : (v64 < tile_out ? (int)v64 : tile_out);
    const int nwrite = w64 <= 0 ? 0 : (w64 < tile_out ? (int)w64 : tile_out);
    const int warp = tid >> 5, lane = tid & 31;
    for (int blk = warp; blk < FeRsCfg::QB; blk += FeRsCfg::NW) {
        const float* src = ost + blk * p.OP;
        const int ib = blk * p.new_e;
        for (int pe = lane; pe < p.new_e; pe += 32) {
            const int i = ib + pe;
            if (i < nwrite) orow[i] = (i < nvalid) ? src[pe] : 0.0f;
        }
    }
}

#if defined(__CUDACC__)
template <bool IN16, int CH, bool TAPS_SMEM>
__global__ void __launch_bounds__(FeRsCfg::NT, 2) fe_resample_kernel(const FeResampleParams p) {
    extern __shared__ float4 fe_rs_smem[];
    float4* tsm = fe_rs_smem;                                   // taps first: 16 B aligned
    float* xs = (float*)(tsm + (TAPS_SMEM ? p.G * p.J : 0));
    float* ost = xs + p.xs_floats;
    const int tid = threadIdx.x;
    if (TAPS_SMEM)
        for (int i = tid; i < p.G * p.J; i += FeRsCfg::NT) tsm[i] = p.taps[i];
    fe_rs_load<IN16, CH>(p, xs, tid, blockIdx.x);
    __syncthreads();
    fe_rs_compute(p, xs, ost, TAPS_SMEM ? tsm : p.taps, tid);
    __syncthreads();
    fe_rs_store(p, ost, tid, blockIdx.x);
}
#endif
