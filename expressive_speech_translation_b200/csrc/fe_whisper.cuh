// Whisper log-mel featurizer: phases of the fused kernel.
//
// Replaces whisper/audio.py::log_mel_spectrogram (UPSTREAM openai-whisper; reached from
// /root/reference/Backend/services/cascaded_backend.py:158, Backend/speech_translation_pipeline.py:140,
// Backend/gtts_pipeline_fixed.py:64; same arithmetic as
// transformers/models/whisper/feature_extraction_whisper.py:135-164).
//
// One CTA = one tile of FT = 32 consecutive frames of one clip; lane <-> frame, so every warp-level
// task is uniform (same n2 / k1 / mel chunk for all 32 lanes): window and twiddle constants come from the
// constant bank, shared-memory strides per frame are odd (conflict-free), the mel weights are immediates.
//
//   phase L : coalesced load of the (FT-1)*160 + 400 samples the tile touches (reflect / zero pad resolved
//             here), into a padded shared tile (one pad word per 160 samples -> frame stride 161).
//   phase A : real-input radix-20 DFT over n1 for each n2 (n = 20*n1 + n2), window fused into the loads,
//             outputs k1 = 0..10 (Hermitian half), times W_400^(n2*k1)            -> A[frame][k1][n2]
//   phase B : complex radix-20 DFT over n2 for k1 = 1..9 (bins k1 + 20*k2 and, mirrored, 400 - those);
//             one light task does k1 = 0 (real input) and k1 = 10; |X|^2             -> P[bin][frame]
//   phase M : unrolled sparse mel projection (immediates), log10, running max, coalesced store.
// The per-clip (or per-call) max lands in max_keys via one atomicMax per CTA; a second tiny kernel applies
// max(x, gmax - 8), (x + 4) / 4 while the un-normalised wave of clips is still L2-resident.
#pragma once
#include "fe_common.h"

struct FeWCfg {
    static constexpr int N = 400, HOP = 160, PADL = 200, R1 = 20, R2 = 20;
    static constexpr int FT = 32;               // frames per tile == lanes
    static constexpr int NW = 10, NT = NW * 32; // warps / threads per CTA
    static constexpr int RT = (FT - 1) * HOP + N;        // samples a tile touches (5360)
    static constexpr int SW = RT + RT / HOP + 2;         // padded wave tile (floats)
    static constexpr int FS = 221;                       // float2 per frame in A (11*20 -> odd)
    static constexpr int PS = FT;                        // P row stride
    static constexpr int NBIN = 200;                     // bin 200 carries zero mel weight
    static constexpr int R0 = (SW > NBIN * PS) ? SW : NBIN * PS;  // region 0: wave tile, later P
    static constexpr int OFF_A = R0;                     // region 1: A (float2)
    static constexpr int OFF_MISC = OFF_A + FT * FS * 2; // 1 int: tile max key
    static constexpr int SMEM_FLOATS = OFF_MISC + 4;
    static constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
};

struct FeWTile {
    int b;          // clip
    int t0;         // first frame of the tile
    int64_t len;    // valid samples
    int64_t lp;     // len + padding
    int T;          // frames of this clip (<= T_out)
};

FE_HD FE_INLINE FeWTile fe_w_tile(const FeWhisperParams& p, int cta) {
    FeWTile t;
    t.b = p.clip0 + cta / p.tiles_per_clip;
    t.t0 = (cta % p.tiles_per_clip) * FeWCfg::FT;
    t.len = fe_clip_len(p.lengths, t.b, p.L);
    t.lp = t.len + p.padding;
    int64_t T = t.lp / FeWCfg::HOP;
    t.T = (int)(T < p.T_out ? T : p.T_out);
    return t;
}

// ---------------------------------------------------------------- phase L
FE_HD FE_INLINE void fe_w_phase_load(const FeWhisperParams& p, float* sm, int tid, int cta) {
    typedef FeWCfg C;
    const FeWTile t = fe_w_tile(p, cta);
    const float* __restrict__ x = p.wave + (int64_t)t.b * p.wave_stride;
    const int64_t g0 = (int64_t)t.t0 * C::HOP - C::PADL;
    if (tid == 0) ((int*)(sm + C::OFF_MISC))[0] = FE_KEY_INIT;
    for (int r4 = tid * 4; r4 < C::RT; r4 += C::NT * 4) {
        const int64_t g = g0 + r4;
        float v0, v1, v2, v3;
        if ((p.vec_ok & 1) && g >= 0 && g + 3 < t.len) {
            const float4 q = *reinterpret_cast<const float4*>(x + g);
            v0 = q.x; v1 = q.y; v2 = q.z; v3 = q.w;
        } else {
            v0 = fe_fetch_reflect(x, g + 0, t.len, t.lp);
            v1 = fe_fetch_reflect(x, g + 1, t.len, t.lp);
            v2 = fe_fetch_reflect(x, g + 2, t.len, t.lp);
            v3 = fe_fetch_reflect(x, g + 3, t.len, t.lp);
        }
        float* d = sm + r4 + r4 / C::HOP;  // 4 | HOP: the four samples share one 160-block
        d[0] = v0; d[1] = v1; d[2] = v2; d[3] = v3;
    }
}

// ---------------------------------------------------------------- phase A
FE_HD FE_INLINE void fe_w_stage_a(float* sm, int lane, int n2) {
    typedef FeWCfg C;
    const float* __restrict__ s = sm + (C::HOP + 1) * lane + n2;
    float x[C::R1];
#pragma unroll
    for (int n1 = 0; n1 < C::R1; ++n1) {
        // sample n = 20*n1 + n2 of this lane's frame; (20*n1 + n2) / 160 == n1 / 8 for every n2 < 20
        x[n1] = s[C::R2 * n1 + n1 / (C::HOP / C::R2)] * fe_win_whisper[C::R2 * n1 + n2];
    }
    float yr[11], yi[11];
    fe_rdft20_h11(x, yr, yi);
    float2* __restrict__ a = reinterpret_cast<float2*>(sm + C::OFF_A) + C::FS * lane + n2;
    const float* __restrict__ tw = fe_tw_whisper + n2 * 22;
    a[0] = make_float2(yr[0], 0.0f);
#pragma unroll
    for (int k1 = 1; k1 < 10; ++k1) {
        const float wr = tw[2 * k1], wi = tw[2 * k1 + 1];
        a[C::R2 * k1] = make_float2(yr[k1] * wr - yi[k1] * wi, yr[k1] * wi + yi[k1] * wr);
    }
    a[C::R2 * 10] = make_float2(yr[10] * tw[20], yr[10] * tw[21]);
}

FE_HD FE_INLINE void fe_w_phase_a(const FeWhisperParams&, float* sm, int tid, int) {
    typedef FeWCfg C;
    const int lane = tid & 31, warp = tid >> 5;
    for (int n2 = warp; n2 < C::R2; n2 += C::NW) fe_w_stage_a(sm, lane, n2);
}

// ---------------------------------------------------------------- phase B
FE_HD FE_INLINE void fe_w_stage_b_heavy(float* sm, int lane, int k1) {
    typedef FeWCfg C;
    const float2* __restrict__ a = reinterpret_cast<const float2*>(sm + C::OFF_A) + C::FS * lane + C::R2 * k1;
    float xr[20], xi[20], yr[20], yi[20];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) {
        const float2 v = a[n2];
        xr[n2] = v.x; xi[n2] = v.y;
    }
    fe_cdft20(xr, xi, yr, yi);
    float* __restrict__ pd = sm + k1 * C::PS + lane;            // bin k1 + 20*k2, k2 = 0..9
    float* __restrict__ pm = sm + (C::R1 - k1) * C::PS + lane;  // bin (20-k1) + 20*(19-k2), k2 = 10..19
#pragma unroll
    for (int k2 = 0; k2 < 10; ++k2) pd[C::R1 * k2 * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
#pragma unroll
    for (int k2 = 10; k2 < 20; ++k2) pm[C::R1 * (19 - k2) * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
}

FE_HD FE_INLINE void fe_w_stage_b_light(float* sm, int lane) {
    typedef FeWCfg C;
    const float2* __restrict__ a = reinterpret_cast<const float2*>(sm + C::OFF_A) + C::FS * lane;
    {   // k1 = 0: real input, bins 20*k2, k2 = 0..9
        float xr[20], yr[10], yi[10];
#pragma unroll
        for (int n2 = 0; n2 < 20; ++n2) xr[n2] = a[n2].x;
        yi[0] = 0.0f;
        fe_rdft20_h10(xr, yr, yi);
        float* __restrict__ pd = sm + lane;
#pragma unroll
        for (int k2 = 0; k2 < 10; ++k2) pd[C::R1 * k2 * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
    }
    {   // k1 = 10: bins 10 + 20*k2, k2 = 0..9 (k2 >= 10 only repeats them mirrored)
        float xr[20], xi[20], yr[10], yi[10];
#pragma unroll
        for (int n2 = 0; n2 < 20; ++n2) {
            const float2 v = a[C::R2 * 10 + n2];
            xr[n2] = v.x; xi[n2] = v.y;
        }
        fe_cdft20_h10(xr, xi, yr, yi);
        float* __restrict__ pd = sm + 10 * C::PS + lane;
#pragma unroll
        for (int k2 = 0; k2 < 10; ++k2) pd[C::R1 * k2 * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
    }
}

FE_HD FE_INLINE void fe_w_phase_b(const FeWhisperParams&, float* sm, int tid, int) {
    typedef FeWCfg C;
    const int lane = tid & 31, warp = tid >> 5;
    for (int j = warp; j < C::R1 / 2; j += C::NW) {
        if (j == 0) fe_w_stage_b_light(sm, lane);
        else fe_w_stage_b_heavy(sm, lane, j);
    }
}

// ---------------------------------------------------------------- phase M
// returns this thread's max key over the values it stored (FE_KEY_INIT when it stored none)
template <int NMELS>
FE_HD FE_INLINE int fe_w_phase_mel(const FeWhisperParams& p, float* sm, int tid, int cta) {
    typedef FeWCfg C;
    const int lane = tid & 31, warp = tid >> 5;
    const FeWTile t = fe_w_tile(p, cta);
    const int frame = t.t0 + lane;
    const bool live = frame < t.T;
    float* __restrict__ o = p.out + (int64_t)t.b * p.out_batch_stride + frame;
    const int64_t rs = p.T_out;
    float mx = -3.0e38f;
    auto emit = [&](int m, float v) {
        const float lg = FE_LOG2(fmaxf(v, 1e-10f)) * 0.30102999566398120f;  // log10
        if (live) {
            o[(int64_t)m * rs] = lg;
            mx = fmaxf(mx, lg);
        }
    };
    constexpr int NCH = (NMELS == 80) ? FE_MELPROJ_W80_NCHUNK : FE_MELPROJ_W128_NCHUNK;
    for (int c = warp; c < NCH; c += C::NW) {
        if (NMELS == 80) fe_melproj_w80<C::PS>(c, sm + lane, emit);
        else fe_melproj_w128<C::PS>(c, sm + lane, emit);
    }
    return live ? fe_float_key(mx) : FE_KEY_INIT;
}

// ---------------------------------------------------------------- normalisation (second kernel)
// out = (max(x, gmax - 8) + 4) / 4 for frames < T_b, 0 beyond (ragged batches).
FE_HD FE_INLINE float fe_w_normalise(float x, float gmax) {
    return (fmaxf(x, gmax - 8.0f) + 4.0f) * 0.25f;
}
