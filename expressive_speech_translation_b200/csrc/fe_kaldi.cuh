// kaldi-style 80-dim fbank (+ optional per-clip mean subtraction): phases of the fused kernel.
//
// Replaces torchaudio.compliance.kaldi.fbank(waveform, num_mel_bins=80, dither=0,
// sample_frequency=16000) [site-packages/torchaudio/compliance/kaldi.py:514-645; framing :44-83,
// DC removal :183-186, pre-emphasis :193-198, povey window :100, zero pad to 512 :207-211, |rfft|^2
// :616-618, mel banks :436-511, log floor :633] and the CosyVoice frontend's feat - feat.mean(0)
// (speaker-embedding input; model named at /root/reference/Docker/debug_trigger_test_cosyvoice.py:102).
//
// Same organisation as fe_whisper.cuh (lane <-> frame, FT = 32 frames per CTA):
//   phase L : coalesced tile load (no padding: snip_edges=True frames lie fully inside the clip)
//   phase S : per-frame partial sums for the DC removal (8 partials per frame)
//   phase A : gather n = 16*n1 + n2 (n1 < 25; n >= 400 is the zero pad), (x - mu) - 0.97*(x_prev - mu),
//             povey window, real radix-32 DFT over n1 -> k1 = 0..16, times W_512^(n2*k1) -> A[frame][k1][n2]
//   phase B : complex radix-16 DFT over n2 for k1 = 1..15 (bins k1 + 32*k2 and mirrored), light task
//             k1 = 0 and 16; |X|^2                                                    -> P[bin][frame]
//   phase M : unrolled sparse mel projection, ln(max(., eps))                         -> S[frame][mel]
//   phase O : coalesced frame-major store + per-tile column sums (deterministic CMN, no float atomics)
#pragma once
#include "fe_common.h"

struct FeKCfg {
    static constexpr int WIN = 400, HOP = 160, NFFT = 512, R1 = 32, R2 = 16, NZ1 = 25;
    static constexpr int FT = 32;
    static constexpr int NW = 8, NT = NW * 32;
    static constexpr int RT = (FT - 1) * HOP + WIN;      // 5360
    static constexpr int SW = RT + RT / HOP + 2;
    static constexpr int FS = 273;                       // float2 per frame in A (17*16 = 272 -> odd)
    static constexpr int PS = FT;
    static constexpr int NBIN = 256;                     // Nyquist bin carries zero weight
    static constexpr int NMEL = 80;
    static constexpr int SS = 81;                        // staging row stride (odd)
    static constexpr int R0 = (SW > NBIN * PS) ? SW : NBIN * PS;
    static constexpr int OFF_A = R0;                     // A; later the staging tile S[FT][SS]
    static constexpr int OFF_PART = OFF_A + FT * FS * 2; // partial sums [NW][FT]
    static constexpr int SMEM_FLOATS = OFF_PART + NW * FT;
    static constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
    static constexpr float EPS = 1.1920928955078125e-07f;
    static constexpr float PREEMPH = 0.97f;
};

struct FeKTile {
    int b, tile, t0, T;
    int64_t len;
};

FE_HD FE_INLINE FeKTile fe_k_tile(const FeKaldiParams& p, int cta) {
    FeKTile t;
    t.b = p.clip0 + cta / p.tiles_per_clip;
    t.tile = cta % p.tiles_per_clip;
    t.t0 = t.tile * FeKCfg::FT;
    t.len = fe_clip_len(p.lengths, t.b, p.L);
    int64_t T = t.len < FeKCfg::WIN ? 0 : 1 + (t.len - FeKCfg::WIN) / FeKCfg::HOP;
    t.T = (int)(T < p.T_out ? T : p.T_out);
    return t;
}

// ---------------------------------------------------------------- phase L
// Coalesced tile load in two halves: fe_k_load_issue puts the (at most six) 128-bit loads of a thread in flight,
// fe_k_load_store writes them to W.  The kernel issues the loads of the CTA's next tile before it starts the
// phases of the current one, so the HBM latency is off the critical path (it was 41 % of the stall samples).
struct FeKStage {
    static constexpr int U = (FeKCfg::RT + FeKCfg::NT * 4 - 1) / (FeKCfg::NT * 4);
    float4 v[U];
};

FE_HD FE_INLINE void fe_k_load_issue(const FeKaldiParams& p, int tid, int cta, FeKStage& st) {
    typedef FeKCfg C;
    const FeKTile t = fe_k_tile(p, cta);
    const float* __restrict__ x = p.wave + (int64_t)t.b * p.wave_stride;
    const int64_t g0 = (int64_t)t.t0 * C::HOP;
#pragma unroll
    for (int j = 0; j < FeKStage::U; ++j) {
        const int r4 = tid * 4 + j * C::NT * 4;
        if (r4 >= C::RT) continue;
        const int64_t g = g0 + r4;
        if ((p.vec_ok & 1) && g + 3 < t.len) {
            st.v[j] = *reinterpret_cast<const float4*>(x + g);
        } else {
            st.v[j].x = (g + 0 < t.len) ? x[g + 0] : 0.0f;
            st.v[j].y = (g + 1 < t.len) ? x[g + 1] : 0.0f;
            st.v[j].z = (g + 2 < t.len) ? x[g + 2] : 0.0f;
            st.v[j].w = (g + 3 < t.len) ? x[g + 3] : 0.0f;
        }
    }
}

FE_HD FE_INLINE void fe_k_load_store(float* sm, int tid, const FeKStage& st) {
    typedef FeKCfg C;
#pragma unroll
    for (int j = 0; j < FeKStage::U; ++j) {
        const int r4 = tid * 4 + j * C::NT * 4;
        if (r4 >= C::RT) continue;
        float* d = sm + r4 + r4 / C::HOP;
        d[0] = st.v[j].x; d[1] = st.v[j].y; d[2] = st.v[j].z; d[3] = st.v[j].w;
    }
}

FE_HD FE_INLINE void fe_k_phase_load(const FeKaldiParams& p, float* sm, int tid, int cta) {
    FeKStage st;
    fe_k_load_issue(p, tid, cta, st);
    fe_k_load_store(sm, tid, st);
}

// ---------------------------------------------------------------- phase S (frame sums for DC removal)
FE_HD FE_INLINE void fe_k_phase_sum(const FeKaldiParams&, float* sm, int tid, int) {
    typedef FeKCfg C;
    const int lane = tid & 31, warp = tid >> 5;
    constexpr int PER = C::WIN / C::NW;  // 50 samples per partial
    const float* __restrict__ s = sm + (C::HOP + 1) * lane;
    float acc = 0.0f;
#pragma unroll 10
    for (int i = 0; i < PER; ++i) {
        const int n = warp * PER + i;
        acc += s[n + n / C::HOP];
    }
    sm[C::OFF_PART + warp * C::FT + lane] = acc;
}

// ---------------------------------------------------------------- phase A
// One instantiation for every n2 (this codelet is a seventh of the kernel's code, and the kernel is fetch-bound):
// n2 == 0 differs only in where the sample before 16*n1 sits -- one pad slot lower when n1 is a multiple of 10,
// and x[-1] := x[0] for the very first sample.
FE_HD FE_INLINE void fe_k_stage_a(float* sm, int lane, int n2) {
    typedef FeKCfg C;
    float mu = 0.0f;
#pragma unroll
    for (int w = 0; w < C::NW; ++w) mu += sm[C::OFF_PART + w * C::FT + lane];
    mu *= (1.0f / C::WIN);
    const float* __restrict__ s = sm + (C::HOP + 1) * lane + n2;
    const int z = n2 == 0 ? 1 : 0;
    float x[C::NZ1];
#pragma unroll
    for (int n1 = 0; n1 < C::NZ1; ++n1) {
        // sample n = 16*n1 + n2; (16*n1 + n2) / 160 == n1 / 10 for every n2 < 16
        const float cur = s[C::R2 * n1 + n1 / (C::HOP / C::R2)];
        float prev;
        if (n1 == 0) {
            prev = s[z - 1];                             // n = 0 replicates x[0] (n2 == 0 reads s[0] again)
        } else if (n1 % (C::HOP / C::R2) == 0) {
            prev = s[C::R2 * n1 + n1 / (C::HOP / C::R2) - 1 - z];   // 16*n1 - 1 sits in the previous 160-block
        } else {
            prev = s[C::R2 * n1 + n1 / (C::HOP / C::R2) - 1];
        }
        const float a = cur - mu, b = prev - mu;
        x[n1] = (a - C::PREEMPH * b) * fe_win_kaldi[C::R2 * n1 + n2];
    }
    float yr[17], yi[17];
    fe_rdft32z25_h17(x, yr, yi);
    float2* __restrict__ a = reinterpret_cast<float2*>(sm + C::OFF_A) + C::FS * lane + n2;
    const float* __restrict__ tw = fe_tw_kaldi + n2 * 34;
    a[0] = make_float2(yr[0], 0.0f);
#pragma unroll
    for (int k1 = 1; k1 < 16; ++k1) {
        const float wr = tw[2 * k1], wi = tw[2 * k1 + 1];
        a[C::R2 * k1] = make_float2(yr[k1] * wr - yi[k1] * wi, yr[k1] * wi + yi[k1] * wr);
    }
    a[C::R2 * 16] = make_float2(yr[16] * tw[32], yr[16] * tw[33]);
}

FE_HD FE_INLINE void fe_k_phase_a(const FeKaldiParams&, float* sm, int tid, int) {
    typedef FeKCfg C;
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll 1
    for (int n2 = warp; n2 < C::R2; n2 += C::NW) fe_k_stage_a(sm, lane, n2);
}

// ---------------------------------------------------------------- phase B
FE_HD FE_INLINE void fe_k_stage_b_heavy(float* sm, int lane, int k1) {
    typedef FeKCfg C;
    const float2* __restrict__ a = reinterpret_cast<const float2*>(sm + C::OFF_A) + C::FS * lane + C::R2 * k1;
    float xr[16], xi[16], yr[16], yi[16];
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) {
        const float2 v = a[n2];
        xr[n2] = v.x; xi[n2] = v.y;
    }
    fe_cdft16(xr, xi, yr, yi);
    float* __restrict__ pd = sm + k1 * C::PS + lane;            // bin k1 + 32*k2, k2 = 0..7
    float* __restrict__ pm = sm + (C::R1 - k1) * C::PS + lane;  // bin (32-k1) + 32*(15-k2), k2 = 8..15
#pragma unroll
    for (int k2 = 0; k2 < 8; ++k2) pd[C::R1 * k2 * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
#pragma unroll
    for (int k2 = 8; k2 < 16; ++k2) pm[C::R1 * (15 - k2) * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
}

FE_HD FE_INLINE void fe_k_stage_b_light(float* sm, int lane) {
    typedef FeKCfg C;
    const float2* __restrict__ a = reinterpret_cast<const float2*>(sm + C::OFF_A) + C::FS * lane;
    {   // k1 = 0: real input, bins 32*k2, k2 = 0..7
        float xr[16], yr[8], yi[8];
#pragma unroll
        for (int n2 = 0; n2 < 16; ++n2) xr[n2] = a[n2].x;
        yi[0] = 0.0f;
        fe_rdft16_h8(xr, yr, yi);
        float* __restrict__ pd = sm + lane;
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) pd[C::R1 * k2 * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
    }
    {   // k1 = 16: bins 16 + 32*k2, k2 = 0..7
        float xr[16], xi[16], yr[8], yi[8];
#pragma unroll
        for (int n2 = 0; n2 < 16; ++n2) {
            const float2 v = a[C::R2 * 16 + n2];
            xr[n2] = v.x; xi[n2] = v.y;
        }
        fe_cdft16_h8(xr, xi, yr, yi);
        float* __restrict__ pd = sm + 16 * C::PS + lane;
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) pd[C::R1 * k2 * C::PS] = yr[k2] * yr[k2] + yi[k2] * yi[k2];
    }
}

FE_HD FE_INLINE void fe_k_phase_b(const FeKaldiParams&, float* sm, int tid, int) {
    typedef FeKCfg C;
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll 1
    for (int j = warp; j < C::R1 / 2; j += C::NW) {
        if (j == 0) fe_k_stage_b_light(sm, lane);
        else fe_k_stage_b_heavy(sm, lane, j);
    }
}

// ---------------------------------------------------------------- phase M
FE_HD FE_INLINE void fe_k_phase_mel(const FeKaldiParams&, float* sm, int tid, int) {
    typedef FeKCfg C;
    const int lane = tid & 31, warp = tid >> 5;
    float* __restrict__ st = sm + C::OFF_A + lane * C::SS;
    auto emit = [&](int m, float v) {
        st[m] = FE_LOG2(fmaxf(v, C::EPS)) * 0.69314718055994531f;  // ln
    };
    // unrolled (immediate weights): the rolled form (fe_mel_rolled) was measured 11 % slower here -- six taps per
    // filter on average do not amortise the loop overhead
    for (int c = warp; c < FE_MELPROJ_KALDI_NCHUNK; c += C::NW) fe_melproj_kaldi<float>(c, [&](int k) { return sm[k * C::PS + lane]; }, emit);
}

// ---------------------------------------------------------------- phase O
FE_HD FE_INLINE void fe_k_phase_store(const FeKaldiParams& p, float* sm, int tid, int cta) {
    typedef FeKCfg C;
    const FeKTile t = fe_k_tile(p, cta);
    int rows = t.T - t.t0;
    if (rows > C::FT) rows = C::FT;
    if (rows < 0) rows = 0;
    const float* __restrict__ st = sm + C::OFF_A;
    if (p.subtract_mean && tid < C::NMEL) {
        float acc = 0.0f;
        for (int r = 0; r < rows; ++r) acc += st[r * C::SS + tid];
        p.colsum[((int64_t)t.b * p.tiles_per_clip + t.tile) * C::NMEL + tid] = acc;
    }
    float* __restrict__ o = p.out + (int64_t)t.b * p.out_batch_stride + (int64_t)t.t0 * C::NMEL;
    int64_t rows_out = p.T_out - t.t0;               // rows of this tile that exist in `out`
    if (rows_out > C::FT) rows_out = C::FT;
    const int total = (int)rows_out * C::NMEL;       // rows in [rows, rows_out) are zero (ragged batches)
    const bool vec = (p.vec_ok & 2) != 0;
    if (vec) {
        for (int i = tid * 4; i < total; i += C::NT * 4) {
            const int r = i / C::NMEL, m = i - r * C::NMEL;
            const float* q = st + r * C::SS + m;
            float4 v; v.x = 0.0f; v.y = 0.0f; v.z = 0.0f; v.w = 0.0f;
            if (r < rows) { v.x = q[0]; v.y = q[1]; v.z = q[2]; v.w = q[3]; }
            *reinterpret_cast<float4*>(o + i) = v;
        }
    } else {
        for (int i = tid; i < total; i += C::NT) {
            const int r = i / C::NMEL, m = i - r * C::NMEL;
            o[i] = (r < rows) ? st[r * C::SS + m] : 0.0f;
        }
    }
}

// ---------------------------------------------------------------- CMN (second kernel)
// mean[m] = (sum over the clip's tiles of colsum) / T ;  out[t][m] -= mean[m]
FE_HD FE_INLINE float fe_k_clip_mean(const FeKaldiParams& p, int b, int m, int T) {
    const int ntiles = (T + FeKCfg::FT - 1) / FeKCfg::FT;
    const float* __restrict__ cs = p.colsum + (int64_t)b * p.tiles_per_clip * FeKCfg::NMEL + m;
    float acc = 0.0f;
    for (int i = 0; i < ntiles; ++i) acc += cs[(int64_t)i * FeKCfg::NMEL];
    return T > 0 ? acc / (float)T : 0.0f;
}
