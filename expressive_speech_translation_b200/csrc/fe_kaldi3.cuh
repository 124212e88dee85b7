// kaldi-style 80-dim fbank, Tensor-Memory formulation: phases of the persistent fused kernel.
//
// Replaces torchaudio.compliance.kaldi.fbank(waveform, num_mel_bins=80, dither=0, sample_frequency=16000)
// [site-packages/torchaudio/compliance/kaldi.py:514-645; framing :44-83, DC removal :183-186, pre-emphasis :193-198,
// povey window :100, zero pad to 512 :207-211, |rfft|^2 :616-618, mel banks :436-511, log floor :633] and the CosyVoice
// frontend's feat - feat.mean(0) (speaker-embedding input; model named at
// /root/reference/Docker/debug_trigger_test_cosyvoice.py:102,258).
//
// Same skeleton as fe_whisper3.cuh (lane = frame, team = SM sub-partition = TMEM lane quadrant, 128-frame CTA units,
// waveform rows by TMA box copy, phase-A output in Tensor Memory, mbarrier pipeline A(u) | M(u-1) | B(u)):
//   W[q]    : 34 rows of 160 samples (row r = samples 160 (f0 + r) ..), pitch 164 floats; snip_edges framing: no padding
//   sums    : the team's four warps sum rows in four column chunks -> per-frame mean (DC removal) from three row sums
//   phase A : N = 512 = 32 (n1) x 16 (n2), n = 16 n1 + n2, n1 < 25 (400 samples, the rest is the zero pad).  8 half-units
//             per team (sub-sequences 2 hu, 2 hu + 1 packed): (x - 0.97 x_prev - 0.03 mean) * povey, real 32-point DFT
//             with zero-input pruning -> k1 = 0..16, times W_512^(n2 k1) -> Tensor Memory, 64 columns per half-unit
//             (all 512 columns are used: rows 0 and 16 are stored real, row 16's twiddle is applied in phase B)
//   phase B : 9 tasks per team: rows (1,2) .. (13,14) as packed complex 16-point DFTs, row 15 scalar, rows (0,16) packed
//             half-output; |X|^2 -> P[bin][4 * lane + (team ^ (lane >> 3))]  (bins k1 + 32 k2 and mirrored 32 - k1 + ..)
//   phase M : a lane owns 4 frames (one per team) x 4 consecutive filters; kaldi's triangles overlap only with their
//             neighbours, so a bin feeds <= 2 filters: segments of bins between filter centres, one LDS.128 and <= 4
//             FFMA2 per bin; ln(max(., eps)); ONE 16-byte store per frame and filter group straight into the
//             frame-major (T, 80) output (no staging tile: P and W leave no shared memory for one)
// CMN (feat - mean over frames) is a second kernel: one CTA per clip holds the clip's (T, 80) tile in shared memory.
#pragma once
#include "fe_common.h"
#include "fe_whisper3.cuh"
#include "fe_kaldi.cuh"

struct FeK3Cfg {
    static constexpr int HOP = 160, WIN = 400, NMEL = 80, NBIN = 256;
    static constexpr int TEAMS = 4, MEMBERS = 4, NW = TEAMS * MEMBERS, NT = NW * 32;
    static constexpr int FT = 32, UF = TEAMS * FT;
    static constexpr int ROWS = 34, PITCH = 164;
    static constexpr int WT = (ROWS * PITCH + 31) / 32 * 32;
    static constexpr int TMEM_COLS = 512, HUCOLS = 64, NHU = 8;
    static constexpr int PARTW = 36;                                    // partial row sums per member (34 rows + pad)
    static constexpr int OFF_W = 0;
    static constexpr int OFF_P = OFF_W + TEAMS * WT;                    // P[bin][UF]
    static constexpr int OFF_MBAR = OFF_P + NBIN * UF;                  // 14 mbarriers + TMEM slot + counters
    static constexpr int NMBAR = 14;
    static constexpr int OFF_PART = OFF_MBAR + 2 * NMBAR + 8;           // [TEAMS][PARTS][PARTW]
    static constexpr int PARTS = MEMBERS + 1;                           // + the raw first sample of every row
    static constexpr int OFF_SEG = OFF_PART + TEAMS * PARTS * PARTW;    // ints: fe_k3_mel_seg (82, padded to 84)
    static constexpr int OFF_MW = OFF_SEG + 84;                         // fe_k3_mel_w (256 x (wf, wr))
    static constexpr int SMEM_FLOATS = OFF_MW + 2 * NBIN;
    static constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
    static_assert(OFF_P % 4 == 0 && OFF_MBAR % 2 == 0 && OFF_MW % 2 == 0, "alignment of the carve-up");
    static_assert(SMEM_BYTES <= 232448, "more than 227 KB of shared memory");
    static constexpr float EPS = 1.1920928955078125e-07f;
    static constexpr float LN2 = 0.69314718055994531f;
};

struct FeK3Unit {
    int b, t0, T, n_in;   // clip, first frame, frames of the clip (<= T_out), output rows this unit covers
    int64_t len, o0;      // valid samples; element offset in `out` of (frame t0, mel 0)
};

FE_HD FE_INLINE FeK3Unit fe_k3_unit_at(const FeKaldiParams& p, int b, int within) {
    typedef FeK3Cfg C;
    FeK3Unit t;
    t.b = b;
    t.len = fe_clip_len(p.lengths, b, p.L);
    int64_t T = 0;
    if (t.len >= C::WIN)
        T = t.len < ((int64_t)1 << 32) ? (int64_t)(1u + (uint32_t)(t.len - C::WIN) / (uint32_t)C::HOP) : 1 + (t.len - C::WIN) / C::HOP;
    t.T = (int)(T < p.T_out ? T : p.T_out);
    t.t0 = within * C::UF;
    const int64_t rest = p.T_out - t.t0;
    t.n_in = (int)(rest < 0 ? 0 : (rest < C::UF ? rest : C::UF));
    t.o0 = (int64_t)b * p.out_batch_stride + (int64_t)t.t0 * C::NMEL;
    return t;
}
// unit order: position-major, u = within * B + clip (see fe_whisper3.cuh)
FE_HD FE_INLINE FeK3Unit fe_k3_unit(const FeKaldiParams& p, int u) {
    const int within = u / p.B;
    return fe_k3_unit_at(p, u - within * p.B, within);
}
// one step = gridDim.x units = (step_w positions, step_b clips), carry when the clip index wraps
FE_HD FE_INLINE FeW3Pos fe_k3_pos_next(const FeKaldiParams& p, FeW3Pos q) {
    q.b += p.step_b; q.within += p.step_w;
    if (q.b >= p.B) { q.b -= p.B; q.within += 1; }
    return q;
}
FE_HD FE_INLINE bool fe_k3_team_live(const FeK3Unit& t, int q) { return t.t0 + FeK3Cfg::FT * q < t.T; }

// ---------------------------------------------------------------- loader
// The box copy (b200feat.cu) fetches rows f0 .. f0 + 33 of the clip's tensor map (rows of 164 floats every 160 floats,
// `tmap_rows` of them inside the L samples of the row buffer; rows outside arrive as zeros).  Live frames of the team need
// rows up to min(f0 + 31, T - 1) + 2; those beyond the map are patched element-wise by the team (one row per clip when
// L is not 4 floats longer than the last row).
FE_HD FE_INLINE int fe_k3_need_row(const FeK3Unit& t, int f0) {     // last clip row a live frame of the team reads
    const int fl = f0 + FeK3Cfg::FT - 1 < t.T - 1 ? f0 + FeK3Cfg::FT - 1 : t.T - 1;
    return fl + 2;
}
FE_HD FE_INLINE bool fe_k3_box_complete(const FeKaldiParams& p, const FeK3Unit& t, int q) {
    return fe_k3_need_row(t, t.t0 + FeK3Cfg::FT * q) < p.tmap_rows;
}
FE_HD FE_INLINE void fe_k3_patch_rows(const FeKaldiParams& p, const FeK3Unit& t, int q, float* wq, int member, int lane) {
    typedef FeK3Cfg C;
    const int f0 = t.t0 + C::FT * q;
    const int need = fe_k3_need_row(t, f0);
    const float* __restrict__ x = p.wave + (int64_t)t.b * p.wave_stride;
#pragma unroll 1
    for (int r = member; r < C::ROWS; r += C::MEMBERS) {
        const int row = f0 + r;
        if (row < p.tmap_rows || row > need) continue;
        const int64_t g = (int64_t)row * C::HOP + lane;
#pragma unroll
        for (int j = 0; j < 5; ++j) wq[r * C::PITCH + lane + 32 * j] = g + 32 * j < t.len ? x[g + 32 * j] : 0.0f;
    }
}
// host emulation of the box copy
FE_HD FE_INLINE void fe_k3_box_host(const FeKaldiParams& p, const FeK3Unit& t, int q, float* wq) {
    typedef FeK3Cfg C;
    const int f0 = t.t0 + C::FT * q;
    const float* x = p.wave + (int64_t)t.b * p.wave_stride;
    for (int r = 0; r < C::ROWS; ++r)
        for (int c = 0; c < C::PITCH; ++c)
            wq[r * C::PITCH + c] = (f0 + r < p.tmap_rows) ? x[(int64_t)(f0 + r) * C::HOP + c] : 0.0f;
}

// every thread of the CTA calls it, a barrier must follow
FE_HD FE_INLINE void fe_k3_copy_tables(float* sm, int tid, int nthreads) {
    typedef FeK3Cfg C;
    int* seg = reinterpret_cast<int*>(sm + C::OFF_SEG);
    for (int i = tid; i < 82; i += nthreads) seg[i] = fe_k3_mel_seg[i];
    for (int i = tid; i < 2 * C::NBIN; i += nthreads) sm[C::OFF_MW + i] = fe_k3_mel_w[i];
}

// ---------------------------------------------------------------- pre-pass: row sums + pre-emphasis in place
// DC removal and pre-emphasis (kaldi.py:183-198): e[j] = (x[j] - m) - 0.97 (x[j-1] - m) = g[j] - 0.03 m for j >= 1 with
// g[t] = x[t] - 0.97 x[t-1], which does not depend on the frame: the team's four warps sweep W once (member j: columns
// [40 j, 40 j + 40) of row `lane`, then of rows 32, 33 by lanes 0, 1), sum the raw samples (-> frame mean m from three
// row sums) and replace x by g in place.  e[0] = 0.03 (x[0] - m) needs the RAW first sample of the frame's first row:
// kept aside.  All loads (fe_k3_pre_load) happen before a team barrier, all stores (fe_k3_pre_store) after it, and a
// second team barrier precedes phase A.  Conflict-free 128-bit accesses; phase A then needs no x[j-1] loads
// (they were 4-way bank-conflicted 32-bit loads: 41 % of the kernel's shared-memory wavefronts, profiles/r02_k3_v1).
struct FeK3Pre { float4 v[2][10]; float left[2]; };

// which (row, column chunk) a thread handles in pass 0 / pass 1: pass 0 = row `lane`, chunk `member`; pass 1 (rows 32, 33)
// is done by member 0 alone, lanes 0..7 = (row 32 + lane / 4, chunk lane % 4) -- the other members skip it
FE_HD FE_INLINE bool fe_k3_pre_item(int pass, int member, int lane, int& row, int& chunk) {
    if (pass == 0) { row = lane; chunk = member; return true; }
    row = 32 + (lane >> 2); chunk = lane & 3;
    return member == 0 && lane < 8;
}
FE_HD FE_INLINE void fe_k3_pre_load(const float* wq, int member, int lane, FeK3Pre& r) {
    typedef FeK3Cfg C;
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 1 && member != 0) break;                  // warp-uniform
        int row, chunk;
        if (!fe_k3_pre_item(pass, member, lane, row, chunk)) continue;
        const float* __restrict__ x = wq + C::PITCH * row + 40 * chunk;
        const float4* __restrict__ x4 = reinterpret_cast<const float4*>(x);
#pragma unroll
        for (int i = 0; i < 10; ++i) r.v[pass][i] = x4[i];
        // the sample before the chunk: previous column, or the last sample of the previous row (rows are consecutive hops)
        r.left[pass] = chunk > 0 ? x[-1] : (row > 0 ? wq[C::PITCH * (row - 1) + C::HOP - 1] : x[0]);
    }
}
FE_HD FE_INLINE void fe_k3_pre_store(float* wq, float* part_team, int member, int lane, const FeK3Pre& r) {
    typedef FeK3Cfg C;
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 1 && member != 0) break;
        int row, chunk;
        if (!fe_k3_pre_item(pass, member, lane, row, chunk)) continue;
        float4* __restrict__ x4 = reinterpret_cast<float4*>(wq + C::PITCH * row + 40 * chunk);
        fe2 s0 = fe_zero<fe2>(), s1 = fe_zero<fe2>();
        float prev = r.left[pass];
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            const float4 a = r.v[pass][i];
            const fe2 t = fe_add(fe_mk2(a.x, a.y), fe_mk2(a.z, a.w));
            if (i & 1) s1 = fe_add(s1, t); else s0 = fe_add(s0, t);
            float4 g;
            g.x = fmaf(-0.97f, prev, a.x); g.y = fmaf(-0.97f, a.x, a.y); g.z = fmaf(-0.97f, a.y, a.z); g.w = fmaf(-0.97f, a.z, a.w);
            prev = a.w;
            x4[i] = g;
        }
        const fe2 st = fe_add(s0, s1);
        part_team[chunk * C::PARTW + row] = st.x + st.y;
        if (chunk == 0) part_team[C::MEMBERS * C::PARTW + row] = r.v[pass][0].x;
    }
}
// frame `lane` = rows lane, lane + 1 and the first 80 samples of row lane + 2
FE_HD FE_INLINE float fe_k3_frame_mean(const float* part_team, int lane) {
    typedef FeK3Cfg C;
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < C::MEMBERS; ++j) s += part_team[j * C::PARTW + lane] + part_team[j * C::PARTW + lane + 1];
    s += part_team[lane + 2] + part_team[C::PARTW + lane + 2];
    return s * (1.0f / C::WIN);
}
FE_HD FE_INLINE float fe_k3_frame_x0(const float* part_team, int lane) { return part_team[FeK3Cfg::MEMBERS * FeK3Cfg::PARTW + lane]; }

// ---------------------------------------------------------------- task maps
FE_HD FE_INLINE int fe_k3_b_count(int member) { return member < 3 ? 2 : 3; }
FE_HD FE_INLINE int fe_k3_b_task(int member, int k) { return member < 3 ? 2 * member + k : 6 + k; }   // 0..6 pairs, 7 row 15, 8 rows 0/16

// ---------------------------------------------------------------- phase A
// Tensor Memory columns of half-unit hu (64): [0, 1] row 0 (c0, c1) | [2, 3] row 16 (real, before its twiddle) |
// [4..7] row 15 (re c0, re c1, im c0, im c1) | [8 + 8 j ..] rows (2 j + 1, 2 j + 2): re_a(c0, c1) im_a re_b im_b
template <class PreStore>
FE_HD FE_INLINE void fe_k3_stage_a(const float* wq, const FeTm& tm, int lane, int hu, float mean, float x0, PreStore pre_store) {
    typedef FeK3Cfg C;
    // W holds g[t] = x[t] - 0.97 x[t-1] (pre-pass).  Sample j0 = 16 n1 + 2 hu (+ 1) of this lane's frame sits at
    // j0 + 4 (n1 / 10): (16 n1 + 2 hu + 1) / 160 == n1 / 10 for every hu < 8
    const fe2* __restrict__ s = reinterpret_cast<const fe2*>(wq + C::PITCH * lane) + hu;
    const fe2* __restrict__ win = reinterpret_cast<const fe2*>(fe_k3_win + 50 * hu);
    const fe2 c03 = fe_mk2(-0.03f * mean, -0.03f * mean);
    fe2 x[25];
#pragma unroll
    for (int n1 = 0; n1 < 25; ++n1) x[n1] = fe_mul(win[n1], fe_add(s[8 * n1 + 2 * (n1 / 10)], c03));
    if (hu == 0) x[0].x = fe_k3_win[0] * (0.03f * (x0 - mean));     // j = 0: x[-1] := x[0] (kaldi.py:193-198)
    fe2 yr[17], yi[17];
    fe_rdft32z25_h17(x, yr, yi);
    pre_store();   // Tensor Memory may still be read by phase B of the previous unit
    const fe2* __restrict__ tw = reinterpret_cast<const fe2*>(fe_k3_tw + 60 * hu);
    const int col = C::HUCOLS * hu;
    fe2 re15, im15;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        fe2 re[2], im[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int k1 = 2 * j + 1 + h;
            if (k1 > 15) continue;
            const fe2 wr = tw[2 * (k1 - 1)], wi = tw[2 * (k1 - 1) + 1];
            re[h] = fe_fma(fe_neg(wi), yi[k1], fe_mul(wr, yr[k1]));
            im[h] = fe_fma(wi, yr[k1], fe_mul(wr, yi[k1]));
        }
        if (j < 7) fe_tm_st8(tm, col + 8 + 8 * j, re[0], im[0], re[1], im[1]);
        else { re15 = re[0]; im15 = im[0]; }
    }
    fe_tm_st8(tm, col, yr[0], yr[16], re15, im15);
}

// ---------------------------------------------------------------- phase B
// P[bin][slot], slot = 4 * lane + (team ^ (lane >> 3)): the 32 lanes of a warp hit 32 different banks, and phase M reads
// the four teams' values of one lane index as ONE 16-byte word
FE_HD FE_INLINE int fe_k3_p_slot(int lane, int team) { return 4 * lane + (team ^ (lane >> 3)); }

// row k1 (1..15) of the 32 x 16 decomposition: outputs k2 < 8 are bins k1 + 32 k2, outputs k2 >= 8 mirror to bins
// (32 - k1) + 32 (15 - k2)  (X[512 - k] = conj X[k]; bin 256 carries no mel weight)
FE_HD FE_INLINE int fe_k3_bin(int k1, int k2) { return k2 < 8 ? k1 + 32 * k2 : (32 - k1) + 32 * (15 - k2); }

template <class PostLoad, class PreStore>
FE_HD FE_INLINE void fe_k3_stage_b_pair(float* P, const FeTm& tm, int lane, int team, int pr, PostLoad post_load, PreStore pre_store) {
    typedef FeK3Cfg C;
    fe2 xr[16], xi[16], yr[16], yi[16];
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) {
        const int c = C::HUCOLS * (n2 / 2) + 8 + (n2 % 2);
        xr[n2].x = fe_tm_ld1(tm, c + 8 * pr); xi[n2].x = fe_tm_ld1(tm, c + 8 * pr + 2);
        xr[n2].y = fe_tm_ld1(tm, c + 8 * pr + 4); xi[n2].y = fe_tm_ld1(tm, c + 8 * pr + 6);
    }
    fe_tm_wait_ld();
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) { fe_tm_tie(xr[n2]); fe_tm_tie(xi[n2]); }
    post_load();
    fe_cdft16(xr, xi, yr, yi);
    fe2 pw[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) pw[k] = fe_fma(yr[k], yr[k], fe_mul(yi[k], yi[k]));
    pre_store();
    float* __restrict__ pl = P + fe_k3_p_slot(lane, team);
    const int ka = 2 * pr + 1;     // rows ka and ka + 1; their bins are ka + 32 k2 / (32 - ka) + 32 (15 - k2) and one above / below
    float* __restrict__ pa = pl + C::UF * ka;
    float* __restrict__ pm = pl + C::UF * (32 - ka);
#pragma unroll
    for (int k2 = 0; k2 < 8; ++k2) {
        pa[C::UF * 32 * k2] = pw[k2].x;
        pa[C::UF * 32 * k2 + C::UF] = pw[k2].y;
        pm[C::UF * 32 * k2] = pw[15 - k2].x;
        pm[C::UF * 32 * k2 - C::UF] = pw[15 - k2].y;
    }
}

// row 15 alone (scalar)
template <class PostLoad, class PreStore>
FE_HD FE_INLINE void fe_k3_stage_b_15(float* P, const FeTm& tm, int lane, int team, PostLoad post_load, PreStore pre_store) {
    typedef FeK3Cfg C;
    float xr[16], xi[16], yr[16], yi[16];
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) {
        const int c = C::HUCOLS * (n2 / 2) + 4 + (n2 % 2);
        xr[n2] = fe_tm_ld1(tm, c); xi[n2] = fe_tm_ld1(tm, c + 2);
    }
    fe_tm_wait_ld();
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) { fe_tm_tie(xr[n2]); fe_tm_tie(xi[n2]); }
    post_load();
    fe_cdft16(xr, xi, yr, yi);
    pre_store();
    float* __restrict__ pl = P + fe_k3_p_slot(lane, team);
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) pl[C::UF * fe_k3_bin(15, k2)] = fmaf(yr[k2], yr[k2], yi[k2] * yi[k2]);
}

// rows 0 (real input) and 16 (its twiddle W_512^(16 n2) is applied here) packed: bins 32 k2 and 16 + 32 k2, k2 < 8
template <class PostLoad, class PreStore>
FE_HD FE_INLINE void fe_k3_stage_b_0_16(float* P, const FeTm& tm, int lane, int team, PostLoad post_load, PreStore pre_store) {
    typedef FeK3Cfg C;
    fe2 xr[16], xi[16], yr[8], yi[8];
    float y16[16];
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) {
        const int c = C::HUCOLS * (n2 / 2) + (n2 % 2);
        xr[n2].x = fe_tm_ld1(tm, c); y16[n2] = fe_tm_ld1(tm, c + 2);
    }
    fe_tm_wait_ld();
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) { fe_tm_tie(xr[n2].x); fe_tm_tie(y16[n2]); }
    post_load();
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) {
        xi[n2].x = 0.0f;
        xr[n2].y = fe_k3_tw16[2 * n2] * y16[n2];
        xi[n2].y = fe_k3_tw16[2 * n2 + 1] * y16[n2];
    }
    fe_cdft16_h8(xr, xi, yr, yi);
    pre_store();
    float* __restrict__ pl = P + fe_k3_p_slot(lane, team);
#pragma unroll
    for (int k2 = 0; k2 < 8; ++k2) {
        const fe2 pw = fe_fma(yr[k2], yr[k2], fe_mul(yi[k2], yi[k2]));
        pl[C::UF * 32 * k2] = pw.x;
        pl[C::UF * (32 * k2 + 16)] = pw.y;
    }
}

template <class PostLoad, class PreStore>
FE_HD FE_INLINE void fe_k3_stage_b(float* P, const FeTm& tm, int lane, int team, int task, PostLoad post_load, PreStore pre_store) {
    if (task < 7) fe_k3_stage_b_pair(P, tm, lane, team, task, post_load, pre_store);
    else if (task == 7) fe_k3_stage_b_15(P, tm, lane, team, post_load, pre_store);
    else fe_k3_stage_b_0_16(P, tm, lane, team, post_load, pre_store);
}

// ---------------------------------------------------------------- phase M
FE_HD FE_INLINE void fe_k3_stg4(float* q, float a, float b, float c, float d, bool vec) {
#if defined(__CUDA_ARCH__)
    if (vec) asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(q), "f"(a), "f"(b), "f"(c), "f"(d));
    else { q[0] = a; q[1] = b; q[2] = c; q[3] = d; }
#else
    (void)vec; q[0] = a; q[1] = b; q[2] = c; q[3] = d;
#endif
}

// lane l, component c <-> frame 32 * (c ^ (l >> 3)) + l of the unit.  Writes ln(max(mel, eps)) for the frames below T,
// zeros for the other rows the unit covers in `out` (ragged clips).
FE_HD FE_INLINE void fe_k3_phase_mel(const FeKaldiParams& p, const FeK3Unit& t, float* sm, int lane, int warp) {
    typedef FeK3Cfg C;
    const int* __restrict__ seg = reinterpret_cast<const int*>(sm + C::OFF_SEG);
    const float2* __restrict__ w2 = reinterpret_cast<const float2*>(sm + C::OFF_MW);
    const float4* __restrict__ Pl = reinterpret_cast<const float4*>(sm + C::OFF_P) + lane;     // + 32 * bin
    const bool vec = (p.vec_ok & 2) != 0;
    int fr[4];
    bool live[4], inb[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        fr[c] = C::FT * (c ^ (lane >> 3)) + lane;
        live[c] = t.t0 + fr[c] < t.T;
        inb[c] = fr[c] < t.n_in;
    }
#pragma unroll 1
    for (int gi = 0; gi < 2; ++gi) {
        const int g = fe_k3_mel_groups[2 * warp + gi];
        if (g < 0) break;
        fe2 a01[4], a23[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { a01[i] = fe_zero<fe2>(); a23[i] = fe_zero<fe2>(); }
#pragma unroll
        for (int s = 0; s < 5; ++s) {
            const int k0 = seg[4 * g + s], k1 = seg[4 * g + s + 1];      // segment j = 4 g - 1 + s: seg[j + 1], seg[j + 2]
#pragma unroll 2
            for (int k = k0; k < k1; ++k) {
                const float4 v = Pl[32 * k];
                const float2 w = w2[k];
                if (s > 0) { a01[s - 1] = fe_fmac(w.x, fe_mk2(v.x, v.y), a01[s - 1]); a23[s - 1] = fe_fmac(w.x, fe_mk2(v.z, v.w), a23[s - 1]); }
                if (s < 4) { a01[s] = fe_fmac(w.y, fe_mk2(v.x, v.y), a01[s]); a23[s] = fe_fmac(w.y, fe_mk2(v.z, v.w), a23[s]); }
            }
        }
        float y[4][4];     // [filter][component]
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            y[i][0] = FE_LOG2(fmaxf(a01[i].x, C::EPS)) * C::LN2;
            y[i][1] = FE_LOG2(fmaxf(a01[i].y, C::EPS)) * C::LN2;
            y[i][2] = FE_LOG2(fmaxf(a23[i].x, C::EPS)) * C::LN2;
            y[i][3] = FE_LOG2(fmaxf(a23[i].y, C::EPS)) * C::LN2;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (!inb[c]) continue;
            float* q = p.out + t.o0 + (int64_t)fr[c] * C::NMEL + 4 * g;
            if (live[c]) fe_k3_stg4(q, y[0][c], y[1][c], y[2][c], y[3][c], vec);
            else fe_k3_stg4(q, 0.0f, 0.0f, 0.0f, 0.0f, vec);
        }
    }
}

// a unit without live frames that still covers rows of `out` (ragged batches): zeros
FE_HD FE_INLINE void fe_k3_zero_unit(const FeKaldiParams& p, const FeK3Unit& t, int lane, int warp) {
    typedef FeK3Cfg C;
    const bool vec = (p.vec_ok & 2) != 0;
    for (int r = warp; r < t.n_in; r += C::NW)
        if (lane < C::NMEL / 4) fe_k3_stg4(p.out + t.o0 + (int64_t)r * C::NMEL + 4 * lane, 0.0f, 0.0f, 0.0f, 0.0f, vec);
}

