// Whisper log-mel featurizer, third formulation: FFT scratch in Tensor Memory, waveform staged by the TMA engine.
//
// Replaces whisper/audio.py::log_mel_spectrogram (UPSTREAM openai-whisper; reached from
// /root/reference/Backend/services/cascaded_backend.py:158, Backend/speech_translation_pipeline.py:140,
// Backend/gtts_pipeline_fixed.py:64; same arithmetic as transformers/models/whisper/feature_extraction_whisper.py:135-164).
//
// Why a third kernel: the frame-pair kernel (fe_whisper2.cuh) is bound by the shared-memory pipe -- every 400-point
// FFT makes two trips through shared memory (4.9 k wavefronts per 64 frames, 62 % of the LSU pipe at 36 % of the HBM
// roofline).  Measured on B200 (profiles/r02_ubench_tmem_mma.txt): tcgen05.st / tcgen05.ld move 300-960 / 150-460
// bytes per cycle and SM -- 2-7 x the shared-memory pipe -- and run beside it.  So the inter-pass buffer A moves to
// Tensor Memory (420 of the 512 columns, lane = frame), the waveform is staged by cp.async.bulk (no loader warps), and
// shared memory only carries the raw samples (read once per frame) and the power spectrum P.
//
//   CTA     : persistent, 16 warps.  Warp w belongs to team q = w % 4 (= its SM sub-partition = its TMEM lane quadrant)
//             and is member j = w / 4 of it.  A CTA unit = 128 consecutive frames of one clip; team q owns frames
//             t0 + 32 q + lane (ONE frame per lane).  Packed FP32 (FADD2 / FMUL2 / FFMA2) now pairs two independent
//             sub-transforms of the SAME frame: two sub-sequences n2 in phase A, two rows k1 in phase B, two frames of
//             two different teams in phase M.
//   W[q]    : the team's samples, 34 rows of 160 (row r = samples of hop t0 + 32 q + r, starting 200 before the frame
//             centre), row pitch 164 floats: a lane's frame starts at its own row, and LDS.128 with lane stride 164
//             (= 4 x 41) is conflict-free.  Filled by 34 cp.async.bulk copies of 640 bytes (interior units) or by the
//             issuing warp element-wise (reflect padding, ragged ends: 2 team-units of ~94 per 30-s clip).
//   phase A : 5 tasks per team; task i loads x[20 n1 + 4 i .. + 3] (n1 = 0..19) with 20 LDS.128, windows them, runs two
//             packed real 20-point DFTs (sub-sequences (4i, 4i+1) and (4i+2, 4i+3)), twiddles by W_400^(n2 k1) and stores
//             Y[n2][k1] to Tensor Memory: 84 columns per task (tcgen05.st x4 / x8).
//   phase B : 6 tasks per team; rows (k1, k1 + 1) = (1,2) (3,4) (5,6) (7,8) (9,10) as one packed complex 20-point DFT
//             over n2 (80 tcgen05.ld x1: the load does the 2 x 2 transposition between the two packings for free),
//             row k1 = 0 as a scalar real-input DFT; |X|^2 -> P[bin][4 * lane + (q ^ (lane >> 3))] (conflict-free
//             32-bit stores; a lane's four teams sit in one 16-byte word).
//   phase M : all 16 warps, chunk = warp: one LDS.128 per tap fetches the power of FOUR frames (one per team), two
//             FFMA2 per tap; log10, (x + 4) / 4, four 32-byte-sector stores per filter and lane group, running min / max.
//   sync    : five kinds of mbarrier, no CTA-wide barrier in the steady state (fe_whisper3_kernel in b200feat.cu).
// Floor / clamp / ragged zero-fill: the same second, tiny fix-up kernel as before, on 128-frame units.
#pragma once
#include "fe_common.h"
#include "fe_whisper2.cuh"

struct FeW3Cfg {
    static constexpr int N = 400, HOP = 160, PADL = 200, R = 20;
    static constexpr int TEAMS = 4, MEMBERS = 4, NW = TEAMS * MEMBERS, NT = NW * 32;
    static constexpr int FT = 32;                        // frames per team and unit
    static constexpr int UF = TEAMS * FT;                // frames per CTA unit (128)
    static constexpr int ROWS = 34, PITCH = 164;         // W[q]: rows of HOP samples, pitch in floats
    static constexpr int WT = (ROWS * PITCH + 31) / 32 * 32;   // floats per team, padded: every W[q] is 128-byte aligned (tensor copies)
    static constexpr int ROW_BYTES = HOP * 4;
    static constexpr int TMAP_ROW0 = 2;                  // first frame whose row lies inside the clip (160 * 2 - 200 >= 0)
    static constexpr int NBINS = 200;
    static constexpr int ATASKS = 5, ACOLS = 84;         // Tensor Memory columns per phase-A task
    static constexpr int TMEM_COLS = 512;                // allocation (power of two >= ATASKS * ACOLS = 420)
    static constexpr int OFF_W = 0;
    static constexpr int OFF_P = OFF_W + TEAMS * WT;                 // P[bin][UF]
    static constexpr int OFF_MBAR = OFF_P + NBINS * UF;              // 14 mbarriers (8 bytes each) + the TMEM base slot
    static constexpr int NMBAR = 14;
    static constexpr int OFF_MDESC = OFF_MBAR + 2 * NMBAR + 8;       // (+ TMEM base slot, pad, four W-request counters)       // mel descriptors (2 ints per filter, <= 128 filters)
    static constexpr int OFF_MW = OFF_MDESC + 256;                   // padded taps
    static constexpr int MW_MAX = 576;
    static constexpr int SMEM_FLOATS = OFF_MW + MW_MAX;
    static constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
    static_assert(OFF_P % 4 == 0 && OFF_MBAR % 2 == 0 && OFF_MDESC % 4 == 0 && OFF_MW % 4 == 0, "alignment of the carve-up");
    static_assert(WT % 32 == 0 && PITCH % 4 == 0, "cp.async.bulk needs 16-byte aligned rows, cp.async.bulk.tensor a 128-byte aligned box");
    static_assert(SMEM_BYTES <= 232448, "more than 227 KB of shared memory");
    static constexpr float LOG_SCALE = FeW2Cfg::LOG_SCALE;
};

// ---------------------------------------------------------------- Tensor Memory as per-lane scratch
// Device: `base` = address of column 0 in the warp's own lane quadrant; every access is warp-wide (.sync.aligned), lane l
// touching its own lane row.  Host emulation: `row` = the 512 columns of one (team, lane).
struct FeTm {
#if defined(__CUDACC__)
    uint32_t base;
#else
    float* row;
#endif
};

FE_HD FE_INLINE void fe_tm_st4(const FeTm& tm, int col, float a, float b, float c, float d) {
#if defined(__CUDACC__)
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tm.base + col), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
#else
    tm.row[col] = a; tm.row[col + 1] = b; tm.row[col + 2] = c; tm.row[col + 3] = d;
#endif
}
FE_HD FE_INLINE void fe_tm_st2(const FeTm& tm, int col, float a, float b) {
#if defined(__CUDACC__)
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(tm.base + col), "f"(a), "f"(b) : "memory");
#else
    tm.row[col] = a; tm.row[col + 1] = b;
#endif
}
FE_HD FE_INLINE void fe_tm_st8(const FeTm& tm, int col, fe2 a, fe2 b, fe2 c, fe2 d) {
#if defined(__CUDACC__)
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(tm.base + col), "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y), "f"(d.x), "f"(d.y) : "memory");
#else
    float* r = tm.row + col;
    r[0] = a.x; r[1] = a.y; r[2] = b.x; r[3] = b.y; r[4] = c.x; r[5] = c.y; r[6] = d.x; r[7] = d.y;
#endif
}
// asynchronous on the device: the value is defined only after fe_tm_wait_ld() AND fe_tm_tie() on it
FE_HD FE_INLINE float fe_tm_ld1(const FeTm& tm, int col) {
#if defined(__CUDACC__)
    float v;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=f"(v) : "r"(tm.base + col) : "memory");
    return v;
#else
    return tm.row[col];
#endif
}
FE_HD FE_INLINE void fe_tm_wait_ld() {
#if defined(__CUDACC__)
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#endif
}
FE_HD FE_INLINE void fe_tm_wait_st() {
#if defined(__CUDACC__)
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
#endif
}
// ties a loaded register to the preceding wait: the compiler may not move a use of `v` above it
FE_HD FE_INLINE void fe_tm_tie(float& v) {
#if defined(__CUDACC__)
    asm volatile("" : "+f"(v));
#else
    (void)v;
#endif
}
FE_HD FE_INLINE void fe_tm_tie(fe2& v) { fe_tm_tie(v.x); fe_tm_tie(v.y); }

// ---------------------------------------------------------------- units
struct FeW3Unit {
    int b;          // clip
    int t0;         // first frame of the CTA unit
    int T;          // frames t0 <= f < T are computed and feed the clip's maximum
    int Ts;         // frames f < Ts are stored (== T except for the tail unit of a windowed call)
    int n_in;       // output columns the unit covers in `out` (0 for the tail unit)
    int pitch;      // output row pitch in elements
    int64_t o0;     // element offset in `out` of (mel row 0, frame t0)
    int64_t len;    // valid samples
    int64_t lp;     // len + padding
    int silent;     // every frame of the unit reads only the zeros of `padding`: nothing to transform (see fe_w3_silent)
};

// A unit whose first frame starts at or behind the clip's last sample (t0 * HOP - PADL >= len) sees nothing but the
// zeros whisper.audio pads with (padding = 480000 in whisper.transcribe: the second half of every 30-s clip's mel, more
// for short clips).  Its power spectrum is exactly 0, its log-mel -inf before the floor, and the value the reference
// ends up with is the clip's clamp threshold max(-10, max - 8) -- known only once the whole clip is done.  The kernel
// skips such units (no copy, no transform, no store) and the fix-up kernel, which knows the threshold, fills them with
// one write; before, they were transformed, stored as -inf, read back and rewritten (36 % of the kernels' time on
// ragged padded speech, profiles/r02_bench_realistic.txt).  padding >= PADL keeps the reflect padding at the right
// edge (it mirrors at most HOP + PADL samples before len + padding) inside the zeros as well.
FE_HD FE_INLINE int fe_w3_silent(const FeWhisper2Params& p, const FeW3Unit& t) {
    typedef FeW3Cfg C;
    return p.padding >= 2 * C::PADL && t.t0 < t.T && (int64_t)t.t0 * C::HOP - C::PADL >= t.len;
}

// Plain call: clip b has units_per_clip units of 128 frames, rows of T_out.  Windowed call (p.win_frames > 0): window w of
// clip b is its own (n_mels, win_frames) block; a unit never straddles two windows (units per window =
// ceil(win_frames / 128), the last one shorter), and one extra TAIL unit per clip computes the frames just behind the last
// window without storing them: they can hold the clip's maximum (whisper takes the maximum over the whole padded mel).
// the decode proper, from (clip, unit within the clip): no division by a run-time value on the plain path -- the
// kernel advances (clip, within) incrementally (FeW3Pos), because a full decode per warp and unit was 7 % of the
// kernel's time (profiles/r02_w3_trace_v7.txt: 850 cycles of every 12.6 k-cycle unit)
// Plain calls: the first frame of clip b behind which every 128-frame unit is either pure padding or lies behind the
// clip's last frame -- the part of each output row the fix-up writes as ONE contiguous run (threshold up to frame Tc,
// 0 behind).  Returns T_out when there is no such part.
FE_HD FE_INLINE int64_t fe_w3_tail_start(const FeWhisper2Params& p, int b, int64_t* Tc_out) {
    typedef FeW3Cfg C;
    const int64_t len = fe_clip_len(p.lengths, b, p.L), lp = len + p.padding;
    int64_t Tc = lp / C::HOP;
    if (Tc > p.T_out) Tc = p.T_out;
    int64_t f0 = (Tc + C::UF - 1) / C::UF * C::UF;                              // first unit behind the last frame
    if (p.padding >= 2 * C::PADL) {
        const int64_t fs = (len + C::PADL + C::HOP - 1) / C::HOP;                // first frame with t * HOP - PADL >= len
        const int64_t s0 = (fs + C::UF - 1) / C::UF * C::UF;
        if (s0 < f0) f0 = s0;
    }
    *Tc_out = Tc;
    return f0 < p.T_out ? f0 : p.T_out;
}

// FIXUP = false (the featurizer kernel, the emulation of its phases): a silent unit is returned with T = Ts = t0, i.e. as
// a unit without frames to compute, and the kernel never looks at `silent`.  FIXUP = true: T / Ts as they are.
// PAD = false: the caller knows that the call has no padding to speak of (padding < 2 * PADL), so no unit is silent; the
// featurizer kernel is instantiated both ways, because the test costs it 1 % (1.040 against 1.029 ms per 1024 x 30 s,
// a/b a/b on one box) that a call without padding need not pay.
template <bool FIXUP = false, bool PAD = true>
FE_HD FE_INLINE FeW3Unit fe_w3_unit_at(const FeWhisper2Params& p, int b, int within) {
    typedef FeW3Cfg C;
    FeW3Unit t;
    t.b = b;
    t.len = fe_clip_len(p.lengths, t.b, p.L);
    t.lp = t.len + p.padding;
    const int64_t Tc = t.lp < ((int64_t)1 << 32) ? (int64_t)((uint32_t)t.lp / (uint32_t)C::HOP) : t.lp / C::HOP;
    if (p.win_frames > 0) {
        const int W = p.win_frames, upw = (W + C::UF - 1) / C::UF;
        if (within < p.n_windows * upw) {
            const int w = within / upw, j = within - w * upw;
            t.t0 = w * W + j * C::UF;
            const int64_t wend = (int64_t)(w + 1) * W;
            t.T = (int)(Tc < wend ? Tc : wend);
            t.Ts = t.T;
            t.n_in = W - j * C::UF < C::UF ? W - j * C::UF : C::UF;
            t.pitch = W;
            t.o0 = (int64_t)t.b * p.out_batch_stride + (int64_t)w * p.n_mels * W + j * C::UF;
        } else {
            t.t0 = p.n_windows * W;
            const int64_t e = (int64_t)t.t0 + C::UF;
            t.T = (int)(Tc < e ? Tc : e);
            t.Ts = t.t0; t.n_in = 0; t.pitch = W; t.o0 = 0;
        }
        t.silent = PAD ? fe_w3_silent(p, t) : 0;
        if (PAD && !FIXUP && t.silent) t.T = t.Ts = t.t0;
        return t;
    }
    t.t0 = within * C::UF;
    t.T = (int)(Tc < p.T_out ? Tc : p.T_out);
    t.Ts = t.T;
    t.n_in = (int)(p.T_out - t.t0 < C::UF ? p.T_out - t.t0 : C::UF);
    t.pitch = (int)p.T_out;
    t.o0 = (int64_t)t.b * p.out_batch_stride + t.t0;
    t.silent = PAD ? fe_w3_silent(p, t) : 0;
    if (PAD && !FIXUP && t.silent) t.T = t.Ts = t.t0;
    return t;
}
// Unit order: POSITION-major, u = within * B + clip.  With clip-major order and gridDim.x = 148, a CTA would only ever
// see 6 of a 30-s clip's 24 unit positions (gcd(148, 24) = 4), and the CTAs that keep drawing a clip's first / last unit
// (edge rows, a 56-frame unit) finish 10 % apart (profiles/r02_w3_v8: 12 % of the SM cycles idle at the end).
template <bool FIXUP = false>
FE_HD FE_INLINE FeW3Unit fe_w3_unit(const FeWhisper2Params& p, int u) {
    const int within = u / p.B;
    return fe_w3_unit_at<FIXUP>(p, u - within * p.B, within);
}
// position of a CTA's current unit; one step = gridDim.x units = (step_w positions, step_b clips) with a carry
struct FeW3Pos { int b, within; };
FE_HD FE_INLINE FeW3Pos fe_w3_pos_next(const FeWhisper2Params& p, FeW3Pos q) {
    q.b += p.step_b; q.within += p.step_w;
    if (q.b >= p.B) { q.b -= p.B; q.within += 1; }
    return q;
}

// How W[q] of a unit is filled, row by row: rows that lie completely inside the clip's samples (and are 16-byte
// aligned in global memory) are bulk copies; the others -- reflect padding at the clip's edges, the zeros of `padding`
// and of ragged ends -- are written element-wise by the issuing warp (2 + 3 rows of a 30-s clip's 3196).
// fe_w3_team_live: the team has at least one frame to compute.
FE_HD FE_INLINE bool fe_w3_team_live(const FeW3Unit& t, int q) { return t.t0 + FeW3Cfg::FT * q < t.T; }
// (Stepping over such units in the kernel's loop -- every warp seeking its CTA's next unit with work instead of walking
// the empty ones through the mbarrier pipeline -- was measured and dropped: 0.517 against 0.530 ms on the ragged
// padded leg, but 1.18 against 1.03 ms on the headline, the loop being sensitive to its shape.)
FE_HD FE_INLINE int64_t fe_w3_row0(const FeW3Unit& t, int q) {     // global sample index of row 0, column 0
    return ((int64_t)t.t0 + FeW3Cfg::FT * q) * FeW3Cfg::HOP - FeW3Cfg::PADL;
}
FE_HD FE_INLINE bool fe_w3_row_bulk(const FeWhisper2Params& p, const FeW3Unit& t, int64_t g0, int r) {
    const int64_t g = g0 + (int64_t)r * FeW3Cfg::HOP;
    return (p.vec_ok & 1) && g >= 0 && g + FeW3Cfg::HOP <= t.len;
}

// element-wise rows of one team, by ONE warp: every row outside [r_lo, r_hi), four rows (20 independent loads per lane)
// in flight at a time
FE_HD FE_INLINE void fe_w3_fill_rows_manual(const float* __restrict__ x, int64_t g0, int64_t len, int64_t lp, int r_lo, int r_hi,
                                            float* wq, int lane) {
    typedef FeW3Cfg C;
    constexpr int NB = 4;
    int rows[NB];
    int n = 0;
#pragma unroll 1
    for (int r = 0; r <= C::ROWS; ++r) {
        if (r < C::ROWS && !(r >= r_lo && r < r_hi)) rows[n++] = r;
        if (n == NB || (r == C::ROWS && n > 0)) {
            float v[NB][5];
#pragma unroll
            for (int k = 0; k < NB; ++k)
#pragma unroll
                for (int j = 0; j < 5; ++j)
                    v[k][j] = k < n ? fe_fetch_reflect(x, g0 + (int64_t)rows[k] * C::HOP + lane + 32 * j, len, lp) : 0.0f;
#pragma unroll
            for (int k = 0; k < NB; ++k)
#pragma unroll
                for (int j = 0; j < 5; ++j)
                    if (k < n) wq[rows[k] * C::PITCH + lane + 32 * j] = v[k][j];
            n = 0;
        }
    }
}
// Tensor-map launches: the box copy is issued for EVERY live team (rows outside the map are zero-filled by the TMA), and
// the rows it cannot get right are patched by the team's own four warps after the copy has landed: a row is right when
// it lies inside the map and inside the clip's valid samples, or outside the map and inside the clip's zero padding.
FE_HD FE_INLINE bool fe_w3_row_from_box(const FeWhisper2Params& p, const FeW3Unit& t, int f0, int r) {
    typedef FeW3Cfg C;
    const int64_t g = ((int64_t)f0 + r) * C::HOP - C::PADL;
    const int ri = f0 - C::TMAP_ROW0 + r;
    const bool inb = ri >= 0 && ri < p.tmap_rows;
    if (inb) return g >= 0 && g + C::HOP <= t.len;
    return ri >= p.tmap_rows && g >= t.len && g + C::HOP <= t.lp;
}
// rows [0, ROWS) of a box are right from row r_lo on and up to r_hi, or (zero padding) all the way: returns true when
// every row is right (the common case, decided from the first and the last row)
FE_HD FE_INLINE bool fe_w3_box_complete(const FeWhisper2Params& p, const FeW3Unit& t, int q) {
    typedef FeW3Cfg C;
    const int f0 = t.t0 + C::FT * q;
    const int64_t g0 = (int64_t)f0 * C::HOP - C::PADL, g1 = g0 + (int64_t)C::ROWS * C::HOP;
    const int r0 = f0 - C::TMAP_ROW0;
    if (r0 >= 0 && r0 + C::ROWS <= p.tmap_rows && g1 <= t.len) return true;          // inside the map and the clip
    if (r0 >= p.tmap_rows && g0 >= t.len && g1 <= t.lp) return true;                 // all zero padding
    return false;
}
// member `member` patches rows member, member + 4, ...; a team barrier must follow
FE_HD FE_INLINE void fe_w3_patch_rows(const FeWhisper2Params& p, const FeW3Unit& t, int q, float* wq, int member, int lane) {
    typedef FeW3Cfg C;
    const int f0 = t.t0 + C::FT * q;
    const float* __restrict__ x = p.wave + (int64_t)t.b * p.wave_stride;
    const int64_t g0 = (int64_t)f0 * C::HOP - C::PADL;
#pragma unroll 1
    for (int r = member; r < C::ROWS; r += C::MEMBERS) {
        if (fe_w3_row_from_box(p, t, f0, r)) continue;
        float v[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) v[j] = fe_fetch_reflect(x, g0 + (int64_t)r * C::HOP + lane + 32 * j, t.len, t.lp);
#pragma unroll
        for (int j = 0; j < 5; ++j) wq[r * C::PITCH + lane + 32 * j] = v[j];
    }
}
// host emulation of the box copy: rows inside the map are copied (PITCH floats each), the others zero-filled
FE_HD FE_INLINE void fe_w3_box_host(const FeWhisper2Params& p, const FeW3Unit& t, int q, float* wq) {
    typedef FeW3Cfg C;
    const int f0 = t.t0 + C::FT * q;
    const float* x = p.wave + (int64_t)t.b * p.wave_stride;
    for (int r = 0; r < C::ROWS; ++r) {
        const int ri = f0 - C::TMAP_ROW0 + r;
        const int64_t g = ((int64_t)f0 + r) * C::HOP - C::PADL;
        for (int c = 0; c < C::PITCH; ++c) wq[r * C::PITCH + c] = (ri >= 0 && ri < p.tmap_rows) ? x[g + c] : 0.0f;
    }
}

// host emulation: the same split into bulk rows and element-wise rows as fe_w3_issue_w (b200feat.cu)
FE_HD FE_INLINE void fe_w3_fill_manual(const FeWhisper2Params& p, const FeW3Unit& t, int q, float* wq, int lane) {
    typedef FeW3Cfg C;
    const int64_t g0 = fe_w3_row0(t, q);
    int r_lo = 0, r_hi = 0;
    for (int r = 0; r < C::ROWS; ++r)
        if (fe_w3_row_bulk(p, t, g0, r)) { if (r_hi == r_lo) r_lo = r; r_hi = r + 1; }
    fe_w3_fill_rows_manual(p.wave + (int64_t)t.b * p.wave_stride, g0, t.len, t.lp, r_lo, r_hi, wq, lane);
}

// host emulation of the bulk copies (the device issues cp.async.bulk, b200feat.cu)
FE_HD FE_INLINE void fe_w3_fill_rows_host(const FeWhisper2Params& p, const FeW3Unit& t, int q, float* wq) {
    typedef FeW3Cfg C;
    const int64_t g0 = fe_w3_row0(t, q);
    const float* x = p.wave + (int64_t)t.b * p.wave_stride;
    for (int r = 0; r < C::ROWS; ++r)
        if (fe_w3_row_bulk(p, t, g0, r))
            for (int c = 0; c < C::HOP; ++c) wq[r * C::PITCH + c] = x[g0 + (int64_t)r * C::HOP + c];
}

template <int NMELS>
FE_HD FE_INLINE void fe_w3_copy_tables(float* sm, int tid, int nthreads) {
    typedef FeW3Cfg C;
    static_assert(FE_MELLOOP3_W80_NW <= C::MW_MAX && FE_MELLOOP3_W128_NW <= C::MW_MAX, "MW_MAX too small");
    const int* __restrict__ desc = NMELS == 80 ? fe_melloop3_w80_desc : fe_melloop3_w128_desc;
    const float* __restrict__ w = NMELS == 80 ? fe_melloop3_w80_w : fe_melloop3_w128_w;
    const int nw = NMELS == 80 ? FE_MELLOOP3_W80_NW : FE_MELLOOP3_W128_NW;
    int* d = reinterpret_cast<int*>(sm + C::OFF_MDESC);
    for (int i = tid; i < 2 * NMELS; i += nthreads) d[i] = desc[i];
    for (int i = tid; i < nw; i += nthreads) sm[C::OFF_MW + i] = w[i];
}

// ---------------------------------------------------------------- task maps
// Phase A is cut into ten HALF-UNITS (hu = 0..9: sub-sequences n2 = 2 hu, 2 hu + 1, one packed real 20-point DFT);
// phase B into six tasks (pairs p = 0..4 = rows (2p+1, 2p+2); 5 = row 0).  Members 0 and 1 run three half-units and
// one light B task, members 2 and 3 two half-units and two pair tasks: A(u) is complete after 1.5 task times (not 2),
// and the mel projection is shared equally (chunk = warp).
FE_HD FE_INLINE int fe_w3_a_single(int member) { return member < 2 ? 8 + member : -1; }     // besides the pair i = member
FE_HD FE_INLINE int fe_w3_b_count(int member) { return member < 2 ? 1 : 2; }
FE_HD FE_INLINE int fe_w3_b_task(int member, int k) { return member == 0 ? 4 : (member == 1 ? 5 : 2 * (member - 2) + k); }

// ---------------------------------------------------------------- phase A
// fe_w3_tab: per pair of half-units (i = hu / 2) 160 floats = window w[20 n1 + 4 i + c] as [n1][c] (80), then for
// k1 = 1..10 the twiddles W_400^((4 i + c) k1) as [k1][wr_c0, wr_c1, wi_c0, wi_c1, wr_c2, wr_c3, wi_c2, wi_c3] (80);
// half-unit hu uses c = 2 (hu % 2), 2 (hu % 2) + 1
// the part after the windowed loads: real 20-point DFT of the packed pair of sub-sequences, twiddles, Tensor Memory stores
template <class PreStore>
FE_HD FE_INLINE void fe_w3_a_core(const fe2* x, const FeTm& tm, int hu, PreStore pre_store) {
    typedef FeW3Cfg C;
    const fe2* __restrict__ tw = reinterpret_cast<const fe2*>(fe_w3_tab + 160 * (hu >> 1)) + 40 + 2 * (hu & 1);
    fe2 yr[11], yi[11];
    fe_rdft20_h11(x, yr, yi);
    pre_store();   // Tensor Memory may still be read by phase B of the previous unit
    const int col = C::ACOLS * (hu >> 1) + 8 * (hu & 1);
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        fe2 re[2], im[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int k1 = 2 * j + 1 + h;
            const fe2 wr = tw[4 * (k1 - 1)], wi = tw[4 * (k1 - 1) + 1];
            if (k1 < 10) {
                re[h] = fe_fma(fe_neg(wi), yi[k1], fe_mul(wr, yr[k1]));     // wr*yr - wi*yi
                im[h] = fe_fma(wi, yr[k1], fe_mul(wr, yi[k1]));             // wr*yi + wi*yr
            } else {                                                        // k1 = 10: the real DFT's Nyquist row
                re[h] = fe_mul(wr, yr[10]);
                im[h] = fe_mul(wi, yr[10]);
            }
        }
        fe_tm_st8(tm, col + 4 + 16 * j, re[0], im[0], re[1], im[1]);
    }
    fe_tm_st2(tm, C::ACOLS * (hu >> 1) + 2 * (hu & 1), yr[0].x, yr[0].y);
}

// ONE half-unit: 64-bit loads at lane stride 164 floats hit every bank pair twice (the 16-byte row alignment the tensor
// copy needs leaves no conflict-free pitch for them): used for half-units 8 and 9 only
template <class PreStore>
FE_HD FE_INLINE void fe_w3_stage_a(const float* wq, const FeTm& tm, int lane, int hu, PreStore pre_store) {
    typedef FeW3Cfg C;
    // sample n = 20 n1 + 2 hu + c of this lane's frame: row lane + n / 160, column n % 160; n / 160 == n1 / 8
    const fe2* __restrict__ s = reinterpret_cast<const fe2*>(wq + C::PITCH * lane) + hu;
    const fe2* __restrict__ win = reinterpret_cast<const fe2*>(fe_w3_tab + 160 * (hu >> 1)) + (hu & 1);
    fe2 x[C::R];
#pragma unroll
    for (int n1 = 0; n1 < C::R; ++n1) x[n1] = fe_mul(win[2 * n1], s[10 * n1 + 2 * (n1 / 8)]);
    fe_w3_a_core(x, tm, hu, pre_store);
}

// the PAIR of half-units (2 i, 2 i + 1): their samples 20 n1 + 4 i .. + 3 are one 16-byte word per n1, and 128-bit loads
// at lane stride 164 floats (= 4 x 41) are conflict-free -- half the shared-memory wavefronts of two single half-units
// (profiles/r02_whisper3_ncu_summary.txt: 31 % of the kernel's wavefronts were these conflicts)
template <class PreStore>
FE_HD FE_INLINE void fe_w3_stage_a_pair(const float* wq, const FeTm& tm, int lane, int i, PreStore pre_store) {
    typedef FeW3Cfg C;
    const float4* __restrict__ s = reinterpret_cast<const float4*>(wq + C::PITCH * lane) + i;
    const float4* __restrict__ win = reinterpret_cast<const float4*>(fe_w3_tab + 160 * i);
    fe2 xa[C::R], xb[C::R];
#pragma unroll
    for (int n1 = 0; n1 < C::R; ++n1) {
        const float4 v = s[5 * n1 + n1 / 8];
        const float4 w = win[n1];
        xa[n1] = fe_mul(fe_mk2(w.x, w.y), fe_mk2(v.x, v.y));
        xb[n1] = fe_mul(fe_mk2(w.z, w.w), fe_mk2(v.z, v.w));
    }
    fe_w3_a_core(xa, tm, 2 * i, pre_store);
    fe_w3_a_core(xb, tm, 2 * i + 1, pre_store);
}

// Tensor Memory column of Y[n2][k1]: task n2 / 4, sub-sequence c = n2 % 4
FE_HD FE_INLINE constexpr int fe_w3_col_re(int n2, int k1) {
    return FeW3Cfg::ACOLS * (n2 / 4) + (k1 == 0 ? (n2 % 4)
                                                : 4 + 16 * ((k1 - 1) / 2) + 8 * ((n2 % 4) / 2) + 4 * ((k1 - 1) % 2) + ((n2 % 4) % 2));
}
FE_HD FE_INLINE constexpr int fe_w3_col_im(int n2, int k1) { return fe_w3_col_re(n2, k1) + 2; }

// ---------------------------------------------------------------- phase B
// P[bin][frame of the unit], frame = 32 * team + lane: a warp's 32-bit stores are conflict-free, and phase M reads four
// CONSECUTIVE frames with one LDS.128
FE_HD FE_INLINE int fe_w3_p_slot(int lane, int team) { return FeW3Cfg::FT * team + lane; }

// rows (k1a, k1b) = (2 pr + 1, 2 pr + 2): one packed complex 20-point DFT over n2
template <class PostLoad, class PreStore>
FE_HD FE_INLINE void fe_w3_stage_b_pair(float* P, const FeTm& tm, int lane, int team, int pr, PostLoad post_load, PreStore pre_store) {
    typedef FeW3Cfg C;
    FeTm t = tm;
#if defined(__CUDACC__)
    t.base += 16 * pr;
#else
    t.row += 16 * pr;
#endif
    fe2 xr[20], xi[20], yr[20], yi[20];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) {
        const int c = fe_w3_col_re(n2, 1);
        xr[n2].x = fe_tm_ld1(t, c); xi[n2].x = fe_tm_ld1(t, c + 2);
        xr[n2].y = fe_tm_ld1(t, c + 4); xi[n2].y = fe_tm_ld1(t, c + 6);
    }
    fe_tm_wait_ld();
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) { fe_tm_tie(xr[n2]); fe_tm_tie(xi[n2]); }
    post_load();
    fe_cdft20(xr, xi, yr, yi);
    fe2 pw[20];
#pragma unroll
    for (int k = 0; k < 20; ++k) pw[k] = fe_fma(yr[k], yr[k], fe_mul(yi[k], yi[k]));
    pre_store();
    const int k1a = 2 * pr + 1;
    float* __restrict__ pa = P + fe_w3_p_slot(lane, team) + C::UF * k1a;            // bins k1a + 20 j  (and k1a + 1 + 20 j)
    float* __restrict__ pb = P + fe_w3_p_slot(lane, team) + C::UF * (C::R - k1a);   // bins (20 - k1a) + 20 j  (one lower for k1b)
#pragma unroll
    for (int j = 0; j < 10; ++j) {
        pa[C::UF * C::R * j] = pw[j].x;
        pa[C::UF * C::R * j + C::UF] = pw[j].y;
        pb[C::UF * C::R * j] = pw[19 - j].x;
    }
    if (pr < 4) {      // k1b = 10 is its own mirror: bins 10 + 20 j were just written
#pragma unroll
        for (int j = 0; j < 10; ++j) pb[C::UF * C::R * j - C::UF] = pw[19 - j].y;
    }
}

// row k1 = 0: real input, bins 20 j
template <class PostLoad, class PreStore>
FE_HD FE_INLINE void fe_w3_stage_b_k0(float* P, const FeTm& tm, int lane, int team, PostLoad post_load, PreStore pre_store) {
    typedef FeW3Cfg C;
    float xr[20], yr[10], yi[10];
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) xr[n2] = fe_tm_ld1(tm, fe_w3_col_re(n2, 0));
    fe_tm_wait_ld();
#pragma unroll
    for (int n2 = 0; n2 < 20; ++n2) fe_tm_tie(xr[n2]);
    post_load();
    yi[0] = 0.0f;
    fe_rdft20_h10(xr, yr, yi);
    pre_store();
    float* __restrict__ pa = P + fe_w3_p_slot(lane, team);
#pragma unroll
    for (int j = 0; j < 10; ++j) pa[C::UF * C::R * j] = fmaf(yr[j], yr[j], yi[j] * yi[j]);
}

template <class PostLoad, class PreStore>
FE_HD FE_INLINE void fe_w3_stage_b(float* P, const FeTm& tm, int lane, int team, int task, PostLoad post_load, PreStore pre_store) {
    if (task < 5) fe_w3_stage_b_pair(P, tm, lane, team, task, post_load, pre_store);
    else fe_w3_stage_b_k0(P, tm, lane, team, post_load, pre_store);
}

// ---------------------------------------------------------------- phase M
// A lane handles the four consecutive frames t0 + 4 * lane + c, c = 0..3 (one 16-byte word of every P row): one LDS.128
// and two FFMA2 per tap, and -- when the whole unit is stored and rows are 16-byte aligned (VEC) -- ONE 128-bit store
// per filter (a warp writes 512 contiguous bytes of the mel row).
struct FeW3MelCtx {
    const char* Pl;      // P + 16 * lane bytes
    const char* wtab;    // padded weights (shared memory)
    const int2* desc;    // descriptors (shared memory)
    char* ob;            // out + offset of (mel row 0, frame t0 + 4 * lane)
    unsigned rs;         // output row pitch in bytes
    fe2 one01, one23;    // 1 for live frames, NaN for dead ones
    bool lv[4];          // frame is stored (scalar path)
    float mn, mx;
};

template <int BYTE_OFF>
FE_HD FE_INLINE void fe_stg_v4(char* q, fe2 a, fe2 b) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.v4.f32 [%0+%1], {%2, %3, %4, %5};" ::"l"(q), "n"(BYTE_OFF), "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
#else
    float* o = reinterpret_cast<float*>(q + BYTE_OFF);
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
#endif
}
template <int BYTE_OFF>
FE_HD FE_INLINE void fe_stg_h4(char* q, fe2 a, fe2 b) {
    const unsigned lo = (unsigned)fe_f2h_rn(a.x) | ((unsigned)fe_f2h_rn(a.y) << 16);
    const unsigned hi = (unsigned)fe_f2h_rn(b.x) | ((unsigned)fe_f2h_rn(b.y) << 16);
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.v2.b32 [%0+%1], {%2, %3};" ::"l"(q), "n"(BYTE_OFF), "r"(lo), "r"(hi));
#else
    unsigned* o = reinterpret_cast<unsigned*>(q + BYTE_OFF);
    o[0] = lo; o[1] = hi;
#endif
}

template <int CAP, bool OUT16, bool VEC>
FE_HD FE_INLINE void fe_w3_mel_class(FeW3MelCtx& c, int i0, int i1) {
    typedef FeW3Cfg C;
    // two filters in flight (four for the short classes): one filter is a dependent chain descriptor -> taps -> FFMA2
    // -> lg2 -> store of ~300 cycles, and a warp has only eight of them per unit
#pragma unroll(!VEC ? 1 : (CAP <= 3 ? 4 : 2))
    for (int i = i0; i < i1; ++i) {
        const int2 d = c.desc[i];      // d.x = 512 * k0 | mel row, d.y = byte offset of the weights
        const float4* __restrict__ P = reinterpret_cast<const float4*>(c.Pl + (d.x & ~0x1ff));
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
        float4 v = P[0];
        fe2 a01 = fe_mulc(w[0], fe_mk2(v.x, v.y)), a23 = fe_mulc(w[0], fe_mk2(v.z, v.w));
#pragma unroll
        for (int j = 1; j < CAP; ++j) {
            v = P[(C::UF / 4) * j];
            a01 = fe_fmac(w[j], fe_mk2(v.x, v.y), a01);
            a23 = fe_fmac(w[j], fe_mk2(v.z, v.w), a23);
        }
        const fe2 y01 = fe_fmac(C::LOG_SCALE, fe_mk2(fe_w2_lg2(a01.x), fe_w2_lg2(a01.y)), c.one01);   // (log10(v) + 4) / 4
        const fe2 y23 = fe_fmac(C::LOG_SCALE, fe_mk2(fe_w2_lg2(a23.x), fe_w2_lg2(a23.y)), c.one23);
        char* o = fe_row_ptr(c.ob, c.rs, (unsigned)d.x & 0x1ffu);
        if (VEC) {     // min / max below stay on the unrounded values: the fix-up decides in float32 like the reference
            if (OUT16) fe_stg_h4<0>(o, y01, y23);
            else fe_stg_v4<0>(o, y01, y23);
        } else if (OUT16) {
            if (c.lv[0]) fe_stg_h<0>(o, y01.x);
            if (c.lv[1]) fe_stg_h<2>(o, y01.y);
            if (c.lv[2]) fe_stg_h<4>(o, y23.x);
            if (c.lv[3]) fe_stg_h<6>(o, y23.y);
        } else {
            if (c.lv[0]) fe_stg<0>(o, y01.x);
            if (c.lv[1]) fe_stg<4>(o, y01.y);
            if (c.lv[2]) fe_stg<8>(o, y23.x);
            if (c.lv[3]) fe_stg<12>(o, y23.y);
        }
        c.mx = fe_max3(c.mx, y01.x, y01.y); c.mx = fe_max3(c.mx, y23.x, y23.y);
        c.mn = fe_min3(c.mn, y01.x, y01.y); c.mn = fe_min3(c.mn, y23.x, y23.y);
    }
}

template <int NMELS> struct FeW3MelCaps;
template <> struct FeW3MelCaps<80> {
    static constexpr int N = FE_MELLOOP3_W80_NCLS;
    static FE_HD constexpr int cap(int k) { constexpr int c[N] = {FE_MELLOOP3_W80_CAPS}; return c[k]; }
};
template <> struct FeW3MelCaps<128> {
    static constexpr int N = FE_MELLOOP3_W128_NCLS;
    static FE_HD constexpr int cap(int k) { constexpr int c[N] = {FE_MELLOOP3_W128_CAPS}; return c[k]; }
};
template <int NMELS, bool OUT16, bool VEC, int K, bool END = (K == FeW3MelCaps<NMELS>::N)>
struct FeW3MelRun {
    static FE_HD FE_INLINE void run(FeW3MelCtx& c, const int* r) {
        fe_w3_mel_class<FeW3MelCaps<NMELS>::cap(K), OUT16, VEC>(c, r[K], r[K + 1]);
        FeW3MelRun<NMELS, OUT16, VEC, K + 1>::run(c, r);
    }
};
template <int NMELS, bool OUT16, bool VEC, int K>
struct FeW3MelRun<NMELS, OUT16, VEC, K, true> {
    static FE_HD FE_INLINE void run(FeW3MelCtx&, const int*) {}
};

// returns this thread's (min, max) over its live outputs
template <int NMELS, bool OUT16>
FE_HD FE_INLINE fe2 fe_w3_phase_mel(const FeWhisper2Params& p, const FeW3Unit& t, float* sm, int lane, int warp) {
    typedef FeW3Cfg C;
    constexpr int NCLS = FeW3MelCaps<NMELS>::N;
    const int* __restrict__ ranges = (NMELS == 80 ? fe_melloop3_w80_ranges : fe_melloop3_w128_ranges) + (NCLS + 1) * warp;
    FeW3MelCtx c;
    c.Pl = reinterpret_cast<const char*>(sm + C::OFF_P + 4 * lane);
    c.wtab = reinterpret_cast<const char*>(sm + C::OFF_MW);
    c.desc = reinterpret_cast<const int2*>(sm + C::OFF_MDESC);
    constexpr int ESZ = OUT16 ? 2 : 4;
    c.rs = (unsigned)t.pitch * (unsigned)ESZ;     // row pitch in bytes; the host guarantees pitch < 2^24
    union { unsigned u; float f; } qnan;
    qnan.u = 0x7fc00000u;
    float one[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int f = t.t0 + 4 * lane + k;
        c.lv[k] = f < t.Ts;
        one[k] = f < t.T ? 1.0f : qnan.f;      // dead frames carry NaN through min / max
    }
    c.ob = reinterpret_cast<char*>(p.out) + (t.o0 + 4 * lane) * ESZ;
    c.one01 = fe_mk2(one[0], one[1]); c.one23 = fe_mk2(one[2], one[3]);
    c.mn = 3.0e38f; c.mx = -3.0e38f;
    int r[NCLS + 1];
#pragma unroll
    for (int k = 0; k <= NCLS; ++k) r[k] = ranges[k];
    // whole unit stored and 16-byte (float16: 8-byte) aligned rows: one vector store per filter
    const bool vec = t.Ts - t.t0 >= C::UF && (p.vec_ok & 2) != 0 && (t.pitch & 3) == 0 && (t.o0 & 3) == 0;
    if (vec) FeW3MelRun<NMELS, OUT16, true, 0>::run(c, r);
    else FeW3MelRun<NMELS, OUT16, false, 0>::run(c, r);
    return fe_mk2(c.mn, c.mx);
}
