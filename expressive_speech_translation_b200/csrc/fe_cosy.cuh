// CosyVoice 24 kHz prompt mel (matcha mel_spectrogram): phases of the fused kernel.
//
// Replaces matcha/utils/audio.py::mel_spectrogram(y, n_fft=1920, num_mels=80, sampling_rate=24000,
// hop_size=480, win_size=1920, fmin=0, fmax=8000, center=False) -- UPSTREAM Matcha-TTS, parameters bound
// at /root/reference/greek_sft.yaml:55-63, used by the CosyVoice frontend behind
// /root/reference/Docker/cosyvoice_api.py:128.  reflect-pad 720 | periodic Hann STFT | sqrt(re^2+im^2+1e-9)
// | Slaney mel (bins >= 640 carry zero weight) | ln(max(., 1e-5)).
//
// N = 1920 = 16 x 12 x 10, real input, three shared-memory stages.  A frame in flight needs 8.6 KB, so a
// CTA holds FT = 16 frames: a warp is two task slots x 16 frames (lane & 15 = frame, lane >> 4 = slot).
//   n = 120*n1 + m,  m = 10*n2 + n3 ;  k = k1 + 16*k',  k' = k2 + 12*k3
//   phase A : real radix-16 DFT over n1 for each m, window fused, k1 = 0..8, times W_1920^(m*k1) -> A[f][k1][m]
//   phase B : complex radix-12 DFT over n2 for each (k1, n3), times W_120^(n3*k2), in place   -> A[f][k1][10*k2+n3]
//   phase C : complex radix-10 DFT over n3 for each (k1, k2); magnitude of the bins <= 639 (direct, and
//             mirrored 1920 - k for k1 not in {0, 8})                                        -> P[bin][f]
//   phase M : unrolled sparse mel projection, ln(max(., 1e-5)), store.
#pragma once
#include "fe_common.h"

// The same three-stage kernel serves two featurizers; the phases below are templates over a configuration:
//   FeCCfg  CosyVoice / Matcha mel          N = 1920 = 16 x 12 x 10, hop 480, reflect pad 720 (center=False + manual pad)
//   FeLCfg  Wav2Lip / Diff2Lip melspectrogram N = 800 = 16 x 5 x 10, hop 200, librosa center pad 400, pre-emphasis 0.97
//           (/root/reference/Backend/diff2lip/audio/audio.py:45-51, hparams.py:22-65)
// Both have HOP / M = 4 (M = R2 * R3), which the staging layout of phase A relies on.
template <class P>
struct FeS3Cfg : P {
    static constexpr int R1 = 16;
    static constexpr int M = P::R2 * P::R3;
    static constexpr int NK1 = R1 / 2 + 1;               // 9
    // Frames per tile: half a warp, two task slots per warp.  (Measured and dropped in round 2: FT = 8 with four slots per
    // warp and conflict-free strides, so that two CTAs of the N = 1920 instance fit an SM -- 2.04 ms against 0.90 ms per
    // 256 x 10 s, and 0.51 against 0.44 ms for N = 800: the window / twiddle / mel-weight reads are indexed constant loads
    // with one address per slot, four slots replay them four times, and the kernel is MIO-bound to begin with --
    // stall mio_throttle 3.0 per issue, profiles/r02_stft3_ncu_summary.txt.)
    static constexpr int FT = 16;
    static constexpr int NW = 16, NT = NW * 32;
    static constexpr int RT = (FT - 1) * P::HOP + P::N;  // samples a tile reads
    static constexpr int SW = RT + RT / P::HOP + 2;
    static constexpr int FS = NK1 * M + 1;               // float2 per frame (odd)
    static constexpr int PS = 17;                        // P row stride (slots of a warp differ by 16 banks)
    static constexpr int NMEL = 80;
    static constexpr int R0 = (SW > P::NBIN * PS) ? SW : P::NBIN * PS;
    static constexpr int OFF_A = (R0 + 3) / 4 * 4;       // A holds float2: keep it 16-byte aligned
    static_assert(OFF_A % 4 == 0, "A must be aligned for float2 access");
    // A/B build only (-DFE_S3_SMEM_TABLES): window and stage-A / stage-B twiddles copied to shared memory by the CTA.  Measured
    // 1.006 against 0.898 ms for N = 1920 and 0.428 against 0.440 ms for N = 800 (profiles/r02_stft3_ncu_summary.txt): the
    // constant bank stays.
    static constexpr int OFF_TAB = OFF_A + FT * FS * 2;
    static constexpr int TAB_WIN = 0, TAB_TWA = P::N, TAB_TWB = P::N + M * 18, TAB_FLOATS = P::N + M * 18 + P::R3 * 2 * P::R2;
#ifdef FE_S3_SMEM_TABLES
    static constexpr int SMEM_FLOATS = OFF_TAB + TAB_FLOATS;
#else
    static constexpr int SMEM_FLOATS = OFF_A + FT * FS * 2;
#endif
    static constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
    static constexpr int TASKS_A = 16 * ((M + 31) / 32); // warp-tasks: slots (m, m + 16), m = 32*j + i
    static constexpr int TASKS_B = (NK1 * P::R3 + 1) / 2;
    static constexpr int TASKS_C = (NK1 * P::R2 + 1) / 2;
    static_assert(P::HOP == 4 * M && P::N == R1 * M, "staging layout of phase A");
};

struct FeCosyTraits {
    static constexpr int N = 1920, HOP = 480, PADL = 720, R2 = 12, R3 = 10, NBIN = 640;
    static constexpr bool PREEMPH = false;
    static FE_HD FE_INLINE const float* win() { return fe_win_cosy; }
    static FE_HD FE_INLINE const float* tw_a() { return fe_tw_cosy_a; }
    static FE_HD FE_INLINE const float* tw_b() { return fe_tw_cosy_b; }
    static FE_HD FE_INLINE const int* mel_desc() { return fe_melroll_cosy_desc; }
    static FE_HD FE_INLINE const float* mel_w() { return fe_melroll_cosy_w; }
    static FE_HD FE_INLINE void dft_b(const float* xr, const float* xi, float* yr, float* yi) { fe_cdft12(xr, xi, yr, yi); }
    static FE_HD FE_INLINE float mag(float s) { return FE_SQRT(s + 1e-9f); }
    static FE_HD FE_INLINE float epilogue(float v) { return FE_LOG2(fmaxf(v, 1e-5f)) * 0.69314718055994531f; }   // ln
    // matcha: reflect-pad (n_fft - hop) / 2 each side, stft(center=False)
    static FE_HD FE_INLINE int64_t frames(int64_t len, int) { return (len + 2 * PADL < N) ? 0 : 1 + (len + 2 * PADL - N) / HOP; }
};
struct FeW2LTraits {
    static constexpr int N = 800, HOP = 200, PADL = 400, R2 = 5, R3 = 10, NBIN = 401;
    static constexpr bool PREEMPH = true;
    static FE_HD FE_INLINE const float* win() { return fe_win_w2l; }
    static FE_HD FE_INLINE const float* tw_a() { return fe_tw_w2l_a; }
    static FE_HD FE_INLINE const float* tw_b() { return fe_tw_w2l_b; }
    static FE_HD FE_INLINE const int* mel_desc() { return fe_melroll_w2l_desc; }
    static FE_HD FE_INLINE const float* mel_w() { return fe_melroll_w2l_w; }
    static FE_HD FE_INLINE void dft_b(const float* xr, const float* xi, float* yr, float* yi) { fe_cdft5(xr, xi, yr, yi); }
    static FE_HD FE_INLINE float mag(float s) { return FE_SQRT(s); }                       // audio.py:47 np.abs(D)
    // audio.py:105-118 with hparams: clip(8 * ((20 log10(max(1e-5, v)) - 20 + 100) / 100) - 4, -4, 4) = clip(1.6 log10(.) + 2.4)
    static FE_HD FE_INLINE float epilogue(float v) {
        const float y = fmaf(FE_LOG2(fmaxf(v, 1e-5f)), 0.48164799306236994f, 2.4f);       // 1.6 * log10(2)
        return fminf(fmaxf(y, -4.0f), 4.0f);
    }
    // librosa.stft(center=True): 1 + len // hop; reflect padding (pad_mode 0) needs more than n_fft / 2 samples
    static FE_HD FE_INLINE int64_t frames(int64_t len, int pad_mode) {
        return (pad_mode == 0 ? len > PADL : len > 0) ? 1 + len / HOP : 0;
    }
};
typedef FeS3Cfg<FeCosyTraits> FeCCfg;
typedef FeS3Cfg<FeW2LTraits> FeLCfg;

#ifdef FE_S3_SMEM_TABLES
#define FE_S3_WIN(C, sm) ((sm) + C::OFF_TAB + C::TAB_WIN)
#define FE_S3_TWA(C, sm) ((sm) + C::OFF_TAB + C::TAB_TWA)
#define FE_S3_TWB(C, sm) ((sm) + C::OFF_TAB + C::TAB_TWB)
#else
#define FE_S3_WIN(C, sm) C::win()
#define FE_S3_TWA(C, sm) C::tw_a()
#define FE_S3_TWB(C, sm) C::tw_b()
#endif
template <class C>
FE_HD FE_INLINE void fe_c_copy_tables(float* sm, int tid) {
#ifdef FE_S3_SMEM_TABLES
    float* t = sm + C::OFF_TAB;
    for (int i = tid; i < C::N; i += C::NT) t[C::TAB_WIN + i] = C::win()[i];
    for (int i = tid; i < C::M * 18; i += C::NT) t[C::TAB_TWA + i] = C::tw_a()[i];
    for (int i = tid; i < C::R3 * 2 * C::R2; i += C::NT) t[C::TAB_TWB + i] = C::tw_b()[i];
#else
    (void)sm; (void)tid;
#endif
}

struct FeCTile {
    int b, t0, T;
    int64_t len;
};

template <class C>
FE_HD FE_INLINE FeCTile fe_c_tile(const FeCosyParams& p, int cta) {
    FeCTile t;
    t.b = p.clip0 + cta / p.tiles_per_clip;
    t.t0 = (cta % p.tiles_per_clip) * C::FT;
    t.len = fe_clip_len(p.lengths, t.b, p.L);
    int64_t T = C::frames(t.len, p.pad_mode);
    t.T = (int)(T < p.T_out ? T : p.T_out);
    return t;
}

// ---------------------------------------------------------------- phase L
// sample g of the padded (and, for Wav2Lip, pre-emphasised: the reference pads AFTER the filter, audio.py:46) signal.
// pad_mode 0 = reflect, 1 = zeros (librosa >= 0.10's default; the reference pins 0.9.2 = reflect)
template <class C>
FE_HD FE_INLINE float fe_c_fetch(const float* __restrict__ x, int64_t g, int64_t len, int pad_mode) {
    if (!C::PREEMPH) return fe_fetch_reflect(x, g, len, len);
    if (pad_mode != 0 && (g < 0 || g >= len)) return 0.0f;
    if (g < 0) g = -g;
    if (g >= len) g = 2 * (len - 1) - g;
    if (g < 0) g = 0;
    if (g >= len) return 0.0f;
    return g > 0 ? fmaf(-0.97f, x[g - 1], x[g]) : x[g];
}

template <class C>
FE_HD FE_INLINE void fe_c_phase_load(const FeCosyParams& p, float* sm, int tid, int cta) {
    const FeCTile t = fe_c_tile<C>(p, cta);
    const float* __restrict__ x = p.wave + (int64_t)t.b * p.wave_stride;
    const int64_t g0 = (int64_t)t.t0 * C::HOP - C::PADL;
    for (int r4 = tid * 4; r4 < C::RT; r4 += C::NT * 4) {
        const int64_t g = g0 + r4;
        float v0, v1, v2, v3;
        if ((p.vec_ok & 1) && g >= 0 && g + 3 < t.len) {
            const float4 q = *reinterpret_cast<const float4*>(x + g);
            if (C::PREEMPH) {       // y[n] = x[n] - 0.97 x[n - 1], y[0] = x[0] (scipy.signal.lfilter([1, -k], [1], wav), audio.py:19-22)
                const float prev = g > 0 ? x[g - 1] : 0.0f;
                v0 = fmaf(-0.97f, prev, q.x); v1 = fmaf(-0.97f, q.x, q.y); v2 = fmaf(-0.97f, q.y, q.z); v3 = fmaf(-0.97f, q.z, q.w);
            } else {
                v0 = q.x; v1 = q.y; v2 = q.z; v3 = q.w;
            }
        } else {
            v0 = fe_c_fetch<C>(x, g + 0, t.len, p.pad_mode);
            v1 = fe_c_fetch<C>(x, g + 1, t.len, p.pad_mode);
            v2 = fe_c_fetch<C>(x, g + 2, t.len, p.pad_mode);
            v3 = fe_c_fetch<C>(x, g + 3, t.len, p.pad_mode);
        }
        float* d = sm + r4 + r4 / C::HOP;
        d[0] = v0; d[1] = v1; d[2] = v2; d[3] = v3;
    }
}

// ---------------------------------------------------------------- phase A
template <class C>
FE_HD FE_INLINE void fe_c_stage_a(float* sm, int f, int m) {
    const float* __restrict__ s = sm + (C::HOP + 1) * f + m;
    float x[C::R1];
#pragma unroll
    for (int n1 = 0; n1 < C::R1; ++n1) {
        // sample n = M*n1 + m; (M*n1 + m) / HOP == n1 / 4 for every m < M (HOP == 4 M)
        x[n1] = s[C::M * n1 + n1 / (C::HOP / C::M)] * FE_S3_WIN(C, sm)[C::M * n1 + m];
    }
    float yr[9], yi[9];
    fe_rdft16_h9(x, yr, yi);
    float2* __restrict__ a = reinterpret_cast<float2*>(sm + C::OFF_A) + C::FS * f + m;
    const float* __restrict__ tw = FE_S3_TWA(C, sm) + m * 18;
    a[0] = make_float2(yr[0], 0.0f);
#pragma unroll
    for (int k1 = 1; k1 < 8; ++k1) {
        const float wr = tw[2 * k1], wi = tw[2 * k1 + 1];
        a[C::M * k1] = make_float2(yr[k1] * wr - yi[k1] * wi, yr[k1] * wi + yi[k1] * wr);
    }
    a[C::M * 8] = make_float2(yr[8] * tw[16], yr[8] * tw[17]);
}

template <class C>
FE_HD FE_INLINE void fe_c_phase_a(const FeCosyParams&, float* sm, int tid, int) {
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    for (int u = warp; u < C::TASKS_A; u += C::NW) {
        const int m = 32 * (u >> 4) + (u & 15) + 16 * slot;  // slots 16 apart: disjoint banks
        if (m < C::M) fe_c_stage_a<C>(sm, f, m);
    }
}

// ---------------------------------------------------------------- phase B
template <class C>
FE_HD FE_INLINE void fe_c_stage_b(float* sm, int f, int k1, int n3) {
    float2* __restrict__ a = reinterpret_cast<float2*>(sm + C::OFF_A) + C::FS * f + C::M * k1 + n3;
    float xr[C::R2], xi[C::R2], yr[C::R2], yi[C::R2];
#pragma unroll
    for (int n2 = 0; n2 < C::R2; ++n2) {
        const float2 v = a[C::R3 * n2];
        xr[n2] = v.x; xi[n2] = v.y;
    }
    C::dft_b(xr, xi, yr, yi);
    const float* __restrict__ tw = FE_S3_TWB(C, sm) + n3 * (2 * C::R2);
    a[0] = make_float2(yr[0], yi[0]);
#pragma unroll
    for (int k2 = 1; k2 < C::R2; ++k2) {
        const float wr = tw[2 * k2], wi = tw[2 * k2 + 1];
        a[C::R3 * k2] = make_float2(yr[k2] * wr - yi[k2] * wi, yr[k2] * wi + yi[k2] * wr);
    }
}

template <class C>
FE_HD FE_INLINE void fe_c_phase_b(const FeCosyParams&, float* sm, int tid, int) {
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    for (int u = warp; u < C::TASKS_B; u += C::NW) {
        const int q = 2 * u + slot;
        if (q < C::NK1 * C::R3) fe_c_stage_b<C>(sm, f, q / C::R3, q % C::R3);
    }
}

// ---------------------------------------------------------------- phase C
template <class C>
FE_HD FE_INLINE void fe_c_stage_c(float* sm, int f, int k1, int k2) {
    const float2* __restrict__ a = reinterpret_cast<const float2*>(sm + C::OFF_A) + C::FS * f + C::M * k1 + C::R3 * k2;
    float xr[10], xi[10], yr[10], yi[10];
#pragma unroll
    for (int n3 = 0; n3 < 10; ++n3) {
        const float2 v = a[n3];
        xr[n3] = v.x; xi[n3] = v.y;
    }
    fe_cdft10(xr, xi, yr, yi);
    const int kb = k1 + C::R1 * k2;                 // bin = kb + 192*k3
    const bool mirror = (k1 != 0) && (k1 != C::R1 / 2);
    float* __restrict__ P = sm + f;
#pragma unroll
    for (int k3 = 0; k3 < 10; ++k3) {
        const int bin = kb + C::R1 * C::R2 * k3;
        const float mag = C::mag(yr[k3] * yr[k3] + yi[k3] * yi[k3]);
        if (bin < C::NBIN) P[bin * C::PS] = mag;
        else if (mirror && (C::N - bin) < C::NBIN) P[(C::N - bin) * C::PS] = mag;
    }
}

template <class C>
FE_HD FE_INLINE void fe_c_phase_c(const FeCosyParams&, float* sm, int tid, int) {
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    for (int u = warp; u < C::TASKS_C; u += C::NW) {
        const int q = 2 * u + slot;
        if (q < C::NK1 * C::R2) fe_c_stage_c<C>(sm, f, q / C::R2, q % C::R2);
    }
}

// ---------------------------------------------------------------- phase M
// Both half-warps work: lane & 15 = frame, lane >> 4 = which of two interleaved filter sequences (filter 2 * warp + slot,
// then every 32nd).  (Until round 2 the upper half-warp left this phase at once.)
template <class C>
FE_HD FE_INLINE void fe_c_phase_mel(const FeCosyParams& p, float* sm, int tid, int cta) {
    const int lane = tid & 31, warp = tid >> 5;
    const int fl = lane & (C::FT - 1), slot = lane >> 4;
    const FeCTile t = fe_c_tile<C>(p, cta);
    const int frame = t.t0 + fl;
    const bool live = frame < t.T;
    // mel-major: out[b][m][frame]; frame-major: out[b][frame][m] (16 frames x 320 B apart per store -- the output is
    // 14 % of this kernel's bytes and the kernel is nowhere near HBM-bound)
    float* __restrict__ o = p.out + (int64_t)t.b * p.out_batch_stride + (p.frame_major ? (int64_t)frame * C::NMEL : (int64_t)frame);
    const int64_t rs = p.frame_major ? 1 : p.T_out;
    auto emit = [&](int m, float v) {
        const float y = C::epilogue(v);
        if (frame < p.T_out) o[(int64_t)m * rs] = live ? y : 0.0f;  // ragged batches: zeros past T_b
    };
    fe_mel_rolled(C::mel_desc(), C::mel_w(), 2 * warp + slot, C::NMEL, 2 * C::NW, [&](int k) { return sm[k * C::PS + fl]; }, emit);
}
