// Spectral-gate denoiser of the reference's AudioProcessor (second half).
//
// Replaces /root/reference/Backend/services/audio_processor.py:322-358, per clip:
//     spec  = torch.stft(audio, n_fft=1024, hop_length=256, win_length=1024, window=hann_window(1024), return_complex=True)
//     mag, phase = |spec|, angle(spec)                                                     (513 bins, T = 1 + L // 256 frames)
//     noise_floor = torch.quantile(mag, 0.1, dim=1, keepdim=True)                         per frequency bin, over time
//     gain_mask   = clamp((mag - 1.5 * noise_floor) / (mag + 1e-8), 0, 1)
//     freq_bins = mag.shape[1]; gain_mask[:, int(freq_bins * 0.0625):int(freq_bins * 0.25)] *= 1.2
//         -- mag is (513, T), so shape[1] is the number of FRAMES: the reference boosts a range of frames, not of
//            frequencies.  "Results identical to the reference's" means the same here: frames [int(T/16), int(T/4)).
//     audio = torch.istft(mag * gain_mask * exp(1j * phase), 1024, 256, 1024, hann_window(1024))      length 256 * (T - 1)
//     audio = audio / audio.abs().max() * 0.95                                             (if the maximum is > 0)
//
// Four kernels:
//   1  fe_gate_stft_kernel      the three-stage STFT of fe_cosy.cuh with N = 1024 = 16 x 8 x 8, hop 256, reflect pad 512;
//                               its last stage stores the complex spectrum spec[b][bin][t] instead of a mel projection
//   2  fe_gate_quantile_kernel  one warp per (clip, bin): 4-pass radix select of the two order statistics torch.quantile
//                               interpolates between
//   3  fe_gate_inverse_kernel   per tile of 16 frames: gain x spectrum, complex FFT 16 x 8 x 8 of the conjugate (= N times the
//                               inverse, real part), window, overlap-add of the 4 frames over each hop, division by the window
//                               envelope, store; tiles overlap by 3 frames so that every output hop is summed inside ONE CTA in
//                               a fixed order (deterministic, no atomics on samples); per-clip peak by atomicMax
//   4  fe_gate_scale_kernel     x 0.95 / peak
#pragma once
#include "fe_common.h"
#include "fe_cosy.cuh"

struct FeGateParams {
    const float* wave;          // (B, wave_stride) float32
    const int64_t* lengths;
    int64_t wave_stride, L;
    int32_t B;
    float2* spec;               // (B, 513, Ts) complex
    int64_t Ts;                 // 1 + L / 256
    float* nf;                  // (B, 513) noise floor
    int32_t* peak;              // (B) float bits of max |out|
    float* out;                 // (B, out_stride), L_out = 256 * (L / 256) samples per row
    int64_t out_stride, L_out;
    int32_t tiles_per_clip;     // inverse kernel: ceil(hops / 13)
};

struct FeGateTraits {
    static constexpr int N = 1024, HOP = 256, PADL = 512, R2 = 8, R3 = 8, NBIN = 513;
    static constexpr bool PREEMPH = false;
    static FE_HD FE_INLINE const float* win() { return fe_win_gate; }
    static FE_HD FE_INLINE const float* tw_a() { return fe_tw_gate_a; }
    static FE_HD FE_INLINE const float* tw_b() { return fe_tw_gate_b; }
    static FE_HD FE_INLINE const int* mel_desc() { return nullptr; }
    static FE_HD FE_INLINE const float* mel_w() { return nullptr; }
    static FE_HD FE_INLINE void dft_b(const float* xr, const float* xi, float* yr, float* yi) { fe_cdft8(xr, xi, yr, yi); }
    static FE_HD FE_INLINE float mag(float s) { return FE_SQRT(s); }
    static FE_HD FE_INLINE float epilogue(float v) { return v; }
    // torch.stft(center=True, pad_mode="reflect"): 1 + len // hop frames; the reflect padding needs more than 512 samples
    static FE_HD FE_INLINE int64_t frames(int64_t len, int) { return len > PADL ? 1 + len / HOP : 0; }
};
typedef FeS3Cfg<FeGateTraits> FeGCfg;

// ---------------------------------------------------------------- forward: last stage, spectrum to global memory
FE_HD FE_INLINE void fe_g_stage_c_spec(const FeGateParams& g, const FeCTile& t, float* sm, int f, int k1, int k2) {
    typedef FeGCfg C;
    const float2* __restrict__ a = reinterpret_cast<const float2*>(sm + C::OFF_A) + C::FS * f + C::M * k1 + C::R3 * k2;
    float xr[8], xi[8], yr[8], yi[8];
#pragma unroll
    for (int n3 = 0; n3 < 8; ++n3) {
        const float2 v = a[n3];
        xr[n3] = v.x; xi[n3] = v.y;
    }
    fe_cdft8(xr, xi, yr, yi);
    const int frame = t.t0 + f;
    if (frame >= t.T) return;
    const int kb = k1 + C::R1 * k2;                 // bin = kb + 128 * k3
    const bool mirror = (k1 != 0) && (k1 != C::R1 / 2);
    float2* __restrict__ s = g.spec + (int64_t)t.b * C::NBIN * g.Ts + frame;
#pragma unroll
    for (int k3 = 0; k3 < 8; ++k3) {
        const int bin = kb + C::R1 * C::R2 * k3;
        float2 v;
        if (bin < C::NBIN) {
            v.x = yr[k3]; v.y = yi[k3];
            s[(int64_t)bin * g.Ts] = v;
        } else if (mirror && (C::N - bin) < C::NBIN) {
            v.x = yr[k3]; v.y = -yi[k3];             // X[N - k] = conj(X[k]) for a real signal
            s[(int64_t)(C::N - bin) * g.Ts] = v;
        }
    }
}
FE_HD FE_INLINE void fe_g_phase_c_spec(const FeGateParams& g, const FeCosyParams& p, float* sm, int tid, int cta) {
    typedef FeGCfg C;
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    const FeCTile t = fe_c_tile<C>(p, cta);
    for (int u = warp; u < C::TASKS_C; u += C::NW) {
        const int q = 2 * u + slot;
        if (q < C::NK1 * C::R2) fe_g_stage_c_spec(g, t, sm, f, q / C::R2, q % C::R2);
    }
}

// ---------------------------------------------------------------- inverse
struct FeGICfg {
    static constexpr int N = 1024, HOP = 256, FT = 16, HALO = 3, OUT_HOPS = FT - HALO;    // 13 output hops per tile
    static constexpr int NW = 16, NT = NW * 32;
    static constexpr int FS = N + 1;                 // float2 per frame in A (odd: frames on different bank pairs)
    static constexpr int XS = N + 1;                 // floats per frame in X
    static constexpr int OFF_X = FT * FS * 2;
    static constexpr int SMEM_FLOATS = OFF_X + FT * XS;
    static constexpr int SMEM_BYTES = SMEM_FLOATS * 4;
};
struct FeGITile { int b, t0, T, b0, b1; int64_t len; };    // frames t0 .. t0 + 15 (t0 may be < 0); boost frames [b0, b1)
FE_HD FE_INLINE FeGITile fe_gi_tile(const FeGateParams& g, int cta) {
    FeGITile t;
    t.b = cta / g.tiles_per_clip;
    const int i = cta - t.b * g.tiles_per_clip;
    t.t0 = FeGICfg::OUT_HOPS * i - FeGICfg::HALO;
    t.len = fe_clip_len(g.lengths, t.b, g.L);
    t.T = (int)FeGateTraits::frames(t.len, 0);
    t.b0 = (int)((float)t.T * 0.0625f);              // int(freq_bins * 0.0625) with freq_bins = mag.shape[1] = T (see the header)
    t.b1 = (int)((float)t.T * 0.25f);
    return t;
}

// A[f][n] = conj(gain * spec[n]) for n <= 512, gain * spec[1024 - n] for n > 512 (the conjugate of the Hermitian extension)
FE_HD FE_INLINE void fe_gi_phase_load(const FeGateParams& g, float* sm, int tid, int cta) {
    typedef FeGICfg C;
    const FeGITile t = fe_gi_tile(g, cta);
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    const int frame = t.t0 + f;
    const bool live = frame >= 0 && frame < t.T;
    const float boost = (frame >= t.b0 && frame < t.b1) ? 1.2f : 1.0f;
    float2* __restrict__ a = reinterpret_cast<float2*>(sm) + C::FS * f;
    const float2* __restrict__ s = g.spec + (int64_t)t.b * 513 * g.Ts + frame;
    const float* __restrict__ nf = g.nf + (int64_t)t.b * 513;
    for (int bin = 2 * warp + slot; bin <= 512; bin += 2 * C::NW) {
        float2 v;
        v.x = 0.0f; v.y = 0.0f;
        if (live) {
            const float2 x = s[(int64_t)bin * g.Ts];
            const float mag = FE_SQRT(x.x * x.x + x.y * x.y);
            float gain = (mag - 1.5f * nf[bin]) / (mag + 1e-8f);
            gain = fminf(fmaxf(gain, 0.0f), 1.0f) * boost;
            v.x = x.x * gain; v.y = x.y * gain;
        }
        float2 c;
        c.x = v.x; c.y = -v.y;
        a[bin] = c;
        if (bin > 0 && bin < 512) a[C::N - bin] = v;
    }
}
// stage 1: complex radix-16 over n1 (stride 64), times W_1024^(m k1), in place
FE_HD FE_INLINE void fe_gi_phase_1(float* sm, int tid) {
    typedef FeGICfg C;
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    for (int u = warp; u < 32; u += C::NW) {
        const int m = 32 * (u >> 4) + (u & 15) + 16 * slot;
        float2* __restrict__ a = reinterpret_cast<float2*>(sm) + C::FS * f + m;
        float xr[16], xi[16], yr[16], yi[16];
#pragma unroll
        for (int n1 = 0; n1 < 16; ++n1) { const float2 v = a[64 * n1]; xr[n1] = v.x; xi[n1] = v.y; }
        fe_cdft16(xr, xi, yr, yi);
        const float* __restrict__ tw = fe_tw_gate_a16 + m * 32;
        float2 o;
        o.x = yr[0]; o.y = yi[0];
        a[0] = o;
#pragma unroll
        for (int k1 = 1; k1 < 16; ++k1) {
            const float wr = tw[2 * k1], wi = tw[2 * k1 + 1];
            o.x = yr[k1] * wr - yi[k1] * wi; o.y = yr[k1] * wi + yi[k1] * wr;
            a[64 * k1] = o;
        }
    }
}
// stage 2: complex radix-8 over n2 for each (k1, n3), times W_64^(n3 k2), in place
FE_HD FE_INLINE void fe_gi_phase_2(float* sm, int tid) {
    typedef FeGICfg C;
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    for (int u = warp; u < 64; u += C::NW) {
        const int q = 2 * u + slot, k1 = q >> 3, n3 = q & 7;
        float2* __restrict__ a = reinterpret_cast<float2*>(sm) + C::FS * f + 64 * k1 + n3;
        float xr[8], xi[8], yr[8], yi[8];
#pragma unroll
        for (int n2 = 0; n2 < 8; ++n2) { const float2 v = a[8 * n2]; xr[n2] = v.x; xi[n2] = v.y; }
        fe_cdft8(xr, xi, yr, yi);
        const float* __restrict__ tw = fe_tw_gate_b + n3 * 16;
        float2 o;
        o.x = yr[0]; o.y = yi[0];
        a[0] = o;
#pragma unroll
        for (int k2 = 1; k2 < 8; ++k2) {
            const float wr = tw[2 * k2], wi = tw[2 * k2 + 1];
            o.x = yr[k2] * wr - yi[k2] * wi; o.y = yr[k2] * wi + yi[k2] * wr;
            a[8 * k2] = o;
        }
    }
}
// stage 3: complex radix-8 over n3 for each (k1, k2); sample k = k1 + 16 k2 + 128 k3 = Re(.) / 1024, times the window -> X[f][k]
FE_HD FE_INLINE void fe_gi_phase_3(float* sm, int tid) {
    typedef FeGICfg C;
    const int lane = tid & 31, warp = tid >> 5;
    const int f = lane & 15, slot = lane >> 4;
    for (int u = warp; u < 64; u += C::NW) {
        const int q = 2 * u + slot, k1 = q >> 3, k2 = q & 7;
        const float2* __restrict__ a = reinterpret_cast<const float2*>(sm) + C::FS * f + 64 * k1 + 8 * k2;
        float xr[8], xi[8], yr[8], yi[8];
#pragma unroll
        for (int n3 = 0; n3 < 8; ++n3) { const float2 v = a[n3]; xr[n3] = v.x; xi[n3] = v.y; }
        fe_cdft8(xr, xi, yr, yi);
        float* __restrict__ X = sm + C::OFF_X + C::XS * f;
#pragma unroll
        for (int k3 = 0; k3 < 8; ++k3) {
            const int k = k1 + 16 * k2 + 128 * k3;
            X[k] = yr[k3] * (1.0f / 1024.0f) * fe_win_gate[k];
        }
    }
}
// overlap-add of the tile's 13 output hops, division by the window envelope, store; returns this thread's max |out|
FE_HD FE_INLINE float fe_gi_phase_ola(const FeGateParams& g, const float* sm, int tid, int cta) {
    typedef FeGICfg C;
    const FeGITile t = fe_gi_tile(g, cta);
    const int64_t lout = t.T > 0 ? (int64_t)C::HOP * (t.T - 1) : 0;       // torch.istft: hop * (T - 1) samples
    float* __restrict__ o = g.out + (int64_t)t.b * g.out_stride;
    float mx = 0.0f;
    for (int idx = tid; idx < C::OUT_HOPS * C::HOP; idx += C::NT) {
        const int hl = idx / C::HOP, j = idx - hl * C::HOP;
        const int fl = hl + C::HALO;                  // local frame whose first quarter lands on this hop
        const int64_t p = ((int64_t)t.t0 + fl) * C::HOP + j;               // position in the padded signal
        const int64_t n = p - C::N / 2;
        if (n < 0 || n >= g.L_out) continue;
        float acc = 0.0f, env = 0.0f;
#pragma unroll
        for (int d = 3; d >= 0; --d) {                // frames h - 3 .. h, oldest first (torch's fold order is not observable)
            const int frame = t.t0 + fl - d;
            if (frame >= 0 && frame < t.T) {
                const float w = fe_win_gate[C::HOP * d + j];
                acc += sm[C::OFF_X + C::XS * (fl - d) + C::HOP * d + j];
                env += w * w;
            }
        }
        const float v = n < lout ? acc / env : 0.0f;
        o[n] = v;
        mx = fmaxf(mx, fabsf(v));
    }
    return mx;
}
