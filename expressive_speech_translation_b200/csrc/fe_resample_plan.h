// Host-side plan of the resampling kernel (fe_resample.cuh): the polyphase tap table of
// torchaudio.functional.resample (site-packages/torchaudio/functional/functional.py `_get_sinc_resample_kernel`:
// float64 arithmetic, windowed sinc, hann or kaiser window, result rounded to float32), cut down to the window's
// support and packed as groups of four phases.  Host C++ only; shared by libb200feat.so and the CPU emulation.
#pragma once
#include <cmath>
#include <cstdint>
#include <numeric>
#include <string>
#include <vector>

struct FeRsPlan {
    int orig = 0, newf = 0;        // reduced by the gcd
    int m = 1, orig_e = 0, new_e = 0;
    int width = 0;
    int G = 0, J = 0;
    int P = 0, OP = 0, nx = 0, xs_floats = 0;
    bool taps_smem = false;
    size_t smem_bytes = 0;
    std::vector<float> taps;       // [G][J][4]
    std::vector<int32_t> kstart;   // [G]
    std::string error;             // non-empty: the plan could not be built
};

enum { FE_RS_HANN = 0, FE_RS_KAISER = 1 };

// one tap of phase p (0 <= p < newf), k in [0, 2*width + orig); 0 outside the window's support
static inline double fe_rs_tap(int orig, int newf, int width, int lpfw, double rolloff, int method, double beta,
                               int p, int k) {
    const double base = std::min(orig, newf) * rolloff;
    // functional.py:1377 (dtype=None, as transforms.Resample calls it): the phase offset -p/new is a float32 quotient
    // (int64 tensor / int), the sample grid is float64; the deployed taps carry that rounding, so does this table
    const double phase = (double)((float)(-p) / (float)newf);
    double t = (phase + (double)(k - width) / orig) * base;
    if (t <= -(double)lpfw || t >= (double)lpfw) return 0.0;    // the reference clamps t: window (hann) or sinc (kaiser) is 0 there
    const double pi = 3.14159265358979323846;
    double window;
    if (method == FE_RS_HANN) {
        const double c = std::cos(t * pi / lpfw / 2.0);
        window = c * c;
    } else {
        const double r = t / lpfw;
        window = std::cyl_bessel_i(0.0, beta * std::sqrt(1.0 - r * r)) / std::cyl_bessel_i(0.0, beta);
    }
    const double scale = base / orig;
    const double tp = t * pi;
    const double s = (tp == 0.0) ? 1.0 : std::sin(tp) / tp;
    return s * window * scale;
}

static inline FeRsPlan fe_rs_make_plan(int orig_freq, int new_freq, int lpfw, double rolloff, int method, double beta,
                                       size_t smem_two_ctas = 113 * 1024, size_t smem_one_cta = 226 * 1024) {
    FeRsPlan pl;
    const int QB = 32, R = 4;
    if (orig_freq <= 0 || new_freq <= 0) { pl.error = "orig_freq and new_freq must be positive"; return pl; }
    if (lpfw <= 0) { pl.error = "lowpass_filter_width must be positive"; return pl; }
    if (!(rolloff > 0.0 && rolloff <= 1.0)) { pl.error = "rolloff must be in (0, 1]"; return pl; }
    if (method != FE_RS_HANN && method != FE_RS_KAISER) { pl.error = "unknown resampling method"; return pl; }
    const int g = std::gcd(orig_freq, new_freq);
    pl.orig = orig_freq / g;
    pl.newf = new_freq / g;
    const double base = std::min(pl.orig, pl.newf) * rolloff;
    pl.width = (int)std::ceil((double)lpfw * pl.orig / base);
    const int K = 2 * pl.width + pl.orig;

    // the float32 tap table, as the reference builds it, and the support of every phase
    std::vector<float> h((size_t)pl.newf * K);
    std::vector<int> lo(pl.newf), hi(pl.newf);
    for (int p = 0; p < pl.newf; ++p) {
        int l = K, u = -1;
        for (int k = 0; k < K; ++k) {
            const float v = (float)fe_rs_tap(pl.orig, pl.newf, pl.width, lpfw, rolloff, method, beta, p, k);
            h[(size_t)p * K + k] = v;
            if (v != 0.0f) { if (k < l) l = k; u = k; }
        }
        if (u < 0) { l = 0; u = 0; }
        lo[p] = l; hi[p] = u;
    }

    // blocks run m at a time so that a 32-block tile is a useful amount of work and still fits shared memory
    // (an odd orig stays odd under an odd m: no pitch padding, and the tap loop takes its two-stream form)
    int m = std::max(1, (128 + pl.newf - 1) / pl.newf);
    for (;; m = (m + 1) / 2) {
        if ((pl.orig & 1) && !(m & 1)) m += (m == 2) ? -1 : 1;
        pl.m = m;
        pl.orig_e = pl.orig * m;
        pl.new_e = pl.newf * m;
        const int Ke = pl.orig_e + 2 * pl.width;
        pl.G = (pl.new_e + R - 1) / R;
        // union support of each group of four effective phases pe = c * newf + p (tap index k' = c * orig + k)
        std::vector<int> glo(pl.G), ghi(pl.G);
        int J = 1;
        for (int gi = 0; gi < pl.G; ++gi) {
            int l = Ke, u = -1;
            for (int r = 0; r < R; ++r) {
                const int pe = gi * R + r;
                if (pe >= pl.new_e) break;
                const int c = pe / pl.newf, p = pe % pl.newf;
                l = std::min(l, c * pl.orig + lo[p]);
                u = std::max(u, c * pl.orig + hi[p]);
            }
            glo[gi] = l; ghi[gi] = u;
            J = std::max(J, u - l + 1);
        }
        J = (J + 3) & ~3;                       // the tap loop is unrolled by four
        if (J > Ke) J = Ke;
        pl.J = J;
        pl.P = pl.orig_e | 1;
        pl.OP = pl.new_e | 1;
        pl.nx = QB * pl.orig_e + 2 * pl.width;
        const int rows = (pl.nx + pl.orig_e - 1) / pl.orig_e;
        pl.xs_floats = rows * pl.P;
        const size_t fixed = ((size_t)pl.xs_floats + (size_t)QB * pl.OP) * 4;
        const size_t tbytes = (size_t)pl.G * J * 16;
        if (fixed <= smem_two_ctas || (m == 1 && fixed <= smem_one_cta)) {   // two CTAs per SM unless one block row is that long
            pl.taps_smem = fixed + tbytes <= smem_two_ctas;
            pl.smem_bytes = fixed + (pl.taps_smem ? tbytes : 0);
            pl.kstart.resize(pl.G);
            pl.taps.assign((size_t)pl.G * J * 4, 0.0f);
            for (int gi = 0; gi < pl.G; ++gi) {
                const int k0 = std::max(0, std::min(glo[gi], Ke - J));   // window stays inside the tile
                pl.kstart[gi] = k0;
                for (int r = 0; r < R; ++r) {
                    const int pe = gi * R + r;
                    if (pe >= pl.new_e) break;
                    const int c = pe / pl.newf, p = pe % pl.newf;
                    for (int j = 0; j < J; ++j) {
                        const int k = k0 + j - c * pl.orig;
                        if (k >= 0 && k < K) pl.taps[((size_t)gi * J + j) * 4 + r] = h[(size_t)p * K + k];
                    }
                }
            }
            return pl;
        }
        if (m == 1) break;
    }
    pl.error = "resampling ratio " + std::to_string(pl.orig) + " -> " + std::to_string(pl.newf) +
               " needs more shared memory per tile than the SM has";
    return pl;
}

// output samples of a clip of n input frames: ceil(new * n / orig) (functional.py `_apply_sinc_resample_kernel`)
static inline int64_t fe_rs_out_len(int64_t n, int orig_freq, int new_freq) {
    if (n <= 0 || orig_freq <= 0 || new_freq <= 0) return 0;
    const int g = std::gcd(orig_freq, new_freq);
    const int64_t o = orig_freq / g, w = new_freq / g;
    return (n * w + o - 1) / o;
}

#ifdef FE_RS_HAVE_PARAMS
// everything of FeResampleParams that depends on the plan and the call's sizes (table pointers are the caller's)
static inline void fe_rs_fill_params(FeResampleParams& p, const FeRsPlan& pl, int64_t L_in, int64_t L_out,
                                     int channels, bool in_s16, const float* mix) {
    p.orig = pl.orig; p.newf = pl.newf; p.orig_e = pl.orig_e; p.new_e = pl.new_e; p.width = pl.width;
    p.P = pl.P; p.OP = pl.OP; p.G = pl.G; p.J = pl.J; p.nx = pl.nx; p.xs_floats = pl.xs_floats;
    p.L_in = L_in; p.L_out = L_out;
    const int64_t per_tile = (int64_t)32 * pl.new_e;
    p.tiles_per_clip = (int32_t)((L_out + per_tile - 1) / per_tile);
    float w0 = channels == 1 ? 1.0f : 0.5f, w1 = channels == 1 ? 0.0f : 0.5f;
    if (mix) { w0 = mix[0]; w1 = channels > 1 ? mix[1] : 0.0f; }
    const float s = in_s16 ? (1.0f / 32768.0f) : 1.0f;
    p.w0 = w0 * s; p.w1 = w1 * s;
    p.vec_ok = 0;   // set by the caller once it knows the row pointers
}
#endif
