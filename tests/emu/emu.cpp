// Host-side emulation of the featurizer kernels' phases (TEST INFRASTRUCTURE, never shipped).
// Compiles the very same phase functions the CUDA kernels call, with g++, and runs them thread after
// thread over a plain array standing in for shared memory.  Built by tests/emu/build.py.
#include <vector>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include "fe_common.h"
#include "generated/fe_codelets.h"
#include "generated/fe_tables.h"
#include "fe_whisper2.cuh"
#include "fe_whisper3.cuh"

// one unit at a time: loader threads, then the compute warps phase by phase (what the named barriers order on the GPU)
template <int NMELS, bool OUT16>
static void run_whisper(const FeWhisper2Params& p) {
    typedef FeW2Cfg C;
    std::vector<float> sm(C::SMEM_FLOATS);
    fe_w2_copy_tables<NMELS>(sm.data(), 0, 1);
    for (int u = 0; u < p.total_units; ++u) {
        const FeW2Unit t = fe_w2_unit(p, u);
        float mn = 3.0e38f, mx = -3.0e38f;
        if (t.t0 < t.T) {
            for (int l = 0; l < C::NLT; ++l) {   // the loader's three paths, as the kernel picks them
                FeW2Stage st;
                if (p.in_s16) {
                    const int mode = fe_w2_load_mode<true>(p, t);
                    fe_w2_load_ahead<true>(p, t, l, mode, st);
                    fe_w2_load_finish<true>(p, t, sm.data(), l, mode, st);
                } else {
                    const int mode = fe_w2_load_mode<false>(p, t);
                    fe_w2_load_ahead<false>(p, t, l, mode, st);
                    fe_w2_load_finish<false>(p, t, sm.data(), l, mode, st);
                }
            }
            for (int w = 0; w < C::NCW; ++w) for (int l = 0; l < 32; ++l) fe_w2_phase_a(sm.data(), l, w);
            for (int w = 0; w < C::NCW; ++w) for (int l = 0; l < 32; ++l) fe_w2_phase_b(sm.data(), l, w);
            for (int w = 0; w < C::NCW; ++w) for (int l = 0; l < 32; ++l) {
                const fe2 mm = fe_w2_phase_mel<NMELS, OUT16>(p, t, sm.data(), l, w);
                mn = std::min(mn, mm.x); mx = std::max(mx, mm.y);
            }
        }
        p.unit_minmax[u] = make_float2(mn, mx);
        int* slot = p.max_keys + (p.norm_mode == FE_NORM_PER_CALL ? 0 : t.b);
        *slot = std::max(*slot, fe_float_key(mx));
    }
}

static bool use_tmap_env() { const char* e = getenv("EMU_NO_TMAP"); return !(e && e[0] == '1'); }
// Tensor-Memory kernel (fe_whisper3.cuh): Tensor Memory is a plain [team][lane][512] array here; per CTA unit the teams'
// rows are filled (bulk-copy path or element-wise path, as the kernel picks them), then phases A, B and M run in the
// order the mbarriers enforce on the GPU.  unit_minmax holds (max key, -min key) like the kernel's atomics.
template <int NMELS, bool OUT16>
static void run_whisper3(const FeWhisper2Params& p) {
    typedef FeW3Cfg C;
    std::vector<float> sm(C::SMEM_FLOATS, 0.0f);
    std::vector<float> tmem((size_t)C::TEAMS * 32 * C::TMEM_COLS, 0.0f);
    fe_w3_copy_tables<NMELS>(sm.data(), 0, 1);
    float* P = sm.data() + C::OFF_P;
    for (int u = 0; u < p.total_units; ++u) {
        const FeW3Unit t = fe_w3_unit(p, u);
        float mn = 3.0e38f, mx = -3.0e38f;
        if (t.silent) {
            // the kernel steps over the unit (T == t0 in its view); the fix-up fills its live columns with the threshold.
            // Here: -inf, which the element-wise fix-up pass below turns into exactly that.
            const FeW3Unit f = fe_w3_unit<true>(p, u);
            const int n_live = std::max(0, std::min(f.Ts - f.t0, f.n_in));
            for (int m = 0; m < NMELS; ++m)
                for (int c = 0; c < n_live; ++c) {
                    const int64_t i = t.o0 + (int64_t)m * t.pitch + c;
                    if (OUT16) static_cast<unsigned short*>(p.out)[i] = 0xfc00u;
                    else static_cast<float*>(p.out)[i] = -INFINITY;
                }
        } else if (t.t0 < t.T) {
            for (int q = 0; q < C::TEAMS; ++q) {
                float* wq = sm.data() + C::OFF_W + q * C::WT;
                if (!fe_w3_team_live(t, q)) continue;
                if (p.use_tmap) {       // box copy, then the team's patch of the rows the copy cannot get right
                    fe_w3_box_host(p, t, q, wq);
                    if (!fe_w3_box_complete(p, t, q))
                        for (int member = 0; member < C::MEMBERS; ++member)
                            for (int l = 0; l < 32; ++l) fe_w3_patch_rows(p, t, q, wq, member, l);
                } else {
                    fe_w3_fill_rows_host(p, t, q, wq);
                    for (int l = 0; l < 32; ++l) fe_w3_fill_manual(p, t, q, wq, l);
                }
                for (int member = 0; member < C::MEMBERS; ++member)
                    for (int l = 0; l < 32; ++l) {
                        FeTm tm; tm.row = tmem.data() + ((size_t)q * 32 + l) * C::TMEM_COLS;
                        if (getenv("EMU_W3_PAIR")) {     // the A/B variant of the kernel (-DFE_W3_PAIR)
                            fe_w3_stage_a_pair(wq, tm, l, member, FeNoHook());
                            if (fe_w3_a_single(member) >= 0) fe_w3_stage_a(wq, tm, l, fe_w3_a_single(member), FeNoHook());
                        } else {
                            for (int k = 0; k < (fe_w3_a_single(member) >= 0 ? 3 : 2); ++k)
                                fe_w3_stage_a(wq, tm, l, k < 2 ? 2 * member + k : fe_w3_a_single(member), FeNoHook());
                        }
                    }
                for (int member = 0; member < C::MEMBERS; ++member)
                    for (int k = 0; k < fe_w3_b_count(member); ++k)
                        for (int l = 0; l < 32; ++l) {
                            FeTm tm; tm.row = tmem.data() + ((size_t)q * 32 + l) * C::TMEM_COLS;
                            fe_w3_stage_b(P, tm, l, q, fe_w3_b_task(member, k), FeNoHook(), FeNoHook());
                        }
            }
            for (int w = 0; w < C::NW; ++w) for (int l = 0; l < 32; ++l) {
                const fe2 mm = fe_w3_phase_mel<NMELS, OUT16>(p, t, sm.data(), l, w);
                mn = std::min(mn, mm.x); mx = std::max(mx, mm.y);
            }
        }
        int* uk = reinterpret_cast<int*>(p.unit_minmax + u);
        uk[0] = fe_float_key(mx); uk[1] = fe_float_key(-mn);
        int* slot = p.max_keys + (p.norm_mode == FE_NORM_PER_CALL ? 0 : t.b);
        *slot = std::max(*slot, fe_float_key(mx));
    }
}

// float32 -> binary16 -> float32 (host form of fe_f2h_rn; the inverse is exact)
static float half_bits_to_float(unsigned short h) {
    const unsigned sign = (unsigned)(h & 0x8000u) << 16, e = (h >> 10) & 31u, m = h & 1023u;
    union { unsigned u; float f; } c;
    if (e == 0) {
        c.f = (float)m * 5.9604644775390625e-08f;            // m * 2^-24
        c.u |= sign;
    } else if (e == 31) {
        c.u = sign | 0x7f800000u | (m << 13);
    } else {
        c.u = sign | ((e + 112u) << 23) | (m << 13);
    }
    return c.f;
}

extern "C" void emu_f2h(const float* in, unsigned short* out, int64_t n) {
    for (int64_t i = 0; i < n; ++i) out[i] = fe_f2h_rn(in[i]);
}

static int emu_whisper_any(const void* wave, bool in_s16, const int64_t* lengths, int B, int64_t L, int64_t stride,
                           int n_mels, int padding, int norm_mode, void* out, bool out16, int64_t T_out, int version = 2) {
    typedef FeW2Cfg C;
    std::vector<int> keys(B > 0 ? B : 1, FE_KEY_INIT);
    FeWhisper2Params p;
    memset(&p, 0, sizeof(p));
    p.wave = static_cast<const float*>(wave); p.in_s16 = in_s16 ? 1 : 0;
    p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B; p.padding = padding;
    p.n_mels = n_mels; p.norm_mode = norm_mode; p.out = out; p.out_batch_stride = (int64_t)n_mels * T_out;
    p.T_out = T_out; p.units_per_clip = (int)((T_out + C::FT - 1) / C::FT); p.total_units = B * p.units_per_clip;
    p.vec_ok = ((((uintptr_t)wave & 15) == 0) && (stride % 4 == 0)) ? 1 : 0;
    if ((((uintptr_t)out & 15) == 0) && (p.out_batch_stride % 4 == 0)) p.vec_ok |= 2;
    std::vector<float2> mm((size_t)std::max(p.total_units, 1));
    p.max_keys = keys.data(); p.unit_minmax = mm.data();
    if (version == 3) {
        if (in_s16) return -2;
        p.units_per_clip = (int)((T_out + FeW3Cfg::UF - 1) / FeW3Cfg::UF); p.total_units = B * p.units_per_clip;
        {   // the tensor map the library would build (whisper3_make_tmap in b200feat.cu)
            const int64_t first = (int64_t)FeW3Cfg::TMAP_ROW0 * 160 - 200;
            const int64_t rows = L - first - FeW3Cfg::PITCH < 0 ? 0 : (L - first - FeW3Cfg::PITCH) / 160 + 1;
            p.use_tmap = ((p.vec_ok & 1) && rows >= FeW3Cfg::ROWS && use_tmap_env()) ? 1 : 0;
            p.tmap_rows = p.use_tmap ? (int)rows : 0;
        }
        if (n_mels == 80 && !out16) run_whisper3<80, false>(p);
        else if (n_mels == 80) run_whisper3<80, true>(p);
        else if (n_mels == 128 && !out16) run_whisper3<128, false>(p);
        else if (n_mels == 128) run_whisper3<128, true>(p);
        else return -1;
    } else
    if (n_mels == 80 && !out16) run_whisper<80, false>(p);
    else if (n_mels == 80) run_whisper<80, true>(p);
    else if (n_mels == 128 && !out16) run_whisper<128, false>(p);
    else if (n_mels == 128) run_whisper<128, true>(p);
    else return -1;
    // the fix-up kernel's arithmetic, applied to every element (float16: on the rounded values, against the
    // rounded threshold, as the kernel does)
    for (int b = 0; b < B; ++b) {
        const float thr = fe_w2_threshold(fe_key_float(keys[norm_mode == FE_NORM_PER_CALL ? 0 : b]));
        const int64_t len = fe_clip_len(lengths, b, L);
        const int64_t Tb = std::min<int64_t>((len + padding) / 160, T_out);
        for (int m = 0; m < n_mels; ++m)
            for (int64_t t = 0; t < T_out; ++t) {
                const int64_t i = (int64_t)b * p.out_batch_stride + m * T_out + t;
                if (out16) {
                    unsigned short& h = static_cast<unsigned short*>(out)[i];
                    const unsigned short th = fe_f2h_rn(thr);
                    h = t < Tb ? (half_bits_to_float(h) < half_bits_to_float(th) ? th : h) : (unsigned short)0;
                } else {
                    float& v = static_cast<float*>(out)[i];
                    v = t < Tb ? (v < thr ? thr : v) : 0.0f;
                }
            }
    }
    return 0;
}

extern "C" int emu_whisper(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                           int n_mels, int padding, int norm_mode, float* out, int64_t T_out) {
    return emu_whisper_any(wave, false, lengths, B, L, stride, n_mels, padding, norm_mode, out, false, T_out);
}

extern "C" int emu_whisper3(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                            int n_mels, int padding, int norm_mode, void* out, int out16, int64_t T_out) {
    return emu_whisper_any(wave, false, lengths, B, L, stride, n_mels, padding, norm_mode, out, out16 != 0, T_out, 3);
}

// b200feat_whisper_logmel_windows: the Tensor-Memory kernel's windowed unit map + the fix-up's arithmetic per element
extern "C" int emu_whisper3_windows(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride, int n_mels,
                                    int padding, int norm_mode, int win_frames, int n_windows, void* out, int out16) {
    typedef FeW3Cfg C;
    std::vector<int> keys(B > 0 ? B : 1, FE_KEY_INIT);
    FeWhisper2Params p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B; p.padding = padding;
    p.n_mels = n_mels; p.norm_mode = norm_mode; p.out = out;
    p.out_batch_stride = (int64_t)n_windows * n_mels * win_frames; p.T_out = (int64_t)n_windows * win_frames;
    p.win_frames = win_frames; p.n_windows = n_windows;
    p.units_per_clip = n_windows * ((win_frames + C::UF - 1) / C::UF) + 1; p.total_units = B * p.units_per_clip;
    p.vec_ok = ((((uintptr_t)wave & 15) == 0) && (stride % 4 == 0)) ? 1 : 0;
    if (!(p.vec_ok & 1)) return -2;
    if ((((uintptr_t)out & 15) == 0) && (p.out_batch_stride % 4 == 0)) p.vec_ok |= 2;
    std::vector<float2> mm((size_t)std::max(p.total_units, 1));
    p.max_keys = keys.data(); p.unit_minmax = mm.data();
    {
        const int64_t first = (int64_t)C::TMAP_ROW0 * 160 - 200;
        const int64_t rows = L - first - C::PITCH < 0 ? 0 : (L - first - C::PITCH) / 160 + 1;
        p.use_tmap = (rows >= C::ROWS && use_tmap_env()) ? 1 : 0;
        p.tmap_rows = p.use_tmap ? (int)rows : 0;
    }
    if (n_mels == 80 && !out16) run_whisper3<80, false>(p);
    else if (n_mels == 80) run_whisper3<80, true>(p);
    else if (n_mels == 128 && !out16) run_whisper3<128, false>(p);
    else if (n_mels == 128) run_whisper3<128, true>(p);
    else return -1;
    for (int b = 0; b < B; ++b) {
        const float thr = fe_w2_threshold(fe_key_float(keys[norm_mode == FE_NORM_PER_CALL ? 0 : b]));
        const int64_t Tb = (fe_clip_len(lengths, b, L) + padding) / 160;
        for (int w = 0; w < n_windows; ++w)
            for (int m = 0; m < n_mels; ++m)
                for (int c = 0; c < win_frames; ++c) {
                    const int64_t i = (int64_t)b * p.out_batch_stride + ((int64_t)w * n_mels + m) * win_frames + c;
                    const int64_t t = (int64_t)w * win_frames + c;
                    if (out16) {
                        unsigned short& h = static_cast<unsigned short*>(out)[i];
                        const unsigned short th = fe_f2h_rn(thr);
                        h = t < Tb ? (half_bits_to_float(h) < half_bits_to_float(th) ? th : h) : (unsigned short)0;
                    } else {
                        float& v = static_cast<float*>(out)[i];
                        v = t < Tb ? (v < thr ? thr : v) : 0.0f;
                    }
                }
    }
    return 0;
}

extern "C" int emu_whisper_f16(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                               int n_mels, int padding, int norm_mode, unsigned short* out, int64_t T_out) {
    return emu_whisper_any(wave, false, lengths, B, L, stride, n_mels, padding, norm_mode, out, true, T_out);
}

extern "C" int emu_whisper_s16(const short* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                               int n_mels, int padding, int norm_mode, float* out, int64_t T_out) {
    return emu_whisper_any(wave, true, lengths, B, L, stride, n_mels, padding, norm_mode, out, false, T_out);
}

// ------------------------------------------------------------------------------------------ kaldi
#include "fe_kaldi.cuh"
extern "C" int emu_kaldi(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                         int subtract_mean, float* out, int64_t T_out) {
    typedef FeKCfg C;
    FeKaldiParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B;
    p.subtract_mean = subtract_mean; p.out = out; p.out_batch_stride = T_out * 80; p.T_out = T_out;
    p.tiles_per_clip = (int)((T_out + 31) / 32);
    p.vec_ok = ((((uintptr_t)wave & 15) == 0) && (stride % 4 == 0)) ? 1 : 0;
    if ((((uintptr_t)out & 15) == 0) && (p.out_batch_stride % 4 == 0)) p.vec_ok |= 2;
    std::vector<float> colsum((size_t)B * p.tiles_per_clip * 80, 0.0f);
    p.colsum = colsum.data(); p.clip0 = 0; p.nclips = B;
    std::vector<float> sm(C::SMEM_FLOATS);
    for (int cta = 0; cta < B * p.tiles_per_clip; ++cta) {
        for (int tid = 0; tid < C::NT; ++tid) fe_k_phase_load(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_k_phase_sum(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_k_phase_a(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_k_phase_b(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_k_phase_mel(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_k_phase_store(p, sm.data(), tid, cta);
    }
    if (subtract_mean) {
        for (int b = 0; b < B; ++b) {
            const int64_t len = fe_clip_len(lengths, b, L);
            int64_t T = len < 400 ? 0 : 1 + (len - 400) / 160;
            if (T > T_out) T = T_out;
            for (int m = 0; m < 80; ++m) {
                const float mean = fe_k_clip_mean(p, b, m, (int)T);
                for (int64_t t = 0; t < T; ++t) out[(int64_t)b * p.out_batch_stride + t * 80 + m] -= mean;
            }
        }
    }
    return 0;
}

// Tensor-Memory kaldi kernel (fe_kaldi3.cuh): box copy + patch, row sums, phases A / B / M in the order the mbarriers
// enforce on the GPU, then the CMN kernel's arithmetic (4 row groups per column, fixed order)
#include "fe_kaldi3.cuh"
extern "C" int emu_kaldi3(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                          int subtract_mean, float* out, int64_t T_out) {
    typedef FeK3Cfg C;
    FeKaldiParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B;
    p.subtract_mean = subtract_mean; p.out = out; p.out_batch_stride = T_out * 80; p.T_out = T_out;
    p.vec_ok = ((((uintptr_t)wave & 15) == 0) && (stride % 4 == 0)) ? 1 : 0;
    if ((((uintptr_t)out & 15) == 0) && (p.out_batch_stride % 4 == 0)) p.vec_ok |= 2;
    const int64_t rows = L < C::PITCH ? 0 : (L - C::PITCH) / C::HOP + 1;
    if (!(p.vec_ok & 1) || rows < C::ROWS) return -2;      // the library falls back to the shared-memory kernel
    p.tmap_rows = (int)rows;
    p.units_per_clip = (int)((T_out + C::UF - 1) / C::UF);
    p.total_units = B * p.units_per_clip;
    std::vector<float> sm(C::SMEM_FLOATS, 0.0f);
    std::vector<float> tmem((size_t)C::TEAMS * 32 * C::TMEM_COLS, 0.0f);
    fe_k3_copy_tables(sm.data(), 0, 1);
    float* P = sm.data() + C::OFF_P;
    for (int u = 0; u < p.total_units; ++u) {
        const FeK3Unit t = fe_k3_unit(p, u);
        if (!(t.t0 < t.T)) {
            for (int w = 0; w < C::NW; ++w) for (int l = 0; l < 32; ++l) fe_k3_zero_unit(p, t, l, w);
            continue;
        }
        for (int q = 0; q < C::TEAMS; ++q) {
            if (!fe_k3_team_live(t, q)) continue;
            float* wq = sm.data() + C::OFF_W + q * C::WT;
            float* part = sm.data() + C::OFF_PART + q * C::PARTS * C::PARTW;
            fe_k3_box_host(p, t, q, wq);
            if (!fe_k3_box_complete(p, t, q))
                for (int m = 0; m < C::MEMBERS; ++m) for (int l = 0; l < 32; ++l) fe_k3_patch_rows(p, t, q, wq, m, l);
            {   // pre-pass: every thread of the team loads, (team barrier), every thread stores
                std::vector<FeK3Pre> regs(C::MEMBERS * 32);
                for (int m = 0; m < C::MEMBERS; ++m) for (int l = 0; l < 32; ++l) fe_k3_pre_load(wq, m, l, regs[m * 32 + l]);
                for (int m = 0; m < C::MEMBERS; ++m) for (int l = 0; l < 32; ++l) fe_k3_pre_store(wq, part, m, l, regs[m * 32 + l]);
            }
            for (int m = 0; m < C::MEMBERS; ++m)
                for (int k = 0; k < 2; ++k)
                    for (int l = 0; l < 32; ++l) {
                        FeTm tm; tm.row = tmem.data() + ((size_t)q * 32 + l) * C::TMEM_COLS;
                        fe_k3_stage_a(wq, tm, l, 2 * m + k, fe_k3_frame_mean(part, l), fe_k3_frame_x0(part, l), FeNoHook());
                    }
            for (int m = 0; m < C::MEMBERS; ++m)
                for (int k = 0; k < fe_k3_b_count(m); ++k)
                    for (int l = 0; l < 32; ++l) {
                        FeTm tm; tm.row = tmem.data() + ((size_t)q * 32 + l) * C::TMEM_COLS;
                        fe_k3_stage_b(P, tm, l, q, fe_k3_b_task(m, k), FeNoHook(), FeNoHook());
                    }
        }
        for (int w = 0; w < C::NW; ++w) for (int l = 0; l < 32; ++l) fe_k3_phase_mel(p, t, sm.data(), l, w);
    }
    if (subtract_mean) {
        for (int b = 0; b < B; ++b) {
            const int T = fe_k3_unit_at(p, b, 0).T;
            float* o = out + (int64_t)b * p.out_batch_stride;
            for (int m = 0; m < 80 && T > 0; ++m) {
                float part[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
                for (int r = 0; r < T; ++r) part[r & 7] += o[(int64_t)r * 80 + m];
                const float mean = (((part[0] + part[1]) + (part[2] + part[3])) + ((part[4] + part[5]) + (part[6] + part[7]))) / (float)T;
                for (int r = 0; r < T; ++r) o[(int64_t)r * 80 + m] -= mean;
            }
        }
    }
    return 0;
}

// ------------------------------------------------------------------------------------------ cosy
#include "fe_cosy.cuh"
template <class C>
static int emu_stft3(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride, float* out, int64_t T_out,
                     int frame_major, int pad_mode) {
    FeCosyParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B;
    p.out = out; p.out_batch_stride = T_out * 80; p.T_out = T_out; p.frame_major = frame_major; p.pad_mode = pad_mode;
    p.tiles_per_clip = (int)((T_out + C::FT - 1) / C::FT);
    p.vec_ok = (((uintptr_t)wave & 15) == 0) && (stride % 4 == 0);
    p.clip0 = 0; p.nclips = B;
    std::vector<float> sm(C::SMEM_FLOATS);
    for (int cta = 0; cta < B * p.tiles_per_clip; ++cta) {
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_load<C>(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_a<C>(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_b<C>(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_c<C>(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_mel<C>(p, sm.data(), tid, cta);
    }
    return 0;
}
extern "C" int emu_cosy(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                        float* out, int64_t T_out) {
    return emu_stft3<FeCCfg>(wave, lengths, B, L, stride, out, T_out, 0, 0);
}
// Wav2Lip / Diff2Lip mel (FeLCfg): out (B, 80, T_out) or, frame_major, (B, T_out, 80)
extern "C" int emu_wav2lip(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                           float* out, int64_t T_out, int frame_major, int pad_mode) {
    return emu_stft3<FeLCfg>(wave, lengths, B, L, stride, out, T_out, frame_major, pad_mode);
}

// ------------------------------------------------------------------------------------------ spectral gate
#include "fe_gate.cuh"
#include <algorithm>
extern "C" int emu_gate(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride, float* out) {
    typedef FeGCfg C;
    const int64_t Ts = 1 + L / 256, L_out = 256 * (L / 256);
    std::vector<float2> spec((size_t)B * 513 * Ts);
    std::vector<float> nf((size_t)B * 513, 0.0f);
    std::vector<int32_t> peak(B > 0 ? B : 1, 0);
    FeGateParams g;
    memset(&g, 0, sizeof(g));
    g.wave = wave; g.lengths = lengths; g.wave_stride = stride; g.L = L; g.B = B;
    g.spec = spec.data(); g.Ts = Ts; g.nf = nf.data(); g.peak = peak.data(); g.out = out; g.out_stride = L_out; g.L_out = L_out;
    g.tiles_per_clip = (int)((L_out / FeGICfg::HOP + 2 + FeGICfg::OUT_HOPS - 1) / FeGICfg::OUT_HOPS);
    FeCosyParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B; p.T_out = Ts;
    p.tiles_per_clip = (int)((Ts + C::FT - 1) / C::FT);
    p.vec_ok = (((uintptr_t)wave & 15) == 0) && (stride % 4 == 0);
    p.clip0 = 0; p.nclips = B;
    {   // kernel 1: STFT
        std::vector<float> sm(C::SMEM_FLOATS);
        for (int cta = 0; cta < B * p.tiles_per_clip; ++cta) {
            for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_load<C>(p, sm.data(), tid, cta);
            for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_a<C>(p, sm.data(), tid, cta);
            for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_b<C>(p, sm.data(), tid, cta);
            for (int tid = 0; tid < C::NT; ++tid) fe_g_phase_c_spec(g, p, sm.data(), tid, cta);
        }
    }
    // kernel 2 restated (the GPU kernel selects by radix; here: sort): torch.quantile's linear interpolation
    for (int b = 0; b < B; ++b) {
        const int n = (int)FeGateTraits::frames(fe_clip_len(lengths, b, L), 0);
        for (int bin = 0; bin < 513 && n > 0; ++bin) {
            std::vector<float> m(n);
            for (int t = 0; t < n; ++t) { const float2 x = spec[((size_t)b * 513 + bin) * Ts + t]; m[t] = sqrtf(x.x * x.x + x.y * x.y); }
            std::sort(m.begin(), m.end());
            const float pos = 0.1f * (float)(n - 1);
            const int lo = (int)floorf(pos);
            const float frac = pos - (float)lo, v_lo = m[lo], v_hi = m[std::min(lo + 1, n - 1)], d = v_hi - v_lo;
            nf[(size_t)b * 513 + bin] = frac < 0.5f ? v_lo + frac * d : v_hi - d * (1.0f - frac);
        }
    }
    {   // kernel 3: inverse
        std::vector<float> sm(FeGICfg::SMEM_FLOATS);
        for (int cta = 0; cta < B * g.tiles_per_clip; ++cta) {
            for (int tid = 0; tid < FeGICfg::NT; ++tid) fe_gi_phase_load(g, sm.data(), tid, cta);
            for (int tid = 0; tid < FeGICfg::NT; ++tid) fe_gi_phase_1(sm.data(), tid);
            for (int tid = 0; tid < FeGICfg::NT; ++tid) fe_gi_phase_2(sm.data(), tid);
            for (int tid = 0; tid < FeGICfg::NT; ++tid) fe_gi_phase_3(sm.data(), tid);
            float mx = 0.0f;
            for (int tid = 0; tid < FeGICfg::NT; ++tid) mx = std::max(mx, fe_gi_phase_ola(g, sm.data(), tid, cta));
            union { float f; int32_t i; } u; u.f = mx;
            int32_t& pk = peak[cta / g.tiles_per_clip];
            pk = std::max(pk, u.i);
        }
    }
    for (int b = 0; b < B; ++b) {   // kernel 4
        union { float f; int32_t i; } u; u.i = peak[b];
        if (!(u.f > 0.0f)) continue;
        for (int64_t i = 0; i < L_out; ++i) out[(size_t)b * L_out + i] = out[(size_t)b * L_out + i] / u.f * 0.95f;
    }
    return 0;
}

// ------------------------------------------------------------------------------------------ resample
#include "fe_resample.cuh"
#define FE_RS_HAVE_PARAMS
#include "fe_resample_plan.h"
template <bool IN16, int CH>
static void emu_rs_run(FeResampleParams& p, const FeRsPlan& pl, int B) {
    std::vector<float> xs(pl.xs_floats), ost((size_t)32 * pl.OP);
    for (int cta = 0; cta < B * p.tiles_per_clip; ++cta) {
        std::fill(xs.begin(), xs.end(), 1e30f);      // anything read but not loaded shows up
        std::fill(ost.begin(), ost.end(), 1e30f);
        for (int tid = 0; tid < FeRsCfg::NT; ++tid) fe_rs_load<IN16, CH>(p, xs.data(), tid, cta);
        for (int tid = 0; tid < FeRsCfg::NT; ++tid) fe_rs_compute(p, xs.data(), ost.data(), p.taps, tid);
        for (int tid = 0; tid < FeRsCfg::NT; ++tid) fe_rs_store(p, ost.data(), tid, cta);
    }
}

// in: (B, L_in, channels) interleaved float32 or int16; out: (B, L_out) with L_out = ceil(new * L_in / orig);
// info (optional, 8 ints): orig, new, m, width, G, J, smem_bytes, taps_smem
extern "C" int emu_resample(const void* in, int in_s16, int channels, const float* mix, const int64_t* lengths, int B,
                            int64_t L_in, int orig_freq, int new_freq, int lpfw, double rolloff, int method,
                            double beta, float* out, int64_t L_out, int* info) {
    FeRsPlan pl = fe_rs_make_plan(orig_freq, new_freq, lpfw, rolloff, method, beta);
    if (!pl.error.empty()) return -1;
    if (info) {
        info[0] = pl.orig; info[1] = pl.newf; info[2] = pl.m; info[3] = pl.width; info[4] = pl.G; info[5] = pl.J;
        info[6] = (int)pl.smem_bytes; info[7] = pl.taps_smem;
    }
    FeResampleParams p;
    memset(&p, 0, sizeof(p));
    fe_rs_fill_params(p, pl, L_in, L_out, channels, in_s16 != 0, mix);
    p.in = in; p.lengths = lengths; p.out = out; p.in_stride = L_in * channels; p.out_stride = L_out;
    p.taps = (const float4*)pl.taps.data(); p.kstart = pl.kstart.data(); p.clip0 = 0;
    p.vec_ok = ((uintptr_t)in % (size_t)(channels * (in_s16 ? 2 : 4)) == 0) ? 1 : 0;
    if (in_s16) { if (channels == 1) emu_rs_run<true, 1>(p, pl, B); else emu_rs_run<true, 2>(p, pl, B); }
    else        { if (channels == 1) emu_rs_run<false, 1>(p, pl, B); else emu_rs_run<false, 2>(p, pl, B); }
    return 0;
}
