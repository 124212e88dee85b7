// Host-side emulation of the featurizer kernels' phases (TEST INFRASTRUCTURE, never shipped).
// Compiles the very same phase functions the CUDA kernels call, with g++, and runs them thread after
// thread over a plain array standing in for shared memory.  Built by tests/emu/build.py.
#include <vector>
#include <cstring>
#include <algorithm>
#include "fe_common.h"
#include "generated/fe_codelets.h"
#include "generated/fe_tables.h"
#include "fe_whisper.cuh"

template <int NMELS>
static void run_whisper(FeWhisperParams p) {
    typedef FeWCfg C;
    std::vector<float> sm(C::SMEM_FLOATS);
    const int ncta = p.nclips * p.tiles_per_clip;
    for (int cta = 0; cta < ncta; ++cta) {
        for (int tid = 0; tid < C::NT; ++tid) fe_w_phase_load(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_w_phase_a(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_w_phase_b(p, sm.data(), tid, cta);
        int key = FE_KEY_INIT;
        for (int tid = 0; tid < C::NT; ++tid) key = std::max(key, fe_w_phase_mel<NMELS>(p, sm.data(), tid, cta));
        const FeWTile t = fe_w_tile(p, cta);
        int* slot = p.max_keys + (p.norm_mode == FE_NORM_PER_CALL ? 0 : t.b);
        *slot = std::max(*slot, key);
    }
}

extern "C" int emu_whisper(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                           int n_mels, int padding, int norm_mode, float* out, int64_t T_out) {
    std::vector<int> keys(B, FE_KEY_INIT);
    FeWhisperParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B; p.padding = padding;
    p.n_mels = n_mels; p.norm_mode = norm_mode; p.out = out; p.out_batch_stride = (int64_t)n_mels * T_out;
    p.T_out = T_out; p.tiles_per_clip = (int)((T_out + 31) / 32);
    p.vec_ok = (((uintptr_t)wave & 15) == 0) && (stride % 4 == 0);
    p.max_keys = keys.data(); p.clip0 = 0; p.nclips = B;
    if (n_mels == 80) run_whisper<80>(p); else if (n_mels == 128) run_whisper<128>(p); else return -1;
    for (int b = 0; b < B; ++b) {
        const float gmax = fe_key_float(keys[norm_mode == FE_NORM_PER_CALL ? 0 : b]);
        const int64_t len = fe_clip_len(lengths, b, L);
        const int64_t Tb = std::min<int64_t>((len + padding) / 160, T_out);
        for (int m = 0; m < n_mels; ++m)
            for (int64_t t = 0; t < T_out; ++t) {
                float& v = out[(int64_t)b * p.out_batch_stride + m * T_out + t];
                v = t < Tb ? fe_w_normalise(v, gmax) : 0.0f;
            }
    }
    return 0;
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

// ------------------------------------------------------------------------------------------ cosy
#include "fe_cosy.cuh"
extern "C" int emu_cosy(const float* wave, const int64_t* lengths, int B, int64_t L, int64_t stride,
                        float* out, int64_t T_out) {
    typedef FeCCfg C;
    FeCosyParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = stride; p.L = L; p.B = B;
    p.out = out; p.out_batch_stride = T_out * 80; p.T_out = T_out;
    p.tiles_per_clip = (int)((T_out + C::FT - 1) / C::FT);
    p.vec_ok = (((uintptr_t)wave & 15) == 0) && (stride % 4 == 0);
    p.clip0 = 0; p.nclips = B;
    std::vector<float> sm(C::SMEM_FLOATS);
    for (int cta = 0; cta < B * p.tiles_per_clip; ++cta) {
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_load(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_a(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_b(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_c(p, sm.data(), tid, cta);
        for (int tid = 0; tid < C::NT; ++tid) fe_c_phase_mel(p, sm.data(), tid, cta);
    }
    return 0;
}
