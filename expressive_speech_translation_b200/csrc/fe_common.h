// Shared definitions of the featurizer kernels (sm_100a) and of their host-side emulation build.
//
// Every kernel is written as a sequence of PHASES separated by block barriers.  A phase is a plain
// function of (params, shared-memory base, thread id, CTA id) that uses no warp intrinsics, so the
// identical source also compiles with g++ (tests/emu/): there the phases run thread after thread over
// an ordinary array.  That build exists to check index maps, codelets and tables on the GPU-less
// authoring box; it is test infrastructure and is never linked into libb200feat.so.
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define FE_HD __device__
#define FE_INLINE __forceinline__
#define FE_CONSTANT __constant__ const
#define FE_TABLE __device__ const       // tables that do not fit the 64 KB constant bank: global memory, read through L1
// lg2.approx.ftz: one MUFU, max abs error 2^-22 on log2; every caller floors its argument at a
// normal number first (1e-10, 1e-5, FLT_EPSILON), so flushing subnormals never triggers
static __device__ __forceinline__ float fe_lg2_ftz(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
#define FE_LOG2(x) fe_lg2_ftz(x)
#define FE_SQRT(x) sqrtf(x)
#else
#define FE_HD
#define FE_INLINE inline __attribute__((always_inline))
#define FE_CONSTANT static const
#define FE_TABLE static const
#define FE_LOG2(x) log2f(x)
#define FE_SQRT(x) sqrtf(x)
struct float2 { float x, y; };
struct float4 { float x, y, z, w; };
struct int2 { int x, y; };
static inline float2 make_float2(float a, float b) { float2 r; r.x = a; r.y = b; return r; }
static inline float4 make_float4(float a, float b, float c, float d) { float4 r; r.x = a; r.y = b; r.z = c; r.w = d; return r; }
#endif

#define FE_WARP 32

// ---------------------------------------------------------------------------------------------------
// Value types of the generated codelets.  `float` = one frame per lane; `fe2` = a PAIR of frames per
// lane, every add / mul / fma issued once for both as a packed sm_100 instruction (FADD2 / FMUL2 /
// FFMA2: same FP32 pipe time as two scalar instructions, half the issue slots, scalar constants are
// broadcast by the instruction itself).  nvcc never contracts the packed intrinsics, so the generator
// emits fused multiply-adds explicitly (tools/fftgen.py).
typedef float2 fe2;

template <class V> FE_HD FE_INLINE V fe_zero();
template <> FE_HD FE_INLINE float fe_zero<float>() { return 0.0f; }
template <> FE_HD FE_INLINE fe2 fe_zero<fe2>() { fe2 r; r.x = 0.0f; r.y = 0.0f; return r; }

FE_HD FE_INLINE float fe_add(float a, float b) { return a + b; }
FE_HD FE_INLINE float fe_sub(float a, float b) { return a - b; }
FE_HD FE_INLINE float fe_neg(float a) { return -a; }
FE_HD FE_INLINE float fe_mulc(float c, float a) { return c * a; }
FE_HD FE_INLINE float fe_fmac(float c, float x, float y) { return fmaf(c, x, y); }     // c*x + y
FE_HD FE_INLINE float fe_fmsc(float c, float x, float y) { return fmaf(c, x, -y); }    // c*x - y
FE_HD FE_INLINE float fe_fnmac(float c, float x, float y) { return fmaf(-c, x, y); }   // y - c*x
FE_HD FE_INLINE float fe_mul(float a, float b) { return a * b; }
FE_HD FE_INLINE float fe_fma(float a, float b, float c) { return fmaf(a, b, c); }

FE_HD FE_INLINE fe2 fe_mk2(float a, float b) { fe2 r; r.x = a; r.y = b; return r; }
#if defined(__CUDACC__)
FE_HD FE_INLINE fe2 fe_add(fe2 a, fe2 b) { return __fadd2_rn(a, b); }
FE_HD FE_INLINE fe2 fe_sub(fe2 a, fe2 b) { return __fadd2_rn(a, fe_mk2(-b.x, -b.y)); }
FE_HD FE_INLINE fe2 fe_neg(fe2 a) { return fe_mk2(-a.x, -a.y); }
FE_HD FE_INLINE fe2 fe_mulc(float c, fe2 a) { return __fmul2_rn(a, fe_mk2(c, c)); }
FE_HD FE_INLINE fe2 fe_fmac(float c, fe2 x, fe2 y) { return __ffma2_rn(x, fe_mk2(c, c), y); }
FE_HD FE_INLINE fe2 fe_fmsc(float c, fe2 x, fe2 y) { return __ffma2_rn(x, fe_mk2(c, c), fe_mk2(-y.x, -y.y)); }
FE_HD FE_INLINE fe2 fe_fnmac(float c, fe2 x, fe2 y) { return __ffma2_rn(x, fe_mk2(-c, -c), y); }
FE_HD FE_INLINE fe2 fe_mul(fe2 a, fe2 b) { return __fmul2_rn(a, b); }
FE_HD FE_INLINE fe2 fe_fma(fe2 a, fe2 b, fe2 c) { return __ffma2_rn(a, b, c); }
#else
FE_HD FE_INLINE fe2 fe_add(fe2 a, fe2 b) { return fe_mk2(a.x + b.x, a.y + b.y); }
FE_HD FE_INLINE fe2 fe_sub(fe2 a, fe2 b) { return fe_mk2(a.x - b.x, a.y - b.y); }
FE_HD FE_INLINE fe2 fe_neg(fe2 a) { return fe_mk2(-a.x, -a.y); }
FE_HD FE_INLINE fe2 fe_mulc(float c, fe2 a) { return fe_mk2(c * a.x, c * a.y); }
FE_HD FE_INLINE fe2 fe_fmac(float c, fe2 x, fe2 y) { return fe_mk2(fmaf(c, x.x, y.x), fmaf(c, x.y, y.y)); }
FE_HD FE_INLINE fe2 fe_fmsc(float c, fe2 x, fe2 y) { return fe_mk2(fmaf(c, x.x, -y.x), fmaf(c, x.y, -y.y)); }
FE_HD FE_INLINE fe2 fe_fnmac(float c, fe2 x, fe2 y) { return fe_mk2(fmaf(-c, x.x, y.x), fmaf(-c, x.y, y.y)); }
FE_HD FE_INLINE fe2 fe_mul(fe2 a, fe2 b) { return fe_mk2(a.x * b.x, a.y * b.y); }
FE_HD FE_INLINE fe2 fe_fma(fe2 a, fe2 b, fe2 c) { return fe_mk2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
#endif


// Rolled sparse mel projection of the one-frame-per-lane kernels (tables: generated/fe_tables.h, fe_melroll_*):
// filters m0, m0 + mstep, ... < m1; ld(k) returns this lane's power at bin k, emit(m, value) is called once per
// filter.  Four partial sums per filter, taps zero-padded to a multiple of four.  (The fully unrolled form with
// immediate weights is thousands of instructions of which every warp runs a different part: the kernels then
// starve on instruction fetch.)
template <class Load, class Emit>
FE_HD FE_INLINE void fe_mel_rolled(const int* __restrict__ desc, const float* __restrict__ w, int m0, int m1, int mstep,
                                   Load ld, Emit emit) {
#pragma unroll 1
    for (int m = m0; m < m1; m += mstep) {
        const int k0 = desc[3 * m], n4 = desc[3 * m + 1];
        const float* __restrict__ wm = w + desc[3 * m + 2];
        float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll 2
        for (int j = 0; j < n4; j += 4) {
            a0 = fmaf(wm[j], ld(k0 + j), a0);
            a1 = fmaf(wm[j + 1], ld(k0 + j + 1), a1);
            a2 = fmaf(wm[j + 2], ld(k0 + j + 2), a2);
            a3 = fmaf(wm[j + 3], ld(k0 + j + 3), a3);
        }
        emit(m, (a0 + a1) + (a2 + a3));
    }
}

// order-preserving float <-> int key (signed compare of keys == float compare)
FE_HD FE_INLINE int fe_float_key(float f) {
    union { float f; int i; } u;
    u.f = f;
    return u.i >= 0 ? u.i : (u.i ^ 0x7fffffff);
}
FE_HD FE_INLINE float fe_key_float(int k) {
    union { float f; int i; } u;
    u.i = k >= 0 ? k : (k ^ 0x7fffffff);
    return u.f;
}
// key lower than the key of every finite value the kernels produce (memset byte 0x80)
#define FE_KEY_INIT_BYTE 0x80
#define FE_KEY_INIT ((int)0x80808080)

// reflect-pad fetch: index g into a signal of `lp` samples (the first `len` real, the rest zeros),
// reflected without repeating the edge sample (torch pad_mode="reflect").
FE_HD FE_INLINE float fe_fetch_reflect(const float* __restrict__ x, int64_t g, int64_t len, int64_t lp) {
    if (g < 0) g = -g;
    if (g >= lp) g = 2 * (lp - 1) - g;
    if (g < 0) g = 0;               // degenerate clips shorter than the pad: stay in bounds
    if (g >= lp) g = lp - 1;
    return (g < len) ? x[g] : 0.0f;
}

enum { FE_NORM_PER_CLIP = 0, FE_NORM_PER_CALL = 1 };

struct FeWhisperParams {
    const float* wave;         // (B, wave_stride) fp32
    const int64_t* lengths;    // nullable, device: valid samples per clip
    int64_t wave_stride;
    int64_t L;                 // samples per clip when lengths == nullptr (also the clamp for lengths)
    int32_t B;
    int32_t padding;           // zeros appended to every clip before framing
    int32_t n_mels;            // 80 | 128
    int32_t norm_mode;         // FE_NORM_*
    float* out;                // (B, n_mels, T_out)
    int64_t out_batch_stride;  // floats between clips in `out`
    int64_t T_out;             // frames per output row (row stride)
    int32_t tiles_per_clip;    // ceil(T_out / 32)
    int32_t vec_ok;            // wave base 16B-aligned and wave_stride % 4 == 0
    int32_t* max_keys;         // (B) or (1): order-preserving keys of the running max
    int32_t clip0;             // first clip of this launch (clips are processed in waves)
    int32_t nclips;            // clips in this launch
};

struct FeCosyParams {
    const float* wave;         // (B, wave_stride) fp32 @ 24 kHz
    const int64_t* lengths;
    int64_t wave_stride;
    int64_t L;
    int32_t B;
    float* out;                // (B, 80, T_out), or (B, T_out, 80) when frame_major
    int64_t out_batch_stride;
    int64_t T_out;
    int32_t frame_major;       // 1: the CosyVoice frontend's layout (speech_feat.transpose(0, 1)): frames outermost
    int32_t tiles_per_clip;    // ceil(T_out / 16)
    int32_t vec_ok;
    int32_t clip0, nclips;     // clips [clip0, clip0 + nclips) are processed by this launch
    int32_t pad_mode;          // Wav2Lip mel only: 0 = reflect (librosa 0.9.2, the reference's pin), 1 = zeros (librosa >= 0.10)
};

struct FeKaldiParams {
    const float* wave;         // (B, wave_stride) fp32 @ 16 kHz
    const int64_t* lengths;
    int64_t wave_stride;
    int64_t L;
    int32_t B;
    int32_t subtract_mean;
    float* out;                // (B, T_out, 80) frame-major
    int64_t out_batch_stride;
    int64_t T_out;
    int32_t tiles_per_clip;    // ceil(T_out / 32)
    int32_t vec_ok;            // bit 0: wave float4-loadable, bit 1: out float4-storable
    float* colsum;             // workspace (B, tiles_per_clip, 80): per-tile column sums for CMN
    int32_t clip0, nclips;
    // Tensor-Memory kernel (fe_kaldi3.cuh) only:
    int32_t tmap_rows;         // rows per clip in the launch's tensor map
    int32_t units_per_clip;    // ceil(T_out / 128)
    int32_t total_units;       // B * units_per_clip
    int32_t step_b, step_w;    // gridDim.x units = step_w unit positions + step_b clips
};

FE_HD FE_INLINE int64_t fe_clip_len(const int64_t* lengths, int b, int64_t L) {
    if (lengths == nullptr) return L;
    int64_t v = lengths[b];
    if (v < 0) v = 0;
    if (v > L) v = L;
    return v;
}
