// Band-limited resampling with LONG filters on the 5th-generation tensor cores (tcgen05.mma kind::tf32, 3 x TF32 split).
//
// Replaces torchaudio.transforms.Resample(orig, 16000, lowpass_filter_width=128, rolloff=0.95,
// resampling_method='kaiser_window', beta=14.769656459379492) as /root/reference/Backend/services/audio_processor.py:296-304
// builds it: 44.1 kHz -> 16 kHz has 160 phases of 1185 taps (743 inside the window).  On CUDA cores that filter ran at
// 0.58 M audio-s/s (profiles/r01_resample_timing.txt), 35 x slower than the featurizer it feeds.
//
// The polyphase sum IS a GEMM:   out[q * new + p] = sum_k  X[q][k] * H[p][k],   X[q][k] = xpad[q * orig + k]
// (xpad = the clip with `width` zeros in front; torchaudio functional.py `_apply_sinc_resample_kernel`), with
//   M = 128 consecutive blocks q per tile, N = new phases (a multiple of 16, <= 256), K = 2 * width + orig padded to 32.
// Rows of X overlap (row q + 1 is row q shifted by `orig` samples, and orig * 4 bytes is not a multiple of 16, so no
// tensor map can describe them): producer warps build the A operand, one row per thread, from 128-bit loads shifted in
// registers, and write it straight into TENSOR MEMORY (tcgen05.st, lane = row): the MMA takes A from Tensor Memory and only
// H from shared memory.  (With A in shared memory as well the kernel was bound by the shared-memory pipe: producer stores +
// H copies + the operand reads of 12 MMAs = 1.4 k wavefront-cycles per chunk against 960 cycles of math.)  Every value is split  v = hi + lo  (both rounded to TF32, cvt.rna), H likewise on the
// host, and each K step issues three MMAs, hi*hi + hi*lo + lo*hi, into one float32 accumulator in Tensor Memory: float32
// accuracy (the dropped lo*lo term is 2^-22 relative) at a third of the TF32 rate.
//
// BANDED form (short filters: torchaudio's default hann / width-6 filter has 33 live taps of 475 at 44.1 -> 16 kHz): the phases are
// cut into blocks of 32; a block only multiplies the K range its phases' windows cover (4 chunks of 32 taps instead of 15), into
// its own 32 accumulator columns.  The kernel runs a table of chunks {first tap, accumulator column, N, H offset}; the dense form
// is the table with one block.
//
// Roles of the 18 warps of a persistent CTA (one per SM):
//   0-11  producers   three groups of four warps, chunks in turn: A(hi, lo) chunk of 128 rows x 32 k -> Tensor Memory
//   12-15 epilogue    Tensor Memory -> registers -> out (one row of `new` samples per lane), masks ragged ends
//   16    MMA issuer  (one lane) 12 tcgen05.mma per chunk, tcgen05.commit to the stage's / the accumulator's mbarrier
//   17    H loader    (one lane) one cp.async.bulk per chunk: H(hi, lo) is stored in global memory in the operand layout
// Ring of NST stages (A: 2 x 32 Tensor Memory columns; H: 2 * N * 128 bytes of shared memory), two accumulators (tile
// i + 1's MMAs run while tile i drains).  Canonical layout of H (cute make_umma_desc<Major::K>, SWIZZLE_NONE): 8 rows x
// 16 bytes contiguous (one core matrix), the two K halves of an MMA LBO = 128 bytes apart, 8-row groups SBO = (KC / 4) * 128.
#pragma once
#include "fe_common.h"

struct FeRsTcCfg {
    static constexpr int M = 128, KC = 32, NST = 3;
    static constexpr int PGROUPS = 3;                          // producer groups of four warps (they take the chunks in turn).  Must not
                                                               // exceed NST: a group more than NST chunks ahead would wait on a stage's
                                                               // barrier two phases early, which a parity wait cannot tell apart
    static constexpr int MMA_WARP = 4 * PGROUPS + 4, NTHREADS = 32 * (MMA_WARP + 2);
    static constexpr int LBO = 128, SBO = (KC / 4) * 128;      // H in shared memory: K-major core matrices, no swizzle
    // Tensor Memory columns (512): two accumulators of N <= 160 columns at 0 and 256; the A operand's NST stages of KC
    // columns each, hi in [160, 256), lo in [416, 512)
    static constexpr int ACC_STRIDE = 256, MAX_N = 160;
    static FE_HD constexpr uint32_t a_hi_col(uint32_t s) { return 160u + (uint32_t)KC * s; }
    static FE_HD constexpr uint32_t a_lo_col(uint32_t s) { return 416u + (uint32_t)KC * s; }
    static_assert(160 + KC * NST <= 256, "A stages must fit between the accumulators");
    static constexpr int BAND = 32;                            // phases per block of the banded form
    static int smem_bytes(int stage_floats) { return NST * stage_floats * 4 + 1024; }
};

struct FeRsTcParams {
    const float* x;            // (B, x_stride) mono float32 -- or int16 stereo frames read as 4-byte elements --, rows 16-byte aligned
    int64_t x_stride;          // in 4-byte elements
    float w0, w1;              // channel weights of the int16 stereo form
    const int64_t* lengths;    // input samples per clip or nullptr
    int64_t L_in;
    float* out;                // (B, out_stride)
    int64_t out_stride, L_out;
    const float* hpack;        // per chunk: [hi | lo][n / 8][KC / 4][8][4]
    const int4* vtab;          // per chunk: x = first tap (column of X), y = first accumulator column | 0x10000 on the first chunk of a
                               // phase block, z = phases n of the block (the MMA's N), w = offset of the chunk's H in hpack (floats)
    int32_t B, orig, newf, width, nch, stage_floats;
    int32_t k0_dense;          // >= 0: one block, chunk c starts at tap k0_dense + 32 c (the producers need not read vtab); -1: banded
    int32_t tiles_per_clip, total_tiles;
    int32_t out_vec;           // out rows 16-byte aligned
};

#if defined(__CUDACC__)
static __device__ __forceinline__ uint64_t fe_umma_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffffu) >> 4);
    d |= (uint64_t)((FeRsTcCfg::LBO >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((FeRsTcCfg::SBO >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;                                    // descriptor version of sm_100
    return d;                                                  // layout type 0: no swizzle; both operands K-major
}
static __device__ __forceinline__ uint32_t fe_umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);   // D f32, A / B tf32
}
static __device__ __forceinline__ void fe_umma_tf32(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// A operand from Tensor Memory (lane = row, one 32-bit column per K element), B from shared memory
static __device__ __forceinline__ void fe_umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
static __device__ __forceinline__ void fe_umma_commit(uint32_t mbar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar) : "memory");
}
static __device__ __forceinline__ float fe_tf32_rna(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}
#endif
