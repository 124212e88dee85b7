// Micro-benchmark (round 2): can Tensor Memory take FFT scratch traffic off the shared-memory pipe,
// and what does a tiny-N tcgen05.mma cost?  Three measurements on one B200, every SM busy (grid = #SMs):
//   1. tcgen05.st / tcgen05.ld .32x32b.xN throughput per SM (bytes per cycle) for 4 / 8 / 16 warps, N = 1, 4, 16, 32
//   2. the same tcgen05.ld stream with a conflict-free LDS.64 stream running in the other half of the warps: are the
//      two pipes additive?
//   3. issue rate of tcgen05.mma kind::tf32, M = 128, K = 8, N = 32 / 64 / 256, A and B from shared memory
//      (the shape a "first FFT pass on the tensor cores" would use)
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o bin/tmem_bw tmem_bw.cu ; run on a B200.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_slot);
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(a), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int N> struct TmemIO;
template <> struct TmemIO<1> {
    static __device__ __forceinline__ void st(uint32_t t, const uint32_t* r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(t), "r"(r[0]) : "memory");
    }
    static __device__ __forceinline__ void ld(uint32_t t, uint32_t* r) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r[0]) : "r"(t) : "memory");
    }
};
template <> struct TmemIO<4> {
    static __device__ __forceinline__ void st(uint32_t t, const uint32_t* r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(t), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
    }
    static __device__ __forceinline__ void ld(uint32_t t, uint32_t* r) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(t) : "memory");
    }
};
template <> struct TmemIO<16> {
    static __device__ __forceinline__ void st(uint32_t t, const uint32_t* r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                     ::"r"(t), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                       "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
    }
    static __device__ __forceinline__ void ld(uint32_t t, uint32_t* r) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(t) : "memory");
    }
};
template <> struct TmemIO<32> {
    static __device__ __forceinline__ void st(uint32_t t, const uint32_t* r) { TmemIO<16>::st(t, r); TmemIO<16>::st(t + 16, r + 16); }
    static __device__ __forceinline__ void ld(uint32_t t, uint32_t* r) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                       "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                       "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(t) : "memory");
    }
};

// MODE 0: stores only, 1: loads only, 2: even warp-quads load TMEM while odd ones run LDS.64, 3: LDS.64 only (odd quads)
template <int N, int MODE>
__global__ void __launch_bounds__(512) tmem_rw(long long* cycles, uint32_t* sink, int iters) {
    __shared__ uint32_t slot;
    __shared__ float2 buf[32 * 65];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 32 * 65; i += blockDim.x) buf[i] = make_float2((float)i, 1.0f);
    if (warp == 0) tmem_alloc(&slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = slot + ((uint32_t)(warp & 3) * 32u << 16);
    uint32_t r[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = threadIdx.x * 37u + i;
    // every warp first fills its column window so that loads read defined data
    const int win = (warp >> 2) * 128;   // 4 warps share a lane quadrant: 128 columns each
    for (int c = 0; c < 128; c += 32) TmemIO<32>::st(base + win + c, r);
    tmem_wait_st();
    __syncthreads();
    const long long t0 = clock64();
    float2 acc = make_float2(0.f, 0.f);
    const bool lds_warp = ((warp >> 2) & 1) != 0;
    if (MODE == 0) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int c = 0; c < 128; c += N) TmemIO<N>::st(base + win + c, r);
        }
        tmem_wait_st();
    } else if (MODE == 1 || (MODE == 2 && !lds_warp)) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int c = 0; c < 128; c += 64) {      // 64 words in flight per lane, then one wait
                uint32_t v[64];
#pragma unroll
                for (int k = 0; k < 64; k += N) TmemIO<N>::ld(base + win + c + k, v + k);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 64; ++j) r[j & 31] ^= v[j];
            }
        }
    } else if (MODE == 2 || (MODE == 3 && lds_warp)) {
        // 128 / N * N words per iteration and warp, like the TMEM stream: 64 LDS.64 (lane stride 65 pairs: conflict-free)
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int c = 0; c < 64; ++c) {
                const float2 v = buf[lane * 65 + ((c + it) & 63)];
                acc.x += v.x; acc.y += v.y;
            }
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    uint32_t x = __float_as_uint(acc.x + acc.y);
#pragma unroll
    for (int i = 0; i < 32; ++i) x ^= r[i];
    if (x == 0x12345678u) sink[0] = x;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(slot, 512);
}

// ---- tcgen05.mma issue rate
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffff) >> 4);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;          // descriptor version (sm_100)
    return d;                        // layout type 0: no swizzle, K-major
}
__device__ __forceinline__ uint32_t idesc_tf32(int M, int N) {
    uint32_t d = 0;
    d |= 1u << 4;                    // D format: f32
    d |= 2u << 7;                    // A format: tf32
    d |= 2u << 10;                   // B format: tf32
    d |= (uint32_t)(N >> 3) << 17;
    d |= (uint32_t)(M >> 4) << 24;
    return d;
}

__global__ void __launch_bounds__(128) mma_rate(long long* cycles, int N, int n_mma, int* status) {
    extern __shared__ __align__(1024) float smf[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long mbar;
    const int warp = threadIdx.x >> 5;
    float* A = smf;                  // 128 x 8 tf32, K-major core matrices: 4 KB
    float* B = smf + 1024;           // N x 8: N * 32 bytes
    for (int i = threadIdx.x; i < 1024 + 256 * 8; i += blockDim.x) smf[i] = 1.0f / (1 + (i & 15));
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(&slot, 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes of A / B -> async proxy (UMMA)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t d_tmem = slot;
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0) {
        const uint64_t da = smem_desc((uint32_t)__cvta_generic_to_shared(A), 128, 256);
        const uint64_t db = smem_desc((uint32_t)__cvta_generic_to_shared(B), 128, 256);
        const uint32_t id = idesc_tf32(128, N);
        t0 = clock64();
        for (int i = 0; i < n_mma; ++i) {
            const uint32_t acc = i > 0 ? 1u : 0u;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                ::"r"(d_tmem + (uint32_t)((i & 1) * 256)), "l"(da), "l"(db), "r"(id), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mb) : "memory");
        // bounded wait: a wrong descriptor must not hang the box
        int ok = 0;
        for (int spin = 0; spin < (1 << 24) && !ok; ++spin) {
            asm volatile(
                "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                : "=r"(ok) : "r"(mb) : "memory");
        }
        t1 = clock64();
        cycles[blockIdx.x] = t1 - t0;
        if (!ok) status[0] = 1;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0) tmem_dealloc(slot, 512);
}

template <int N, int MODE>
static double run_rw(int warps, int iters, long long* d_cyc, uint32_t* d_sink, int sms) {
    tmem_rw<N, MODE><<<sms, warps * 32>>>(d_cyc, d_sink, iters);   // warm-up
    tmem_rw<N, MODE><<<sms, warps * 32>>>(d_cyc, d_sink, iters);
    CK(cudaDeviceSynchronize());
    static long long h[1024];
    CK(cudaMemcpy(h, d_cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
    double s = 0;
    for (int i = 0; i < sms; ++i) s += (double)h[i];
    return s / sms;
}

int main() {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    long long* d_cyc; uint32_t* d_sink; int* d_status;
    CK(cudaMalloc(&d_cyc, sizeof(long long) * 1024));
    CK(cudaMalloc(&d_sink, 64));
    CK(cudaMalloc(&d_status, 4));
    CK(cudaMemset(d_status, 0, 4));
    const int iters = 2000;
    printf("# SMs %d; bytes/cycle/SM = warps_active * iters * 128 columns * 128 B / cycles\n", sms);
    for (int warps : {4, 8, 16}) {
        const double bytes = (double)warps * iters * 128.0 * 128.0;
#define ROW(N) { \
        const double cs = run_rw<N, 0>(warps, iters, d_cyc, d_sink, sms); \
        const double cl = run_rw<N, 1>(warps, iters, d_cyc, d_sink, sms); \
        printf("tmem warps=%2d x%-2d  st %7.1f B/cyc/SM   ld %7.1f B/cyc/SM\n", warps, N, bytes / cs, bytes / cl); }
        ROW(1) ROW(4) ROW(16) ROW(32)
#undef ROW
    }
    {   // 16 warps: quads 0, 2 stream TMEM loads (x4), quads 1, 3 stream LDS.64
        const double half = 8.0 * iters * 128.0 * 128.0;
        const double both = run_rw<4, 2>(16, iters, d_cyc, d_sink, sms);
        const double lds_only = run_rw<4, 3>(16, iters, d_cyc, d_sink, sms);
        const double ld8 = run_rw<4, 1>(8, iters, d_cyc, d_sink, sms);
        printf("mix: 8 warps LDTM.x4 alone %.0f cyc (%.1f B/cyc), 8 warps LDS.64 alone %.0f cyc (%.1f B/cyc), together %.0f cyc\n",
               ld8, half / ld8, lds_only, half / lds_only, both);
    }
    for (int N : {32, 64, 256}) {
        const int n_mma = 4096;
        CK(cudaFuncSetAttribute(mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
        mma_rate<<<sms, 128, 16384>>>(d_cyc, N, n_mma, d_status);
        mma_rate<<<sms, 128, 16384>>>(d_cyc, N, n_mma, d_status);
        CK(cudaDeviceSynchronize());
        static long long h[1024];
        int st = 0;
        CK(cudaMemcpy(h, d_cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(&st, d_status, 4, cudaMemcpyDeviceToHost));
        double s = 0;
        for (int i = 0; i < sms; ++i) s += (double)h[i];
        s /= sms;
        printf("mma tf32 M=128 N=%3d K=8 (A, B in smem): %.1f cycles per MMA, %.1f flop/cycle/SM%s\n", N, s / n_mma,
               2.0 * 128 * N * 8 * n_mma / s, st ? "  [TIMEOUT]" : "");
    }
    return 0;
}
