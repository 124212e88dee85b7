// Micro-benchmark: issue rate of scalar FFMA/FADD vs packed FFMA2/FADD2 (sm_100a), alone and mixed with LDS.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o f32x2 f32x2.cu ; run on a B200.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define ITERS 4096
#define ILP 8

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float seed, int lds_every) {
    __shared__ float2 sm[256 * 4];
    float2 a[ILP], b, c;
    b = make_float2(seed, seed * 0.5f);
    c = make_float2(0.25f, 0.125f);
    for (int i = 0; i < ILP; ++i) a[i] = make_float2(threadIdx.x + i, i);
    for (int i = threadIdx.x; i < 256 * 4; i += blockDim.x) sm[i] = make_float2(i, -i);
    __syncthreads();
    float2 acc = make_float2(0.f, 0.f);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (MODE == 0) { a[i].x = fmaf(a[i].x, b.x, c.x); a[i].y = fmaf(a[i].y, b.y, c.y); }        // 2 FFMA
            if (MODE == 1) { a[i] = __ffma2_rn(a[i], b, c); }                                            // 1 FFMA2
            if (MODE == 2) { a[i].x = a[i].x + b.x; a[i].y = a[i].y + b.y; }                             // 2 FADD
            if (MODE == 3) { a[i] = __fadd2_rn(a[i], b); }                                               // 1 FADD2
            if (MODE == 4) { a[i] = __fmul2_rn(a[i], b); }                                               // 1 FMUL2
            if (MODE == 5) {                                                                             // FFMA2 + LDS.64
                a[i] = __ffma2_rn(a[i], b, c);
                if ((i % lds_every) == 0) { float2 v = sm[(threadIdx.x + i * 33 + it) & 1023]; acc.x += v.x; acc.y += v.y; }
            }
            if (MODE == 6) {                                                                             // 2 FFMA + 2 LDS.32
                a[i].x = fmaf(a[i].x, b.x, c.x); a[i].y = fmaf(a[i].y, b.y, c.y);
                if ((i % lds_every) == 0) {
                    const float* s = reinterpret_cast<const float*>(sm);
                    acc.x += s[(threadIdx.x + i * 33 + it) & 2047]; acc.y += s[(threadIdx.x + i * 35 + it + 7) & 2047];
                }
            }
        }
    }
    float s = acc.x + acc.y;
    for (int i = 0; i < ILP; ++i) s += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, int warps_per_sm, int lds_every, double flop_per_inner) {
    float* out;
    int blocks = 148 * warps_per_sm / 8;
    cudaMalloc(&out, blocks * 256 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(out, 1.0001f, lds_every);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, 1.0001f, lds_every);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double lanes = (double)blocks * 256 * ITERS * ILP;          // inner-body executions (thread level)
    printf("%-28s warps/SM %2d  %.3f ms  %.1f G body/s  %.2f TFLOP/s\n", name, warps_per_sm, ms, lanes / ms / 1e6,
           lanes * flop_per_inner / ms / 1e9);
    cudaFree(out);
}

int main() {
    for (int w : {8, 16, 32, 64}) {
        run<0>("2xFFMA", w, 1, 4);
        run<1>("FFMA2", w, 1, 4);
        run<2>("2xFADD", w, 1, 2);
        run<3>("FADD2", w, 1, 2);
        run<4>("FMUL2", w, 1, 2);
        run<5>("FFMA2+LDS64 every 1", w, 1, 4);
        run<5>("FFMA2+LDS64 every 2", w, 2, 4);
        run<6>("2xFFMA+2xLDS32 every 1", w, 1, 4);
        run<6>("2xFFMA+2xLDS32 every 2", w, 2, 4);
    }
    return 0;
}
