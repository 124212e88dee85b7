// libb200feat.so -- kernels, launch logic and the C ABI declared in include/b200feat.h.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared
// (see expressive_speech_translation_b200/build.py).  sm_100a only; no other architecture, no CPU path.
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200feat.h"
#include "fe_common.h"
#include "generated/fe_codelets.h"
#include "generated/fe_tables.h"
#include "fe_whisper.cuh"
#include "fe_kaldi.cuh"
#include "fe_cosy.cuh"

// =====================================================================================================
// kernels
// =====================================================================================================
template <int NMELS>
__global__ void __launch_bounds__(FeWCfg::NT, 2) fe_whisper_kernel(const FeWhisperParams p) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x, cta = blockIdx.x;
    fe_w_phase_load(p, sm, tid, cta);
    __syncthreads();
    fe_w_phase_a(p, sm, tid, cta);
    __syncthreads();
    fe_w_phase_b(p, sm, tid, cta);
    __syncthreads();
    int key = fe_w_phase_mel<NMELS>(p, sm, tid, cta);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) key = max(key, __shfl_xor_sync(0xffffffffu, key, o));
    int* smax = reinterpret_cast<int*>(sm + FeWCfg::OFF_MISC);
    if ((tid & 31) == 0) atomicMax(smax, key);
    __syncthreads();
    if (tid == 0) {
        const int b = p.clip0 + cta / p.tiles_per_clip;
        atomicMax(p.max_keys + (p.norm_mode == FE_NORM_PER_CALL ? 0 : b), *smax);
    }
}

// one CTA per (clip, mel row): out = (max(x, gmax - 8) + 4) / 4, zeros past the clip's own frames
__global__ void __launch_bounds__(256) fe_whisper_norm_kernel(const FeWhisperParams p) {
    const int row = blockIdx.x;
    const int b = p.clip0 + row / p.n_mels;
    const int m = row % p.n_mels;
    const float gmax = fe_key_float(p.max_keys[p.norm_mode == FE_NORM_PER_CALL ? 0 : b]);
    const int64_t len = fe_clip_len(p.lengths, b, p.L);
    int64_t T = (len + p.padding) / FeWCfg::HOP;
    if (T > p.T_out) T = p.T_out;
    float* __restrict__ o = p.out + (int64_t)b * p.out_batch_stride + (int64_t)m * p.T_out;
    if (p.vec_ok & 2) {
        float4* o4 = reinterpret_cast<float4*>(o);
        const int64_t n4 = p.T_out >> 2;
        for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
            const int64_t t = i << 2;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (t < T) {
                v = __ldcg(o4 + i);  // written by the previous kernel: served from L2
                v.x = fe_w_normalise(v.x, gmax);
                v.y = (t + 1 < T) ? fe_w_normalise(v.y, gmax) : 0.f;
                v.z = (t + 2 < T) ? fe_w_normalise(v.z, gmax) : 0.f;
                v.w = (t + 3 < T) ? fe_w_normalise(v.w, gmax) : 0.f;
            }
            o4[i] = v;
        }
    } else {
        for (int64_t t = threadIdx.x; t < p.T_out; t += blockDim.x)
            o[t] = (t < T) ? fe_w_normalise(__ldcg(o + t), gmax) : 0.f;
    }
}

__global__ void __launch_bounds__(FeKCfg::NT, 2) fe_kaldi_kernel(const FeKaldiParams p) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x, cta = blockIdx.x;
    fe_k_phase_load(p, sm, tid, cta);
    __syncthreads();
    fe_k_phase_sum(p, sm, tid, cta);
    __syncthreads();
    fe_k_phase_a(p, sm, tid, cta);
    __syncthreads();
    fe_k_phase_b(p, sm, tid, cta);
    __syncthreads();
    fe_k_phase_mel(p, sm, tid, cta);
    __syncthreads();
    fe_k_phase_store(p, sm, tid, cta);
}

// CMN: one CTA per (clip, slab of 64 frames); 80 threads rebuild the clip mean from the per-tile sums
__global__ void __launch_bounds__(256) fe_kaldi_cmn_kernel(const FeKaldiParams p, int slabs_per_clip) {
    __shared__ float mean[FeKCfg::NMEL];
    const int b = p.clip0 + blockIdx.x / slabs_per_clip;
    const int slab = blockIdx.x % slabs_per_clip;
    const int64_t len = fe_clip_len(p.lengths, b, p.L);
    int64_t T = len < FeKCfg::WIN ? 0 : 1 + (len - FeKCfg::WIN) / FeKCfg::HOP;
    if (T > p.T_out) T = p.T_out;
    const int64_t r0 = (int64_t)slab * 64;
    if (r0 >= T) return;
    if (threadIdx.x < FeKCfg::NMEL) mean[threadIdx.x] = fe_k_clip_mean(p, b, threadIdx.x, (int)T);
    __syncthreads();
    int64_t r1 = r0 + 64;
    if (r1 > T) r1 = T;
    float* __restrict__ o = p.out + (int64_t)b * p.out_batch_stride + r0 * FeKCfg::NMEL;
    const int total = (int)(r1 - r0) * FeKCfg::NMEL;
    if (p.vec_ok & 2) {
        for (int i = threadIdx.x * 4; i < total; i += blockDim.x * 4) {
            const int m = i % FeKCfg::NMEL;
            float4 v = __ldcg(reinterpret_cast<const float4*>(o + i));
            v.x -= mean[m]; v.y -= mean[m + 1]; v.z -= mean[m + 2]; v.w -= mean[m + 3];
            *reinterpret_cast<float4*>(o + i) = v;
        }
    } else {
        for (int i = threadIdx.x; i < total; i += blockDim.x) o[i] = __ldcg(o + i) - mean[i % FeKCfg::NMEL];
    }
}

__global__ void __launch_bounds__(FeCCfg::NT, 1) fe_cosy_kernel(const FeCosyParams p) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x, cta = blockIdx.x;
    fe_c_phase_load(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_a(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_b(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_c(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_mel(p, sm, tid, cta);
}

// =====================================================================================================
// host side
// =====================================================================================================
namespace {

thread_local std::string g_err;
std::atomic<int64_t> g_launches{0};

// optional per-launch CUDA-event timing (bench.py's roofline leg): events are recorded on the launching
// stream right before and after each kernel; b200feat_profile_end() synchronises and sums them.
struct ProfRec { cudaEvent_t a, b; int kind; };   // kind 0 = featurizer kernel, 1 = epilogue (norm / CMN)
std::atomic<int> g_prof_on{0};
std::mutex g_prof_mu;
std::vector<ProfRec> g_prof;

struct ProfScope {
    cudaStream_t st; int kind; cudaEvent_t a = nullptr, b = nullptr; bool on;
    ProfScope(cudaStream_t s, int k) : st(s), kind(k), on(g_prof_on.load() != 0) {
        if (on && cudaEventCreate(&a) == cudaSuccess && cudaEventCreate(&b) == cudaSuccess) cudaEventRecord(a, st);
        else on = false;
    }
    ~ProfScope() {
        if (!on) return;
        cudaEventRecord(b, st);
        std::lock_guard<std::mutex> lk(g_prof_mu);
        g_prof.push_back({a, b, kind});
    }
};

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define FE_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(B200FEAT_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_),    \
                        __FILE__, __LINE__);                                                       \
    } while (0)

int64_t l2_wave_bytes() {
    static const int64_t v = [] {
        const char* s = getenv("B200FEAT_L2_WAVE_MB");
        int64_t mb = s ? atoll(s) : 48;
        if (mb < 1) mb = 1;
        return mb << 20;
    }();
    return v;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int check_device_arch() {
    // one query per device per process; the library carries sm_100a SASS only
    static std::mutex mu;
    static std::vector<int> state;  // 0 unknown, 1 ok, -1 bad
    int dev = 0;
    FE_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    if ((int)state.size() <= dev) state.resize(dev + 1, 0);
    if (state[dev] == 0) {
        int major = 0;
        FE_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        state[dev] = (major == 10) ? 1 : -1;
        if (state[dev] == 1) {
            FE_CUDA(cudaFuncSetAttribute(fe_whisper_kernel<80>, cudaFuncAttributeMaxDynamicSharedMemorySize, FeWCfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_whisper_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, FeWCfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_kaldi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeKCfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_cosy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeCCfg::SMEM_BYTES));
        }
    }
    if (state[dev] != 1) return fail(B200FEAT_EARCH, "device %d is not compute capability 10.x (sm_100a build)", dev);
    return 0;
}

int common_checks(const char* who, const void* wave, const void* out, int32_t B, int64_t L, int64_t wave_stride) {
    if (B < 0) return fail(B200FEAT_EINVAL, "%s: B = %d < 0", who, B);
    if (B > 0 && (wave == nullptr || out == nullptr)) return fail(B200FEAT_EINVAL, "%s: null wave/out pointer", who);
    if (L < 0 || L > (int64_t)1 << 40) return fail(B200FEAT_EINVAL, "%s: bad L = %lld", who, (long long)L);
    if (wave_stride < L) return fail(B200FEAT_EINVAL, "%s: wave_stride %lld < L %lld", who, (long long)wave_stride, (long long)L);
    return 0;
}

}  // namespace

extern "C" {

int b200feat_version(void) { return B200FEAT_VERSION; }
const char* b200feat_last_error(void) { return g_err.c_str(); }
int64_t b200feat_launch_count(void) { return g_launches.load(); }

int b200feat_profile_begin(void) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (auto& r : g_prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    g_prof.clear();
    g_prof_on.store(1);
    return 0;
}

int b200feat_profile_end(double* feature_ms, int64_t* feature_launches, double* epilogue_ms, int64_t* epilogue_launches) {
    g_prof_on.store(0);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    double ms[2] = {0.0, 0.0};
    int64_t n[2] = {0, 0};
    int rc = 0;
    for (auto& r : g_prof) {
        float t = 0.f;
        cudaError_t e = cudaEventSynchronize(r.b);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, r.a, r.b);
        if (e != cudaSuccess && rc == 0) rc = fail(B200FEAT_ECUDA, "profile_end: %s", cudaGetErrorString(e));
        ms[r.kind] += t;
        n[r.kind] += 1;
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    g_prof.clear();
    if (feature_ms) *feature_ms = ms[0];
    if (feature_launches) *feature_launches = n[0];
    if (epilogue_ms) *epilogue_ms = ms[1];
    if (epilogue_launches) *epilogue_launches = n[1];
    return rc;
}

int64_t b200feat_num_frames(int32_t kind, int64_t n, int32_t padding) {
    switch (kind) {
    case B200FEAT_KIND_WHISPER: return (n + padding) / FeWCfg::HOP;
    case B200FEAT_KIND_COSY: return (n + 2 * FeCCfg::PADL < FeCCfg::N) ? 0 : 1 + (n + 2 * FeCCfg::PADL - FeCCfg::N) / FeCCfg::HOP;
    case B200FEAT_KIND_KALDI: return n < FeKCfg::WIN ? 0 : 1 + (n - FeKCfg::WIN) / FeKCfg::HOP;
    default: return -1;
    }
}

int64_t b200feat_workspace_bytes(int32_t kind, int32_t B, int64_t L) {
    if (B < 0 || L < 0) return -1;
    switch (kind) {
    case B200FEAT_KIND_WHISPER: return 16 + (int64_t)B * 4;
    case B200FEAT_KIND_COSY: return 16;
    case B200FEAT_KIND_KALDI: {
        const int64_t T = b200feat_num_frames(kind, L, 0);
        const int64_t tiles = (T + FeKCfg::FT - 1) / FeKCfg::FT;
        return 16 + (int64_t)B * tiles * FeKCfg::NMEL * 4;
    }
    default: return -1;
    }
}

// ------------------------------------------------------------------------------------------ Whisper
int b200feat_whisper_logmel(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                            int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                            float* out, int64_t out_batch_stride, int64_t T_out,
                            void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    if (int rc = common_checks("whisper_logmel", wave, out, B, L, wave_stride)) return rc;
    if (n_mels != 80 && n_mels != 128) return fail(B200FEAT_EINVAL, "whisper_logmel: Unsupported n_mels: %d", n_mels);
    if (padding < 0) return fail(B200FEAT_EINVAL, "whisper_logmel: padding %d < 0", padding);
    if (norm_mode != FE_NORM_PER_CLIP && norm_mode != FE_NORM_PER_CALL) return fail(B200FEAT_EINVAL, "whisper_logmel: bad norm_mode %d", norm_mode);
    const int64_t T_full = (L + padding) / FeWCfg::HOP;
    if (T_out < T_full) return fail(B200FEAT_EINVAL, "whisper_logmel: T_out %lld < %lld frames", (long long)T_out, (long long)T_full);
    if (out_batch_stride < (int64_t)n_mels * T_out) return fail(B200FEAT_EINVAL, "whisper_logmel: out_batch_stride too small");
    if (L + padding <= FeWCfg::PADL && B > 0 && T_out > 0)
        return fail(B200FEAT_EINVAL, "whisper_logmel: reflect padding (200) needs more than 200 samples, got %lld", (long long)(L + padding));
    if (workspace_bytes < b200feat_workspace_bytes(B200FEAT_KIND_WHISPER, B, L) || workspace == nullptr)
        return fail(B200FEAT_EWORKSPACE, "whisper_logmel: workspace too small (%lld bytes)", (long long)workspace_bytes);
    if (B == 0 || T_out == 0) return 0;
    if (int rc = check_device_arch()) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

    FeWhisperParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = wave_stride; p.L = L; p.B = B;
    p.padding = padding; p.n_mels = n_mels; p.norm_mode = norm_mode;
    p.out = out; p.out_batch_stride = out_batch_stride; p.T_out = T_out;
    const int64_t tiles = (T_out + FeWCfg::FT - 1) / FeWCfg::FT;
    if (tiles > 0x7fffffff / 2) return fail(B200FEAT_EINVAL, "whisper_logmel: clip too long");
    p.tiles_per_clip = (int32_t)tiles;
    p.vec_ok = (aligned16(wave) && wave_stride % 4 == 0) ? 1 : 0;
    if (aligned16(out) && out_batch_stride % 4 == 0 && T_out % 4 == 0) p.vec_ok |= 2;
    p.max_keys = reinterpret_cast<int32_t*>(workspace);
    FE_CUDA(cudaMemsetAsync(workspace, FE_KEY_INIT_BYTE, 16 + (size_t)B * 4, st));

    // clips go through in waves small enough for the un-normalised log-mel to stay in L2 until the
    // normalisation kernel of the same wave has consumed it
    const int64_t clip_bytes = (int64_t)n_mels * T_out * 4;
    int64_t per_wave = l2_wave_bytes() / clip_bytes;
    if (per_wave < 1) per_wave = 1;
    const int64_t max_ctas = 0x7fffffff;
    if (per_wave * tiles > max_ctas) per_wave = max_ctas / tiles;
    if (per_wave * n_mels > max_ctas) per_wave = max_ctas / n_mels;
    for (int pass = 0; pass < 2; ++pass) {
        // per-clip: feature+norm interleaved per wave (single pass); per-call: all features, then all norms
        for (int64_t c0 = 0; c0 < B; c0 += per_wave) {
            const int64_t nc = (B - c0 < per_wave) ? (B - c0) : per_wave;
            p.clip0 = (int32_t)c0; p.nclips = (int32_t)nc;
            const bool do_feat = (norm_mode == FE_NORM_PER_CLIP) ? (pass == 0) : (pass == 0);
            const bool do_norm = (norm_mode == FE_NORM_PER_CLIP) ? (pass == 0) : (pass == 1);
            if (do_feat) {
                const unsigned grid = (unsigned)(nc * tiles);
                ProfScope ps(st, 0);
                if (n_mels == 80) fe_whisper_kernel<80><<<grid, FeWCfg::NT, FeWCfg::SMEM_BYTES, st>>>(p);
                else fe_whisper_kernel<128><<<grid, FeWCfg::NT, FeWCfg::SMEM_BYTES, st>>>(p);
                g_launches++;
            }
            if (do_norm) {
                ProfScope ps(st, 1);
                fe_whisper_norm_kernel<<<(unsigned)(nc * n_mels), 256, 0, st>>>(p);
                g_launches++;
            }
        }
        if (norm_mode == FE_NORM_PER_CLIP) break;
    }
    FE_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------ CosyVoice
int b200feat_cosy_mel(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                      int64_t wave_stride, float* out, int64_t out_batch_stride, int64_t T_out,
                      void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    (void)workspace; (void)workspace_bytes;
    if (int rc = common_checks("cosy_mel", wave, out, B, L, wave_stride)) return rc;
    const int64_t T_full = b200feat_num_frames(B200FEAT_KIND_COSY, L, 0);
    if (T_out < T_full) return fail(B200FEAT_EINVAL, "cosy_mel: T_out %lld < %lld frames", (long long)T_out, (long long)T_full);
    if (out_batch_stride < (int64_t)FeCCfg::NMEL * T_out) return fail(B200FEAT_EINVAL, "cosy_mel: out_batch_stride too small");
    if (L <= FeCCfg::PADL && B > 0 && T_out > 0)
        return fail(B200FEAT_EINVAL, "cosy_mel: reflect padding (720) needs more than 720 samples, got %lld", (long long)L);
    if (B == 0 || T_out == 0) return 0;
    if (int rc = check_device_arch()) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    FeCosyParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = wave_stride; p.L = L; p.B = B;
    p.out = out; p.out_batch_stride = out_batch_stride; p.T_out = T_out;
    const int64_t tiles = (T_out + FeCCfg::FT - 1) / FeCCfg::FT;
    p.tiles_per_clip = (int32_t)tiles;
    p.vec_ok = (aligned16(wave) && wave_stride % 4 == 0) ? 1 : 0;
    int64_t per_launch = 0x7fffffff / tiles;
    if (per_launch < 1) return fail(B200FEAT_EINVAL, "cosy_mel: clip too long");
    for (int64_t c0 = 0; c0 < B; c0 += per_launch) {
        const int64_t nc = (B - c0 < per_launch) ? (B - c0) : per_launch;
        p.clip0 = (int32_t)c0; p.nclips = (int32_t)nc;
        ProfScope ps(st, 0);
        fe_cosy_kernel<<<(unsigned)(nc * tiles), FeCCfg::NT, FeCCfg::SMEM_BYTES, st>>>(p);
        g_launches++;
    }
    FE_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------ kaldi
int b200feat_kaldi_fbank(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                         int64_t wave_stride, int32_t subtract_mean, float* out,
                         int64_t out_batch_stride, int64_t T_out,
                         void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    if (int rc = common_checks("kaldi_fbank", wave, out, B, L, wave_stride)) return rc;
    const int64_t T_full = b200feat_num_frames(B200FEAT_KIND_KALDI, L, 0);
    if (T_out < T_full) return fail(B200FEAT_EINVAL, "kaldi_fbank: T_out %lld < %lld frames", (long long)T_out, (long long)T_full);
    if (out_batch_stride < (int64_t)FeKCfg::NMEL * T_out) return fail(B200FEAT_EINVAL, "kaldi_fbank: out_batch_stride too small");
    const int64_t tiles = (T_out + FeKCfg::FT - 1) / FeKCfg::FT;
    if (subtract_mean && (workspace == nullptr || workspace_bytes < 16 + (int64_t)B * tiles * FeKCfg::NMEL * 4))
        return fail(B200FEAT_EWORKSPACE, "kaldi_fbank: workspace too small (%lld bytes)", (long long)workspace_bytes);
    if (B == 0 || T_out == 0) return 0;
    if (int rc = check_device_arch()) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    FeKaldiParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = wave_stride; p.L = L; p.B = B;
    p.subtract_mean = subtract_mean ? 1 : 0;
    p.out = out; p.out_batch_stride = out_batch_stride; p.T_out = T_out;
    p.tiles_per_clip = (int32_t)tiles;
    p.vec_ok = (aligned16(wave) && wave_stride % 4 == 0) ? 1 : 0;
    if (aligned16(out) && out_batch_stride % 4 == 0) p.vec_ok |= 2;
    p.colsum = workspace ? reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + 16) : nullptr;
    const int64_t clip_bytes = (int64_t)FeKCfg::NMEL * T_out * 4;
    int64_t per_wave = l2_wave_bytes() / clip_bytes;
    if (per_wave < 1) per_wave = 1;
    const int64_t slabs = (T_out + 63) / 64;
    if (per_wave * tiles > 0x7fffffff) per_wave = 0x7fffffff / tiles;
    if (per_wave < 1) return fail(B200FEAT_EINVAL, "kaldi_fbank: clip too long");
    for (int64_t c0 = 0; c0 < B; c0 += per_wave) {
        const int64_t nc = (B - c0 < per_wave) ? (B - c0) : per_wave;
        p.clip0 = (int32_t)c0; p.nclips = (int32_t)nc;
        {
            ProfScope ps(st, 0);
            fe_kaldi_kernel<<<(unsigned)(nc * tiles), FeKCfg::NT, FeKCfg::SMEM_BYTES, st>>>(p);
            g_launches++;
        }
        if (subtract_mean) {
            ProfScope ps(st, 1);
            fe_kaldi_cmn_kernel<<<(unsigned)(nc * slabs), 256, 0, st>>>(p, (int)slabs);
            g_launches++;
        }
    }
    FE_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

#include "b200feat_host.inl"
