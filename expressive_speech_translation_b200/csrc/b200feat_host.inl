// Host-buffer entry points (the end-to-end path): clips are streamed host -> device arena -> kernels ->
// host in chunks, round-robin over NSLOT streams so the H2D copy of chunk i+1, the kernels of chunk i and
// the D2H copy of chunk i-1 overlap (two copy engines + SMs).  Included at the end of b200feat.cu.
namespace {

constexpr int NSLOT = 4;

// One Arena = a device buffer + NSLOT streams, used by ONE host call at a time.  Calls lease an arena from the device's
// pool and give it back when they are done: concurrent calls (Flask runs threaded=True, /root/reference/Backend/app.py:523)
// get different arenas and overlap on the GPU instead of queueing behind one whole-call lock; the pool only grows with
// the number of calls that are really in flight at once.
struct Arena {
    int dev = -1;
    char* base = nullptr;
    size_t bytes = 0;
    cudaStream_t streams[NSLOT] = {nullptr, nullptr, nullptr, nullptr};
    bool streams_ok = false;
};
struct ArenaPool {
    std::mutex mu;                 // held only while an arena is taken from / returned to the lists
    std::vector<Arena*> idle, all;
};

std::mutex g_pools_mu;
std::vector<ArenaPool*> g_pools;   // index = device ordinal

ArenaPool* pool_for(int dev) {
    std::lock_guard<std::mutex> lk(g_pools_mu);
    if ((int)g_pools.size() <= dev) g_pools.resize(dev + 1, nullptr);
    if (!g_pools[dev]) g_pools[dev] = new ArenaPool();
    return g_pools[dev];
}

// the idle arena with the largest buffer, or a new one
Arena* arena_lease(int dev) {
    ArenaPool* pool = pool_for(dev);
    std::lock_guard<std::mutex> lk(pool->mu);
    if (!pool->idle.empty()) {
        size_t best = 0;
        for (size_t i = 1; i < pool->idle.size(); ++i)
            if (pool->idle[i]->bytes > pool->idle[best]->bytes) best = i;
        Arena* a = pool->idle[best];
        pool->idle.erase(pool->idle.begin() + best);
        return a;
    }
    Arena* a = new Arena();
    a->dev = dev;
    pool->all.push_back(a);
    return a;
}
void arena_return(Arena* a) {
    ArenaPool* pool = pool_for(a->dev);
    std::lock_guard<std::mutex> lk(pool->mu);
    pool->idle.push_back(a);
}
struct ArenaLease {
    Arena* a;
    explicit ArenaLease(int dev) : a(arena_lease(dev)) {}
    ~ArenaLease() { arena_return(a); }
};

// the arena is leased: its streams are idle (every call synchronises them before it returns the arena)
int arena_reserve(Arena* a, size_t need) {
    if (!a->streams_ok) {
        for (int i = 0; i < NSLOT; ++i) FE_CUDA(cudaStreamCreateWithFlags(&a->streams[i], cudaStreamNonBlocking));
        a->streams_ok = true;
    }
    if (a->bytes >= need) return 0;
    if (a->base) {
        FE_CUDA(cudaFree(a->base));
        a->base = nullptr;
        a->bytes = 0;
    }
    FE_CUDA(cudaMalloc(reinterpret_cast<void**>(&a->base), need));
    a->bytes = need;
    return 0;
}

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int64_t host_chunk_bytes() {
    static const int64_t v = [] {
        const char* s = getenv("B200FEAT_HOST_CHUNK_MB");
        int64_t mb = s ? atoll(s) : 64;   // measured on B200 / PCIe 5 x16 (tools/e2e_sweep.py): 16 MB 46.8 ms, 64 MB 38.6 ms per
                                          // 1024 x 30 s step; smaller chunks leave the kernels with too few units per launch
        if (mb < 1) mb = 1;
        return mb << 20;
    }();
    return v;
}

// launch(d_wave (void*), d_len_or_null, nclips, d_out (void*), d_ws, ws_bytes, stream) -> rc
template <class Launch>
int host_pipeline(int kind, const void* wave_host_v, int64_t in_esz, const int64_t* lengths_host, int32_t B, int64_t L, int64_t ws_L,
                  int64_t out_per_clip, void* out_host_v, int64_t out_esz, int device, bool single_chunk, Launch launch,
                  int64_t extra_ws_per_clip = 0, int64_t extra_ws_fixed = 0) {
    const char* wave_host = static_cast<const char*>(wave_host_v);
    char* out_host = static_cast<char*>(out_host_v);
    if (B < 0 || L < 0) return fail(B200FEAT_EINVAL, "host call: negative B or L");
    if (B == 0 || out_per_clip == 0) return 0;
    if (!wave_host || !out_host) return fail(B200FEAT_EINVAL, "host call: null wave/out pointer");
    int prev = 0;
    FE_CUDA(cudaGetDevice(&prev));
    if (device < 0) device = prev;
    FE_CUDA(cudaSetDevice(device));
    ArenaLease lease(device);
    Arena* a = lease.a;

    const int64_t in_bytes = L * in_esz, out_bytes = out_per_clip * out_esz;
    int64_t chunk = host_chunk_bytes() / (in_bytes + out_bytes);
    if (chunk < 1) chunk = 1;
    if (chunk > B || single_chunk) chunk = B;
    const int nslot = (chunk >= B) ? 1 : NSLOT;
    const size_t s_wave = round_up((size_t)chunk * in_bytes, 256);
    const size_t s_out = round_up((size_t)chunk * out_bytes, 256);
    const size_t s_len = round_up((size_t)chunk * 8, 256);
    const size_t s_ws = round_up((size_t)(b200feat_workspace_bytes(kind, (int32_t)chunk, ws_L) + chunk * extra_ws_per_clip + extra_ws_fixed), 256);
    const size_t slot_bytes = s_wave + s_out + s_len + s_ws;
    int rc = arena_reserve(a, slot_bytes * nslot);
    if (rc) { cudaSetDevice(prev); return rc; }

    int64_t i = 0;
    for (int64_t c0 = 0; c0 < B; c0 += chunk, ++i) {
        const int64_t nc = (B - c0 < chunk) ? (B - c0) : chunk;
        const int slot = (int)(i % nslot);
        cudaStream_t st = a->streams[slot];
        char* sb = a->base + slot_bytes * slot;
        void* d_wave = sb;
        void* d_out = sb + s_wave;
        int64_t* d_len = reinterpret_cast<int64_t*>(sb + s_wave + s_out);
        void* d_ws = sb + s_wave + s_out + s_len;
        cudaError_t e = cudaMemcpyAsync(d_wave, wave_host + c0 * in_bytes, (size_t)nc * in_bytes, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess && lengths_host)
            e = cudaMemcpyAsync(d_len, lengths_host + c0, (size_t)nc * 8, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) { rc = fail(B200FEAT_ECUDA, "host call: H2D copy failed: %s", cudaGetErrorString(e)); break; }
        rc = launch(d_wave, lengths_host ? d_len : nullptr, (int32_t)nc, d_out, d_ws, (int64_t)s_ws, st);
        if (rc) break;
        e = cudaMemcpyAsync(out_host + c0 * out_bytes, d_out, (size_t)nc * out_bytes, cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) { rc = fail(B200FEAT_ECUDA, "host call: D2H copy failed: %s", cudaGetErrorString(e)); break; }
    }
    for (int s = 0; s < nslot; ++s) {
        cudaError_t e = cudaStreamSynchronize(a->streams[s]);
        if (e != cudaSuccess && rc == 0) rc = fail(B200FEAT_ECUDA, "host call: stream sync failed: %s", cudaGetErrorString(e));
    }
    cudaSetDevice(prev);
    return rc;
}

}  // namespace

extern "C" {

static int whisper_logmel_host_impl(const void* wave_host, bool in_s16, const int64_t* lengths_host, int32_t B, int64_t L,
                                    int32_t n_mels, int32_t padding, int32_t norm_mode, void* out_host, bool out16,
                                    int32_t device) {
    if (n_mels != 80 && n_mels != 128) return fail(B200FEAT_EINVAL, "whisper_logmel_host: Unsupported n_mels: %d", n_mels);
    if (padding < 0) return fail(B200FEAT_EINVAL, "whisper_logmel_host: padding < 0");
    const int64_t T = (L + padding) / FeW2Cfg::HOP;
    return host_pipeline(B200FEAT_KIND_WHISPER, wave_host, in_s16 ? 2 : 4, lengths_host, B, L, L + padding, (int64_t)n_mels * T,
                         out_host, out16 ? 2 : 4, device, norm_mode == FE_NORM_PER_CALL,
                         [=](const void* dw, const int64_t* dl, int32_t nc, void* dout, void* ws, int64_t wsb, cudaStream_t st) {
                             return whisper_logmel_impl(dw, in_s16, dl, nc, L, L, n_mels, padding, norm_mode, dout, out16,
                                                        (int64_t)n_mels * T, T, ws, wsb, st);
                         });
}

int b200feat_whisper_logmel_host(const float* wave_host, const int64_t* lengths_host, int32_t B, int64_t L,
                                 int32_t n_mels, int32_t padding, int32_t norm_mode, float* out_host,
                                 int32_t device) {
    return whisper_logmel_host_impl(wave_host, false, lengths_host, B, L, n_mels, padding, norm_mode, out_host, false, device);
}

int b200feat_whisper_logmel_ex_host(const void* wave_host, int32_t wave_dtype, const int64_t* lengths_host, int32_t B,
                                    int64_t L, int32_t n_mels, int32_t padding, int32_t norm_mode, void* out_host,
                                    int32_t out_dtype, int32_t device) {
    if (int rc = whisper_ex_dtypes("whisper_logmel_ex_host", wave_dtype, out_dtype)) return rc;
    return whisper_logmel_host_impl(wave_host, wave_dtype == B200FEAT_S16, lengths_host, B, L, n_mels, padding, norm_mode,
                                    out_host, out_dtype == B200FEAT_F16, device);
}

// PCM in -> log-mel out through the chunked pipeline: the H2D copy carries int16 PCM at the capture rate (44.1 kHz stereo:
// 5.3 KB per audio-ms ... 176 KB per audio-second), the resampled float32 waveform lives only in the arena.
int b200feat_pcm_whisper_logmel_host(const void* pcm_host, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                                     const int64_t* lengths_host, int32_t B, int64_t L_in, int32_t orig_freq, int32_t new_freq,
                                     int32_t lowpass_filter_width, double rolloff, int32_t method, double beta,
                                     int32_t n_mels, int32_t padding, int32_t norm_mode, void* out_host, int32_t out_dtype,
                                     int32_t device) {
    if (int rc = whisper_ex_dtypes("pcm_whisper_logmel_host", B200FEAT_F32, out_dtype)) return rc;
    if (pcm_dtype != B200FEAT_F32 && pcm_dtype != B200FEAT_S16)
        return fail(B200FEAT_EINVAL, "pcm_whisper_logmel_host: pcm_dtype must be B200FEAT_F32 or B200FEAT_S16, got %d", pcm_dtype);
    if (channels != 1 && channels != 2) return fail(B200FEAT_EINVAL, "pcm_whisper_logmel_host: channels must be 1 or 2, got %d", channels);
    if (n_mels != 80 && n_mels != 128) return fail(B200FEAT_EINVAL, "pcm_whisper_logmel_host: Unsupported n_mels: %d", n_mels);
    if (padding < 0 || orig_freq <= 0 || new_freq <= 0) return fail(B200FEAT_EINVAL, "pcm_whisper_logmel_host: negative padding or frequency");
    const int64_t L = fe_rs_out_len(L_in, orig_freq, new_freq), Ls = (L + 3) / 4 * 4;
    const int64_t T = (L + padding) / FeW2Cfg::HOP;
    const int64_t frame_bytes = (int64_t)channels * (pcm_dtype == B200FEAT_S16 ? 2 : 4);
    return host_pipeline(B200FEAT_KIND_WHISPER, pcm_host, frame_bytes, lengths_host, B, L_in, L + padding, (int64_t)n_mels * T,
                         out_host, out_dtype == B200FEAT_F16 ? 2 : 4, device, norm_mode == FE_NORM_PER_CALL,
                         [=](const void* dw, const int64_t* dl, int32_t nc, void* dout, void* ws, int64_t wsb, cudaStream_t st) {
                             return b200feat_pcm_whisper_logmel(dw, pcm_dtype, channels, mix_host, dl, nc, L_in, L_in * channels,
                                                                orig_freq, new_freq, lowpass_filter_width, rolloff, method, beta,
                                                                n_mels, padding, norm_mode, dout, out_dtype, (int64_t)n_mels * T, T,
                                                                ws, wsb, st);
                         },
                         Ls * 4 + 8, 1024);
}

// PCM in -> resampled float32 waveform out through the chunked pipeline (the host-buffer twin of b200feat_resample)
int b200feat_resample_host(const void* pcm_host, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                           const int64_t* lengths_host, int32_t B, int64_t L_in, int32_t orig_freq, int32_t new_freq,
                           int32_t lowpass_filter_width, double rolloff, int32_t method, double beta, float* out_host, int32_t device) {
    if (pcm_dtype != B200FEAT_F32 && pcm_dtype != B200FEAT_S16)
        return fail(B200FEAT_EINVAL, "resample_host: pcm_dtype must be B200FEAT_F32 or B200FEAT_S16, got %d", pcm_dtype);
    if (channels != 1 && channels != 2) return fail(B200FEAT_EINVAL, "resample_host: channels must be 1 or 2, got %d", channels);
    if (orig_freq <= 0 || new_freq <= 0) return fail(B200FEAT_EINVAL, "resample_host: Original frequency and desired frequecy should be positive");
    const int64_t L = fe_rs_out_len(L_in, orig_freq, new_freq);
    const int64_t frame_bytes = (int64_t)channels * (pcm_dtype == B200FEAT_S16 ? 2 : 4);
    return host_pipeline(B200FEAT_KIND_COSY /* 16-byte workspace */, pcm_host, frame_bytes, lengths_host, B, L_in, L_in, L, out_host, 4, device, false,
                         [=](const void* dw, const int64_t* dl, int32_t nc, void* dout, void*, int64_t, cudaStream_t st) {
                             return b200feat_resample(dw, pcm_dtype, channels, mix_host, dl, nc, L_in, L_in * channels, orig_freq, new_freq,
                                                      lowpass_filter_width, rolloff, method, beta, static_cast<float*>(dout), L, L, st);
                         });
}

int b200feat_cosy_mel_host(const float* wave_host, const int64_t* lengths_host, int32_t B, int64_t L,
                           float* out_host, int32_t device) {
    const int64_t T = b200feat_num_frames(B200FEAT_KIND_COSY, L, 0);
    return host_pipeline(B200FEAT_KIND_COSY, wave_host, 4, lengths_host, B, L, L, (int64_t)FeCCfg::NMEL * T, out_host, 4, device, false,
                         [=](const void* dw, const int64_t* dl, int32_t nc, void* dout, void* ws, int64_t wsb, cudaStream_t st) {
                             return b200feat_cosy_mel(static_cast<const float*>(dw), dl, nc, L, L, static_cast<float*>(dout), (int64_t)FeCCfg::NMEL * T, T, ws, wsb, st);
                         });
}

int b200feat_kaldi_fbank_host(const float* wave_host, const int64_t* lengths_host, int32_t B, int64_t L,
                              int32_t subtract_mean, float* out_host, int32_t device) {
    const int64_t T = b200feat_num_frames(B200FEAT_KIND_KALDI, L, 0);
    return host_pipeline(B200FEAT_KIND_KALDI, wave_host, 4, lengths_host, B, L, L, (int64_t)FeKCfg::NMEL * T, out_host, 4, device, false,
                         [=](const void* dw, const int64_t* dl, int32_t nc, void* dout, void* ws, int64_t wsb, cudaStream_t st) {
                             return b200feat_kaldi_fbank(static_cast<const float*>(dw), dl, nc, L, L, subtract_mean, static_cast<float*>(dout), (int64_t)FeKCfg::NMEL * T, T, ws, wsb, st);
                         });
}

void b200feat_release(void) {      // frees every IDLE arena (arenas leased by calls in flight are left alone)
    std::lock_guard<std::mutex> lk(g_pools_mu);
    for (ArenaPool* pool : g_pools) {
        if (!pool) continue;
        std::lock_guard<std::mutex> lk2(pool->mu);
        int prev = 0;
        if (cudaGetDevice(&prev) != cudaSuccess) continue;
        for (Arena* a : pool->idle) {
            cudaSetDevice(a->dev);
            if (a->base) { cudaFree(a->base); a->base = nullptr; a->bytes = 0; }
            if (a->streams_ok) {
                for (int i = 0; i < NSLOT; ++i) cudaStreamDestroy(a->streams[i]);
                a->streams_ok = false;
            }
        }
        cudaSetDevice(prev);
    }
}

}  // extern "C"
