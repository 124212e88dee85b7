// libb200feat.so -- kernels, launch logic and the C ABI declared in include/b200feat.h.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared
// (see expressive_speech_translation_b200/build.py).  sm_100a only; no other architecture, no CPU path.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
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
#include "fe_whisper2.cuh"
#include "fe_whisper3.cuh"
#include "fe_kaldi.cuh"
#include "fe_kaldi3.cuh"
#include "fe_resample_tc.cuh"
#include "fe_cosy.cuh"
#include "fe_gate.cuh"
#include "fe_resample.cuh"
#define FE_RS_HAVE_PARAMS
#include "fe_resample_plan.h"
#include <map>
#include <tuple>

// =====================================================================================================
// kernels
// =====================================================================================================
// named barriers (id 0 is __syncthreads)
enum { FE_BAR_W_FULL = 1, FE_BAR_W_FREE = 2, FE_BAR_COMPUTE = 3 };
#ifdef FE_W2_NOSYNC   // timing experiment only (free-running warps, garbage results): tools/trace_w2.py
static __device__ __forceinline__ void fe_bar_sync(int, int) { asm volatile("" ::: "memory"); }
static __device__ __forceinline__ void fe_bar_arrive(int, int) { asm volatile("" ::: "memory"); }
#else
static __device__ __forceinline__ void fe_bar_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
static __device__ __forceinline__ void fe_bar_arrive(int id, int count) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
#endif

#ifdef FE_W2_TRACE
// debug build only (tools/trace_w2.py): clock64 stamps of CTA 0's warps at the phase boundaries
#define FE_TR_UNITS 12
#define FE_TR_EV 8
__device__ long long fe_w2_trace_buf[FE_TR_UNITS * 20 * FE_TR_EV];
#define FE_TR(iter, ev) do { if (blockIdx.x == 0 && (unsigned)(iter) < FE_TR_UNITS && lane == 0) \
    fe_w2_trace_buf[((iter) * 20 + warp) * FE_TR_EV + (ev)] = clock64(); } while (0)
#else
#define FE_TR(iter, ev) do { } while (0)
#endif

// mbarriers in shared memory (32-bit shared-window addresses)
static __device__ __forceinline__ void fe_mbar_init(unsigned addr, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count) : "memory");
}
static __device__ __forceinline__ void fe_mbar_arrive(unsigned addr) {   // release.cta: orders the thread's prior accesses
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(addr) : "memory");
}
static __device__ __forceinline__ void fe_mbar_wait(unsigned addr, unsigned parity) {
#ifdef FE_W2_NOSYNC
    asm volatile("" ::: "memory");
    return;
#endif
    asm volatile(
        "{\n\t.reg .pred p;\n"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n"
        "DONE_%=:\n\t}" ::"r"(addr), "r"(parity) : "memory");
}
// one arrival per warp, after every lane's prior shared-memory accesses
struct FeWarpArrive {
    unsigned addr; int lane;
    __device__ __forceinline__ void operator()() const {
        __syncwarp();
        if (lane == 0) fe_mbar_arrive(addr);
    }
};
struct FeMbarWait {
    unsigned addr, parity;
    __device__ __forceinline__ void operator()() const { fe_mbar_wait(addr, parity); }
};

// Persistent, warp-specialised: NLW loader warps run one unit ahead of NCW compute warps.
//   loader : [wait W_FREE] fill W(u), publish the unit's (clip, first frame, frame count) -> arrive W_FULL
//   compute: ONE CTA-wide barrier per unit.  Between two barriers every compute warp runs, in its own order,
//            its task of phase B(u), its tasks of phase A(u+1) and its chunk of phase M(u); two mbarriers carry the
//            dependencies inside the interval ("B(u) has read the A planes" before A(u+1) stores into them,
//            "P(u) complete" before M(u) reads it).  Warps of one SM sub-partition are in different phases at any
//            time, so the shared-memory bursts of one overlap the FP32 work of another.
template <int NMELS, bool OUT16, bool IN16>
__global__ void __launch_bounds__(FeW2Cfg::NT, 1) fe_whisper2_kernel(const FeWhisper2Params p) {
    typedef FeW2Cfg C;
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler
    int* misc = reinterpret_cast<int*>(sm + C::OFF_MISC);     // [0] max key, [1] key of -min, [4 + 4*par ..] unit info
    const unsigned mb_bl = (unsigned)__cvta_generic_to_shared(sm + C::OFF_MBAR);   // B(u) loads done
    const unsigned mb_pb = mb_bl + 8;                                               // P(u) complete
    if (tid == 0) {
        misc[0] = FE_KEY_INIT; misc[1] = FE_KEY_INIT;
        fe_mbar_init(mb_bl, C::NCW);
        fe_mbar_init(mb_pb, C::NCW);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    fe_w2_copy_tables<NMELS>(sm, tid, C::NT);
    __syncthreads();
    if (warp >= C::NCW) {
        const int ltid = tid - C::NCT;
        int par = 0;
        int it = 0;
        FeW2Stage stage;
        FeW2Unit t = fe_w2_unit(p, blockIdx.x);
        int mode = fe_w2_load_mode<IN16>(p, t);
        fe_w2_load_ahead<IN16>(p, t, ltid, mode, stage);
        for (int u = blockIdx.x; u < p.total_units; u += gridDim.x, par ^= 1, ++it) {
            FE_TR(it, 0);
            if (u != (int)blockIdx.x) fe_bar_sync(FE_BAR_W_FREE, C::NT);
            FE_TR(it, 1);
            fe_w2_load_finish<IN16>(p, t, sm, ltid, mode, stage);
            if (ltid == 0) { misc[4 + 4 * par] = t.b; misc[5 + 4 * par] = t.t0; misc[6 + 4 * par] = t.T; }
            FE_TR(it, 2);
            fe_bar_arrive(FE_BAR_W_FULL, C::NT);
            const int un = u + (int)gridDim.x;
            if (un < p.total_units) {   // next unit: loads go in flight now, stores wait for W_FREE
                t = fe_w2_unit(p, un);
                mode = fe_w2_load_mode<IN16>(p, t);
                fe_w2_load_ahead<IN16>(p, t, ltid, mode, stage);
            }
            FE_TR(it, 3);
        }
        return;
    }
    // ---- compute warps
    const int u0 = blockIdx.x, du = gridDim.x;
    const int n2b = fe_w2_a_second(warp);
    // Order of the steps after B and the first A task, by warp: warps 4-7 run M(u) before their second A(u+1) task,
    // the others after it.  Every SM sub-partition then holds warps in different phases, so the shared-memory
    // bursts of one overlap the FP32 work of another and the latency-bound mel loop fills the gaps.  (M before
    // the first A task was measured slower: it delays the release of W and with it the loader.)
    const bool m_before_a2 = (warp & 4) != 0;
    FeW2Unit cur;
    cur.b = 0; cur.t0 = 0; cur.T = 0; cur.len = 0; cur.lp = 0;
    // iteration -1 is the prologue: only the A(u0) tasks run (same code as in the steady state)
    int it = -1;
    for (int u = u0 - du; u < p.total_units; u += du, ++it) {
        FE_TR(it, 0);
        const bool first = it < 0;
        const bool live = cur.t0 < cur.T;
        const int un = u + du;
        const bool has_next = un < p.total_units;
        const unsigned parity = (unsigned)(it & 1);                // -1 -> 1: passes on the fresh mbarrier
        FeW2Unit nxt = cur;
        bool live_n = false;
        fe2 mm = fe_mk2(3.0e38f, -3.0e38f);
        // ---- B(u)
        if (!first) {
            const FeWarpArrive loads_done{mb_bl, lane};
            if (live) fe_w2_phase_b(sm, lane, warp, loads_done);
            else loads_done();
            const FeWarpArrive p_done{mb_pb, lane};
            p_done();
        }
        FE_TR(it, 1);
        // ---- first A(u+1) task
        if (has_next) {
            fe_bar_sync(FE_BAR_W_FULL, C::NT);                     // W(next) and its descriptor are in place
            const int np = (it + 1) & 1;
            nxt.b = misc[4 + 4 * np]; nxt.t0 = misc[5 + 4 * np]; nxt.T = misc[6 + 4 * np];
#ifdef FE_W2_NOSYNC
            nxt = fe_w2_unit(p, un);
#endif
            live_n = nxt.t0 < nxt.T;
            if (live_n) fe_w2_stage_a(sm, lane, warp, FeMbarWait{mb_bl, parity});
        }
        FE_TR(it, 2);
        // ---- M(u) and the second A(u+1) task, in the warp's order (one call site each)
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
            if ((k == 1) == m_before_a2 && has_next) {
                if (live_n && n2b >= 0) fe_w2_stage_a(sm, lane, n2b, FeMbarWait{mb_bl, parity});
                if (un + du < p.total_units) fe_bar_arrive(FE_BAR_W_FREE, C::NT);   // last read of W(next)
            }
            if (k == 0 && !first) {
                fe_mbar_wait(mb_pb, parity);
                if (live) mm = fe_w2_phase_mel<NMELS, OUT16>(p, cur, sm, lane, warp);
            }
            FE_TR(it, 3 + k);
        }
        int kmax = fe_float_key(mm.y), kmin = fe_float_key(-mm.x);
        kmax = __reduce_max_sync(0xffffffffu, kmax);
        kmin = __reduce_max_sync(0xffffffffu, kmin);
        if (lane == 0) { atomicMax(misc + 0, kmax); atomicMax(misc + 1, kmin); }
        FE_TR(it, 5);
        fe_bar_sync(FE_BAR_COMPUTE, C::NCT);                   // P(u) consumed, A(next) complete, keys final
        FE_TR(it, 6);
        if (tid == 0 && !first) {
            const float ymax = fe_key_float(misc[0]), ymin = -fe_key_float(misc[1]);
            p.unit_minmax[u] = make_float2(ymin, ymax);
            if (live) atomicMax(p.max_keys + (p.norm_mode == FE_NORM_PER_CALL ? 0 : cur.b), misc[0]);
            misc[0] = FE_KEY_INIT; misc[1] = FE_KEY_INIT;
        }
        cur = nxt;
    }
}


// Fix-up: clamps the units whose minimum lies below max - 8 decades (y < ymax - 2) and zero-fills the
// frames past a ragged clip's own end.  Everything else is left untouched (one trip to HBM).
// One thread checks one unit (coalesced read of its (min, max)); the units that need work -- few or none for
// noise-like audio of full length -- are then rewritten by the whole CTA, one after the other.
template <int NMELS, bool OUT16>
__global__ void __launch_bounds__(256) fe_whisper2_fixup_kernel(const FeWhisper2Params p) {
    typedef FeW2Cfg C;
    __shared__ int list[256];
    __shared__ int count;
    for (int base = blockIdx.x * blockDim.x; base < p.total_units; base += gridDim.x * blockDim.x) {
        if (threadIdx.x == 0) count = 0;
        __syncthreads();
        const int u = base + threadIdx.x;
        if (u < p.total_units) {
            const FeW2Unit t = fe_w2_unit(p, u);
            const int n_in = (int)(p.T_out - t.t0 < C::FT ? p.T_out - t.t0 : C::FT);
            const int n_live = t.T - t.t0;
            const float2 mm = p.unit_minmax[u];
            const float thr = fe_w2_threshold(fe_key_float(p.max_keys[p.norm_mode == FE_NORM_PER_CALL ? 0 : t.b]));
            if ((n_live > 0 && mm.x < thr) || n_live < n_in) list[atomicAdd(&count, 1)] = u;
        }
        __syncthreads();
        const int n = count;
        for (int j = 0; j < n; ++j) {
            const int uu = list[j];
            const FeW2Unit t = fe_w2_unit(p, uu);
            const int n_in = (int)(p.T_out - t.t0 < C::FT ? p.T_out - t.t0 : C::FT);
            int n_live = t.T - t.t0;
            n_live = n_live < 0 ? 0 : (n_live > C::FT ? C::FT : n_live);
            const float2 mm = p.unit_minmax[uu];
            const float thr = fe_w2_threshold(fe_key_float(p.max_keys[p.norm_mode == FE_NORM_PER_CALL ? 0 : t.b]));
            const bool clampu = n_live > 0 && mm.x < thr;
            // a thread's column c is fixed (blockDim.x is a multiple of FT); its rows are m = m0, m0 + RSTEP, ...
            // Batches of eight independent loads: one HBM round trip per batch instead of one per element.
            // float16 output: rounding is monotone, so clamping the rounded values against the rounded threshold
            // equals rounding the clamped float32 values.
            constexpr int RSTEP = 256 / C::FT;
            const int c = threadIdx.x % C::FT, m0 = threadIdx.x / C::FT;
            if (c < n_in) {
                if (OUT16) {
                    __half* __restrict__ q0 = reinterpret_cast<__half*>(p.out) + (int64_t)t.b * p.out_batch_stride + t.t0 + c;
                    const __half thr_h = __ushort_as_half(fe_f2h_rn(thr));
                    if (c >= n_live) {
                        for (int m = m0; m < NMELS; m += RSTEP) q0[(int64_t)m * p.T_out] = __ushort_as_half(0);
                    } else if (clampu) {
                        for (int mb = m0; mb < NMELS; mb += 8 * RSTEP) {
                            __half v[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const int m = mb + j * RSTEP;
                                v[j] = m < NMELS ? q0[(int64_t)m * p.T_out] : thr_h;
                            }
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const int m = mb + j * RSTEP;
                                if (m < NMELS && __half2float(v[j]) < __half2float(thr_h)) q0[(int64_t)m * p.T_out] = thr_h;
                            }
                        }
                    }
                } else {
                    float* __restrict__ q0 = reinterpret_cast<float*>(p.out) + (int64_t)t.b * p.out_batch_stride + t.t0 + c;
                    if (c >= n_live) {
                        for (int m = m0; m < NMELS; m += RSTEP) q0[(int64_t)m * p.T_out] = 0.0f;
                    } else if (clampu) {
                        for (int mb = m0; mb < NMELS; mb += 8 * RSTEP) {
                            float v[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const int m = mb + j * RSTEP;
                                v[j] = m < NMELS ? q0[(int64_t)m * p.T_out] : thr;
                            }
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const int m = mb + j * RSTEP;
                                if (m < NMELS && v[j] < thr) q0[(int64_t)m * p.T_out] = thr;
                            }
                        }
                    }
                }
            }
        }
        __syncthreads();
    }
}


// =====================================================================================================
// Whisper, Tensor-Memory formulation (fe_whisper3.cuh)
// =====================================================================================================
#ifdef FE_W3_TRACE
// debug build only (tools/trace_w3.py): clock64 stamps of CTA 0's warps at the phase boundaries
#define FE_T3_UNITS 16
#define FE_T3_EV 10
__device__ long long fe_w3_trace_buf[FE_T3_UNITS * 16 * FE_T3_EV];
#define FE_T3(iter, ev) do { if (blockIdx.x == 0 && (unsigned)(iter) < FE_T3_UNITS && lane == 0) \
    fe_w3_trace_buf[((iter) * 16 + warp) * FE_T3_EV + (ev)] = clock64(); } while (0)
#else
#define FE_T3(iter, ev) do { } while (0)
#endif

static __device__ __forceinline__ void fe_tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
static __device__ __forceinline__ void fe_tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// mbarrier wait that cannot hang the GPU: a protocol error traps (the launch fails) instead of spinning forever
static __device__ __forceinline__ void fe_mbar_wait_bounded(unsigned addr, unsigned parity) {
    unsigned ok = 0;
#pragma unroll 1
    for (int spin = 0; spin < (1 << 20); ++spin) {
        // the suspend-time hint lets the hardware park the warp instead of returning at once: without it the spin loops
        // were 23 % of all executed instructions (profiles/r02_w3_v1)
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.b32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(addr), "r"(parity), "r"(200u) : "memory");
        if (ok) return;
    }
    __trap();
}
static __device__ __forceinline__ void fe_mbar_expect_tx(unsigned addr, unsigned bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(addr), "r"(bytes) : "memory");
}
// one arrival per warp, after every lane's prior accesses (shared memory: __syncwarp + release; Tensor Memory: the
// caller runs the tcgen05 wait + fence first)
static __device__ __forceinline__ void fe_warp_arrive(unsigned addr, int lane) {
    __syncwarp();
    if (lane == 0) fe_mbar_arrive(addr);
}

struct FeW3Sync {          // shared-window addresses of the mbarriers
    unsigned wfull, afull, bloaded;   // of this warp's team
    unsigned pfull, pfree;
};
struct FeW3WaitBLoaded {   // phase A, before its first Tensor Memory store of a unit: B(previous unit) has loaded everything
    unsigned addr, parity; bool on; bool* done;
    __device__ __forceinline__ void operator()() const {
        if (on && !*done) { fe_mbar_wait_bounded(addr, parity); fe_tc_fence_after(); *done = true; }
    }
};
struct FeW3ArriveBLoaded { // phase B, after the loads of the member's LAST task of the unit have landed
    unsigned addr; int lane; bool on;
    __device__ __forceinline__ void operator()() const {
        if (on) { fe_tc_fence_before(); fe_warp_arrive(addr, lane); }
    }
};
struct FeW3WaitPFree {     // phase B, before its first store into P of a unit: M(previous unit) has read everything
    unsigned addr, parity; bool on; bool* done;
    __device__ __forceinline__ void operator()() const {
        if (on && !*done) { fe_mbar_wait_bounded(addr, parity); *done = true; }
    }
};

// W[q] of unit `t`: the issuing warp first writes the edge rows element-wise, then lane 0 arms the team's mbarrier with
// the byte count of the bulk rows and issues one 640-byte cp.async.bulk per row (they complete on that barrier); a dead
// team's phase is completed by the bare arrival
struct FeW3Issue {      // what the copy issuer needs of the launch parameters, by value: passing the parameter block by
    const float* wave;  // reference to a non-inlined function moves it to local memory for the WHOLE kernel
    int64_t wave_stride;
    int32_t vec_ok, use_tmap, tmap_rows;
};
static __device__ __noinline__ void fe_w3_issue_w(const FeW3Issue p, const CUtensorMap* tmap, int b, int f0, int64_t len, int64_t lp,
                                                  float* wq, unsigned mb_wfull, int lane) {
    typedef FeW3Cfg C;
    const int64_t g0 = (int64_t)f0 * C::HOP - C::PADL;
    // interior units: ONE tensor copy of the whole (34 rows x 164 floats) box.  The tensor map describes clip b as rows of
    // 164 floats that start every 160 floats (overlapping rows: the four extra floats per row are the pitch padding), so
    // the box lands in shared memory exactly in the W[q] layout.
    if (p.use_tmap) {      // always the whole box; rows outside the map arrive as zeros, the team patches what is not right
        if (lane == 0) {
            fe_mbar_expect_tx(mb_wfull, C::ROWS * C::PITCH * 4);
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"((unsigned)__cvta_generic_to_shared(wq)), "l"(tmap), "r"(0), "r"(f0 - C::TMAP_ROW0), "r"(b), "r"(mb_wfull) : "memory");
        }
        return;
    }
    // launches without a tensor map: the issuing warp first writes the edge rows element-wise, then lane 0 arms the team's
    // mbarrier with the byte count of the bulk rows (one contiguous run [r_lo, r_hi)) and issues one 640-byte copy per row
    int r_lo = 0, r_hi = 0;
    if (p.vec_ok & 1) {
        const int64_t lo = g0 >= 0 ? 0 : (-g0 + C::HOP - 1) / C::HOP;
        const int64_t hi = len - g0 < 0 ? 0 : (len - g0) / C::HOP;
        r_lo = (int)(lo < C::ROWS ? lo : C::ROWS);
        r_hi = (int)(hi < C::ROWS ? hi : C::ROWS);
        if (r_hi < r_lo) r_hi = r_lo;
    }
    const float* __restrict__ x = p.wave + (int64_t)b * p.wave_stride;
    if (r_hi - r_lo < C::ROWS) fe_w3_fill_rows_manual(x, g0, len, lp, r_lo, r_hi, wq, lane);
    __syncwarp();
    if (lane == 0) {
        if (r_hi > r_lo) fe_mbar_expect_tx(mb_wfull, (unsigned)((r_hi - r_lo) * C::ROW_BYTES));
        else fe_mbar_arrive(mb_wfull);
        const float* src = x + g0;
        const unsigned dst = (unsigned)__cvta_generic_to_shared(wq);
#pragma unroll 4
        for (int r = r_lo; r < r_hi; ++r)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst + (unsigned)(r * C::PITCH * 4)), "l"(src + r * C::HOP), "n"(C::ROW_BYTES), "r"(mb_wfull) : "memory");
    }
}
// W[q] of unit `t` (inlined part: a dead team's phase is completed by a bare arrival)
static __device__ __forceinline__ void fe_w3_request_w(const FeWhisper2Params& p, const CUtensorMap* tmap, const FeW3Unit& t, int q,
                                                       float* wq, unsigned mb_wfull, int lane) {
    if (!fe_w3_team_live(t, q)) { fe_warp_arrive(mb_wfull, lane); return; }
    FeW3Issue a;
    a.wave = p.wave; a.wave_stride = p.wave_stride; a.vec_ok = p.vec_ok; a.use_tmap = p.use_tmap; a.tmap_rows = p.tmap_rows;
    fe_w3_issue_w(a, tmap, t.b, t.t0 + FeW3Cfg::FT * q, t.len, t.lp, wq, mb_wfull, lane);
}
// the rows of the unit after next: into L2 now, so that the bulk copies issued one unit later hit there
static __device__ __forceinline__ void fe_w3_prefetch_l2(const FeWhisper2Params& p, const FeW3Unit& t, int q, int lane) {
    typedef FeW3Cfg C;
    const int64_t g0 = fe_w3_row0(t, q);
    if (lane == 0 && fe_w3_team_live(t, q) && (p.vec_ok & 1) && g0 >= 0 && g0 + (int64_t)C::ROWS * C::HOP <= t.len) {
        const float* src = p.wave + (int64_t)t.b * p.wave_stride + g0;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "n"(C::ROWS * C::ROW_BYTES) : "memory");
    }
}

// Per unit `it` of this CTA every warp runs, in program order:
//   A(it)   : wait WFULL(it); its phase-A tasks [first Tensor Memory store waits BLOADED(it-1)]; arrive AFULL(it)
//   M(it-1) : wait PFULL(it-1); its share of the mel projection; arrive PFREE(it-1)
//   B(it)   : wait AFULL(it) [member 1 then issues the copies for W(it+1): every A(it) task has read W];
//             its phase-B tasks [last task's loads landed: arrive BLOADED(it); first store into P waits PFREE(it-1)];
//             arrive PFULL(it)
// A warp arrives on phase k of a barrier only after it has waited for phase k-1 of it (where it waits at all), so no
// barrier can run two phases ahead of a waiter (parity waits would alias).
template <int NMELS, bool OUT16, bool PAD>
__global__ void __launch_bounds__(FeW3Cfg::NT, 1) fe_whisper3_kernel(const FeWhisper2Params p, const __grid_constant__ CUtensorMap tmap) {
    typedef FeW3Cfg C;
    extern __shared__ __align__(1024) float sm3[];
    float* const sm = sm3;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int team = warp & 3, member = warp >> 2;
    float* const wq = sm + C::OFF_W + team * C::WT;
    float* const P = sm + C::OFF_P;
    const unsigned mb0 = (unsigned)__cvta_generic_to_shared(sm + C::OFF_MBAR);
    FeW3Sync sy;
    sy.wfull = mb0 + 8u * team; sy.afull = mb0 + 8u * (4 + team); sy.bloaded = mb0 + 8u * (8 + team);
    sy.pfull = mb0 + 8u * 12; sy.pfree = mb0 + 8u * 13;
    unsigned* const tm_slot = reinterpret_cast<unsigned*>(sm + C::OFF_MBAR + 2 * C::NMBAR);
    int* const issued = reinterpret_cast<int*>(sm + C::OFF_MBAR + 2 * C::NMBAR + 4);    // per team: members that finished phase A, all units
    if (tid < 4) issued[tid] = 0;
    if (tid == 0) {
        for (int q = 0; q < 4; ++q) {
            fe_mbar_init(mb0 + 8u * q, 1);
            fe_mbar_init(mb0 + 8u * (4 + q), C::MEMBERS);
            fe_mbar_init(mb0 + 8u * (8 + q), C::MEMBERS);
        }
        fe_mbar_init(sy.pfull, C::NW);
        fe_mbar_init(sy.pfree, C::NW);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    fe_w3_copy_tables<NMELS>(sm, tid, C::NT);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"((unsigned)__cvta_generic_to_shared(tm_slot)), "n"(C::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fe_tc_fence_before();
    __syncthreads();
    fe_tc_fence_after();
    const unsigned tm_base = *tm_slot;
    FeTm tm;
    tm.base = tm_base + ((unsigned)(team * 32) << 16);

    const int u0 = blockIdx.x, du = gridDim.x;
    const int n_it = u0 < p.total_units ? (p.total_units - u0 + du - 1) / du : 0;
    const int a_single = fe_w3_a_single(member), n_b = fe_w3_b_count(member);
    FeW3Pos pos;
    pos.within = u0 / p.B; pos.b = u0 - pos.within * p.B;      // the one division of the kernel
    if (member == 1 && n_it > 0) {
        fe_w3_request_w(p, &tmap, fe_w3_unit_at<false, PAD>(p, pos.b, pos.within), team, wq, sy.wfull, lane);
        const FeW3Pos p1 = fe_w3_pos_next(p, pos);
        if (n_it > 1) fe_w3_prefetch_l2(p, fe_w3_unit_at<false, PAD>(p, p1.b, p1.within), team, lane);
    }
    FeW3Unit prev;
    memset(&prev, 0, sizeof(prev));
#pragma unroll 1
    for (int it = 0; it <= n_it; ++it) {
        const bool has = it < n_it;
        const unsigned par = (unsigned)(it & 1), ppar = par ^ 1u;
        FeW3Unit cur = prev;
        bool live = false;
        FE_T3(it, 0);
        // ---- A(it)
        if (has) {
            cur = fe_w3_unit_at<false, PAD>(p, pos.b, pos.within);
            live = fe_w3_team_live(cur, team);
            fe_mbar_wait_bounded(sy.wfull, par);
            FE_T3(it, 1);
            if (live) {
                if (p.use_tmap && !fe_w3_box_complete(p, cur, team)) {     // clip edges, ragged ends: patch rows, team barrier
                    fe_w3_patch_rows(p, cur, team, wq, member, lane);
                    fe_bar_sync(1 + team, C::MEMBERS * 32);
                }
                bool waited = false;
                const FeW3WaitBLoaded pre{sy.bloaded, ppar, it > 0, &waited};
#ifdef FE_W3_PAIR         // A/B build only: conflict-free LDS.128 for the member's first two half-units.  Measured on one
                          // B200, same box: 1.046 ms against 1.028 ms for the form below -- the kernel is bound by instruction
                          // issue, not by the shared-memory pipe, and the pair form costs code size and registers.
                fe_w3_stage_a_pair(wq, tm, lane, member, pre);
                if (a_single >= 0) fe_w3_stage_a(wq, tm, lane, a_single, pre);
#else
#pragma unroll 1
                for (int k = 0; k < (a_single >= 0 ? 3 : 2); ++k)      // one call site: the kernel is sensitive to code size
                    fe_w3_stage_a(wq, tm, lane, k < 2 ? 2 * member + k : a_single, pre);
#endif
                fe_tm_wait_st();
            }
            fe_tc_fence_before();
            fe_warp_arrive(sy.afull, lane);
            {   // the member that finishes A(it) LAST refills W for the next unit at once (every A(it) task has read W by
                // then): the copy then has the whole of M(it-1) and B(it) to land
                int last = 0;
                if (lane == 0) {
                    __threadfence_block();
                    last = atomicAdd(issued + team, 1) == 4 * it + 3;
                    __threadfence_block();
                }
                last = __shfl_sync(0xffffffffu, last, 0);
                if (last) {
                    const FeW3Pos p1 = fe_w3_pos_next(p, pos), p2 = fe_w3_pos_next(p, p1);
                    if (it + 1 < n_it) fe_w3_request_w(p, &tmap, fe_w3_unit_at<false, PAD>(p, p1.b, p1.within), team, wq, sy.wfull, lane);
                    if (it + 2 < n_it) fe_w3_prefetch_l2(p, fe_w3_unit_at<false, PAD>(p, p2.b, p2.within), team, lane);
                }
            }
        }
        FE_T3(it, 2);
        // ---- M(it - 1)
        if (it > 0) {
            fe_mbar_wait_bounded(sy.pfull, ppar);
            FE_T3(it, 3);
            if (prev.t0 < prev.T) {
                const fe2 mm = fe_w3_phase_mel<NMELS, OUT16>(p, prev, sm, lane, warp);
                int kmax = fe_float_key(mm.y), kmin = fe_float_key(-mm.x);
                kmax = __reduce_max_sync(0xffffffffu, kmax);
                kmin = __reduce_max_sync(0xffffffffu, kmin);
                if (lane == 0) {
                    int* uk = reinterpret_cast<int*>(p.unit_minmax + (u0 + (it - 1) * du));
                    atomicMax(uk, kmax);                // [0] key of the unit's maximum, [1] key of -minimum
                    atomicMax(uk + 1, kmin);
                    atomicMax(p.max_keys + (p.norm_mode == FE_NORM_PER_CALL ? 0 : prev.b), kmax);
                }
            }
            if (has) fe_warp_arrive(sy.pfree, lane);
        }
        FE_T3(it, 4);
        // ---- B(it)
        if (has) {
            fe_mbar_wait_bounded(sy.afull, par);
            fe_tc_fence_after();
            FE_T3(it, 5);
            bool waited = false;
            const FeW3WaitPFree pre{sy.pfree, ppar, it > 0, &waited};
            if (live) {
#pragma unroll 1
                for (int k = 0; k < n_b; ++k) {
                    fe_w3_stage_b(P, tm, lane, team, fe_w3_b_task(member, k), FeW3ArriveBLoaded{sy.bloaded, lane, k == n_b - 1}, pre);
                    FE_T3(it, 6 + k);
                }
            } else {
                FeW3ArriveBLoaded{sy.bloaded, lane, true}();
            }
            fe_warp_arrive(sy.pfull, lane);
        }
        prev = cur;
        pos = fe_w3_pos_next(p, pos);
    }
    fe_tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm_base), "n"(C::TMEM_COLS) : "memory");
}

// Fix-up of the Tensor-Memory kernel's output: 128-frame units, (max, -min) published as order-preserving keys
template <int NMELS, bool OUT16>
__global__ void __launch_bounds__(256) fe_whisper3_fixup_kernel(const FeWhisper2Params p) {
    typedef FeW3Cfg C;
    __shared__ int list[256];
    __shared__ int count;
    const int2* __restrict__ ukeys = reinterpret_cast<const int2*>(p.unit_minmax);
    const bool plain = p.win_frames == 0;
    if (plain) {
        // The tail of every clip: whole units of padding (the kernel skipped them; their value is the clip's threshold)
        // and whole units behind the clip's last frame (0).  Written as contiguous runs along the output rows, one
        // (clip, 16 rows) item per CTA pass, one row per warp: unit by unit (512 bytes every 4 * T_out bytes) the same
        // bytes went out at 3 TB/s, profiles/r02_bench_realistic.txt.
        constexpr int GROWS = 16, NG = (NMELS + GROWS - 1) / GROWS;
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        const bool vec = (p.vec_ok & 2) && (p.T_out & 3) == 0;
        for (int w = blockIdx.x; w < p.B * NG; w += gridDim.x) {
            const int b = w / NG, g = w - b * NG;
            int64_t Tc;
            const int64_t f0 = fe_w3_tail_start(p, b, &Tc);
            if (f0 >= p.T_out) continue;
            const float thr = f0 < Tc ? fe_w2_threshold(fe_key_float(p.max_keys[p.norm_mode == FE_NORM_PER_CALL ? 0 : b])) : 0.0f;
            const unsigned short thr_h = fe_f2h_rn(thr);
            for (int m = g * GROWS + warp; m < NMELS && m < (g + 1) * GROWS; m += 8) {
                const int64_t row = (int64_t)b * p.out_batch_stride + (int64_t)m * p.T_out;
                if (OUT16) {
                    unsigned short* __restrict__ q = reinterpret_cast<unsigned short*>(p.out) + row;
                    if (vec) {
                        for (int64_t f = f0 + 4 * lane; f < p.T_out; f += 128) {
                            const unsigned a0 = f < Tc ? thr_h : 0u, a1 = f + 1 < Tc ? thr_h : 0u;
                            const unsigned a2 = f + 2 < Tc ? thr_h : 0u, a3 = f + 3 < Tc ? thr_h : 0u;
                            *reinterpret_cast<uint2*>(q + f) = make_uint2(a0 | (a1 << 16), a2 | (a3 << 16));
                        }
                    } else {
                        for (int64_t f = f0 + lane; f < p.T_out; f += 32) q[f] = f < Tc ? thr_h : (unsigned short)0;
                    }
                } else {
                    float* __restrict__ q = reinterpret_cast<float*>(p.out) + row;
                    if (vec) {
                        for (int64_t f = f0 + 4 * lane; f < p.T_out; f += 128)
                            *reinterpret_cast<float4*>(q + f) = make_float4(f < Tc ? thr : 0.0f, f + 1 < Tc ? thr : 0.0f,
                                                                            f + 2 < Tc ? thr : 0.0f, f + 3 < Tc ? thr : 0.0f);
                    } else {
                        for (int64_t f = f0 + lane; f < p.T_out; f += 32) q[f] = f < Tc ? thr : 0.0f;
                    }
                }
            }
        }
    }
    // unit u is checked by CTA u % gridDim.x: neighbouring units (the same position of neighbouring clips, which tend to
    // need the clamp together -- clip starts, silent stretches) go to DIFFERENT CTAs, so the rewrites spread over the
    // whole GPU.  (With contiguous blocks of units per CTA the 37 flagged units of the noise bench sat in a few CTAs
    // that rewrote them one after the other: 57 us with the SMs 95 % idle, profiles/r02_whisper3_ncu_summary.txt.)
    for (int base = 0; base < p.total_units; base += gridDim.x * blockDim.x) {
        if (threadIdx.x == 0) count = 0;
        __syncthreads();
        const int u = base + (int)threadIdx.x * (int)gridDim.x + (int)blockIdx.x;
        if (u < p.total_units) {
            const FeW3Unit t = fe_w3_unit<true>(p, u);
            const int n_in = t.n_in;
            const int n_live = t.Ts - t.t0;
            const float umin = -fe_key_float(ukeys[u].y);
            const float thr = fe_w2_threshold(fe_key_float(p.max_keys[p.norm_mode == FE_NORM_PER_CALL ? 0 : t.b]));
            const bool in_tail = plain && (t.silent || n_live <= 0);       // written above, as part of a row run
            if (!in_tail && (t.silent || (n_live > 0 && umin < thr) || n_live < n_in)) list[atomicAdd(&count, 1)] = u;
        }
        __syncthreads();
        const int n = count;
        for (int j = 0; j < n; ++j) {
            const int uu = list[j];
            const FeW3Unit t = fe_w3_unit<true>(p, uu);
            const int n_in = t.n_in;
            int n_live = t.Ts - t.t0;
            n_live = n_live < 0 ? 0 : (n_live > n_in ? n_in : n_live);
            const int64_t pitch = t.pitch;
            const float umin = -fe_key_float(ukeys[uu].y);
            const float thr = fe_w2_threshold(fe_key_float(p.max_keys[p.norm_mode == FE_NORM_PER_CALL ? 0 : t.b]));
            const bool clampu = n_live > 0 && umin < thr;
            constexpr int RSTEP = 256 / C::UF;
            const int c = threadIdx.x % C::UF, m0 = threadIdx.x / C::UF;
            if ((p.vec_ok & 2) && (pitch & 3) == 0 && n_in == C::UF && ((t.silent && n_live == C::UF) || n_live == 0)) {
                // a whole unit of padding, or a whole unit behind the clip's last frame: (n_mels x 128) times the
                // threshold, or 0 -- 16 bytes (8 for float16) per thread and row
                const int c4 = 4 * (threadIdx.x & 31), r0 = threadIdx.x >> 5;
                const float fill = n_live == 0 ? 0.0f : thr;
                if (OUT16) {
                    const unsigned h = fe_f2h_rn(fill);
                    const uint2 v = make_uint2(h | (h << 16), h | (h << 16));
                    __half* __restrict__ q0 = reinterpret_cast<__half*>(p.out) + t.o0 + c4;
#pragma unroll 4
                    for (int m = r0; m < NMELS; m += 8) *reinterpret_cast<uint2*>(q0 + (int64_t)m * pitch) = v;
                } else {
                    const float4 v = make_float4(fill, fill, fill, fill);
                    float* __restrict__ q0 = reinterpret_cast<float*>(p.out) + t.o0 + c4;
#pragma unroll 4
                    for (int m = r0; m < NMELS; m += 8) *reinterpret_cast<float4*>(q0 + (int64_t)m * pitch) = v;
                }
            } else if (c < n_in) {
                if (OUT16) {
                    __half* __restrict__ q0 = reinterpret_cast<__half*>(p.out) + t.o0 + c;
                    const __half thr_h = __ushort_as_half(fe_f2h_rn(thr));
                    if (c >= n_live) {
                        for (int m = m0; m < NMELS; m += RSTEP) q0[(int64_t)m * pitch] = __ushort_as_half(0);
                    } else if (t.silent) {
                        for (int m = m0; m < NMELS; m += RSTEP) q0[(int64_t)m * pitch] = thr_h;
                    } else if (clampu) {
                        for (int mb = m0; mb < NMELS; mb += 8 * RSTEP) {
                            __half v[8];
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                const int m = mb + k * RSTEP;
                                v[k] = m < NMELS ? q0[(int64_t)m * pitch] : thr_h;
                            }
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                const int m = mb + k * RSTEP;
                                if (m < NMELS && __half2float(v[k]) < __half2float(thr_h)) q0[(int64_t)m * pitch] = thr_h;
                            }
                        }
                    }
                } else {
                    float* __restrict__ q0 = reinterpret_cast<float*>(p.out) + t.o0 + c;
                    if (c >= n_live) {
                        for (int m = m0; m < NMELS; m += RSTEP) q0[(int64_t)m * pitch] = 0.0f;
                    } else if (t.silent) {
                        for (int m = m0; m < NMELS; m += RSTEP) q0[(int64_t)m * pitch] = thr;
                    } else if (clampu) {
                        for (int mb = m0; mb < NMELS; mb += 8 * RSTEP) {
                            float v[8];
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                const int m = mb + k * RSTEP;
                                v[k] = m < NMELS ? q0[(int64_t)m * pitch] : thr;
                            }
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                const int m = mb + k * RSTEP;
                                if (m < NMELS && v[k] < thr) q0[(int64_t)m * pitch] = thr;
                            }
                        }
                    }
                }
            }
        }
        __syncthreads();
    }
}

// =====================================================================================================
// kaldi fbank, Tensor-Memory formulation (fe_kaldi3.cuh) -- the skeleton of fe_whisper3_kernel
// =====================================================================================================
// Per unit `it` every warp runs, in program order:
//   A(it)   : wait WFULL(it); [patch rows the box copy could not deliver; team barrier]; row partial sums; team barrier;
//             frame mean; its two phase-A half-units [first Tensor Memory store waits BLOADED(it-1)]; arrive AFULL(it);
//             the member that finishes last requests W(it+1)
//   M(it-1) : wait PFULL(it-1); its filter groups of the mel projection straight to `out`; arrive PFREE(it-1)
//   B(it)   : wait AFULL(it); its phase-B tasks [last task's loads landed: arrive BLOADED(it); first store into P waits
//             PFREE(it-1)]; arrive PFULL(it)
__global__ void __launch_bounds__(FeK3Cfg::NT, 1) fe_kaldi3_kernel(const FeKaldiParams p, const __grid_constant__ CUtensorMap tmap) {
    typedef FeK3Cfg C;
    extern __shared__ __align__(1024) float smk3[];
    float* const sm = smk3;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int team = warp & 3, member = warp >> 2;
    float* const wq = sm + C::OFF_W + team * C::WT;
    float* const P = sm + C::OFF_P;
    float* const part = sm + C::OFF_PART + team * C::PARTS * C::PARTW;
    const unsigned mb0 = (unsigned)__cvta_generic_to_shared(sm + C::OFF_MBAR);
    FeW3Sync sy;
    sy.wfull = mb0 + 8u * team; sy.afull = mb0 + 8u * (4 + team); sy.bloaded = mb0 + 8u * (8 + team);
    sy.pfull = mb0 + 8u * 12; sy.pfree = mb0 + 8u * 13;
    unsigned* const tm_slot = reinterpret_cast<unsigned*>(sm + C::OFF_MBAR + 2 * C::NMBAR);
    int* const finished = reinterpret_cast<int*>(sm + C::OFF_MBAR + 2 * C::NMBAR + 4);   // per team: members done with phase A
    if (tid < 4) finished[tid] = 0;
    if (tid == 0) {
        for (int q = 0; q < 4; ++q) {
            fe_mbar_init(mb0 + 8u * q, 1);
            fe_mbar_init(mb0 + 8u * (4 + q), C::MEMBERS);
            fe_mbar_init(mb0 + 8u * (8 + q), C::MEMBERS);
        }
        fe_mbar_init(sy.pfull, C::NW);
        fe_mbar_init(sy.pfree, C::NW);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    fe_k3_copy_tables(sm, tid, C::NT);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"((unsigned)__cvta_generic_to_shared(tm_slot)), "n"(C::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fe_tc_fence_before();
    __syncthreads();
    fe_tc_fence_after();
    const unsigned tm_base = *tm_slot;
    FeTm tm;
    tm.base = tm_base + ((unsigned)(team * 32) << 16);

    const int u0 = blockIdx.x, du = gridDim.x;
    const int n_it = u0 < p.total_units ? (p.total_units - u0 + du - 1) / du : 0;
    FeW3Pos pos;
    pos.within = u0 / p.B; pos.b = u0 - pos.within * p.B;
    const int n_b = fe_k3_b_count(member);
    // W[q] of a unit: ONE tensor copy of the (34 rows x 164 floats) box; a dead team's phase is completed by a bare arrival
    auto request_w = [&](const FeK3Unit& t) {
        if (!fe_k3_team_live(t, team)) { fe_warp_arrive(sy.wfull, lane); return; }
        if (lane == 0) {
            fe_mbar_expect_tx(sy.wfull, C::ROWS * C::PITCH * 4);
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"((unsigned)__cvta_generic_to_shared(wq)), "l"(&tmap), "r"(0), "r"(t.t0 + C::FT * team), "r"(t.b), "r"(sy.wfull) : "memory");
        }
    };
    if (member == 1 && n_it > 0) request_w(fe_k3_unit_at(p, pos.b, pos.within));
    FeK3Unit prev;
    memset(&prev, 0, sizeof(prev));
#pragma unroll 1
    for (int it = 0; it <= n_it; ++it) {
        const bool has = it < n_it;
        const unsigned par = (unsigned)(it & 1), ppar = par ^ 1u;
        FeK3Unit cur = prev;
        bool live = false;
        // ---- A(it)
        if (has) {
            cur = fe_k3_unit_at(p, pos.b, pos.within);
            live = fe_k3_team_live(cur, team);
            fe_mbar_wait_bounded(sy.wfull, par);
            if (live) {
                if (!fe_k3_box_complete(p, cur, team)) {
                    fe_k3_patch_rows(p, cur, team, wq, member, lane);
                    fe_bar_sync(1 + team, C::MEMBERS * 32);
                }
                {
                    FeK3Pre pre_regs;
                    fe_k3_pre_load(wq, member, lane, pre_regs);
                    fe_bar_sync(1 + team, C::MEMBERS * 32);          // every member has read its neighbours' samples
                    fe_k3_pre_store(wq, part, member, lane, pre_regs);
                }
                fe_bar_sync(1 + team, C::MEMBERS * 32);
                const float mean = fe_k3_frame_mean(part, lane), x0 = fe_k3_frame_x0(part, lane);
                bool waited = false;
                const FeW3WaitBLoaded pre{sy.bloaded, ppar, it > 0, &waited};
#pragma unroll 1
                for (int k = 0; k < 2; ++k) fe_k3_stage_a(wq, tm, lane, 2 * member + k, mean, x0, pre);
                fe_tm_wait_st();
            }
            fe_tc_fence_before();
            fe_warp_arrive(sy.afull, lane);
            {
                int last = 0;
                if (lane == 0) {
                    __threadfence_block();
                    last = atomicAdd(finished + team, 1) == 4 * it + 3;
                    __threadfence_block();
                }
                last = __shfl_sync(0xffffffffu, last, 0);
                if (last && it + 1 < n_it) {
                    const FeW3Pos p1 = fe_k3_pos_next(p, pos);
                    request_w(fe_k3_unit_at(p, p1.b, p1.within));
                }
            }
        }
        // ---- M(it - 1)
        if (it > 0) {
            fe_mbar_wait_bounded(sy.pfull, ppar);
            if (prev.t0 < prev.T) fe_k3_phase_mel(p, prev, sm, lane, warp);
            else fe_k3_zero_unit(p, prev, lane, warp);
            if (has) fe_warp_arrive(sy.pfree, lane);
        }
        // ---- B(it)
        if (has) {
            fe_mbar_wait_bounded(sy.afull, par);
            fe_tc_fence_after();
            bool waited = false;
            const FeW3WaitPFree pre{sy.pfree, ppar, it > 0, &waited};
            if (live) {
#pragma unroll 1
                for (int k = 0; k < n_b; ++k)
                    fe_k3_stage_b(P, tm, lane, team, fe_k3_b_task(member, k), FeW3ArriveBLoaded{sy.bloaded, lane, k == n_b - 1}, pre);
            } else {
                FeW3ArriveBLoaded{sy.bloaded, lane, true}();
            }
            fe_warp_arrive(sy.pfull, lane);
        }
        prev = cur;
        pos = fe_k3_pos_next(p, pos);
    }
    fe_tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm_base), "n"(C::TMEM_COLS) : "memory");
}

// CMN of the Tensor-Memory path: feat - feat.mean(0) per clip.  One CTA per (clip, half of the 80 columns): the half-tile
// (T, 40) of a clip up to 6.4 s fits 100 KB of shared memory, so TWO CTAs are resident per SM and one's loads overlap the
// other's stores (one CTA per clip with the whole tile: 0.34 ms for 4096 x 5 s; the traffic alone is 0.20 ms).  The tile
// is read ONCE (global -> shared memory -> column means -> subtract -> write); longer clips make two passes over global
// memory.  Fixed summation order: deterministic.
__global__ void __launch_bounds__(320, 2) fe_kaldi3_cmn_kernel(const FeKaldiParams p, int smem_rows) {
    typedef FeK3Cfg C;
    constexpr int HC = C::NMEL / 2, H4 = HC / 4;       // 40 columns = 10 float4 per row and CTA
    extern __shared__ __align__(16) float tile[];      // [T][HC]
    __shared__ float partial[8][HC];
    __shared__ float mean[HC];
    const int b = blockIdx.x >> 1, c0 = (blockIdx.x & 1) * HC;
    const FeK3Unit t = fe_k3_unit_at(p, b, 0);
    const int T = t.T;
    if (T <= 0) return;
    float* __restrict__ o = p.out + (int64_t)b * p.out_batch_stride + c0;
    const bool in_smem = T <= smem_rows;
    const bool vec = (p.vec_ok & 2) != 0;
    const int tid = threadIdx.x;
    const int n4 = T * H4;
    if (in_smem) {
        if (vec) {
            // eight independent 16-byte loads per thread in flight (2 CTAs x 320 threads x 128 B = 80 KB per SM)
            for (int i0 = tid; i0 < n4; i0 += 320 * 8) {
                float4 v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int i = i0 + 320 * k;
                    if (i < n4) v[k] = __ldcs(reinterpret_cast<const float4*>(o + (int64_t)(i / H4) * C::NMEL) + i % H4);
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int i = i0 + 320 * k;
                    if (i < n4) reinterpret_cast<float4*>(tile)[i] = v[k];
                }
            }
        } else {
            for (int i = tid; i < T * HC; i += 320) tile[i] = __ldcs(o + (int64_t)(i / HC) * C::NMEL + i % HC);
        }
        __syncthreads();
    }
    const int m = tid % HC, rg = tid / HC;             // 8 row groups x 40 columns
    float acc = 0.0f;
    if (in_smem) for (int r = rg; r < T; r += 8) acc += tile[r * HC + m];
    else for (int r = rg; r < T; r += 8) acc += __ldcg(o + (int64_t)r * C::NMEL + m);
    partial[rg][m] = acc;
    __syncthreads();
    if (tid < HC)
        mean[tid] = (((partial[0][tid] + partial[1][tid]) + (partial[2][tid] + partial[3][tid])) +
                     ((partial[4][tid] + partial[5][tid]) + (partial[6][tid] + partial[7][tid]))) / (float)T;
    __syncthreads();
    if (vec) {
        for (int i = tid; i < n4; i += 320) {
            const int r = i / H4, q = i % H4;
            float4* dst = reinterpret_cast<float4*>(o + (int64_t)r * C::NMEL) + q;
            float4 v = in_smem ? reinterpret_cast<const float4*>(tile)[i] : __ldcg(dst);
            v.x -= mean[4 * q]; v.y -= mean[4 * q + 1]; v.z -= mean[4 * q + 2]; v.w -= mean[4 * q + 3];
            __stcs(dst, v);
        }
    } else {
        for (int i = tid; i < T * HC; i += 320) {
            float* dst = o + (int64_t)(i / HC) * C::NMEL + i % HC;
            *dst = (in_smem ? tile[i] : __ldcg(dst)) - mean[i % HC];
        }
    }
}

static int tiles_per_cta_env(const char* name, int dflt) {
    const char* s = getenv(name);
    const int t = s ? atoi(s) : dflt;
    return t < 1 ? 1 : t;
}

// B200, 4096 x 5 s: 1 tile per CTA 2.45 ms, 2 -> 2.42, 4 -> 2.45, 8 -> 2.76 (tail of the L2-sized waves)
static int kaldi_tiles_per_cta() {
    static const int v = tiles_per_cta_env("B200FEAT_KALDI_TPC", 2);
    return v;
}

// One CTA runs `tiles_per_cta` consecutive tiles (32 frames each) of the launch; the global loads of the next tile
// are in flight while the phases of the current one run.
__global__ void __launch_bounds__(FeKCfg::NT, 2) fe_kaldi_kernel(const FeKaldiParams p, int total_tiles, int tiles_per_cta) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x;
    const int first = blockIdx.x * tiles_per_cta;
    const int last = first + tiles_per_cta < total_tiles ? first + tiles_per_cta : total_tiles;
    FeKStage st;
    if (first < last) fe_k_load_issue(p, tid, first, st);
    for (int cta = first; cta < last; ++cta) {
        fe_k_load_store(sm, tid, st);
        if (cta + 1 < last) fe_k_load_issue(p, tid, cta + 1, st);
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
        // no barrier needed here: the next pass writes W (the region the dead P occupied) and every thread meets two
        // barriers before anything writes the staging tile that phase O is still reading
    }
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

// One tile (16 frames) per CTA.  A multi-tile CTA with the next tile's loads in flight (as in the kaldi kernel) was
// measured: 0.926 ms against 0.905 ms for 256 x 10 s, the tile load is not what this kernel waits for.
template <class C, int MINB>
static __device__ __forceinline__ void fe_stft3_body(const FeCosyParams& p) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x, cta = blockIdx.x;
    fe_c_copy_tables<C>(sm, tid);
    fe_c_phase_load<C>(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_a<C>(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_b<C>(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_c<C>(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_mel<C>(p, sm, tid, cta);
}
__global__ void __launch_bounds__(FeCCfg::NT, 1) fe_cosy_kernel(const FeCosyParams p) { fe_stft3_body<FeCCfg, 1>(p); }
// Wav2Lip / Diff2Lip mel: the same three stages with N = 800 = 16 x 5 x 10; 85 KB of shared
// memory per CTA, two CTAs per SM
__global__ void __launch_bounds__(FeLCfg::NT, 2) fe_wav2lip_kernel(const FeCosyParams p) { fe_stft3_body<FeLCfg, 2>(p); }

// =====================================================================================================
// resampler, long filters: tensor-core GEMM (fe_resample_tc.cuh)
// =====================================================================================================
// interleaved PCM -> mono float32 rows (the A-operand producers read 16-byte aligned float32 rows)
template <bool IN16, int CH>
__global__ void __launch_bounds__(256) fe_pcm_mono_kernel(const void* __restrict__ pcm, int64_t pcm_stride, float w0, float w1,
                                                          float* __restrict__ out, int64_t out_stride, int64_t L) {
    const int b = blockIdx.y;
    float* __restrict__ o = out + (int64_t)b * out_stride;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < L; i += (int64_t)gridDim.x * blockDim.x) {
        float v;
        if (IN16) {
            const short* __restrict__ s = static_cast<const short*>(pcm) + (int64_t)b * pcm_stride + i * CH;
            v = CH == 2 ? fmaf(w1, (float)s[1] * (1.0f / 32768.0f), w0 * ((float)s[0] * (1.0f / 32768.0f))) : w0 * ((float)s[0] * (1.0f / 32768.0f));
        } else {
            const float* __restrict__ s = static_cast<const float*>(pcm) + (int64_t)b * pcm_stride + i * CH;
            v = CH == 2 ? fmaf(w1, s[1], w0 * s[0]) : w0 * s[0];
        }
        o[i] = v;
    }
}

// bounded spin on the barrier's phase WITHOUT a suspend-time hint: the roles of the GEMM kernel hand stages to each other
// every few hundred cycles, a parked warp wakes too late for that
static __device__ __forceinline__ void fe_mbar_wait_spin(unsigned addr, unsigned parity) {
    unsigned ok = 0;
#pragma unroll 1
    for (int spin = 0; spin < (1 << 26); ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.b32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
        if (ok) return;
    }
    __trap();
}

// S16X2: the rows are interleaved int16 stereo PCM; a 4-byte frame takes the place of a float32 sample in the producers'
// loads and is mixed to mono there (w0 * l / 32768 + w1 * r / 32768), so the usual capture format needs no pre-pass
template <bool S16X2>
__global__ void __launch_bounds__(FeRsTcCfg::NTHREADS, 1) fe_resample_tc_kernel(const FeRsTcParams p) {
    typedef FeRsTcCfg C;
    extern __shared__ __align__(1024) unsigned char tc_smem[];
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int N = p.newf;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(tc_smem);
    const uint32_t stage_bytes = (uint32_t)p.stage_floats * 4u;            // room for the H(hi, lo) of the largest chunk
    const uint32_t bar0 = sbase + (uint32_t)C::NST * stage_bytes;          // full[NST], empty[NST], accf[2], acce[2], tmem slot
    const uint32_t full0 = bar0, empty0 = bar0 + 8u * C::NST, accf0 = bar0 + 16u * C::NST, acce0 = accf0 + 16u;
    uint32_t* const tm_slot = reinterpret_cast<uint32_t*>(tc_smem + (size_t)C::NST * stage_bytes + 16 * C::NST + 32);
    if (threadIdx.x == 0) {
        for (int s = 0; s < C::NST; ++s) {
            fe_mbar_init(full0 + 8u * s, 5);        // four producer warps + the loader's arrive.expect_tx
            fe_mbar_init(empty0 + 8u * s, 1);       // tcgen05.commit of the chunk's MMAs
        }
        for (int a = 0; a < 2; ++a) {
            fe_mbar_init(accf0 + 8u * a, 1);        // tcgen05.commit of the tile's last MMA
            fe_mbar_init(acce0 + 8u * a, 4);        // four epilogue warps
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == C::MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                     ::"r"((uint32_t)__cvta_generic_to_shared(tm_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fe_tc_fence_before();
    __syncthreads();
    fe_tc_fence_after();
    const uint32_t tmem = *tm_slot;
    const int first = blockIdx.x, step = gridDim.x;

    if (warp < 4 * C::PGROUPS) {
        // ------------------------------------------------------------------ producers: one row of the A operand per thread,
        // written straight into Tensor Memory (lane = row, 32 columns of hi and 32 of lo per stage).  PGROUPS groups of four
        // warps take the chunks in turn: a group's global loads (an L2 / HBM round trip of 1.5-3 k cycles) overlap the other
        // groups' conversion and stores (profiles/r02_resample_tc.txt).
        const int group = warp >> 2, r = threadIdx.x & 127;
        const uint32_t lane_base = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
        uint32_t it = 0, s = 0, ph = 0;
        for (int tile = first; tile < p.total_tiles; tile += step) {
            const int b = tile / p.tiles_per_clip, tq = tile - b * p.tiles_per_clip;
            const int64_t len = fe_clip_len(p.lengths, b, p.L_in);
            const float* __restrict__ x = p.x + (int64_t)b * p.x_stride;
            const int64_t g0 = ((int64_t)tq * C::M + r) * p.orig - p.width;
#pragma unroll 1
            for (int c = 0; c < p.nch; ++c, ++it, s = (s + 1 == C::NST ? 0 : s + 1), ph ^= (s == 0 ? 1u : 0u)) {
                if (it % (uint32_t)C::PGROUPS != (uint32_t)group) continue;
                const int64_t g = g0 + (p.k0_dense >= 0 ? p.k0_dense + c * C::KC : p.vtab[c].x);
                float e[C::KC];      // 4-byte elements: float32 samples, or (S16X2) int16 stereo frames as raw bits
                if (g >= 0 && g + C::KC + 4 <= len) {
                    // 36 elements from the 16-byte boundary below g, shifted by g & 3 in registers
                    const float4* __restrict__ q4 = reinterpret_cast<const float4*>(x + (g & ~(int64_t)3));
                    float v[C::KC + 4];
#pragma unroll
                    for (int j = 0; j < C::KC / 4 + 1; ++j) {
                        const float4 t = q4[j];
                        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
                    }
                    const int sh = (int)(g & 3);
#pragma unroll
                    for (int j = 0; j < C::KC; ++j) {
                        const float a01 = (sh & 1) ? v[j + 1] : v[j], a23 = (sh & 1) ? v[j + 3] : v[j + 2];
                        e[j] = (sh & 2) ? a23 : a01;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < C::KC; ++j) {
                        const int64_t i = g + j;
                        e[j] = (i >= 0 && i < len) ? x[i] : 0.0f;      // bits 0 = a frame of two zero samples
                    }
                }
                if (S16X2) {
#pragma unroll
                    for (int j = 0; j < C::KC; ++j) {
                        const uint32_t f = __float_as_uint(e[j]);
                        const float l = (float)(short)(f & 0xffffu) * (1.0f / 32768.0f), rr = (float)(short)(f >> 16) * (1.0f / 32768.0f);
                        e[j] = fmaf(p.w1, rr, p.w0 * l);
                    }
                }
                float hi[C::KC], lo[C::KC];
#pragma unroll
                for (int j = 0; j < C::KC; ++j) { hi[j] = fe_tf32_rna(e[j]); lo[j] = fe_tf32_rna(e[j] - hi[j]); }
                fe_mbar_wait_spin(empty0 + 8u * s, ph ^ 1u);
                fe_tc_fence_after();
                const uint32_t t_hi = lane_base + C::a_hi_col(s), t_lo = lane_base + C::a_lo_col(s);
#pragma unroll
                for (int j = 0; j < C::KC; j += 8) {
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                                 ::"r"(t_hi + j), "f"(hi[j]), "f"(hi[j + 1]), "f"(hi[j + 2]), "f"(hi[j + 3]), "f"(hi[j + 4]), "f"(hi[j + 5]),
                                   "f"(hi[j + 6]), "f"(hi[j + 7]) : "memory");
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                                 ::"r"(t_lo + j), "f"(lo[j]), "f"(lo[j + 1]), "f"(lo[j + 2]), "f"(lo[j + 3]), "f"(lo[j + 4]), "f"(lo[j + 5]),
                                   "f"(lo[j + 6]), "f"(lo[j + 7]) : "memory");
                }
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                fe_tc_fence_before();
                fe_warp_arrive(full0 + 8u * s, lane);
            }
        }
    } else if (warp < 4 * C::PGROUPS + 4) {
        // ------------------------------------------------------------------ epilogue: one output row (block q) per lane
        const int w = warp - 4 * C::PGROUPS, r = 32 * w + lane;
        uint32_t i = 0;
        for (int tile = first; tile < p.total_tiles; tile += step, ++i) {
            const uint32_t a = i & 1u, aph = (i >> 1) & 1u;
            const int b = tile / p.tiles_per_clip, tq = tile - b * p.tiles_per_clip;
            const int64_t len = fe_clip_len(p.lengths, b, p.L_in);
            int64_t olen = (len * p.newf + p.orig - 1) / p.orig;               // samples this clip yields; zeros behind
            if (olen > p.L_out) olen = p.L_out;
            const int64_t o0 = ((int64_t)tq * C::M + r) * N;                   // first output sample of this lane's row
            float* __restrict__ orow = p.out + (int64_t)b * p.out_stride + o0;
            fe_mbar_wait_bounded(accf0 + 8u * a, aph);                        // a whole tile away: park the warp
            fe_tc_fence_after();
            const uint32_t taddr = tmem + ((uint32_t)(32 * w) << 16) + a * (uint32_t)C::ACC_STRIDE;
#pragma unroll 1
            for (int c0 = 0; c0 < N; c0 += 16) {
                uint32_t v[16];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                               "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                             : "r"(taddr + (uint32_t)c0) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                const int64_t i0 = o0 + c0;
                if (p.out_vec && i0 + 16 <= olen) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        *reinterpret_cast<float4*>(orow + c0 + 4 * j) = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                                                                    __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
                } else {
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (i0 + j < p.L_out) orow[c0 + j] = i0 + j < olen ? __uint_as_float(v[j]) : 0.0f;
                }
            }
            fe_tc_fence_before();
            fe_warp_arrive(acce0 + 8u * a, lane);
        }
    } else if (warp == C::MMA_WARP) {
        // ------------------------------------------------------------------ MMA issuer: A from Tensor Memory, H from shared memory
        if (lane == 0) {
            uint32_t s = 0, ph = 0, i = 0;
            for (int tile = first; tile < p.total_tiles; tile += step, ++i) {
                const uint32_t a = i & 1u, aph = (i >> 1) & 1u;
                fe_mbar_wait_spin(acce0 + 8u * a, aph ^ 1u);                   // the epilogue has drained this accumulator
                fe_tc_fence_after();
                const uint32_t d = tmem + a * (uint32_t)C::ACC_STRIDE;
                int4 vnext = p.vtab[0];
#pragma unroll 1
                for (int c = 0; c < p.nch; ++c) {
                    const int4 vc = vnext;
                    if (c + 1 < p.nch) vnext = p.vtab[c + 1];              // the next entry's load overlaps this chunk's wait and issue
                    const uint32_t n = (uint32_t)vc.z, dcol = d + ((uint32_t)vc.y & 0xffffu), first = ((uint32_t)vc.y >> 16) & 1u;
                    const uint32_t idesc = fe_umma_idesc_tf32(C::M, (int)n);
                    fe_mbar_wait_spin(full0 + 8u * s, ph);
                    fe_tc_fence_after();
                    const uint32_t t_hi = tmem + C::a_hi_col(s), t_lo = tmem + C::a_lo_col(s);
                    const uint64_t dbh = fe_umma_desc(sbase + s * stage_bytes), dbl = fe_umma_desc(sbase + s * stage_bytes + n * C::KC * 4u);
#pragma unroll
                    for (int ks = 0; ks < C::KC / 8; ++ks) {
                        const uint64_t ko = (uint64_t)(ks * 2 * C::LBO >> 4);  // the descriptor's address field counts 16-byte units
                        fe_umma_tf32_ts(dcol, t_hi + 8 * ks, dbh + ko, idesc, (first && ks == 0) ? 0u : 1u);
                        fe_umma_tf32_ts(dcol, t_hi + 8 * ks, dbl + ko, idesc, 1u);
                        fe_umma_tf32_ts(dcol, t_lo + 8 * ks, dbh + ko, idesc, 1u);
                    }
                    fe_umma_commit(empty0 + 8u * s);       // arrives when the MMAs issued so far have read this stage
                    if (++s == C::NST) { s = 0; ph ^= 1u; }
                }
                fe_umma_commit(accf0 + 8u * a);
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------------ H loader
        if (lane == 0) {
            uint32_t s = 0, ph = 0;
            for (int tile = first; tile < p.total_tiles; tile += step) {
#pragma unroll 1
                for (int c = 0; c < p.nch; ++c) {
                    const int4 vc = p.vtab[c];
                    const uint32_t bytes = (uint32_t)vc.z * (uint32_t)(2 * C::KC * 4);
                    fe_mbar_wait_spin(empty0 + 8u * s, ph ^ 1u);
                    fe_mbar_expect_tx(full0 + 8u * s, bytes);
                    const float* src = p.hpack + (size_t)vc.w;
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(sbase + s * stage_bytes), "l"(src), "r"(bytes), "r"(full0 + 8u * s) : "memory");
                    if (++s == C::NST) { s = 0; ph ^= 1u; }
                }
            }
        }
        __syncwarp();
    }
    fe_tc_fence_before();
    __syncthreads();
    if (warp == C::MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

// =====================================================================================================
// waveform conditioning: the reference's stereo -> mono rule, and the glue of the one-call PCM -> log-mel path
// =====================================================================================================
// /root/reference/Backend/services/audio_processor.py:307-319, applied to the RESAMPLED channels r0 / r1 of a clip:
//     correlation = sum(r0 * r1) / sqrt(sum(r0 ** 2) * sum(r1 ** 2))
//     audio = mean(r0, r1)                 if correlation > 0.5
//           = mid + 0.3 * side             otherwise   (mid = (r0 + r1) / 2, side = (r0 - r1) / 2)
// One CTA per clip: pass 1 = the three sums (float64 accumulators, fixed reduction order: deterministic), pass 2 = the mix,
// written over r0, with the reference's own float32 operations (no FMA contraction).  r1's second read comes from L2.
__global__ void __launch_bounds__(1024) fe_mono_auto_kernel(float* __restrict__ r0, const float* __restrict__ r1, int64_t stride,
                                                            int64_t n, int32_t* __restrict__ decision, float* __restrict__ corr_out) {
    __shared__ double red[3][32];
    __shared__ int use_mean;
    float* __restrict__ a = r0 + (int64_t)blockIdx.x * stride;
    const float* __restrict__ b = r1 + (int64_t)blockIdx.x * stride;
    double s01 = 0.0, s00 = 0.0, s11 = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const double x = a[i], y = b[i];
        s01 += x * y; s00 += x * x; s11 += y * y;
    }
    for (int o = 16; o > 0; o >>= 1) {
        s01 += __shfl_xor_sync(0xffffffffu, s01, o);
        s00 += __shfl_xor_sync(0xffffffffu, s00, o);
        s11 += __shfl_xor_sync(0xffffffffu, s11, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { red[0][warp] = s01; red[1][warp] = s00; red[2][warp] = s11; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t01 = 0.0, t00 = 0.0, t11 = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { t01 += red[0][w]; t00 += red[1][w]; t11 += red[2][w]; }
        // 0 / 0 = NaN for a silent channel, and NaN > 0.5 is false: the mid / side branch, as in the reference
        const float c = (float)(t01 / sqrt(t00 * t11));
        use_mean = c > 0.5f ? 1 : 0;
        if (decision) decision[blockIdx.x] = use_mean;
        if (corr_out) corr_out[blockIdx.x] = c;
    }
    __syncthreads();
    const bool mean = use_mean != 0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float x = a[i], y = b[i];
        const float mid = __fmul_rn(__fadd_rn(x, y), 0.5f);                       // torch.mean over 2 rows == (x + y) / 2
        a[i] = mean ? mid : __fadd_rn(mid, __fmul_rn(0.3f, __fmul_rn(__fsub_rn(x, y), 0.5f)));
    }
}

// input frames per clip -> resampled samples per clip, ceil(new * len / orig) with the reduced ratio (fe_rs_out_len)
__global__ void fe_rs_lengths_kernel(const int64_t* __restrict__ in_len, int64_t* __restrict__ out_len, int32_t B, int64_t L_in,
                                     int64_t orig_r, int64_t new_r, int64_t L_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    int64_t n = in_len[i];
    n = n < 0 ? 0 : (n > L_in ? L_in : n);
    int64_t o = (new_r * n + orig_r - 1) / orig_r;
    out_len[i] = o > L_out ? L_out : o;
}

// =====================================================================================================
// spectral gate (fe_gate.cuh)
// =====================================================================================================
__global__ void __launch_bounds__(FeGCfg::NT, 2) fe_gate_stft_kernel(const FeCosyParams p, const FeGateParams g) {
    typedef FeGCfg C;
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x, cta = blockIdx.x;
    fe_c_phase_load<C>(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_a<C>(p, sm, tid, cta);
    __syncthreads();
    fe_c_phase_b<C>(p, sm, tid, cta);
    __syncthreads();
    fe_g_phase_c_spec(g, p, sm, tid, cta);
}

// One warp per (clip, bin): the order statistics lo = floor(0.1 (n - 1)) and lo + 1 of the bin's n magnitudes by a 4-pass
// radix select on the float bits (magnitudes are >= 0, so the bit patterns order like the values), then torch.quantile's
// linear interpolation (torch.lerp's two-sided form).
// (A bit-by-bit select without atomics was measured: 4.1 ms against 2.2 ms per 256 x 30 s for this histogram form -- 31 passes
// over the row cost more than the serialised atomics of the top-byte pass.)
#define FE_GATE_Q_CACHE 2048      // magnitudes of a row kept in shared memory (30-s clips have 1876 frames); longer rows re-read the spectrum
__global__ void __launch_bounds__(256) fe_gate_quantile_kernel(const FeGateParams g) {
    extern __shared__ __align__(16) float qsm[];          // [8 warps][FE_GATE_Q_CACHE] magnitudes
    __shared__ int hist[8][256];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t row = (int64_t)blockIdx.x * 8 + w;
    if (row >= (int64_t)g.B * 513) return;
    const int b = (int)(row / 513);
    const int n = (int)FeGateTraits::frames(fe_clip_len(g.lengths, b, g.L), 0);
    if (n <= 0) { if (lane == 0) g.nf[row] = 0.0f; return; }
    const float2* __restrict__ s = g.spec + row * g.Ts;
    float* const mags = qsm + w * FE_GATE_Q_CACHE;
    const bool cached = n <= FE_GATE_Q_CACHE;
    if (cached) {
        for (int t = lane; t < n; t += 32) { const float2 x = s[t]; mags[t] = sqrtf(x.x * x.x + x.y * x.y); }
        __syncwarp();
    }
    auto mag_at = [&](int t) -> float {
        if (cached) return mags[t];
        const float2 x = s[t];
        return sqrtf(x.x * x.x + x.y * x.y);
    };
    const float pos = 0.1f * (float)(n - 1);
    const int lo = (int)floorf(pos);
    const float frac = pos - (float)lo;
    unsigned prefix = 0, mask = 0;
    int k = lo;                                           // rank still to find among the elements that match `prefix` under `mask`
    int* const h = hist[w];
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = lane; i < 256; i += 32) h[i] = 0;
        __syncwarp();
        for (int t = lane; t < n; t += 32) {
            const unsigned u = __float_as_uint(mag_at(t));
            if ((u & mask) == prefix) atomicAdd(&h[(u >> shift) & 255u], 1);
        }
        __syncwarp();
        // bucket holding rank k: prefix sums over the 256 counters, 8 per lane
        int c[8], tot = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) { c[j] = h[8 * lane + j]; tot += c[j]; }
        int incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        int before = incl - tot, found = -1, kk = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (found < 0 && k < before + c[j]) { found = 8 * lane + j; kk = k - before; }
            before += c[j];
        }
        const unsigned vote = __ballot_sync(0xffffffffu, found >= 0);
        const int src = __ffs(vote) - 1;
        found = __shfl_sync(0xffffffffu, found, src);
        kk = __shfl_sync(0xffffffffu, kk, src);
        prefix |= (unsigned)found << shift;
        mask |= 255u << shift;
        k = kk;
        __syncwarp();
    }
    const float v_lo = __uint_as_float(prefix);
    // v_hi = element of rank lo + 1: v_lo again if enough copies of it remain above rank lo, else the smallest larger element
    int le = 0;
    float nxt = 3.0e38f;
    for (int t = lane; t < n; t += 32) {
        const float m = mag_at(t);
        if (m <= v_lo) ++le; else nxt = fminf(nxt, m);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { le += __shfl_xor_sync(0xffffffffu, le, o); nxt = fminf(nxt, __shfl_xor_sync(0xffffffffu, nxt, o)); }
    if (lane == 0) {
        const float v_hi = (lo + 1 >= n || le > lo + 1) ? v_lo : nxt;
        const float d = v_hi - v_lo;
        g.nf[row] = frac < 0.5f ? v_lo + frac * d : v_hi - d * (1.0f - frac);
    }
}

__global__ void __launch_bounds__(FeGICfg::NT, 1) fe_gate_inverse_kernel(const FeGateParams g) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float red[FeGICfg::NW];
    const int tid = threadIdx.x, cta = blockIdx.x;
    fe_gi_phase_load(g, sm, tid, cta);
    __syncthreads();
    fe_gi_phase_1(sm, tid);
    __syncthreads();
    fe_gi_phase_2(sm, tid);
    __syncthreads();
    fe_gi_phase_3(sm, tid);
    __syncthreads();
    float mx = fe_gi_phase_ola(g, sm, tid, cta);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((tid & 31) == 0) red[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) {
        for (int i = 1; i < FeGICfg::NW; ++i) mx = fmaxf(mx, red[i]);
        atomicMax(g.peak + cta / g.tiles_per_clip, __float_as_int(mx));      // mx >= 0: the int order is the float order
    }
}

__global__ void __launch_bounds__(256) fe_gate_scale_kernel(const FeGateParams g) {
    const int b = blockIdx.y;
    const float peak = __int_as_float(g.peak[b]);
    if (!(peak > 0.0f)) return;                                              // audio_processor.py:356 `if max_amp > 0`
    float* __restrict__ o = g.out + (int64_t)b * g.out_stride;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g.L_out; i += (int64_t)gridDim.x * blockDim.x)
        o[i] = o[i] / peak * 0.95f;                                          // `audio / max_amp * 0.95`, in that order
}

// the eight instantiations (mel bins x output dtype x input dtype) behind one switch
template <class F>
static int whisper2_for_each(F f) {
    if (int rc = f(fe_whisper2_kernel<80, false, false>, 0)) return rc;
    if (int rc = f(fe_whisper2_kernel<80, false, true>, 1)) return rc;
    if (int rc = f(fe_whisper2_kernel<80, true, false>, 2)) return rc;
    if (int rc = f(fe_whisper2_kernel<80, true, true>, 3)) return rc;
    if (int rc = f(fe_whisper2_kernel<128, false, false>, 4)) return rc;
    if (int rc = f(fe_whisper2_kernel<128, false, true>, 5)) return rc;
    if (int rc = f(fe_whisper2_kernel<128, true, false>, 6)) return rc;
    if (int rc = f(fe_whisper2_kernel<128, true, true>, 7)) return rc;
    return 0;
}
static int whisper2_smem_attr() {
    return whisper2_for_each([](void (*k)(const FeWhisper2Params), int) {
        return cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, FeW2Cfg::SMEM_BYTES) == cudaSuccess ? 0 : 1;
    });
}
static void launch_whisper2(int n_mels, bool out16, bool in16, unsigned grid, cudaStream_t st, const FeWhisper2Params& p) {
    const int want = (n_mels == 128 ? 4 : 0) + (out16 ? 2 : 0) + (in16 ? 1 : 0);
    whisper2_for_each([&](void (*k)(const FeWhisper2Params), int id) {
        if (id == want) k<<<grid, FeW2Cfg::NT, FeW2Cfg::SMEM_BYTES, st>>>(p);
        return 0;
    });
}

// the Tensor-Memory kernel: four instantiations (mel bins x output dtype); float32 input only
typedef void (*FeW3Kernel)(const FeWhisper2Params, const CUtensorMap);
template <class F>
static int whisper3_for_each(F f) {
    if (int rc = f(fe_whisper3_kernel<80, false, false>, 0)) return rc;
    if (int rc = f(fe_whisper3_kernel<80, true, false>, 1)) return rc;
    if (int rc = f(fe_whisper3_kernel<128, false, false>, 2)) return rc;
    if (int rc = f(fe_whisper3_kernel<128, true, false>, 3)) return rc;
    if (int rc = f(fe_whisper3_kernel<80, false, true>, 4)) return rc;
    if (int rc = f(fe_whisper3_kernel<80, true, true>, 5)) return rc;
    if (int rc = f(fe_whisper3_kernel<128, false, true>, 6)) return rc;
    if (int rc = f(fe_whisper3_kernel<128, true, true>, 7)) return rc;
    return 0;
}
static int whisper3_smem_attr() {
    return whisper3_for_each([](FeW3Kernel k, int) {
        return cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, FeW3Cfg::SMEM_BYTES) == cudaSuccess ? 0 : 1;
    });
}
static void launch_whisper3(int n_mels, bool out16, unsigned grid, cudaStream_t st, const FeWhisper2Params& p, const CUtensorMap& tmap) {
    // padding >= 2 * PADL: units of pure padding exist and are skipped (fe_w3_silent); the variant without that test otherwise
    const int want = (n_mels == 128 ? 2 : 0) + (out16 ? 1 : 0) + (p.padding >= 2 * FeW3Cfg::PADL ? 4 : 0);
    whisper3_for_each([&](FeW3Kernel k, int id) {
        if (id == want) k<<<grid, FeW3Cfg::NT, FeW3Cfg::SMEM_BYTES, st>>>(p, tmap);
        return 0;
    });
}
// Tensor map of the waveform batch for the kernel's box copies: clip b = rows of PITCH floats starting every HOP floats
// (dimension 1) from sample TMAP_ROW0 * HOP - PADL of the clip, clips `wave_stride` floats apart (dimension 2).  The driver entry
// point is resolved at run time (no link against libcuda).  Returns false when the map cannot be built: the kernel then
// falls back to its row-by-row copies.
static bool whisper3_make_tmap(const float* wave, int64_t wave_stride, int64_t L, int32_t B, CUtensorMap* map, int32_t* rows_out) {
    typedef FeW3Cfg C;
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) fn = nullptr;
        return reinterpret_cast<EncodeFn>(fn);
    }();
    memset(map, 0, sizeof(*map));
    *rows_out = 0;
    if (!encode || getenv("B200FEAT_NO_TMAP")) return false;
    const int64_t first = (int64_t)C::TMAP_ROW0 * C::HOP - C::PADL;           // clip sample of (row 0, column 0)
    const int64_t rows = L - first - C::PITCH < 0 ? 0 : (L - first - C::PITCH) / C::HOP + 1;   // rows that lie inside the clip
    if (rows < C::ROWS || rows > 0x7fffffff || B < 1) return false;
    cuuint64_t gdim[3] = {(cuuint64_t)C::PITCH, (cuuint64_t)rows, (cuuint64_t)B};
    cuuint64_t gstr[2] = {(cuuint64_t)C::HOP * 4, (cuuint64_t)wave_stride * 4};
    cuuint32_t box[3] = {(cuuint32_t)C::PITCH, (cuuint32_t)C::ROWS, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    const CUresult rc = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(wave + first), gdim, gstr, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) return false;
    *rows_out = (int32_t)rows;
    return true;
}
// the same kind of map for the kaldi kernel: rows start at sample 0 of the clip (snip_edges framing, no padding)
static bool kaldi3_make_tmap(const float* wave, int64_t wave_stride, int64_t L, int32_t B, CUtensorMap* map, int32_t* rows_out) {
    typedef FeK3Cfg C;
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) fn = nullptr;
        return reinterpret_cast<EncodeFn>(fn);
    }();
    memset(map, 0, sizeof(*map));
    *rows_out = 0;
    if (!encode || getenv("B200FEAT_NO_TMAP")) return false;
    const int64_t rows = L < C::PITCH ? 0 : (L - C::PITCH) / C::HOP + 1;      // rows whose 164 floats lie inside the clip's buffer
    if (rows < C::ROWS || rows > 0x7fffffff || B < 1) return false;
    cuuint64_t gdim[3] = {(cuuint64_t)C::PITCH, (cuuint64_t)rows, (cuuint64_t)B};
    cuuint64_t gstr[2] = {(cuuint64_t)C::HOP * 4, (cuuint64_t)wave_stride * 4};
    cuuint32_t box[3] = {(cuuint32_t)C::PITCH, (cuuint32_t)C::ROWS, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    const CUresult rc = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(wave), gdim, gstr, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) return false;
    *rows_out = (int32_t)rows;
    return true;
}
// B200FEAT_KALDI_KERNEL=1 forces the shared-memory kernel of round 1 (A/B timing)
static int kaldi_kernel_choice() {      // read per call: tests switch it inside one process
    const char* s = getenv("B200FEAT_KALDI_KERNEL");
    return s ? atoi(s) : 3;
}
static constexpr int FE_K3_CMN_SMEM_ROWS = 640;      // (T, 80) float32 tiles up to 6.4 s stay in shared memory (200 KB)

// B200FEAT_WHISPER_KERNEL=2 forces the shared-memory frame-pair kernel (A/B timing, profiles/); default: the
// Tensor-Memory kernel wherever it applies (float32 input, 16-byte aligned rows)
static int whisper_kernel_choice() {
    static const int v = [] { const char* s = getenv("B200FEAT_WHISPER_KERNEL"); return s ? atoi(s) : 3; }();
    return v;
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

int g_num_sms[64];

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
        if (major != 10) {
            state[dev] = -1;
        } else {
            // state stays 0 until every attribute call succeeded: a failure here is reported now AND retried by the next
            // call (it used to be latched as "ok", and later launches failed with an opaque "invalid argument")
            if (whisper2_smem_attr()) return fail(B200FEAT_ECUDA, "cudaFuncSetAttribute(fe_whisper2_kernel, max dynamic shared memory) failed");
            if (whisper3_smem_attr()) return fail(B200FEAT_ECUDA, "cudaFuncSetAttribute(fe_whisper3_kernel, max dynamic shared memory) failed");
            FE_CUDA(cudaDeviceGetAttribute(&g_num_sms[dev < 64 ? dev : 0], cudaDevAttrMultiProcessorCount, dev));
            FE_CUDA(cudaFuncSetAttribute(fe_kaldi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeKCfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_kaldi3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeK3Cfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_kaldi3_cmn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         FE_K3_CMN_SMEM_ROWS * (FeK3Cfg::NMEL / 2) * 4));
            FE_CUDA(cudaFuncSetAttribute(fe_cosy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeCCfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_wav2lip_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeLCfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_gate_stft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeGCfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_gate_inverse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FeGICfg::SMEM_BYTES));
            FE_CUDA(cudaFuncSetAttribute(fe_gate_quantile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * FE_GATE_Q_CACHE * 4));
            state[dev] = 1;
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

#ifdef FE_W3_TRACE
int b200feat_debug_trace3(long long* host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, fe_w3_trace_buf, sizeof(long long) * FE_T3_UNITS * 16 * FE_T3_EV);
}
#endif
#ifdef FE_W2_TRACE
int b200feat_debug_trace(long long* host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, fe_w2_trace_buf, sizeof(long long) * FE_TR_UNITS * 20 * FE_TR_EV);
}
#endif

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
    case B200FEAT_KIND_WHISPER: return (n + padding) / FeW2Cfg::HOP;
    case B200FEAT_KIND_COSY: return (n + 2 * FeCCfg::PADL < FeCCfg::N) ? 0 : 1 + (n + 2 * FeCCfg::PADL - FeCCfg::N) / FeCCfg::HOP;
    case B200FEAT_KIND_KALDI: return n < FeKCfg::WIN ? 0 : 1 + (n - FeKCfg::WIN) / FeKCfg::HOP;
    default: return -1;
    }
}

int64_t b200feat_workspace_bytes(int32_t kind, int32_t B, int64_t L) {
    if (B < 0 || L < 0) return -1;
    switch (kind) {
    case B200FEAT_KIND_WHISPER: {
        // L here = samples INCLUDING the padding of the call (or T_out * 160 if the caller's rows are longer)
        const int64_t units = (L / FeW2Cfg::HOP + 1 + FeW2Cfg::FT - 1) / FeW2Cfg::FT;
        return 16 + (((int64_t)B * 4 + 15) & ~(int64_t)15) + (int64_t)B * units * 8;
    }
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
static int whisper_logmel_impl(const void* wave, bool in_s16, const int64_t* lengths, int32_t B, int64_t L,
                               int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                               void* out, bool out16, int64_t out_batch_stride, int64_t T_out,
                               void* workspace, int64_t workspace_bytes, void* cuda_stream,
                               int32_t win_frames = 0, int32_t n_windows = 0) {
    if (int rc = common_checks("whisper_logmel", wave, out, B, L, wave_stride)) return rc;
    if (n_mels != 80 && n_mels != 128) return fail(B200FEAT_EINVAL, "whisper_logmel: Unsupported n_mels: %d", n_mels);
    if (padding < 0) return fail(B200FEAT_EINVAL, "whisper_logmel: padding %d < 0", padding);
    if (norm_mode != FE_NORM_PER_CLIP && norm_mode != FE_NORM_PER_CALL) return fail(B200FEAT_EINVAL, "whisper_logmel: bad norm_mode %d", norm_mode);
    const int64_t T_full = (L + padding) / FeW2Cfg::HOP;
    if (win_frames == 0 && T_out < T_full) return fail(B200FEAT_EINVAL, "whisper_logmel: T_out %lld < %lld frames", (long long)T_out, (long long)T_full);
    if (out_batch_stride < (int64_t)n_mels * T_out) return fail(B200FEAT_EINVAL, "whisper_logmel: out_batch_stride too small");
    if (L + padding <= FeW2Cfg::PADL && B > 0 && T_out > 0)
        return fail(B200FEAT_EINVAL, "whisper_logmel: reflect padding (200) needs more than 200 samples, got %lld", (long long)(L + padding));
    int64_t units = (T_out + FeW2Cfg::FT - 1) / FeW2Cfg::FT;
    if (win_frames > 0) {
        // windowed output (b200feat_whisper_logmel_windows): complete slices of the padded mel, all of the longest clip's
        // content covered (the frames behind the last window are transformed for the clip's maximum, not stored)
        if (T_out != (int64_t)win_frames * n_windows) return fail(B200FEAT_EINVAL, "whisper_logmel_windows: internal T_out");
        if ((int64_t)n_windows * win_frames > T_full)
            return fail(B200FEAT_EINVAL, "whisper_logmel_windows: %d windows of %d frames exceed the %lld frames of the padded clip "
                        "(whisper.transcribe pads with padding = N_SAMPLES)", n_windows, win_frames, (long long)T_full);
        if ((int64_t)n_windows * win_frames < L / FeW2Cfg::HOP)
            return fail(B200FEAT_EINVAL, "whisper_logmel_windows: %d windows of %d frames do not cover the clip's %lld content frames",
                        n_windows, win_frames, (long long)(L / FeW2Cfg::HOP));
        units = (int64_t)n_windows * ((win_frames + FeW3Cfg::UF - 1) / FeW3Cfg::UF) + 1;
    }
    const int64_t keys_bytes = 16 + (((int64_t)B * 4 + 15) & ~(int64_t)15);
    if (workspace == nullptr || workspace_bytes < keys_bytes + (int64_t)B * units * 8)
        return fail(B200FEAT_EWORKSPACE, "whisper_logmel: workspace too small (%lld bytes, need %lld)",
                    (long long)workspace_bytes, (long long)(keys_bytes + (int64_t)B * units * 8));
    if (B == 0 || T_out == 0) return 0;
    if (int rc = check_device_arch()) return rc;
    if ((int64_t)B * units > 0x7fffffff) return fail(B200FEAT_EINVAL, "whisper_logmel: batch too large");
    if (T_out >= (1 << 24)) return fail(B200FEAT_EINVAL, "whisper_logmel: T_out %lld >= 2^24", (long long)T_out);
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    int dev = 0;
    FE_CUDA(cudaGetDevice(&dev));

    FeWhisper2Params p;
    memset(&p, 0, sizeof(p));
    p.wave = static_cast<const float*>(wave); p.in_s16 = in_s16 ? 1 : 0;
    p.lengths = lengths; p.wave_stride = wave_stride; p.L = L; p.B = B;
    p.padding = padding; p.n_mels = n_mels; p.norm_mode = norm_mode;
    p.out = out; p.out_batch_stride = out_batch_stride; p.T_out = T_out;
    p.units_per_clip = (int32_t)units;
    p.total_units = (int32_t)(B * units);
    p.vec_ok = (aligned16(wave) && wave_stride % 4 == 0) ? 1 : 0;
    if (aligned16(out) && out_batch_stride % 4 == 0) p.vec_ok |= 2;     // 128-bit (float16: 64-bit) row stores possible
    p.max_keys = reinterpret_cast<int32_t*>(workspace);
    p.unit_minmax = reinterpret_cast<float2*>(reinterpret_cast<char*>(workspace) + keys_bytes);
    const int sms = g_num_sms[dev < 64 ? dev : 0] > 0 ? g_num_sms[dev < 64 ? dev : 0] : 148;
    const bool v3 = !in_s16 && (p.vec_ok & 1) && (whisper_kernel_choice() != 2 || win_frames > 0);
    if (win_frames > 0 && !v3)
        return fail(B200FEAT_EINVAL, "whisper_logmel_windows: needs float32 samples in 16-byte aligned rows (wave pointer and wave_stride * 4 multiples of 16)");
    if (v3) {
        // Tensor-Memory kernel: 128-frame units; their (max, -min) keys are accumulated with atomicMax, so the unit
        // table is initialised together with the per-clip keys
        const int64_t units3 = win_frames > 0 ? units : (T_out + FeW3Cfg::UF - 1) / FeW3Cfg::UF;
        p.units_per_clip = (int32_t)units3;
        p.total_units = (int32_t)(B * units3);
        p.win_frames = win_frames; p.n_windows = n_windows;
        FE_CUDA(cudaMemsetAsync(workspace, FE_KEY_INIT_BYTE, (size_t)(keys_bytes + (int64_t)B * units3 * 8), st));
        {
            const unsigned grid = (unsigned)(p.total_units < sms ? p.total_units : sms);
            p.step_w = (int32_t)(grid / (unsigned)B); p.step_b = (int32_t)(grid % (unsigned)B);
            CUtensorMap tmap;
            p.use_tmap = whisper3_make_tmap(p.wave, wave_stride, L, B, &tmap, &p.tmap_rows) ? 1 : 0;
            ProfScope ps(st, 0);
            launch_whisper3(n_mels, out16, grid, st, p, tmap);
            g_launches++;
        }
        {
            const int want = sms * 4;
            const unsigned grid = (unsigned)(p.total_units < want ? p.total_units : want);
            ProfScope ps(st, 1);
            if (n_mels == 80 && !out16) fe_whisper3_fixup_kernel<80, false><<<grid, 256, 0, st>>>(p);
            else if (n_mels == 80) fe_whisper3_fixup_kernel<80, true><<<grid, 256, 0, st>>>(p);
            else if (!out16) fe_whisper3_fixup_kernel<128, false><<<grid, 256, 0, st>>>(p);
            else fe_whisper3_fixup_kernel<128, true><<<grid, 256, 0, st>>>(p);
            g_launches++;
        }
        FE_CUDA(cudaGetLastError());
        return 0;
    }
    FE_CUDA(cudaMemsetAsync(workspace, FE_KEY_INIT_BYTE, (size_t)keys_bytes, st));
    {
        const unsigned grid = (unsigned)(p.total_units < sms ? p.total_units : sms);
        ProfScope ps(st, 0);
        launch_whisper2(n_mels, out16, in_s16, grid, st, p);
        g_launches++;
    }
    {
        // 64 units per CTA-pass would leave most SMs idle for small batches: spread the checks over all SMs
        const int need = (p.total_units + 255) / 256;
        const unsigned grid = (unsigned)(need < sms * 8 ? need : sms * 8);
        ProfScope ps(st, 1);
        if (n_mels == 80 && !out16) fe_whisper2_fixup_kernel<80, false><<<grid, 256, 0, st>>>(p);
        else if (n_mels == 80) fe_whisper2_fixup_kernel<80, true><<<grid, 256, 0, st>>>(p);
        else if (!out16) fe_whisper2_fixup_kernel<128, false><<<grid, 256, 0, st>>>(p);
        else fe_whisper2_fixup_kernel<128, true><<<grid, 256, 0, st>>>(p);
        g_launches++;
    }
    FE_CUDA(cudaGetLastError());
    return 0;
}

int b200feat_whisper_logmel(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                            int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                            float* out, int64_t out_batch_stride, int64_t T_out,
                            void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    return whisper_logmel_impl(wave, false, lengths, B, L, wave_stride, n_mels, padding, norm_mode, out, false,
                               out_batch_stride, T_out, workspace, workspace_bytes, cuda_stream);
}

static int whisper_ex_dtypes(const char* who, int32_t wave_dtype, int32_t out_dtype) {
    if (wave_dtype != B200FEAT_F32 && wave_dtype != B200FEAT_S16)
        return fail(B200FEAT_EINVAL, "%s: wave_dtype %d (B200FEAT_F32 or B200FEAT_S16)", who, wave_dtype);
    if (out_dtype != B200FEAT_F32 && out_dtype != B200FEAT_F16)
        return fail(B200FEAT_EINVAL, "%s: out_dtype %d (B200FEAT_F32 or B200FEAT_F16)", who, out_dtype);
    return 0;
}

int b200feat_whisper_logmel_ex(const void* wave, int32_t wave_dtype, const int64_t* lengths, int32_t B, int64_t L,
                               int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                               void* out, int32_t out_dtype, int64_t out_batch_stride, int64_t T_out,
                               void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    if (int rc = whisper_ex_dtypes("whisper_logmel_ex", wave_dtype, out_dtype)) return rc;
    return whisper_logmel_impl(wave, wave_dtype == B200FEAT_S16, lengths, B, L, wave_stride, n_mels, padding, norm_mode, out,
                               out_dtype == B200FEAT_F16, out_batch_stride, T_out, workspace, workspace_bytes, cuda_stream);
}

int b200feat_whisper_logmel_windows(const float* wave, const int64_t* lengths, int32_t B, int64_t L, int64_t wave_stride,
                                    int32_t n_mels, int32_t padding, int32_t norm_mode, int32_t win_frames, int32_t n_windows,
                                    void* out, int32_t out_dtype, int64_t out_batch_stride,
                                    void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    if (int rc = whisper_ex_dtypes("whisper_logmel_windows", B200FEAT_F32, out_dtype)) return rc;
    if (win_frames <= 0 || n_windows <= 0 || (int64_t)win_frames * n_windows >= (1 << 24))
        return fail(B200FEAT_EINVAL, "whisper_logmel_windows: win_frames %d, n_windows %d must be positive (and their product < 2^24)", win_frames, n_windows);
    if (out_batch_stride < (int64_t)n_windows * n_mels * win_frames)
        return fail(B200FEAT_EINVAL, "whisper_logmel_windows: out_batch_stride < n_windows * n_mels * win_frames");
    return whisper_logmel_impl(wave, false, lengths, B, L, wave_stride, n_mels, padding, norm_mode, out, out_dtype == B200FEAT_F16,
                               out_batch_stride, (int64_t)win_frames * n_windows, workspace, workspace_bytes, cuda_stream,
                               win_frames, n_windows);
}

// ------------------------------------------------------------------------------------------ Whisper, no allocation
// The workspace of b200feat_whisper_logmel (per-clip maximum keys + per-unit (max, -min) keys) as a file-scope device
// array, for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every
// GPU that loads the library holds these bytes until exit.  In HOST code the name below is only a stand-in for the
// symbol: the device address comes from cudaGetSymbolAddress, once per device, with that device current.
#define FE_NOALLOC_WS_BYTES (1 << 20)
static __device__ __align__(256) unsigned char fe_noalloc_ws[FE_NOALLOC_WS_BYTES];

static int noalloc_ws_ptr(void** out) {
    static std::mutex mu;
    static void* ptrs[64];
    int dev = 0;
    FE_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(B200FEAT_EINVAL, "whisper_logmel_noalloc: device index %d not supported (< 64)", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (ptrs[dev] == nullptr) {
        void* q = nullptr;
        FE_CUDA(cudaGetSymbolAddress(&q, fe_noalloc_ws));      // `dev` is current: the address is this device's copy
        if (q == nullptr) return fail(B200FEAT_ECUDA, "whisper_logmel_noalloc: cudaGetSymbolAddress returned a null address");
        ptrs[dev] = q;
    }
    *out = ptrs[dev];
    return 0;
}

int64_t b200feat_noalloc_workspace_max(void) { return FE_NOALLOC_WS_BYTES; }

int b200feat_whisper_logmel_noalloc(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                                    int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                                    float* out, int64_t out_batch_stride, int64_t T_out, void* cuda_stream) {
    if (B < 0 || L < 0 || padding < 0) return fail(B200FEAT_EINVAL, "whisper_logmel_noalloc: negative B, L or padding");
    const int64_t need = b200feat_workspace_bytes(B200FEAT_KIND_WHISPER, B, (T_out > 0 ? T_out : 0) * FeW2Cfg::HOP);
    const int64_t need_l = b200feat_workspace_bytes(B200FEAT_KIND_WHISPER, B, L + padding);
    const int64_t want = need > need_l ? need : need_l;
    if (want < 0 || want > FE_NOALLOC_WS_BYTES)
        return fail(B200FEAT_EWORKSPACE, "whisper_logmel_noalloc: this call needs %lld bytes of scratch, the built-in array holds %d "
                    "(B x ceil(frames / 64) x 8 + 4 x B + 32 <= %d); split the batch or use b200feat_whisper_logmel with a "
                    "caller-owned workspace", (long long)want, FE_NOALLOC_WS_BYTES, FE_NOALLOC_WS_BYTES);
    void* ws = nullptr;
    if (B > 0 && T_out > 0) {
        if (int rc = check_device_arch()) return rc;
        if (int rc = noalloc_ws_ptr(&ws)) return rc;
    } else {
        static unsigned char dummy;      // nothing is launched for an empty call; the shared path only checks non-null
        ws = &dummy;
    }
    // the shared path clears the part of the array it uses (cudaMemsetAsync on cuda_stream) before the kernels run, so
    // the previous call's contents never leak into this one; it allocates nothing
    return whisper_logmel_impl(wave, false, lengths, B, L, wave_stride, n_mels, padding, norm_mode, out, false,
                               out_batch_stride, T_out, ws, FE_NOALLOC_WS_BYTES, cuda_stream);
}

// ------------------------------------------------------------------------------------------ CosyVoice
static int cosy_mel_impl(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                         int64_t wave_stride, float* out, bool frame_major, int64_t out_batch_stride, int64_t T_out,
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
    p.out = out; p.out_batch_stride = out_batch_stride; p.T_out = T_out; p.frame_major = frame_major ? 1 : 0;
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

int b200feat_cosy_mel(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                      int64_t wave_stride, float* out, int64_t out_batch_stride, int64_t T_out,
                      void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    return cosy_mel_impl(wave, lengths, B, L, wave_stride, out, false, out_batch_stride, T_out, workspace, workspace_bytes, cuda_stream);
}

int b200feat_cosy_mel_frames(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                             int64_t wave_stride, float* out, int64_t out_batch_stride, int64_t T_out,
                             void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    return cosy_mel_impl(wave, lengths, B, L, wave_stride, out, true, out_batch_stride, T_out, workspace, workspace_bytes, cuda_stream);
}

// ------------------------------------------------------------------------------------------ Wav2Lip / Diff2Lip mel
int64_t b200feat_wav2lip_num_frames(int64_t n_samples) { return n_samples < 0 ? -1 : 1 + n_samples / FeLCfg::HOP; }

int b200feat_wav2lip_mel(const float* wave, const int64_t* lengths, int32_t B, int64_t L, int64_t wave_stride,
                         int32_t pad_mode, int32_t frame_major, float* out, int64_t out_batch_stride, int64_t T_out,
                         void* cuda_stream) {
    if (int rc = common_checks("wav2lip_mel", wave, out, B, L, wave_stride)) return rc;
    if (pad_mode != B200FEAT_PAD_REFLECT && pad_mode != B200FEAT_PAD_ZEROS)
        return fail(B200FEAT_EINVAL, "wav2lip_mel: pad_mode %d (B200FEAT_PAD_REFLECT or B200FEAT_PAD_ZEROS)", pad_mode);
    const int64_t T_full = b200feat_wav2lip_num_frames(L);
    if (T_out < T_full) return fail(B200FEAT_EINVAL, "wav2lip_mel: T_out %lld < %lld frames", (long long)T_out, (long long)T_full);
    if (out_batch_stride < (int64_t)FeLCfg::NMEL * T_out) return fail(B200FEAT_EINVAL, "wav2lip_mel: out_batch_stride too small");
    if (pad_mode == B200FEAT_PAD_REFLECT && L <= FeLCfg::PADL && B > 0)
        return fail(B200FEAT_EINVAL, "wav2lip_mel: reflect padding (400) needs more than 400 samples, got %lld", (long long)L);
    if (B == 0 || T_out == 0) return 0;
    if (int rc = check_device_arch()) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    FeCosyParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = wave_stride; p.L = L; p.B = B;
    p.out = out; p.out_batch_stride = out_batch_stride; p.T_out = T_out; p.frame_major = frame_major ? 1 : 0;
    p.pad_mode = pad_mode;
    const int64_t tiles = (T_out + FeLCfg::FT - 1) / FeLCfg::FT;
    p.tiles_per_clip = (int32_t)tiles;
    p.vec_ok = (aligned16(wave) && wave_stride % 4 == 0) ? 1 : 0;
    const int64_t per_launch = 0x7fffffff / tiles;
    if (per_launch < 1) return fail(B200FEAT_EINVAL, "wav2lip_mel: clip too long");
    for (int64_t c0 = 0; c0 < B; c0 += per_launch) {
        const int64_t nc = (B - c0 < per_launch) ? (B - c0) : per_launch;
        p.clip0 = (int32_t)c0; p.nclips = (int32_t)nc;
        ProfScope ps(st, 0);
        fe_wav2lip_kernel<<<(unsigned)(nc * tiles), FeLCfg::NT, FeLCfg::SMEM_BYTES, st>>>(p);
        g_launches++;
    }
    FE_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------ spectral gate
int64_t b200feat_spectral_gate_out_len(int64_t n_samples) { return n_samples < 0 ? -1 : FeGICfg::HOP * (n_samples / FeGICfg::HOP); }

int64_t b200feat_spectral_gate_workspace_bytes(int32_t B, int64_t L) {
    if (B < 0 || L < 0) return -1;
    const int64_t Ts = 1 + L / FeGICfg::HOP;
    auto r = [](int64_t v) { return (v + 255) & ~(int64_t)255; };
    return 256 + r((int64_t)B * 513 * Ts * 8) + r((int64_t)B * 513 * 4) + r((int64_t)B * 4);
}

int b200feat_spectral_gate(const float* wave, const int64_t* lengths, int32_t B, int64_t L, int64_t wave_stride,
                           float* out, int64_t out_stride, void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    if (int rc = common_checks("spectral_gate", wave, out, B, L, wave_stride)) return rc;
    const int64_t L_out = b200feat_spectral_gate_out_len(L);
    if (out_stride < L_out) return fail(B200FEAT_EINVAL, "spectral_gate: out_stride %lld < %lld samples", (long long)out_stride, (long long)L_out);
    if (B > 0 && L <= FeGCfg::PADL)
        return fail(B200FEAT_EINVAL, "spectral_gate: reflect padding (512) needs more than 512 samples, got %lld", (long long)L);
    const int64_t need = b200feat_spectral_gate_workspace_bytes(B, L);
    if (B > 0 && (workspace == nullptr || workspace_bytes < need))
        return fail(B200FEAT_EWORKSPACE, "spectral_gate: workspace too small (%lld bytes, need %lld)", (long long)workspace_bytes, (long long)need);
    if (B == 0 || L_out == 0) return 0;
    if (int rc = check_device_arch()) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int64_t Ts = 1 + L / FeGICfg::HOP;
    auto r256 = [](int64_t v) { return (v + 255) & ~(int64_t)255; };
    char* w = static_cast<char*>(workspace);
    w += (256 - (reinterpret_cast<uintptr_t>(w) & 255)) & 255;
    FeGateParams g;
    memset(&g, 0, sizeof(g));
    g.wave = wave; g.lengths = lengths; g.wave_stride = wave_stride; g.L = L; g.B = B;
    g.spec = reinterpret_cast<float2*>(w); w += r256((int64_t)B * 513 * Ts * 8);
    g.nf = reinterpret_cast<float*>(w); w += r256((int64_t)B * 513 * 4);
    g.peak = reinterpret_cast<int32_t*>(w);
    g.Ts = Ts; g.out = out; g.out_stride = out_stride; g.L_out = L_out;
    g.tiles_per_clip = (int32_t)((L_out / FeGICfg::HOP + 2 + FeGICfg::OUT_HOPS - 1) / FeGICfg::OUT_HOPS);
    FeCosyParams p;
    memset(&p, 0, sizeof(p));
    p.wave = wave; p.lengths = lengths; p.wave_stride = wave_stride; p.L = L; p.B = B; p.T_out = Ts;
    const int64_t tiles = (Ts + FeGCfg::FT - 1) / FeGCfg::FT;
    p.tiles_per_clip = (int32_t)tiles;
    p.vec_ok = (aligned16(wave) && wave_stride % 4 == 0) ? 1 : 0;
    p.clip0 = 0; p.nclips = B;
    if ((int64_t)B * tiles > 0x7fffffff || (int64_t)B * g.tiles_per_clip > 0x7fffffff || B > 65535)
        return fail(B200FEAT_EINVAL, "spectral_gate: batch too large");
    FE_CUDA(cudaMemsetAsync(g.peak, 0, (size_t)B * 4, st));
    {
        ProfScope ps(st, 0);
        fe_gate_stft_kernel<<<(unsigned)(B * tiles), FeGCfg::NT, FeGCfg::SMEM_BYTES, st>>>(p, g);
        g_launches++;
    }
    {
        ProfScope ps(st, 1);
        fe_gate_quantile_kernel<<<(unsigned)(((int64_t)B * 513 + 7) / 8), 256, 8 * FE_GATE_Q_CACHE * 4, st>>>(g);
        g_launches++;
    }
    {
        ProfScope ps(st, 0);
        fe_gate_inverse_kernel<<<(unsigned)((int64_t)B * g.tiles_per_clip), FeGICfg::NT, FeGICfg::SMEM_BYTES, st>>>(g);
        g_launches++;
    }
    {
        int64_t gx = (L_out + 1023) / 1024;
        if (gx > 1024) gx = 1024;
        ProfScope ps(st, 1);
        fe_gate_scale_kernel<<<dim3((unsigned)gx, (unsigned)B), 256, 0, st>>>(g);
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
    {   // Tensor-Memory kernel wherever the waveform can be described by a tensor map (16-byte aligned rows, >= 34 hops)
        const int64_t units3 = (T_out + FeK3Cfg::UF - 1) / FeK3Cfg::UF;
        CUtensorMap tmap;
        if ((p.vec_ok & 1) && kaldi_kernel_choice() != 1 && (int64_t)B * units3 <= 0x7fffffff &&
            kaldi3_make_tmap(wave, wave_stride, L, B, &tmap, &p.tmap_rows)) {
            int dev = 0;
            FE_CUDA(cudaGetDevice(&dev));
            const int sms = g_num_sms[dev < 64 ? dev : 0] > 0 ? g_num_sms[dev < 64 ? dev : 0] : 148;
            p.clip0 = 0; p.nclips = B;
            p.units_per_clip = (int32_t)units3;
            p.total_units = (int32_t)(B * units3);
            // (CMN fused into this kernel -- a CTA takes whole clips and re-reads each from L2 after its last unit -- was
            // measured: 1.47 ms per 4096 x 5 s against 1.38 ms with the separate kernel below, the CTA-wide L2 round trips
            // per clip cost more than the second pass; it also made a clip's result depend on the batch it is in.)
            const unsigned grid = (unsigned)(p.total_units < sms ? p.total_units : sms);
            p.step_w = (int32_t)(grid / (unsigned)B); p.step_b = (int32_t)(grid % (unsigned)B);
            {
                ProfScope ps(st, 0);
                fe_kaldi3_kernel<<<grid, FeK3Cfg::NT, FeK3Cfg::SMEM_BYTES, st>>>(p, tmap);
                g_launches++;
            }
            if (subtract_mean) {
                const int rows = (int)(T_out < FE_K3_CMN_SMEM_ROWS ? T_out : FE_K3_CMN_SMEM_ROWS);
                const bool fits = T_out <= FE_K3_CMN_SMEM_ROWS;
                ProfScope ps(st, 1);
                fe_kaldi3_cmn_kernel<<<2u * (unsigned)B, 320, fits ? (size_t)rows * (FeK3Cfg::NMEL / 2) * 4 : 0, st>>>(p, fits ? rows : 0);
                g_launches++;
            }
            FE_CUDA(cudaGetLastError());
            return 0;
        }
    }
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
            const int total_tiles = (int)(nc * tiles);
            const int tpc = kaldi_tiles_per_cta();
            fe_kaldi_kernel<<<(unsigned)((total_tiles + tpc - 1) / tpc), FeKCfg::NT, FeKCfg::SMEM_BYTES, st>>>(p, total_tiles, tpc);
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

// ------------------------------------------------------------------------------------------ resample
namespace {

// tap tables are library-owned constants: built once per (device, ratio, filter) and kept until unload
struct RsDevicePlan {
    FeRsPlan plan;
    float4* taps = nullptr;
    int32_t* kstart = nullptr;
    // tensor-core path (fe_resample_tc.cuh): the dense tap matrix, split hi / lo, in the B-operand layout
    float* hpack = nullptr;
    int4* vtab = nullptr;
    int nch = 0, tc_m = 0, tc_stage_floats = 0, tc_k0_dense = -1;      // tc_m blocks of the plan form one row of the GEMM
};


// The resampler's GEMM on the tensor cores (fe_resample_tc.cuh).  m consecutive blocks are taken as one (orig * m input samples ->
// new * m outputs, the "phases" of the GEMM), m the largest value with N = new * m a multiple of 16 and <= 160: 44.1 -> 16 kHz
// is 441 -> 160 with m = 1, 48 -> 16 kHz is 3 -> 1 with m = 160 (480 -> 160).  The plan is a table of chunks of 32 taps:
//   dense   one block of all N phases over the whole K = 2 * width + orig * m       (long filters: the reference's kaiser / width 128)
//   banded  blocks of 32 phases, each over the K range its phases' windows cover      (short filters: torchaudio's default)
// whichever multiplies fewer zeros.  Ratios whose phase count does not fit stay on the CUDA-core kernel.
struct RsTcPlan {
    int m = 0, N = 0, orig_e = 0, stage_floats = 0, k0_dense = -1;
    std::vector<int4> vtab;
    std::vector<float> hpack;
};
static float tf32_rna_host(float v) {            // cvt.rna.tf32.f32: nearest, ties away from zero
    uint32_t u;
    memcpy(&u, &v, 4);
    u = (u + 0x1000u) & 0xffffe000u;
    memcpy(&v, &u, 4);
    return v;
}
static bool rs_tc_make(const FeRsPlan& pl, int lpfw, double rolloff, int method, double beta, RsTcPlan* out) {
    typedef FeRsTcCfg C;
    const char* e = getenv("B200FEAT_RESAMPLE_TC");
    const int mode = e ? atoi(e) : 3;                // 0 = off, 1 = dense only, 2 = banded only, 3 = both (default)
    if (mode == 0) return false;
    int m = C::MAX_N / pl.newf;
    while (m > 0 && (pl.newf * m) % 16 != 0) --m;
    if (m <= 0) return false;
    const int N = pl.newf * m, K = 2 * pl.width + pl.orig, Ke = K + pl.orig * (m - 1);
    // the regrouped tap matrix He[pe][ke] and every row's support
    std::vector<float> He((size_t)N * Ke, 0.0f);
    std::vector<int> lo(N, Ke), hi(N, -1);
    for (int p = 0; p < pl.newf; ++p)
        for (int k = 0; k < K; ++k) {
            const float v = (float)fe_rs_tap(pl.orig, pl.newf, pl.width, lpfw, rolloff, method, beta, p, k);
            if (v == 0.0f) continue;
            for (int c = 0; c < m; ++c) {             // output pe = c * new + p of the block group reads the window shifted by c * orig
                const int pe = c * pl.newf + p, ke = k + c * pl.orig;
                He[(size_t)pe * Ke + ke] = v;
                lo[pe] = std::min(lo[pe], ke); hi[pe] = std::max(hi[pe], ke);
            }
        }
    struct Blk { int p0, n, k0, nch; };
    auto blocks = [&](int band) {
        std::vector<Blk> b;
        for (int p0 = 0; p0 < N; p0 += band) {
            const int n = std::min(band, N - p0);
            int l = Ke, u = -1;
            for (int p = p0; p < p0 + n; ++p) { l = std::min(l, lo[p]); u = std::max(u, hi[p]); }
            if (u < l) { l = 0; u = 0; }
            b.push_back({p0, n, l, (u - l + C::KC) / C::KC});
        }
        return b;
    };
    // cost of a tile of 128 rows in cycles: the producers take about 900 per chunk whatever its N, the MMA thread 12 MMAs per chunk
    // at about 30 + 0.35 N cycles each (46 at N = 32, 86 at N = 160: profiles/r02_ubench_tmem_mma.txt, r02_resample_tc.txt)
    auto cost = [](const std::vector<Blk>& b) {
        double chunks = 0.0, mma = 0.0;
        for (const Blk& x : b) { chunks += x.nch; mma += x.nch * 12.0 * (30.0 + 0.35 * x.n); }
        return std::max(900.0 * chunks, mma);
    };
    const std::vector<Blk> dense = blocks(N), banded = blocks(C::BAND);
    const bool use_banded = mode == 2 || (mode == 3 && cost(banded) < cost(dense));
    const std::vector<Blk>& bl = use_banded ? banded : dense;
    out->m = m; out->N = N; out->orig_e = pl.orig * m;
    out->vtab.clear(); out->hpack.clear();
    int max_n = 0;
    for (const Blk& b : bl) {
        max_n = std::max(max_n, b.n);
        for (int c = 0; c < b.nch; ++c) {
            int4 v;
            v.x = b.k0 + c * C::KC; v.y = b.p0 | (c == 0 ? 0x10000 : 0); v.z = b.n; v.w = (int)out->hpack.size();
            out->vtab.push_back(v);
            const size_t base = out->hpack.size();
            out->hpack.resize(base + (size_t)2 * b.n * C::KC, 0.0f);
            for (int pl_ = 0; pl_ < b.n; ++pl_)
                for (int kk = 0; kk < C::KC; ++kk) {
                    const int ke = v.x + kk;
                    const float t = ke < Ke ? He[(size_t)(b.p0 + pl_) * Ke + ke] : 0.0f;
                    const float h = tf32_rna_host(t), l = tf32_rna_host(t - h);
                    const size_t off = (size_t)(pl_ / 8) * (C::SBO / 4) + (size_t)(kk / 4) * (C::LBO / 4) + (pl_ % 8) * 4 + (kk % 4);
                    out->hpack[base + off] = h;
                    out->hpack[base + (size_t)b.n * C::KC + off] = l;
                }
        }
    }
    out->stage_floats = 2 * max_n * C::KC;
    out->k0_dense = use_banded ? -1 : bl[0].k0;
    return true;
}
typedef std::tuple<int, int, int, int, double, int, double> RsKey;
std::mutex g_rs_mu;
std::map<RsKey, RsDevicePlan*> g_rs_plans;

template <bool IN16, int CH, bool TS>
int rs_launch(unsigned grid, size_t smem, cudaStream_t st, const FeResampleParams& p) {
    static std::atomic<size_t> attr_bytes[64];
    int dev = 0;
    FE_CUDA(cudaGetDevice(&dev));
    std::atomic<size_t>& have = attr_bytes[dev < 64 ? dev : 0];
    if (smem > have.load()) {
        FE_CUDA(cudaFuncSetAttribute(fe_resample_kernel<IN16, CH, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        have.store(smem);
    }
    fe_resample_kernel<IN16, CH, TS><<<grid, FeRsCfg::NT, smem, st>>>(p);
    return 0;
}

int rs_dispatch(bool in16, int ch, bool ts, unsigned grid, size_t smem, cudaStream_t st, const FeResampleParams& p) {
#define RS_CASE(A, C, T) if (in16 == A && ch == C && ts == T) return rs_launch<A, C, T>(grid, smem, st, p);
    RS_CASE(false, 1, false) RS_CASE(false, 1, true) RS_CASE(false, 2, false) RS_CASE(false, 2, true)
    RS_CASE(true, 1, false) RS_CASE(true, 1, true) RS_CASE(true, 2, false) RS_CASE(true, 2, true)
#undef RS_CASE
    return fail(B200FEAT_EINVAL, "resample: unsupported input format");
}

int rs_get_plan(int orig_freq, int new_freq, int lpfw, double rolloff, int method, double beta, RsDevicePlan** out) {
    int dev = 0;
    FE_CUDA(cudaGetDevice(&dev));
    const RsKey key(dev, orig_freq, new_freq, lpfw, rolloff, method, beta);
    std::lock_guard<std::mutex> lk(g_rs_mu);
    auto it = g_rs_plans.find(key);
    if (it != g_rs_plans.end()) { *out = it->second; return 0; }
    RsDevicePlan* dp = new RsDevicePlan();
    dp->plan = fe_rs_make_plan(orig_freq, new_freq, lpfw, rolloff, method, beta);
    if (!dp->plan.error.empty()) {
        const int rc = fail(B200FEAT_EINVAL, "resample: %s", dp->plan.error.c_str());
        delete dp;
        return rc;
    }
    // first call for this filter only: the table upload is synchronous.  Nothing leaks when a step fails.
    cudaError_t e = cudaMalloc(&dp->taps, dp->plan.taps.size() * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&dp->kstart, dp->plan.kstart.size() * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMemcpy(dp->taps, dp->plan.taps.data(), dp->plan.taps.size() * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dp->kstart, dp->plan.kstart.data(), dp->plan.kstart.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
    RsTcPlan tc;
    if (e == cudaSuccess && rs_tc_make(dp->plan, lpfw, rolloff, method, beta, &tc)) {
        dp->tc_m = tc.m; dp->nch = (int)tc.vtab.size(); dp->tc_stage_floats = tc.stage_floats; dp->tc_k0_dense = tc.k0_dense;
        e = cudaMalloc(&dp->hpack, tc.hpack.size() * sizeof(float));
        if (e == cudaSuccess) e = cudaMalloc(&dp->vtab, tc.vtab.size() * sizeof(int4));
        if (e == cudaSuccess) e = cudaMemcpy(dp->hpack, tc.hpack.data(), tc.hpack.size() * sizeof(float), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(dp->vtab, tc.vtab.data(), tc.vtab.size() * sizeof(int4), cudaMemcpyHostToDevice);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(fe_resample_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, FeRsTcCfg::smem_bytes(2 * FeRsTcCfg::MAX_N * FeRsTcCfg::KC));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(fe_resample_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FeRsTcCfg::smem_bytes(2 * FeRsTcCfg::MAX_N * FeRsTcCfg::KC));
    }
    if (e != cudaSuccess) {
        if (dp->taps) cudaFree(dp->taps);
        if (dp->kstart) cudaFree(dp->kstart);
        if (dp->hpack) cudaFree(dp->hpack);
        if (dp->vtab) cudaFree(dp->vtab);
        delete dp;
        return fail(B200FEAT_ECUDA, "resample: uploading the tap table failed: %s", cudaGetErrorString(e));
    }
    g_rs_plans[key] = dp;
    *out = dp;
    return 0;
}

}  // namespace

extern "C" {

int64_t b200feat_resample_out_len(int64_t n_frames, int32_t orig_freq, int32_t new_freq) {
    return fe_rs_out_len(n_frames, orig_freq, new_freq);
}

int b200feat_resample(const void* pcm, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                      const int64_t* lengths, int32_t B, int64_t L_in, int64_t pcm_stride,
                      int32_t orig_freq, int32_t new_freq, int32_t lowpass_filter_width, double rolloff,
                      int32_t method, double beta, float* out, int64_t out_stride, int64_t L_out, void* cuda_stream) {
    if (pcm_dtype != B200FEAT_F32 && pcm_dtype != B200FEAT_S16)
        return fail(B200FEAT_EINVAL, "resample: pcm_dtype must be B200FEAT_F32 or B200FEAT_S16, got %d", pcm_dtype);
    if (channels != 1 && channels != 2) return fail(B200FEAT_EINVAL, "resample: channels must be 1 or 2, got %d", channels);
    if (orig_freq <= 0 || new_freq <= 0)
        return fail(B200FEAT_EINVAL, "resample: Original frequency and desired frequecy should be positive");
    if (B < 0) return fail(B200FEAT_EINVAL, "resample: B = %d < 0", B);
    if (L_in < 0 || L_in > (int64_t)1 << 40) return fail(B200FEAT_EINVAL, "resample: bad L_in = %lld", (long long)L_in);
    if (pcm_stride < L_in * channels) return fail(B200FEAT_EINVAL, "resample: pcm_stride %lld < L_in * channels", (long long)pcm_stride);
    const int64_t full = fe_rs_out_len(L_in, orig_freq, new_freq);
    if (L_out < 0 || L_out > full) return fail(B200FEAT_EINVAL, "resample: L_out %lld not in [0, %lld]", (long long)L_out, (long long)full);
    if (out_stride < L_out) return fail(B200FEAT_EINVAL, "resample: out_stride %lld < L_out %lld", (long long)out_stride, (long long)L_out);
    if (B > 0 && L_out > 0 && (pcm == nullptr || out == nullptr)) return fail(B200FEAT_EINVAL, "resample: null pcm/out pointer");
    if (B == 0 || L_out == 0) return 0;
    if (int rc = check_device_arch()) return rc;
    if (method == FE_RS_KAISER) beta = (double)(float)beta;    // the reference holds beta in a float32 tensor
    else beta = 0.0;
    RsDevicePlan* dp = nullptr;
    if (int rc = rs_get_plan(orig_freq, new_freq, lowpass_filter_width, rolloff, method, beta, &dp)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    if (dp->hpack != nullptr && B <= 65535) {
        // ---- tensor-core path: (mono mix to aligned float32 rows, unless the input is that already) + the GEMM kernel
        typedef FeRsTcCfg C;
        const bool in16 = pcm_dtype == B200FEAT_S16;
        const float w0 = mix_host ? mix_host[0] : 1.0f / channels, w1 = channels == 2 ? (mix_host ? mix_host[1] : 0.5f) : 0.0f;
        const bool direct = channels == 1 && !in16 && w0 == 1.0f && aligned16(pcm) && pcm_stride % 4 == 0;
        const bool s16x2 = channels == 2 && in16 && aligned16(pcm) && pcm_stride % 8 == 0;     // frames read as 4-byte elements
        float* mono = nullptr;
        const int64_t Ls = (L_in + 3) / 4 * 4;
        if (!direct && !s16x2) {
            {   // stream-ordered scratch for the mono rows: keep the pool's memory across calls (the default releases it at
                // every synchronisation and the next call maps it again -- milliseconds for a 5 GB batch)
                static std::atomic<int> pool_set[64];
                int dev0 = 0;
                FE_CUDA(cudaGetDevice(&dev0));
                if (dev0 >= 0 && dev0 < 64 && pool_set[dev0].exchange(1) == 0) {
                    cudaMemPool_t pool;
                    if (cudaDeviceGetDefaultMemPool(&pool, dev0) == cudaSuccess) {
                        uint64_t keep = ~(uint64_t)0;
                        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
                    }
                }
            }
            FE_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&mono), (size_t)B * Ls * sizeof(float), st));
            int64_t gx = (L_in + 255) / 256;
            if (gx > 2048) gx = 2048;
            const dim3 grid((unsigned)gx, (unsigned)B);
            ProfScope ps(st, 1);
            if (in16 && channels == 2) fe_pcm_mono_kernel<true, 2><<<grid, 256, 0, st>>>(pcm, pcm_stride, w0, w1, mono, Ls, L_in);
            else if (in16) fe_pcm_mono_kernel<true, 1><<<grid, 256, 0, st>>>(pcm, pcm_stride, w0, w1, mono, Ls, L_in);
            else if (channels == 2) fe_pcm_mono_kernel<false, 2><<<grid, 256, 0, st>>>(pcm, pcm_stride, w0, w1, mono, Ls, L_in);
            else fe_pcm_mono_kernel<false, 1><<<grid, 256, 0, st>>>(pcm, pcm_stride, w0, w1, mono, Ls, L_in);
            g_launches++;
        }
        FeRsTcParams q;
        memset(&q, 0, sizeof(q));
        q.x = (direct || s16x2) ? static_cast<const float*>(pcm) : mono;
        q.x_stride = direct ? pcm_stride : (s16x2 ? pcm_stride / 2 : Ls);
        q.w0 = w0; q.w1 = w1;
        q.lengths = lengths; q.L_in = L_in; q.out = out; q.out_stride = out_stride; q.L_out = L_out;
        q.hpack = dp->hpack; q.B = B; q.orig = dp->plan.orig * dp->tc_m; q.newf = dp->plan.newf * dp->tc_m; q.width = dp->plan.width;
        q.nch = dp->nch; q.vtab = dp->vtab; q.stage_floats = dp->tc_stage_floats; q.k0_dense = dp->tc_k0_dense;
        const int64_t nblocks = (L_out + q.newf - 1) / q.newf;
        q.tiles_per_clip = (int32_t)((nblocks + C::M - 1) / C::M);
        const int64_t total = (int64_t)B * q.tiles_per_clip;
        if (total > 0x7fffffff) return fail(B200FEAT_EINVAL, "resample: batch too large");
        q.total_tiles = (int32_t)total;
        q.out_vec = (aligned16(out) && out_stride % 4 == 0) ? 1 : 0;
        int dev = 0;
        FE_CUDA(cudaGetDevice(&dev));
        const int sms = g_num_sms[dev < 64 ? dev : 0] > 0 ? g_num_sms[dev < 64 ? dev : 0] : 148;
        {
            ProfScope ps(st, 0);
            if (s16x2) fe_resample_tc_kernel<true><<<(unsigned)(total < sms ? total : sms), C::NTHREADS, C::smem_bytes(q.stage_floats), st>>>(q);
            else fe_resample_tc_kernel<false><<<(unsigned)(total < sms ? total : sms), C::NTHREADS, C::smem_bytes(q.stage_floats), st>>>(q);
            g_launches++;
        }
        if (mono) FE_CUDA(cudaFreeAsync(mono, st));
        FE_CUDA(cudaGetLastError());
        return 0;
    }
    FeResampleParams p;
    memset(&p, 0, sizeof(p));
    fe_rs_fill_params(p, dp->plan, L_in, L_out, channels, pcm_dtype == B200FEAT_S16, mix_host);
    p.in = pcm; p.lengths = lengths; p.out = out; p.in_stride = pcm_stride; p.out_stride = out_stride;
    p.taps = dp->taps; p.kstart = dp->kstart;
    const int64_t frame_bytes = (int64_t)channels * (pcm_dtype == B200FEAT_S16 ? 2 : 4);
    p.vec_ok = (reinterpret_cast<uintptr_t>(pcm) % frame_bytes == 0 && pcm_stride % channels == 0) ? 1 : 0;
    const int64_t tiles = p.tiles_per_clip;
    int64_t per_launch = 0x7fffffff / tiles;
    if (per_launch < 1) return fail(B200FEAT_EINVAL, "resample: clip too long");
    for (int64_t c0 = 0; c0 < B; c0 += per_launch) {
        const int64_t nc = (B - c0 < per_launch) ? (B - c0) : per_launch;
        p.clip0 = (int32_t)c0;
        ProfScope ps(st, 0);
        if (int rc = rs_dispatch(pcm_dtype == B200FEAT_S16, channels, dp->plan.taps_smem, (unsigned)(nc * tiles),
                                 dp->plan.smem_bytes, st, p)) return rc;
        g_launches++;
    }
    FE_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------ conditioning, reference rule
int b200feat_condition_stereo_auto(const void* pcm, int32_t pcm_dtype, const int64_t* lengths, int32_t B, int64_t L_in,
                                   int64_t pcm_stride, int32_t orig_freq, int32_t new_freq, int32_t lowpass_filter_width,
                                   double rolloff, int32_t method, double beta, float* out, int64_t out_stride, int64_t L_out,
                                   float* scratch, int64_t scratch_floats, int32_t* decisions, float* correlations,
                                   void* cuda_stream) {
    if (B > 0 && L_out > 0 && scratch == nullptr) return fail(B200FEAT_EINVAL, "condition_stereo_auto: null scratch");
    if (scratch_floats < (int64_t)B * out_stride)
        return fail(B200FEAT_EWORKSPACE, "condition_stereo_auto: scratch holds %lld floats, need B * out_stride = %lld",
                    (long long)scratch_floats, (long long)((int64_t)B * out_stride));
    const float only0[2] = {1.0f, 0.0f}, only1[2] = {0.0f, 1.0f};
    if (int rc = b200feat_resample(pcm, pcm_dtype, 2, only0, lengths, B, L_in, pcm_stride, orig_freq, new_freq, lowpass_filter_width,
                                   rolloff, method, beta, out, out_stride, L_out, cuda_stream)) return rc;
    if (int rc = b200feat_resample(pcm, pcm_dtype, 2, only1, lengths, B, L_in, pcm_stride, orig_freq, new_freq, lowpass_filter_width,
                                   rolloff, method, beta, scratch, out_stride, L_out, cuda_stream)) return rc;
    if (B == 0 || L_out == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    {
        ProfScope ps(st, 1);
        fe_mono_auto_kernel<<<(unsigned)B, 1024, 0, st>>>(out, scratch, out_stride, L_out, decisions, correlations);
        g_launches++;
    }
    FE_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------ PCM -> log-mel in one call
int64_t b200feat_pcm_whisper_workspace_bytes(int32_t B, int64_t L_in, int32_t orig_freq, int32_t new_freq, int32_t padding) {
    if (B < 0 || L_in < 0 || orig_freq <= 0 || new_freq <= 0 || padding < 0) return -1;
    const int64_t L = fe_rs_out_len(L_in, orig_freq, new_freq);
    const int64_t Ls = (L + 3) / 4 * 4;                          // waveform rows 16-byte aligned (the Tensor-Memory kernel's copies)
    return 256 + (((int64_t)B * 8 + 255) & ~(int64_t)255) + (((int64_t)B * Ls * 4 + 255) & ~(int64_t)255) +
           b200feat_workspace_bytes(B200FEAT_KIND_WHISPER, B, L + padding);
}

int b200feat_pcm_whisper_logmel(const void* pcm, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                                const int64_t* lengths, int32_t B, int64_t L_in, int64_t pcm_stride,
                                int32_t orig_freq, int32_t new_freq, int32_t lowpass_filter_width, double rolloff,
                                int32_t method, double beta, int32_t n_mels, int32_t padding, int32_t norm_mode,
                                void* out, int32_t out_dtype, int64_t out_batch_stride, int64_t T_out,
                                void* workspace, int64_t workspace_bytes, void* cuda_stream) {
    if (int rc = whisper_ex_dtypes("pcm_whisper_logmel", B200FEAT_F32, out_dtype)) return rc;
    if (B < 0 || L_in < 0 || orig_freq <= 0 || new_freq <= 0 || padding < 0)
        return fail(B200FEAT_EINVAL, "pcm_whisper_logmel: negative size, padding or frequency");
    const int64_t need = b200feat_pcm_whisper_workspace_bytes(B, L_in, orig_freq, new_freq, padding);
    if (B > 0 && (workspace == nullptr || workspace_bytes < need))
        return fail(B200FEAT_EWORKSPACE, "pcm_whisper_logmel: workspace too small (%lld bytes, need %lld)", (long long)workspace_bytes, (long long)need);
    if (B == 0) return 0;
    const int64_t L = fe_rs_out_len(L_in, orig_freq, new_freq), Ls = (L + 3) / 4 * 4;
    char* w = static_cast<char*>(workspace);
    w += (256 - (reinterpret_cast<uintptr_t>(w) & 255)) & 255;
    int64_t* d_len = reinterpret_cast<int64_t*>(w);
    w += ((int64_t)B * 8 + 255) & ~(int64_t)255;
    float* wave = reinterpret_cast<float*>(w);
    w += ((int64_t)B * Ls * 4 + 255) & ~(int64_t)255;
    const int64_t ws_left = workspace_bytes - (w - static_cast<char*>(workspace));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    if (int rc = b200feat_resample(pcm, pcm_dtype, channels, mix_host, lengths, B, L_in, pcm_stride, orig_freq, new_freq,
                                   lowpass_filter_width, rolloff, method, beta, wave, Ls, L, cuda_stream)) return rc;
    if (lengths) {
        int64_t g = orig_freq, h = new_freq;
        while (h) { const int64_t t = g % h; g = h; h = t; }
        fe_rs_lengths_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(lengths, d_len, B, L_in, orig_freq / g, new_freq / g, L);
        g_launches++;
    }
    return whisper_logmel_impl(wave, false, lengths ? d_len : nullptr, B, L, Ls, n_mels, padding, norm_mode, out,
                               out_dtype == B200FEAT_F16, out_batch_stride, T_out, w, ws_left, cuda_stream);
}

}  // extern "C"

#include "b200feat_host.inl"
