// Fused compute + collective kernels over NVLink 5 / NVSwitch peer memory (one process per GPU).
//
//   idcb_allreduce_optim   : bucketed gradient all-reduce fused with loss-unscale, the optimizer update
//                            (RMSprop / SGD / Adam on the owned shard, optimizer state sharded N ways),
//                            the bf16 cast of the new weights and their all-gather      
//   idcb_fedavg_reduce     : example-weighted FedAvg delta mean over clients resident on all GPUs with the
//                            non-finite-client guard, server SGD apply and broadcast     
//   idcb_secagg_masked_sum : secure aggregation: fixed-point quantise, pairwise Philox masks generated in
//                            registers (masks never exist in memory), integer sum, dequantise, mean,
//                            broadcast                                                    
//
// All buffers a peer touches live in symmetric memory (same size on every rank, mapped into every
// process); a kernel gets the table of peer base pointers and, when the fabric supports it, the NVLS
// multicast alias of the same allocation. Data path per owned shard (rank r owns the r-th 1/N of every
// range): reduce = `multimem.ld_reduce` on the multicast address (the switch adds the N copies) or N
// peer `ld.relaxed.sys` loads; result broadcast = `multimem.st` or N peer stores. Cross-rank ordering
// uses epoch-stamped flags in each rank's signal pad (st.release.sys / ld.acquire.sys); the epoch lives
// in device memory so a captured CUDA graph can be replayed.
//
// Reference behaviour replaced: tf.distribute.MirroredStrategy's NcclAllReduce of the packed gradient
// buffer followed by one optimizer kernel per variable (SURVEY 2.B B12, 2.E C1/K13;
// /root/reference/dist_model_tf_mobile.py:115-132), TFF's federated_mean + server optimizer
// (fed_model.py:207-227) and the Paillier aggregate of secure_fed_model.py:160-168
// (the north star replaces Paillier by pairwise masks).
#include <mutex>

#include "common.cuh"
#include "philox.cuh"
#include "../include/idcb200.h"

namespace idcb {

constexpr int kMaxWorld = 8;
constexpr int kCommThreads = 256;

// ---- system-scope memory helpers ------------------------------------------------------------------
__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint4 ld_sys_u4(const uint32_t* p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys_f4(float* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_sys_u2(void* p, uint2 v) {
    asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" :: "l"(p), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// NVLS: one load that returns the sum over every rank's copy / one store that lands on every rank
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st_f4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void mc_st_u2(void* mc, uint2 v) {
    // bf16x2 pairs travel as two 32-bit words (multimem.st is type-agnostic for plain stores)
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};"
                 :: "l"(mc), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)) : "memory");
}
__device__ __forceinline__ uint32_t mc_ld_reduce_u32(const uint32_t* mc) {
    uint32_t v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v) : "l"(mc) : "memory");
    return v;
}

// ---- cross-rank barrier ----------------------------------------------------------------------------
// Signal pad layout (uint32 words, per rank): [0..7] arrival flags written by peer p at slot p,
// [8] the rank's own epoch counter, [16] local CTA arrival counter, [17] local release flag.
struct Peers {
    int world, rank;
    uint32_t* pad[kMaxWorld];   // signal pad of every rank (peer-mapped)
    int use_barrier;            // 0 when one process emulates the ranks sequentially (tests)
};

__device__ __forceinline__ void st_relaxed_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_gpu_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// All CTAs of the local grid arrive; the last one to arrive handshakes with every peer; everyone is released.
// `epoch` distinguishes the barriers inside one launch (the pad epoch advances by 2 per launch).
// Cost model (measured: the first version, with a system-scope release per flag store, took ~11 us per barrier
// on 2 GPUs): system-scope fences are the expensive part, so there is exactly ONE per CTA (making its own data
// writes visible everywhere before it is counted) and one in the handshaking thread; the flags themselves are
// relaxed stores issued by the lanes of one warp in parallel (lane p talks to rank p), and the local release of
// the other CTAs is GPU scope.
// A peer never arrived: say who is missing (per-rank flag state) before trapping, so that a protocol bug or a
// crashed rank shows up as a diagnosable launch failure instead of a silent hang.
__device__ __noinline__ void barrier_timeout(uint32_t* mypad, int rank, int world, uint32_t epoch, int waiting_for) {
    printf("idcb comm barrier timeout: rank %d of %d waited ~10 s for rank %d at epoch %u; flags seen:", rank, world,
           waiting_for, epoch);
    for (int p = 0; p < world; ++p) printf(" [%d]=%u", p, ld_acquire_sys_u32(mypad + p));
    printf(" (local arrivals %u, released %u)\n", mypad[16], mypad[17]);
    __trap();
}

// `publish`: this kernel has written data that peers will read after the barrier (peer / multicast stores): every
// CTA then makes its writes visible system-wide before it is counted. An opening barrier that only waits for the
// peers' PREVIOUS kernels (whose writes the kernel boundary already completed) passes publish = false.
__device__ void grid_and_world_barrier(const Peers& pr, uint32_t epoch, bool publish = true) {
    if (!pr.use_barrier) return;
    __syncthreads();
    uint32_t* mypad = pr.pad[pr.rank];
    __shared__ int s_last;
    if (threadIdx.x == 0) {
        if (publish) __threadfence_system();
        const uint32_t arrived = atomicAdd(mypad + 16, 1u) + 1u;
        s_last = arrived == gridDim.x;
        if (s_last) {
            atomicExch(mypad + 16, 0u);
            __threadfence();   // the other CTAs' (already published) writes, seen through the counter, precede the flags
        }
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x < 32) {
            const int p = threadIdx.x;
            if (p < pr.world) {
                st_relaxed_sys_u32(pr.pad[p] + pr.rank, epoch);
                long long t0 = clock64();
                while (static_cast<int32_t>(ld_acquire_sys_u32(mypad + p) - epoch) < 0) {
                    if (clock64() - t0 > 20000000000LL) barrier_timeout(mypad, pr.rank, pr.world, epoch, p);  // ~10 s
                }
            }
            __syncwarp();
            if (p == 0) st_release_gpu_u32(mypad + 17, epoch);
        }
    } else if (threadIdx.x == 0) {
        long long t0 = clock64();
        while (static_cast<int32_t>(ld_acquire_gpu_u32(mypad + 17) - epoch) < 0) {
            if (clock64() - t0 > 20000000000LL) __trap();
        }
    }
    __syncthreads();
}

__device__ __forceinline__ uint32_t load_epoch(const Peers& pr) {
    // written by the previous launch on this GPU (kernel boundary orders it): a plain volatile load is enough
    return pr.use_barrier ? *reinterpret_cast<volatile const uint32_t*>(pr.pad[pr.rank] + 8) : 0u;
}
// the LAST CTA to finish bumps the stored epoch (all CTAs have read it long before: they passed barriers)
__device__ __forceinline__ void store_epoch(const Peers& pr, uint32_t next, uint32_t* done_counter) {
    if (!pr.use_barrier) return;
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t d = atomicAdd(done_counter, 1u) + 1u;
        if (d == gridDim.x) {
            atomicExch(done_counter, 0u);
            *reinterpret_cast<volatile uint32_t*>(pr.pad[pr.rank] + 8) = next;   // read by the next launch only
        }
    }
}

__device__ __forceinline__ void shard_bounds(long long n4, int world, int rank, long long& lo, long long& hi) {
    const long long per = (n4 + world - 1) / world;
    lo = per * rank;
    hi = lo + per;
    if (lo > n4) lo = n4;
    if (hi > n4) hi = n4;
}

// =====================================================================================================
// K15: gradient all-reduce + optimizer + cast + all-gather
// =====================================================================================================
struct AllreduceOptArgs {
    Peers pr;
    float* grad[kMaxWorld];
    float* master[kMaxWorld];
    __nv_bfloat16* shadow[kMaxWorld];   // may be null
    float* mc_grad;                     // multicast aliases (null => peer loads/stores)
    float* mc_master;
    __nv_bfloat16* mc_shadow;
    float* s1; float* s2;               // local optimizer state (only the owned shard is touched)
    long long n;                        // elements in the range (multiple of 4)
    int kind;
    float lr, p1, p2, eps, grad_scale;
    int owner;                          // -1: every rank owns 1/N (mirrored); r: rank r owns everything
                                        //     (central storage: reduce-to-owner + broadcast)
    const float* scal[kMaxWorld];       // optional: per-rank 4-float slots (loss sum, correct, count, -)
    float* scal_out;                    //           their sum over ranks, piggy-backed on the barrier (8.3)
};

__device__ __forceinline__ float ld_sys_f32(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ float opt_step(int kind, float w, float g, float& s1, float& s2,
                                          float lr, float p1, float p2, float eps) {
    if (kind == IDCB_OPT_RMSPROP) {
        s1 = p1 * s1 + (1.f - p1) * g * g;
        return w - lr * g / (sqrtf(s1) + eps);
    } else if (kind == IDCB_OPT_SGD) {
        s1 = p1 * s1 - lr * g;
        return w + s1;
    }
    s1 = p1 * s1 + (1.f - p1) * g;
    s2 = p2 * s2 + (1.f - p2) * g * g;
    return w - lr * s1 / (sqrtf(s2) + eps);
}

__global__ void __launch_bounds__(kCommThreads) allreduce_optim_kernel(AllreduceOptArgs a) {
    const Peers& pr = a.pr;
    const uint32_t e0 = load_epoch(pr);
    // every rank's backward has finished once its kernel has started (stream order) -> safe to read
    grid_and_world_barrier(pr, e0 + 1, /*publish=*/false);
    long long lo, hi;
    shard_bounds(a.n >> 2, pr.world, pr.rank, lo, hi);
    if (a.owner >= 0) {
        lo = 0;
        hi = pr.rank == a.owner ? (a.n >> 2) : 0;
    }
    if (a.scal_out != nullptr && blockIdx.x == 0 && threadIdx.x < 4) {
        float t = 0.f;
        for (int p = 0; p < pr.world; ++p) t += ld_sys_f32(a.scal[p] + threadIdx.x);
        a.scal_out[threadIdx.x] = t;
    }
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    // Software pipeline over chunks of kU 16-byte reductions per thread: the reductions of chunk c+1 are issued
    // BEFORE chunk c is updated and broadcast, so the two NVLink directions work at the same time (reduce =
    // this GPU pulling, broadcast = this GPU pushing) and 2*kU reductions per thread hide the ~2-4 us round trip
    // through the switch. (A single pass with everything in flight leaves the push phase waiting for the pull.)
    // (kU = 4 measured on 8 GPUs: 116.6 us for the DenseNet-121 bucket against 108.8 us with kU = 2)
    constexpr int kU = 2;
    const long long tstride = (long long)gridDim.x * blockDim.x;
    const long long cstride = tstride * kU;
    auto pull = [&](long long i0, float4 (&gg)[kU]) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long i = i0 + u * tstride;
            gg[u] = zero;
            if (i >= hi) continue;
            const long long e = i << 2;
            if (a.mc_grad != nullptr) {
                gg[u] = mc_ld_reduce_f4(a.mc_grad + e);
            } else {
                for (int q = 0; q < pr.world; ++q) {
                    const int p = (pr.rank + q) % pr.world;   // stagger peers so ranks do not gang up on one
                    const float4 v = ld_sys_f4(a.grad[p] + e);
                    gg[u].x += v.x; gg[u].y += v.y; gg[u].z += v.z; gg[u].w += v.w;
                }
            }
        }
    };
    auto update_and_push = [&](long long i0, const float4 (&gg)[kU]) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long i = i0 + u * tstride;
            if (i >= hi) continue;
            const long long e = i << 2;
            float4 g = gg[u];
            // The consumed bucket must be zero again before the next backward accumulates into it. With live
            // barriers every rank clears its OWN bucket after the closing barrier (HBM speed, nothing on NVLink);
            // broadcasting zeros from here would add (N-1)/N of the bucket to every GPU's inbound traffic.
            // Only the single-process emulation (no barriers, ranks run one after another) clears per shard.
            if (!pr.use_barrier) {
                for (int q = 0; q < pr.world; ++q) st_sys_f4(a.grad[(pr.rank + q) % pr.world] + e, zero);
            }
            g.x *= a.grad_scale; g.y *= a.grad_scale; g.z *= a.grad_scale; g.w *= a.grad_scale;
            float4 w = *reinterpret_cast<const float4*>(a.master[pr.rank] + e);
            float4 s1 = a.s1 ? *reinterpret_cast<const float4*>(a.s1 + e) : zero;
            float4 s2 = a.s2 ? *reinterpret_cast<const float4*>(a.s2 + e) : zero;
            w.x = opt_step(a.kind, w.x, g.x, s1.x, s2.x, a.lr, a.p1, a.p2, a.eps);
            w.y = opt_step(a.kind, w.y, g.y, s1.y, s2.y, a.lr, a.p1, a.p2, a.eps);
            w.z = opt_step(a.kind, w.z, g.z, s1.z, s2.z, a.lr, a.p1, a.p2, a.eps);
            w.w = opt_step(a.kind, w.w, g.w, s1.w, s2.w, a.lr, a.p1, a.p2, a.eps);
            if (a.s1) *reinterpret_cast<float4*>(a.s1 + e) = s1;
            if (a.s2) *reinterpret_cast<float4*>(a.s2 + e) = s2;
            const uint2 wb = make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
            if (a.mc_master != nullptr) {
                mc_st_f4(a.mc_master + e, w);
                if (a.mc_shadow) mc_st_u2(a.mc_shadow + e, wb);
            } else {
                for (int q = 0; q < pr.world; ++q) {
                    const int p = (pr.rank + q) % pr.world;
                    st_sys_f4(a.master[p] + e, w);
                    if (a.shadow[p]) st_sys_u2(a.shadow[p] + e, wb);
                }
            }
        }
    };
    {
        float4 cur[kU], nxt[kU];
        long long i0 = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
        if (i0 < hi) pull(i0, cur);
        for (; i0 < hi; i0 += cstride) {
            const bool more = i0 + cstride < hi;
            if (more) pull(i0 + cstride, nxt);
            update_and_push(i0, cur);
#pragma unroll
            for (int u = 0; u < kU; ++u) cur[u] = nxt[u];
        }
    }
    // every rank must have received all shards (and had its bucket re-zeroed) before anyone goes on
    grid_and_world_barrier(pr, e0 + 2);
    if (pr.use_barrier) {
        float4* mine = reinterpret_cast<float4*>(a.grad[pr.rank]);
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (a.n >> 2); i += tstride) mine[i] = zero;
    }
    store_epoch(pr, e0 + 2, pr.pad[pr.rank] + 18);
}

// =====================================================================================================
// BatchNorm moving statistics: MEAN over replicas when they are read (checkpoint / evaluate). Under
// MirroredStrategy they are ON_READ variables: every replica updates its own copy from its own batches and
// the copies are averaged on read. Owner-computes like the kernels above: rank r
// averages its 1/N of the buffer out of every peer and writes the result back into every peer.
// =====================================================================================================
struct StatMeanArgs {
    Peers pr;
    float* stats[kMaxWorld];
    float* mc;
    long long n;
    float scale;     // 1/world: mean over replicas (BatchNorm moving statistics); 1: plain sum (metric scalars)
};

__global__ void __launch_bounds__(kCommThreads) stat_mean_kernel(StatMeanArgs a) {
    const Peers& pr = a.pr;
    const uint32_t e0 = load_epoch(pr);
    grid_and_world_barrier(pr, e0 + 1, /*publish=*/false);
    long long lo, hi;
    shard_bounds(a.n >> 2, pr.world, pr.rank, lo, hi);
    const float inv = a.scale;
    const long long tstride = (long long)gridDim.x * blockDim.x;
    for (long long i = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < hi; i += tstride) {
        const long long e = i << 2;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.mc != nullptr) {
            acc = mc_ld_reduce_f4(a.mc + e);
        } else {
            for (int q = 0; q < pr.world; ++q) {
                const float4 v = ld_sys_f4(a.stats[(pr.rank + q) % pr.world] + e);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
        }
        acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
        if (a.mc != nullptr) {
            mc_st_f4(a.mc + e, acc);
        } else {
            for (int q = 0; q < pr.world; ++q) st_sys_f4(a.stats[(pr.rank + q) % pr.world] + e, acc);
        }
    }
    grid_and_world_barrier(pr, e0 + 2);
    store_epoch(pr, e0 + 2, pr.pad[pr.rank] + 18);
}

// =====================================================================================================
// K16: FedAvg  w <- w + lr_s * sum_k n_k (w_k - w) / sum_k n_k     over all clients on all GPUs
// =====================================================================================================
struct FedAvgArgs {
    Peers pr;
    const float* clients[kMaxWorld];   // per rank: [k_local][stride] trained client weights (symmetric)
    const float* weights[kMaxWorld];   // per rank: [k_local] example counts (0 => dropped / non-finite)
    float* global[kMaxWorld];          // per rank: the server model (identical everywhere)
    __nv_bfloat16* shadow[kMaxWorld];
    float* mc_global; __nv_bfloat16* mc_shadow;
    int k_local;
    long long stride;                  // elements between consecutive clients
    long long offset, n;               // the aggregated range inside a model (multiple of 4)
    float server_lr;
};

// pre-pass (local): a client whose delta contains a non-finite value contributes nothing (TFF zeroes
// its update and its weight). flags[k] is cleared to 0 if client k has any non-finite entry.
__global__ void fedavg_finite_kernel(const float* clients, long long stride, long long offset, long long n,
                                     int k_local, float* weights) {
    const int k = blockIdx.y;
    if (k >= k_local) return;
    const float* w = clients + k * stride + offset;
    bool bad = false;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        bad |= !isfinite(w[i]);
    if (__syncthreads_or(bad) && threadIdx.x == 0) weights[k] = 0.f;
}

__global__ void __launch_bounds__(kCommThreads) fedavg_reduce_kernel(FedAvgArgs a) {
    const Peers& pr = a.pr;
    const uint32_t e0 = load_epoch(pr);
    grid_and_world_barrier(pr, e0 + 1, /*publish=*/false);
    // total weight (tiny: world * k_local scalars, read by every thread once)
    float wsum = 0.f;
    for (int p = 0; p < pr.world; ++p)
        for (int k = 0; k < a.k_local; ++k) wsum += a.weights[p][k];
    const float inv = wsum > 0.f ? a.server_lr / wsum : 0.f;
    long long lo, hi;
    shard_bounds(a.n >> 2, pr.world, pr.rank, lo, hi);
    for (long long i = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < hi;
         i += (long long)gridDim.x * blockDim.x) {
        const long long e = a.offset + (i << 2);
        const float4 w0 = *reinterpret_cast<const float4*>(a.global[pr.rank] + e);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int q = 0; q < pr.world; ++q) {
            const int p = (pr.rank + q) % pr.world;
            for (int k = 0; k < a.k_local; ++k) {
                const float nk = a.weights[p][k];
                if (nk == 0.f) continue;   // dropped client: its (possibly non-finite) values are never read
                const float4 v = ld_sys_f4(a.clients[p] + k * a.stride + e);
                acc.x += nk * (v.x - w0.x); acc.y += nk * (v.y - w0.y);
                acc.z += nk * (v.z - w0.z); acc.w += nk * (v.w - w0.w);
            }
        }
        float4 w = make_float4(w0.x + inv * acc.x, w0.y + inv * acc.y, w0.z + inv * acc.z, w0.w + inv * acc.w);
        const uint2 wb = make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
        if (a.mc_global != nullptr) {
            mc_st_f4(a.mc_global + e, w);
            if (a.mc_shadow) mc_st_u2(a.mc_shadow + e, wb);
        } else {
            for (int q = 0; q < pr.world; ++q) {
                const int p = (pr.rank + q) % pr.world;
                st_sys_f4(a.global[p] + e, w);
                if (a.shadow[p]) st_sys_u2(a.shadow[p] + e, wb);
            }
        }
    }
    grid_and_world_barrier(pr, e0 + 2);
    store_epoch(pr, e0 + 2, pr.pad[pr.rank] + 18);
}

// One launch for a whole FedAvg aggregation (every trainable range of the model): the non-finite-client scan of
// the LOCAL clients, the weighted delta mean of every range (sharded over the ranks, result multicast into every
// rank's symmetric buffer), and the assembly of this rank's private copy of the complete new server model
// (aggregated ranges from the symmetric buffer after the closing barrier, everything else copied from the
// round's server model). Replaces: one finite-flag launch and one reduce launch per range plus two host-issued
// device copies.
constexpr int kMaxFedRanges = 16;
struct FedAvgRoundArgs {
    FedAvgArgs f;                       // offset / n unused
    int n_ranges;
    long long off[kMaxFedRanges], len[kMaxFedRanges];   // ascending, disjoint, multiples of 4
    const float* base;                  // local: this round's server model [total] (read only)
    float* out;                         // local, private: the new server model [total]
    long long total;
};

__device__ __forceinline__ bool finite4(const float4& v) {
    return isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w);
}

__global__ void __launch_bounds__(kCommThreads) fedavg_round_kernel(FedAvgRoundArgs a) {
    const Peers& pr = a.f.pr;
    const uint32_t e0 = load_epoch(pr);
    const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    // phase 0 (local): drop clients with a non-finite weight anywhere in the aggregated ranges; copy the parts of
    // the model that are not aggregated
    {
        float* wl = const_cast<float*>(a.f.weights[pr.rank]);
        for (int k = 0; k < a.f.k_local; ++k) {
            bool bad = false;
            const float* w = a.f.clients[pr.rank] + k * a.f.stride;
            for (int r = 0; r < a.n_ranges; ++r) {
                const float4* src = reinterpret_cast<const float4*>(w + a.off[r]);
                for (long long i = tid; i < (a.len[r] >> 2); i += nth) bad |= !finite4(src[i]);
            }
            if (__syncthreads_or(bad) && threadIdx.x == 0) wl[k] = 0.f;
        }
        long long prev = 0;
        for (int r = 0; r <= a.n_ranges; ++r) {
            const long long end = r < a.n_ranges ? a.off[r] : a.total;
            const float4* src = reinterpret_cast<const float4*>(a.base + prev);
            float4* dst = reinterpret_cast<float4*>(a.out + prev);
            for (long long i = tid; i < ((end - prev) >> 2); i += nth) dst[i] = src[i];
            if (r < a.n_ranges) prev = a.off[r] + a.len[r];
        }
    }
    __syncthreads();                      // (with the barriers off the launch is a single CTA: see the entry point)
    grid_and_world_barrier(pr, e0 + 1);   // publishes the cleared weights; peers' previous kernels are done
    float wsum = 0.f;
    for (int p = 0; p < pr.world; ++p)
        for (int k = 0; k < a.f.k_local; ++k) wsum += ld_sys_f32(a.f.weights[p] + k);
    const float inv = wsum > 0.f ? a.f.server_lr / wsum : 0.f;
    for (int r = 0; r < a.n_ranges; ++r) {
        long long lo, hi;
        shard_bounds(a.len[r] >> 2, pr.world, pr.rank, lo, hi);
        for (long long i = lo + tid; i < hi; i += nth) {
            const long long e = a.off[r] + (i << 2);
            const float4 w0 = *reinterpret_cast<const float4*>(a.base + e);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int q = 0; q < pr.world; ++q) {
                const int p = (pr.rank + q) % pr.world;
                for (int k = 0; k < a.f.k_local; ++k) {
                    const float nk = ld_sys_f32(a.f.weights[p] + k);
                    if (nk == 0.f) continue;   // dropped client: its (possibly non-finite) values are never read
                    const float4 v = ld_sys_f4(a.f.clients[p] + k * a.f.stride + e);
                    acc.x += nk * (v.x - w0.x); acc.y += nk * (v.y - w0.y);
                    acc.z += nk * (v.z - w0.z); acc.w += nk * (v.w - w0.w);
                }
            }
            const float4 w = make_float4(w0.x + inv * acc.x, w0.y + inv * acc.y, w0.z + inv * acc.z, w0.w + inv * acc.w);
            const uint2 wb = make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
            if (a.f.mc_global != nullptr) {
                mc_st_f4(a.f.mc_global + e, w);
                if (a.f.mc_shadow) mc_st_u2(a.f.mc_shadow + e, wb);
            } else {
                for (int q = 0; q < pr.world; ++q) {
                    const int p = (pr.rank + q) % pr.world;
                    st_sys_f4(a.f.global[p] + e, w);
                    if (a.f.shadow[p]) st_sys_u2(a.f.shadow[p] + e, wb);
                }
            }
        }
    }
    __syncthreads();
    grid_and_world_barrier(pr, e0 + 2);
    // every rank's symmetric buffer now holds the aggregated ranges: assemble the private copy
    for (int r = 0; r < a.n_ranges; ++r) {
        const float* src = a.f.global[pr.rank] + a.off[r];
        float4* dst = reinterpret_cast<float4*>(a.out + a.off[r]);
        for (long long i = tid; i < (a.len[r] >> 2); i += nth) dst[i] = ld_sys_f4(src + (i << 2));
    }
    store_epoch(pr, e0 + 2, pr.pad[pr.rank] + 18);
}

// =====================================================================================================
// K17: secure aggregation by pairwise masks in a 2^32 fixed-point ring
//   client k sends  y_k = Q(w_k) + sum_{j>k} PRG(k,j) - sum_{j<k} PRG(j,k)   (mod 2^32)
//   sum_k y_k = sum_k Q(w_k) exactly; the aggregator never sees an individual Q(w_k).
// Phase A (client side, local): every GPU writes the ring-sum of ITS clients' masked vectors (the masks
// are produced by Philox4x32-10 in registers; pairs resident on the same GPU cancel and are skipped).
// Phase B (aggregator, sharded): integer sum over GPUs (multimem.ld_reduce.add.u32 or peer loads),
// dequantise, divide by the client count (unweighted mean like Server.aggregate), broadcast.
// =====================================================================================================
struct SecAggArgs {
    Peers pr;
    const float* clients[kMaxWorld];   // per rank: [k_local][stride]
    uint32_t* ybuf[kMaxWorld];         // per rank: [n] masked partial sums (symmetric)
    uint32_t* mc_ybuf;
    float* out[kMaxWorld];             // per rank: [stride] averaged weights (written at offset)
    float* mc_out;
    int k_local;
    long long stride, offset, n;       // protected range (multiple of 4)
    int frac_bits;
    unsigned long long session_seed;
};

__device__ __forceinline__ uint32_t quantise(float w, float scale) {
    return static_cast<uint32_t>(__float2int_rn(w * scale));
}

__global__ void __launch_bounds__(kCommThreads) secagg_kernel(SecAggArgs a) {
    const Peers& pr = a.pr;
    const uint32_t e0 = load_epoch(pr);
    const int total = pr.world * a.k_local;
    const float scale = static_cast<float>(1u << a.frac_bits);
    // ---- phase A: masked ring-sum of the local clients ------------------------------------------------
    const long long n4 = a.n >> 2;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
         i += (long long)gridDim.x * blockDim.x) {
        uint4 acc = make_uint4(0u, 0u, 0u, 0u);
        for (int kl = 0; kl < a.k_local; ++kl) {
            const int k = pr.rank * a.k_local + kl;   // global client id
            const float4 w = *reinterpret_cast<const float4*>(a.clients[pr.rank] + kl * a.stride + a.offset + (i << 2));
            acc.x += quantise(w.x, scale); acc.y += quantise(w.y, scale);
            acc.z += quantise(w.z, scale); acc.w += quantise(w.w, scale);
            // Every pair (k, j) with j on ANOTHER GPU needs its own mask block (pairs on this GPU cancel inside this
            // partial sum and are skipped). Walking the peers rank by rank keeps the division out of the loop and makes
            // the sign a per-peer constant: clients of a higher rank have higher ids (k < j: add), of a lower rank
            // lower ids (subtract).
            for (int pq = 0; pq < pr.world; ++pq) {
                if (pq == pr.rank) continue;
                const bool add = pq > pr.rank;
                for (int jl = 0; jl < a.k_local; ++jl) {
                    const int j = pq * a.k_local + jl;
                    const uint4 m = philox4x32_10(static_cast<uint32_t>(i), static_cast<uint32_t>(i >> 32),
                                                  static_cast<uint32_t>(add ? k : j), static_cast<uint32_t>(add ? j : k),
                                                  static_cast<uint32_t>(a.session_seed),
                                                  static_cast<uint32_t>(a.session_seed >> 32));
                    if (add) { acc.x += m.x; acc.y += m.y; acc.z += m.z; acc.w += m.w; }
                    else     { acc.x -= m.x; acc.y -= m.y; acc.z -= m.z; acc.w -= m.w; }
                }
            }
        }
        *reinterpret_cast<uint4*>(a.ybuf[pr.rank] + (i << 2)) = acc;
    }
    grid_and_world_barrier(pr, e0 + 1);   // publishes the partial sums written above
    // ---- phase B: sharded integer sum, dequantise, mean, broadcast ------------------------------------
    long long lo, hi;
    shard_bounds(n4, pr.world, pr.rank, lo, hi);
    const float inv = 1.f / (scale * static_cast<float>(total));
    for (long long i = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < hi;
         i += (long long)gridDim.x * blockDim.x) {
        uint4 s = make_uint4(0u, 0u, 0u, 0u);
        if (a.mc_ybuf != nullptr) {
            s.x = mc_ld_reduce_u32(a.mc_ybuf + (i << 2) + 0);
            s.y = mc_ld_reduce_u32(a.mc_ybuf + (i << 2) + 1);
            s.z = mc_ld_reduce_u32(a.mc_ybuf + (i << 2) + 2);
            s.w = mc_ld_reduce_u32(a.mc_ybuf + (i << 2) + 3);
        } else {
            for (int q = 0; q < pr.world; ++q) {
                const int p = (pr.rank + q) % pr.world;
                const uint4 v = ld_sys_u4(a.ybuf[p] + (i << 2));
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
        }
        const float4 w = make_float4(static_cast<int32_t>(s.x) * inv, static_cast<int32_t>(s.y) * inv,
                                     static_cast<int32_t>(s.z) * inv, static_cast<int32_t>(s.w) * inv);
        const long long e = a.offset + (i << 2);
        if (a.mc_out != nullptr) {
            mc_st_f4(a.mc_out + e, w);
        } else {
            for (int q = 0; q < pr.world; ++q) st_sys_f4(a.out[(pr.rank + q) % pr.world] + e, w);
        }
    }
    grid_and_world_barrier(pr, e0 + 2);
    store_epoch(pr, e0 + 2, pr.pad[pr.rank] + 18);
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int fill_peers(Peers& pr, const idcb_comm_world* w) {
    IDCB_REQUIRE(w != nullptr && w->world >= 1 && w->world <= kMaxWorld && w->rank >= 0 && w->rank < w->world,
                 "comm: bad world description");
    pr.world = w->world;
    pr.rank = w->rank;
    pr.use_barrier = w->use_barrier;
    for (int p = 0; p < w->world; ++p) {
        pr.pad[p] = static_cast<uint32_t*>(w->signal_pads[p]);
        IDCB_REQUIRE(!w->use_barrier || pr.pad[p] != nullptr, "comm: missing signal pad for rank %d", p);
    }
    return 0;
}

// Optional cap on the CTAs of the communication kernels (0 = one per SM). A bucket that is reduced while
// the backward pass is still running should leave most SM issue slots and memory ports to the convolutions.
static int g_comm_max_ctas = 0;

static int comm_grid(long long n4, int world) {
    long long per = ceil_div_ll(n4, world);
    long long g = ceil_div_ll(per, kCommThreads);
    int cap = num_sms();    // persistent and co-resident: the in-kernel barriers need every CTA live
    if (g_comm_max_ctas > 0 && g_comm_max_ctas < cap) cap = g_comm_max_ctas;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}

}  // namespace idcb

using namespace idcb;

extern "C" int idcb_comm_set_max_ctas(int max_ctas) {
    g_comm_max_ctas = max_ctas > 0 ? max_ctas : 0;
    return 0;
}

extern "C" int idcb_allreduce_optim(const idcb_comm_world* w, void* const* grad, void* const* master,
                                    void* const* shadow, void* mc_grad, void* mc_master, void* mc_shadow,
                                    float* s1, float* s2, long long n, int kind, float lr, float p1, float p2,
                                    float eps, float grad_scale, int owner, void* const* scal, float* scal_out,
                                    void* stream) {
    AllreduceOptArgs a{};
    if (fill_peers(a.pr, w)) return 1;
    IDCB_REQUIRE(owner >= -1 && owner < w->world, "allreduce_optim: bad owner rank %d", owner);
    IDCB_REQUIRE((scal == nullptr) == (scal_out == nullptr), "allreduce_optim: scalar slots and output go together");
    a.owner = owner;
    a.scal_out = scal_out;
    for (int p = 0; p < w->world && scal != nullptr; ++p) a.scal[p] = static_cast<const float*>(scal[p]);
    IDCB_REQUIRE(n % 4 == 0, "allreduce_optim: range length must be a multiple of 4");
    IDCB_REQUIRE(kind >= 0 && kind <= 2, "allreduce_optim: unknown optimizer kind %d", kind);
    for (int p = 0; p < w->world; ++p) {
        a.grad[p] = static_cast<float*>(grad[p]);
        a.master[p] = static_cast<float*>(master[p]);
        a.shadow[p] = shadow ? static_cast<__nv_bfloat16*>(shadow[p]) : nullptr;
        IDCB_REQUIRE(a.grad[p] && a.master[p], "allreduce_optim: null peer pointer for rank %d", p);
    }
    a.mc_grad = static_cast<float*>(mc_grad);
    a.mc_master = static_cast<float*>(mc_master);
    a.mc_shadow = static_cast<__nv_bfloat16*>(mc_shadow);
    IDCB_REQUIRE((a.mc_grad == nullptr) == (a.mc_master == nullptr), "allreduce_optim: multicast aliases go together");
    a.s1 = s1; a.s2 = s2; a.n = n; a.kind = kind;
    a.lr = lr; a.p1 = p1; a.p2 = p2; a.eps = eps; a.grad_scale = grad_scale;
    if (n == 0) return 0;
    allreduce_optim_kernel<<<comm_grid(n / 4, w->world), kCommThreads, 0, (cudaStream_t)stream>>>(a);
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_stat_reduce(const idcb_comm_world* w, void* const* stats, void* mc_stats, long long n,
                                float scale, void* stream) {
    StatMeanArgs a{};
    if (fill_peers(a.pr, w)) return 1;
    IDCB_REQUIRE(n % 4 == 0, "stat_reduce: length must be a multiple of 4");
    for (int p = 0; p < w->world; ++p) {
        a.stats[p] = static_cast<float*>(stats[p]);
        IDCB_REQUIRE(a.stats[p] != nullptr, "stat_reduce: null peer pointer for rank %d", p);
    }
    a.mc = static_cast<float*>(mc_stats);
    a.n = n;
    a.scale = scale;
    if (n == 0) return 0;
    stat_mean_kernel<<<comm_grid(n / 4, w->world), kCommThreads, 0, (cudaStream_t)stream>>>(a);
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_stat_mean(const idcb_comm_world* w, void* const* stats, void* mc_stats, long long n,
                              void* stream) {
    IDCB_REQUIRE(w != nullptr && w->world >= 1, "stat_mean: bad world description");
    return idcb_stat_reduce(w, stats, mc_stats, n, 1.f / static_cast<float>(w->world), stream);
}

extern "C" int idcb_fedavg_finite_flags(const float* clients, long long stride, long long offset, long long n,
                                        int k_local, float* weights, void* stream) {
    IDCB_REQUIRE(k_local >= 1 && n >= 0, "fedavg_finite_flags: bad sizes");
    if (n == 0) return 0;
    int gx = static_cast<int>(ceil_div_ll(n, 256 * 8));
    if (gx > 4 * num_sms()) gx = 4 * num_sms();
    fedavg_finite_kernel<<<dim3(gx, k_local), 256, 0, (cudaStream_t)stream>>>(clients, stride, offset, n, k_local,
                                                                             weights);
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_fedavg_reduce(const idcb_comm_world* w, void* const* clients, void* const* weights,
                                  void* const* global, void* const* shadow, void* mc_global, void* mc_shadow,
                                  int k_local, long long stride, long long offset, long long n, float server_lr,
                                  void* stream) {
    FedAvgArgs a{};
    if (fill_peers(a.pr, w)) return 1;
    IDCB_REQUIRE(n % 4 == 0 && offset % 4 == 0 && stride % 4 == 0, "fedavg_reduce: sizes must be multiples of 4");
    IDCB_REQUIRE(k_local >= 1 && k_local <= 64, "fedavg_reduce: 1..64 clients per GPU");
    for (int p = 0; p < w->world; ++p) {
        a.clients[p] = static_cast<const float*>(clients[p]);
        a.weights[p] = static_cast<const float*>(weights[p]);
        a.global[p] = static_cast<float*>(global[p]);
        a.shadow[p] = shadow ? static_cast<__nv_bfloat16*>(shadow[p]) : nullptr;
        IDCB_REQUIRE(a.clients[p] && a.weights[p] && a.global[p], "fedavg_reduce: null peer pointer for rank %d", p);
    }
    a.mc_global = static_cast<float*>(mc_global);
    a.mc_shadow = static_cast<__nv_bfloat16*>(mc_shadow);
    a.k_local = k_local; a.stride = stride; a.offset = offset; a.n = n; a.server_lr = server_lr;
    if (n == 0) return 0;
    fedavg_reduce_kernel<<<comm_grid(n / 4, w->world), kCommThreads, 0, (cudaStream_t)stream>>>(a);
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_fedavg_round(const idcb_comm_world* w, void* const* clients, void* const* weights,
                                 void* const* global, void* const* shadow, void* mc_global, void* mc_shadow,
                                 int k_local, long long stride, int n_ranges, const long long* offsets,
                                 const long long* lengths, float server_lr, const float* base, float* out,
                                 long long total, void* stream) {
    FedAvgRoundArgs a{};
    if (fill_peers(a.f.pr, w)) return 1;
    IDCB_REQUIRE(n_ranges >= 0 && n_ranges <= kMaxFedRanges, "fedavg_round: at most %d ranges per launch (got %d)",
                 kMaxFedRanges, n_ranges);
    IDCB_REQUIRE(k_local >= 1 && k_local <= 64, "fedavg_round: 1..64 clients per GPU");
    IDCB_REQUIRE(total >= 0 && total % 4 == 0 && stride % 4 == 0 && total <= stride,
                 "fedavg_round: model size %lld / client stride %lld must be multiples of 4, size <= stride", total, stride);
    IDCB_REQUIRE(base != nullptr && out != nullptr && aligned16(base) && aligned16(out) && base != out,
                 "fedavg_round: base and out must be distinct 16-byte aligned device buffers");
    long long prev = 0, work = 0;
    for (int r = 0; r < n_ranges; ++r) {
        IDCB_REQUIRE(offsets[r] % 4 == 0 && lengths[r] % 4 == 0 && lengths[r] >= 0 && offsets[r] >= prev &&
                         offsets[r] + lengths[r] <= total,
                     "fedavg_round: range %d [%lld, +%lld) must be 4-aligned, ascending and inside the model of %lld",
                     r, offsets[r], lengths[r], total);
        a.off[r] = offsets[r]; a.len[r] = lengths[r];
        prev = offsets[r] + lengths[r];
        work += lengths[r];
    }
    for (int p = 0; p < w->world; ++p) {
        a.f.clients[p] = static_cast<const float*>(clients[p]);
        a.f.weights[p] = static_cast<const float*>(weights[p]);
        a.f.global[p] = static_cast<float*>(global[p]);
        a.f.shadow[p] = shadow ? static_cast<__nv_bfloat16*>(shadow[p]) : nullptr;
        IDCB_REQUIRE(a.f.clients[p] && a.f.weights[p] && a.f.global[p], "fedavg_round: null peer pointer for rank %d", p);
        IDCB_REQUIRE(aligned16(a.f.clients[p]) && aligned16(a.f.global[p]), "fedavg_round: peer buffers of rank %d must be 16-byte aligned", p);
    }
    a.f.mc_global = static_cast<float*>(mc_global);
    a.f.mc_shadow = static_cast<__nv_bfloat16*>(mc_shadow);
    a.f.k_local = k_local; a.f.stride = stride; a.f.server_lr = server_lr;
    a.n_ranges = n_ranges; a.base = base; a.out = out; a.total = total;
    if (total == 0) return 0;
    // sized for the local phases (finite scan of k_local clients, copy of the model), which touch more than a shard
    // the phases depend on each other across CTAs (cleared weights, the assembled copy), so without the in-kernel
    // barriers (sequentially emulated ranks; world of one with the barrier switched off) the launch is one CTA
    const long long per_rank = work > total - work ? work : total - work;
    const int grid = a.f.pr.use_barrier ? comm_grid(per_rank / 4 + 1, 1) : 1;
    IDCB_REQUIRE(a.f.pr.use_barrier || w->world == 1, "fedavg_round: ranks emulated one after another cannot run the "
                 "one-launch round (rank r needs every rank's finite flags first)");
    fedavg_round_kernel<<<grid, kCommThreads, 0, (cudaStream_t)stream>>>(a);
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_secagg_masked_sum(const idcb_comm_world* w, void* const* clients, void* const* ybuf,
                                      void* mc_ybuf, void* const* out, void* mc_out, int k_local,
                                      long long stride, long long offset, long long n, int frac_bits,
                                      unsigned long long session_seed, void* stream) {
    SecAggArgs a{};
    if (fill_peers(a.pr, w)) return 1;
    IDCB_REQUIRE(n % 4 == 0 && offset % 4 == 0 && stride % 4 == 0, "secagg: sizes must be multiples of 4");
    IDCB_REQUIRE(n >= 0 && offset >= 0 && offset + n <= stride,
                 "secagg: range [offset=%lld, +n=%lld) does not fit in a client row of %lld elements", offset, n, stride);
    IDCB_REQUIRE(k_local >= 1 && k_local <= 64, "secagg: 1..64 clients per GPU");
    IDCB_REQUIRE(frac_bits >= 1 && frac_bits <= 30, "secagg: frac_bits must be in 1..30");
    for (int p = 0; p < w->world; ++p) {
        a.clients[p] = static_cast<const float*>(clients[p]);
        a.ybuf[p] = static_cast<uint32_t*>(ybuf[p]);
        a.out[p] = static_cast<float*>(out[p]);
        IDCB_REQUIRE(a.ybuf[p] && a.out[p], "secagg: null peer pointer for rank %d", p);
        // the kernel moves 16 bytes at a time
        IDCB_REQUIRE(aligned16(a.ybuf[p]) && aligned16(a.out[p]) && aligned16(a.clients[p]),
                     "secagg: pointers of rank %d must be 16-byte aligned", p);
    }
    IDCB_REQUIRE(a.clients[w->rank] != nullptr, "secagg: local clients pointer is null");
    a.mc_ybuf = static_cast<uint32_t*>(mc_ybuf);
    a.mc_out = static_cast<float*>(mc_out);
    a.k_local = k_local; a.stride = stride; a.offset = offset; a.n = n; a.frac_bits = frac_bits;
    a.session_seed = session_seed;
    if (n == 0) return 0;
    // phase A covers the whole range on every rank: size the grid for that
    long long g = ceil_div_ll(n / 4, kCommThreads);
    if (g > num_sms()) g = num_sms();
    secagg_kernel<<<(int)g, kCommThreads, 0, (cudaStream_t)stream>>>(a);
    IDCB_LAUNCH_CHECK();
    return 0;
}

// =====================================================================================================
// Allocation-free single-GPU secure mean
//
// Some hosts embed this library where device allocation is off the table: a serving loop that has
// handed the whole GPU to a fixed arena, a process that runs under a memory quota set at start-up, or
// code that is already inside a region where cudaMalloc would synchronise the device. For them the
// masked-sum kernel above is offered with its staging memory (the masked partial sums and the barrier
// flags) placed in file-scope __device__ storage that the module itself owns: 4 MiB of uint32 plus a
// 256-byte flag pad per GPU that loads the library, held until the process exits.
//
// The symbol names below are host-side handles, not device addresses; the device address is fetched
// once per device with cudaGetSymbolAddress (with that device current) and cached under a mutex, so only
// the first call on a device pays for the lookup (make that first call outside any stream capture).
// The cache is keyed by device ordinal: after cudaDeviceReset the module is reloaded and the cached
// addresses are stale, so a process that resets a device must not call this entry point on it again.
// Nothing here calls cudaMalloc, cudaFree or an allocating library routine. All calls on one GPU share
// the same staging memory, which keeps whatever the previous call left in it: phase A rewrites every
// word it later reads, and the flag pad is a monotonically increasing epoch by design, so nothing needs
// clearing — but two calls whose kernels overlap on the GPU (different streams) would race on it;
// serialise them.
//
// What "secure" means with one GPU: pairwise masks are only generated between clients that live on
// DIFFERENT GPUs (a pair on one GPU cancels inside that GPU's partial sum and is skipped). Here every
// client is on the same GPU, so no mask is generated at all: the staging buffer holds the plain
// fixed-point ring sum of the quantised weights and the result is the fixed-point mean. The call is the
// arithmetic of the masked sum (same quantisation, same rounding), not a privacy mechanism.
//
// The kernel has a grid-wide barrier between its two phases, so it is launched cooperatively: a grid
// that cannot be co-resident (another tenant holds the SMs' shared memory, a smaller MIG slice) is
// refused with a launch error instead of spinning. A cooperative launch allocates nothing.
// =====================================================================================================
constexpr long long kNoAllocMaxElems = 1LL << 20;   // protected range limit: 1,048,576 weights (4 MiB)
static __device__ __align__(256) uint32_t g_noalloc_ybuf[kNoAllocMaxElems];
static __device__ __align__(256) uint32_t g_noalloc_pad[64];

namespace {
struct NoAllocPtrs {
    uint32_t* ybuf;
    uint32_t* pad;
};
NoAllocPtrs g_noalloc_ptrs[64] = {};
std::mutex g_noalloc_mu;

// once per device, with that device current; concurrent first calls are serialised
int noalloc_lookup(int dev, NoAllocPtrs& out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoAllocPtrs& np = g_noalloc_ptrs[dev];
    if (np.ybuf == nullptr || np.pad == nullptr) {
        void* p = nullptr;
        IDCB_CHECK(cudaGetSymbolAddress(&p, g_noalloc_ybuf));
        void* q = nullptr;
        IDCB_CHECK(cudaGetSymbolAddress(&q, g_noalloc_pad));
        IDCB_REQUIRE(p != nullptr && q != nullptr, "secagg_mean_noalloc: symbol lookup returned a null address");
        np.pad = static_cast<uint32_t*>(q);
        np.ybuf = static_cast<uint32_t*>(p);
    }
    out = np;
    return 0;
}
}  // namespace

extern "C" long long idcb_secagg_noalloc_max_elems(void) { return kNoAllocMaxElems; }

extern "C" int idcb_secagg_mean_noalloc(const float* clients, float* out, int k_clients, long long stride,
                                        long long offset, long long n, int frac_bits,
                                        unsigned long long session_seed, void* stream) {
    IDCB_REQUIRE(clients != nullptr && out != nullptr, "secagg_mean_noalloc: null input/output");
    IDCB_REQUIRE(n % 4 == 0 && offset % 4 == 0 && stride % 4 == 0 && n >= 0, "secagg_mean_noalloc: sizes must be multiples of 4");
    IDCB_REQUIRE(n <= kNoAllocMaxElems,
                 "secagg_mean_noalloc: range of %lld elements exceeds the built-in staging area (%lld); "
                 "use idcb_secagg_masked_sum with caller-provided buffers", n, kNoAllocMaxElems);
    IDCB_REQUIRE(offset >= 0 && offset + n <= stride,
                 "secagg_mean_noalloc: range [offset=%lld, +n=%lld) does not fit in a client row of %lld elements",
                 offset, n, stride);
    IDCB_REQUIRE(aligned16(clients) && aligned16(out),
                 "secagg_mean_noalloc: clients and out must be 16-byte aligned (the kernel moves 16 bytes at a time)");
    IDCB_REQUIRE(k_clients >= 1 && k_clients <= 64, "secagg_mean_noalloc: 1..64 clients");
    IDCB_REQUIRE(frac_bits >= 1 && frac_bits <= 30, "secagg_mean_noalloc: frac_bits must be in 1..30");
    int dev = 0;
    IDCB_CHECK(cudaGetDevice(&dev));
    IDCB_REQUIRE(dev >= 0 && dev < 64, "secagg_mean_noalloc: device index %d out of range", dev);
    NoAllocPtrs np{};
    if (noalloc_lookup(dev, np)) return 1;
    if (n == 0) return 0;
    SecAggArgs a{};
    a.pr.world = 1;
    a.pr.rank = 0;
    a.pr.use_barrier = 1;          // the grid-wide barrier between the summing and the averaging phase
    a.pr.pad[0] = np.pad;
    a.clients[0] = clients;
    a.ybuf[0] = np.ybuf;
    a.out[0] = out;
    a.mc_ybuf = nullptr;
    a.mc_out = nullptr;
    a.k_local = k_clients; a.stride = stride; a.offset = offset; a.n = n; a.frac_bits = frac_bits;
    a.session_seed = session_seed;
    long long g = ceil_div_ll(n / 4, kCommThreads);
    if (g > num_sms()) g = num_sms();
    void* kargs[] = {&a};
    IDCB_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(secagg_kernel), dim3((unsigned)g),
                                           dim3(kCommThreads), kargs, 0, (cudaStream_t)stream));
    return 0;
}
