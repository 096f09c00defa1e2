// tcgen05 / TMEM / TMA implicit-GEMM convolutions for NHWC bf16 activations (sm_100a).
//
// One tile engine, three entry points:
//   idcb_conv_fprop : y = epilogue( conv( prologue(x), W ) )
//   idcb_conv_dgrad : dx = epilogue( conv^T( dy, W ) )            (stride 1 only)
//   idcb_conv_wgrad : dW += prologue(x)^T (*) dy                   (split over positions)
//
// Data flow per CTA:
//   warps 0-7  : workers. dgrad: issue the asynchronous copies (cp.async, zero-fill for padding) of the
//                im2col rows straight into the 128-byte-swizzled operand tile. fprop / wgrad with a BatchNorm
//                scale/shift + ReLU/ReLU6 prologue: read raw chunks from a staging ring, apply the prologue in
//                registers, store the operand tile and arrive on the stage's "full" mbarrier. After the main
//                loop the same eight warps are the epilogue: tcgen05.ld the fp32 accumulators out of TMEM
//                (warp w owns TMEM lanes 32*(w%4).. and column half w/4), apply the fused epilogue and store.
//   warps 8-11 : (fprop, wgrad) loaders: nothing but cp.async + cp.async.mbarrier.arrive, into the operand
//                tile (no prologue) or the raw ring (prologue). No thread that waits, fences or release-arrives
//                ever owns copies in flight (see GemmCfg).
//   next warp  : TMA producer for the second operand (weights, or dy for wgrad).
//   last warp  : allocates TMEM, then one lane issues the generic->async proxy fence, tcgen05.mma
//                (UMMA 128 x N x 16, bf16 -> fp32) and tcgen05.commit to recycle the stage / publish the
//                accumulator.
//
// The problems of this workload are tiny (DenseNet-121 at 50x50: M = 128..21632 rows, K = 64..1152),
// so a CTA lives for a few microseconds and what matters is the length of the per-thread dependent
// instruction chain, not MMA throughput: index math uses precomputed multiply-high division, rows are
// spread over eight warps, and the epilogue code is specialised per mode at compile time.
//
// The activation tile layout [rows = positions][64 channels = 128 B], SWIZZLE_128B, is read as a
// K-major operand by fprop/dgrad (K = channels) and as an MN-major operand by wgrad (K = positions),
// so no transposed copies of activations, gradients or weights ever exist in memory.
//
// The GEMM K axis of fprop and the M axis of wgrad are the LINEARISED (tap, channel) index, so a conv
// with few input channels (the RGB stems, Cin padded to 8) packs several filter taps into one 64-wide
// k-block instead of wasting 7/8 of every MMA.
//
// Replaces the cuDNN convolution calls TensorFlow makes for the reference's Keras models
// (/root/reference/dist_model_tf_vgg.py:119, dist_model_tf_dense.py:131,
// dist_model_tf_mobile.py:119).
#include "common.cuh"
#include "../include/idcb200.h"

namespace idcb {

constexpr int kBM = 128;      // UMMA M (rows of the accumulator = TMEM lanes)
constexpr int kBK = 64;       // K elements per pipeline stage (one 128-byte swizzle row)
constexpr int kMaxStages = 8;
#ifndef IDCB_XFORM_GROUPS
#define IDCB_XFORM_GROUPS 1
#endif
constexpr int kXformGroups = IDCB_XFORM_GROUPS;   // worker-warp groups that transform alternate k-blocks (prologue path); 2 measured slower: DenseNet-121 3.86 vs 3.78 ms
constexpr int kRaw = 6;                          // k-blocks of raw activations in flight per CTA (transform path)
constexpr int kBarBytes = 256;                    // barrier block: full/empty/acc/raw_full/raw_empty + TMEM slot
constexpr int kRawBytes = kBM * 128;             // one k-block of raw 16-byte chunks, [chunk row i][thread]
constexpr int kATileBytes = kBM * 128;  // 16 KiB

// n / d for n*d < 2^64 with one 64-bit multiply-high (magic == 0 encodes d == 1)
struct FastDiv {
    unsigned long long magic;
};
static FastDiv make_fdiv(unsigned d) {
    FastDiv f;
    f.magic = d <= 1 ? 0ULL : (~0ULL) / d + 1ULL;
    return f;
}
__device__ __forceinline__ unsigned fdiv(unsigned n, const FastDiv& f) {
    return f.magic ? static_cast<unsigned>(__umul64hi(static_cast<unsigned long long>(n), f.magic)) : n;
}

struct ConvArgs {
    // activations (A operand source)
    const __nv_bfloat16* x;
    int x_pitch, x_coff;
    int B, H, W, Cin;
    int KH, KW, stride, pad_t, pad_l, Ho, Wo;
    int M;     // B * Ho * Wo
    int Cout;  // N extent
    int flip;  // dgrad: tap index into the weight tensor is mirrored
    int stages;
    FastDiv d_hw, d_wo, d_cin, d_kw, d_cb;
    // prologue on A: a = act(scale[c] * x + shift[c]) (null scale => identity)
    const float* pro_scale;
    const float* pro_shift;
    int pro_act;
    // ... or BatchNorm finalised in-kernel from batch sums (training mode): every CTA derives the
    // per-channel scale/shift itself; CTA (0,0) also publishes them and updates the moving statistics
    const float* bn_sum;
    const float* bn_sq;
    const float* bn_gamma;
    const float* bn_beta;
    float bn_count, bn_eps, bn_momentum;
    float* bn_scale_out;
    float* bn_shift_out;
    float* bn_mean_out;
    float* bn_rstd_out;
    float* bn_mmean;
    float* bn_mvar;
    // epilogue (mode is a template parameter)
    const float* epi_scale;
    const float* epi_shift;
    int epi_act;
    const __nv_bfloat16* aux;  // residual (mode 0) / relu-mask source (mode 1) / raw x (mode 2)
    int aux_pitch, aux_coff;
    const float* bwd_scale;  // mode 2: the forward prologue's scale/shift/act of the consumer
    const float* bwd_shift;
    int bwd_act;
    __nv_bfloat16* y;  // may be null in mode 2 when only G is accumulated
    int y_pitch, y_coff;
    float* gacc;  // mode 2: G[m, n] += bwd_scale[n] * dpre
    int g_pitch, g_coff;
    float* stat_a;  // per-N column sums: mode 0/1: sum(y); mode 2: sum(dpre)
    float* stat_b;  //                    mode 0: sum(y^2);  mode 2: sum(dpre * x_raw)
    int w_cols_per_tap;  // dgrad: column stride between taps in the 2-D weight view
    long long* dbg;      // optional: per-phase clock64 stamps of CTA (0,0) (profiling aid)
    // mode 2, optional: the backward of the BatchNorm whose output gradient this kernel produces, finished in
    // the same launch. tail = 1 (training BN): after every CTA has added its column sums to stat_a / stat_b the
    // grid meets at a barrier (the accumulators stay in TMEM meanwhile), then each CTA turns its own dpre tile
    // into dx = s*(dpre - sa/N - xhat*dgamma/N) and the first row of CTAs emits dgamma / dbeta.
    // tail = 2 (frozen BN): dx = s * dpre, no barrier. Needs the whole grid co-resident (checked on the host).
    // A operand through TMA im2col (one cp.async.bulk.tensor...im2col per k-block issued by one thread) instead
    // of per-thread cp.async gathers. Possible when every 64-wide k-block lies inside one filter tap
    // (single tap, or Cin a multiple of 64). Both operands are then written by the async proxy, so the plain path
    // needs no generic->async fence at all; on the prologue path the transforming workers fence their own stores.
    int a_tma;
    // fprop only: the K loop split over a thread-block cluster of `ksplit` CTAs along grid z (1 = off). CTA r of
    // the cluster runs k-blocks [r * ceil(nkb / ksplit), ...); ranks > 0 push their fp32 partial tile into rank 0's
    // shared memory (st.shared::cluster, buffer at byte offset red_off of the aligned dynamic shared memory,
    // laid out [source][column group of 4][row][4 floats]); after a cluster barrier rank 0 adds them to its own
    // accumulator in the ordinary epilogue. Small-M layers (DenseNet blocks 3-4: 1 to 9 row tiles) otherwise run
    // their whole K loop, BatchNorm prologue included, on a handful of SMs.
    int ksplit, red_off;
    int tail;
    __nv_bfloat16* tail_out;
    int tail_pitch, tail_coff;
    const float* tail_mean;
    const float* tail_rstd;
    float tail_count;
    float* tail_dgamma;
    float* tail_dbeta;
    unsigned int* tail_counter;   // zero at launch
    // tail = 3 (DenseNet block buffer; mode 2 accumulating into G, N tile = 32): besides the BatchNorm in front of
    // THIS convolution (tail_mean / tail_rstd / tail_dgamma / tail_dbeta describe it), the gradient of the 32
    // buffer channels [sl_c0, sl_c0 + 32) — the slice the next-lower layer produced — is final once this kernel's
    // reductions have landed. After a grid barrier the CTAs of that N tile read G back, fold this BatchNorm's share
    // of the batch-statistics correction into S1 / S2 on the fly and write the slice gradient
    //   dy = G - S1/N - xhat * S2/N   (bf16, to tail_out),
    // and one of them does the bookkeeping idcb_bn_bwd_coeffs would do (dgamma, dbeta, S1/S2 of the other channels).
    float* sl_acc_s1;
    float* sl_acc_s2;
    const float* sl_mean;         // the block buffer's batch statistics, indexed by block channel
    const float* sl_rstd;
    int sl_c0;
};

#define IDCB_STAMP(i)                                                                     \
    do {                                                                                  \
        if (p.dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0) p.dbg[i] = clock64(); \
    } while (0)

// Butterfly transpose-reduce over kW (32 or 16) per-lane values: on return lane l holds the warp-wide
// sum of column (l % kW).
template <int kW>
__device__ __forceinline__ float warp_column_sums(float (&v)[kW], int lane) {
    if (kW == 16) {
#pragma unroll
        for (int i = 0; i < kW; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], 16);
    }
#pragma unroll
    for (int n = (kW == 32 ? 16 : 8); n >= 1; n >>= 1) {
        const bool up = (lane & n) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            float send = up ? v[i] : v[i + n];
            float keep = up ? v[i + n] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, n);
        }
    }
    return v[0];
}

template <int kAct>
__device__ __forceinline__ void transform8_act(uint4& q, const float4& s0, const float4& s1,
                                               const float4& h0, const float4& h1) {
    q.x = pack_act_bf16x2<kAct>(fmaf(bf16_lo(q.x), s0.x, h0.x), fmaf(bf16_hi(q.x), s0.y, h0.y));
    q.y = pack_act_bf16x2<kAct>(fmaf(bf16_lo(q.y), s0.z, h0.z), fmaf(bf16_hi(q.y), s0.w, h0.w));
    q.z = pack_act_bf16x2<kAct>(fmaf(bf16_lo(q.z), s1.x, h1.x), fmaf(bf16_hi(q.z), s1.y, h1.y));
    q.w = pack_act_bf16x2<kAct>(fmaf(bf16_lo(q.w), s1.z, h1.z), fmaf(bf16_hi(q.w), s1.w, h1.w));
}
// a = act(scale[c] * x + shift[c]) on 8 packed bf16 channels (act is warp-uniform)
__device__ __forceinline__ void transform8(uint4& q, const float4& s0, const float4& s1,
                                           const float4& h0, const float4& h1, int act) {
    if (act == 1) transform8_act<1>(q, s0, s1, h0, h1);
    else if (act == 2) transform8_act<2>(q, s0, s1, h0, h1);
    else transform8_act<0>(q, s0, s1, h0, h1);
}

template <int kW>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, float (&v)[kW]) {
    if (kW == 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr, r);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(r[e]);
    } else {
        uint32_t r[16];
        tmem_ld_32x32b_x16(taddr, r);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; ++e) v[e] = __uint_as_float(r[e]);
    }
}

// bf16 row segment of kW channels -> float (zero where n >= limit or the row is invalid)
template <int kW>
__device__ __forceinline__ void load_row_bf16(const __nv_bfloat16* rowptr, bool rowvalid, int nb,
                                              int limit, float (&out)[kW]) {
#pragma unroll
    for (int g = 0; g < kW / 8; ++g) {
        uint4 q = make_uint4(0u, 0u, 0u, 0u);
        if (rowvalid && nb + g * 8 < limit) q = *reinterpret_cast<const uint4*>(rowptr + g * 8);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(&q);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float2 f = unpack_bf16x2(w[e]);
            out[g * 8 + 2 * e] = f.x;
            out[g * 8 + 2 * e + 1] = f.y;
        }
    }
}

// kW bf16 values held in registers as 16-byte groups -> float
template <int kW>
__device__ __forceinline__ void unpack_row_bf16(const uint4 (&q)[kW / 8], float (&out)[kW]) {
#pragma unroll
    for (int g = 0; g < kW / 8; ++g) {
        const uint32_t w[4] = {q[g].x, q[g].y, q[g].z, q[g].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            out[g * 8 + 2 * e] = bf16_lo(w[e]);
            out[g * 8 + 2 * e + 1] = bf16_hi(w[e]);
        }
    }
}
// the raw 16-byte groups of a bf16 row segment (zeros where n >= limit or the row is invalid); issued early so that
// the global latency overlaps the main loop
template <int kW>
__device__ __forceinline__ void prefetch_row_bf16(const __nv_bfloat16* rowptr, bool rowvalid, int nb, int limit,
                                                  uint4 (&q)[kW / 8]) {
#pragma unroll
    for (int g = 0; g < kW / 8; ++g) {
        q[g] = make_uint4(0u, 0u, 0u, 0u);
        if (rowvalid && nb + g * 8 < limit) q[g] = *reinterpret_cast<const uint4*>(rowptr + g * 8);
    }
}

// One kW-column slab of the fused epilogue for the row owned by this thread.
// es_scale / es_shift (mode 2): the consumer's forward scale/shift for this slab's columns, staged in shared
// memory by the CTA (zero beyond Cout).
// `auxq`: the slab's aux row segment, already in registers (kEpi 1 / 2: prefetched before the main loop ended).
template <int kEpi, int kW>
__device__ __forceinline__ void epilogue_slab(const ConvArgs& p, uint32_t taddr, int m, bool mvalid,
                                              int nb, int lane, const float* es_scale, const float* es_shift,
                                              const uint4 (&auxq)[kW / 8], uint32_t red = 0u, int red_n = 0,
                                              uint32_t red_stride = 0u) {
    float vv[kW];
    tmem_ld_cols<kW>(taddr, vv);
    if (red != 0u) {
        // K split over a cluster: add the partial tiles the other ranks pushed into this CTA's shared memory
        for (int r = 0; r < red_n; ++r) {
#pragma unroll
            for (int g4 = 0; g4 < kW / 4; ++g4) {
                const float4 t = ld_shared_f4(red + static_cast<uint32_t>(r) * red_stride + static_cast<uint32_t>(g4) * 2048u);
                vv[g4 * 4 + 0] += t.x; vv[g4 * 4 + 1] += t.y; vv[g4 * 4 + 2] += t.z; vv[g4 * 4 + 3] += t.w;
            }
        }
    }
    float xr[kW];
    const bool need_aux = (kEpi != 0) || (p.aux != nullptr);
    if (kEpi != 0)
        unpack_row_bf16<kW>(auxq, xr);
    else if (need_aux)
        load_row_bf16<kW>(p.aux + static_cast<size_t>(m) * p.aux_pitch + p.aux_coff + nb, mvalid, nb,
                          p.Cout, xr);
    if (kEpi == 0) {
        if (p.epi_scale != nullptr) {
#pragma unroll
            for (int e = 0; e < kW; ++e) vv[e] *= (nb + e < p.Cout) ? __ldg(p.epi_scale + nb + e) : 0.f;
        }
        if (p.epi_shift != nullptr) {
#pragma unroll
            for (int e = 0; e < kW; ++e) vv[e] += (nb + e < p.Cout) ? __ldg(p.epi_shift + nb + e) : 0.f;
        }
        if (p.epi_act == 1) {
#pragma unroll
            for (int e = 0; e < kW; ++e) vv[e] = fmaxf(vv[e], 0.f);
        } else if (p.epi_act == 2) {
#pragma unroll
            for (int e = 0; e < kW; ++e) vv[e] = fminf(fmaxf(vv[e], 0.f), 6.f);
        }
        if (need_aux) {
#pragma unroll
            for (int e = 0; e < kW; ++e) vv[e] += xr[e];
        }
    } else if (kEpi == 1) {
#pragma unroll
        for (int e = 0; e < kW; ++e) vv[e] = xr[e] > 0.f ? vv[e] : 0.f;
    } else {
        // BN + activation backward: dpre = acc * act'(scale*x + shift); xr becomes the G scale
        const int act = p.bwd_act;
#pragma unroll
        for (int g4 = 0; g4 < kW / 4; ++g4) {
            const float4 gs = *reinterpret_cast<const float4*>(es_scale + g4 * 4);
            const float4 gh = *reinterpret_cast<const float4*>(es_shift + g4 * 4);
            // columns beyond Cout carry scale = shift = 0 and a zero accumulator (zero-filled weight tile)
            vv[g4 * 4 + 0] *= act_mask(fmaf(gs.x, xr[g4 * 4 + 0], gh.x), act);
            vv[g4 * 4 + 1] *= act_mask(fmaf(gs.y, xr[g4 * 4 + 1], gh.y), act);
            vv[g4 * 4 + 2] *= act_mask(fmaf(gs.z, xr[g4 * 4 + 2], gh.z), act);
            vv[g4 * 4 + 3] *= act_mask(fmaf(gs.w, xr[g4 * 4 + 3], gh.w), act);
        }
    }
    uint32_t packed[kW / 2];
#pragma unroll
    for (int e = 0; e < kW / 2; ++e) packed[e] = pack_bf16x2(vv[2 * e], vv[2 * e + 1]);
    if (kEpi == 0 && p.stat_a != nullptr) {
        // BatchNorm statistics must describe exactly what the consumers will read back (bf16)
#pragma unroll
        for (int e = 0; e < kW / 2; ++e) {
            const float2 f = unpack_bf16x2(packed[e]);
            vv[2 * e] = f.x;
            vv[2 * e + 1] = f.y;
        }
    }
    if (mvalid && p.y != nullptr) {
        __nv_bfloat16* yrow = p.y + static_cast<size_t>(m) * p.y_pitch + p.y_coff + nb;
#pragma unroll
        for (int g = 0; g < kW / 8; ++g)
            if (nb + g * 8 < p.Cout)
                *reinterpret_cast<uint4*>(yrow + g * 8) =
                    make_uint4(packed[4 * g], packed[4 * g + 1], packed[4 * g + 2], packed[4 * g + 3]);
    }
    if (kEpi == 2 && mvalid && p.gacc != nullptr) {
        float* grow = p.gacc + static_cast<size_t>(m) * p.g_pitch + p.g_coff + nb;
#pragma unroll
        for (int g = 0; g < kW / 4; ++g) {
            if (nb + g * 4 < p.Cout) {
                // fire-and-forget vector reduction: no load of G, nothing to wait for (each element of G is
                // touched once per kernel, so the result does not depend on arrival order)
                const float4 gs = *reinterpret_cast<const float4*>(es_scale + g * 4);
                red_add_f32x4(grow + g * 4, gs.x * vv[g * 4 + 0], gs.y * vv[g * 4 + 1], gs.z * vv[g * 4 + 2],
                              gs.w * vv[g * 4 + 3]);
            }
        }
    }
    // ---- column statistics (BatchNorm sums / bias gradient / BN-backward sums) ----
    if (p.stat_a != nullptr) {
        float sb[kW];
        const bool want_b = (kEpi != 1) && p.stat_b != nullptr;
        if (want_b) {
#pragma unroll
            for (int e = 0; e < kW; ++e) sb[e] = mvalid ? (kEpi == 2 ? vv[e] * xr[e] : vv[e] * vv[e]) : 0.f;
        }
        if (!mvalid) {
#pragma unroll
            for (int e = 0; e < kW; ++e) vv[e] = 0.f;
        }
        const float ca = warp_column_sums<kW>(vv, lane);
        const bool owner = lane < kW && nb + lane < p.Cout;
        if (owner) atomicAdd(p.stat_a + nb + lane, ca);
        if (want_b) {
            const float cb = warp_column_sums<kW>(sb, lane);
            if (owner) atomicAdd(p.stat_b + nb + lane, cb);
        }
    }
}

// Second pass of a mode-2 epilogue with a BatchNorm tail: the same dpre as the first pass (recomputed from the
// accumulator that is still in TMEM and the raw activation row), turned into dx = A*dpre + B*x + C.
template <int kW>
__device__ __forceinline__ void tail_slab(const ConvArgs& p, uint32_t taddr, int m, bool mvalid, int nb,
                                          const float* es_scale, const float* es_shift, const float* es_b,
                                          const float* es_c, const uint4 (&auxq)[kW / 8]) {
    float vv[kW];
    tmem_ld_cols<kW>(taddr, vv);
    float xr[kW];
    unpack_row_bf16<kW>(auxq, xr);
    const int act = p.bwd_act;
    uint32_t packed[kW / 2];
#pragma unroll
    for (int g4 = 0; g4 < kW / 4; ++g4) {
        const float4 gs = *reinterpret_cast<const float4*>(es_scale + g4 * 4);
        const float4 gh = *reinterpret_cast<const float4*>(es_shift + g4 * 4);
        const float4 cb = *reinterpret_cast<const float4*>(es_b + g4 * 4);
        const float4 cc = *reinterpret_cast<const float4*>(es_c + g4 * 4);
        const float d0 = vv[g4 * 4 + 0] * act_mask(fmaf(gs.x, xr[g4 * 4 + 0], gh.x), act);
        const float d1 = vv[g4 * 4 + 1] * act_mask(fmaf(gs.y, xr[g4 * 4 + 1], gh.y), act);
        const float d2 = vv[g4 * 4 + 2] * act_mask(fmaf(gs.z, xr[g4 * 4 + 2], gh.z), act);
        const float d3 = vv[g4 * 4 + 3] * act_mask(fmaf(gs.w, xr[g4 * 4 + 3], gh.w), act);
        packed[g4 * 2 + 0] = pack_bf16x2(fmaf(gs.x, d0, fmaf(cb.x, xr[g4 * 4 + 0], cc.x)),
                                         fmaf(gs.y, d1, fmaf(cb.y, xr[g4 * 4 + 1], cc.y)));
        packed[g4 * 2 + 1] = pack_bf16x2(fmaf(gs.z, d2, fmaf(cb.z, xr[g4 * 4 + 2], cc.z)),
                                         fmaf(gs.w, d3, fmaf(cb.w, xr[g4 * 4 + 3], cc.w)));
    }
    if (mvalid) {
        __nv_bfloat16* orow = p.tail_out + static_cast<size_t>(m) * p.tail_pitch + p.tail_coff + nb;
#pragma unroll
        for (int g = 0; g < kW / 8; ++g)
            if (nb + g * 8 < p.Cout)
                *reinterpret_cast<uint4*>(orow + g * 8) =
                    make_uint4(packed[4 * g], packed[4 * g + 1], packed[4 * g + 2], packed[4 * g + 3]);
    }
}

// tail = 3: the finished gradient of a block-buffer slice for the row owned by this thread:
// out = G + B*x + C (G read back from L2; x = the raw buffer row, already in registers)
template <int kW>
__device__ __forceinline__ void slice_slab(const ConvArgs& p, int m, bool mvalid, int nb, const float* es_b,
                                           const float* es_c, const uint4 (&auxq)[kW / 8]) {
    if (!mvalid) return;
    float xr[kW];
    unpack_row_bf16<kW>(auxq, xr);
    const float* grow = p.gacc + static_cast<size_t>(m) * p.g_pitch + p.g_coff + nb;
    uint32_t packed[kW / 2];
#pragma unroll
    for (int g4 = 0; g4 < kW / 4; ++g4) {
        const float4 g = __ldcg(reinterpret_cast<const float4*>(grow + g4 * 4));
        const float4 cb = *reinterpret_cast<const float4*>(es_b + g4 * 4);
        const float4 cc = *reinterpret_cast<const float4*>(es_c + g4 * 4);
        packed[g4 * 2 + 0] = pack_bf16x2(g.x + fmaf(cb.x, xr[g4 * 4 + 0], cc.x), g.y + fmaf(cb.y, xr[g4 * 4 + 1], cc.y));
        packed[g4 * 2 + 1] = pack_bf16x2(g.z + fmaf(cb.z, xr[g4 * 4 + 2], cc.z), g.w + fmaf(cb.w, xr[g4 * 4 + 3], cc.w));
    }
    __nv_bfloat16* orow = p.tail_out + static_cast<size_t>(m) * p.tail_pitch + p.tail_coff + (nb - p.sl_c0);
#pragma unroll
    for (int g = 0; g < kW / 8; ++g)
        *reinterpret_cast<uint4*>(orow + g * 8) =
            make_uint4(packed[4 * g], packed[4 * g + 1], packed[4 * g + 2], packed[4 * g + 3]);
}

template <int kBN, bool kBMN>
struct FwdSmem {
    static constexpr int kBTileBytes = kBMN ? ((kBN + 63) / 64) * 8192 : kBN * 128;
    static constexpr int kStageBytes = kATileBytes + kBTileBytes;
};

// =============================================================================================
// fprop / dgrad kernel
// =============================================================================================
// Thread layout of conv_gemm_kernel.
//   warps 0-7   workers: epilogue, and (a) dgrad: issue the asynchronous A copies themselves,
//                        (b) fprop with a BatchNorm prologue: transform raw chunks into operand tiles
//   warps 8-11  fprop only: loaders. They do nothing but issue cp.async and let the copies signal an
//               mbarrier on completion, so no thread that must later wait, fence or release-arrive ever has
//               copies in flight (measured: an mbarrier.arrive with release semantics, like a MEMBAR, does
//               not retire until the issuing thread's outstanding copies have landed, which turned every
//               k-block of a register- or cp.async-pipelined producer into a full memory round trip).
//   then the TMA warp (weights) and the MMA warp.
template <bool kBMN>
struct GemmCfg {
    static constexpr int kWork = 256;                      // worker threads
    static constexpr int kLoad = kBMN ? 0 : 128;           // dedicated loader threads
    static constexpr int kNT = kWork + kLoad + 64;         // + TMA warp + MMA warp
    static constexpr int kTmaWarp = (kWork + kLoad) / 32, kMmaWarp = kTmaWarp + 1;
    static constexpr int kIssuers = kBMN ? kWork : kLoad;  // threads whose copies feed a k-block
};

// One thread's share of the implicit-GEMM A operand: 16-byte chunk j of rows rb + i * kStep (i < kR) of
// every k-block. Holds, per row, the element offset of the window origin and a bit mask over the taps whose
// input position is inside the image (zero padding and rows past M are simply cleared bits), plus a
// (tap, channel) cursor that advances one k-block per call — so the main loop does no divisions, coordinate
// arithmetic or bounds checks.
template <int kR, int kStep, bool kBMN>
struct Gather {
    int roff[kR];
    uint32_t vlo[kR], vhi[kR];   // tap-validity bits 0..31 / 32..48
    int cur_c, cur_tap, cur_kw, cur_off, row_wrap, span;

    // kb_first: first k-block this CTA runs (fprop K split over a cluster; 0 otherwise)
    __device__ __forceinline__ void init(const ConvArgs& p, int m0, int t, int kb_first = 0) {
        const int j = t & 7, rb = t >> 3;
#pragma unroll
        for (int i = 0; i < kR; ++i) {
            const int m = m0 + i * kStep + rb;
            roff[i] = 0;
            unsigned long long vm = 0ull;
            if (m < p.M) {
                const int b = fdiv(m, p.d_hw);
                const int rem = m - b * (p.Ho * p.Wo);
                const int ho = fdiv(rem, p.d_wo);
                const int wo = rem - ho * p.Wo;
                const int hb = ho * p.stride - p.pad_t;
                const int wb = wo * p.stride - p.pad_l;
                roff[i] = ((b * p.H + hb) * p.W + wb) * p.x_pitch + p.x_coff;
                // valid kw form one contiguous run [kw_lo, kw_hi); likewise kh
                const int kw_lo = max(0, -wb), kw_hi = min(p.KW, p.W - wb);
                const int kh_lo = max(0, -hb), kh_hi = min(p.KH, p.H - hb);
                if (kw_hi > kw_lo) {
                    const unsigned long long wm = ((1ull << kw_hi) - 1ull) & ~((1ull << kw_lo) - 1ull);
#pragma unroll 1
                    for (int kh = kh_lo; kh < kh_hi; ++kh) vm |= wm << (kh * p.KW);
                }
            }
            vlo[i] = static_cast<uint32_t>(vm);
            vhi[i] = static_cast<uint32_t>(vm >> 32);
        }
        const int klin = (kBMN ? 0 : kb_first * kBK) + j * 8;
        cur_tap = kBMN ? 0 : static_cast<int>(fdiv(klin, p.d_cin));
        cur_c = klin - cur_tap * p.Cin;
        const int kh0 = fdiv(cur_tap, p.d_kw);
        cur_kw = cur_tap - kh0 * p.KW;
        cur_off = (kh0 * p.W + cur_kw) * p.x_pitch;
        row_wrap = (p.W - p.KW) * p.x_pitch;
        span = ((p.Cin + kBK - 1) / kBK) * kBK;
    }
    // Row mask of the next k-block (bit i: row i's chunk holds data rather than zeros); `c` = the chunk's first
    // channel, `base` = element offset to add to roff[i]. Advances the cursor.
    __device__ __forceinline__ uint32_t next(const ConvArgs& p, int taps, int& c, int& base) {
        c = cur_c;
        const int tap = cur_tap;
        const bool cvalid = tap < taps && c < p.Cin;
        base = cur_off + c;
        cur_c += kBK;
        if (kBMN) {
            // dgrad walks whole 64-wide blocks per tap (the tail block of a tap is zero padded)
            if (cur_c >= span) {
                cur_c -= span;
                ++cur_tap;
                cur_off += p.x_pitch;
                if (++cur_kw == p.KW) { cur_kw = 0; cur_off += row_wrap; }
            }
        } else {
            while (cur_c >= p.Cin) {
                cur_c -= p.Cin;
                ++cur_tap;
                cur_off += p.x_pitch;
                if (++cur_kw == p.KW) { cur_kw = 0; cur_off += row_wrap; }
            }
        }
        const uint32_t tb_lo = (cvalid && tap < 32) ? 1u << tap : 0u;
        const uint32_t tb_hi = (cvalid && tap >= 32) ? 1u << (tap - 32) : 0u;
        uint32_t okmask = 0;
#pragma unroll
        for (int i = 0; i < kR; ++i)
            if (((vlo[i] & tb_lo) | (vhi[i] & tb_hi)) != 0u) okmask |= 1u << i;
        return okmask;
    }
    // Start the asynchronous copies (LDGSTS) of the next k-block to dst0 + i * dstep; chunks without data are
    // zero-filled.
    __device__ __forceinline__ void issue(const ConvArgs& p, int taps, uint32_t dst0, uint32_t dstep) {
        int c, base;
        const uint32_t ok = next(p, taps, c, base);
#pragma unroll
        for (int i = 0; i < kR; ++i) {
            const bool v = (ok >> i) & 1u;
            cp_async16(dst0 + i * dstep, p.x + (v ? base + roff[i] : 0), v ? 16u : 0u);
        }
    }
};

template <int kBN, bool kBMN, int kEpi>
__global__ void __launch_bounds__(GemmCfg<kBMN>::kNT, (!kBMN || kBN > 64 || (kEpi == 2 && kBN > 32)) ? 1 : 2)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_a,
                 const __grid_constant__ ConvArgs p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    using L = FwdSmem<kBN, kBMN>;
    using G = GemmCfg<kBMN>;
    constexpr int kWork = G::kWork, kNT = G::kNT;
    constexpr int kTmaWarp = G::kTmaWarp, kMmaWarp = G::kMmaWarp;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    const int nstages = p.stages;
    uint8_t* a_tiles = smem;
    uint8_t* b_tiles = smem + nstages * kATileBytes;
    const bool has_pro = !kBMN && (p.pro_scale != nullptr || p.bn_sum != nullptr);
    uint8_t* raw_tiles = b_tiles + nstages * L::kBTileBytes;   // raw staging ring, present only with a prologue
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(raw_tiles + (has_pro ? kRaw * kRawBytes : 0));
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* acc_bar = empty_bar + kMaxStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);
    uint64_t* raw_full = acc_bar + 2;
    uint64_t* raw_empty = raw_full + kRaw;
    float* s_scale = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(full_bar) + kBarBytes);
    float* s_shift = s_scale + (kEpi == 2 ? 128 : ((p.Cin + 3) & ~3));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * kBM;
    const int n0 = blockIdx.y * kBN;
    const int taps = p.KH * p.KW;
    const int cblocks = (p.Cin + kBK - 1) / kBK;                 // dgrad: k-blocks per tap
    const int ktot = taps * p.Cin;                              // fprop: linearised K extent
    const int num_kb_total = kBMN ? taps * cblocks : (ktot + kBK - 1) / kBK;
    // K split over the cluster (fprop): this CTA's k-blocks are [kb0, kb0 + num_kb) of the layer's
    const int ksplit = kBMN ? 1 : p.ksplit;
    const int krank = ksplit > 1 ? static_cast<int>(blockIdx.z) : 0;
    const int kb_per = (num_kb_total + ksplit - 1) / ksplit;
    const int kb0 = krank * kb_per;
    const int num_kb = (kb0 + kb_per > num_kb_total ? num_kb_total : kb0 + kb_per) - kb0;   // >= 1 (host)
    pdl_launch_dependents();

    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            // transform path: one arrival per warp of the owning worker group; copy path: one per issuing thread; + TMA
            // (with TMA im2col on the plain path one thread announces both operands' bytes)
            mbar_init(&full_bar[s], has_pro ? kWork / 32 / kXformGroups + 1 : (p.a_tma ? 1 : G::kIssuers + 1));
            mbar_init(&empty_bar[s], 1);  // one tcgen05.commit
        }
        mbar_init(acc_bar, 1);
        if (has_pro) {
            for (int s = 0; s < kRaw; ++s) {
                mbar_init(&raw_full[s], p.a_tma ? 1 : G::kLoad);   // every loader thread's copies / one TMA issue
                mbar_init(&raw_empty[s], kWork / 32 / kXformGroups);   // every warp of the group that owns the k-block has read it
            }
        }
        mbar_fence_init();
    }
    if (warp == kTmaWarp && lane == 0) {
        tma_prefetch_desc(&tmap_w);
        if (p.a_tma) tma_prefetch_desc(&tmap_a);
    }
    if (warp == kMmaWarp) tmem_alloc<kBN < 32 ? 32 : kBN>(tmem_slot);
    // Everything above touched only this CTA's shared memory, TMEM and the kernel parameters, so it (and the
    // CTA-wide barrier that publishes it) runs while the predecessor kernel is still finishing.
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // from here on the kernel reads what its predecessors wrote
    pdl_wait();
    if (threadIdx.x == 0) IDCB_STAMP(0);
    if (warp < kWork / 32) {
        // Per-channel prologue / epilogue constants, staged by the worker warps only: the TMA, loader and MMA
        // threads are already issuing the first operand loads while these global reads are in flight.
        // a K-split CTA of a single-tap convolution only needs the channels of its own k-blocks (the CTA that
        // publishes the coefficients at the end needs them all)
        int c_lo = 0, c_hi = p.Cin;
        if (ksplit > 1 && taps == 1 && !(blockIdx.x == 0 && blockIdx.y == 0 && krank == 0)) {
            c_lo = kb0 * kBK;
            c_hi = (kb0 + num_kb) * kBK < p.Cin ? (kb0 + num_kb) * kBK : p.Cin;
        }
        if (p.bn_sum != nullptr) {
            // BatchNorm finalised from the batch sums by every CTA for itself (the publication of the
            // coefficients and the moving-statistics update happen at the very end of the kernel)
            const float inv = 1.f / p.bn_count;
            for (int c = c_lo + threadIdx.x; c < c_hi; c += kWork) {
                const float mean = p.bn_sum[c] * inv;
                const float var = fmaxf(p.bn_sq[c] * inv - mean * mean, 0.f);
                const float sc = p.bn_gamma[c] * rsqrtf(var + p.bn_eps);
                s_scale[c] = sc;
                s_shift[c] = p.bn_beta[c] - mean * sc;
            }
        } else if (has_pro) {
            for (int c = c_lo + threadIdx.x; c < c_hi; c += kWork) {
                s_scale[c] = p.pro_scale[c];
                s_shift[c] = p.pro_shift[c];
            }
        }
        if (kEpi == 2) {
            // the consumer's forward scale/shift for this CTA's columns (BN + activation backward in the epilogue)
            for (int c = threadIdx.x; c < kBN; c += kWork) {
                const bool nv = n0 + c < p.Cout;
                s_scale[c] = nv ? p.bwd_scale[n0 + c] : 0.f;
                s_shift[c] = nv ? p.bwd_shift[n0 + c] : 0.f;
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kWork) : "memory");    // the workers among themselves
    }
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) IDCB_STAMP(1);
    // TMA im2col: window origin of the tile's first row; k-block kb = (tap kb / cblocks, channels 64 * (kb % cblocks))
    int t_w = 0, t_h = 0, t_n = 0;
    if (p.a_tma) {
        t_n = fdiv(m0, p.d_hw);
        const int rem = m0 - t_n * (p.Ho * p.Wo);
        const int ho = fdiv(rem, p.d_wo);
        t_h = ho * p.stride - p.pad_t;
        t_w = (rem - ho * p.Wo) * p.stride - p.pad_l;
    }
    auto issue_a_tma = [&](uint32_t dst, uint32_t bar, int kb_local) {
        const int kb = kb0 + kb_local;
        const int tap = fdiv(kb, p.d_cb);
        const int cb = kb - tap * cblocks;
        if (p.a_tma == 2) {
            // 1x1 stride-1 convolution: the operand is a plain [M, channels] matrix -> tiled 2-D load
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                ::"r"(dst), "l"(reinterpret_cast<uint64_t>(&tmap_a)), "r"(bar), "r"(cb * kBK), "r"(m0) : "memory");
            return;
        }
        const int kh = fdiv(tap, p.d_kw);
        tma_load_im2col(dst, &tmap_a, bar, cb * kBK, t_w, t_h, t_n, tap - kh * p.KW, kh);
    };

    // epilogue geometry of a worker thread: warp w owns TMEM lanes 32*(w%4).. and column half w/4
    constexpr int kHalf = kBN / 2;               // columns per warp
    constexpr int kW = (kHalf >= 32 && (kBMN || kBN < 128)) ? 32 : 16;    // slab width (register budget)
    constexpr int kSlabs = kHalf / kW;
    const int quad = warp & 3;
    const int half = (warp >> 2) & 1;
    const int m = m0 + quad * 32 + lane;
    const bool mvalid = m < p.M;
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    uint4 auxq[kSlabs][kW / 8];      // modes 1 / 2: this thread's aux row (also reused by the BatchNorm tail pass)
    const uint32_t trow = static_cast<uint32_t>(quad * 32 + lane);          // this thread's row of the tile
    const uint32_t red_base = smem_u32(smem) + static_cast<uint32_t>(p.red_off);   // K-split partial tiles (rank 0)
    auto run_epilogue = [&]() {
        const uint32_t red = ksplit > 1 ? red_base + trow * 16u : 0u;
#pragma unroll
        for (int it = 0; it < kSlabs; ++it) {
            const int cc = half * kHalf + it * kW;
            const int nb = n0 + cc;
            if (nb < p.Cout)   // warp-uniform
                epilogue_slab<kEpi, kW>(p, t_row + cc, m, mvalid, nb, lane, s_scale + cc, s_shift + cc, auxq[it],
                                        red != 0u ? red + static_cast<uint32_t>(cc / 4) * 2048u : 0u, ksplit - 1,
                                        kBN * 512u);
        }
        tc_fence_before();
        // Publication of the BatchNorm coefficients for the backward pass + moving-statistics update: one CTA,
        // after everything latency-critical (global stores in flight would stall any later release/fence).
        if (!kBMN && p.bn_sum != nullptr && blockIdx.x == 0 && blockIdx.y == 0) {
            const float inv = 1.f / p.bn_count;
            for (int c = threadIdx.x; c < p.Cin; c += kWork) {
                const float mean = p.bn_sum[c] * inv;
                const float var = fmaxf(p.bn_sq[c] * inv - mean * mean, 0.f);
                const float rstd = rsqrtf(var + p.bn_eps);
                p.bn_scale_out[c] = s_scale[c];
                p.bn_shift_out[c] = s_shift[c];
                p.bn_mean_out[c] = mean;
                p.bn_rstd_out[c] = rstd;
                if (p.bn_mmean != nullptr) {
                    const float unbiased = p.bn_count > 1.f ? var * (p.bn_count / (p.bn_count - 1.f)) : var;
                    p.bn_mmean[c] = p.bn_mmean[c] * p.bn_momentum + mean * (1.f - p.bn_momentum);
                    p.bn_mvar[c] = p.bn_mvar[c] * p.bn_momentum + unbiased * (1.f - p.bn_momentum);
                }
            }
        }
    };

    if (warp < kWork / 32) {
        const int j = threadIdx.x & 7;       // 16-byte chunk inside the 128-byte row
        const int rbase = threadIdx.x >> 3;  // this thread owns tile rows rbase + 32 * i
        const uint32_t full0 = smem_u32(full_bar), empty0 = smem_u32(empty_bar);
        const uint32_t tile_off = sw128_offset(rbase, j);   // row i adds i * 4096 (raw slots use the same layout)
        if (kBMN && p.a_tma) {
            // (dgrad, TMA im2col: the TMA warp feeds both operands; the workers go straight to the epilogue)
        } else if (kBMN) {
            // ===================== dgrad: the workers issue the A copies =====================
            // The chunks land directly in the swizzled operand tile; the stage's "full" barrier hears from
            // every thread when its copies have completed. Nothing here waits for data, so up to `nstages`
            // k-blocks of global traffic are in flight per CTA.
            Gather<4, 32, kBMN> g;
            g.init(p, m0, threadIdx.x);
            const uint32_t a_base = smem_u32(a_tiles) + tile_off;
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                if (kb >= nstages) mbar_wait_s(empty0 + stage * 8, phase ^ 1);   // first pass: slots start empty
                g.issue(p, taps, a_base + stage * kATileBytes, 4096);
                cp_async_mbar_arrive(full0 + stage * 8);
                if (threadIdx.x == 0 && kb == 0) IDCB_STAMP(2);
                if (++stage == nstages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        } else if (has_pro) {
            // ===================== fprop with a prologue: transform raw chunks into operand tiles ==========
            // a = act(scale[c] * x + shift[c]) in registers between the raw ring (filled by the loader warps)
            // and the swizzled operand tile. These threads never have global traffic in flight, so their
            // release-arrive on the "full" barrier retires at once.
            // The worker warps can be split into kGroups groups that take alternate k-blocks (more time per
            // group for its dependent wait -> load -> transform -> store -> arrive chain); one group is fastest.
            constexpr int kGroups = kXformGroups, kGT = kWork / kGroups, kGR = 1024 / kGT;
            const int grp = threadIdx.x / kGT, gt = threadIdx.x % kGT;
            Gather<kGR, kGT / 8, kBMN> g;   // used for its (tap, channel) cursor and padding masks only
            g.init(p, m0, gt, kb0);
            const uint32_t goff = sw128_offset(gt >> 3, gt & 7);   // row i adds i * kGT * 16
            const uint32_t a_base = smem_u32(a_tiles) + goff;
            const uint32_t raw_base = smem_u32(raw_tiles) + goff;
            const uint32_t rfull0 = smem_u32(raw_full), rempty0 = smem_u32(raw_empty);
            const int pro_act = p.pro_act;
            int stage = 0, slot = 0;
            uint32_t phase = 0, rphase = 0;
            auto advance = [&]() {
                if (++stage == nstages) { stage = 0; phase ^= 1; }
                if (++slot == kRaw) { slot = 0; rphase ^= 1; }
            };
            for (int kb = 0; kb < num_kb; ++kb) {
                int c, base;
                const uint32_t ok = g.next(p, taps, c, base);
                if ((kb & (kGroups - 1)) != grp) {   // the other group's k-block
                    advance();
                    continue;
                }
                mbar_wait_s(rfull0 + slot * 8, rphase);
                uint4 v[kGR];
#pragma unroll
                for (int i = 0; i < kGR; ++i) v[i] = ld_shared_v4(raw_base + slot * kRawBytes + i * (kGT * 16));
                if (ok != 0) {
                    const float4 s0 = *reinterpret_cast<const float4*>(s_scale + c);
                    const float4 s1 = *reinterpret_cast<const float4*>(s_scale + c + 4);
                    const float4 h0 = *reinterpret_cast<const float4*>(s_shift + c);
                    const float4 h1 = *reinterpret_cast<const float4*>(s_shift + c + 4);
                    if (pro_act == 1) {
#pragma unroll
                        for (int i = 0; i < kGR; ++i)
                            if (ok & (1u << i)) transform8_act<1>(v[i], s0, s1, h0, h1);
                    } else if (pro_act == 2) {
#pragma unroll
                        for (int i = 0; i < kGR; ++i)
                            if (ok & (1u << i)) transform8_act<2>(v[i], s0, s1, h0, h1);
                    } else {
#pragma unroll
                        for (int i = 0; i < kGR; ++i)
                            if (ok & (1u << i)) transform8_act<0>(v[i], s0, s1, h0, h1);
                    }
                }
                if (threadIdx.x == 0 && kb == 0) IDCB_STAMP(2);
                if (kb >= nstages) mbar_wait_s(empty0 + stage * 8, phase ^ 1);
                const uint32_t a_dst = a_base + stage * kATileBytes;
#pragma unroll
                for (int i = 0; i < kGR; ++i) st_shared_v4(a_dst + i * (kGT * 16), v[i]);
                // legacy path: the generic->async proxy fence for these stores is issued by the MMA thread after
                // its wait; TMA path: every worker fences its own stores (they have nothing else in flight), which
                // takes the fence off the single MMA thread's serial chain
                if (p.a_tma) fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive_s(full0 + stage * 8);
                    mbar_arrive_s(rempty0 + slot * 8);   // the raw slot may be refilled
                }
                advance();
            }
        }
        // (fprop without a prologue: the loader warps feed the operand tiles directly; nothing to do here)

        // ===================== epilogue =====================
        if (threadIdx.x == 0) IDCB_STAMP(3);
        // the aux rows of modes 1 / 2 (ReLU mask source / raw activation) are fetched now, under the main loop
        if (kEpi != 0) {
#pragma unroll
            for (int it = 0; it < kSlabs; ++it)
                prefetch_row_bf16<kW>(p.aux + static_cast<size_t>(m) * p.aux_pitch + p.aux_coff + n0 + half * kHalf + it * kW,
                                      mvalid, n0 + half * kHalf + it * kW, p.Cout, auxq[it]);
        }
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        if (threadIdx.x == 0) IDCB_STAMP(6);
        if (krank != 0) {
            // K split, rank > 0: the fp32 partial tile goes into rank 0's shared memory
            const uint32_t dst0 = cluster_map_shared(red_base, 0) + static_cast<uint32_t>(krank - 1) * (kBN * 512u) + trow * 16u;
#pragma unroll
            for (int it = 0; it < kSlabs; ++it) {
                const int cc = half * kHalf + it * kW;
                if (n0 + cc < p.Cout) {
                    float vv[kW];
                    tmem_ld_cols<kW>(t_row + cc, vv);
#pragma unroll
                    for (int g4 = 0; g4 < kW / 4; ++g4)
                        st_cluster_f4(dst0 + static_cast<uint32_t>(cc / 4 + g4) * 2048u, vv[g4 * 4 + 0], vv[g4 * 4 + 1],
                                      vv[g4 * 4 + 2], vv[g4 * 4 + 3]);
                }
            }
            tc_fence_before();
        }
        // (the epilogue proper follows the role split: a K-split rank 0 first waits for its peers' partial tiles)
    } else if (warp < kTmaWarp) {
        // ===================== fprop loader warps: asynchronous copies only =====================
        const int lt = threadIdx.x - kWork;
        if (p.a_tma) {
            // one thread: raw activations of the prologue path into the raw ring (the plain path's A tiles are
            // issued by the TMA warp together with the weights)
            if (has_pro && lt == 0) {
                const uint32_t raw0 = smem_u32(raw_tiles);
                const uint32_t rfull0 = smem_u32(raw_full), rempty0 = smem_u32(raw_empty);
                int slot = 0;
                uint32_t rphase = 0;
                for (int kb = 0; kb < num_kb; ++kb) {
                    if (kb >= kRaw) mbar_wait_s(rempty0 + slot * 8, rphase ^ 1);
                    mbar_arrive_expect_tx_s(rfull0 + slot * 8, kATileBytes);
                    issue_a_tma(raw0 + slot * kRawBytes, rfull0 + slot * 8, kb);
                    if (++slot == kRaw) {
                        slot = 0;
                        rphase ^= 1;
                    }
                }
            }
        } else {
        Gather<8, 16, kBMN> g;
        g.init(p, m0, lt, kb0);
        const uint32_t tile_off = sw128_offset(lt >> 3, lt & 7);   // row i adds i * 2048
        if (has_pro) {
            const uint32_t raw_base = smem_u32(raw_tiles) + tile_off;
            const uint32_t rfull0 = smem_u32(raw_full), rempty0 = smem_u32(raw_empty);
            int slot = 0;
            uint32_t rphase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                if (kb >= kRaw) mbar_wait_s(rempty0 + slot * 8, rphase ^ 1);
                g.issue(p, taps, raw_base + slot * kRawBytes, 2048);
                cp_async_mbar_arrive(rfull0 + slot * 8);
                if (++slot == kRaw) {
                    slot = 0;
                    rphase ^= 1;
                }
            }
        } else {
            const uint32_t a_base = smem_u32(a_tiles) + tile_off;
            const uint32_t full0 = smem_u32(full_bar), empty0 = smem_u32(empty_bar);
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                if (kb >= nstages) mbar_wait_s(empty0 + stage * 8, phase ^ 1);
                g.issue(p, taps, a_base + stage * kATileBytes, 2048);
                cp_async_mbar_arrive(full0 + stage * 8);
                if (++stage == nstages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
        }   // !a_tma
    } else if (warp == kTmaWarp) {
        // ===================== TMA producer for the weight operand =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* b_tile = b_tiles + stage * L::kBTileBytes;
                const bool with_a = p.a_tma && !has_pro;
                mbar_arrive_expect_tx(&full_bar[stage], L::kBTileBytes + (with_a ? kATileBytes : 0));
                if (with_a) issue_a_tma(smem_u32(a_tiles + stage * kATileBytes), smem_u32(&full_bar[stage]), kb);
                if (!kBMN) {
                    // K-major: rows = output channels, 64 linearised (tap, cin) K elements per row
                    tma_load_2d(b_tile, &tmap_w, &full_bar[stage], (kb0 + kb) * kBK, n0);
                } else {
                    // MN-major: rows = K (the original Cout), 64 N elements per row, per 64-wide chunk
                    const int tap = kb / cblocks;
                    const int cb = kb - tap * cblocks;
                    const int wtap = p.flip ? (taps - 1 - tap) : tap;
#pragma unroll
                    for (int ch = 0; ch < (kBN + 63) / 64; ++ch)
                        tma_load_2d(b_tile + ch * 8192, &tmap_w, &full_bar[stage],
                                    wtap * p.w_cols_per_tap + n0 + ch * 64, cb * kBK);
                }
                if (++stage == nstages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kBM, kBN, 0, kBMN ? 1 : 0);
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                if (kb == 0) IDCB_STAMP(12);
                // The producers' generic-proxy st.shared are ordered before this point by their release
                // arrive; the generic->async proxy fence that must sit on that causality path is issued
                // here, by the thread whose tcgen05.mma reads the tile (it has no loads in flight to drain).
                if (!p.a_tma) fence_proxy_async_smem();
                tc_fence_after();
                if (kb == 0) IDCB_STAMP(4);
                const uint32_t a_addr = smem_u32(a_tiles + stage * kATileBytes);
                const uint32_t b_addr = smem_u32(b_tiles + stage * L::kBTileBytes);
#pragma unroll
                for (int k = 0; k < kBK / 16; ++k) {
                    const uint64_t ad = umma_desc_sw128(a_addr + k * 32, 16, 1024);
                    const uint64_t bd = kBMN ? umma_desc_sw128(b_addr + k * 2048, 8192, 1024)
                                             : umma_desc_sw128(b_addr + k * 32, 16, 1024);
                    umma_bf16(tmem_base, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
                }
                umma_commit(&empty_bar[stage]);
                if (++stage == nstages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(acc_bar);
            IDCB_STAMP(5);
        }
        __syncwarp();
    }
    if (threadIdx.x == 0) IDCB_STAMP(7);
    if (ksplit > 1) {
        // every thread of every CTA of the cluster: the partial tiles are in rank 0's shared memory after this
        cluster_arrive_release();
        cluster_wait_acquire();
    }
    if (krank == 0 && warp < kWork / 32) run_epilogue();
    __syncthreads();
    if (threadIdx.x == 0) IDCB_STAMP(8);
    if (kEpi == 2 && kBN == 32 && p.tail == 3) {
        float* s_cb = s_scale + 256;
        float* s_cc = s_scale + 384;
        // every CTA: its G reductions and column sums (issued before the __syncthreads above) are published and it
        // is counted; only the CTAs of the slice's N tile wait for the rest of the grid
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(p.tail_counter, 1u);
        }
        if (n0 == p.sl_c0) {
            if (threadIdx.x == 0) {
                const unsigned int total = gridDim.x * gridDim.y;
                unsigned int seen;
                long long t0 = clock64();
                do {
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.tail_counter) : "memory");
                    if (seen < total && clock64() - t0 > 8000000000LL) __trap();   // a grid that is not co-resident
                } while (seen < total);
            }
            __syncthreads();
            const float invn = 1.f / p.tail_count;
            if (threadIdx.x < kBN) {
                const int c = threadIdx.x, n = n0 + c;
                const float a = __ldcg(p.stat_a + n), b = __ldcg(p.stat_b + n), sc = s_scale[c];
                const float dg = p.tail_rstd[n] * (b - p.tail_mean[n] * a);
                const float s1 = __ldcg(p.sl_acc_s1 + n) + sc * a, s2 = __ldcg(p.sl_acc_s2 + n) + sc * dg;
                const float rt = p.sl_rstd[n], mt = p.sl_mean[n];
                s_cb[c] = -rt * s2 * invn;
                s_cc[c] = -s1 * invn + rt * mt * s2 * invn;
            }
            if (blockIdx.x == 0) {
                // this BatchNorm's parameter gradients and its S1 / S2 share for every channel outside the slice
                for (int pc = threadIdx.x; pc < p.Cout; pc += kNT) {
                    const float a = __ldcg(p.stat_a + pc), b = __ldcg(p.stat_b + pc), sc = p.bwd_scale[pc];
                    const float dg = p.tail_rstd[pc] * (b - p.tail_mean[pc] * a);
                    if (p.tail_dgamma != nullptr) p.tail_dgamma[pc] += dg;
                    if (p.tail_dbeta != nullptr) p.tail_dbeta[pc] += a;
                    if (pc < p.sl_c0 || pc >= p.sl_c0 + kBN) {
                        p.sl_acc_s1[pc] += sc * a;
                        p.sl_acc_s2[pc] += sc * dg;
                    }
                }
            }
            __syncthreads();
            if (warp < kWork / 32) {
#pragma unroll
                for (int it = 0; it < kSlabs; ++it) {
                    const int cc = half * kHalf + it * kW;
                    slice_slab<kW>(p, m, mvalid, n0 + cc, s_cb + cc, s_cc + cc, auxq[it]);
                }
            }
            __syncthreads();
        }
    } else if (kEpi == 2 && p.tail != 0) {
        float* s_cb = s_scale + 256;
        float* s_cc = s_scale + 384;
        if (p.tail == 1) {
            // grid barrier: this CTA's column sums (atomics issued before the __syncthreads above) are made
            // visible, it is counted, and it waits for everybody else's
            if (threadIdx.x == 0) {
                __threadfence();
                atomicAdd(p.tail_counter, 1u);
                const unsigned int total = gridDim.x * gridDim.y;
                unsigned int seen;
                long long t0 = clock64();
                do {
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.tail_counter) : "memory");
                    if (seen < total && clock64() - t0 > 8000000000LL) __trap();   // a grid that is not co-resident
                } while (seen < total);
            }
            __syncthreads();
        }
        const float invn = 1.f / p.tail_count;
        for (int c = threadIdx.x; c < kBN; c += kNT) {
            const int n = n0 + c;
            float cb = 0.f, cc = 0.f;
            if (p.tail == 1 && n < p.Cout) {
                const float a = __ldcg(p.stat_a + n), b = __ldcg(p.stat_b + n);
                const float m = p.tail_mean[n], r = p.tail_rstd[n], sc = s_scale[c];
                const float dg = r * (b - m * a);
                cb = -sc * r * dg * invn;
                cc = -sc * a * invn + sc * r * m * dg * invn;
                if (blockIdx.x == 0) {
                    if (p.tail_dgamma != nullptr) p.tail_dgamma[n] += dg;
                    if (p.tail_dbeta != nullptr) p.tail_dbeta[n] += a;
                }
            }
            s_cb[c] = cb;
            s_cc[c] = cc;
        }
        __syncthreads();
        tc_fence_after();
        if (warp < kWork / 32) {
#pragma unroll
            for (int it = 0; it < kSlabs; ++it) {
                const int cc = half * kHalf + it * kW;
                const int nb = n0 + cc;
                if (nb < p.Cout)   // warp-uniform
                    tail_slab<kW>(p, t_row + cc, m, mvalid, nb, s_scale + cc, s_shift + cc, s_cb + cc, s_cc + cc,
                                  auxq[it]);
            }
            tc_fence_before();
        }
        __syncthreads();
    }
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc<kBN < 32 ? 32 : kBN>(tmem_base);
    }
}

// =============================================================================================
// wgrad kernel: dW[cout][tap*Cin + cin] += sum_pos  act(s*x+t)[pos+tap][cin] * dy[pos][cout]
//   D (TMEM) = [128 linearised (tap,cin) rows  x  kBN cout], both operands MN-major, K = positions.
//   grid = (ceil(taps*Cin/128), cout tiles, splits over positions)
// =============================================================================================
struct WgradArgs {
    const __nv_bfloat16* x;
    int x_pitch, x_coff;
    int B, H, W, Cin;
    int KH, KW, stride, pad_t, pad_l, Ho, Wo;
    int M, Cout;
    FastDiv d_hw, d_wo, d_cin, d_kw;
    const float* pro_scale;
    const float* pro_shift;
    int pro_act;
    float* dw;  // fp32 [Cout][KH*KW*Cin] (row pitch dw_pitch), accumulated with red.add
    int dw_pitch;
    int splits;
    int kb_per_split;  // 64-position blocks per split
    int stages;
};

// One thread's share of the wgrad A operand: 16-byte chunk j16 (8 of the tile's 128 linearised (tap, cin) columns,
// fixed for the thread) of positions rb + i * kStep (i < kR) of every 64-position k-block. A cursor per row —
// (ho, wo), position index and the element offset of the input pixel this thread's tap reads — advances by 64
// positions per k-block with two carries: no divisions and no coordinate rebuild in the main loop.
template <int kR, int kStep>
struct WgradRows {
    int r_ho[kR], r_wo[kR], r_off[kR], r_m[kR];
    int s_ho, s_wo, step_off, wrapw_off, wraph_off, ch, cw, st;
    bool cvalid;

    __device__ __forceinline__ void init(const WgradArgs& p, int kb_begin, int rb, int tap, int c, bool cv) {
        const int hw = p.Ho * p.Wo;
        const int kh = fdiv(tap, p.d_kw);
        const int kw = tap - kh * p.KW;
        const int s_b = 64 / hw;
        s_ho = (64 - s_b * hw) / p.Wo;
        s_wo = 64 - s_b * hw - s_ho * p.Wo;
        st = p.stride;
        step_off = ((s_b * p.H + s_ho * st) * p.W + s_wo * st) * p.x_pitch;
        wrapw_off = (st * p.W - p.Wo * st) * p.x_pitch;
        wraph_off = (p.H * p.W - p.Ho * st * p.W) * p.x_pitch;
        ch = kh - p.pad_t;
        cw = kw - p.pad_l;
        cvalid = cv;
#pragma unroll
        for (int i = 0; i < kR; ++i) {
            const int m = kb_begin * 64 + i * kStep + rb;
            const int b = fdiv(m, p.d_hw);
            const int rem = m - b * hw;
            r_ho[i] = fdiv(rem, p.d_wo);
            r_wo[i] = rem - r_ho[i] * p.Wo;
            r_m[i] = m;
            r_off[i] = ((b * p.H + r_ho[i] * st + ch) * p.W + r_wo[i] * st + cw) * p.x_pitch + p.x_coff + c;
        }
    }
    // validity mask of the next k-block's rows (bit i) and their element offsets; advances the cursor
    __device__ __forceinline__ uint32_t next(const WgradArgs& p, int (&off)[kR]) {
        uint32_t okmask = 0;
#pragma unroll
        for (int i = 0; i < kR; ++i) {
            const unsigned hi = static_cast<unsigned>(r_ho[i] * st + ch);
            const unsigned wi = static_cast<unsigned>(r_wo[i] * st + cw);
            const bool ok = cvalid && r_m[i] < p.M && hi < static_cast<unsigned>(p.H) && wi < static_cast<unsigned>(p.W);
            if (ok) okmask |= 1u << i;
            off[i] = r_off[i];
            r_m[i] += 64;
            r_wo[i] += s_wo;
            r_ho[i] += s_ho;
            r_off[i] += step_off;
            if (r_wo[i] >= p.Wo) { r_wo[i] -= p.Wo; ++r_ho[i]; r_off[i] += wrapw_off; }
            if (r_ho[i] >= p.Ho) { r_ho[i] -= p.Ho; r_off[i] += wraph_off; }
        }
        return okmask;
    }
};

// Thread layout (see conv_gemm_kernel): warps 0-7 workers (prologue transform + epilogue), warps 8-11 loaders
// (asynchronous copies only), warp 12 TMA (dy tiles), warp 13 MMA.
constexpr int kWgWork = 256, kWgLoad = 128, kWgThreads = kWgWork + kWgLoad + 64;
constexpr int kWgTmaWarp = (kWgWork + kWgLoad) / 32, kWgMmaWarp = kWgTmaWarp + 1;

template <int kBN>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ WgradArgs p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    constexpr int kBTile = ((kBN + 63) / 64) * 8192;
    constexpr int kTmaWarp = kWgTmaWarp, kMmaWarp = kWgMmaWarp;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    const int nstages = p.stages;
    uint8_t* a_tiles = smem;  // [64 pos][128 k'] as two 8 KiB chunks
    uint8_t* b_tiles = smem + nstages * kATileBytes;
    const bool has_pro = p.pro_scale != nullptr;
    uint8_t* raw_tiles = b_tiles + nstages * kBTile;   // raw staging ring, present only with a prologue
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(raw_tiles + (has_pro ? kRaw * kRawBytes : 0));
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* acc_bar = empty_bar + kMaxStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);
    uint64_t* raw_full = acc_bar + 2;
    uint64_t* raw_empty = raw_full + kRaw;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int k0 = blockIdx.x * 128;  // first linearised (tap, cin) row of this tile
    const int n0 = blockIdx.y * kBN;
    const int ktot = p.KH * p.KW * p.Cin;
    const int split = blockIdx.z;
    const int total_kb = (p.M + 63) / 64;
    const int kb_begin = split * p.kb_per_split;
    int kb_end = kb_begin + p.kb_per_split;
    if (kb_end > total_kb) kb_end = total_kb;
    const int num_kb = kb_end > kb_begin ? kb_end - kb_begin : 0;
    pdl_launch_dependents();

    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&full_bar[s], has_pro ? kWgWork / 32 + 1 : kWgLoad + 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(acc_bar, 1);
        if (has_pro) {
            for (int s = 0; s < kRaw; ++s) {
                mbar_init(&raw_full[s], kWgLoad);
                mbar_init(&raw_empty[s], kWgWork / 32);
            }
        }
        mbar_fence_init();
    }
    if (warp == kTmaWarp && lane == 0) tma_prefetch_desc(&tmap_dy);
    if (warp == kMmaWarp) tmem_alloc<kBN < 32 ? 32 : kBN>(tmem_slot);
    pdl_wait();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (num_kb == 0) {  // uniform: nothing to do for this split
        __syncthreads();
        if (warp == kMmaWarp) tmem_dealloc<kBN < 32 ? 32 : kBN>(tmem_base);
        return;
    }

    if (warp < kWgTmaWarp) {
        // ---- A side: 16 threads cover one position row (128 k' = 256 B); workers own rows rb + 16 i, loaders rb + 8 i
        const bool is_loader = warp >= kWgWork / 32;
        const int t = is_loader ? threadIdx.x - kWgWork : threadIdx.x;
        const int j16 = t & 15;
        const int half = j16 >> 3;  // which 64-wide chunk
        const int j = j16 & 7;
        const int rbase = t >> 4;   // workers 0..15, loaders 0..7
        const int kidx = k0 + j16 * 8;
        const bool cvalid = kidx < ktot;
        const int tap = cvalid ? fdiv(kidx, p.d_cin) : 0;
        const int c = cvalid ? kidx - tap * p.Cin : 0;
        const uint32_t tile_off = half * 8192 + sw128_offset(rbase, j);
        const uint32_t full0 = smem_u32(full_bar), empty0 = smem_u32(empty_bar);
        const uint32_t rfull0 = smem_u32(raw_full), rempty0 = smem_u32(raw_empty);
        if (is_loader) {
            // ===================== loaders: asynchronous copies only =====================
            WgradRows<8, 8> rows;
            rows.init(p, kb_begin, rbase, tap, c, cvalid);
            const uint32_t dst_base = smem_u32(has_pro ? raw_tiles : a_tiles) + tile_off;   // row i adds i * 1024
            const uint32_t slot_bytes = has_pro ? kRawBytes : kATileBytes;
            const int ring = has_pro ? kRaw : nstages;
            const uint32_t ready0 = has_pro ? rfull0 : full0, free0 = has_pro ? rempty0 : empty0;
            int slot = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                if (kb >= ring) mbar_wait_s(free0 + slot * 8, phase ^ 1);
                int off[8];
                const uint32_t ok = rows.next(p, off);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const bool v = (ok >> i) & 1u;
                    cp_async16(dst_base + slot * slot_bytes + i * 1024, p.x + (v ? off[i] : 0), v ? 16u : 0u);
                }
                cp_async_mbar_arrive(ready0 + slot * 8);
                if (++slot == ring) {
                    slot = 0;
                    phase ^= 1;
                }
            }
        } else {
            if (has_pro) {
                // ===================== workers: BatchNorm/activation transform raw ring -> operand tile ==========
                float4 s0 = make_float4(1.f, 1.f, 1.f, 1.f), s1 = s0;
                float4 h0 = make_float4(0.f, 0.f, 0.f, 0.f), h1 = h0;
                if (cvalid) {
                    s0 = make_float4(p.pro_scale[c], p.pro_scale[c + 1], p.pro_scale[c + 2], p.pro_scale[c + 3]);
                    s1 = make_float4(p.pro_scale[c + 4], p.pro_scale[c + 5], p.pro_scale[c + 6], p.pro_scale[c + 7]);
                    h0 = make_float4(p.pro_shift[c], p.pro_shift[c + 1], p.pro_shift[c + 2], p.pro_shift[c + 3]);
                    h1 = make_float4(p.pro_shift[c + 4], p.pro_shift[c + 5], p.pro_shift[c + 6], p.pro_shift[c + 7]);
                }
                WgradRows<4, 16> rows;   // validity masks only
                rows.init(p, kb_begin, rbase, tap, c, cvalid);
                const uint32_t a_base = smem_u32(a_tiles) + tile_off;       // row i adds i * 2048
                const uint32_t raw_base = smem_u32(raw_tiles) + tile_off;
                const int pro_act = p.pro_act;
                int stage = 0, slot = 0;
                uint32_t phase = 0, rphase = 0;
                for (int kb = 0; kb < num_kb; ++kb) {
                    int off[4];
                    const uint32_t ok = rows.next(p, off);
                    mbar_wait_s(rfull0 + slot * 8, rphase);
                    uint4 v[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[i] = ld_shared_v4(raw_base + slot * kRawBytes + i * 2048);
                    if (ok != 0) {
                        if (pro_act == 1) {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if (ok & (1u << i)) transform8_act<1>(v[i], s0, s1, h0, h1);
                        } else if (pro_act == 2) {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if (ok & (1u << i)) transform8_act<2>(v[i], s0, s1, h0, h1);
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if (ok & (1u << i)) transform8_act<0>(v[i], s0, s1, h0, h1);
                        }
                    }
                    if (kb >= nstages) mbar_wait_s(empty0 + stage * 8, phase ^ 1);
                    const uint32_t a_dst = a_base + stage * kATileBytes;
#pragma unroll
                    for (int i = 0; i < 4; ++i) st_shared_v4(a_dst + i * 2048, v[i]);
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive_s(full0 + stage * 8);
                        mbar_arrive_s(rempty0 + slot * 8);
                    }
                    if (++stage == nstages) { stage = 0; phase ^= 1; }
                    if (++slot == kRaw) { slot = 0; rphase ^= 1; }
                }
            }
        // ---- epilogue: red.add the fp32 tile into dW ----
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        constexpr int kHalf = kBN / 2;
        constexpr int kW = kHalf >= 32 ? 32 : 16;
        const int quad = warp & 3;
        const int hf = warp >> 2;
        const int krow = k0 + quad * 32 + lane;
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
#pragma unroll 1
        for (int cc = hf * kHalf; cc < (hf + 1) * kHalf; cc += kW) {
            if (n0 + cc >= p.Cout) break;
            float vv[kW];
            tmem_ld_cols<kW>(t_row + cc, vv);
            if (krow < ktot) {
#pragma unroll
                for (int e = 0; e < kW; ++e) {
                    const int n = n0 + cc + e;
                    if (n < p.Cout) atomicAdd(p.dw + static_cast<size_t>(n) * p.dw_pitch + krow, vv[e]);
                }
            }
        }
        tc_fence_before();
        }   // workers
    } else if (warp == kTmaWarp) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* b_tile = b_tiles + stage * kBTile;
                mbar_arrive_expect_tx(&full_bar[stage], kBTile);
#pragma unroll
                for (int ch = 0; ch < (kBN + 63) / 64; ++ch)
                    tma_load_2d(b_tile + ch * 8192, &tmap_dy, &full_bar[stage], n0 + ch * 64, kb * 64);
                if (++stage == nstages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kBM, kBN, 1, 1);
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                // The producers' generic-proxy st.shared are ordered before this point by their release
                // arrive; the generic->async proxy fence that must sit on that causality path is issued
                // here, by the thread whose tcgen05.mma reads the tile (it has no loads in flight to drain).
                fence_proxy_async_smem();
                tc_fence_after();
                const uint32_t a_addr = smem_u32(a_tiles + stage * kATileBytes);
                const uint32_t b_addr = smem_u32(b_tiles + stage * kBTile);
#pragma unroll
                for (int k = 0; k < 4; ++k) {  // 64 positions = 4 x UMMA_K(16)
                    const uint64_t ad = umma_desc_sw128(a_addr + k * 2048, 8192, 1024);
                    const uint64_t bd = umma_desc_sw128(b_addr + k * 2048, 8192, 1024);
                    umma_bf16(tmem_base, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
                }
                umma_commit(&empty_bar[stage]);
                if (++stage == nstages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(acc_bar);
        }
        __syncwarp();
    }
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc<kBN < 32 ? 32 : kBN>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------
// Host launchers
// ---------------------------------------------------------------------------------------------
static long long* g_dbg = nullptr;

static int current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    return dev;
}

// N-tile width: the widest tile that still leaves enough CTAs to spread a small problem over the chip
// (a 128-row x 128-column tile of a tiny layer would otherwise run its whole MMA chain on one SM).
static int pick_bn(int M, int N) {
    if (N <= 32) return 32;
    const int mt = ceil_div(M, kBM);
    static const int div = [] { const char* e = getenv("IDCB_CONV_TILE_DIV"); int v = e ? atoi(e) : 2; return v > 0 ? v : 2; }();
    const int target = num_sms() / div;
    if (N > 64 && mt * ceil_div(N, 128) >= target) return 128;
    if (mt * ceil_div(N, 64) >= target) return 64;
    return N <= 64 && mt * 1 >= target ? 64 : 32;
}

// Pipeline depth. Global->shared copy latency is ~2k cycles here, so the number of k-blocks in flight sets
// the per-k-block rate: the copy path takes as many operand stages as fit (they ARE its copies in flight),
// the transform path keeps its depth in the raw staging ring (kRaw) and needs only a short operand ring.
static int pick_stages(int num_kb, bool has_pro, int stage_bytes, int other_bytes, long long ctas = 0) {
    // measured: a deeper operand ring helps the copy path of small grids (fewer, latency-bound CTAs); large
    // grids do better with 4 stages and two co-resident CTAs per SM; the transform path gains nothing past 4
    int cap = (has_pro || ctas > 2 * num_sms()) ? 4 : kMaxStages;
    const int fit = (220 * 1024 - other_bytes) / stage_bytes;
    if (cap > fit) cap = fit;
    int s = num_kb < cap ? num_kb : cap;
    return s < 1 ? 1 : s;
}

// IDCB_CONV_KSPLIT=0 turns the cluster K split of small fprop launches off (A/B measurements)
static bool ksplit_enabled() {
    static const int v = [] { const char* e = getenv("IDCB_CONV_KSPLIT"); return e ? atoi(e) : 1; }();
    return v != 0;
}

// IDCB_CONV_TMA=0 keeps the per-thread cp.async gathers for the activation operand (A/B measurements)
static bool tma_a_enabled() {
    static const int v = [] { const char* e = getenv("IDCB_CONV_TMA"); return e ? atoi(e) : 1; }();
    return v != 0;
}

// Decide whether the activation operand of this fprop / dgrad launch can come through TMA im2col and, if so,
// encode its tensor map (box = 128 output positions x 64 channels of one filter tap).
// `per_tap_blocks`: the kernel walks whole 64-wide channel blocks per tap (dgrad), so any channel count works —
// TMA zero-fills the channels past Cin; fprop's linearised (tap, channel) K axis needs Cin % 64 == 0 or one tap.
static int prep_a_tma(CUtensorMap* tm, ConvArgs& a, bool per_tap_blocks) {
    a.a_tma = 0;
    const int taps = a.KH * a.KW;
    if (!tma_a_enabled() || !(per_tap_blocks || taps == 1 || a.Cin % 64 == 0) || a.stride > 8) return 0;
    const int pad_b = (a.Ho - 1) * a.stride + a.KH - a.H - a.pad_t, pad_r = (a.Wo - 1) * a.stride + a.KW - a.W - a.pad_l;
    const int c[4] = {-a.pad_l, -a.pad_t, pad_r - (a.KW - 1), pad_b - (a.KH - 1)};
    for (int i = 0; i < 4; ++i)
        if (c[i] < -128 || c[i] > 127) return 0;
    static const int tiled_1x1 = [] { const char* e = getenv("IDCB_CONV_TMA_TILED_1X1"); return e ? atoi(e) : 1; }();
    if (tiled_1x1 && taps == 1 && a.stride == 1 && a.pad_t == 0 && a.pad_l == 0 && a.Ho == a.H && a.Wo == a.W) {
        // a 1x1 stride-1 convolution reads a plain [M, Cin] matrix (row pitch x_pitch): tiled 2-D boxes of 128 rows x 64
        // channels, rows past M and channels past Cin zero-filled
        if (make_tmap_2d_bf16(tm, a.x + a.x_coff, (uint64_t)a.M, (uint64_t)a.Cin, (uint64_t)a.x_pitch, kBM, kBK)) return 1;
        a.a_tma = 2;
        return 0;
    }
    if (make_tmap_im2col_bf16(tm, a.x + a.x_coff, a.B, a.H, a.W, a.Cin, a.x_pitch, a.KH, a.KW, a.stride, a.pad_t,
                              a.pad_l, a.Ho, a.Wo, kBM))
        return 1;
    a.a_tma = 1;
    return 0;
}

template <int kBN, bool kBMN, int kEpi>
static int launch_gemm(const CUtensorMap& tm, ConvArgs& a, cudaStream_t st) {
    using L = FwdSmem<kBN, kBMN>;
    const int taps = a.KH * a.KW;
    const int num_kb = kBMN ? taps * ceil_div(a.Cin, kBK) : ceil_div(taps * a.Cin, kBK);
    const int extra = kBarBytes + 2 * (int)sizeof(float) * ((a.Cin > 128 ? a.Cin : 128) + 4) + 64 +
                      (kEpi == 2 ? 1024 : 0);   // mode 2 with a BatchNorm tail: two more 128-float tables
    const bool has_pro = !kBMN && (a.pro_scale != nullptr || a.bn_sum != nullptr);
    // K split over a cluster (fprop of layers with few tiles): as many ranks as keep the grid within the SM count
    // and leave every rank enough k-blocks to pay for the cluster launch, the partial-tile push and the barrier
    int ksplit = 1;
    if (!kBMN && ksplit_enabled()) {
        const long long tiles = 1LL * ceil_div(a.M, kBM) * ceil_div(a.Cout, kBN);
        static const int min_kb = [] { const char* e = getenv("IDCB_CONV_KSPLIT_MINKB"); int v = e ? atoi(e) : 4; return v > 0 ? v : 4; }();
        for (int cand = 4; cand >= 2; cand /= 2)
            if (tiles * cand <= num_sms() && num_kb >= min_kb * cand && (cand - 1) * ceil_div(num_kb, cand) < num_kb) {
                ksplit = cand;
                break;
            }
    }
    const int red_bytes = ksplit > 1 ? (ksplit - 1) * kBN * 512 : 0;
    a.ksplit = ksplit;
    a.stages = pick_stages(ceil_div(num_kb, ksplit), has_pro, L::kStageBytes,
                           1024 + (has_pro ? kRaw * kRawBytes : 0) + extra + red_bytes + 16,
                           kBMN ? 1LL * ceil_div(a.M, kBM) * ceil_div(a.Cout, kBN) : 0);
    a.red_off = (a.stages * L::kStageBytes + (has_pro ? kRaw * kRawBytes : 0) + extra + 15) & ~15;
    const size_t smem = 1024 + a.red_off + red_bytes;
    IDCB_REQUIRE(smem <= 227 * 1024, "conv: Cin=%d needs %zu bytes of shared memory", a.Cin, smem);
    IDCB_REQUIRE(1LL * a.B * a.H * a.W * a.x_pitch < (1LL << 31), "conv: input tensor too large for 32-bit offsets");
    a.d_hw = make_fdiv(a.Ho * a.Wo);
    a.d_wo = make_fdiv(a.Wo);
    a.d_cin = make_fdiv(a.Cin);
    a.d_kw = make_fdiv(a.KW);
    a.d_cb = make_fdiv(ceil_div(a.Cin, kBK));
    auto kern = conv_gemm_kernel<kBN, kBMN, kEpi>;
    static bool attr_set[64] = {};      // the attribute belongs to the (function, device) pair
    const int dev = current_device();
    if (!attr_set[dev]) {
        IDCB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
        // the whole unified array as shared memory: these kernels do not use L1, and co-residency of several
        // small-tile CTAs per SM (what the grid-barrier tails rely on) must not depend on the carve-out the
        // driver would otherwise pick per launch
        IDCB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                        cudaSharedmemCarveoutMaxShared));
        attr_set[dev] = true;
    }
    dim3 grid(ceil_div(a.M, kBM), ceil_div(a.Cout, kBN), ksplit);
    if (kEpi == 2 && (a.tail == 1 || a.tail == 3)) {
        // CTAs of this launch that fit on one SM at the same time. Computed from the kernel's own resource use
        // (registers in 512-register warp granules out of 64 K, shared memory incl. the 1 KiB the driver reserves per
        // CTA out of 228 KiB, 2048 threads): cudaOccupancyMaxActiveBlocksPerMultiprocessor answers for the carve-out
        // it assumes rather than the one the launch gets, and reported 1 where ncu shows 2 resident CTAs.
        cudaFuncAttributes fa;
        IDCB_CHECK(cudaFuncGetAttributes(&fa, kern));
        const int threads = GemmCfg<kBMN>::kNT;
        const int regs_per_cta = ((fa.numRegs * 32 + 511) / 512) * 512 * (threads / 32);
        int per_sm = 65536 / (regs_per_cta > 0 ? regs_per_cta : 1);
        const int by_smem = static_cast<int>((228 * 1024) / (smem + fa.sharedSizeBytes + 1024));
        if (by_smem < per_sm) per_sm = by_smem;
        if (2048 / threads < per_sm) per_sm = 2048 / threads;
        IDCB_REQUIRE(static_cast<long long>(grid.x) * grid.y <= static_cast<long long>(per_sm) * num_sms(),
                     "conv_dgrad: a %u x %u grid with a grid-barrier tail cannot be co-resident (%d CTAs of %zu bytes per SM)",
                     grid.x, grid.y, per_sm, smem);
    }
    a.dbg = g_dbg;
    CUtensorMap tma = tm;      // placeholder when the gather path is used
    if (prep_a_tma(&tma, a, kBMN)) return 1;
    IDCB_CHECK(launch_pdl_cluster(kern, grid, dim3(GemmCfg<kBMN>::kNT), smem, st, ksplit, tm, tma, a));
    return 0;
}

template <int kBN>
static int launch_wgrad(const CUtensorMap& tm, WgradArgs& a, cudaStream_t st) {
    constexpr int kBTile = ((kBN + 63) / 64) * 8192;
    const int extra = kBarBytes + 64;
    a.stages = pick_stages(a.kb_per_split, a.pro_scale != nullptr, kATileBytes + kBTile,
                           1024 + (a.pro_scale != nullptr ? kRaw * kRawBytes : 0) + extra);
    const size_t smem = 1024 + a.stages * (kATileBytes + kBTile) + (a.pro_scale != nullptr ? kRaw * kRawBytes : 0) + extra;
    IDCB_REQUIRE(1LL * a.B * a.H * a.W * a.x_pitch < (1LL << 31), "conv: input tensor too large for 32-bit offsets");
    a.d_hw = make_fdiv(a.Ho * a.Wo);
    a.d_wo = make_fdiv(a.Wo);
    a.d_cin = make_fdiv(a.Cin);
    a.d_kw = make_fdiv(a.KW);
    auto kern = conv_wgrad_kernel<kBN>;
    static bool attr_set[64] = {};
    const int dev = current_device();
    if (!attr_set[dev]) {
        IDCB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
        attr_set[dev] = true;
    }
    dim3 grid(ceil_div(a.KH * a.KW * a.Cin, 128), ceil_div(a.Cout, kBN), a.splits);
    IDCB_CHECK(launch_pdl(kern, grid, dim3(kWgThreads), smem, st, tm, a));
    return 0;
}

static int check_geom(const idcb_conv_desc* d) {
    IDCB_REQUIRE(d != nullptr, "conv: null descriptor");
    IDCB_REQUIRE(d->cin % 8 == 0 && d->cout % 8 == 0,
                 "conv: channel counts must be multiples of 8 (cin=%d cout=%d)", d->cin, d->cout);
    IDCB_REQUIRE(d->x_pitch % 8 == 0 && d->x_coff % 8 == 0 && d->y_pitch % 8 == 0 &&
                     d->y_coff % 8 == 0,
                 "conv: pitches/offsets must be multiples of 8 channels");
    IDCB_REQUIRE(d->kh >= 1 && d->kw >= 1 && d->kh * d->kw <= 49, "conv: bad kernel size");
    IDCB_REQUIRE(d->stride >= 1, "conv: bad stride");
    IDCB_REQUIRE(d->w_pitch >= 0 && d->w_off >= 0 && d->w_pitch % 8 == 0 && d->w_off % 8 == 0 &&
                     (d->w_pitch == 0 || d->w_pitch >= d->kh * d->kw * d->cin),
                 "conv: bad weight window (pitch %d, offset %d)", d->w_pitch, d->w_off);
    const long long m = 1LL * d->batch * d->ho * d->wo;
    IDCB_REQUIRE(m > 0 && m < (1LL << 30), "conv: bad problem size");
    return 0;
}

}  // namespace idcb

using namespace idcb;

// profiling aid: when set (device pointer to >= 16 int64), CTA (0,0) of fprop/dgrad kernels records
// clock64() at its phase boundaries
extern "C" int idcb_conv_set_debug(void* dev_ptr) {
    g_dbg = static_cast<long long*>(dev_ptr);
    return 0;
}

extern "C" int idcb_conv_fprop(const idcb_conv_desc* d, const void* x, const void* w, void* y,
                               const idcb_conv_fused* f, void* stream) {
    if (check_geom(d)) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ConvArgs a{};
    a.x = static_cast<const __nv_bfloat16*>(x);
    a.x_pitch = d->x_pitch;
    a.x_coff = d->x_coff;
    a.B = d->batch; a.H = d->h; a.W = d->w; a.Cin = d->cin;
    a.KH = d->kh; a.KW = d->kw; a.stride = d->stride; a.pad_t = d->pad_t; a.pad_l = d->pad_l;
    a.Ho = d->ho; a.Wo = d->wo;
    a.M = d->batch * d->ho * d->wo;
    a.Cout = d->cout;
    a.flip = 0;
    a.y = static_cast<__nv_bfloat16*>(y);
    a.y_pitch = d->y_pitch; a.y_coff = d->y_coff;
    a.w_cols_per_tap = d->cin;
    if (f) {
        a.pro_scale = f->pro_scale; a.pro_shift = f->pro_shift; a.pro_act = f->pro_act;
        a.epi_scale = f->epi_scale; a.epi_shift = f->epi_shift; a.epi_act = f->epi_act;
        a.aux = static_cast<const __nv_bfloat16*>(f->aux);
        a.aux_pitch = f->aux_pitch; a.aux_coff = f->aux_coff;
        a.stat_a = f->stat_a; a.stat_b = f->stat_b;
        IDCB_REQUIRE((a.pro_scale == nullptr) == (a.pro_shift == nullptr),
                     "conv_fprop: prologue scale and shift must be given together");
        if (f->bn != nullptr) {  // training-mode BatchNorm finalised in-kernel
            const idcb_bn_inline* b = f->bn;
            IDCB_REQUIRE(a.pro_scale == nullptr, "conv_fprop: give either pro_scale/shift or bn, not both");
            IDCB_REQUIRE(b->sum && b->sq && b->gamma && b->beta && b->scale_out && b->shift_out &&
                             b->mean_out && b->rstd_out && b->count > 0.f,
                         "conv_fprop: incomplete inline BatchNorm description");
            IDCB_REQUIRE((b->moving_mean == nullptr) == (b->moving_var == nullptr),
                         "conv_fprop: moving mean/var go together");
            a.bn_sum = b->sum; a.bn_sq = b->sq; a.bn_gamma = b->gamma; a.bn_beta = b->beta;
            a.bn_count = b->count; a.bn_eps = b->eps; a.bn_momentum = b->momentum;
            a.bn_scale_out = b->scale_out; a.bn_shift_out = b->shift_out;
            a.bn_mean_out = b->mean_out; a.bn_rstd_out = b->rstd_out;
            a.bn_mmean = b->moving_mean; a.bn_mvar = b->moving_var;
        }
    }
    CUtensorMap tm;
    const int kbn = pick_bn(a.M, d->cout);
    const uint64_t wk = (uint64_t)d->kh * d->kw * d->cin;
    if (make_tmap_2d_bf16(&tm, static_cast<const __nv_bfloat16*>(w) + d->w_off, d->cout, wk,
                          d->w_pitch ? (uint64_t)d->w_pitch : wk, kbn, 64))
        return 1;
    if (kbn == 32) return launch_gemm<32, false, 0>(tm, a, st);
    if (kbn == 64) return launch_gemm<64, false, 0>(tm, a, st);
    return launch_gemm<128, false, 0>(tm, a, st);
}

// dgrad: the descriptor describes the FORWARD convolution; dy has the forward output geometry and
// dx the forward input geometry (x_pitch/x_coff address dx, y_pitch/y_coff address dy).
extern "C" int idcb_conv_dgrad(const idcb_conv_desc* d, const void* dy, const void* w, void* dx,
                               const idcb_conv_bwd_fused* f, void* stream) {
    if (check_geom(d)) return 1;
    IDCB_REQUIRE(d->stride == 1, "conv_dgrad: only stride 1 is implemented on the tensor-core path");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    ConvArgs a{};
    a.x = static_cast<const __nv_bfloat16*>(dy);
    a.x_pitch = d->y_pitch; a.x_coff = d->y_coff;
    a.B = d->batch; a.H = d->ho; a.W = d->wo; a.Cin = d->cout;
    a.KH = d->kh; a.KW = d->kw; a.stride = 1;
    a.pad_t = d->kh - 1 - d->pad_t;
    a.pad_l = d->kw - 1 - d->pad_l;
    a.Ho = d->h; a.Wo = d->w;
    a.M = d->batch * d->h * d->w;
    a.Cout = d->cin;
    a.flip = 1;
    a.y = static_cast<__nv_bfloat16*>(dx);
    a.y_pitch = d->x_pitch; a.y_coff = d->x_coff;
    a.w_cols_per_tap = d->cin;
    int mode = 0;
    if (f) {
        mode = f->mode;
        a.aux = static_cast<const __nv_bfloat16*>(f->aux);
        a.aux_pitch = f->aux_pitch; a.aux_coff = f->aux_coff;
        a.bwd_scale = f->bwd_scale; a.bwd_shift = f->bwd_shift; a.bwd_act = f->bwd_act;
        a.gacc = f->gacc; a.g_pitch = f->g_pitch; a.g_coff = f->g_coff;
        a.stat_a = f->stat_a; a.stat_b = f->stat_b;
        IDCB_REQUIRE(f->mode >= 0 && f->mode <= 2, "conv_dgrad: bad epilogue mode %d", f->mode);
        IDCB_REQUIRE(f->mode == 0 || f->aux != nullptr, "conv_dgrad: epilogue mode %d needs aux", f->mode);
        IDCB_REQUIRE(f->mode != 2 || (f->bwd_scale && f->bwd_shift),
                     "conv_dgrad: mode 2 needs bwd_scale/bwd_shift");
        IDCB_REQUIRE(f->gacc == nullptr || (f->g_pitch % 4 == 0 && f->g_coff % 4 == 0),
                     "conv_dgrad: G pitch/offset must be multiples of 4");
    }
    int kbn = pick_bn(a.M, d->cin);
    bool split_tail = false;
    if (f && f->tail != nullptr) {
        const idcb_bn_tail* t = f->tail;
        IDCB_REQUIRE(mode == 2 && a.stat_a && a.stat_b, "conv_dgrad: a BatchNorm tail needs mode 2 and both column sums");
        IDCB_REQUIRE(t->out && t->mean && t->rstd && t->count > 0.f && (t->frozen || t->counter),
                     "conv_dgrad: incomplete BatchNorm tail");
        IDCB_REQUIRE(t->out_pitch % 8 == 0 && t->out_coff % 8 == 0, "conv_dgrad: tail pitch/offset must be multiples of 8");
        // mode-2 CTAs take an SM each (launch bounds, ~200 KB of shared memory)
        const long long tiles = 1LL * ceil_div(a.M, kBM) * ceil_div(d->cin, kbn);
        if (t->frozen || tiles <= num_sms()) {
            a.tail = t->frozen ? 2 : 1;
            a.tail_out = static_cast<__nv_bfloat16*>(t->out);
            a.tail_pitch = t->out_pitch; a.tail_coff = t->out_coff;
            a.tail_mean = t->mean; a.tail_rstd = t->rstd; a.tail_count = t->count;
            a.tail_dgamma = t->dgamma; a.tail_dbeta = t->dbeta; a.tail_counter = t->counter;
            a.y = nullptr;     // dpre itself is not needed in memory
        } else {
            IDCB_REQUIRE(a.y != nullptr, "conv_dgrad: the tail does not fit on the device and no dx buffer was given");
            split_tail = true;
        }
    }
    if (f && f->slice != nullptr) {
        const idcb_slice_tail* t = f->slice;
        IDCB_REQUIRE(mode == 2 && a.gacc && a.stat_a && a.stat_b && a.tail == 0 && !split_tail && a.g_coff == 0,
                     "conv_dgrad: a slice tail needs mode 2 into G (channel offset 0), both column sums and no BatchNorm tail");
        IDCB_REQUIRE(t->out && t->acc_s1 && t->acc_s2 && t->mean && t->rstd && t->bn_mean && t->bn_rstd && t->counter &&
                         t->count > 0.f && t->out_pitch % 8 == 0 && t->out_coff % 8 == 0,
                     "conv_dgrad: incomplete slice tail");
        IDCB_REQUIRE(t->c0 % 32 == 0 && t->c0 + 32 <= d->cin, "conv_dgrad: the slice must be one 32-channel N tile inside the output");
        kbn = 32;
        a.tail = 3;
        a.tail_out = static_cast<__nv_bfloat16*>(t->out);
        a.tail_pitch = t->out_pitch; a.tail_coff = t->out_coff;
        a.tail_mean = t->bn_mean; a.tail_rstd = t->bn_rstd; a.tail_count = t->count;
        a.tail_dgamma = t->dgamma; a.tail_dbeta = t->dbeta; a.tail_counter = t->counter;
        a.sl_acc_s1 = t->acc_s1; a.sl_acc_s2 = t->acc_s2; a.sl_mean = t->mean; a.sl_rstd = t->rstd; a.sl_c0 = t->c0;
    }
    IDCB_REQUIRE(a.y != nullptr || a.gacc != nullptr || a.tail != 0, "conv_dgrad: no output given");
    CUtensorMap tm;
    // weight view [Cout rows][taps*Cin cols]; MN-major boxes are 64 cols x 64 rows
    const uint64_t wk = (uint64_t)d->kh * d->kw * d->cin;
    if (make_tmap_2d_bf16(&tm, static_cast<const __nv_bfloat16*>(w) + d->w_off, d->cout, wk,
                          d->w_pitch ? (uint64_t)d->w_pitch : wk, 64, 64))
        return 1;
    int rc;
    if (kbn == 32) {
        rc = mode == 0 ? launch_gemm<32, true, 0>(tm, a, st)
             : mode == 1 ? launch_gemm<32, true, 1>(tm, a, st) : launch_gemm<32, true, 2>(tm, a, st);
    } else if (kbn == 64) {
        rc = mode == 0 ? launch_gemm<64, true, 0>(tm, a, st)
             : mode == 1 ? launch_gemm<64, true, 1>(tm, a, st) : launch_gemm<64, true, 2>(tm, a, st);
    } else {
        rc = mode == 0 ? launch_gemm<128, true, 0>(tm, a, st)
             : mode == 1 ? launch_gemm<128, true, 1>(tm, a, st) : launch_gemm<128, true, 2>(tm, a, st);
    }
    if (rc != 0 || !split_tail) return rc;
    // more tiles than SMs: the BatchNorm backward runs as a launch of its own over the stored dpre
    const idcb_bn_tail* t = f->tail;
    return idcb_bn_bwd_apply(0, dx, 0, d->x_pitch, d->x_coff, f->aux, f->aux_pitch, f->aux_coff, t->out, t->out_pitch,
                             t->out_coff, f->stat_a, f->stat_b, t->mean, t->rstd, f->bwd_scale, t->count, t->dgamma,
                             t->dbeta, a.M, d->cin, stream);
}

extern "C" int idcb_conv_wgrad(const idcb_conv_desc* d, const void* x, const void* dy, float* dw,
                               const float* pro_scale, const float* pro_shift, int pro_act,
                               void* stream) {
    if (check_geom(d)) return 1;
    IDCB_REQUIRE((pro_scale == nullptr) == (pro_shift == nullptr),
                 "conv_wgrad: prologue scale and shift must be given together");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    WgradArgs a{};
    a.x = static_cast<const __nv_bfloat16*>(x);
    a.x_pitch = d->x_pitch; a.x_coff = d->x_coff;
    a.B = d->batch; a.H = d->h; a.W = d->w; a.Cin = d->cin;
    a.KH = d->kh; a.KW = d->kw; a.stride = d->stride; a.pad_t = d->pad_t; a.pad_l = d->pad_l;
    a.Ho = d->ho; a.Wo = d->wo;
    a.M = d->batch * d->ho * d->wo;
    a.Cout = d->cout;
    a.pro_scale = pro_scale; a.pro_shift = pro_shift; a.pro_act = pro_act;
    a.dw = dw + d->w_off;
    a.dw_pitch = d->w_pitch ? d->w_pitch : d->kh * d->kw * d->cin;
    const int kbn = d->cout <= 32 ? 32 : (d->cout <= 64 ? 64 : 128);
    const int total_kb = ceil_div(a.M, 64);
    const int base_ctas = ceil_div(a.KH * a.KW * a.Cin, 128) * ceil_div(a.Cout, kbn);
    int splits = ceil_div(num_sms(), base_ctas);
    if (splits > total_kb) splits = total_kb;
    if (splits < 1) splits = 1;
    a.kb_per_split = ceil_div(total_kb, splits);
    a.splits = ceil_div(total_kb, a.kb_per_split);
    CUtensorMap tm;
    // dy viewed as [M rows][y_pitch cols] starting at channel offset y_coff; boxes 64 x 64
    if (make_tmap_2d_bf16(&tm, static_cast<const __nv_bfloat16*>(dy) + d->y_coff, a.M,
                          (uint64_t)(d->y_pitch - d->y_coff), d->y_pitch, 64, 64))
        return 1;
    if (kbn == 32) return launch_wgrad<32>(tm, a, st);
    if (kbn == 64) return launch_wgrad<64>(tm, a, st);
    return launch_wgrad<128>(tm, a, st);
}
