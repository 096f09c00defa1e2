// Depthwise 3x3 convolution (stride 1 / 2) for MobileNetV2, NHWC bf16, weights [KH][KW][C] bf16.
//   fprop : y = dw( act(scale*x+shift) )                + batch statistics of y
//   dgrad : dpre = dw^T(dy) * act'(scale*x+shift)       + BN-backward sums
//   wgrad : dw += sum_pos act(scale*x+shift) * dy
// Bandwidth-bound (~6 % of MobileNetV2 MACs), so these are direct kernels, one
// thread per 8 channels, 16-byte accesses laid out so that a warp touches whole 128-byte lines (DwLane),
// fused BN prologue/epilogue bookkeeping instead of separate BN kernels. Reference: DepthwiseConv2D layers inside tf.keras.applications.MobileNetV2
// (/root/reference/dist_model_tf_mobile.py:119), executed by TF's depthwise CUDA kernels.
#include <stdlib.h>

#include "common.cuh"
#include "../include/idcb200.h"

namespace idcb {

struct DwArgs {
    const __nv_bfloat16* x; int x_pitch, x_coff;
    const __nv_bfloat16* w;             // [k][k][C]
    __nv_bfloat16* y; int y_pitch, y_coff;
    const __nv_bfloat16* dy; int dy_pitch, dy_coff;
    __nv_bfloat16* dx; int dx_pitch, dx_coff;
    float* dw;                          // fp32 [k][k][C]
    int B, H, W, C, Ho, Wo, k, stride, pad_t, pad_l;
    const float* scale; const float* shift; int act;
    float* stat_a; float* stat_b;
};

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 q = *reinterpret_cast<const uint4*>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        float2 f = unpack_bf16x2(w[e]);
        v[2 * e] = f.x;
        v[2 * e + 1] = f.y;
    }
}
__device__ __forceinline__ void ldw8(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 q;
    uint32_t* w = reinterpret_cast<uint32_t*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) w[e] = pack_bf16x2(v[2 * e], v[2 * e + 1]);
    *reinterpret_cast<uint4*>(p) = q;
}

// Thread layout shared by the three kernels: consecutive threads own consecutive channel octets of one
// position (so a warp reads whole 128-byte lines of the NHWC tensor), then the next position:
//   oct = tid % nOct, pl = tid / nOct, ppb = 256 / nOct positions per block pass; threads with pl >= ppb idle.
// A thread's octet never changes, so its 3x3 weights and BN coefficients live in registers.
struct DwLane {
    int nOct, ppb, oct, pl, c;
    bool active;
    __device__ __forceinline__ DwLane(int C) {
        nOct = C >> 3;
        ppb = blockDim.x / nOct;
        oct = threadIdx.x % nOct;
        pl = threadIdx.x / nOct;
        c = oct * 8;
        active = pl < ppb;
    }
};

// block-level reduction of 8 per-thread values per channel octet into global atomics: dst[oct * 8 + e]
__device__ __forceinline__ void block_reduce8_atomic(const float (&v)[8], float* dst, float (*sh)[9], const DwLane& L) {
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 8; ++e) sh[threadIdx.x][e] = L.active ? v[e] : 0.f;
    __syncthreads();
    for (int i = threadIdx.x; i < L.nOct * 8; i += blockDim.x) {
        const int o = i >> 3, e = i & 7;
        float t = 0.f;
        for (int r = 0; r < L.ppb; ++r) t += sh[r * L.nOct + o][e];
        atomicAdd(dst + i, t);
    }
}

template <bool kSmemW>
__global__ void __launch_bounds__(256) dw_fprop_kernel(DwArgs p) {
    pdl_enter();
    __shared__ float sh[256][9];
    const DwLane L(p.C);
    const int c = L.c;
    float sc[8], sf[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sc[e] = p.scale ? p.scale[c + e] : 1.f;
        sf[e] = p.scale ? p.shift[c + e] : 0.f;
    }
    // kSmemW (layers with many positions per thread): the 3x3 weights of every channel, fp32, in shared memory —
    // 72 registers per thread less than holding this thread's 9 x 8 values, which is the difference between one
    // and three resident blocks per SM. Small layers keep them in registers (no fill, no barrier).
    extern __shared__ float wsm[];
    float wk[kSmemW ? 1 : 9][8];
    if (kSmemW) {
        for (int i = threadIdx.x; i < 9 * p.C; i += blockDim.x) wsm[i] = __bfloat162float(p.w[i]);
        __syncthreads();
    } else {
#pragma unroll
        for (int t = 0; t < (kSmemW ? 1 : 9); ++t) load8(p.w + static_cast<size_t>(t) * p.C + c, wk[t]);
    }
    float sa[8], sb[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) sa[e] = sb[e] = 0.f;
    const long long npos = 1LL * p.B * p.Ho * p.Wo;
    for (long long pos = blockIdx.x * (long long)L.ppb + L.pl; L.active && pos < npos;
         pos += (long long)gridDim.x * L.ppb) {
        const int wo = static_cast<int>(pos % p.Wo);
        const long long t2 = pos / p.Wo;
        const int ho = static_cast<int>(t2 % p.Ho);
        const int b = static_cast<int>(t2 / p.Ho);
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int hi = ho * p.stride - p.pad_t + kh;
            if (hi < 0 || hi >= p.H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int wi = wo * p.stride - p.pad_l + kw;
                if (wi < 0 || wi >= p.W) continue;
                float xv[8];
                load8(p.x + (static_cast<size_t>(b * p.H + hi) * p.W + wi) * p.x_pitch + p.x_coff + c, xv);
                float wv[8];
                if (kSmemW) {
                    ldw8(wsm + (kh * 3 + kw) * p.C + c, wv);
                } else {
#pragma unroll
                    for (int e = 0; e < 8; ++e) wv[e] = wk[kSmemW ? 0 : kh * 3 + kw][e];
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float a = xv[e];
                    if (p.scale) a = __bfloat162float(__float2bfloat16_rn(apply_act(fmaf(a, sc[e], sf[e]), p.act)));
                    acc[e] = fmaf(a, wv[e], acc[e]);
                }
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            acc[e] = __bfloat162float(__float2bfloat16_rn(acc[e]));
            sa[e] += acc[e];
            sb[e] += acc[e] * acc[e];
        }
        store8(p.y + pos * p.y_pitch + p.y_coff + c, acc);
    }
    if (p.stat_a) block_reduce8_atomic(sa, p.stat_a, sh, L);
    if (p.stat_b) block_reduce8_atomic(sb, p.stat_b, sh, L);
}

template <bool kSmemW>
__global__ void __launch_bounds__(256) dw_dgrad_kernel(DwArgs p) {
    pdl_enter();
    __shared__ float sh[256][9];
    const DwLane L(p.C);
    const int c = L.c;
    float sc[8], sf[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sc[e] = p.scale ? p.scale[c + e] : 1.f;
        sf[e] = p.scale ? p.shift[c + e] : 0.f;
    }
    // kSmemW (layers with many positions per thread): the 3x3 weights of every channel, fp32, in shared memory —
    // 72 registers per thread less than holding this thread's 9 x 8 values, which is the difference between one
    // and three resident blocks per SM. Small layers keep them in registers (no fill, no barrier).
    extern __shared__ float wsm[];
    float wk[kSmemW ? 1 : 9][8];
    if (kSmemW) {
        for (int i = threadIdx.x; i < 9 * p.C; i += blockDim.x) wsm[i] = __bfloat162float(p.w[i]);
        __syncthreads();
    } else {
#pragma unroll
        for (int t = 0; t < (kSmemW ? 1 : 9); ++t) load8(p.w + static_cast<size_t>(t) * p.C + c, wk[t]);
    }
    float sa[8], sb[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) sa[e] = sb[e] = 0.f;
    const long long npos = 1LL * p.B * p.H * p.W;
    for (long long pos = blockIdx.x * (long long)L.ppb + L.pl; L.active && pos < npos;
         pos += (long long)gridDim.x * L.ppb) {
        const int wi = static_cast<int>(pos % p.W);
        const long long t2 = pos / p.W;
        const int hi = static_cast<int>(t2 % p.H);
        const int b = static_cast<int>(t2 / p.H);
        float g[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = 0.f;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int hn = hi + p.pad_t - kh;
            if (hn < 0 || hn % p.stride) continue;
            const int ho = hn / p.stride;
            if (ho >= p.Ho) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int wn = wi + p.pad_l - kw;
                if (wn < 0 || wn % p.stride) continue;
                const int wo = wn / p.stride;
                if (wo >= p.Wo) continue;
                float dv[8];
                load8(p.dy + (static_cast<size_t>(b * p.Ho + ho) * p.Wo + wo) * p.dy_pitch + p.dy_coff + c, dv);
                float wv[8];
                if (kSmemW) {
                    ldw8(wsm + (kh * 3 + kw) * p.C + c, wv);
                } else {
#pragma unroll
                    for (int e = 0; e < 8; ++e) wv[e] = wk[kSmemW ? 0 : kh * 3 + kw][e];
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) g[e] = fmaf(dv[e], wv[e], g[e]);
            }
        }
        float xv[8];
        if (p.scale || p.stat_b) load8(p.x + pos * p.x_pitch + p.x_coff + c, xv);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (p.scale) g[e] *= act_mask(fmaf(xv[e], sc[e], sf[e]), p.act);
            g[e] = __bfloat162float(__float2bfloat16_rn(g[e]));
            sa[e] += g[e];
            if (p.stat_b) sb[e] += g[e] * xv[e];
        }
        store8(p.dx + pos * p.dx_pitch + p.dx_coff + c, g);
    }
    if (p.stat_a) block_reduce8_atomic(sa, p.stat_a, sh, L);
    if (p.stat_b) block_reduce8_atomic(sb, p.stat_b, sh, L);
}

__global__ void __launch_bounds__(256) dw_wgrad_kernel(DwArgs p) {
    pdl_enter();
    __shared__ float sh[256][9];
    const DwLane L(p.C);
    const int c = L.c;
    float sc[8], sf[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sc[e] = p.scale ? p.scale[c + e] : 1.f;
        sf[e] = p.scale ? p.shift[c + e] : 0.f;
    }
    float acc[9][8];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[t][e] = 0.f;
    const long long npos = 1LL * p.B * p.Ho * p.Wo;
    for (long long pos = blockIdx.x * (long long)L.ppb + L.pl; L.active && pos < npos;
         pos += (long long)gridDim.x * L.ppb) {
        const int wo = static_cast<int>(pos % p.Wo);
        const long long t2 = pos / p.Wo;
        const int ho = static_cast<int>(t2 % p.Ho);
        const int b = static_cast<int>(t2 / p.Ho);
        float dv[8];
        load8(p.dy + pos * p.dy_pitch + p.dy_coff + c, dv);
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int hi = ho * p.stride - p.pad_t + kh;
            if (hi < 0 || hi >= p.H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int wi = wo * p.stride - p.pad_l + kw;
                if (wi < 0 || wi >= p.W) continue;
                float xv[8];
                load8(p.x + (static_cast<size_t>(b * p.H + hi) * p.W + wi) * p.x_pitch + p.x_coff + c, xv);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float a = xv[e];
                    if (p.scale) a = __bfloat162float(__float2bfloat16_rn(apply_act(fmaf(a, sc[e], sf[e]), p.act)));
                    acc[kh * 3 + kw][e] = fmaf(a, dv[e], acc[kh * 3 + kw][e]);
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) block_reduce8_atomic(acc[t], p.dw + static_cast<size_t>(t) * p.C, sh, L);
}

static int fill(DwArgs& p, const idcb_conv_desc* d) {
    IDCB_REQUIRE(d != nullptr, "depthwise: null descriptor");
    IDCB_REQUIRE(d->cin == d->cout && d->cin % 8 == 0 && d->cin <= 2048, "depthwise: C must be a multiple of 8, at most 2048");
    IDCB_REQUIRE(d->kh == 3 && d->kw == 3, "depthwise: only 3x3 is implemented");
    IDCB_REQUIRE(d->x_pitch % 8 == 0 && d->x_coff % 8 == 0 && d->y_pitch % 8 == 0 && d->y_coff % 8 == 0,
                 "depthwise: alignment");
    p.B = d->batch; p.H = d->h; p.W = d->w; p.C = d->cin; p.Ho = d->ho; p.Wo = d->wo;
    p.k = 3; p.stride = d->stride; p.pad_t = d->pad_t; p.pad_l = d->pad_l;
    return 0;
}

// the shared-memory-weights variant pays a fill + barrier per block: worth it once every thread has several positions
static bool dw_use_smem_weights(long long npos, int C) {
    static const long long thr = [] { const char* e = getenv("IDCB_DW_SMEM_MIN_ITEMS"); return e ? atoll(e) : 1LL; }();
    return npos * (C / 8) >= thr * 3 * 256 * num_sms();
}

static dim3 dw_grid(long long npos, int C, int pos_per_thread = 1, int blocks_per_sm = 0) {
    const int ppb = 256 / (C / 8);                          // positions per block pass (see DwLane)
    long long gx = ceil_div_ll(npos, 1LL * ppb * pos_per_thread);
    // every block ends with per-channel atomics (statistics / weight gradient): the number of blocks, not the
    // data volume, sets the contention on those few hundred addresses
    static const int cap_fd = [] { const char* e = getenv("IDCB_DW_BLOCKS_PER_SM"); int v = e ? atoi(e) : 2; return v > 0 ? v : 2; }();
    static const int cap_w = [] { const char* e = getenv("IDCB_DW_WGRAD_BLOCKS_PER_SM"); int v = e ? atoi(e) : 2; return v > 0 ? v : 2; }();
    const long long cap = 1LL * (blocks_per_sm > 0 ? blocks_per_sm : (pos_per_thread > 1 ? cap_w : cap_fd)) * num_sms();
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    return dim3(static_cast<unsigned>(gx), 1);
}

}  // namespace idcb

using namespace idcb;

extern "C" int idcb_dwconv_fprop(const idcb_conv_desc* d, const void* x, const void* w, void* y,
                                 const float* scale, const float* shift, int act, float* stat_sum,
                                 float* stat_sq, void* stream) {
    DwArgs p{};
    if (fill(p, d)) return 1;
    p.x = (const __nv_bfloat16*)x; p.x_pitch = d->x_pitch; p.x_coff = d->x_coff;
    p.w = (const __nv_bfloat16*)w;
    p.y = (__nv_bfloat16*)y; p.y_pitch = d->y_pitch; p.y_coff = d->y_coff;
    p.scale = scale; p.shift = shift; p.act = act; p.stat_a = stat_sum; p.stat_b = stat_sq;
    const long long npos = 1LL * p.B * p.Ho * p.Wo;
    if (npos == 0) return 0;
    if (dw_use_smem_weights(npos, p.C))
        IDCB_CHECK(launch_pdl(dw_fprop_kernel<true>, dw_grid(npos, p.C, 1, 3), 256, 9 * p.C * sizeof(float), (cudaStream_t)stream, p));
    else
        IDCB_CHECK(launch_pdl(dw_fprop_kernel<false>, dw_grid(npos, p.C), 256, 0, (cudaStream_t)stream, p));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_dwconv_dgrad(const idcb_conv_desc* d, const void* dy, const void* w, const void* x,
                                 void* dx, const float* scale, const float* shift, int act,
                                 float* stat_a, float* stat_b, void* stream) {
    DwArgs p{};
    if (fill(p, d)) return 1;
    p.x = (const __nv_bfloat16*)x; p.x_pitch = d->x_pitch; p.x_coff = d->x_coff;
    p.w = (const __nv_bfloat16*)w;
    p.dy = (const __nv_bfloat16*)dy; p.dy_pitch = d->y_pitch; p.dy_coff = d->y_coff;
    p.dx = (__nv_bfloat16*)dx; p.dx_pitch = d->x_pitch; p.dx_coff = d->x_coff;
    p.scale = scale; p.shift = shift; p.act = act; p.stat_a = stat_a; p.stat_b = stat_b;
    IDCB_REQUIRE(!(scale || stat_b) || x != nullptr, "dwconv_dgrad: x needed for the fused backward");
    const long long npos = 1LL * p.B * p.H * p.W;
    if (npos == 0) return 0;
    if (dw_use_smem_weights(npos, p.C))
        IDCB_CHECK(launch_pdl(dw_dgrad_kernel<true>, dw_grid(npos, p.C, 1, 3), 256, 9 * p.C * sizeof(float), (cudaStream_t)stream, p));
    else
        IDCB_CHECK(launch_pdl(dw_dgrad_kernel<false>, dw_grid(npos, p.C), 256, 0, (cudaStream_t)stream, p));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_dwconv_wgrad(const idcb_conv_desc* d, const void* x, const void* dy, float* dw,
                                 const float* scale, const float* shift, int act, void* stream) {
    DwArgs p{};
    if (fill(p, d)) return 1;
    p.x = (const __nv_bfloat16*)x; p.x_pitch = d->x_pitch; p.x_coff = d->x_coff;
    p.dy = (const __nv_bfloat16*)dy; p.dy_pitch = d->y_pitch; p.dy_coff = d->y_coff;
    p.dw = dw;
    p.scale = scale; p.shift = shift; p.act = act;
    const long long npos = 1LL * p.B * p.Ho * p.Wo;
    if (npos == 0) return 0;
    IDCB_CHECK(launch_pdl(dw_wgrad_kernel, dw_grid(npos, p.C, 8), 256, 0, (cudaStream_t)stream, p));
    IDCB_LAUNCH_CHECK();
    return 0;
}
