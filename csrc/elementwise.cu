// Bandwidth-bound helper kernels: BatchNorm bookkeeping (finalize / fold / backward coefficients),
// per-channel affine transforms, channel statistics, pooling forward/backward, casts.
//
// These replace the cuDNN BatchNorm / pooling calls and Eigen element-wise kernels TensorFlow
// launches for the reference's Keras layers (e.g.
// /root/reference/dist_model_tf_dense.py:131 DenseNet BN/pool layers, dist_model_tf_vgg.py:119
// VGG16 MaxPooling2D). Activations are NHWC bf16 addressed as (base, pitch, coff).
#include "common.cuh"
#include "../include/idcb200.h"

namespace idcb {

// ---------------------------------------------------------------------------------------------
// BatchNorm per-channel kernels (C is small: <= 2048)
// ---------------------------------------------------------------------------------------------
__global__ void bn_finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sq,
                                   float count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float eps, float* scale,
                                   float* shift, float* mean_out, float* rstd_out,
                                   float* moving_mean, float* moving_var, float momentum, int C) {
    pdl_enter();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float m = sum[c] / count;
    float v = sq[c] / count - m * m;
    v = fmaxf(v, 0.f);
    const float r = rsqrtf(v + eps);
    const float s = gamma[c] * r;
    scale[c] = s;
    shift[c] = beta[c] - m * s;
    if (mean_out) mean_out[c] = m;
    if (rstd_out) rstd_out[c] = r;
    if (moving_mean) {
        const float unbiased = count > 1.f ? v * (count / (count - 1.f)) : v;
        moving_mean[c] = moving_mean[c] * momentum + m * (1.f - momentum);
        moving_var[c] = moving_var[c] * momentum + unbiased * (1.f - momentum);
    }
}

__global__ void bn_fold_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                               const float* __restrict__ mean, const float* __restrict__ var,
                               float eps, float* scale, float* shift, int C) {
    pdl_enter();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float s = gamma[c] * rsqrtf(var[c] + eps);
    scale[c] = s;
    shift[c] = beta[c] - mean[c] * s;
}

// BN backward coefficients from the epilogue sums of the consumer's dgrad (mode 2):
//   sa = sum dpre, sb = sum dpre * x_raw
//   dbeta = sa ; dgamma = rstd * (sb - mean * sa)
// standalone BN (acc_s1 == null):  dx = A*g + B*x + C0 with
//   A = scale, B = -scale*rstd*dgamma/N, C0 = -scale*(dbeta/N) + scale*rstd*mean*dgamma/N
// shared-buffer BN (DenseNet): S1 += scale*dbeta ; S2 += scale*dgamma (coefficients come later from
// bn_shared_coeffs once every consumer has contributed).
__global__ void bn_bwd_coeffs_kernel(const float* __restrict__ sa, const float* __restrict__ sb,
                                     const float* __restrict__ mean, const float* __restrict__ rstd,
                                     const float* __restrict__ scale, float count, float* dgamma,
                                     float* dbeta, float* A, float* B, float* C0, float* acc_s1,
                                     float* acc_s2, int C) {
    pdl_enter();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float a = sa[c], b = sb[c], m = mean[c], r = rstd[c], s = scale[c];
    const float dg = r * (b - m * a);
    if (dgamma) dgamma[c] += dg;
    if (dbeta) dbeta[c] += a;
    if (acc_s1) {
        acc_s1[c] += s * a;
        acc_s2[c] += s * dg;
    }
    if (A) {
        A[c] = s;
        B[c] = -s * r * dg / count;
        C0[c] = -s * a / count + s * r * m * dg / count;
    }
}

// DenseNet shared-buffer channels: dx = G - S1/N - xhat*S2/N = 1*G + B*x + C0
__global__ void bn_shared_coeffs_kernel(const float* __restrict__ s1, const float* __restrict__ s2,
                                        const float* __restrict__ mean,
                                        const float* __restrict__ rstd, float count, float* A,
                                        float* B, float* C0, int C) {
    pdl_enter();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float r = rstd[c], m = mean[c];
    A[c] = 1.f;
    B[c] = -r * s2[c] / count;
    C0[c] = -s1[c] / count + r * m * s2[c] / count;
}

// ---------------------------------------------------------------------------------------------
// y = act(scale*x + shift) (+ residual)     [bn apply / fold]
// ---------------------------------------------------------------------------------------------
__global__ void affine_act_kernel(const __nv_bfloat16* __restrict__ x, int x_pitch, int x_coff,
                                  __nv_bfloat16* y, int y_pitch, int y_coff,
                                  const float* __restrict__ scale, const float* __restrict__ shift,
                                  int act, const __nv_bfloat16* __restrict__ res, int r_pitch,
                                  int r_coff, long long M, int C) {
    pdl_enter();
    const int c8 = C >> 3;
    const long long total = M * c8;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / c8;
        const int c = static_cast<int>(i - m * c8) * 8;
        uint4 q = *reinterpret_cast<const uint4*>(x + m * x_pitch + x_coff + c);
        uint4 rr = make_uint4(0, 0, 0, 0);
        if (res) rr = *reinterpret_cast<const uint4*>(res + m * r_pitch + r_coff + c);
        uint32_t* w = reinterpret_cast<uint32_t*>(&q);
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(&rr);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float2 f = unpack_bf16x2(w[e]);
            const float s0 = scale ? scale[c + 2 * e] : 1.f, s1 = scale ? scale[c + 2 * e + 1] : 1.f;
            const float t0 = shift ? shift[c + 2 * e] : 0.f, t1 = shift ? shift[c + 2 * e + 1] : 0.f;
            f.x = apply_act(fmaf(f.x, s0, t0), act);
            f.y = apply_act(fmaf(f.y, s1, t1), act);
            if (res) {
                float2 r2 = unpack_bf16x2(rw[e]);
                f.x += r2.x;
                f.y += r2.y;
            }
            w[e] = pack_bf16x2(f.x, f.y);
        }
        *reinterpret_cast<uint4*>(y + m * y_pitch + y_coff + c) = q;
    }
}

// out(bf16) = A[c]*g + B[c]*x + C0[c];  g is fp32 (G buffer) or bf16 (dpre)
template <bool kGFloat>
__global__ void affine2_kernel(const void* __restrict__ g, int g_pitch, int g_coff,
                               const __nv_bfloat16* __restrict__ x, int x_pitch, int x_coff,
                               __nv_bfloat16* out, int o_pitch, int o_coff,
                               const float* __restrict__ A, const float* __restrict__ B,
                               const float* __restrict__ C0, long long M, int C) {
    pdl_enter();
    const int c8 = C >> 3;
    const long long total = M * c8;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / c8;
        const int c = static_cast<int>(i - m * c8) * 8;
        float gv[8];
        if (kGFloat) {
            const float* gp = static_cast<const float*>(g) + m * g_pitch + g_coff + c;
            float4 a = *reinterpret_cast<const float4*>(gp);
            float4 b = *reinterpret_cast<const float4*>(gp + 4);
            gv[0] = a.x; gv[1] = a.y; gv[2] = a.z; gv[3] = a.w;
            gv[4] = b.x; gv[5] = b.y; gv[6] = b.z; gv[7] = b.w;
        } else {
            uint4 q = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(g) +
                                                      m * g_pitch + g_coff + c);
            const uint32_t* w = reinterpret_cast<const uint32_t*>(&q);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float2 f = unpack_bf16x2(w[e]);
                gv[2 * e] = f.x;
                gv[2 * e + 1] = f.y;
            }
        }
        uint4 xq = *reinterpret_cast<const uint4*>(x + m * x_pitch + x_coff + c);
        const uint32_t* xw = reinterpret_cast<const uint32_t*>(&xq);
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float2 xf = unpack_bf16x2(xw[e]);
            const int c0 = c + 2 * e;
            const float r0 = fmaf(A[c0], gv[2 * e], fmaf(B[c0], xf.x, C0[c0]));
            const float r1 = fmaf(A[c0 + 1], gv[2 * e + 1], fmaf(B[c0 + 1], xf.y, C0[c0 + 1]));
            ow[e] = pack_bf16x2(r0, r1);
        }
        *reinterpret_cast<uint4*>(out + m * o_pitch + o_coff + c) = o;
    }
}

// ---------------------------------------------------------------------------------------------
// Per-channel sum / sum of squares of a bf16 NHWC slice (for BN stats of tensors that are not
// produced by a GEMM epilogue). grid.x = channel groups of 64, grid.y = row splits.
// ---------------------------------------------------------------------------------------------
__global__ void channel_stats_kernel(const __nv_bfloat16* __restrict__ x, int pitch, int coff,
                                     long long M, int C, float* sum, float* sq) {
    pdl_enter();
    // block = 256 threads = 8 channel-octets x 32 row lanes
    const int oct = threadIdx.x & 7;
    const int rl = threadIdx.x >> 3;
    const int c = blockIdx.x * 64 + oct * 8;
    float s[8], q[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) s[e] = q[e] = 0.f;
    if (c < C) {
        for (long long m = blockIdx.y * 32 + rl; m < M; m += (long long)gridDim.y * 32) {
            uint4 v = *reinterpret_cast<const uint4*>(x + m * pitch + coff + c);
            const uint32_t* w = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float2 f = unpack_bf16x2(w[e]);
                s[2 * e] += f.x; q[2 * e] += f.x * f.x;
                s[2 * e + 1] += f.y; q[2 * e + 1] += f.y * f.y;
            }
        }
    }
    __shared__ float sh[2][32][65];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sh[0][rl][oct * 8 + e] = s[e];
        sh[1][rl][oct * 8 + e] = q[e];
    }
    __syncthreads();
    if (threadIdx.x < 128) {
        const int which = threadIdx.x >> 6;
        const int ch = threadIdx.x & 63;
        float t = 0.f;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) t += sh[which][r][ch];
        if (blockIdx.x * 64 + ch < C) atomicAdd((which ? sq : sum) + blockIdx.x * 64 + ch, t);
    }
}

// sum_g[c] += sum_m g ; sum_gx[c] += sum_m g*x   (BN backward sums when dpre is already in memory)
__global__ void channel_stats2_kernel(const __nv_bfloat16* __restrict__ g, int g_pitch, int g_coff,
                                      const __nv_bfloat16* __restrict__ x, int x_pitch, int x_coff,
                                      long long M, int C, float* sum_g, float* sum_gx) {
    pdl_enter();
    const int oct = threadIdx.x & 7;
    const int rl = threadIdx.x >> 3;
    const int c = blockIdx.x * 64 + oct * 8;
    float s[8], q[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) s[e] = q[e] = 0.f;
    if (c < C) {
        for (long long m = blockIdx.y * 32 + rl; m < M; m += (long long)gridDim.y * 32) {
            uint4 gv = *reinterpret_cast<const uint4*>(g + m * g_pitch + g_coff + c);
            uint4 xv = *reinterpret_cast<const uint4*>(x + m * x_pitch + x_coff + c);
            const uint32_t* gw = reinterpret_cast<const uint32_t*>(&gv);
            const uint32_t* xw = reinterpret_cast<const uint32_t*>(&xv);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float2 a = unpack_bf16x2(gw[e]);
                float2 b = unpack_bf16x2(xw[e]);
                s[2 * e] += a.x; q[2 * e] += a.x * b.x;
                s[2 * e + 1] += a.y; q[2 * e + 1] += a.y * b.y;
            }
        }
    }
    __shared__ float sh[2][32][65];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sh[0][rl][oct * 8 + e] = s[e];
        sh[1][rl][oct * 8 + e] = q[e];
    }
    __syncthreads();
    if (threadIdx.x < 128) {
        const int which = threadIdx.x >> 6;
        const int ch = threadIdx.x & 63;
        float t = 0.f;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) t += sh[which][r][ch];
        if (blockIdx.x * 64 + ch < C) atomicAdd((which ? sum_gx : sum_g) + blockIdx.x * 64 + ch, t);
    }
}

// ---------------------------------------------------------------------------------------------
// Pooling. One thread handles 8 channels of one output position.
//   max: y = max over window of act(scale*x+shift) (prologue optional); window positions outside the
//        input (pad) are skipped (equivalent to Keras ZeroPadding2D + VALID pooling on ReLU outputs).
//   avg: y = mean over the k x k window (no padding).
// ---------------------------------------------------------------------------------------------
struct PoolArgs {
    const __nv_bfloat16* x; int x_pitch, x_coff;
    __nv_bfloat16* y; int y_pitch, y_coff;
    int B, H, W, C, Ho, Wo, k, stride, pad;
    const float* scale; const float* shift; int act;
    int is_max;
    float* stat_sum; float* stat_sq;
    unsigned char* argmax;   // optional (max): window index kh*k+kw of the first maximum, one byte per output element
};

__global__ void pool_fwd_kernel(PoolArgs p) {
    pdl_enter();
    const int c8 = p.C >> 3;
    const long long total = 1LL * p.B * p.Ho * p.Wo * c8;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = static_cast<int>(i % c8) * 8;
        long long pos = i / c8;
        const int wo = static_cast<int>(pos % p.Wo);
        pos /= p.Wo;
        const int ho = static_cast<int>(pos % p.Ho);
        const int b = static_cast<int>(pos / p.Ho);
        float acc[8];
        unsigned char am[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) { acc[e] = p.is_max ? -3.0e38f : 0.f; am[e] = 0; }
        bool any = false;
        for (int kh = 0; kh < p.k; ++kh) {
            const int hi = ho * p.stride - p.pad + kh;
            if (hi < 0 || hi >= p.H) continue;
            for (int kw = 0; kw < p.k; ++kw) {
                const int wi = wo * p.stride - p.pad + kw;
                if (wi < 0 || wi >= p.W) continue;
                any = true;
                uint4 v = *reinterpret_cast<const uint4*>(
                    p.x + (static_cast<long long>(b * p.H + hi) * p.W + wi) * p.x_pitch + p.x_coff + c);
                const uint32_t* w = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float2 f = unpack_bf16x2(w[e]);
                    if (p.scale) {
                        f.x = apply_act(fmaf(f.x, p.scale[c + 2 * e], p.shift[c + 2 * e]), p.act);
                        f.y = apply_act(fmaf(f.y, p.scale[c + 2 * e + 1], p.shift[c + 2 * e + 1]), p.act);
                        // the pooled tensor is stored in bf16; compare in that precision
                        f.x = __bfloat162float(__float2bfloat16_rn(f.x));
                        f.y = __bfloat162float(__float2bfloat16_rn(f.y));
                    }
                    if (p.is_max) {
                        // strict '>' in row-major window order: ties go to the first maximum (cuDNN / TF / PyTorch)
                        if (f.x > acc[2 * e]) { acc[2 * e] = f.x; am[2 * e] = static_cast<unsigned char>(kh * p.k + kw); }
                        if (f.y > acc[2 * e + 1]) { acc[2 * e + 1] = f.y; am[2 * e + 1] = static_cast<unsigned char>(kh * p.k + kw); }
                    } else {
                        acc[2 * e] += f.x;
                        acc[2 * e + 1] += f.y;
                    }
                }
            }
        }
        const float inv = 1.f / (p.k * p.k);
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float a0 = acc[2 * e], a1 = acc[2 * e + 1];
            if (!p.is_max) { a0 *= inv; a1 *= inv; }
            if (!any) { a0 = 0.f; a1 = 0.f; }
            ow[e] = pack_bf16x2(a0, a1);
        }
        const long long opos = (static_cast<long long>(b * p.Ho + ho) * p.Wo + wo);
        *reinterpret_cast<uint4*>(p.y + opos * p.y_pitch + p.y_coff + c) = o;
        if (p.argmax != nullptr) {
            uint2 pk;
            pk.x = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
            pk.y = am[4] | (am[5] << 8) | (am[6] << 16) | (am[7] << 24);
            *reinterpret_cast<uint2*>(p.argmax + opos * p.C + c) = pk;
        }
    }
}

// Backward, gather form: one thread = 8 channels of one INPUT position.
//   max : dx = sum over windows containing the position of dy * [a == y_window]  (a = act(scale*x+shift));
//         then the activation/affine backward: dx *= act'(pre) (dpre w.r.t. scale*x+shift; the caller
//         handles the BN part) ; relu_mask_src (optional): dx *= (x > 0) when x itself is a ReLU output.
//   avg : dx = sum over windows of dy / k^2.
// stat_a[c] += sum dx ; stat_b[c] += sum dx * x_raw   (bias gradient / BN backward sums)
struct PoolBwdArgs {
    const __nv_bfloat16* x; int x_pitch, x_coff;     // forward input (raw)
    const __nv_bfloat16* y; int y_pitch, y_coff;     // forward output (pooled), max only
    const __nv_bfloat16* dy; int dy_pitch, dy_coff;
    __nv_bfloat16* dx; int dx_pitch, dx_coff;
    int B, H, W, C, Ho, Wo, k, stride, pad;
    const float* scale; const float* shift; int act;
    int is_max;
    int relu_mask;   // dx *= (x > 0)
    float* stat_a; float* stat_b;
    const unsigned char* argmax;   // optional (max): what the forward pass recorded; replaces the value comparison
};

__global__ void pool_bwd_kernel(PoolBwdArgs p) {
    pdl_enter();
    const int c8 = p.C >> 3;
    const long long total = 1LL * p.B * p.H * p.W * c8;
    __shared__ float s_a[256][9];   // padded to dodge bank conflicts on the final reduce
    float sa[8], sb[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) sa[e] = sb[e] = 0.f;
    // each block walks a fixed channel octet so the statistics can be reduced per block
    const int oct = blockIdx.y;
    const int c = oct * 8;
    const long long npos = 1LL * p.B * p.H * p.W;
    (void)total;
    for (long long pos = blockIdx.x * (long long)blockDim.x + threadIdx.x; pos < npos;
         pos += (long long)gridDim.x * blockDim.x) {
        const int wi = static_cast<int>(pos % p.W);
        const long long t = pos / p.W;
        const int hi = static_cast<int>(t % p.H);
        const int b = static_cast<int>(t / p.H);
        uint4 xv = *reinterpret_cast<const uint4*>(p.x + pos * p.x_pitch + p.x_coff + c);
        const uint32_t* xw = reinterpret_cast<const uint32_t*>(&xv);
        float xr[8], a[8], pre[8];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float2 f = unpack_bf16x2(xw[e]);
            xr[2 * e] = f.x; xr[2 * e + 1] = f.y;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            pre[e] = p.scale ? fmaf(xr[e], p.scale[c + e], p.shift[c + e]) : xr[e];
            a[e] = p.scale ? __bfloat162float(__float2bfloat16_rn(apply_act(pre[e], p.act))) : xr[e];
        }
        float g[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = 0.f;
        // windows (ho, wo) with ho*stride - pad <= hi < ho*stride - pad + k
        const int ho_lo = max(0, (hi + p.pad - p.k + p.stride) / p.stride);
        const int ho_hi = min(p.Ho - 1, (hi + p.pad) / p.stride);
        const int wo_lo = max(0, (wi + p.pad - p.k + p.stride) / p.stride);
        const int wo_hi = min(p.Wo - 1, (wi + p.pad) / p.stride);
        for (int ho = ho_lo; ho <= ho_hi; ++ho)
            for (int wo = wo_lo; wo <= wo_hi; ++wo) {
                const long long op = (static_cast<long long>(b * p.Ho + ho) * p.Wo + wo);
                uint4 dv = *reinterpret_cast<const uint4*>(p.dy + op * p.dy_pitch + p.dy_coff + c);
                const uint32_t* dw = reinterpret_cast<const uint32_t*>(&dv);
                if (p.is_max && p.argmax != nullptr) {
                    // the forward pass recorded which window element won: no re-evaluation of the activation, no
                    // tie-breaking scan over the earlier window elements
                    const uint2 pk = *reinterpret_cast<const uint2*>(p.argmax + op * p.C + c);
                    const unsigned me = static_cast<unsigned>((hi - (ho * p.stride - p.pad)) * p.k + (wi - (wo * p.stride - p.pad)));
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 d = unpack_bf16x2(dw[e]);
                        const unsigned a0 = ((e < 2 ? pk.x : pk.y) >> (16 * (e & 1))) & 0xFFu;
                        const unsigned a1 = ((e < 2 ? pk.x : pk.y) >> (16 * (e & 1) + 8)) & 0xFFu;
                        if (a0 == me) g[2 * e] += d.x;
                        if (a1 == me) g[2 * e + 1] += d.y;
                    }
                } else if (p.is_max) {
                    uint4 yv = *reinterpret_cast<const uint4*>(p.y + op * p.y_pitch + p.y_coff + c);
                    const uint32_t* yw = reinterpret_cast<const uint32_t*>(&yv);
                    float ymax[8], dd[8];
                    bool hit[8];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float2 d = unpack_bf16x2(dw[e]);
                        float2 yy = unpack_bf16x2(yw[e]);
                        ymax[2 * e] = yy.x; ymax[2 * e + 1] = yy.y;
                        dd[2 * e] = d.x; dd[2 * e + 1] = d.y;
                    }
                    bool any_hit = false;
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        hit[e] = a[e] == ymax[e];
                        any_hit |= hit[e];
                    }
                    if (any_hit) {
                        // cuDNN / TF / PyTorch route the gradient to the FIRST maximum of the window
                        // (row-major scan): drop the hit if an earlier window element ties.
                        const int h0 = ho * p.stride - p.pad, w0 = wo * p.stride - p.pad;
                        const int qme = (hi - h0) * p.k + (wi - w0);
                        for (int q = 0; q < qme; ++q) {
                            const int h2 = h0 + q / p.k, w2 = w0 + q % p.k;
                            if (h2 < 0 || h2 >= p.H || w2 < 0 || w2 >= p.W) continue;
                            uint4 ov = *reinterpret_cast<const uint4*>(
                                p.x + (static_cast<long long>(b * p.H + h2) * p.W + w2) * p.x_pitch +
                                p.x_coff + c);
                            const uint32_t* ow2 = reinterpret_cast<const uint32_t*>(&ov);
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                float2 f = unpack_bf16x2(ow2[e]);
                                if (p.scale) {
                                    f.x = __bfloat162float(__float2bfloat16_rn(apply_act(
                                        fmaf(f.x, p.scale[c + 2 * e], p.shift[c + 2 * e]), p.act)));
                                    f.y = __bfloat162float(__float2bfloat16_rn(apply_act(
                                        fmaf(f.y, p.scale[c + 2 * e + 1], p.shift[c + 2 * e + 1]), p.act)));
                                }
                                if (f.x == ymax[2 * e]) hit[2 * e] = false;
                                if (f.y == ymax[2 * e + 1]) hit[2 * e + 1] = false;
                            }
                        }
#pragma unroll
                        for (int e = 0; e < 8; ++e)
                            if (hit[e]) g[e] += dd[e];
                    }
                } else {
                    const float inv = 1.f / (p.k * p.k);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float2 d = unpack_bf16x2(dw[e]);
                        g[2 * e] += d.x * inv;
                        g[2 * e + 1] += d.y * inv;
                    }
                }
            }
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (p.scale) g[e] *= act_mask(pre[e], p.act);
            if (p.relu_mask && !(xr[e] > 0.f)) g[e] = 0.f;
            g[e] = __bfloat162float(__float2bfloat16_rn(g[e]));
            sa[e] += g[e];
            sb[e] += g[e] * xr[e];
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) ow[e] = pack_bf16x2(g[2 * e], g[2 * e + 1]);
        *reinterpret_cast<uint4*>(p.dx + pos * p.dx_pitch + p.dx_coff + c) = o;
    }
    if (p.stat_a == nullptr) return;
    // block-wide sums of the 8 channels: warp shuffles, then one partial per warp through shared memory
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sa[e] = warp_sum(sa[e]);
        sb[e] = warp_sum(sb[e]);
    }
    if (lane == 0) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            s_a[wid][e] = sa[e];
            s_a[8 + wid][e] = sb[e];
        }
    }
    __syncthreads();
    if (threadIdx.x < 16) {
        const int which = threadIdx.x >> 3, e = threadIdx.x & 7;
        float t = 0.f;
        for (int w2 = 0; w2 < static_cast<int>(blockDim.x >> 5); ++w2) t += s_a[which * 8 + w2][e];
        if (which == 0) atomicAdd(p.stat_a + c + e, t);
        else if (p.stat_b != nullptr) atomicAdd(p.stat_b + c + e, t);
    }
}

// ---------------------------------------------------------------------------------------------
// casts / input packing
// ---------------------------------------------------------------------------------------------
// fp32 NHWC [.., Cin] -> bf16 NHWC [.., Cpad] with zero fill of the pad channels.
__global__ void pack_input_kernel(const float* __restrict__ x, __nv_bfloat16* y, long long npos,
                                  int cin, int cpad) {
    pdl_enter();
    if (cpad == 8 && cin <= 8) {
        // one thread per position: cin scalar loads, one 16-byte store
        for (long long pos = blockIdx.x * (long long)blockDim.x + threadIdx.x; pos < npos;
             pos += (long long)gridDim.x * blockDim.x) {
            float v[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = c < cin ? x[pos * cin + c] : 0.f;
            *reinterpret_cast<uint4*>(y + pos * 8) =
                make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
        }
        return;
    }
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < npos * cpad;
         i += (long long)gridDim.x * blockDim.x) {
        const long long pos = i / cpad;
        const int c = static_cast<int>(i - pos * cpad);
        y[i] = __float2bfloat16_rn(c < cin ? x[pos * cin + c] : 0.f);
    }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* y, long long n) {
    pdl_enter();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        y[i] = __float2bfloat16_rn(x[i]);
}

static inline int ew_grid(long long work_items, int threads) {
    long long g = ceil_div_ll(work_items, threads);
    const long long cap = 8LL * num_sms();
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}

// ---------------------------------------------------------------------------------------------
// BatchNorm backward applied in ONE launch: every thread derives the coefficients of its 8 channels
// from the per-channel sums (what bn_bwd_coeffs / bn_shared_coeffs compute) and then streams
//     out = A*g + B*x + C0
// over its rows. mode 0: standalone training BN (also emits dgamma/dbeta), mode 1: frozen BN
// (out = scale*g), mode 2: DenseNet shared-buffer slice (A = 1, correction from S1/S2).
// grid = (row splits, ceil(C/64)); block = 256 = 8 channel octets x 32 row lanes.
// ---------------------------------------------------------------------------------------------
struct BnApplyArgs {
    const void* g; int g_is_float, g_pitch, g_coff;
    const __nv_bfloat16* x; int x_pitch, x_coff;
    __nv_bfloat16* out; int o_pitch, o_coff;
    const float* sa; const float* sb;      // mode 0: sum dpre, sum dpre*x ; mode 2: S1, S2
    const float* mean; const float* rstd; const float* scale;
    float count;
    float* dgamma; float* dbeta;
    int mode;
    long long M; int C;
    // mode 2 only, optional: the BatchNorm-backward bookkeeping of the LAST consumer of this block buffer that
    // has not been folded into S1/S2 yet (what bn_bwd_coeffs_kernel would have done in a launch of its own).
    // The pending BatchNorm read channels [0, p_C) of the buffer; this launch's slice starts at channel p_off of
    // that range. Every CTA adds the pending share of ITS channels on the fly; CTA (0,0) also writes the
    // pending layer's dgamma / dbeta and folds the other channels' shares into the S1 / S2 accumulators
    // (acc_s1 / acc_s2 point at channel 0 of the block). The slice's own accumulator entries are not
    // updated: they are final with this launch and never read again.
    const float* p_sa; const float* p_sb; const float* p_mean; const float* p_rstd; const float* p_scale;
    float* p_dgamma; float* p_dbeta; float* acc_s1; float* acc_s2;
    int p_off, p_C;
};

template <bool kGFloat>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(BnApplyArgs p) {
    pdl_enter();
    const int oct = threadIdx.x & 7;
    const int rl = threadIdx.x >> 3;
    const int c = blockIdx.y * 64 + oct * 8;
    if (p.p_sa != nullptr && blockIdx.x == 0 && blockIdx.y == 0) {
        for (int pc = threadIdx.x; pc < p.p_C; pc += blockDim.x) {
            const float a = p.p_sa[pc], s = p.p_scale[pc];
            const float dg = p.p_rstd[pc] * (p.p_sb[pc] - p.p_mean[pc] * a);
            if (p.p_dgamma) p.p_dgamma[pc] += dg;
            if (p.p_dbeta) p.p_dbeta[pc] += a;
            if (pc < p.p_off || pc >= p.p_off + p.C) {
                p.acc_s1[pc] += s * a;
                p.acc_s2[pc] += s * dg;
            }
        }
    }
    if (c >= p.C) return;
    float A[8], B[8], C0[8];
    const float invn = 1.f / p.count;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const int ch = c + e;
        if (p.mode == 1) {
            A[e] = p.scale[ch]; B[e] = 0.f; C0[e] = 0.f;
        } else if (p.mode == 0) {
            const float a = p.sa[ch], b = p.sb[ch], m = p.mean[ch], r = p.rstd[ch], s = p.scale[ch];
            const float dg = r * (b - m * a);
            A[e] = s;
            B[e] = -s * r * dg * invn;
            C0[e] = -s * a * invn + s * r * m * dg * invn;
            if (blockIdx.x == 0 && rl == 0) {
                if (p.dgamma) p.dgamma[ch] += dg;
                if (p.dbeta) p.dbeta[ch] += a;
            }
        } else {
            const float r = p.rstd[ch], m = p.mean[ch];
            float s1 = p.sa[ch], s2 = p.sb[ch];
            if (p.p_sa != nullptr) {
                const int pc = p.p_off + ch;
                const float a = p.p_sa[pc], s = p.p_scale[pc];
                s1 += s * a;
                s2 += s * (p.p_rstd[pc] * (p.p_sb[pc] - p.p_mean[pc] * a));
            }
            A[e] = 1.f;
            B[e] = -r * s2 * invn;
            C0[e] = -s1 * invn + r * m * s2 * invn;
        }
    }
    for (long long m = blockIdx.x * 32 + rl; m < p.M; m += (long long)gridDim.x * 32) {
        float gv[8];
        if (kGFloat) {
            const float* gp = static_cast<const float*>(p.g) + m * p.g_pitch + p.g_coff + c;
            const float4 a = *reinterpret_cast<const float4*>(gp);
            const float4 b = *reinterpret_cast<const float4*>(gp + 4);
            gv[0] = a.x; gv[1] = a.y; gv[2] = a.z; gv[3] = a.w;
            gv[4] = b.x; gv[5] = b.y; gv[6] = b.z; gv[7] = b.w;
        } else {
            const uint4 q = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.g) +
                                                            m * p.g_pitch + p.g_coff + c);
            const uint32_t* w = reinterpret_cast<const uint32_t*>(&q);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = unpack_bf16x2(w[e]);
                gv[2 * e] = f.x;
                gv[2 * e + 1] = f.y;
            }
        }
        const uint4 xq = *reinterpret_cast<const uint4*>(p.x + m * p.x_pitch + p.x_coff + c);
        const uint32_t* xw = reinterpret_cast<const uint32_t*>(&xq);
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 xf = unpack_bf16x2(xw[e]);
            ow[e] = pack_bf16x2(fmaf(A[2 * e], gv[2 * e], fmaf(B[2 * e], xf.x, C0[2 * e])),
                                fmaf(A[2 * e + 1], gv[2 * e + 1], fmaf(B[2 * e + 1], xf.y, C0[2 * e + 1])));
        }
        *reinterpret_cast<uint4*>(p.out + m * p.o_pitch + p.o_coff + c) = o;
    }
}

}  // namespace idcb

using namespace idcb;
typedef const __nv_bfloat16* cbf;
typedef __nv_bfloat16* bf;

extern "C" int idcb_bn_finalize(const float* sum, const float* sq, float count, const float* gamma,
                                const float* beta, float eps, float* scale, float* shift,
                                float* mean_out, float* rstd_out, float* moving_mean,
                                float* moving_var, float momentum, int C, void* stream) {
    IDCB_REQUIRE(C > 0 && count > 0.f, "bn_finalize: bad sizes");
    IDCB_CHECK(launch_pdl(bn_finalize_kernel, ceil_div(C, 128), 128, 0, (cudaStream_t)stream, 
        sum, sq, count, gamma, beta, eps, scale, shift, mean_out, rstd_out, moving_mean, moving_var,
        momentum, C));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_bn_fold(const float* gamma, const float* beta, const float* mean,
                            const float* var, float eps, float* scale, float* shift, int C,
                            void* stream) {
    IDCB_CHECK(launch_pdl(bn_fold_kernel, ceil_div(C, 128), 128, 0, (cudaStream_t)stream, gamma, beta, mean, var, eps,
                                                                       scale, shift, C));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_bn_bwd_coeffs(const float* sa, const float* sb, const float* mean,
                                  const float* rstd, const float* scale, float count, float* dgamma,
                                  float* dbeta, float* A, float* B, float* C0, float* acc_s1,
                                  float* acc_s2, int C, void* stream) {
    IDCB_REQUIRE((acc_s1 == nullptr) == (acc_s2 == nullptr), "bn_bwd_coeffs: S1/S2 go together");
    IDCB_CHECK(launch_pdl(bn_bwd_coeffs_kernel, ceil_div(C, 128), 128, 0, (cudaStream_t)stream, 
        sa, sb, mean, rstd, scale, count, dgamma, dbeta, A, B, C0, acc_s1, acc_s2, C));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_bn_shared_coeffs(const float* s1, const float* s2, const float* mean,
                                     const float* rstd, float count, float* A, float* B, float* C0,
                                     int C, void* stream) {
    IDCB_CHECK(launch_pdl(bn_shared_coeffs_kernel, ceil_div(C, 128), 128, 0, (cudaStream_t)stream, s1, s2, mean, rstd,
                                                                                count, A, B, C0, C));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_affine_act(const void* x, int x_pitch, int x_coff, void* y, int y_pitch,
                               int y_coff, const float* scale, const float* shift, int act,
                               const void* res, int r_pitch, int r_coff, long long M, int C,
                               void* stream) {
    IDCB_REQUIRE(C % 8 == 0 && x_pitch % 8 == 0 && y_pitch % 8 == 0 && x_coff % 8 == 0 &&
                     y_coff % 8 == 0 && r_pitch % 8 == 0 && r_coff % 8 == 0,
                 "affine_act: channel counts/pitches must be multiples of 8");
    if (M == 0) return 0;
    IDCB_CHECK(launch_pdl(affine_act_kernel, ew_grid(M * (C / 8), 256), 256, 0, (cudaStream_t)stream, 
        (cbf)x, x_pitch, x_coff, (bf)y, y_pitch, y_coff, scale, shift, act, (cbf)res, r_pitch, r_coff,
        M, C));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_affine2(const void* g, int g_is_float, int g_pitch, int g_coff, const void* x,
                            int x_pitch, int x_coff, void* out, int o_pitch, int o_coff,
                            const float* A, const float* B, const float* C0, long long M, int C,
                            void* stream) {
    IDCB_REQUIRE(C % 8 == 0 && g_pitch % 8 == 0 && g_coff % 8 == 0 && x_pitch % 8 == 0 &&
                     x_coff % 8 == 0 && o_pitch % 8 == 0 && o_coff % 8 == 0,
                 "affine2: channel counts/pitches must be multiples of 8");
    if (M == 0) return 0;
    const int grid = ew_grid(M * (C / 8), 256);
    if (g_is_float)
        IDCB_CHECK(launch_pdl(affine2_kernel<true>, grid, 256, 0, (cudaStream_t)stream, 
            g, g_pitch, g_coff, (cbf)x, x_pitch, x_coff, (bf)out, o_pitch, o_coff, A, B, C0, M, C));
    else
        IDCB_CHECK(launch_pdl(affine2_kernel<false>, grid, 256, 0, (cudaStream_t)stream, 
            g, g_pitch, g_coff, (cbf)x, x_pitch, x_coff, (bf)out, o_pitch, o_coff, A, B, C0, M, C));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_bn_bwd_apply(int mode, const void* g, int g_is_float, int g_pitch, int g_coff,
                                 const void* x, int x_pitch, int x_coff, void* out, int o_pitch,
                                 int o_coff, const float* sa, const float* sb, const float* mean,
                                 const float* rstd, const float* scale, float count, float* dgamma,
                                 float* dbeta, long long M, int C, void* stream) {
    IDCB_REQUIRE(mode >= 0 && mode <= 2, "bn_bwd_apply: bad mode %d", mode);
    IDCB_REQUIRE(C % 8 == 0 && g_pitch % 8 == 0 && g_coff % 8 == 0 && x_pitch % 8 == 0 &&
                     x_coff % 8 == 0 && o_pitch % 8 == 0 && o_coff % 8 == 0,
                 "bn_bwd_apply: channel counts/pitches must be multiples of 8");
    IDCB_REQUIRE(mode == 1 ? scale != nullptr : (sa && sb && mean && rstd && count > 0.f),
                 "bn_bwd_apply: missing per-channel inputs for mode %d", mode);
    IDCB_REQUIRE(mode != 0 || scale != nullptr, "bn_bwd_apply: mode 0 needs scale");
    if (M == 0) return 0;
    BnApplyArgs a{};
    a.g = g; a.g_is_float = g_is_float; a.g_pitch = g_pitch; a.g_coff = g_coff;
    a.x = (cbf)x; a.x_pitch = x_pitch; a.x_coff = x_coff;
    a.out = (bf)out; a.o_pitch = o_pitch; a.o_coff = o_coff;
    a.sa = sa; a.sb = sb; a.mean = mean; a.rstd = rstd; a.scale = scale; a.count = count;
    a.dgamma = dgamma; a.dbeta = dbeta; a.mode = mode; a.M = M; a.C = C;
    const int ygroups = ceil_div(C, 64);
    int splits = static_cast<int>(ceil_div_ll(M, 32 * 4));
    const int cap = ceil_div(4 * num_sms(), ygroups);
    if (splits > cap) splits = cap;
    if (splits < 1) splits = 1;
    dim3 grid(splits, ygroups);
    if (g_is_float)
        IDCB_CHECK(launch_pdl(bn_bwd_apply_kernel<true>, grid, 256, 0, (cudaStream_t)stream, a));
    else
        IDCB_CHECK(launch_pdl(bn_bwd_apply_kernel<false>, grid, 256, 0, (cudaStream_t)stream, a));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_channel_stats(const void* x, int pitch, int coff, long long M, int C, float* sum,
                                  float* sq, void* stream) {
    IDCB_REQUIRE(C % 8 == 0 && pitch % 8 == 0 && coff % 8 == 0, "channel_stats: alignment");
    if (M == 0) return 0;
    int splits = static_cast<int>(ceil_div_ll(M, 32 * 16));
    const int cap = ceil_div(4 * num_sms(), ceil_div(C, 64));
    if (splits > cap) splits = cap;
    if (splits < 1) splits = 1;
    dim3 grid(ceil_div(C, 64), splits);
    IDCB_CHECK(launch_pdl(channel_stats_kernel, grid, 256, 0, (cudaStream_t)stream, (cbf)x, pitch, coff, M, C, sum, sq));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_channel_stats2(const void* g, int g_pitch, int g_coff, const void* x, int x_pitch,
                                   int x_coff, long long M, int C, float* sum_g, float* sum_gx,
                                   void* stream) {
    IDCB_REQUIRE(C % 8 == 0 && g_pitch % 8 == 0 && g_coff % 8 == 0 && x_pitch % 8 == 0 && x_coff % 8 == 0,
                 "channel_stats2: alignment");
    if (M == 0) return 0;
    int splits = static_cast<int>(ceil_div_ll(M, 32 * 16));
    const int cap = ceil_div(4 * num_sms(), ceil_div(C, 64));
    if (splits > cap) splits = cap;
    if (splits < 1) splits = 1;
    dim3 grid(ceil_div(C, 64), splits);
    IDCB_CHECK(launch_pdl(channel_stats2_kernel, grid, 256, 0, (cudaStream_t)stream, (cbf)g, g_pitch, g_coff, (cbf)x, x_pitch,
                                                               x_coff, M, C, sum_g, sum_gx));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_pool_fwd(const idcb_pool_desc* d, const void* x, void* y, const float* scale,
                             const float* shift, int act, unsigned char* argmax, void* stream) {
    IDCB_REQUIRE(d && d->c % 8 == 0 && d->x_pitch % 8 == 0 && d->y_pitch % 8 == 0 &&
                     d->x_coff % 8 == 0 && d->y_coff % 8 == 0, "pool_fwd: alignment");
    PoolArgs p{};
    p.x = (cbf)x; p.x_pitch = d->x_pitch; p.x_coff = d->x_coff;
    p.y = (bf)y; p.y_pitch = d->y_pitch; p.y_coff = d->y_coff;
    p.B = d->batch; p.H = d->h; p.W = d->w; p.C = d->c; p.Ho = d->ho; p.Wo = d->wo;
    p.k = d->k; p.stride = d->stride; p.pad = d->pad;
    p.scale = scale; p.shift = shift; p.act = act; p.is_max = d->is_max;
    p.argmax = d->is_max ? argmax : nullptr;
    const long long items = 1LL * p.B * p.Ho * p.Wo * (p.C / 8);
    if (items == 0) return 0;
    IDCB_CHECK(launch_pdl(pool_fwd_kernel, ew_grid(items, 256), 256, 0, (cudaStream_t)stream, p));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_pool_bwd(const idcb_pool_desc* d, const void* x, const void* y, const void* dy,
                             int dy_pitch, int dy_coff, void* dx, int dx_pitch, int dx_coff,
                             const float* scale, const float* shift, int act, int relu_mask,
                             float* stat_a, float* stat_b, const unsigned char* argmax, void* stream) {
    IDCB_REQUIRE(d && d->c % 8 == 0 && d->x_pitch % 8 == 0 && d->y_pitch % 8 == 0 &&
                     dy_pitch % 8 == 0 && dx_pitch % 8 == 0 && dy_coff % 8 == 0 && dx_coff % 8 == 0,
                 "pool_bwd: alignment");
    PoolBwdArgs p{};
    p.x = (cbf)x; p.x_pitch = d->x_pitch; p.x_coff = d->x_coff;
    p.y = (cbf)y; p.y_pitch = d->y_pitch; p.y_coff = d->y_coff;
    p.dy = (cbf)dy; p.dy_pitch = dy_pitch; p.dy_coff = dy_coff;
    p.dx = (bf)dx; p.dx_pitch = dx_pitch; p.dx_coff = dx_coff;
    p.B = d->batch; p.H = d->h; p.W = d->w; p.C = d->c; p.Ho = d->ho; p.Wo = d->wo;
    p.k = d->k; p.stride = d->stride; p.pad = d->pad;
    p.scale = scale; p.shift = shift; p.act = act; p.is_max = d->is_max; p.relu_mask = relu_mask;
    p.stat_a = stat_a; p.stat_b = stat_b;
    p.argmax = d->is_max ? argmax : nullptr;
    const long long npos = 1LL * p.B * p.H * p.W;
    if (npos == 0) return 0;
    int gx = static_cast<int>(ceil_div_ll(npos, 256));
    const int cap = ceil_div(8 * num_sms(), p.C / 8);
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    dim3 grid(gx, p.C / 8);
    IDCB_CHECK(launch_pdl(pool_bwd_kernel, grid, 256, 0, (cudaStream_t)stream, p));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_pack_input(const float* x, void* y, long long npos, int cin, int cpad,
                               void* stream) {
    if (npos == 0) return 0;
IDCB_CHECK(launch_pdl(pack_input_kernel, ew_grid(cpad == 8 && cin <= 8 ? npos : npos * cpad, 256), 256, 0, (cudaStream_t)stream, x, (bf)y, npos, cin, cpad));
    IDCB_LAUNCH_CHECK();
    return 0;
}

extern "C" int idcb_cast_f32_bf16(const float* x, void* y, long long n, void* stream) {
    if (n == 0) return 0;
    IDCB_CHECK(launch_pdl(cast_f32_bf16_kernel, ew_grid(n, 256), 256, 0, (cudaStream_t)stream, x, (bf)y, n));
    IDCB_LAUNCH_CHECK();
    return 0;
}

// DenseNet block-buffer slice: dx = G - S1/N - xhat * S2/N, with the bookkeeping of the block's most recent
// consumer BatchNorm (dgamma / dbeta / its S1,S2 share) folded into the same launch.
extern "C" int idcb_bn_bwd_slice(const float* g, int g_pitch, int g_coff, const void* x, int x_pitch, int x_coff,
                                 void* out, int o_pitch, int o_coff, float* acc_s1, float* acc_s2, int c_lo,
                                 const float* mean, const float* rstd, float count, long long M, int C,
                                 const float* p_sa, const float* p_sb, const float* p_mean, const float* p_rstd,
                                 const float* p_scale, float* p_dgamma, float* p_dbeta, int p_C, void* stream) {
    IDCB_REQUIRE(C % 8 == 0 && g_pitch % 8 == 0 && g_coff % 8 == 0 && x_pitch % 8 == 0 && x_coff % 8 == 0 &&
                     o_pitch % 8 == 0 && o_coff % 8 == 0 && c_lo % 8 == 0,
                 "bn_bwd_slice: channel counts/pitches must be multiples of 8");
    IDCB_REQUIRE(acc_s1 && acc_s2 && mean && rstd && count > 0.f, "bn_bwd_slice: missing per-channel inputs");
    IDCB_REQUIRE(p_sa == nullptr || (p_sb && p_mean && p_rstd && p_scale && c_lo >= 0 && c_lo + C <= p_C),
                 "bn_bwd_slice: the slice [%d, %d) must lie inside the pending BatchNorm's %d channels", c_lo,
                 c_lo + C, p_C);
    if (M == 0) return 0;
    BnApplyArgs a{};
    a.g = g; a.g_is_float = 1; a.g_pitch = g_pitch; a.g_coff = g_coff;
    a.x = (cbf)x; a.x_pitch = x_pitch; a.x_coff = x_coff;
    a.out = (bf)out; a.o_pitch = o_pitch; a.o_coff = o_coff;
    a.sa = acc_s1 + c_lo; a.sb = acc_s2 + c_lo; a.mean = mean; a.rstd = rstd; a.count = count;
    a.mode = 2; a.M = M; a.C = C;
    a.p_sa = p_sa; a.p_sb = p_sb; a.p_mean = p_mean; a.p_rstd = p_rstd; a.p_scale = p_scale;
    a.p_dgamma = p_dgamma; a.p_dbeta = p_dbeta; a.acc_s1 = acc_s1; a.acc_s2 = acc_s2;
    a.p_off = c_lo; a.p_C = p_sa ? p_C : 0;
    const int ygroups = ceil_div(C, 64);
    int splits = static_cast<int>(ceil_div_ll(M, 32 * 4));
    const int cap = ceil_div(4 * num_sms(), ygroups);
    if (splits > cap) splits = cap;
    if (splits < 1) splits = 1;
    IDCB_CHECK(launch_pdl(bn_bwd_apply_kernel<true>, dim3(splits, ygroups), 256, 0, (cudaStream_t)stream, a));
    return 0;
}
