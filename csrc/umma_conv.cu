// tcgen05 / TMEM / TMA implicit-GEMM convolutions for NHWC bf16 activations (sm_100a).
//
// One tile engine, three entry points:
//   idcb_conv_fprop : y = epilogue( conv( prologue(x), W ) )
//   idcb_conv_dgrad : dx = epilogue( conv^T( dy, W ) )            (stride 1 only)
//   idcb_conv_wgrad : dW += prologue(x)^T (*) dy                   (split over positions)
//
// Data flow per CTA (192 threads):
//   warps 0-3  : A-operand producers. They gather the im2col rows straight from global memory
//                (zero padding, channel-prefix views of a DenseNet concat buffer, and the fused
//                BatchNorm scale/shift + ReLU/ReLU6 prologue happen here, in registers), write the
//                128-byte-swizzled operand tile into shared memory, fence it to the async proxy and
//                arrive on the stage's "full" mbarrier. The loads of k-block i+1 are issued before
//                k-block i is transformed and stored, so two k-blocks of global traffic are in flight
//                per CTA. After the main loop the same four warps are the epilogue: tcgen05.ld the fp32
//                accumulators out of TMEM, apply the fused epilogue and store.
//   warp 4     : TMA producer for the second operand (weights, or dy for wgrad).
//   warp 5     : allocates TMEM, then one lane issues tcgen05.mma (UMMA 128 x N x 16, bf16 -> fp32)
//                and tcgen05.commit to recycle the stage / publish the accumulator.
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
constexpr int kMaxStages = 4;
constexpr int kThreads = 192;
constexpr int kATileBytes = kBM * 128;  // 16 KiB

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
    // prologue on A: a = act(scale[c] * x + shift[c]) (null scale => identity)
    const float* pro_scale;
    const float* pro_shift;
    int pro_act;
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
};

// Butterfly transpose-reduce: on return lane l holds the sum over the warp of column l.
__device__ __forceinline__ float warp_column_sums(float (&v)[32], int lane) {
#pragma unroll
    for (int n = 16; n >= 1; n >>= 1) {
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

__device__ __forceinline__ void transform8(uint4& q, const float* sc, const float* sh, int act) {
    uint32_t* w = reinterpret_cast<uint32_t*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        float2 f = unpack_bf16x2(w[e]);
        f.x = apply_act(fmaf(f.x, sc[2 * e], sh[2 * e]), act);
        f.y = apply_act(fmaf(f.y, sc[2 * e + 1], sh[2 * e + 1]), act);
        w[e] = pack_bf16x2(f.x, f.y);
    }
}

template <int kBN, bool kBMN>
struct FwdSmem {
    static constexpr int kBTileBytes = kBMN ? ((kBN + 63) / 64) * 8192 : kBN * 128;
    static constexpr int kStageBytes = kATileBytes + kBTileBytes;
};

// =============================================================================================
// fprop / dgrad kernel
// =============================================================================================
template <int kBN, bool kBMN, int kEpi>
__global__ void __launch_bounds__(kThreads, 2)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ ConvArgs p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    using L = FwdSmem<kBN, kBMN>;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    const int nstages = p.stages;
    uint8_t* a_tiles = smem;
    uint8_t* b_tiles = smem + nstages * kATileBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(b_tiles + nstages * L::kBTileBytes);
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* acc_bar = empty_bar + kMaxStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);
    float* s_scale = reinterpret_cast<float*>(tmem_slot + 4);
    float* s_shift = s_scale + p.Cin;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * kBM;
    const int n0 = blockIdx.y * kBN;
    const int taps = p.KH * p.KW;
    const int cblocks = (p.Cin + kBK - 1) / kBK;                 // dgrad: k-blocks per tap
    const int ktot = taps * p.Cin;                              // fprop: linearised K extent
    const int num_kb = kBMN ? taps * cblocks : (ktot + kBK - 1) / kBK;
    const bool has_pro = p.pro_scale != nullptr;

    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&full_bar[s], 4 + 1);  // 4 producer warps + the TMA thread
            mbar_init(&empty_bar[s], 1);     // one tcgen05.commit
        }
        mbar_init(acc_bar, 1);
        mbar_fence_init();
    }
    if (warp == 4 && lane == 0) tma_prefetch_desc(&tmap_w);
    if (warp == 5) tmem_alloc<kBN < 32 ? 32 : kBN>(tmem_slot);
    if (has_pro)
        for (int c = threadIdx.x; c < p.Cin; c += kThreads) {
            s_scale[c] = p.pro_scale[c];
            s_shift[c] = p.pro_shift[c];
        }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4) {
        // ===================== A producers =====================
        const int j = threadIdx.x & 7;       // 16-byte chunk inside the 128-byte row
        const int rbase = threadIdx.x >> 3;  // 0..15
        int hb[8], wb[8], ib[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int m = m0 + i * 16 + rbase;
            if (m < p.M) {
                const int hw = p.Ho * p.Wo;
                const int b = m / hw;
                const int rem = m - b * hw;
                const int ho = rem / p.Wo;
                const int wo = rem - ho * p.Wo;
                hb[i] = ho * p.stride - p.pad_t;
                wb[i] = wo * p.stride - p.pad_l;
                ib[i] = b * p.H * p.W;
            } else {
                hb[i] = -100000;
                wb[i] = -100000;
                ib[i] = 0;
            }
        }
        // issue the 8 row loads of k-block kb; returns the channel (or -1) and a validity mask
        auto issue = [&](int kb, uint4 (&v)[8], uint32_t& okmask) -> int {
            int tap, c;
            if (kBMN) {
                tap = kb / cblocks;
                c = (kb - tap * cblocks) * kBK + j * 8;
                if (c >= p.Cin) c = -1;
            } else {
                const int kidx = kb * kBK + j * 8;
                if (kidx < ktot) {
                    tap = kidx / p.Cin;
                    c = kidx - tap * p.Cin;
                } else {
                    tap = 0;
                    c = -1;
                }
            }
            const int kh = tap / p.KW;
            const int kw = tap - kh * p.KW;
            okmask = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int hi = hb[i] + kh;
                const int wi = wb[i] + kw;
                const bool ok = c >= 0 && hi >= 0 && hi < p.H && wi >= 0 && wi < p.W;
                v[i] = make_uint4(0u, 0u, 0u, 0u);
                if (ok) {
                    okmask |= 1u << i;
                    const size_t off =
                        (static_cast<size_t>(ib[i] + hi * p.W + wi)) * p.x_pitch + p.x_coff + c;
                    v[i] = __ldg(reinterpret_cast<const uint4*>(p.x + off));
                }
            }
            return c;
        };
        int stage = 0;
        uint32_t phase = 0;
        uint4 v[8], vn[8];
        uint32_t ok = 0, okn = 0;
        int c = issue(0, v, ok), cn = -1;
        for (int kb = 0; kb < num_kb; ++kb) {
            if (kb + 1 < num_kb) cn = issue(kb + 1, vn, okn);
            if (has_pro && c >= 0) {
                float sc[8], sh[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    sc[e] = s_scale[c + e];
                    sh[e] = s_shift[c + e];
                }
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (ok & (1u << i)) transform8(v[i], sc, sh, p.pro_act);
            }
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* a_tile = a_tiles + stage * kATileBytes;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = i * 16 + rbase;
                *reinterpret_cast<uint4*>(a_tile + sw128_offset(r, j)) = v[i];
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&full_bar[stage]);
            if (++stage == nstages) {
                stage = 0;
                phase ^= 1;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = vn[i];
            ok = okn;
            c = cn;
        }

        // ===================== epilogue =====================
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        const int row = warp * 32 + lane;
        const int m = m0 + row;
        const bool mvalid = m < p.M;
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
#pragma unroll 1
        for (int cc = 0; cc < kBN; cc += 32) {
            const int nb = n0 + cc;
            if (nb >= p.Cout) break;  // uniform across the CTA
            uint32_t raw[32];
            tmem_ld_32x32b_x32(t_row + cc, raw);
            tmem_ld_wait();
            float vv[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) vv[e] = __uint_as_float(raw[e]);
            float xr[32];  // mode 0: residual / mode 1: mask source / mode 2: raw x
            const bool need_aux = (kEpi != 0) || (p.aux != nullptr);
            if (need_aux) {
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    uint4 q = make_uint4(0u, 0u, 0u, 0u);
                    if (mvalid && nb + g * 8 < p.Cout)
                        q = *reinterpret_cast<const uint4*>(
                            p.aux + static_cast<size_t>(m) * p.aux_pitch + p.aux_coff + nb + g * 8);
                    const uint32_t* w = reinterpret_cast<const uint32_t*>(&q);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float2 f = unpack_bf16x2(w[e]);
                        xr[g * 8 + 2 * e] = f.x;
                        xr[g * 8 + 2 * e + 1] = f.y;
                    }
                }
            }
            float sa[32], sb[32];
            if (kEpi == 0) {
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    const int n = nb + e;
                    float o = vv[e];
                    if (n < p.Cout) {
                        if (p.epi_scale) o *= p.epi_scale[n];
                        if (p.epi_shift) o += p.epi_shift[n];
                        o = apply_act(o, p.epi_act);
                        if (need_aux) o += xr[e];
                    } else {
                        o = 0.f;
                    }
                    o = __bfloat162float(__float2bfloat16_rn(o));
                    vv[e] = o;
                    sa[e] = mvalid ? o : 0.f;
                    sb[e] = mvalid ? o * o : 0.f;
                }
            } else if (kEpi == 1) {
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    float o = (nb + e < p.Cout && xr[e] > 0.f) ? vv[e] : 0.f;
                    o = __bfloat162float(__float2bfloat16_rn(o));
                    vv[e] = o;
                    sa[e] = mvalid ? o : 0.f;
                    sb[e] = 0.f;
                }
            } else {
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    const int n = nb + e;
                    float o = 0.f;
                    float gs = 0.f;
                    if (n < p.Cout) {
                        gs = p.bwd_scale[n];
                        const float pre = fmaf(gs, xr[e], p.bwd_shift[n]);
                        o = vv[e] * act_mask(pre, p.bwd_act);
                    }
                    sa[e] = mvalid ? o : 0.f;
                    sb[e] = mvalid ? o * xr[e] : 0.f;
                    vv[e] = o;
                    xr[e] = gs;  // reuse as the G scale
                }
            }
            // ---- stores ----
            if (mvalid && p.y != nullptr) {
                __nv_bfloat16* yrow = p.y + static_cast<size_t>(m) * p.y_pitch + p.y_coff + nb;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    if (nb + g * 8 < p.Cout) {
                        uint4 q;
                        q.x = pack_bf16x2(vv[g * 8 + 0], vv[g * 8 + 1]);
                        q.y = pack_bf16x2(vv[g * 8 + 2], vv[g * 8 + 3]);
                        q.z = pack_bf16x2(vv[g * 8 + 4], vv[g * 8 + 5]);
                        q.w = pack_bf16x2(vv[g * 8 + 6], vv[g * 8 + 7]);
                        *reinterpret_cast<uint4*>(yrow + g * 8) = q;
                    }
                }
            }
            if (kEpi == 2 && mvalid && p.gacc != nullptr) {
                float* grow = p.gacc + static_cast<size_t>(m) * p.g_pitch + p.g_coff + nb;
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    if (nb + g * 4 < p.Cout) {
                        float4 q = *reinterpret_cast<float4*>(grow + g * 4);
                        q.x = fmaf(xr[g * 4 + 0], vv[g * 4 + 0], q.x);
                        q.y = fmaf(xr[g * 4 + 1], vv[g * 4 + 1], q.y);
                        q.z = fmaf(xr[g * 4 + 2], vv[g * 4 + 2], q.z);
                        q.w = fmaf(xr[g * 4 + 3], vv[g * 4 + 3], q.w);
                        *reinterpret_cast<float4*>(grow + g * 4) = q;
                    }
                }
            }
            // ---- column statistics (BatchNorm sums / bias gradient / BN-backward sums) ----
            if (p.stat_a != nullptr) {
                const float ca = warp_column_sums(sa, lane);
                if (nb + lane < p.Cout) atomicAdd(p.stat_a + nb + lane, ca);
                if (kEpi != 1 && p.stat_b != nullptr) {
                    const float cb2 = warp_column_sums(sb, lane);
                    if (nb + lane < p.Cout) atomicAdd(p.stat_b + nb + lane, cb2);
                }
            }
        }
        tc_fence_before();
    } else if (warp == 4) {
        // ===================== TMA producer for the weight operand =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* b_tile = b_tiles + stage * L::kBTileBytes;
                mbar_arrive_expect_tx(&full_bar[stage], L::kBTileBytes);
                if (!kBMN) {
                    // K-major: rows = output channels, 64 linearised (tap, cin) K elements per row
                    tma_load_2d(b_tile, &tmap_w, &full_bar[stage], kb * kBK, n0);
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
                tc_fence_after();
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
        }
        __syncwarp();
    }
    __syncthreads();
    if (warp == 5) {
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
    const float* pro_scale;
    const float* pro_shift;
    int pro_act;
    float* dw;  // fp32 [Cout][KH*KW*Cin], accumulated with red.add
    int splits;
    int kb_per_split;  // 64-position blocks per split
    int stages;
};

template <int kBN>
__global__ void __launch_bounds__(kThreads, 2)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ WgradArgs p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    constexpr int kBTile = ((kBN + 63) / 64) * 8192;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    const int nstages = p.stages;
    uint8_t* a_tiles = smem;  // [64 pos][128 k'] as two 8 KiB chunks
    uint8_t* b_tiles = smem + nstages * kATileBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(b_tiles + nstages * kBTile);
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* acc_bar = empty_bar + kMaxStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

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

    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&full_bar[s], 4 + 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(acc_bar, 1);
        mbar_fence_init();
    }
    if (warp == 4 && lane == 0) tma_prefetch_desc(&tmap_dy);
    if (warp == 5) tmem_alloc<kBN < 32 ? 32 : kBN>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (num_kb == 0) {  // uniform: nothing to do for this split
        __syncthreads();
        if (warp == 5) tmem_dealloc<kBN < 32 ? 32 : kBN>(tmem_base);
        return;
    }

    if (warp < 4) {
        // ---- A producers: 16 threads cover one position row (128 k' = 256 B) ----
        const int j16 = threadIdx.x & 15;
        const int half = j16 >> 3;  // which 64-wide chunk
        const int j = j16 & 7;
        const int rbase = threadIdx.x >> 4;  // 0..7
        const int kidx = k0 + j16 * 8;
        const bool cvalid = kidx < ktot;
        const int tap = cvalid ? kidx / p.Cin : 0;
        const int c = cvalid ? kidx - tap * p.Cin : 0;
        const int kh = tap / p.KW;
        const int kw = tap - kh * p.KW;
        const bool has_pro = p.pro_scale != nullptr;
        float sc[8], sh[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            sc[e] = (has_pro && cvalid) ? p.pro_scale[c + e] : 1.f;
            sh[e] = (has_pro && cvalid) ? p.pro_shift[c + e] : 0.f;
        }
        const int hw = p.Ho * p.Wo;
        auto issue = [&](int kb, uint4 (&v)[8], uint32_t& okmask) {
            okmask = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int m = kb * 64 + i * 8 + rbase;
                v[i] = make_uint4(0u, 0u, 0u, 0u);
                if (cvalid && m < p.M) {
                    const int b = m / hw;
                    const int rem = m - b * hw;
                    const int ho = rem / p.Wo;
                    const int wo = rem - ho * p.Wo;
                    const int hi = ho * p.stride - p.pad_t + kh;
                    const int wi = wo * p.stride - p.pad_l + kw;
                    if (hi >= 0 && hi < p.H && wi >= 0 && wi < p.W) {
                        okmask |= 1u << i;
                        const size_t off =
                            (static_cast<size_t>((b * p.H + hi) * p.W + wi)) * p.x_pitch + p.x_coff + c;
                        v[i] = __ldg(reinterpret_cast<const uint4*>(p.x + off));
                    }
                }
            }
        };
        int stage = 0;
        uint32_t phase = 0;
        uint4 v[8], vn[8];
        uint32_t ok = 0, okn = 0;
        issue(kb_begin, v, ok);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            if (kb + 1 < kb_end) issue(kb + 1, vn, okn);
            if (has_pro) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (ok & (1u << i)) transform8(v[i], sc, sh, p.pro_act);
            }
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* a_tile = a_tiles + stage * kATileBytes + half * 8192;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = i * 8 + rbase;
                *reinterpret_cast<uint4*>(a_tile + sw128_offset(r, j)) = v[i];
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&full_bar[stage]);
            if (++stage == nstages) {
                stage = 0;
                phase ^= 1;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = vn[i];
            ok = okn;
        }
        // ---- epilogue: red.add the fp32 tile into dW ----
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        const int krow = k0 + warp * 32 + lane;
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
#pragma unroll 1
        for (int cc = 0; cc < kBN; cc += 32) {
            if (n0 + cc >= p.Cout) break;
            uint32_t raw[32];
            tmem_ld_32x32b_x32(t_row + cc, raw);
            tmem_ld_wait();
            if (krow < ktot) {
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    const int n = n0 + cc + e;
                    if (n < p.Cout)
                        atomicAdd(p.dw + static_cast<size_t>(n) * ktot + krow, __uint_as_float(raw[e]));
                }
            }
        }
        tc_fence_before();
    } else if (warp == 4) {
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
    if (warp == 5) {
        tc_fence_after();
        tmem_dealloc<kBN < 32 ? 32 : kBN>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------
// Host launchers
// ---------------------------------------------------------------------------------------------
static int pick_stages(int num_kb, int stage_bytes, int extra_bytes) {
    int s = num_kb < kMaxStages ? num_kb : kMaxStages;
    if (s < 1) s = 1;
    // keep two CTAs resident per SM when the pipeline depth allows it (latency hiding across CTAs)
    while (s > 2 && 2 * (1024 + s * stage_bytes + extra_bytes + 256) > 227 * 1024) --s;
    return s;
}

template <int kBN, bool kBMN, int kEpi>
static int launch_gemm(const CUtensorMap& tm, ConvArgs& a, cudaStream_t st) {
    using L = FwdSmem<kBN, kBMN>;
    const int taps = a.KH * a.KW;
    const int num_kb = kBMN ? taps * ceil_div(a.Cin, kBK) : ceil_div(taps * a.Cin, kBK);
    const int extra = (2 * kMaxStages + 1) * 8 + 16 + 2 * (int)sizeof(float) * a.Cin + 64;
    a.stages = pick_stages(num_kb, L::kStageBytes, extra);
    const size_t smem = 1024 + a.stages * L::kStageBytes + extra;
    IDCB_REQUIRE(smem <= 227 * 1024, "conv: Cin=%d needs %zu bytes of shared memory", a.Cin, smem);
    auto kern = conv_gemm_kernel<kBN, kBMN, kEpi>;
    static bool attr_set = false;
    if (!attr_set) {
        IDCB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
        attr_set = true;
    }
    dim3 grid(ceil_div(a.M, kBM), ceil_div(a.Cout, kBN), 1);
    kern<<<grid, kThreads, smem, st>>>(tm, a);
    IDCB_LAUNCH_CHECK();
    return 0;
}

template <int kBN>
static int launch_wgrad(const CUtensorMap& tm, WgradArgs& a, cudaStream_t st) {
    constexpr int kBTile = ((kBN + 63) / 64) * 8192;
    const int extra = (2 * kMaxStages + 1) * 8 + 16 + 64;
    a.stages = pick_stages(a.kb_per_split, kATileBytes + kBTile, extra);
    const size_t smem = 1024 + a.stages * (kATileBytes + kBTile) + extra;
    auto kern = conv_wgrad_kernel<kBN>;
    static bool attr_set = false;
    if (!attr_set) {
        IDCB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
        attr_set = true;
    }
    dim3 grid(ceil_div(a.KH * a.KW * a.Cin, 128), ceil_div(a.Cout, kBN), a.splits);
    kern<<<grid, kThreads, smem, st>>>(tm, a);
    IDCB_LAUNCH_CHECK();
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
    const long long m = 1LL * d->batch * d->ho * d->wo;
    IDCB_REQUIRE(m > 0 && m < (1LL << 30), "conv: bad problem size");
    return 0;
}

}  // namespace idcb

using namespace idcb;

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
    }
    CUtensorMap tm;
    const int kbn = d->cout <= 32 ? 32 : (d->cout <= 64 ? 64 : 128);
    if (make_tmap_2d_bf16(&tm, w, d->cout, (uint64_t)d->kh * d->kw * d->cin,
                          (uint64_t)d->kh * d->kw * d->cin, kbn, 64))
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
    IDCB_REQUIRE(a.y != nullptr || a.gacc != nullptr, "conv_dgrad: no output given");
    CUtensorMap tm;
    const int kbn = d->cin <= 64 ? 64 : 128;
    // weight view [Cout rows][taps*Cin cols]; MN-major boxes are 64 cols x 64 rows
    if (make_tmap_2d_bf16(&tm, w, d->cout, (uint64_t)d->kh * d->kw * d->cin,
                          (uint64_t)d->kh * d->kw * d->cin, 64, 64))
        return 1;
    if (kbn == 64) {
        if (mode == 0) return launch_gemm<64, true, 0>(tm, a, st);
        if (mode == 1) return launch_gemm<64, true, 1>(tm, a, st);
        return launch_gemm<64, true, 2>(tm, a, st);
    }
    if (mode == 0) return launch_gemm<128, true, 0>(tm, a, st);
    if (mode == 1) return launch_gemm<128, true, 1>(tm, a, st);
    return launch_gemm<128, true, 2>(tm, a, st);
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
    a.dw = dw;
    const int kbn = d->cout <= 32 ? 32 : (d->cout <= 64 ? 64 : 128);
    const int total_kb = ceil_div(a.M, 64);
    const int base_ctas = ceil_div(a.KH * a.KW * a.Cin, 128) * ceil_div(a.Cout, kbn);
    int splits = ceil_div(2 * num_sms(), base_ctas);
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
