// Classification head in one kernel: (BN+activation prologue) -> GlobalAveragePooling2D -> Dense(k)
// -> loss (BCE / softmax-CE from logits) -> accuracy counters -> full backward of all of the above.
//
// Reference: GlobalAveragePooling2D + Dense(1|10) + BinaryCrossentropy/CategoricalCrossentropy
// (from_logits) + 'accuracy' appended by every transfer script
// (/root/reference/dist_model_tf_mobile.py:123-132, dist_model_tf_dense.py:135-144,
// dist_model_tf_vgg.py:123-132, fed_model.py:117-128). TensorFlow runs a reduction, a cuBLAS GEMM and
// several Eigen kernels for this; the per-step loss/accuracy scalars are
// accumulated into a 4-float buffer that the data-parallel layer reduces.
#include "common.cuh"
#include "../include/idcb200.h"

namespace idcb {

constexpr int kHeadThreads = 256;
constexpr int kMaxClasses = 16;

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x < kHeadThreads / 32) t = red[threadIdx.x];
    if (w == 0) {
        t = warp_sum(t);
        if (l == 0) red[0] = t;
    }
    __syncthreads();
    t = red[0];
    return t;
}

struct HeadArgs {
    const __nv_bfloat16* x; int x_pitch, x_coff;
    int B, HW, C, K;
    const float* pro_scale; const float* pro_shift; int pro_act;
    const float* w; const float* bias;    // fp32 [K][C], [K]
    const float* labels;                  // fp32 [B] (class index for CE)
    int loss_kind;                        // 0 = BCE from logits, 1 = softmax CE
    int acc_on_prob; float threshold;
    float* logits;                        // [B][K]
    float* acc;                           // [0] += sum loss, [1] += correct, [2] += count
    int training;
    float grad_scale;                     // 1 / global batch
    float* dw; float* db;                 // fp32 grads (accumulated)
    __nv_bfloat16* dx; int dx_pitch, dx_coff;   // dpre (w.r.t. scale*x+shift), optional
    float* gacc; int g_pitch, g_coff;           // optional fp32 accumulation, scaled by pro_scale
    float* stat_a; float* stat_b;               // sum dpre, sum dpre*x_raw
};

__global__ void __launch_bounds__(kHeadThreads) head_kernel(HeadArgs p) {
    pdl_enter();
    extern __shared__ float sm[];
    float* feat = sm;                 // [C]
    float* red = feat + p.C;          // [32]
    float* lg = red + 32;             // [K]
    float* dl = lg + kMaxClasses;     // [K]
    const int b = blockIdx.x;
    const float inv_hw = 1.f / p.HW;
    const __nv_bfloat16* xb = p.x + static_cast<size_t>(b) * p.HW * p.x_pitch + p.x_coff;
    for (int c = threadIdx.x; c < p.C; c += kHeadThreads) {
        const float s = p.pro_scale ? p.pro_scale[c] : 1.f;
        const float t = p.pro_scale ? p.pro_shift[c] : 0.f;
        float a = 0.f;
        for (int i = 0; i < p.HW; ++i) {
            float v = __bfloat162float(xb[static_cast<size_t>(i) * p.x_pitch + c]);
            if (p.pro_scale) v = apply_act(fmaf(v, s, t), p.pro_act);
            a += v;
        }
        feat[c] = a * inv_hw;
    }
    __syncthreads();
    for (int k = 0; k < p.K; ++k) {
        float a = 0.f;
        for (int c = threadIdx.x; c < p.C; c += kHeadThreads) a += p.w[static_cast<size_t>(k) * p.C + c] * feat[c];
        a = block_sum(a, red);
        if (threadIdx.x == 0) lg[k] = a + p.bias[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const float z = p.labels[b];
        float loss, correct;
        if (p.loss_kind == 0) {
            const float x = lg[0];
            loss = fmaxf(x, 0.f) - x * z + log1pf(expf(-fabsf(x)));
            const float sg = 1.f / (1.f + expf(-x));
            const float pred = ((p.acc_on_prob ? sg : x) > p.threshold) ? 1.f : 0.f;
            correct = (pred == z) ? 1.f : 0.f;
            dl[0] = (sg - z) * p.grad_scale;
        } else {
            float mx = lg[0];
            int am = 0;
            for (int k = 1; k < p.K; ++k)
                if (lg[k] > mx) { mx = lg[k]; am = k; }
            float se = 0.f;
            for (int k = 0; k < p.K; ++k) se += expf(lg[k] - mx);
            const int zi = static_cast<int>(z);
            loss = logf(se) + mx - lg[zi];
            correct = (am == zi) ? 1.f : 0.f;
            for (int k = 0; k < p.K; ++k)
                dl[k] = (expf(lg[k] - mx) / se - (k == zi ? 1.f : 0.f)) * p.grad_scale;
        }
        for (int k = 0; k < p.K; ++k) p.logits[static_cast<size_t>(b) * p.K + k] = lg[k];
        if (p.acc) {
            atomicAdd(p.acc + 0, loss);
            atomicAdd(p.acc + 1, correct);
            atomicAdd(p.acc + 2, 1.f);
        }
    }
    if (!p.training) return;
    __syncthreads();
    // ---- backward ----
    if (p.dw) {
        for (int k = 0; k < p.K; ++k) {
            const float d = dl[k];
            for (int c = threadIdx.x; c < p.C; c += kHeadThreads)
                atomicAdd(p.dw + static_cast<size_t>(k) * p.C + c, d * feat[c]);
            if (threadIdx.x == 0) atomicAdd(p.db + k, d);
        }
    }
    if (p.dx == nullptr && p.gacc == nullptr) return;
    for (int c = threadIdx.x; c < p.C; c += kHeadThreads) {
        float df = 0.f;
        for (int k = 0; k < p.K; ++k) df += dl[k] * p.w[static_cast<size_t>(k) * p.C + c];
        df *= inv_hw;
        const float s = p.pro_scale ? p.pro_scale[c] : 1.f;
        const float t = p.pro_scale ? p.pro_shift[c] : 0.f;
        float sa = 0.f, sb = 0.f;
        for (int i = 0; i < p.HW; ++i) {
            const size_t pos = static_cast<size_t>(b) * p.HW + i;
            const float xr = __bfloat162float(xb[static_cast<size_t>(i) * p.x_pitch + c]);
            float d = df;
            if (p.pro_scale) d *= act_mask(fmaf(xr, s, t), p.pro_act);
            sa += d;
            sb += d * xr;
            if (p.dx) p.dx[pos * p.dx_pitch + p.dx_coff + c] = __float2bfloat16_rn(d);
            if (p.gacc) p.gacc[pos * p.g_pitch + p.g_coff + c] += s * d;
        }
        if (p.stat_a) atomicAdd(p.stat_a + c, sa);
        if (p.stat_b) atomicAdd(p.stat_b + c, sb);
    }
}

}  // namespace idcb

using namespace idcb;

extern "C" int idcb_head(const idcb_head_desc* d, const void* x, const float* w, const float* bias,
                         const float* labels, float* logits, float* acc, float* dw, float* db, void* dx,
                         float* gacc, float* stat_a, float* stat_b, void* stream) {
    IDCB_REQUIRE(d != nullptr, "head: null descriptor");
    IDCB_REQUIRE(d->k >= 1 && d->k <= kMaxClasses, "head: 1..%d classes supported", kMaxClasses);
    IDCB_REQUIRE(d->c >= 1 && d->c <= 8192, "head: bad channel count");
    IDCB_REQUIRE(d->loss_kind == 0 ? d->k == 1 : true, "head: BCE needs a single logit");
    HeadArgs p{};
    p.x = (const __nv_bfloat16*)x; p.x_pitch = d->x_pitch; p.x_coff = d->x_coff;
    p.B = d->batch; p.HW = d->hw; p.C = d->c; p.K = d->k;
    p.pro_scale = d->pro_scale; p.pro_shift = d->pro_shift; p.pro_act = d->pro_act;
    p.w = w; p.bias = bias; p.labels = labels;
    p.loss_kind = d->loss_kind; p.acc_on_prob = d->acc_on_prob; p.threshold = d->threshold;
    p.logits = logits; p.acc = acc; p.training = d->training; p.grad_scale = d->grad_scale;
    p.dw = dw; p.db = db;
    p.dx = (__nv_bfloat16*)dx; p.dx_pitch = d->dx_pitch; p.dx_coff = d->dx_coff;
    p.gacc = gacc; p.g_pitch = d->g_pitch; p.g_coff = d->g_coff;
    p.stat_a = stat_a; p.stat_b = stat_b;
    if (d->batch == 0) return 0;
    const size_t smem = (d->c + 32 + 2 * kMaxClasses) * sizeof(float);
    IDCB_CHECK(launch_pdl(head_kernel, d->batch, kHeadThreads, smem, (cudaStream_t)stream, p));
    IDCB_LAUNCH_CHECK();
    return 0;
}
