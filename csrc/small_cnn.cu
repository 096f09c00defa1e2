// The whole 10x10 secure-FL CNN in one kernel per direction-pair: forward + loss + metrics + backward for a
// batch, one thread block per example, everything in shared memory / registers (1,937 parameters):
//   Conv2D(32, 3x3, stride 2, VALID, relu) 10x10x3 -> 4x4x32 ; MaxPool 2x2 -> 2x2x32 ; Dropout(.25) ;
//   Flatten (NHWC order, 128) ; Dense(8, relu) ; Dropout(.5) ; Dense(1) ; BCE-from-logits.
// Reference: create_model() in /root/reference/secure_fed_model.py:84-98.
// TensorFlow runs ~20 cuDNN/Eigen kernels per step for this model; at batch 32 each is pure launch
// latency. fp32 throughout (the model is far too small for tensor cores to matter). Many clients can be
// trained concurrently by giving each its own parameter/gradient block: blockIdx.y selects the client.
// Dropout masks come from Philox4x32-10 keyed by (seed, step), counter = (example, feature) — reproducible
// and identical between forward and backward without storing the mask.
#include "common.cuh"
#include "philox.cuh"
#include "../include/idcb200.h"

namespace idcb {

constexpr int kH = 10, kC = 3, kF = 32, kConvOut = 4, kPool = 2, kFlat = 128, kHid = 8;
// The six tensors (conv W [32][3][3][cstride], conv b [32], dense W [8][128], dense b [8], out W [1][8],
// out b [1]) live at caller-given offsets inside one parameter block of at most kMaxBlock floats (the flat
// ParamStore layout: 64-element aligned slots, RGB kernels channel-padded to cstride = 8).
constexpr int kMaxBlock = 4096;

struct SmallCnnArgs {
    const float* x;        // [clients][batch][10][10][3]
    const float* labels;   // [clients][batch]
    const float* params;   // [clients][param_stride]
    float* grads;          // [clients][param_stride]   (accumulated)
    float* logits;         // [clients][batch]
    float* acc;            // [clients][4]: loss sum, correct, count
    int batch, param_stride;
    int oCW, oCB, oDW, oDB, oOW, oOB, cstride, span;
    int training;
    float grad_scale;      // 1 / batch
    float drop1, drop2;
    unsigned long long seed;
    unsigned int step;
    float threshold;
    int acc_on_prob;
    // optional: the batch is a gather out of each client's whole shard (x / labels are then [clients][n_ex][...])
    const long long* idx;      // [clients][idx_stride]: example index of batch element ex at idx[cl * idx_stride + ex]
    long long idx_stride;
    int n_ex;
};

__device__ __forceinline__ bool keep(unsigned long long seed, unsigned step, unsigned ex, unsigned feat, unsigned layer,
                                     float rate) {
    if (rate <= 0.f) return true;
    const uint4 r = philox4x32_10(ex, feat, layer, step, static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    return (r.x >> 8) * (1.0f / 16777216.0f) >= rate;
}

__global__ void __launch_bounds__(128) small_cnn_kernel(SmallCnnArgs a) {
    pdl_enter();
    __shared__ float sx[kH * kH * kC];
    __shared__ float sW[kMaxBlock];
    __shared__ float conv[kConvOut * kConvOut * kF];   // post-relu conv output
    __shared__ float flat[kFlat];                      // pooled + dropout
    __shared__ int amax[kFlat];                        // argmax position inside conv[] (or -1 if dropped)
    __shared__ float hid[kHid], hpre[kHid], dhid[kHid];
    __shared__ float dflat[kFlat];
    __shared__ float dconv[kConvOut * kConvOut * kF];
    __shared__ float s_dl;
    const int ex = blockIdx.x, cl = blockIdx.y, t = threadIdx.x;
    const float* P = a.params + static_cast<size_t>(cl) * a.param_stride;
    // where this example lives: a dense [clients][batch] batch, or an index into the client's resident shard
    const size_t row = a.idx != nullptr
                           ? static_cast<size_t>(cl) * a.n_ex + static_cast<size_t>(a.idx[static_cast<size_t>(cl) * a.idx_stride + ex])
                           : static_cast<size_t>(cl) * a.batch + ex;
    const float* X = a.x + row * (kH * kH * kC);
    const int oCW = a.oCW, oCB = a.oCB, oDW = a.oDW, oDB = a.oDB, oOW = a.oOW, oOB = a.oOB, cs = a.cstride;
    for (int i = t; i < a.span; i += 128) sW[i] = P[i];
    for (int i = t; i < kH * kH * kC; i += 128) sx[i] = X[i];
    __syncthreads();
    // conv 3x3 stride 2 VALID + bias + relu: 4*4*32 = 512 outputs, 4 per thread
    for (int o = t; o < kConvOut * kConvOut * kF; o += 128) {
        const int f = o % kF, pos = o / kF, oh = pos / kConvOut, ow = pos % kConvOut;
        float s = sW[oCB + f];
        for (int kh = 0; kh < 3; ++kh)
            for (int kw = 0; kw < 3; ++kw)
                for (int c = 0; c < kC; ++c)
                    s += sx[((oh * 2 + kh) * kH + (ow * 2 + kw)) * kC + c] * sW[oCW + ((f * 3 + kh) * 3 + kw) * cs + c];
        conv[o] = fmaxf(s, 0.f);
    }
    __syncthreads();
    // maxpool 2x2 + dropout(.25) + flatten in NHWC order: feature = (ph*2+pw)*32 + f
    {
        const int f = t % kF, pp = t / kF, ph = pp / kPool, pw = pp % kPool;
        float best = -1.f;
        int bi = 0;
        for (int dh = 0; dh < 2; ++dh)
            for (int dw = 0; dw < 2; ++dw) {
                const int idx = ((ph * 2 + dh) * kConvOut + (pw * 2 + dw)) * kF + f;
                if (conv[idx] > best) { best = conv[idx]; bi = idx; }   // first max wins
            }
        const bool k1 = !a.training || keep(a.seed, a.step, ex + cl * a.batch, t, 1, a.drop1);
        flat[t] = k1 ? best * (a.training ? 1.f / (1.f - a.drop1) : 1.f) : 0.f;
        amax[t] = k1 ? bi : -1;
    }
    __syncthreads();
    // dense(8) + relu + dropout(.5): one warp per two hidden units
    {
        const int w = t >> 5, lane = t & 31;
        for (int h = w; h < kHid; h += 4) {
            float s = 0.f;
            for (int i = lane; i < kFlat; i += 32) s += flat[i] * sW[oDW + h * kFlat + i];
            s = warp_sum(s);
            if (lane == 0) {
                s += sW[oDB + h];
                hpre[h] = s;
                const bool k2 = !a.training || keep(a.seed, a.step, ex + cl * a.batch, h, 2, a.drop2);
                hid[h] = k2 ? fmaxf(s, 0.f) * (a.training ? 1.f / (1.f - a.drop2) : 1.f) : 0.f;
                dhid[h] = k2 ? (a.training ? 1.f / (1.f - a.drop2) : 1.f) : 0.f;   // dropout factor
            }
        }
    }
    __syncthreads();
    if (t == 0) {
        float x = sW[oOB];
        for (int h = 0; h < kHid; ++h) x += hid[h] * sW[oOW + h];
        const float z = a.labels[row];
        const float loss = fmaxf(x, 0.f) - x * z + log1pf(expf(-fabsf(x)));
        const float sg = 1.f / (1.f + expf(-x));
        const float pred = ((a.acc_on_prob ? sg : x) > a.threshold) ? 1.f : 0.f;
        a.logits[static_cast<size_t>(cl) * a.batch + ex] = x;
        atomicAdd(a.acc + cl * 4 + 0, loss);
        atomicAdd(a.acc + cl * 4 + 1, pred == z ? 1.f : 0.f);
        atomicAdd(a.acc + cl * 4 + 2, 1.f);
        s_dl = (sg - z) * a.grad_scale;
    }
    if (!a.training) return;
    __syncthreads();
    float* G = a.grads + static_cast<size_t>(cl) * a.param_stride;
    const float dl = s_dl;
    // output layer grads + hidden grads
    if (t < kHid) {
        atomicAdd(G + oOW + t, dl * hid[t]);
        float d = dl * sW[oOW + t] * dhid[t];
        d = hpre[t] > 0.f ? d : 0.f;
        dhid[t] = d;                       // d loss / d hpre
        atomicAdd(G + oDB + t, d);
    }
    if (t == 0) atomicAdd(G + oOB, dl);
    __syncthreads();
    // dense(8) weight grads and d flat
    {
        float df = 0.f;
        for (int h = 0; h < kHid; ++h) {
            atomicAdd(G + oDW + h * kFlat + t, dhid[h] * flat[t]);
            df += dhid[h] * sW[oDW + h * kFlat + t];
        }
        dflat[t] = amax[t] >= 0 ? df * (1.f / (1.f - a.drop1)) : 0.f;
    }
    for (int o = t; o < kConvOut * kConvOut * kF; o += 128) dconv[o] = 0.f;
    __syncthreads();
    if (amax[t] >= 0 && conv[amax[t]] > 0.f) dconv[amax[t]] = dflat[t];   // pool windows are disjoint
    __syncthreads();
    // conv grads: 864 weights + 32 biases
    for (int wi = t; wi < 864 + kF; wi += 128) {
        float g = 0.f;
        if (wi < 864) {
            const int c = wi % kC, kw = (wi / kC) % 3, kh = (wi / (kC * 3)) % 3, f = wi / 27;
            for (int oh = 0; oh < kConvOut; ++oh)
                for (int ow = 0; ow < kConvOut; ++ow)
                    g += dconv[(oh * kConvOut + ow) * kF + f] * sx[((oh * 2 + kh) * kH + (ow * 2 + kw)) * kC + c];
            if (g != 0.f) atomicAdd(G + oCW + ((f * 3 + kh) * 3 + kw) * cs + c, g);
        } else {
            const int f = wi - 864;
            for (int p = 0; p < kConvOut * kConvOut; ++p) g += dconv[p * kF + f];
            if (g != 0.f) atomicAdd(G + oCB + f, g);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Device-side synthetic IDC-like patches: Philox noise + a class-dependent tint, the same
// recipe as data/synthetic.py (statistically, not bit-wise), written straight into the training batch.
// ---------------------------------------------------------------------------------------------
__global__ void synth_patches_kernel(float* x, float* labels, long long n_img, int hwc, int channels,
                                     unsigned long long seed, unsigned int step) {
    pdl_enter();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_img * hwc;
         i += (long long)gridDim.x * blockDim.x) {
        const long long img = i / hwc;
        const int c = static_cast<int>(i % channels);
        const uint4 r = philox4x32_10(static_cast<uint32_t>(i), static_cast<uint32_t>(i >> 32), step, 7u,
                                      static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
        const uint4 l = philox4x32_10(static_cast<uint32_t>(img), static_cast<uint32_t>(img >> 32), step, 11u,
                                      static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
        const int label = l.x & 1;
        const float noise = (r.x >> 8) * (1.0f / 16777216.0f);
        const float tint_pos[3] = {0.55f, 0.35f, 0.60f}, tint_neg[3] = {0.85f, 0.70f, 0.80f};
        const float tint = label ? tint_pos[c % 3] : tint_neg[c % 3];
        x[i] = 0.55f * noise + 0.45f * tint;
        if (i % hwc == 0 && labels) labels[img] = static_cast<float>(label);
    }
}

}  // namespace idcb

using namespace idcb;

extern "C" int idcb_small_cnn_step(const float* x, const float* labels, const float* params, float* grads,
                                   float* logits, float* acc, int n_clients, int batch, int param_stride,
                                   const int* offsets6, int conv_cin_stride, int training, float grad_scale, float drop1, float drop2,
                                   unsigned long long seed, unsigned int step, float threshold, int acc_on_prob,
                                   void* stream) {
    IDCB_REQUIRE(n_clients >= 1 && batch >= 0 && offsets6 != nullptr && conv_cin_stride >= 3, "small_cnn_step: bad sizes");
    const int span = offsets6[5] + 1;
    IDCB_REQUIRE(span <= kMaxBlock && param_stride >= span, "small_cnn_step: parameter block of %d floats exceeds %d", span, kMaxBlock);
    IDCB_REQUIRE(drop1 >= 0.f && drop1 < 1.f && drop2 >= 0.f && drop2 < 1.f, "small_cnn_step: bad dropout rates");
    IDCB_REQUIRE(!training || grads != nullptr, "small_cnn_step: training needs a gradient buffer");
    if (batch == 0) return 0;
    SmallCnnArgs a{x, labels, params, grads, logits, acc, batch, param_stride,
                   offsets6[0], offsets6[1], offsets6[2], offsets6[3], offsets6[4], offsets6[5], conv_cin_stride, span,
                   training, grad_scale, drop1, drop2, seed, step, threshold, acc_on_prob, nullptr, 0, 0};
    IDCB_CHECK(launch_pdl(small_cnn_kernel, dim3(batch, n_clients), 128, 0, (cudaStream_t)stream, a));
    IDCB_LAUNCH_CHECK();
    return 0;
}

// The same step with the batch gathered inside the kernel: x_all [clients][n_ex][10][10][3] and labels_all
// [clients][n_ex] are the clients' whole shards, idx [clients][idx_stride] (int64, e.g. a row of torch.randperm per
// client, already offset to the batch's first column) names the n_ex-relative example of every batch slot.
extern "C" int idcb_small_cnn_step_indexed(const float* x_all, const float* labels_all, const long long* idx,
                                           long long idx_stride, int n_ex, const float* params, float* grads,
                                           float* logits, float* acc, int n_clients, int batch, int param_stride,
                                           const int* offsets6, int conv_cin_stride, int training, float grad_scale,
                                           float drop1, float drop2, unsigned long long seed, unsigned int step,
                                           float threshold, int acc_on_prob, void* stream) {
    IDCB_REQUIRE(n_clients >= 1 && batch >= 0 && offsets6 != nullptr && conv_cin_stride >= 3, "small_cnn_step: bad sizes");
    IDCB_REQUIRE(idx != nullptr && n_ex >= 1 && idx_stride >= batch, "small_cnn_step_indexed: bad index description");
    const int span = offsets6[5] + 1;
    IDCB_REQUIRE(span <= kMaxBlock && param_stride >= span, "small_cnn_step: parameter block of %d floats exceeds %d", span, kMaxBlock);
    IDCB_REQUIRE(drop1 >= 0.f && drop1 < 1.f && drop2 >= 0.f && drop2 < 1.f, "small_cnn_step: bad dropout rates");
    IDCB_REQUIRE(!training || grads != nullptr, "small_cnn_step: training needs a gradient buffer");
    if (batch == 0) return 0;
    SmallCnnArgs a{x_all, labels_all, params, grads, logits, acc, batch, param_stride,
                   offsets6[0], offsets6[1], offsets6[2], offsets6[3], offsets6[4], offsets6[5], conv_cin_stride, span,
                   training, grad_scale, drop1, drop2, seed, step, threshold, acc_on_prob, idx, idx_stride, n_ex};
    IDCB_CHECK(launch_pdl(small_cnn_kernel, dim3(batch, n_clients), 128, 0, (cudaStream_t)stream, a));
    return 0;
}

extern "C" int idcb_synth_patches(float* x, float* labels, long long n_img, int h, int w, int channels,
                                  unsigned long long seed, unsigned int step, void* stream) {
    IDCB_REQUIRE(n_img >= 0 && h > 0 && w > 0 && channels > 0, "synth_patches: bad sizes");
    if (n_img == 0) return 0;
    const int hwc = h * w * channels;
    long long g = ceil_div_ll(n_img * hwc, 256);
    if (g > 16LL * num_sms()) g = 16LL * num_sms();
    IDCB_CHECK(launch_pdl(synth_patches_kernel, (int)g, 256, 0, (cudaStream_t)stream, x, labels, n_img, hwc, channels, seed, step));
    IDCB_LAUNCH_CHECK();
    return 0;
}
