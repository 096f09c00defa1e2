// Flat-buffer optimizers: one launch updates a contiguous range of the fp32 master weights from
// the fp32 gradient buffer, refreshes the bf16 shadow the conv kernels read, and re-zeroes the
// gradient range for the next step's accumulating wgrad kernels.
//
// Formulas are Keras' (/root/reference/dist_model_tf_mobile.py:130 RMSprop,
// fed_model.py:208 client RMSprop, TFF server SGD(1.0)):
//   RMSprop : ms = rho*ms + (1-rho)*g^2 ; w -= lr * g / (sqrt(ms) + eps)
//   SGD     : v = mom*v - lr*g ; w += v            (mom = 0 -> w -= lr*g)
//   Adam    : m,v moments ; w -= lr_t * m / (sqrt(v) + eps), lr_t = lr*sqrt(1-b2^t)/(1-b1^t)
// TensorFlow launches one such kernel PER VARIABLE (239 launches for a DenseNet121
// fine-tune step); here it is one.
#include "common.cuh"
#include "../include/idcb200.h"

namespace idcb {

struct OptArgs {
    float* w; float* g; float* s1; float* s2; __nv_bfloat16* shadow;
    long long n;
    int kind;
    float lr, p1, p2, eps, grad_scale;
    int zero_grad;
};

__device__ __forceinline__ float opt_update(int kind, float w, float g, float& s1, float& s2,
                                            const OptArgs& a) {
    if (kind == IDCB_OPT_RMSPROP) {
        s1 = a.p1 * s1 + (1.f - a.p1) * g * g;
        return w - a.lr * g / (sqrtf(s1) + a.eps);
    } else if (kind == IDCB_OPT_SGD) {
        s1 = a.p1 * s1 - a.lr * g;
        return w + s1;
    } else {
        s1 = a.p1 * s1 + (1.f - a.p1) * g;
        s2 = a.p2 * s2 + (1.f - a.p2) * g * g;
        return w - a.lr * s1 / (sqrtf(s2) + a.eps);  // lr already bias-corrected by the host
    }
}

__global__ void optim_kernel(OptArgs a) {
    pdl_enter();
    const long long n4 = a.n >> 2;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
         i += (long long)gridDim.x * blockDim.x) {
        float4 w = reinterpret_cast<float4*>(a.w)[i];
        float4 g = reinterpret_cast<float4*>(a.g)[i];
        float4 s1 = a.s1 ? reinterpret_cast<float4*>(a.s1)[i] : make_float4(0, 0, 0, 0);
        float4 s2 = a.s2 ? reinterpret_cast<float4*>(a.s2)[i] : make_float4(0, 0, 0, 0);
        g.x *= a.grad_scale; g.y *= a.grad_scale; g.z *= a.grad_scale; g.w *= a.grad_scale;
        w.x = opt_update(a.kind, w.x, g.x, s1.x, s2.x, a);
        w.y = opt_update(a.kind, w.y, g.y, s1.y, s2.y, a);
        w.z = opt_update(a.kind, w.z, g.z, s1.z, s2.z, a);
        w.w = opt_update(a.kind, w.w, g.w, s1.w, s2.w, a);
        reinterpret_cast<float4*>(a.w)[i] = w;
        if (a.s1) reinterpret_cast<float4*>(a.s1)[i] = s1;
        if (a.s2) reinterpret_cast<float4*>(a.s2)[i] = s2;
        if (a.shadow) {
            uint2 o;
            o.x = pack_bf16x2(w.x, w.y);
            o.y = pack_bf16x2(w.z, w.w);
            reinterpret_cast<uint2*>(a.shadow)[i] = o;
        }
        if (a.zero_grad) reinterpret_cast<float4*>(a.g)[i] = make_float4(0, 0, 0, 0);
    }
}

}  // namespace idcb

using namespace idcb;

extern "C" int idcb_optim_step(int kind, float* w, float* g, float* s1, float* s2, void* shadow,
                               long long n, float lr, float p1, float p2, float eps, float grad_scale,
                               int zero_grad, void* stream) {
    IDCB_REQUIRE(n % 4 == 0, "optim_step: range length must be a multiple of 4 (got %lld)", n);
    IDCB_REQUIRE(kind >= 0 && kind <= 2, "optim_step: unknown optimizer kind %d", kind);
    IDCB_REQUIRE(((uintptr_t)w & 15) == 0 && ((uintptr_t)g & 15) == 0, "optim_step: alignment");
    if (n == 0) return 0;
    OptArgs a{w, g, s1, s2, (__nv_bfloat16*)shadow, n, kind, lr, p1, p2, eps, grad_scale, zero_grad};
    long long blocks = ceil_div_ll(n / 4, 256);
    if (blocks > 8LL * num_sms()) blocks = 8LL * num_sms();
    IDCB_CHECK(launch_pdl(optim_kernel, (int)blocks, 256, 0, (cudaStream_t)stream, a));
    IDCB_LAUNCH_CHECK();
    return 0;
}
