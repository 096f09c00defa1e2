/* idcb200 — C ABI of the B200 (sm_100a) kernel library behind idc_models_b200.
 *
 * Conventions
 *   - every entry point returns 0 on success; on failure it returns non-zero and
 *     idcb_last_error() describes the problem (thread-local string).
 *   - activations are NHWC bf16; a tensor is addressed as (base, pitch, coff): element
 *     (position p, channel c) lives at base[p * pitch + coff + c]. This lets kernels read a
 *     channel prefix of, and write a channel slice into, a DenseNet concat buffer in place.
 *   - conv weights are bf16 [Cout][KH][KW][Cin] (OHWI); gradients/optimizer state are fp32.
 *   - `stream` is a cudaStream_t passed as void*; all work is asynchronous on that stream.
 *   - pointers are device pointers unless stated otherwise.
 *
 * Each declaration cites the reference behaviour it stands in for (file:line under
 * /root/reference, or the SURVEY.md section for behaviour that lives in TensorFlow/cuDNN/NCCL).
 */
#ifndef IDCB200_H_
#define IDCB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------ */
const char* idcb_last_error(void);
int idcb_version(void);
int idcb_device_sm_count(void);

/* ---- convolution (tcgen05 implicit GEMM) --------------------------------------------------- */
/* Geometry of the FORWARD convolution; dgrad/wgrad take the same descriptor. */
typedef struct idcb_conv_desc {
    int batch, h, w, cin;        /* input  */
    int ho, wo, cout;            /* output */
    int kh, kw, stride;
    int pad_t, pad_l;            /* zero padding before the first row/column (Keras SAME/VALID and
                                    correct_pad asymmetry are expressed through these + ho/wo) */
    int x_pitch, x_coff;         /* addressing of the input  activation (or of dx in dgrad) */
    int y_pitch, y_coff;         /* addressing of the output activation (or of dy in dgrad/wgrad) */
} idcb_conv_desc;

/* Activation codes used by prologues/epilogues. */
#define IDCB_ACT_NONE 0
#define IDCB_ACT_RELU 1
#define IDCB_ACT_RELU6 2

/* Fused prologue/epilogue of fprop. All pointers optional (NULL = off).
 *   a   = pro_act(pro_scale[c] * x + pro_shift[c])          (BatchNorm + ReLU folded into the load)
 *   y   = epi_act(epi_scale[n] * acc + epi_shift[n]) + aux  (bias / frozen BN / ReLU / residual)
 *   stat_a[n] += sum_p y,  stat_b[n] += sum_p y^2           (batch statistics for the next BN)   */
typedef struct idcb_conv_fused {
    const float* pro_scale;
    const float* pro_shift;
    int pro_act;
    const float* epi_scale;
    const float* epi_shift;
    int epi_act;
    const void* aux;             /* bf16 residual, same geometry as y */
    int aux_pitch, aux_coff;
    float* stat_a;
    float* stat_b;
} idcb_conv_fused;

/* Fused epilogue of dgrad.
 *   mode 0: dx = acc
 *   mode 1: dx = acc * (aux > 0)                 ReLU backward of the layer that produced aux (= x)
 *   mode 2: dpre = acc * act'(bwd_scale*aux + bwd_shift)    BN+activation backward, aux = raw x
 *           dx (bf16, optional) = dpre ; gacc (fp32, optional) += bwd_scale * dpre
 *           stat_a[c] += sum_p dpre ; stat_b[c] += sum_p dpre * aux
 *   in modes 0/1 stat_a[c] += sum_p dx (bias gradient of the producer layer).                    */
typedef struct idcb_conv_bwd_fused {
    int mode;
    const void* aux;
    int aux_pitch, aux_coff;
    const float* bwd_scale;
    const float* bwd_shift;
    int bwd_act;
    float* gacc;
    int g_pitch, g_coff;
    float* stat_a;
    float* stat_b;
} idcb_conv_bwd_fused;

/* cuDNN convolution forward as called by Keras Conv2D (SURVEY 2.E K1/K2/K3/K6;
 * dist_model_tf_vgg.py:119, dist_model_tf_dense.py:131, dist_model_tf_mobile.py:119). */
int idcb_conv_fprop(const idcb_conv_desc* d, const void* x, const void* w, void* y,
                    const idcb_conv_fused* fused, void* stream);
/* cuDNN backward-data (stride 1). */
int idcb_conv_dgrad(const idcb_conv_desc* d, const void* dy, const void* w, void* dx,
                    const idcb_conv_bwd_fused* fused, void* stream);
/* cuDNN backward-filter; dw is fp32 [Cout][KH][KW][Cin] and is ACCUMULATED into (zero it first).
 * The optional prologue recomputes the forward's BN+activation on x while loading it. */
int idcb_conv_wgrad(const idcb_conv_desc* d, const void* x, const void* dy, float* dw,
                    const float* pro_scale, const float* pro_shift, int pro_act, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* IDCB200_H_ */
