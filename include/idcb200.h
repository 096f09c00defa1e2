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
/* Programmatic dependent launch of the library's single-GPU kernels (default on, or IDCB_PDL=0):
 * each kernel lets its successor start its set-up early and waits (griddepcontrol.wait) before it
 * touches data. Returns the previous setting. Replaces nothing in the reference (the TF runtime's
 * kernel scheduling, /root/reference/dist_model_tf_dense.py:148). */
int idcb_set_pdl(int on);

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
    int w_pitch, w_off;          /* weight window: elements between consecutive output-channel rows of the weight /
                                    weight-gradient tensor (0 = kh*kw*cin, dense) and element offset of the first
                                    weight used. Lets a kh x kw descriptor address a sub-window of a larger
                                    filter: a 3x3 SAME convolution on a 1x1 image only ever touches its centre tap,
                                    so it runs as a 1x1 convolution over w[:, 1, 1, :] (pitch 9*cin, offset 4*cin). */
} idcb_conv_desc;

/* Activation codes used by prologues/epilogues. */
#define IDCB_ACT_NONE 0
#define IDCB_ACT_RELU 1
#define IDCB_ACT_RELU6 2

/* Fused prologue/epilogue of fprop. All pointers optional (NULL = off).
 *   a   = pro_act(pro_scale[c] * x + pro_shift[c])          (BatchNorm + ReLU folded into the load)
 *   y   = epi_act(epi_scale[n] * acc + epi_shift[n]) + aux  (bias / frozen BN / ReLU / residual)
 *   stat_a[n] += sum_p y,  stat_b[n] += sum_p y^2           (batch statistics for the next BN)   */
/* Training-mode BatchNorm folded into the conv that consumes it: the kernel turns the per-channel
 * batch sums (accumulated by the producer's epilogue) into scale/shift itself, so no separate
 * finalisation launch is needed; CTA (0,0) publishes scale/shift/mean/rstd for the backward pass and
 * applies the Keras moving-statistics update (momentum convention, unbiased variance). */
typedef struct idcb_bn_inline {
    const float* sum;
    const float* sq;
    const float* gamma;
    const float* beta;
    float count, eps, momentum;
    float* scale_out;
    float* shift_out;
    float* mean_out;
    float* rstd_out;
    float* moving_mean;          /* optional (NULL = do not touch) */
    float* moving_var;
} idcb_bn_inline;

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
    const idcb_bn_inline* bn;    /* optional: prologue = BatchNorm(batch statistics) + pro_act */
} idcb_conv_fused;

/* Fused epilogue of dgrad.
 *   mode 0: dx = acc (+ aux, an optional bf16 tensor added in: the residual branch of an Add layer)
 *   mode 1: dx = acc * (aux > 0)                 ReLU backward of the layer that produced aux (= x)
 *   mode 2: dpre = acc * act'(bwd_scale*aux + bwd_shift)    BN+activation backward, aux = raw x
 *           dx (bf16, optional) = dpre ; gacc (fp32, optional) += bwd_scale * dpre
 *           stat_a[c] += sum_p dpre ; stat_b[c] += sum_p dpre * aux
 *   in modes 0/1 stat_a[c] += sum_p dx (bias gradient of the producer layer).                    */
/* Mode 2 only, optional: the backward of the training-mode BatchNorm in front of the activation is finished
 * inside the dgrad launch instead of by idcb_bn_bwd_apply. The kernel adds its column sums to stat_a/stat_b,
 * the grid meets at a barrier (the fp32 tile stays in TMEM), and every CTA writes
 *   out = scale*(dpre - stat_a/N - xhat*dgamma/N),  dgamma = rstd*(stat_b - mean*stat_a)
 * (scale = bwd_scale); the first row of CTAs does dgamma += / dbeta +=. frozen != 0: out = scale*dpre, no
 * barrier. `counter` is one zero-initialised 32-bit word per launch. If the grid cannot be co-resident
 * (more tiles than SMs) the call falls back to the two-launch sequence through dx, which must then be given. */
typedef struct idcb_bn_tail {
    void* out;                   /* bf16, geometry of dx */
    int out_pitch, out_coff;
    const float* mean;
    const float* rstd;
    float count;
    float* dgamma;               /* optional */
    float* dbeta;
    unsigned int* counter;
    int frozen;
} idcb_bn_tail;

/* Mode 2 accumulating into a DenseNet block's gradient buffer G (g_coff = 0), optional: when this dgrad is the last
 * writer of the 32 buffer channels [c0, c0 + 32), their finished gradient
 *   out = G - S1/N - xhat * S2/N        (what idcb_bn_bwd_slice computes in a launch of its own)
 * is produced in the same launch after a grid barrier, with the BatchNorm in front of THIS convolution (bn_mean /
 * bn_rstd, scale = bwd_scale, column sums stat_a / stat_b) folded into S1 / S2 on the fly; one CTA also does that
 * BatchNorm's bookkeeping (dgamma +=, dbeta +=, S1/S2 += for the channels outside the slice). The launch uses
 * 32-wide N tiles and needs ceil(M/128) * ceil(cin/32) CTAs co-resident (2 per SM), else it is refused. */
typedef struct idcb_slice_tail {
    void* out;                   /* bf16 [M, out_pitch], 32 channels at out_coff */
    int out_pitch, out_coff;
    float* acc_s1;               /* S1 / S2 accumulators, indexed by block channel */
    float* acc_s2;
    const float* mean;           /* batch statistics of the block buffer (for xhat), by block channel */
    const float* rstd;
    const float* bn_mean;        /* the BatchNorm in front of this convolution */
    const float* bn_rstd;
    float* dgamma;               /* optional */
    float* dbeta;
    float count;
    unsigned int* counter;       /* one zeroed 32-bit word per launch */
    int c0;
} idcb_slice_tail;

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
    const idcb_bn_tail* tail;    /* optional (NULL = off) */
    const idcb_slice_tail* slice;/* optional (NULL = off) */
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

/* Profiling aid: when dev_ptr is non-null, CTA (0,0) of every following conv launch writes clock64() stamps of its
 * phases (entry, setup done, first tile, producers done, first MMA, last MMA, accumulator ready, ...) into the 16
 * int64 slots it points to (scripts/conv_phases.py). Pass NULL to switch it off. Note that the stamping thread's
 * own global stores delay its next release-arrive, so stamps perturb what they measure by a few microseconds. */
int idcb_conv_set_debug(void* dev_ptr);

/* ---- depthwise 3x3 convolution (direct, bandwidth-bound) ------------------------------------ */
/* TF depthwise kernels behind DepthwiseConv2D in MobileNetV2 (dist_model_tf_mobile.py:119;
 * SURVEY 2.E K7). Weights bf16 [3][3][C]; the descriptor has cin == cout == C.
 *   fprop: y = dw(act(scale*x+shift)), stat_sum/sq += batch statistics of y
 *   dgrad: dpre = dw^T(dy) * act'(scale*x+shift); stat_a += sum dpre, stat_b += sum dpre*x
 *   wgrad: dw (fp32, accumulated) += sum_pos act(scale*x+shift) * dy                          */
int idcb_dwconv_fprop(const idcb_conv_desc* d, const void* x, const void* w, void* y,
                      const float* scale, const float* shift, int act, float* stat_sum,
                      float* stat_sq, void* stream);
int idcb_dwconv_dgrad(const idcb_conv_desc* d, const void* dy, const void* w, const void* x, void* dx,
                      const float* scale, const float* shift, int act, float* stat_a, float* stat_b,
                      void* stream);
int idcb_dwconv_wgrad(const idcb_conv_desc* d, const void* x, const void* dy, float* dw,
                      const float* scale, const float* shift, int act, void* stream);

/* ---- BatchNorm bookkeeping (per-channel, tiny) ---------------------------------------------- */
/* cuDNN BatchNormalizationForwardTraining / Backward as used by Keras BatchNormalization
 * (dist_model_tf_dense.py:131, dist_model_tf_mobile.py:119). The heavy parts
 * (normalise, activation, reductions) are fused into conv prologues/epilogues; these kernels only
 * turn per-channel sums into scale/shift, moving statistics and backward coefficients.           */
int idcb_bn_finalize(const float* sum, const float* sq, float count, const float* gamma,
                     const float* beta, float eps, float* scale, float* shift, float* mean_out,
                     float* rstd_out, float* moving_mean, float* moving_var, float momentum, int C,
                     void* stream);
/* inference-mode (frozen layer) fold of moving statistics into scale/shift */
int idcb_bn_fold(const float* gamma, const float* beta, const float* mean, const float* var, float eps,
                 float* scale, float* shift, int C, void* stream);
int idcb_bn_bwd_coeffs(const float* sa, const float* sb, const float* mean, const float* rstd,
                       const float* scale, float count, float* dgamma, float* dbeta, float* A,
                       float* B, float* C0, float* acc_s1, float* acc_s2, int C, void* stream);
int idcb_bn_shared_coeffs(const float* s1, const float* s2, const float* mean, const float* rstd,
                          float count, float* A, float* B, float* C0, int C, void* stream);
/* y = act(scale*x+shift) (+res)  — BN apply, residual Add of MobileNetV2 */
int idcb_affine_act(const void* x, int x_pitch, int x_coff, void* y, int y_pitch, int y_coff,
                    const float* scale, const float* shift, int act, const void* res, int r_pitch,
                    int r_coff, long long M, int C, void* stream);
/* out = A[c]*g + B[c]*x + C0[c]  — BN backward applied element-wise; g fp32 or bf16 */
int idcb_affine2(const void* g, int g_is_float, int g_pitch, int g_coff, const void* x, int x_pitch,
                 int x_coff, void* out, int o_pitch, int o_coff, const float* A, const float* B,
                 const float* C0, long long M, int C, void* stream);
/* bn_bwd_coeffs / bn_shared_coeffs + affine2 in ONE launch (coefficients derived per thread):
 *   mode 0: training BN backward  out = scale*(g - sa/N - xhat*dgamma/N), emits dgamma += , dbeta +=
 *   mode 1: frozen BN backward    out = scale*g
 *   mode 2: DenseNet shared slice out = g - S1/N - xhat*S2/N   (sa = S1, sb = S2)              */
int idcb_bn_bwd_apply(int mode, const void* g, int g_is_float, int g_pitch, int g_coff, const void* x,
                      int x_pitch, int x_coff, void* out, int o_pitch, int o_coff, const float* sa,
                      const float* sb, const float* mean, const float* rstd, const float* scale,
                      float count, float* dgamma, float* dbeta, long long M, int C, void* stream);
/* DenseNet shared-buffer slice backward (mode 2 of idcb_bn_bwd_apply) that also finishes the bookkeeping of the
 * block's most recent consumer BatchNorm in the same launch: dgamma/dbeta of that layer and its share of the
 * S1/S2 accumulators (idcb_bn_bwd_coeffs would need a launch of its own). acc_s1/acc_s2 point at channel 0 of
 * the block, the slice is channels [c_lo, c_lo + C); p_* describe the pending BatchNorm over channels [0, p_C)
 * (p_sa == NULL: nothing pending). Replaces TF's FusedBatchNormGradV3 for the DenseNet BN layers
 * (/root/reference/dist_model_tf_dense.py:131). */
int idcb_bn_bwd_slice(const float* g, int g_pitch, int g_coff, const void* x, int x_pitch, int x_coff,
                      void* out, int o_pitch, int o_coff, float* acc_s1, float* acc_s2, int c_lo,
                      const float* mean, const float* rstd, float count, long long M, int C,
                      const float* p_sa, const float* p_sb, const float* p_mean, const float* p_rstd,
                      const float* p_scale, float* p_dgamma, float* p_dbeta, int p_C, void* stream);
int idcb_channel_stats(const void* x, int pitch, int coff, long long M, int C, float* sum, float* sq,
                       void* stream);
/* sum_g[c] += sum_p g, sum_gx[c] += sum_p g*x — BN backward sums for a dpre that is already in memory */
int idcb_channel_stats2(const void* g, int g_pitch, int g_coff, const void* x, int x_pitch, int x_coff,
                        long long M, int C, float* sum_g, float* sum_gx, void* stream);

/* ---- pooling ---------------------------------------------------------------------------------- */
/* cuDNN pooling behind MaxPooling2D / AveragePooling2D (VGG16 block pools
 * dist_model_tf_vgg.py:119, DenseNet pool1 and transition pools dist_model_tf_dense.py:131,
 * secure_fed_model.py:88). `pad` models Keras ZeroPadding2D in front of a VALID pool.            */
typedef struct idcb_pool_desc {
    int batch, h, w, c, ho, wo, k, stride, pad, is_max;
    int x_pitch, x_coff, y_pitch, y_coff;
} idcb_pool_desc;
/* `argmax` (max pooling, optional): one byte per output element [B, Ho, Wo, C] = index kh*k+kw of the first
 * maximum of the window; written by the forward pass and, when given to the backward pass, used instead of
 * comparing re-evaluated activations (an overlapping pool otherwise needs a tie-breaking scan per window). */
int idcb_pool_fwd(const idcb_pool_desc* d, const void* x, void* y, const float* scale,
                  const float* shift, int act, unsigned char* argmax, void* stream);
int idcb_pool_bwd(const idcb_pool_desc* d, const void* x, const void* y, const void* dy, int dy_pitch,
                  int dy_coff, void* dx, int dx_pitch, int dx_coff, const float* scale,
                  const float* shift, int act, int relu_mask, float* stat_a, float* stat_b,
                  const unsigned char* argmax, void* stream);

/* ---- input packing / casts --------------------------------------------------------------------- */
/* convert_image_dtype + channel padding to 8 (dist_model_tf_mobile.py:37-40 produces fp32 [0,1]) */
int idcb_pack_input(const float* x, void* y, long long npos, int cin, int cpad, void* stream);
int idcb_cast_f32_bf16(const float* x, void* y, long long n, void* stream);

/* ---- head: GAP + Dense + loss + accuracy, forward and backward in one kernel ------------------- */
/* dist_model_tf_mobile.py:123-132 (GlobalAveragePooling2D, Dense(1), BinaryCrossentropy(from_logits),
 * metrics=['accuracy']); dist_model_tf_dense.py:135-144 (Dense(10), CategoricalCrossentropy).     */
typedef struct idcb_head_desc {
    int batch, hw, c, k;
    int x_pitch, x_coff;
    const float* pro_scale;   /* optional BN(+activation) applied before the pooling */
    const float* pro_shift;
    int pro_act;
    int loss_kind;            /* 0 = BCE from logits, 1 = softmax cross-entropy (integer labels) */
    int acc_on_prob;          /* 0 = threshold the raw logit (Keras quirk), 1 = threshold sigmoid */
    float threshold;
    int training;             /* 0 = forward only */
    float grad_scale;         /* 1 / global batch */
    int dx_pitch, dx_coff;
    int g_pitch, g_coff;
} idcb_head_desc;
int idcb_head(const idcb_head_desc* d, const void* x, const float* w, const float* bias,
              const float* labels, float* logits, float* acc, float* dw, float* db, void* dx,
              float* gacc, float* stat_a, float* stat_b, void* stream);

/* ---- the 10x10 secure-FL CNN: forward + loss + metrics + backward in one launch ----------------------- */
/* create_model() of secure_fed_model.py:84-98. blockIdx.y = client: several clients
 * (each with its own parameter / gradient block of `param_stride` floats) train in the same launch.
 * offsets6 = element offsets of conv W, conv b, dense W, dense b, out W, out b inside a block; conv W is
 * [32][3][3][conv_cin_stride]. Dropout masks are Philox(seed, step, example, feature): reproducible. */
int idcb_small_cnn_step(const float* x, const float* labels, const float* params, float* grads, float* logits,
                        float* acc, int n_clients, int batch, int param_stride, const int* offsets6,
                        int conv_cin_stride, int training, float grad_scale, float drop1, float drop2,
                        unsigned long long seed, unsigned int step, float threshold, int acc_on_prob,
                        void* stream);
/* The same step with the mini-batch gathered inside the kernel out of each client's resident shard
 * (x_all [clients][n_ex][10][10][3], labels_all [clients][n_ex], idx [clients][idx_stride] int64): replaces the
 * per-step index-gather launches of a tf.data shuffle + batch (secure_fed_model.py:56-62, 122-126). */
int idcb_small_cnn_step_indexed(const float* x_all, const float* labels_all, const long long* idx,
                                long long idx_stride, int n_ex, const float* params, float* grads, float* logits,
                                float* acc, int n_clients, int batch, int param_stride, const int* offsets6,
                                int conv_cin_stride, int training, float grad_scale, float drop1, float drop2,
                                unsigned long long seed, unsigned int step, float threshold, int acc_on_prob,
                                void* stream);
/* Device-side synthetic IDC-like patches: Philox noise + class tint, fp32 [n][h][w][c] in [0,1] and 0/1
 * labels (stands in for the tf.data input pipeline of dist_model_tf_mobile.py:34-65). */
int idcb_synth_patches(float* x, float* labels, long long n_img, int h, int w, int channels,
                       unsigned long long seed, unsigned int step, void* stream);

/* ---- optimizers (flat buffers) ------------------------------------------------------------------ */
#define IDCB_OPT_RMSPROP 0
#define IDCB_OPT_SGD 1
#define IDCB_OPT_ADAM 2
/* Keras RMSprop/SGD/Adam update (dist_model_tf_mobile.py:130, fed_model.py:126,208; SURVEY 2.B B7,
 * 2.E K13) over one contiguous range: updates fp32 master, writes the bf16 shadow, optionally
 * re-zeroes the gradient range. For Adam `lr` must already contain the bias correction.           */
int idcb_optim_step(int kind, float* w, float* g, float* s1, float* s2, void* shadow, long long n,
                    float lr, float p1, float p2, float eps, float grad_scale, int zero_grad,
                    void* stream);

/* ---- fused compute + collective kernels over NVLink peer / multicast memory -------------------- */
/* One process per GPU. Every buffer named `void* const* x` is a HOST array of `world` device pointers:
 * the same symmetric allocation as mapped into this process for rank 0..world-1 (x[rank] is local).
 * `mc_*` are the NVLS multicast aliases of the same allocations (NULL = use peer loads/stores).
 * signal_pads[p] points to >= 128 bytes of zero-initialised uint32 flags in rank p's symmetric memory.
 * use_barrier = 0 runs without cross-rank synchronisation: one process may then emulate the ranks by
 * calling the entry point once per rank, in any order (shards are disjoint) — used by the 1-GPU tests. */
typedef struct idcb_comm_world {
    int world, rank, use_barrier;
    void* signal_pads[8];
} idcb_comm_world;

/* MirroredStrategy's packed NcclAllReduce + per-variable optimizer kernels in one launch
 * (dist_model_tf_mobile.py:115-132 under strategy.scope()):
 * rank r reduces the r-th 1/world of grad[*], scales it, applies the optimizer to its shard of the
 * master weights (s1/s2 = local state, only the shard is touched), writes the new fp32 weights and
 * their bf16 cast to every rank, and re-zeroes the consumed gradient shard on every rank.
 * owner = -1: mirrored (sharded ownership). owner = r: CentralStorageStrategy — rank r alone reduces,
 * updates and broadcasts (dist_model_tf_dense.py:24).
 * scal/scal_out (optional): per-rank 4-float metric slots summed over ranks on the kernel's first
 * barrier (loss / correct / count). */
int idcb_allreduce_optim(const idcb_comm_world* w, void* const* grad, void* const* master,
                         void* const* shadow, void* mc_grad, void* mc_master, void* mc_shadow, float* s1,
                         float* s2, long long n, int kind, float lr, float p1, float p2, float eps,
                         float grad_scale, int owner, void* const* scal, float* scal_out, void* stream);

/* TFF federated averaging (fed_model.py:207-227): clients[p] is rank p's
 * [k_local][stride] block of trained client models, weights[p] its [k_local] example counts.
 *   global[offset:offset+n] += server_lr * sum_k n_k (w_k - global) / sum_k n_k        on every rank
 * idcb_fedavg_finite_flags zeroes the weight of a local client whose range holds a non-finite value
 * (TFF drops such a client's update). */
/* BatchNorm moving statistics, MEAN over replicas on read (tf.distribute ON_READ variables under
 * MirroredStrategy, dist_model_tf_mobile.py:115-129): stats[p][0..n) of every rank is
 * replaced by the element-wise mean over ranks. n % 4 == 0. */
int idcb_stat_mean(const idcb_comm_world* w, void* const* stats, void* mc_stats, long long n, void* stream);
/* The same owner-computes kernel with an explicit factor: every rank's `stats` <- scale * sum over ranks.
 * scale = 1 is the sum of a few metric scalars (loss sum, correct, count) over replicas / client groups
 * without a library collective (the per-replica metric reduction of Keras fit/evaluate,
 * /root/reference/dist_model_tf_mobile.py:136, and TFF's federated metric aggregation, fed_model.py:228). */
int idcb_stat_reduce(const idcb_comm_world* w, void* const* stats, void* mc_stats, long long n, float scale,
                     void* stream);
/* Cap on the CTAs used by the communication kernels (0 = one per SM, the default). Used for gradient buckets
 * that are reduced while the backward pass is still running (SURVEY 8.1 "tile by tile as buckets become ready"). */
int idcb_comm_set_max_ctas(int max_ctas);
int idcb_fedavg_finite_flags(const float* clients, long long stride, long long offset, long long n,
                             int k_local, float* weights, void* stream);
int idcb_fedavg_reduce(const idcb_comm_world* w, void* const* clients, void* const* weights,
                       void* const* global, void* const* shadow, void* mc_global, void* mc_shadow,
                       int k_local, long long stride, long long offset, long long n, float server_lr,
                       void* stream);
/* A whole FedAvg aggregation in ONE launch (the state update of TFF's `iterative_process.next`,
 * /root/reference/fed_model.py:219): the non-finite-client scan of this rank's clients over every range, the
 * example-weighted delta mean of every range (sharded over the ranks, multicast into every rank's symmetric
 * `global` / bf16 `shadow`), and this rank's private copy `out` of the complete new model (`base` is this round's
 * server model, read locally; positions outside the ranges are copied from it). At most 16 ascending, disjoint,
 * 4-aligned ranges. */
int idcb_fedavg_round(const idcb_comm_world* w, void* const* clients, void* const* weights,
                      void* const* global, void* const* shadow, void* mc_global, void* mc_shadow,
                      int k_local, long long stride, int n_ranges, const long long* offsets,
                      const long long* lengths, float server_lr, const float* base, float* out,
                      long long total, void* stream);

/* Secure aggregation (stands in for the Paillier path of secure_fed_model.py:109-121,160-168; SURVEY
 * B20, C9/C10, 8.6): pairwise Philox4x32-10 masks in the 2^32 fixed-point ring, generated in registers.
 *   out[offset:offset+n] = mean over all world*k_local clients of clients[.][k][offset:offset+n]
 * ybuf[p] = rank p's [n] uint32 staging buffer for its masked partial sum. Requires
 * |sum_k w_k| * 2^frac_bits < 2^31. */
int idcb_secagg_masked_sum(const idcb_comm_world* w, void* const* clients, void* const* ybuf, void* mc_ybuf,
                           void* const* out, void* mc_out, int k_local, long long stride, long long offset,
                           long long n, int frac_bits, unsigned long long session_seed, void* stream);

/* Single-GPU secure mean that allocates nothing (this entry point has no counterpart in the reference).
 * Same arithmetic as idcb_secagg_masked_sum with world = 1, but the staging memory it needs between
 * its two phases (the masked ring sums and the grid-barrier flags) lives inside the library, in fixed
 * device storage of idcb_secagg_noalloc_max_elems() uint32 words (4 MiB) per GPU, instead of being
 * passed in. It is meant for callers that must not allocate device memory at this point: a service that
 * has given the whole GPU to its own arena, a process under a memory quota fixed at start-up, or code
 * that cannot tolerate the implicit device synchronisation of cudaMalloc/cudaFree.
 *   out[offset:offset+n] = mean over the k_clients rows of clients[k][offset:offset+n]
 * - n above the limit is refused with an error (use the caller-buffer entry point instead); so are a
 *   range that does not fit in a row (offset + n > stride) and pointers that are not 16-byte aligned;
 * - the first call on each device looks the storage up (cudaGetSymbolAddress, serialised by a mutex) and
 *   caches the address per device ordinal: make that first call outside any stream capture, and do not
 *   call again on a device after cudaDeviceReset (the cached address would be stale);
 * - every call on a GPU reuses the same storage: calls whose kernels could overlap on the device
 *   (different streams) race with each other and must be serialised by the caller;
 * - with ONE GPU no pairwise mask is ever generated (masks exist only between clients on different GPUs;
 *   pairs on one GPU cancel and are skipped), so the staging area holds the plain fixed-point ring sum:
 *   this call reproduces the masked sum's arithmetic, it hides nothing from the GPU that runs it;
 * - the kernel contains a grid-wide barrier and is launched cooperatively: if its grid cannot be
 *   co-resident on the device the launch fails with an error instead of waiting. */
long long idcb_secagg_noalloc_max_elems(void);
int idcb_secagg_mean_noalloc(const float* clients, float* out, int k_clients, long long stride,
                             long long offset, long long n, int frac_bits, unsigned long long session_seed,
                             void* stream);

#ifdef __cplusplus
}
#endif
#endif /* IDCB200_H_ */
