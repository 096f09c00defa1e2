// Host runtime of the idcb200 CUDA library: error reporting, device queries, TMA descriptor
// encoding through the driver entry point (no link-time libcuda dependency).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "common.cuh"
#include "../include/idcb200.h"

namespace idcb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return 2;
}

void set_pdl(int on);

int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static int g_pdl = -1;   // -1: not decided yet (IDCB_PDL, default on)
bool pdl_enabled() {
    if (g_pdl < 0) {
        const char* e = getenv("IDCB_PDL");
        g_pdl = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return g_pdl != 0;
}
void set_pdl(int on) { g_pdl = on ? 1 : 0; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
                cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

int make_tmap_2d_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                      uint64_t row_pitch_elems, uint32_t box_rows, uint32_t box_cols) {
    EncodeTiledFn fn = get_encode_fn();
    IDCB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
    IDCB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
    IDCB_REQUIRE((row_pitch_elems * 2) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes");
    IDCB_REQUIRE(box_cols * 2 == 128, "TMA box must be 128 bytes wide (SWIZZLE_128B)");
    IDCB_REQUIRE(box_rows >= 1 && box_rows <= 256, "TMA box rows out of range");
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {row_pitch_elems * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim,
                    gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    IDCB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeIm2colFn get_im2col_fn() {
    static EncodeIm2colFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeIm2colFn>(p);
    });
    return fn;
}

// im2col tensor map over an NHWC bf16 activation [B, H, W, pitch] (channel window of `channels` starting at
// `base`): one cp.async.bulk.tensor.4d...im2col instruction then gathers `pixels` consecutive output positions
// (raster order over w, h, n, with the convolution's stride) of ONE filter tap, 64 channels each, into a
// 128-byte-swizzled [pixels][64] tile; positions whose input pixel falls into the zero padding, channels past
// `channels` and images past B arrive as zeros.
// Coordinates of a load: {first channel, wo*stride - pad_l, ho*stride - pad_t, image}, offsets {kw, kh}.
int make_tmap_im2col_bf16(CUtensorMap* map, const void* base, int B, int H, int W, int channels, int pitch,
                          int KH, int KW, int stride, int pad_t, int pad_l, int Ho, int Wo, int pixels) {
    EncodeIm2colFn fn = get_im2col_fn();
    IDCB_REQUIRE(fn != nullptr, "cuTensorMapEncodeIm2col entry point unavailable");
    IDCB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
    IDCB_REQUIRE(pitch % 8 == 0 && stride >= 1 && stride <= 8 && pixels >= 1 && pixels <= 1024,
                 "im2col tensor map: unsupported pitch/stride/box");
    const int pad_b = (Ho - 1) * stride + KH - H - pad_t;     // implied far-side padding (may be negative)
    const int pad_r = (Wo - 1) * stride + KW - W - pad_l;
    int lower[2] = {-pad_l, -pad_t};                          // {W, H} order, like the dimensions after C
    int upper[2] = {pad_r - (KW - 1), pad_b - (KH - 1)};
    for (int i = 0; i < 2; ++i)
        IDCB_REQUIRE(lower[i] >= -128 && lower[i] <= 127 && upper[i] >= -128 && upper[i] <= 127,
                     "im2col tensor map: padding/kernel size out of the 8-bit corner range");
    cuuint64_t gdim[4] = {(cuuint64_t)channels, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstride[3] = {(cuuint64_t)pitch * 2, (cuuint64_t)W * pitch * 2, (cuuint64_t)H * W * pitch * 2};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, lower, upper,
                    64, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    IDCB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeIm2col failed with CUresult %d", (int)r);
    // Drivers up to 13.1 mis-encode im2col maps of tensors smaller than 128 KiB (one descriptor bit must be
    // cleared; the same correction CUTLASS applies in cute/atom/copy_traits_sm90_im2col.hpp).
    int drv = 0;
    if (cudaDriverGetVersion(&drv) == cudaSuccess && drv <= 13010 &&
        (unsigned long long)B * H * W * pitch * 2ull < 131072ull)
        reinterpret_cast<unsigned long long*>(map)[1] &= ~(1ull << 21);
    return 0;
}

}  // namespace idcb

extern "C" const char* idcb_last_error(void) { return idcb::g_err; }

extern "C" int idcb_version(void) { return 100; }

extern "C" int idcb_device_sm_count(void) { return idcb::num_sms(); }

extern "C" int idcb_set_pdl(int on) {
    const int was = idcb::pdl_enabled() ? 1 : 0;
    idcb::set_pdl(on);
    return was;
}
