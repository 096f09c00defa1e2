// Philox4x32-10 counter-based generator (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3").
// One header, compiled for the device (secure-aggregation masks generated in registers) and for the
// host (csrc/host/secagg_ref.cpp: the CPU restatement the unit tests use as the oracle), so both sides
// produce bit-identical masks.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define IDCB_HD __host__ __device__ __forceinline__
#include <vector_types.h>
#else
#define IDCB_HD inline
struct uint4 {
    uint32_t x, y, z, w;
};
#endif

namespace idcb {

IDCB_HD void philox_mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
    const uint64_t p = static_cast<uint64_t>(a) * static_cast<uint64_t>(b);
    hi = static_cast<uint32_t>(p >> 32);
    lo = static_cast<uint32_t>(p);
}

IDCB_HD uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#ifdef __CUDACC__
#pragma unroll
#endif
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        philox_mulhilo(M0, c0, hi0, lo0);
        philox_mulhilo(M1, c2, hi1, lo1);
        const uint32_t n0 = hi1 ^ c1 ^ k0;
        const uint32_t n1 = lo1;
        const uint32_t n2 = hi0 ^ c3 ^ k1;
        const uint32_t n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0;
        k1 += W1;
    }
    uint4 out;
    out.x = c0; out.y = c1; out.z = c2; out.w = c3;
    return out;
}

}  // namespace idcb
