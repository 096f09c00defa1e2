// CPU restatement of the secure-aggregation arithmetic (same Philox header as the device kernel):
// per-client masked vectors in the 2^32 fixed-point ring, their sum, and the dequantised mean.
// Used by the unit tests as the oracle for csrc/comm.cu:secagg_kernel and by the CPU execution path of
// idc_models_b200.fed.secure (reference behaviour being replaced: the Paillier encrypt / mean / decrypt of
// /root/reference/secure_fed_model.py:109-121,160-168).
#include <cmath>
#include <cstdint>
#include <cstring>

#include "../philox.cuh"
#include "host_common.h"
#include "idcb200_host.h"

extern "C" {

// y[k][i] = Q(w[k][i]) + sum_{j>k} PRG(k,j,i) - sum_{j<k} PRG(j,k,i)   (mod 2^32), n multiple of 4
int idcbh_secagg_mask(const float* w, uint32_t* y, int n_clients, long long n, int frac_bits,
                      unsigned long long seed) {
    if (n % 4 != 0 || n_clients < 1 || frac_bits < 1 || frac_bits > 30) {
        idcbh::set_error("secagg_mask: bad arguments");
        return 1;
    }
    const float scale = static_cast<float>(1u << frac_bits);
    for (int k = 0; k < n_clients; ++k) {
        for (long long g = 0; g < n / 4; ++g) {
            uint32_t acc[4];
            for (int e = 0; e < 4; ++e)
                acc[e] = static_cast<uint32_t>(static_cast<int32_t>(std::nearbyintf(w[k * n + g * 4 + e] * scale)));
            for (int j = 0; j < n_clients; ++j) {
                if (j == k) continue;
                const int lo = k < j ? k : j, hi = k < j ? j : k;
                const uint4 m = idcb::philox4x32_10(static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32),
                                                    static_cast<uint32_t>(lo), static_cast<uint32_t>(hi),
                                                    static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
                const uint32_t mm[4] = {m.x, m.y, m.z, m.w};
                for (int e = 0; e < 4; ++e) acc[e] = k < j ? acc[e] + mm[e] : acc[e] - mm[e];
            }
            for (int e = 0; e < 4; ++e) y[k * n + g * 4 + e] = acc[e];
        }
    }
    return 0;
}

// out[i] = dequant( sum_k y[k][i] mod 2^32 ) / n_clients
int idcbh_secagg_unmask_mean(const uint32_t* y, float* out, int n_clients, long long n, int frac_bits) {
    const float inv = 1.0f / (static_cast<float>(1u << frac_bits) * static_cast<float>(n_clients));
    for (long long i = 0; i < n; ++i) {
        uint32_t s = 0;
        for (int k = 0; k < n_clients; ++k) s += y[k * n + i];
        out[i] = static_cast<float>(static_cast<int32_t>(s)) * inv;
    }
    return 0;
}

// raw generator access for tests
void idcbh_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out4) {
    const uint4 r = idcb::philox4x32_10(c0, c1, c2, c3, k0, k1);
    out4[0] = r.x; out4[1] = r.y; out4[2] = r.z; out4[3] = r.w;
}

}  // extern "C"
