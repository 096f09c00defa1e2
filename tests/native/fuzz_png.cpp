// Sanitizer harness for the native PNG decoder (csrc/host/png.cpp): every prefix of each input file, every single-bit
// flip, and random byte damage behind the header, each from an exact-size heap copy so that AddressSanitizer sees any
// read past the end of the data. Built and run by tests/test_cpu_png.py with -fsanitize=address,undefined; a decoder
// bug shows up as a sanitizer abort (non-zero exit), not as a wrong count.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "idcb200_host.h"

static int decode_copy(const std::vector<uint8_t>& d, long n, long flip_pos, int flip_bit, std::vector<float>& out) {
    uint8_t* p = static_cast<uint8_t*>(std::malloc(n > 0 ? n : 1));
    if (n > 0) std::memcpy(p, d.data(), static_cast<size_t>(n));
    if (flip_pos >= 0) p[flip_pos] ^= static_cast<uint8_t>(1u << flip_bit);
    const int rc = idcbh_png_decode_f32(p, n > 0 ? n : 1, out.data(), 50, 50, nullptr, nullptr);
    std::free(p);
    return rc;
}

int main(int argc, char** argv) {
    long total = 0, decoded = 0;
    const int stride = argc > 1 ? std::atoi(argv[1]) : 1;      // 1 = exhaustive
    std::vector<float> out(50 * 50 * 3);
    for (int a = 2; a < argc; ++a) {
        FILE* f = std::fopen(argv[a], "rb");
        if (!f) return 2;
        std::fseek(f, 0, SEEK_END);
        const long n = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        std::vector<uint8_t> d(static_cast<size_t>(n));
        if (std::fread(d.data(), 1, d.size(), f) != d.size()) return 2;
        std::fclose(f);
        for (long cut = 0; cut <= n; cut += stride) { decoded += decode_copy(d, cut, -1, 0, out) == 0; ++total; }
        for (long pos = 8; pos < n; pos += stride)
            for (int bit = 0; bit < 8; ++bit) { decoded += decode_copy(d, n, pos, bit, out) == 0; ++total; }
        std::srand(7);
        for (int t = 0; t < 3000 / stride; ++t) {
            std::vector<uint8_t> e(d);
            const int k = 1 + std::rand() % 8;
            for (int j = 0; j < k; ++j) e[33 + std::rand() % (n - 33)] = static_cast<uint8_t>(std::rand());
            decoded += decode_copy(e, n, -1, 0, out) == 0;
            ++total;
        }
    }
    std::printf("calls %ld decoded %ld\n", total, decoded);
    return 0;
}
