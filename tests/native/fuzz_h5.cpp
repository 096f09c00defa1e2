// Sanitizer harness for the HDF5 subset reader and the checkpoint reader (csrc/host/hdf5.cpp, ckpt.cpp): every prefix
// (strided) of each input file, single-bit flips and random byte damage; every mutated file is opened, listed, and
// every dataset the listing names is read. Built by tests/test_cpu_io.py with -fsanitize=address,undefined: a parser
// that trusts a length or an offset from the file shows up as a sanitizer abort.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "idcb200_host.h"

static long g_opened = 0, g_reads = 0;

static void exercise(const char* path) {
    void* h = idcbh_h5r_open(path);
    if (h) {
        ++g_opened;
        const long long need = idcbh_h5r_list(h, nullptr, 0);
        if (need > 0 && need < (64 << 20)) {
            std::vector<char> js(static_cast<size_t>(need) + 1);
            idcbh_h5r_list(h, js.data(), need + 1);
            const std::string s(js.data());
            // {"path":"...","dims":[a,b,...],
            for (size_t p = s.find("{\"path\":\""); p != std::string::npos; p = s.find("{\"path\":\"", p + 1)) {
                const size_t q = s.find('"', p + 9);
                const size_t d0 = s.find("\"dims\":[", q);
                const size_t d1 = s.find(']', d0);
                if (q == std::string::npos || d0 == std::string::npos || d1 == std::string::npos) break;
                const std::string name = s.substr(p + 9, q - (p + 9));
                if (name.find('\\') != std::string::npos) continue;       // escaped names: not worth un-escaping here
                unsigned long long cnt = 1;
                bool any = false, huge = false;
                const char* c = s.c_str() + d0 + 8;
                while (c < s.c_str() + d1) {
                    char* e;
                    const unsigned long long v = std::strtoull(c, &e, 10);
                    if (e == c) break;
                    any = true;
                    if (v != 0 && cnt > (1ull << 22) / v) huge = true; else cnt *= v;
                    c = (*e == ',') ? e + 1 : e;
                }
                (void)any;
                if (huge || cnt > (1ull << 22)) continue;
                std::vector<float> dst(static_cast<size_t>(cnt ? cnt : 1));
                idcbh_h5r_read_f32(h, name.c_str(), dst.data(), static_cast<long long>(cnt));
                ++g_reads;
            }
        }
        idcbh_h5r_close(h);
    }
    // the same bytes through the checkpoint reader
    std::vector<char> hdr(1 << 16);
    const long long hl = idcbh_ckpt_header(path, hdr.data(), static_cast<long long>(hdr.size()));
    if (hl > 0) {
        std::vector<float> dst(1024);
        idcbh_ckpt_read(path, static_cast<unsigned long long>(hl), 0, dst.data(), 4096);
    }
}

static void write_file(const char* path, const std::vector<uint8_t>& d, size_t n) {
    FILE* f = std::fopen(path, "wb");
    if (!f) std::exit(2);
    if (n) std::fwrite(d.data(), 1, n, f);
    std::fclose(f);
}

int main(int argc, char** argv) {
    if (argc < 4) return 2;
    const int stride = std::atoi(argv[1]);
    const char* tmp = argv[2];
    long total = 0;
    for (int a = 3; a < argc; ++a) {
        FILE* f = std::fopen(argv[a], "rb");
        if (!f) return 2;
        std::fseek(f, 0, SEEK_END);
        const long n = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        std::vector<uint8_t> d(static_cast<size_t>(n));
        if (std::fread(d.data(), 1, d.size(), f) != d.size()) return 2;
        std::fclose(f);
        for (long cut = 0; cut <= n; cut += stride * 7) { write_file(tmp, d, static_cast<size_t>(cut)); exercise(tmp); ++total; }
        for (long pos = 0; pos < n; pos += stride) {
            std::vector<uint8_t> e(d);
            e[static_cast<size_t>(pos)] ^= static_cast<uint8_t>(1u << (pos % 8));
            write_file(tmp, e, e.size());
            exercise(tmp);
            ++total;
            e[static_cast<size_t>(pos)] = 0xFF;                            // and a saturated byte (lengths, offsets, counts)
            write_file(tmp, e, e.size());
            exercise(tmp);
            ++total;
        }
        std::srand(11);
        for (int t = 0; t < 2000 / stride; ++t) {
            std::vector<uint8_t> e(d);
            const int k = 1 + std::rand() % 6;
            for (int j = 0; j < k; ++j) e[static_cast<size_t>(std::rand() % n)] = static_cast<uint8_t>(std::rand());
            write_file(tmp, e, e.size());
            exercise(tmp);
            ++total;
        }
    }
    std::printf("files %ld opened %ld reads %ld\n", total, g_opened, g_reads);
    return 0;
}
