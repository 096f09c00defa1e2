// PNG decode + float conversion + bilinear resize for the input pipeline, in native threads.
//
// The reference decodes every patch with TensorFlow's C++ kernels inside tf.data:
//   tf.io.read_file -> tf.image.decode_png(channels=3) -> convert_image_dtype(float32) -> tf.image.resize([H, W])
// (/root/reference/dist_model_tf_mobile.py:37-45, fed_model.py:69-78, secure_fed_model.py:54-63; SURVEY.md 2.A A3,
// 2.C "libpng/zlib"). This file is that stage without Python in the loop: a self-contained DEFLATE decoder (RFC 1951
// stored / fixed / dynamic blocks), the zlib and PNG containers, the five PNG row filters, conversion of every
// non-interlaced 1/2/4/8-bit colour type to 8-bit RGB (alpha dropped, as decode_png(channels=3) does), division by
// 255 and a bilinear resize with half-pixel centres (no antialias). A file list is decoded by a pool of threads
// straight into the caller's [n][H][W][3] float32 array (typically the pinned cache the loader batches from).
//
// What is NOT handled is reported per file (status 1) so the caller can fall back to a general decoder:
// Adam7-interlaced and 16-bit images. Corrupt / unreadable files are status 2 with the reason in idcbh_last_error().
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "host_common.h"
#include "idcb200_host.h"

namespace {

// ---------------------------------------------------------------------------------------------------------
// DEFLATE
// ---------------------------------------------------------------------------------------------------------
struct BitReader {
    const uint8_t* p;
    size_t n, pos = 0;
    uint64_t acc = 0;
    int cnt = 0;
    bool overrun = false;      // bits were CONSUMED past the end (looking ahead past the end is normal)
    void fill() {
        while (cnt <= 56 && pos < n) {
            acc |= static_cast<uint64_t>(p[pos++]) << cnt;
            cnt += 8;
        }
    }
    uint32_t peek(int bits) {   // bits past the end of the data read as zero
        if (cnt < bits) fill();
        return static_cast<uint32_t>(acc & ((1ull << bits) - 1));
    }
    void drop(int bits) {
        if (bits > cnt) { overrun = true; acc = 0; cnt = 0; return; }
        acc >>= bits;
        cnt -= bits;
    }
    uint32_t take(int bits) {
        if (bits == 0) return 0;
        const uint32_t v = peek(bits);
        drop(bits);
        return v;
    }
    void align_byte() { drop(cnt & 7); }
};

// canonical Huffman code, decoded with one table lookup of kFast bits and a slow path for longer codes
struct Huffman {
    static constexpr int kFast = 9;
    uint16_t fast[1 << kFast];      // (length << 12) | symbol, 0 = longer than kFast
    uint16_t count[16];
    uint16_t symbol[320];
    bool build(const uint8_t* lengths, int n) {
        std::memset(count, 0, sizeof(count));
        for (int i = 0; i < n; ++i) count[lengths[i]]++;
        count[0] = 0;
        int left = 1;
        for (int len = 1; len < 16; ++len) {
            left <<= 1;
            left -= count[len];
            if (left < 0) return false;          // over-subscribed
        }
        uint16_t offs[16];
        offs[1] = 0;
        for (int len = 1; len < 15; ++len) offs[len + 1] = offs[len] + count[len];
        for (int i = 0; i < n; ++i)
            if (lengths[i]) symbol[offs[lengths[i]]++] = static_cast<uint16_t>(i);
        // fast table: walk the canonical codes in order
        std::memset(fast, 0, sizeof(fast));
        int code = 0, idx = 0;
        for (int len = 1; len <= kFast; ++len) {
            for (int k = 0; k < count[len]; ++k, ++idx, ++code) {
                // DEFLATE packs Huffman codes MSB first into an LSB-first bit stream: reverse the code
                int rev = 0;
                for (int b = 0; b < len; ++b) rev |= ((code >> b) & 1) << (len - 1 - b);
                for (int fill = rev; fill < (1 << kFast); fill += 1 << len)
                    fast[fill] = static_cast<uint16_t>((len << 12) | symbol[idx]);
            }
            code <<= 1;
        }
        return true;
    }
    int decode(BitReader& br) const {
        const uint32_t look = br.peek(15);
        const uint16_t f = fast[look & ((1 << kFast) - 1)];
        if (f) {
            br.drop(f >> 12);
            return f & 0x0fff;
        }
        int code = 0, first = 0, index = 0;
        for (int len = 1; len < 16; ++len) {
            code |= (look >> (len - 1)) & 1;
            const int c = count[len];
            if (code - c < first) {
                br.drop(len);
                return symbol[index + (code - first)];
            }
            index += c;
            first += c;
            first <<= 1;
            code <<= 1;
        }
        return -1;
    }
};

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115,
                               131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537,
                                2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12,
                                13, 13};

// inflate a raw DEFLATE stream into exactly `want` bytes (the PNG header fixes the size). Returns nullptr or a reason.
const char* inflate_raw(const uint8_t* src, size_t n, uint8_t* out, size_t want) {
    BitReader br{src, n};
    size_t o = 0;
    Huffman lit, dist;
    for (;;) {
        const uint32_t last = br.take(1), type = br.take(2);
        if (type == 0) {
            br.align_byte();
            const uint32_t len = br.take(16), nlen = br.take(16);
            if ((len ^ 0xffffu) != nlen) return "stored block length check failed";
            if (o + len > want) return "more data than the image holds";
            for (uint32_t i = 0; i < len; ++i) out[o++] = static_cast<uint8_t>(br.take(8));
        } else if (type == 1 || type == 2) {
            uint8_t lengths[320];
            if (type == 1) {
                for (int i = 0; i < 144; ++i) lengths[i] = 8;
                for (int i = 144; i < 256; ++i) lengths[i] = 9;
                for (int i = 256; i < 280; ++i) lengths[i] = 7;
                for (int i = 280; i < 288; ++i) lengths[i] = 8;
                lit.build(lengths, 288);
                for (int i = 0; i < 30; ++i) lengths[i] = 5;
                dist.build(lengths, 30);
            } else {
                const int hlit = br.take(5) + 257, hdist = br.take(5) + 1, hclen = br.take(4) + 4;
                if (hlit > 286 || hdist > 30) return "bad code counts in a dynamic block";
                static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
                uint8_t cl[19] = {0};
                for (int i = 0; i < hclen; ++i) cl[order[i]] = static_cast<uint8_t>(br.take(3));
                Huffman clh;
                if (!clh.build(cl, 19)) return "bad code-length code";
                int i = 0;
                while (i < hlit + hdist) {
                    const int sym = clh.decode(br);
                    if (sym < 0) return "bad code-length symbol";
                    if (sym < 16) {
                        lengths[i++] = static_cast<uint8_t>(sym);
                    } else {
                        int rep, val = 0;
                        if (sym == 16) {
                            if (i == 0) return "repeat with no previous length";
                            val = lengths[i - 1];
                            rep = 3 + br.take(2);
                        } else if (sym == 17) {
                            rep = 3 + br.take(3);
                        } else {
                            rep = 11 + br.take(7);
                        }
                        if (i + rep > hlit + hdist) return "code lengths overflow";
                        while (rep--) lengths[i++] = static_cast<uint8_t>(val);
                    }
                }
                if (lengths[256] == 0) return "no end-of-block code";
                if (!lit.build(lengths, hlit)) return "bad literal/length code";
                if (!dist.build(lengths + hlit, hdist)) return "bad distance code";
            }
            for (;;) {
                const int sym = lit.decode(br);
                if (sym < 0) return "bad literal/length symbol";
                if (sym < 256) {
                    if (o >= want) return "more data than the image holds";
                    out[o++] = static_cast<uint8_t>(sym);
                } else if (sym == 256) {
                    break;
                } else {
                    if (sym > 285) return "bad length symbol";
                    const int len = kLenBase[sym - 257] + br.take(kLenExtra[sym - 257]);
                    const int ds = dist.decode(br);
                    if (ds < 0 || ds > 29) return "bad distance symbol";
                    const size_t d = kDistBase[ds] + br.take(kDistExtra[ds]);
                    if (d > o) return "distance reaches before the start of the data";
                    if (o + len > want) return "more data than the image holds";
                    for (int k = 0; k < len; ++k, ++o) out[o] = out[o - d];
                }
                if (br.overrun) return "compressed data ends early";
            }
        } else {
            return "reserved block type";
        }
        if (br.overrun) return "compressed data ends early";
        if (last) break;
    }
    return o == want ? nullptr : "less data than the image holds";
}

// ---------------------------------------------------------------------------------------------------------
// PNG container
// ---------------------------------------------------------------------------------------------------------
inline uint32_t be32(const uint8_t* p) { return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | p[3]; }

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

struct Image {
    int w = 0, h = 0;
    std::vector<uint8_t> rgb;   // [h][w][3]
};

// 0 = decoded, 1 = valid PNG of a kind this decoder leaves to the caller's fallback, 2 = not decodable
int decode_png(const uint8_t* d, size_t n, Image& img, std::string& why) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (n < 8 + 25 || std::memcmp(d, sig, 8) != 0) { why = "not a PNG file"; return 2; }
    size_t pos = 8;
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat, plte;
    bool seen_end = false;
    while (pos + 12 <= n && !seen_end) {
        const uint32_t len = be32(d + pos);
        const uint8_t* tag = d + pos + 4;
        if (len > n - pos - 12) { why = "chunk runs past the end of the file"; return 2; }
        const uint8_t* body = d + pos + 8;
        if (!std::memcmp(tag, "IHDR", 4)) {
            if (len != 13) { why = "bad IHDR"; return 2; }
            img.w = static_cast<int>(be32(body));
            img.h = static_cast<int>(be32(body + 4));
            depth = body[8]; ctype = body[9]; interlace = body[12];
            if (body[10] != 0 || body[11] != 0) { why = "unknown compression / filter method"; return 2; }
        } else if (!std::memcmp(tag, "PLTE", 4)) {
            plte.assign(body, body + len);
        } else if (!std::memcmp(tag, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!std::memcmp(tag, "IEND", 4)) {
            seen_end = true;
        }
        pos += 12 + len;
    }
    if (ctype < 0 || img.w <= 0 || img.h <= 0 || img.w > 16384 || img.h > 16384) { why = "missing or bad IHDR"; return 2; }
    if (interlace != 0) { why = "Adam7 interlaced"; return 1; }
    if (depth == 16) { why = "16-bit samples"; return 1; }
    int channels;
    switch (ctype) {
        case 0: channels = 1; break;
        case 2: channels = 3; break;
        case 3: channels = 1; break;
        case 4: channels = 2; break;
        case 6: channels = 4; break;
        default: why = "unknown colour type"; return 2;
    }
    const bool depth_ok = (ctype == 0 && (depth == 1 || depth == 2 || depth == 4 || depth == 8)) ||
                          (ctype == 3 && (depth == 1 || depth == 2 || depth == 4 || depth == 8)) ||
                          ((ctype == 2 || ctype == 4 || ctype == 6) && depth == 8);
    if (!depth_ok) { why = "bit depth not allowed for the colour type"; return 2; }
    if (ctype == 3 && plte.size() < 3) { why = "palette image without a palette"; return 2; }
    if (idat.size() < 6) { why = "no image data"; return 2; }
    // zlib container: CMF, FLG, [deflate], adler32
    if ((idat[0] & 0x0f) != 8 || ((idat[0] << 8) | idat[1]) % 31 != 0 || (idat[1] & 0x20)) { why = "bad zlib header"; return 2; }
    const int bpp_bits = channels * depth;
    const size_t stride = (static_cast<size_t>(img.w) * bpp_bits + 7) / 8;
    const int bpp = bpp_bits >= 8 ? bpp_bits / 8 : 1;   // filter distance in bytes
    std::vector<uint8_t> raw((stride + 1) * img.h);
    if (const char* e = inflate_raw(idat.data() + 2, idat.size() - 6, raw.data(), raw.size())) { why = e; return 2; }
    // adler32 of the filtered data (cheap, and the only integrity check a decoder can make without the CRC table)
    {
        uint32_t a = 1, b = 0;
        size_t i = 0;
        while (i < raw.size()) {
            const size_t blk = std::min<size_t>(raw.size() - i, 5552);
            for (size_t k = 0; k < blk; ++k) { a += raw[i + k]; b += a; }
            a %= 65521; b %= 65521;
            i += blk;
        }
        if (((b << 16) | a) != be32(idat.data() + idat.size() - 4)) { why = "adler32 mismatch"; return 2; }
    }
    // undo the row filters in place
    for (int y = 0; y < img.h; ++y) {
        uint8_t* row = raw.data() + y * (stride + 1);
        const uint8_t ft = row[0];
        uint8_t* cur = row + 1;
        const uint8_t* up = y ? row - stride : nullptr;     // previous row's pixels (its filter byte is at -stride-1)
        switch (ft) {
            case 0: break;
            case 1: for (size_t x = bpp; x < stride; ++x) cur[x] = static_cast<uint8_t>(cur[x] + cur[x - bpp]); break;
            case 2: if (up) for (size_t x = 0; x < stride; ++x) cur[x] = static_cast<uint8_t>(cur[x] + up[x]); break;
            case 3:
                for (size_t x = 0; x < stride; ++x) {
                    const int a = x >= static_cast<size_t>(bpp) ? cur[x - bpp] : 0, b = up ? up[x] : 0;
                    cur[x] = static_cast<uint8_t>(cur[x] + ((a + b) >> 1));
                }
                break;
            case 4:
                for (size_t x = 0; x < stride; ++x) {
                    const int a = x >= static_cast<size_t>(bpp) ? cur[x - bpp] : 0, b = up ? up[x] : 0,
                              c = (up && x >= static_cast<size_t>(bpp)) ? up[x - bpp] : 0;
                    cur[x] = static_cast<uint8_t>(cur[x] + paeth(a, b, c));
                }
                break;
            default: why = "unknown row filter"; return 2;
        }
    }
    // to 8-bit RGB
    img.rgb.resize(static_cast<size_t>(img.w) * img.h * 3);
    const int maxv = (1 << depth) - 1;
    for (int y = 0; y < img.h; ++y) {
        const uint8_t* cur = raw.data() + y * (stride + 1) + 1;
        uint8_t* o = img.rgb.data() + static_cast<size_t>(y) * img.w * 3;
        for (int x = 0; x < img.w; ++x, o += 3) {
            if (ctype == 2 || ctype == 6) {
                const uint8_t* s = cur + x * channels;
                o[0] = s[0]; o[1] = s[1]; o[2] = s[2];
            } else if (ctype == 4) {
                o[0] = o[1] = o[2] = cur[x * 2];
            } else {
                int v;
                if (depth == 8) {
                    v = cur[x];
                } else {
                    const int per = 8 / depth, shift = (per - 1 - (x % per)) * depth;
                    v = (cur[x / per] >> shift) & maxv;
                }
                if (ctype == 0) {
                    o[0] = o[1] = o[2] = static_cast<uint8_t>(v * 255 / maxv);
                } else {
                    if (static_cast<size_t>(v) * 3 + 2 >= plte.size()) { why = "palette index out of range"; return 2; }
                    o[0] = plte[v * 3]; o[1] = plte[v * 3 + 1]; o[2] = plte[v * 3 + 2];
                }
            }
        }
    }
    return 0;
}

// [h][w][3] uint8 -> [H][W][3] float32 in [0, 1]; bilinear, half-pixel centres, edges clamped, no antialias
// (tf.image.resize's default method; the same arithmetic as the Python path's interpolate(align_corners=False))
void to_float_resized(const Image& img, float* out, int H, int W) {
    if (img.h == H && img.w == W) {
        const size_t n = static_cast<size_t>(H) * W * 3;
        for (size_t i = 0; i < n; ++i) out[i] = img.rgb[i] / 255.0f;
        return;
    }
    const float sy = static_cast<float>(img.h) / H, sx = static_cast<float>(img.w) / W;
    std::vector<int> x0(W), x1(W);
    std::vector<float> lx(W);
    for (int x = 0; x < W; ++x) {
        float s = sx * (x + 0.5f) - 0.5f;
        if (s < 0.f) s = 0.f;
        const int i0 = std::min(static_cast<int>(s), img.w - 1);
        x0[x] = i0; x1[x] = std::min(i0 + 1, img.w - 1); lx[x] = s - i0;
    }
    for (int y = 0; y < H; ++y) {
        float s = sy * (y + 0.5f) - 0.5f;
        if (s < 0.f) s = 0.f;
        const int i0 = std::min(static_cast<int>(s), img.h - 1), i1 = std::min(i0 + 1, img.h - 1);
        const float ly = s - i0;
        const uint8_t* r0 = img.rgb.data() + static_cast<size_t>(i0) * img.w * 3;
        const uint8_t* r1 = img.rgb.data() + static_cast<size_t>(i1) * img.w * 3;
        float* o = out + static_cast<size_t>(y) * W * 3;
        for (int x = 0; x < W; ++x)
            for (int c = 0; c < 3; ++c) {
                const float a = r0[x0[x] * 3 + c] / 255.0f, b = r0[x1[x] * 3 + c] / 255.0f;
                const float cc = r1[x0[x] * 3 + c] / 255.0f, d = r1[x1[x] * 3 + c] / 255.0f;
                const float l1 = lx[x], l0 = 1.0f - l1;
                o[x * 3 + c] = (1.0f - ly) * (l0 * a + l1 * b) + ly * (l0 * cc + l1 * d);
            }
    }
}

bool read_file(const char* path, std::vector<uint8_t>& buf) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (sz < 0) { std::fclose(f); return false; }
    buf.resize(static_cast<size_t>(sz));
    const size_t got = sz ? std::fread(buf.data(), 1, buf.size(), f) : 0;
    std::fclose(f);
    return got == buf.size();
}

}  // namespace

extern "C" {

// Decode one PNG held in memory into out[H][W][3] float32. Returns 0 ok, 1 unsupported kind (use a general decoder),
// 2 corrupt (reason in idcbh_last_error()). src_w / src_h (optional) receive the image's own size.
int idcbh_png_decode_f32(const uint8_t* data, long long len, float* out, int H, int W, int* src_w, int* src_h) {
    if (!data || len <= 0 || !out || H <= 0 || W <= 0) { idcbh::set_error("png_decode: bad arguments"); return 2; }
    Image img;
    std::string why;
    const int rc = decode_png(data, static_cast<size_t>(len), img, why);
    if (src_w) *src_w = img.w;
    if (src_h) *src_h = img.h;
    if (rc != 0) { idcbh::set_error("png_decode: %s", why.c_str()); return rc; }
    to_float_resized(img, out, H, W);
    return 0;
}

// Decode n files with `threads` worker threads into out[n][H][W][3]; status[i] = 0 / 1 / 2 as above (2 also for an
// unreadable file). Returns the number of files whose status is not 0.
int idcbh_png_decode_files(const char* const* paths, long long n, float* out, int H, int W, unsigned char* status,
                           int threads) {
    if (!paths || n < 0 || !out || !status || H <= 0 || W <= 0) { idcbh::set_error("png_decode_files: bad arguments"); return -1; }
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    if (threads > n) threads = static_cast<int>(n > 0 ? n : 1);
    std::atomic<long long> next{0};
    std::atomic<int> bad{0};
    std::mutex err_mu;
    std::string first_err;     // the error string is per thread: hand the first failure back through the caller's
    auto work = [&]() {
        std::vector<uint8_t> buf;
        Image img;
        std::string why;
        for (;;) {
            const long long i = next.fetch_add(1);
            if (i >= n) break;
            int rc = 2;
            if (read_file(paths[i], buf)) rc = decode_png(buf.data(), buf.size(), img, why);
            else why = "cannot read the file";
            if (rc == 0) to_float_resized(img, out + static_cast<size_t>(i) * H * W * 3, H, W);
            else {
                bad.fetch_add(1);
                if (rc == 2) {
                    std::lock_guard<std::mutex> lk(err_mu);
                    if (first_err.empty()) first_err = std::string(paths[i]) + ": " + why;
                }
            }
            status[i] = static_cast<unsigned char>(rc);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    if (!first_err.empty()) idcbh::set_error("png_decode_files: %s", first_err.c_str());
    return bad.load();
}

}  // extern "C"
