// Flat weight-checkpoint files ("cp.ckpt"): the native replacement for the TensorFlow checkpoint
// (.index + .data shards) that tf.keras.callbacks.ModelCheckpoint(save_weights_only=True) writes in the
// reference (/root/reference/fed_model.py:100-105,135,138).
//
// Layout (little endian), safetensors-like:
//   bytes 0..7   magic "IDCBCKP1"
//   bytes 8..15  uint64 header length H
//   bytes 16..   H bytes of JSON: {"tensors":[{"name":..,"dtype":"f32","shape":[..],"offset":o,"nbytes":n},..]}
//   then the raw tensor payloads, each 64-byte aligned, offsets relative to the end of the header.
// Writing goes to "<path>.tmp" followed by rename(), so a crash never leaves a torn checkpoint.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "host_common.h"
#include "idcb200_host.h"

namespace {
struct Entry {
    std::string name;
    std::vector<long long> shape;
    const void* data;
    unsigned long long nbytes;
    unsigned long long offset;
};
struct Writer {
    std::string path;
    std::vector<Entry> entries;
};
std::string json_escape(const std::string& s) {
    std::string o;
    for (char c : s) {
        if (c == '"' || c == '\\') { o += '\\'; o += c; }
        else if (static_cast<unsigned char>(c) < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o += c;
    }
    return o;
}
}  // namespace

extern "C" {

void* idcbh_ckpt_begin(const char* path) {
    Writer* w = new Writer();
    w->path = path;
    return w;
}

int idcbh_ckpt_add_f32(void* h, const char* name, const float* data, const long long* shape, int ndim) {
    Writer* w = static_cast<Writer*>(h);
    Entry e;
    e.name = name;
    unsigned long long n = 1;
    for (int i = 0; i < ndim; ++i) { e.shape.push_back(shape[i]); n *= static_cast<unsigned long long>(shape[i]); }
    e.data = data;
    e.nbytes = n * 4ULL;
    e.offset = 0;
    w->entries.push_back(e);
    return 0;
}

int idcbh_ckpt_end(void* h) {
    Writer* w = static_cast<Writer*>(h);
    unsigned long long off = 0;
    std::string js = "{\"format\":\"idcb-ckpt-1\",\"tensors\":[";
    for (size_t i = 0; i < w->entries.size(); ++i) {
        Entry& e = w->entries[i];
        e.offset = off;
        off += (e.nbytes + 63ULL) & ~63ULL;
        if (i) js += ",";
        js += "{\"name\":\"" + json_escape(e.name) + "\",\"dtype\":\"f32\",\"shape\":[";
        for (size_t d = 0; d < e.shape.size(); ++d) { if (d) js += ","; js += std::to_string(e.shape[d]); }
        js += "],\"offset\":" + std::to_string(e.offset) + ",\"nbytes\":" + std::to_string(e.nbytes) + "}";
    }
    js += "]}";
    while ((16 + js.size()) % 64) js += ' ';
    const std::string tmp = w->path + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) { idcbh::set_error("ckpt: cannot open %s for writing", tmp.c_str()); delete w; return 1; }
    const unsigned long long hl = js.size();
    bool ok = fwrite("IDCBCKP1", 1, 8, f) == 8 && fwrite(&hl, 8, 1, f) == 1 && fwrite(js.data(), 1, js.size(), f) == js.size();
    static const char zeros[64] = {0};
    for (const Entry& e : w->entries) {
        if (!ok) break;
        ok = e.nbytes == 0 || fwrite(e.data, 1, e.nbytes, f) == e.nbytes;
        const unsigned long long pad = ((e.nbytes + 63ULL) & ~63ULL) - e.nbytes;
        if (ok && pad) ok = fwrite(zeros, 1, pad, f) == pad;
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok) { idcbh::set_error("ckpt: short write to %s", tmp.c_str()); remove(tmp.c_str()); delete w; return 1; }
    if (rename(tmp.c_str(), w->path.c_str()) != 0) { idcbh::set_error("ckpt: rename to %s failed", w->path.c_str()); delete w; return 1; }
    delete w;
    return 0;
}

// Reader: returns the JSON header (caller parses it) and reads payload ranges.
long long idcbh_ckpt_header(const char* path, char* buf, long long buflen) {
    FILE* f = fopen(path, "rb");
    if (!f) { idcbh::set_error("ckpt: cannot open %s", path); return -1; }
    char magic[8];
    unsigned long long hl = 0;
    if (fread(magic, 1, 8, f) != 8 || memcmp(magic, "IDCBCKP1", 8) != 0 || fread(&hl, 8, 1, f) != 1) {
        fclose(f); idcbh::set_error("ckpt: %s is not an idcb checkpoint", path); return -1;
    }
    // the length comes from the file: it has to fit in the file (and in a long long) before it sizes anything
    fseek(f, 0, SEEK_END);
    const long fsize = ftell(f);
    fseek(f, 16, SEEK_SET);
    if (fsize < 16 || hl > static_cast<unsigned long long>(fsize - 16)) {
        fclose(f); idcbh::set_error("ckpt: header length of %s exceeds the file", path); return -1;
    }
    if (buf == nullptr || buflen < static_cast<long long>(hl) + 1) { fclose(f); return static_cast<long long>(hl) + 1; }
    if (fread(buf, 1, hl, f) != hl) { fclose(f); idcbh::set_error("ckpt: truncated header in %s", path); return -1; }
    buf[hl] = 0;
    fclose(f);
    return static_cast<long long>(hl);
}

int idcbh_ckpt_read(const char* path, unsigned long long header_len, unsigned long long offset, void* dst,
                    unsigned long long nbytes) {
    FILE* f = fopen(path, "rb");
    if (!f) { idcbh::set_error("ckpt: cannot open %s", path); return 1; }
    const unsigned long long kMax = 1ull << 62;     // keeps the sum below inside a long
    if (header_len > kMax || offset > kMax || nbytes > kMax || dst == nullptr) {
        fclose(f); idcbh::set_error("ckpt: bad read range for %s", path); return 1;
    }
    if (fseek(f, static_cast<long>(16 + header_len + offset), SEEK_SET) != 0 || fread(dst, 1, nbytes, f) != nbytes) {
        fclose(f); idcbh::set_error("ckpt: truncated payload in %s", path); return 1;
    }
    fclose(f);
    return 0;
}

}  // extern "C"
