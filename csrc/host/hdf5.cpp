// A minimal HDF5 reader/writer — just enough of the format to import and export Keras weight files
// (`model.save_weights('x.h5')`, the `*_weights_tf_dim_ordering_tf_kernels_notop.h5` application files)
// without libhdf5 / h5py, neither of which exists on this image. The reference
// reaches HDF5 through `tf.keras.applications(weights='imagenet')` (/root/reference/dist_model_tf_mobile.py:119).
//
// Writer: superblock v0, version-1 object headers, "old style" groups (symbol-table message -> v1 B-tree
// -> one symbol node -> local heap), contiguous float32 datasets, version-1 attributes holding fixed-length
// ASCII strings (scalar or 1-D). That is the layout libhdf5 1.8/1.10 itself writes by default, so files
// produced here open in h5py/Keras.
// Reader: superblock v0/v1 (and v2/v3 headers), v1 and v2 object headers with continuation blocks,
// symbol-table groups and compact link messages, contiguous/compact datasets of float32/float64,
// attributes v1-v3 with fixed-length strings, variable-length strings (global heap) and numeric scalars.
// Anything outside the subset (chunked/compressed datasets, dense link storage) is reported as a clear
// error instead of returning garbage.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "host_common.h"
#include "idcb200_host.h"

namespace {

const uint64_t UNDEF = 0xFFFFFFFFFFFFFFFFULL;

// =====================================================================================================
// Writer
// =====================================================================================================
struct WAttr {
    std::string name;
    std::vector<std::string> values;
    bool scalar;
};
struct WDataset {
    std::vector<long long> dims;
    std::vector<float> data;
};
struct WGroup {
    std::map<std::string, std::unique_ptr<WGroup>> groups;   // std::map keeps names strcmp-sorted
    std::map<std::string, WDataset> datasets;
    std::vector<WAttr> attrs;
};
struct Writer {
    std::string path;
    WGroup root;
    std::vector<uint8_t> buf;
    int leaf_k = 4;
};

void put(std::vector<uint8_t>& b, size_t off, const void* p, size_t n) {
    if (b.size() < off + n) b.resize(off + n, 0);
    memcpy(b.data() + off, p, n);
}
template <typename T>
void put_le(std::vector<uint8_t>& b, size_t off, T v) {
    put(b, off, &v, sizeof(T));
}
size_t alloc(std::vector<uint8_t>& b, size_t n) {
    size_t off = (b.size() + 7) & ~size_t(7);
    b.resize(off + n, 0);
    return off;
}
size_t pad8(size_t n) { return (n + 7) & ~size_t(7); }

std::vector<uint8_t> dtype_f32() {
    std::vector<uint8_t> d(24, 0);
    d[0] = 0x11; d[1] = 0x20; d[2] = 0x1F; d[3] = 0x00;
    uint32_t size = 4; memcpy(&d[4], &size, 4);
    uint16_t bit_off = 0, prec = 32; memcpy(&d[8], &bit_off, 2); memcpy(&d[10], &prec, 2);
    d[12] = 23; d[13] = 8; d[14] = 0; d[15] = 23;
    uint32_t bias = 127; memcpy(&d[16], &bias, 4);
    return d;   // 20 bytes used, padded to 24
}
std::vector<uint8_t> dtype_str(uint32_t len) {
    std::vector<uint8_t> d(8, 0);
    d[0] = 0x13; d[1] = 0x01;   // class 3 (string) v1; null-padded, ASCII
    memcpy(&d[4], &len, 4);
    return d;
}
std::vector<uint8_t> dataspace(const std::vector<long long>& dims) {
    std::vector<uint8_t> d(8 + 8 * dims.size(), 0);
    d[0] = 1; d[1] = static_cast<uint8_t>(dims.size());
    for (size_t i = 0; i < dims.size(); ++i) { uint64_t v = dims[i]; memcpy(&d[8 + 8 * i], &v, 8); }
    return d;
}

struct Msg {
    uint16_t type;
    std::vector<uint8_t> data;
};

Msg attr_msg(const WAttr& a) {
    uint32_t len = 1;
    for (const auto& s : a.values) len = std::max<uint32_t>(len, static_cast<uint32_t>(s.size()));
    std::vector<uint8_t> dt = dtype_str(len);
    std::vector<long long> dims;
    if (!a.scalar) dims.push_back(static_cast<long long>(a.values.size()));
    std::vector<uint8_t> ds = dataspace(dims);
    const size_t name_sz = a.name.size() + 1;
    std::vector<uint8_t> m(8, 0);
    m[0] = 1;
    uint16_t v16 = static_cast<uint16_t>(name_sz); memcpy(&m[2], &v16, 2);
    v16 = static_cast<uint16_t>(dt.size()); memcpy(&m[4], &v16, 2);
    v16 = static_cast<uint16_t>(ds.size()); memcpy(&m[6], &v16, 2);
    size_t off = m.size();
    m.resize(off + pad8(name_sz), 0); memcpy(&m[off], a.name.c_str(), name_sz);
    off = m.size(); m.resize(off + pad8(dt.size()), 0); memcpy(&m[off], dt.data(), dt.size());
    off = m.size(); m.resize(off + pad8(ds.size()), 0); memcpy(&m[off], ds.data(), ds.size());
    off = m.size(); m.resize(off + pad8(size_t(len) * a.values.size()), 0);
    for (size_t i = 0; i < a.values.size(); ++i) memcpy(&m[off + i * len], a.values[i].data(), a.values[i].size());
    return Msg{0x000C, m};
}

size_t write_header(std::vector<uint8_t>& b, const std::vector<Msg>& msgs) {
    size_t body = 0;
    for (const auto& m : msgs) body += 8 + pad8(m.data.size());
    const size_t at = alloc(b, 16 + body);
    b[at] = 1;
    put_le<uint16_t>(b, at + 2, static_cast<uint16_t>(msgs.size()));
    put_le<uint32_t>(b, at + 4, 1);
    put_le<uint32_t>(b, at + 8, static_cast<uint32_t>(body));
    size_t off = at + 16;
    for (const auto& m : msgs) {
        put_le<uint16_t>(b, off, m.type);
        put_le<uint16_t>(b, off + 2, static_cast<uint16_t>(pad8(m.data.size())));
        if (!m.data.empty()) put(b, off + 8, m.data.data(), m.data.size());
        off += 8 + pad8(m.data.size());
    }
    return at;
}

size_t write_dataset(std::vector<uint8_t>& b, const WDataset& d) {
    const size_t nbytes = d.data.size() * 4;
    const size_t data_at = alloc(b, nbytes ? nbytes : 8);
    if (nbytes) put(b, data_at, d.data.data(), nbytes);
    std::vector<Msg> msgs;
    msgs.push_back(Msg{0x0001, dataspace(d.dims)});
    msgs.push_back(Msg{0x0003, dtype_f32()});
    std::vector<uint8_t> fill(8, 0);
    fill[0] = 2; fill[1] = 2; fill[2] = 2; fill[3] = 0;   // v2, alloc late, write if set, undefined
    msgs.push_back(Msg{0x0005, fill});
    std::vector<uint8_t> lay(24, 0);
    lay[0] = 3; lay[1] = 1;
    uint64_t a = data_at, s = nbytes; memcpy(&lay[2], &a, 8); memcpy(&lay[10], &s, 8);
    msgs.push_back(Msg{0x0008, lay});
    return write_header(b, msgs);
}

struct GroupAddr {
    size_t header, btree, heap;
};

GroupAddr write_group(Writer& w, const WGroup& g) {
    std::vector<uint8_t>& b = w.buf;
    // children first (their object-header addresses go into this group's symbol node)
    struct Child { std::string name; size_t header; bool is_group; size_t btree, heap; };
    std::vector<Child> kids;
    for (const auto& kv : g.groups) {
        GroupAddr ga = write_group(w, *kv.second);
        kids.push_back(Child{kv.first, ga.header, true, ga.btree, ga.heap});
    }
    for (const auto& kv : g.datasets) kids.push_back(Child{kv.first, write_dataset(b, kv.second), false, 0, 0});
    std::sort(kids.begin(), kids.end(), [](const Child& a, const Child& c) { return a.name < c.name; });
    // local heap: offset 0 = "", then the names
    std::vector<uint8_t> heap_data(8, 0);
    std::vector<uint64_t> name_off;
    for (const auto& k : kids) {
        name_off.push_back(heap_data.size());
        const size_t n = pad8(k.name.size() + 1);
        const size_t o = heap_data.size();
        heap_data.resize(o + n, 0);
        memcpy(&heap_data[o], k.name.c_str(), k.name.size());
    }
    const size_t free_at = heap_data.size();
    heap_data.resize(free_at + 16, 0);                      // one free block closing the segment
    { uint64_t nxt = 1, sz = 16; memcpy(&heap_data[free_at], &nxt, 8); memcpy(&heap_data[free_at + 8], &sz, 8); }
    const size_t seg_at = alloc(b, heap_data.size());
    put(b, seg_at, heap_data.data(), heap_data.size());
    const size_t heap_at = alloc(b, 32);
    put(b, heap_at, "HEAP", 4);
    put_le<uint64_t>(b, heap_at + 8, heap_data.size());
    put_le<uint64_t>(b, heap_at + 16, free_at);
    put_le<uint64_t>(b, heap_at + 24, seg_at);
    // one symbol node holding every entry (leaf K is sized for the largest group of the file)
    const size_t cap = 2 * static_cast<size_t>(w.leaf_k);
    const size_t snod_at = alloc(b, 8 + 40 * cap);
    put(b, snod_at, "SNOD", 4);
    b[snod_at + 4] = 1;
    put_le<uint16_t>(b, snod_at + 6, static_cast<uint16_t>(kids.size()));
    for (size_t i = 0; i < kids.size(); ++i) {
        const size_t e = snod_at + 8 + 40 * i;
        put_le<uint64_t>(b, e, name_off[i]);
        put_le<uint64_t>(b, e + 8, kids[i].header);
        put_le<uint32_t>(b, e + 16, kids[i].is_group ? 1u : 0u);
        if (kids[i].is_group) { put_le<uint64_t>(b, e + 24, kids[i].btree); put_le<uint64_t>(b, e + 32, kids[i].heap); }
    }
    // B-tree root (leaf level 0) with one child
    const int ik = 16;
    const size_t bt_at = alloc(b, 24 + (2 * ik + 1) * 8 + 2 * ik * 8);
    put(b, bt_at, "TREE", 4);
    b[bt_at + 4] = 0; b[bt_at + 5] = 0;
    put_le<uint16_t>(b, bt_at + 6, kids.empty() ? 0 : 1);
    put_le<uint64_t>(b, bt_at + 8, UNDEF);
    put_le<uint64_t>(b, bt_at + 16, UNDEF);
    put_le<uint64_t>(b, bt_at + 24, 0);                          // key 0: ""
    put_le<uint64_t>(b, bt_at + 32, snod_at);                    // child 0
    put_le<uint64_t>(b, bt_at + 40, kids.empty() ? 0 : name_off.back());   // key 1: largest name
    // object header: symbol table message + attributes
    std::vector<Msg> msgs;
    std::vector<uint8_t> st(16, 0);
    { uint64_t a = bt_at, h = heap_at; memcpy(&st[0], &a, 8); memcpy(&st[8], &h, 8); }
    msgs.push_back(Msg{0x0011, st});
    for (const auto& a : g.attrs) msgs.push_back(attr_msg(a));
    return GroupAddr{write_header(b, msgs), bt_at, heap_at};
}

int max_children(const WGroup& g) {
    int m = static_cast<int>(g.groups.size() + g.datasets.size());
    for (const auto& kv : g.groups) m = std::max(m, max_children(*kv.second));
    return m;
}

WGroup* descend(WGroup& root, const std::string& path, std::string* leaf) {
    WGroup* g = &root;
    size_t pos = 0;
    std::vector<std::string> parts;
    while (pos <= path.size()) {
        size_t nx = path.find('/', pos);
        if (nx == std::string::npos) nx = path.size();
        if (nx > pos) parts.push_back(path.substr(pos, nx - pos));
        pos = nx + 1;
    }
    const size_t n = leaf ? (parts.empty() ? 0 : parts.size() - 1) : parts.size();
    for (size_t i = 0; i < n; ++i) {
        auto& slot = g->groups[parts[i]];
        if (!slot) slot.reset(new WGroup());
        g = slot.get();
    }
    if (leaf) *leaf = parts.empty() ? "" : parts.back();
    return g;
}

// =====================================================================================================
// Reader
// =====================================================================================================
struct RAttr {
    std::string name;
    std::vector<std::string> strs;
    std::vector<double> nums;
    bool is_str = false;
};
struct RDataset {
    std::vector<uint64_t> dims;
    int cls = -1;          // 1 = float
    uint32_t elem = 0;     // element size
    uint64_t addr = UNDEF, size = 0;
    std::vector<uint8_t> compact;
    bool ok = false;
    std::string why;
};
struct Reader {
    std::vector<uint8_t> f;
    int so = 8, sl = 8;
    uint64_t base = 0;
    std::map<std::string, RDataset> datasets;
    std::map<std::string, std::vector<RAttr>> attrs;   // by object path ("" = root)
    std::string err;
    int visited = 0;
};

uint64_t rd(const Reader& r, size_t off, int n) {
    uint64_t v = 0;
    if (off > r.f.size() || static_cast<size_t>(n) > r.f.size() - off) return UNDEF;   // (off + n may wrap)
    memcpy(&v, r.f.data() + off, n);
    if (n == 8) return v;
    return (v == ((uint64_t(1) << (8 * n)) - 1)) ? UNDEF : v;
}
uint64_t rdraw(const Reader& r, size_t off, int n) {
    uint64_t v = 0;
    if (off > r.f.size() || static_cast<size_t>(n) > r.f.size() - off) return 0;
    memcpy(&v, r.f.data() + off, n);
    return v;
}

// Every byte the reader looks at goes through these: an offset or a length taken from the file is never trusted
// (a damaged or hostile .h5 must end in an error message, not in a read outside the buffer).
inline uint8_t u8(const Reader& r, size_t off) { return off < r.f.size() ? r.f[off] : 0; }
inline bool in_file(const Reader& r, size_t off, size_t n) { return off <= r.f.size() && n <= r.f.size() - off; }
inline bool tag_at(const Reader& r, size_t off, const char* tag) { return in_file(r, off, 4) && memcmp(&r.f[off], tag, 4) == 0; }
// at most maxn bytes starting at off, cut at the first NUL and at the end of the file
inline std::string str_at(const Reader& r, size_t off, size_t maxn) {
    if (off >= r.f.size()) return "";
    maxn = std::min(maxn, r.f.size() - off);
    const char* s = reinterpret_cast<const char*>(&r.f[off]);
    return std::string(s, strnlen(s, maxn));
}
constexpr size_t kMaxBlocks = 4096;          // continuation blocks of one object header
constexpr uint64_t kMaxAttrValues = 1 << 20;  // elements of one attribute
constexpr int kMaxObjects = 1 << 20;          // objects visited in one file (cycles in a damaged file end here)

struct Message {
    uint16_t type;
    size_t off, size;
    uint8_t flags;
};

bool collect_messages(Reader& r, uint64_t addr, std::vector<Message>& out) {
    const size_t a = static_cast<size_t>(addr + r.base);
    if (!in_file(r, a, 16)) { r.err = "object header outside the file"; return false; }
    if (tag_at(r, a, "OHDR")) {
        // version 2 object header
        const uint8_t flags = u8(r, a + 5);
        size_t p = a + 6;
        if (flags & 0x20) p += 16;
        if (flags & 0x10) p += 4;
        const int szb = 1 << (flags & 3);
        uint64_t chunk0 = rdraw(r, p, szb);
        p += szb;
        std::vector<std::pair<size_t, size_t>> blocks{{p, static_cast<size_t>(chunk0)}};
        const bool track = flags & 0x04;
        for (size_t bi = 0; bi < blocks.size(); ++bi) {
            if (!in_file(r, blocks[bi].first, blocks[bi].second)) { r.err = "object header block outside the file"; return false; }
            size_t q = blocks[bi].first, end = blocks[bi].first + blocks[bi].second;
            while (q + 4 + (track ? 2 : 0) <= end) {
                const uint8_t type = u8(r, q);
                const uint16_t sz = static_cast<uint16_t>(rdraw(r, q + 1, 2));
                const uint8_t mf = u8(r, q + 3);
                q += 4 + (track ? 2 : 0);
                if (q + sz > end) break;
                if (type == 0x10) {
                    const uint64_t co = rd(r, q, r.so), cl = rdraw(r, q + r.so, r.sl);
                    // continuation block: "OCHK" + messages + checksum
                    if (co == UNDEF || cl < 8 || blocks.size() >= kMaxBlocks) { r.err = "bad object header continuation"; return false; }
                    blocks.push_back({static_cast<size_t>(co + r.base) + 4, static_cast<size_t>(cl) - 8});
                } else if (type != 0) {
                    out.push_back(Message{type, q, sz, mf});
                }
                q += sz;
            }
        }
        return true;
    }
    if (u8(r, a) != 1) { r.err = "unsupported object header version"; return false; }
    const uint16_t nmsg = static_cast<uint16_t>(rdraw(r, a + 2, 2));
    const uint32_t hsize = static_cast<uint32_t>(rdraw(r, a + 8, 4));
    std::vector<std::pair<size_t, size_t>> blocks{{a + 16, hsize}};
    size_t seen = 0;
    for (size_t bi = 0; bi < blocks.size() && seen < nmsg; ++bi) {
        if (!in_file(r, blocks[bi].first, blocks[bi].second)) { r.err = "object header block outside the file"; return false; }
        size_t q = blocks[bi].first, end = blocks[bi].first + blocks[bi].second;
        while (q + 8 <= end && seen < nmsg) {
            const uint16_t type = static_cast<uint16_t>(rdraw(r, q, 2));
            const uint16_t sz = static_cast<uint16_t>(rdraw(r, q + 2, 2));
            const uint8_t mf = u8(r, q + 4);
            q += 8;
            if (!in_file(r, q, sz)) { r.err = "message outside the file"; return false; }
            ++seen;
            if (type == 0x0010) {
                const uint64_t co = rd(r, q, r.so), cl = rdraw(r, q + r.so, r.sl);
                if (co == UNDEF || blocks.size() >= kMaxBlocks) { r.err = "bad object header continuation"; return false; }
                blocks.push_back({static_cast<size_t>(co + r.base), static_cast<size_t>(cl)});
            } else if (type != 0) {
                out.push_back(Message{type, q, sz, mf});
            }
            q += sz;
        }
    }
    return true;
}

struct DType {
    int cls = -1;
    uint32_t size = 0;
    bool vlen_str = false;
    size_t consumed = 0;
};
DType parse_dtype(const Reader& r, size_t off) {
    DType d;
    d.cls = u8(r, off) & 0x0F;
    d.size = static_cast<uint32_t>(rdraw(r, off + 4, 4));
    d.consumed = 8;
    if (d.cls == 0) d.consumed += 4;
    else if (d.cls == 1) d.consumed += 12;
    else if (d.cls == 9) {   // variable length: bit field type 1 = string
        d.vlen_str = (u8(r, off + 1) & 0x0F) == 1;
        d.consumed += 8 + 4;  // base type (char) is an 8+4 byte integer type; not needed further
    }
    return d;
}
size_t parse_dataspace(const Reader& r, size_t off, std::vector<uint64_t>& dims) {
    const uint8_t ver = u8(r, off), rank = u8(r, off + 1), flags = u8(r, off + 2);
    size_t p = off + (ver == 1 ? 8 : 4);
    dims.clear();
    for (int i = 0; i < rank; ++i) dims.push_back(rdraw(r, p + size_t(r.sl) * i, r.sl));
    p += size_t(r.sl) * rank;
    if (flags & 1) p += size_t(r.sl) * rank;
    return p - off;
}

std::string global_heap_string(const Reader& r, uint64_t coll, uint32_t index) {
    const size_t a = static_cast<size_t>(coll + r.base);
    if (!in_file(r, a, 16) || !tag_at(r, a, "GCOL")) return "";
    const uint64_t csize = rdraw(r, a + 8, r.sl);
    size_t q = a + 8 + r.sl;
    const size_t end = a + static_cast<size_t>(csize);
    while (q + 8 + r.sl <= end && q + 8 + r.sl <= r.f.size()) {
        const uint16_t idx = static_cast<uint16_t>(rdraw(r, q, 2));
        const uint64_t osz = rdraw(r, q + 8, r.sl);
        const size_t data = q + 8 + r.sl;
        if (idx == 0) break;
        if (idx == index) {
            if (!in_file(r, data, static_cast<size_t>(osz)) || osz > (1u << 24)) return "";
            return std::string(reinterpret_cast<const char*>(&r.f[data]), static_cast<size_t>(osz));
        }
        if (osz > r.f.size()) return "";
        q = data + ((static_cast<size_t>(osz) + 7) & ~size_t(7));
    }
    return "";
}

void parse_attr(Reader& r, const Message& m, std::vector<RAttr>& out) {
    const size_t o = m.off;
    const uint8_t ver = u8(r, o);
    if (ver < 1 || ver > 3) return;
    const uint16_t nsz = static_cast<uint16_t>(rdraw(r, o + 2, 2));
    const uint16_t dtsz = static_cast<uint16_t>(rdraw(r, o + 4, 2));
    const uint16_t dssz = static_cast<uint16_t>(rdraw(r, o + 6, 2));
    size_t p = o + 8 + (ver == 3 ? 1 : 0);
    auto adv = [&](size_t n) { return ver == 1 ? pad8(n) : n; };
    RAttr a;
    a.name = str_at(r, p, nsz);
    p += adv(nsz);
    const DType dt = parse_dtype(r, p);
    p += adv(dtsz);
    std::vector<uint64_t> dims;
    parse_dataspace(r, p, dims);
    p += adv(dssz);
    uint64_t count = 1;
    for (uint64_t d : dims) {
        if (d != 0 && count > kMaxAttrValues / d) return;      // not an attribute a weight file carries
        count *= d;
    }
    if (dt.cls == 3) {
        a.is_str = true;
        if (dt.size == 0 || !in_file(r, p, static_cast<size_t>(count) * dt.size)) return;
        for (uint64_t i = 0; i < count; ++i) a.strs.push_back(str_at(r, p + i * dt.size, dt.size));
    } else if (dt.cls == 9 && dt.vlen_str) {
        a.is_str = true;
        if (!in_file(r, p, static_cast<size_t>(count) * (4 + r.so + 4))) return;
        for (uint64_t i = 0; i < count; ++i) {
            const size_t e = p + i * (4 + r.so + 4);
            const uint64_t coll = rd(r, e + 4, r.so);
            const uint32_t idx = static_cast<uint32_t>(rdraw(r, e + 4 + r.so, 4));
            a.strs.push_back(global_heap_string(r, coll, idx));
        }
    } else if (dt.cls == 1 || dt.cls == 0) {
        for (uint64_t i = 0; i < count && i < 64; ++i) {
            const size_t e = p + i * dt.size;
            if (dt.size == 0 || !in_file(r, e, std::max<uint32_t>(dt.size, 8))) break;
            double v = 0;
            if (dt.cls == 1 && dt.size == 4) { float f; memcpy(&f, &r.f[e], 4); v = f; }
            else if (dt.cls == 1 && dt.size == 8) memcpy(&v, &r.f[e], 8);
            else { int64_t iv = 0; memcpy(&iv, &r.f[e], std::min<uint32_t>(dt.size, 8)); v = static_cast<double>(iv); }
            a.nums.push_back(v);
        }
    }
    out.push_back(a);
}

bool walk(Reader& r, uint64_t header, const std::string& path, int depth);

bool walk_btree(Reader& r, uint64_t bt, uint64_t heap_data, const std::string& path, int depth, int tree_depth = 0) {
    const size_t a = static_cast<size_t>(bt + r.base);
    if (bt == UNDEF || !in_file(r, a, 24) || !tag_at(r, a, "TREE")) { r.err = "bad group B-tree node"; return false; }
    if (tree_depth > 16) { r.err = "group B-tree too deep"; return false; }
    const uint8_t level = u8(r, a + 5);
    const uint16_t used = static_cast<uint16_t>(rdraw(r, a + 6, 2));
    size_t p = a + 8 + 2 * r.so;
    for (int i = 0; i < used; ++i) {
        p += r.sl;   // key
        const uint64_t child = rd(r, p, r.so);
        p += r.so;
        if (level > 0) {
            if (!walk_btree(r, child, heap_data, path, depth, tree_depth + 1)) return false;
            continue;
        }
        const size_t s = static_cast<size_t>(child + r.base);
        if (child == UNDEF || !in_file(r, s, 8) || !tag_at(r, s, "SNOD")) { r.err = "bad symbol node"; return false; }
        const uint16_t nsym = static_cast<uint16_t>(rdraw(r, s + 6, 2));
        for (int k = 0; k < nsym; ++k) {
            const size_t e = s + 8 + size_t(k) * (2 * r.so + 24);
            const uint64_t noff = rdraw(r, e, r.so);
            const uint64_t oh = rd(r, e + r.so, r.so);
            if (heap_data == UNDEF || noff >= r.f.size()) { r.err = "symbol name outside the local heap"; return false; }
            const std::string nm = str_at(r, static_cast<size_t>(heap_data + r.base + noff), 4096);
            const std::string child_path = path.empty() ? nm : path + "/" + nm;
            if (!walk(r, oh, child_path, depth + 1)) return false;
        }
    }
    return true;
}

bool walk(Reader& r, uint64_t header, const std::string& path, int depth) {
    if (depth > 32) { r.err = "group nesting too deep"; return false; }
    if (header == UNDEF) { r.err = "link to an undefined object header"; return false; }
    if (++r.visited > kMaxObjects) { r.err = "too many objects (a cycle in the link structure?)"; return false; }
    std::vector<Message> msgs;
    if (!collect_messages(r, header, msgs)) return false;
    bool is_group = false;
    RDataset ds;
    bool has_space = false, has_type = false, has_layout = false;
    for (const auto& m : msgs) {
        switch (m.type) {
            case 0x0011: {   // symbol table
                is_group = true;
                const uint64_t bt = rd(r, m.off, r.so), hp = rd(r, m.off + r.so, r.so);
                const size_t h = static_cast<size_t>(hp + r.base);
                if (hp == UNDEF || !in_file(r, h, 8 + 2 * r.sl + r.so) || !tag_at(r, h, "HEAP")) { r.err = "bad local heap"; return false; }
                const uint64_t seg = rd(r, h + 8 + 2 * r.sl, r.so);
                if (!walk_btree(r, bt, seg, path, depth)) return false;
                break;
            }
            case 0x0002:     // link info
                is_group = true;
                if (rd(r, m.off + 2 + ((u8(r, m.off + 1) & 1) ? 8 : 0), r.so) != UNDEF) {
                    r.err = "group '" + path + "' uses dense link storage (fractal heap), which is outside the supported subset";
                    return false;
                }
                break;
            case 0x0006: {   // link message (compact new-style group)
                is_group = true;
                size_t p = m.off;
                const uint8_t lf = u8(r, p + 1);
                p += 2;
                uint8_t ltype = 0;
                if (lf & 0x08) ltype = u8(r, p++);
                if (lf & 0x04) p += 8;
                if (lf & 0x10) p += 1;
                const int lsz = 1 << (lf & 3);
                const uint64_t nlen = rdraw(r, p, lsz);
                p += lsz;
                if (nlen > 4096 || !in_file(r, p, static_cast<size_t>(nlen))) { r.err = "link name outside the file"; return false; }
                const std::string nm(reinterpret_cast<const char*>(&r.f[p]), static_cast<size_t>(nlen));
                p += static_cast<size_t>(nlen);
                if (ltype == 0) {
                    const uint64_t oh = rd(r, p, r.so);
                    if (!walk(r, oh, path.empty() ? nm : path + "/" + nm, depth + 1)) return false;
                }
                break;
            }
            case 0x0001: parse_dataspace(r, m.off, ds.dims); has_space = true; break;
            case 0x0003: { DType dt = parse_dtype(r, m.off); ds.cls = dt.cls; ds.elem = dt.size; has_type = true; break; }
            case 0x0008: {
                has_layout = true;
                const uint8_t ver = u8(r, m.off);
                if (ver == 3 || ver == 4) {
                    const uint8_t cls = u8(r, m.off + 1);
                    if (cls == 1) { ds.addr = rd(r, m.off + 2, r.so); ds.size = rdraw(r, m.off + 2 + r.so, r.sl); ds.ok = true; }
                    else if (cls == 0) {
                        const uint16_t sz = static_cast<uint16_t>(rdraw(r, m.off + 2, 2));
                        if (!in_file(r, m.off + 4, sz)) { r.err = "compact dataset outside the file"; return false; }
                        ds.compact.assign(r.f.begin() + (m.off + 4), r.f.begin() + (m.off + 4 + sz)); ds.size = sz; ds.ok = true;
                    } else ds.why = "chunked/compressed storage is outside the supported subset";
                } else if (ver == 1 || ver == 2) {
                    const uint8_t rank = u8(r, m.off + 1), cls = u8(r, m.off + 2);
                    if (cls == 1) {
                        ds.addr = rd(r, m.off + 8, r.so);
                        ds.ok = true;   // size derived from dims below
                        (void)rank;
                    } else ds.why = "only contiguous version-1/2 layouts are supported";
                } else ds.why = "unknown layout message version";
                break;
            }
            case 0x000C: parse_attr(r, m, r.attrs[path]); break;
            default: break;
        }
    }
    if (!is_group && has_space && has_type && has_layout) {
        uint64_t n = 1;
        for (uint64_t d : ds.dims) n *= d;
        if (ds.size == 0) ds.size = n * ds.elem;
        r.datasets[path] = ds;
    }
    return true;
}

std::string jesc(const std::string& s) {
    std::string o;
    for (unsigned char c : s) {
        if (c == '"' || c == '\\') { o += '\\'; o += static_cast<char>(c); }
        else if (c < 0x20 || c >= 0x7F) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o += static_cast<char>(c);
    }
    return o;
}

}  // namespace

extern "C" {

// ---- writer API ---------------------------------------------------------------------------------------
void* idcbh_h5w_open(const char* path) {
    Writer* w = new Writer();
    w->path = path;
    return w;
}

int idcbh_h5w_dataset_f32(void* h, const char* path, const float* data, const long long* dims, int rank) {
    Writer* w = static_cast<Writer*>(h);
    std::string leaf;
    WGroup* g = descend(w->root, path, &leaf);
    if (leaf.empty()) { idcbh::set_error("h5w: empty dataset name"); return 1; }
    WDataset d;
    size_t n = 1;
    for (int i = 0; i < rank; ++i) { d.dims.push_back(dims[i]); n *= static_cast<size_t>(dims[i]); }
    d.data.assign(data, data + n);
    g->datasets[leaf] = std::move(d);
    return 0;
}

int idcbh_h5w_attr_strs(void* h, const char* group_path, const char* name, const char* const* strs, int n, int scalar) {
    Writer* w = static_cast<Writer*>(h);
    WGroup* g = descend(w->root, group_path, nullptr);
    WAttr a;
    a.name = name;
    a.scalar = scalar != 0;
    size_t bytes = 0;
    for (int i = 0; i < n; ++i) { a.values.push_back(strs[i]); bytes = std::max(bytes, a.values.back().size()); }
    if (bytes * static_cast<size_t>(n) > 64000) {
        idcbh::set_error("h5w: attribute '%s' is larger than 64000 bytes; split it (Keras chunks layer_names0, 1, ...)", name);
        return 1;
    }
    g->attrs.push_back(a);
    return 0;
}

int idcbh_h5w_close(void* h) {
    std::unique_ptr<Writer> w(static_cast<Writer*>(h));
    w->leaf_k = std::max(4, (max_children(w->root) + 1) / 2);
    if (w->leaf_k > 32000) { idcbh::set_error("h5w: too many children in one group"); return 1; }
    std::vector<uint8_t>& b = w->buf;
    b.assign(96, 0);
    GroupAddr root = write_group(*w, w->root);
    const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    put(b, 0, sig, 8);
    b[13] = 8; b[14] = 8;
    put_le<uint16_t>(b, 16, static_cast<uint16_t>(w->leaf_k));
    put_le<uint16_t>(b, 18, 16);
    put_le<uint64_t>(b, 24, 0);
    put_le<uint64_t>(b, 32, UNDEF);
    put_le<uint64_t>(b, 40, b.size());
    put_le<uint64_t>(b, 48, UNDEF);
    put_le<uint64_t>(b, 56, 0);
    put_le<uint64_t>(b, 64, root.header);
    put_le<uint32_t>(b, 72, 1);
    put_le<uint64_t>(b, 80, root.btree);
    put_le<uint64_t>(b, 88, root.heap);
    const std::string tmp = w->path + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) { idcbh::set_error("h5w: cannot open %s", tmp.c_str()); return 1; }
    const bool ok = fwrite(b.data(), 1, b.size(), f) == b.size();
    if (fclose(f) != 0 || !ok) { idcbh::set_error("h5w: short write"); remove(tmp.c_str()); return 1; }
    if (rename(tmp.c_str(), w->path.c_str()) != 0) { idcbh::set_error("h5w: rename failed"); return 1; }
    return 0;
}

// ---- reader API ---------------------------------------------------------------------------------------
void* idcbh_h5r_open(const char* path) {
    std::unique_ptr<Reader> r(new Reader());
    FILE* f = fopen(path, "rb");
    if (!f) { idcbh::set_error("h5r: cannot open %s", path); return nullptr; }
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    r->f.resize(static_cast<size_t>(n));
    if (n > 0 && fread(r->f.data(), 1, static_cast<size_t>(n), f) != static_cast<size_t>(n)) { fclose(f); idcbh::set_error("h5r: read error"); return nullptr; }
    fclose(f);
    const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    size_t sb = 0;
    bool found = false;
    for (sb = 0; sb + 96 <= r->f.size(); sb = sb ? sb * 2 : 512) {
        if (memcmp(&r->f[sb], sig, 8) == 0) { found = true; break; }
        if (sb > (1u << 24)) break;
    }
    if (!found) { idcbh::set_error("h5r: %s is not an HDF5 file", path); return nullptr; }
    const uint8_t ver = r->f[sb + 8];
    uint64_t root_header = UNDEF;
    const char* only8 = "h5r: only 8-byte offsets/lengths are supported";
    if (ver == 0 || ver == 1) {
        r->so = r->f[sb + 13];
        r->sl = r->f[sb + 14];
        if (r->so != 8 || r->sl != 8) { idcbh::set_error("%s", only8); return nullptr; }   // before any sized read
        size_t p = sb + 24 + (ver == 1 ? 4 : 0);
        r->base = rdraw(*r, p, r->so);
        p += 4 * size_t(r->so);
        root_header = rd(*r, p + r->so, r->so);
    } else if (ver == 2 || ver == 3) {
        r->so = r->f[sb + 9];
        r->sl = r->f[sb + 10];
        if (r->so != 8 || r->sl != 8) { idcbh::set_error("%s", only8); return nullptr; }
        r->base = rdraw(*r, sb + 12, r->so);
        root_header = rd(*r, sb + 12 + 3 * size_t(r->so), r->so);
    } else { idcbh::set_error("h5r: unsupported superblock version %d", ver); return nullptr; }
    if (r->base > r->f.size()) { idcbh::set_error("h5r: base address outside the file"); return nullptr; }
    if (!walk(*r, root_header, "", 0)) { idcbh::set_error("h5r: %s", r->err.c_str()); return nullptr; }
    return r.release();
}

long long idcbh_h5r_list(void* h, char* buf, long long buflen) {
    Reader* r = static_cast<Reader*>(h);
    std::string js = "{\"datasets\":[";
    bool first = true;
    for (const auto& kv : r->datasets) {
        if (!first) js += ",";
        first = false;
        js += "{\"path\":\"" + jesc(kv.first) + "\",\"dims\":[";
        for (size_t i = 0; i < kv.second.dims.size(); ++i) { if (i) js += ","; js += std::to_string(kv.second.dims[i]); }
        js += "],\"class\":" + std::to_string(kv.second.cls) + ",\"elem\":" + std::to_string(kv.second.elem) +
              ",\"readable\":" + (kv.second.ok ? "true" : "false") + ",\"why\":\"" + jesc(kv.second.why) + "\"}";
    }
    js += "],\"attrs\":[";
    first = true;
    for (const auto& kv : r->attrs)
        for (const auto& a : kv.second) {
            if (!first) js += ",";
            first = false;
            js += "{\"object\":\"" + jesc(kv.first) + "\",\"name\":\"" + jesc(a.name) + "\",\"values\":[";
            if (a.is_str) for (size_t i = 0; i < a.strs.size(); ++i) { if (i) js += ","; js += "\"" + jesc(a.strs[i]) + "\""; }
            else for (size_t i = 0; i < a.nums.size(); ++i) { if (i) js += ","; char b[40]; snprintf(b, sizeof b, "%.17g", a.nums[i]); js += b; }
            js += "]}";
        }
    js += "]}";
    if (buf == nullptr || buflen < static_cast<long long>(js.size()) + 1) return static_cast<long long>(js.size()) + 1;
    memcpy(buf, js.c_str(), js.size() + 1);
    return static_cast<long long>(js.size());
}

int idcbh_h5r_read_f32(void* h, const char* path, float* dst, long long n) {
    Reader* r = static_cast<Reader*>(h);
    auto it = r->datasets.find(path);
    if (it == r->datasets.end()) { idcbh::set_error("h5r: no dataset '%s'", path); return 1; }
    const RDataset& d = it->second;
    if (!d.ok) { idcbh::set_error("h5r: dataset '%s': %s", path, d.why.c_str()); return 1; }
    if (d.cls != 1 || (d.elem != 4 && d.elem != 8)) { idcbh::set_error("h5r: dataset '%s' is not float32/float64", path); return 1; }
    uint64_t cnt = 1;
    for (uint64_t v : d.dims) {
        if (v != 0 && cnt > (uint64_t(1) << 40) / v) { idcbh::set_error("h5r: dataset '%s' has an absurd shape", path); return 1; }
        cnt *= v;
    }
    if (n < 0 || static_cast<uint64_t>(n) != cnt) { idcbh::set_error("h5r: dataset '%s' has %llu elements, caller expects %lld", path, (unsigned long long)cnt, n); return 1; }
    const uint8_t* src = d.compact.empty() ? nullptr : d.compact.data();
    if (src && d.compact.size() < cnt * d.elem) { idcbh::set_error("h5r: compact dataset '%s' is shorter than its shape", path); return 1; }
    if (!src) {
        if (d.addr == UNDEF) { memset(dst, 0, cnt * 4); return 0; }   // never written: fill value
        const size_t a = static_cast<size_t>(d.addr + r->base);
        if (!in_file(*r, a, static_cast<size_t>(cnt * d.elem))) { idcbh::set_error("h5r: dataset '%s' points outside the file", path); return 1; }
        if (cnt == 0) return 0;
        src = &r->f[a];
    }
    if (d.elem == 4) memcpy(dst, src, cnt * 4);
    else for (uint64_t i = 0; i < cnt; ++i) { double v; memcpy(&v, src + 8 * i, 8); dst[i] = static_cast<float>(v); }
    return 0;
}

void idcbh_h5r_close(void* h) { delete static_cast<Reader*>(h); }

}  // extern "C"
