// TensorFlow "V2" checkpoint bundle writer/reader (native, no TensorFlow needed).
//
// tf.train.Saver() in the reference (encoder:138, gan:160, lstm:101) writes
//   <prefix>.index                 a sorted string table: "" -> BundleHeaderProto,
//                                  <variable name> -> BundleEntryProto
//   <prefix>.data-00000-of-00001   the tensors' raw little-endian bytes, back to back
// The table is the LevelDB SSTable layout (prefix-compressed data blocks with restart points,
// per-block masked CRC32C trailer, index block, 48-byte footer with the table magic).  This file
// implements that layout from its public description so that checkpoints written here carry the
// reference's on-disk format as well as its variable names.  Uncompressed blocks
// only.  TensorFlow is not installed in this image, so the format is validated by the reader below
// (block CRCs, footer magic, entry CRCs) and by byte-level tests of known vectors, not by TF itself.
//
// C ABI (ctypes, see psb200/tf_bundle.py):
//   psb_crc32c(data, n)                                            -> unmasked CRC32C
//   psb_bundle_writer_open(prefix) / _add(...) / _finish(w)        -> write one bundle
//   psb_bundle_reader_open(prefix) / _count / _entry / _read / _close
//   psb_bundle_error()                                             -> last error text (thread-local)
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

namespace {

thread_local std::string g_err;
int fail(const std::string& m) { g_err = m; return -1; }

// ---------------------------------------------------------------- CRC32C (Castagnoli)
uint32_t g_tab[8][256];
bool g_tab_ready = false;
void init_tab() {
    if (g_tab_ready) return;
    for (uint32_t i = 0; i < 256; ++i) {
        uint32_t c = i;
        for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);
        g_tab[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
        for (int t = 1; t < 8; ++t) g_tab[t][i] = (g_tab[t - 1][i] >> 8) ^ g_tab[0][g_tab[t - 1][i] & 0xff];
    g_tab_ready = true;
}

uint32_t crc32c_sw(uint32_t crc, const uint8_t* p, size_t n) {
    init_tab();
    crc = ~crc;
    while (n >= 8) {   // slicing-by-8
        uint64_t w;
        memcpy(&w, p, 8);
        w ^= crc;
        crc = g_tab[7][w & 0xff] ^ g_tab[6][(w >> 8) & 0xff] ^ g_tab[5][(w >> 16) & 0xff] ^ g_tab[4][(w >> 24) & 0xff] ^
              g_tab[3][(w >> 32) & 0xff] ^ g_tab[2][(w >> 40) & 0xff] ^ g_tab[1][(w >> 48) & 0xff] ^ g_tab[0][w >> 56];
        p += 8;
        n -= 8;
    }
    while (n--) crc = g_tab[0][(crc ^ *p++) & 0xff] ^ (crc >> 8);
    return ~crc;
}

#if defined(__x86_64__)
__attribute__((target("sse4.2"))) uint32_t crc32c_hw(uint32_t crc, const uint8_t* p, size_t n) {
    uint64_t c = static_cast<uint32_t>(~crc);
    while (n >= 8) {
        uint64_t w;
        memcpy(&w, p, 8);
        c = __builtin_ia32_crc32di(c, w);
        p += 8;
        n -= 8;
    }
    uint32_t c32 = static_cast<uint32_t>(c);
    while (n--) c32 = __builtin_ia32_crc32qi(c32, *p++);
    return ~c32;
}
#endif

uint32_t crc32c(const void* data, size_t n, uint32_t init = 0) {
    const uint8_t* p = static_cast<const uint8_t*>(data);
#if defined(__x86_64__)
    static const bool hw = __builtin_cpu_supports("sse4.2");
    if (hw) return crc32c_hw(init, p, n);
#endif
    return crc32c_sw(init, p, n);
}

// Stored CRCs are rotated and offset so that a CRC of data containing CRCs stays well distributed.
uint32_t mask_crc(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xa282ead8u; }

// ---------------------------------------------------------------- little encoders
void put_fixed32(std::string& s, uint32_t v) { char b[4]; memcpy(b, &v, 4); s.append(b, 4); }
void put_fixed64(std::string& s, uint64_t v) { char b[8]; memcpy(b, &v, 8); s.append(b, 8); }
void put_varint(std::string& s, uint64_t v) {
    while (v >= 0x80) { s.push_back(static_cast<char>(v | 0x80)); v >>= 7; }
    s.push_back(static_cast<char>(v));
}
bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t* v) {
    uint64_t r = 0;
    for (int shift = 0; shift <= 63 && p < end; shift += 7) {
        uint64_t b = *p++;
        r |= (b & 0x7f) << shift;
        if (!(b & 0x80)) { *v = r; return true; }
    }
    return false;
}
uint32_t get_fixed32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }

// protobuf wire helpers: field key = (number << 3) | wire type; 0 varint, 2 length-delimited, 5 fixed32
void pb_varint(std::string& s, int field, uint64_t v) { put_varint(s, (field << 3) | 0); put_varint(s, v); }
void pb_bytes(std::string& s, int field, const std::string& v) { put_varint(s, (field << 3) | 2); put_varint(s, v.size()); s += v; }
void pb_fixed32(std::string& s, int field, uint32_t v) { put_varint(s, (field << 3) | 5); put_fixed32(s, v); }

// ---------------------------------------------------------------- table blocks
constexpr uint64_t kTableMagic = 0xdb4775248b80fb57ull;
constexpr size_t kBlockSize = 262144;      // flush a data block once it is this large
constexpr int kRestartInterval = 16;
constexpr size_t kFooterSize = 48;         // two padded block handles (40) + magic (8)

struct BlockBuilder {
    std::string buf, last_key;
    std::vector<uint32_t> restarts{0};
    int counter = 0, interval;
    explicit BlockBuilder(int iv) : interval(iv) {}
    bool empty() const { return buf.empty(); }
    size_t size_estimate() const { return buf.size() + restarts.size() * 4 + 4; }
    void add(const std::string& key, const std::string& value) {
        size_t shared = 0;
        if (counter < interval) {
            const size_t lim = std::min(last_key.size(), key.size());
            while (shared < lim && last_key[shared] == key[shared]) ++shared;
        } else {
            restarts.push_back(static_cast<uint32_t>(buf.size()));
            counter = 0;
        }
        put_varint(buf, shared);
        put_varint(buf, key.size() - shared);
        put_varint(buf, value.size());
        buf.append(key, shared, std::string::npos);
        buf += value;
        last_key = key;
        ++counter;
    }
    std::string finish() {
        std::string out = buf;
        for (uint32_t r : restarts) put_fixed32(out, r);
        put_fixed32(out, static_cast<uint32_t>(restarts.size()));
        return out;
    }
    void reset() { buf.clear(); last_key.clear(); restarts.assign(1, 0); counter = 0; }
};

struct Handle { uint64_t offset = 0, size = 0; };
std::string encode_handle(const Handle& h) { std::string s; put_varint(s, h.offset); put_varint(s, h.size); return s; }

// Appends block contents + 5-byte trailer (compression type 0, masked CRC of contents+type).
Handle write_block(std::string& file, const std::string& contents) {
    Handle h{file.size(), contents.size()};
    file += contents;
    const char type = 0;
    uint32_t c = crc32c(contents.data(), contents.size());
    c = crc32c(&type, 1, c);
    file.push_back(type);
    put_fixed32(file, mask_crc(c));
    return h;
}

std::string build_table(const std::map<std::string, std::string>& kv) {
    std::string file;
    BlockBuilder data(kRestartInterval), index(1);
    auto flush = [&]() {
        if (data.empty()) return;
        const std::string last = data.last_key;
        Handle h = write_block(file, data.finish());
        index.add(last, encode_handle(h));      // any key >= last key of the block and < the next block's first
        data.reset();
    };
    for (const auto& e : kv) {
        data.add(e.first, e.second);
        if (data.size_estimate() >= kBlockSize) flush();
    }
    flush();
    BlockBuilder meta(kRestartInterval);
    Handle mh = write_block(file, meta.finish());
    Handle ih = write_block(file, index.finish());
    std::string footer = encode_handle(mh) + encode_handle(ih);
    footer.resize(40, '\0');
    put_fixed64(footer, kTableMagic);
    file += footer;
    return file;
}

bool read_block(const std::string& file, const Handle& h, std::string* out) {
    if (h.offset + h.size + 5 > file.size()) { g_err = "block handle out of range"; return false; }
    const char* p = file.data() + h.offset;
    if (p[h.size] != 0) { g_err = "compressed blocks are not supported"; return false; }
    const uint32_t want = get_fixed32(reinterpret_cast<const uint8_t*>(p + h.size + 1));
    if (mask_crc(crc32c(p, h.size + 1)) != want) { g_err = "block checksum mismatch"; return false; }
    out->assign(p, h.size);
    return true;
}

bool parse_block(const std::string& blk, std::vector<std::pair<std::string, std::string>>* out) {
    if (blk.size() < 4) { g_err = "short block"; return false; }
    const uint32_t nrestart = get_fixed32(reinterpret_cast<const uint8_t*>(blk.data() + blk.size() - 4));
    if (4ull * nrestart + 4 > blk.size()) { g_err = "bad restart count"; return false; }
    const uint8_t* p = reinterpret_cast<const uint8_t*>(blk.data());
    const uint8_t* end = p + blk.size() - 4 - 4ull * nrestart;
    std::string key;
    while (p < end) {
        uint64_t shared, nonshared, vlen;
        if (!get_varint(p, end, &shared) || !get_varint(p, end, &nonshared) || !get_varint(p, end, &vlen) ||
            shared > key.size() || nonshared + vlen > static_cast<uint64_t>(end - p)) {
            g_err = "corrupt block entry";
            return false;
        }
        key.resize(shared);
        key.append(reinterpret_cast<const char*>(p), nonshared);
        p += nonshared;
        out->emplace_back(key, std::string(reinterpret_cast<const char*>(p), vlen));
        p += vlen;
    }
    return true;
}

bool decode_handle(const uint8_t*& p, const uint8_t* end, Handle* h) {
    return get_varint(p, end, &h->offset) && get_varint(p, end, &h->size);
}

bool read_table(const std::string& file, std::vector<std::pair<std::string, std::string>>* out) {
    if (file.size() < kFooterSize) { g_err = "index file shorter than a table footer"; return false; }
    const uint8_t* f = reinterpret_cast<const uint8_t*>(file.data() + file.size() - kFooterSize);
    uint64_t magic;
    memcpy(&magic, f + 40, 8);
    if (magic != kTableMagic) { g_err = "not a table file (bad magic)"; return false; }
    const uint8_t* p = f;
    Handle mh, ih;
    if (!decode_handle(p, f + 40, &mh) || !decode_handle(p, f + 40, &ih)) { g_err = "bad footer"; return false; }
    std::string iblk;
    if (!read_block(file, ih, &iblk)) return false;
    std::vector<std::pair<std::string, std::string>> idx;
    if (!parse_block(iblk, &idx)) return false;
    for (const auto& e : idx) {
        const uint8_t* q = reinterpret_cast<const uint8_t*>(e.second.data());
        Handle dh;
        if (!decode_handle(q, q + e.second.size(), &dh)) { g_err = "bad index entry"; return false; }
        std::string dblk;
        if (!read_block(file, dh, &dblk) || !parse_block(dblk, out)) return false;
    }
    return true;
}

bool slurp(const std::string& path, std::string* out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) { g_err = "cannot open " + path; return false; }
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out->resize(n > 0 ? n : 0);
    const bool ok = n <= 0 || fread(&(*out)[0], 1, n, f) == static_cast<size_t>(n);
    fclose(f);
    if (!ok) g_err = "short read on " + path;
    return ok;
}

// ---------------------------------------------------------------- bundle protos
// BundleHeaderProto { int32 num_shards = 1; Endianness endianness = 2 (LITTLE = 0); VersionDef version = 3 }
// VersionDef { int32 producer = 1; int32 min_consumer = 2 }
// BundleEntryProto { DataType dtype = 1; TensorShapeProto shape = 2; int32 shard_id = 3; int64 offset = 4;
//                    int64 size = 5; fixed32 crc32c = 6 }
// TensorShapeProto { repeated Dim dim = 2 { int64 size = 1 } }
struct Entry {
    std::string name;
    int dtype = 0;
    std::vector<int64_t> dims;
    int64_t offset = 0, size = 0;
    uint32_t crc = 0;       // masked, as stored
};

std::string encode_header() {
    std::string ver, h;
    pb_varint(ver, 1, 1);          // producer = tensor-bundle version 1
    pb_varint(h, 1, 1);            // one shard
    pb_bytes(h, 3, ver);
    return h;
}

std::string encode_entry(const Entry& e) {
    std::string shape, s;
    for (int64_t d : e.dims) {
        std::string dim;
        pb_varint(dim, 1, static_cast<uint64_t>(d));
        pb_bytes(shape, 2, dim);
    }
    if (e.dtype) pb_varint(s, 1, e.dtype);
    pb_bytes(s, 2, shape);
    if (e.offset) pb_varint(s, 4, static_cast<uint64_t>(e.offset));
    if (e.size) pb_varint(s, 5, static_cast<uint64_t>(e.size));
    pb_fixed32(s, 6, e.crc);
    return s;
}

// Generic walk over one message; calls back per field.  Returns false on malformed input.
template <class F>
bool pb_walk(const uint8_t* p, const uint8_t* end, F&& on_field) {
    while (p < end) {
        uint64_t key;
        if (!get_varint(p, end, &key)) return false;
        const int field = static_cast<int>(key >> 3), wt = static_cast<int>(key & 7);
        uint64_t v = 0;
        const uint8_t* body = nullptr;
        if (wt == 0) { if (!get_varint(p, end, &v)) return false; }
        else if (wt == 1) { if (end - p < 8) return false; memcpy(&v, p, 8); p += 8; }
        else if (wt == 5) { if (end - p < 4) return false; v = get_fixed32(p); p += 4; }
        else if (wt == 2) { if (!get_varint(p, end, &v) || v > static_cast<uint64_t>(end - p)) return false; body = p; p += v; }
        else return false;
        on_field(field, wt, v, body);
    }
    return true;
}

bool decode_entry(const std::string& s, Entry* e) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(s.data());
    bool ok = true;
    ok &= pb_walk(p, p + s.size(), [&](int field, int wt, uint64_t v, const uint8_t* body) {
        if (field == 1 && wt == 0) e->dtype = static_cast<int>(v);
        else if (field == 4 && wt == 0) e->offset = static_cast<int64_t>(v);
        else if (field == 5 && wt == 0) e->size = static_cast<int64_t>(v);
        else if (field == 6 && wt == 5) e->crc = static_cast<uint32_t>(v);
        else if (field == 3 && wt == 0 && v != 0) ok = false;                 // multi-shard bundles: not supported
        else if (field == 7) ok = false;                                      // partitioned-variable slices: not supported
        else if (field == 2 && wt == 2)
            ok &= pb_walk(body, body + v, [&](int f2, int wt2, uint64_t v2, const uint8_t* b2) {
                if (f2 == 2 && wt2 == 2)
                    ok &= pb_walk(b2, b2 + v2, [&](int f3, int wt3, uint64_t v3, const uint8_t*) {
                        if (f3 == 1 && wt3 == 0) e->dims.push_back(static_cast<int64_t>(v3));
                    });
            });
    });
    return ok;
}

struct Writer {
    std::string prefix;
    FILE* data = nullptr;
    int64_t offset = 0;
    std::map<std::string, std::string> table;
};

struct Reader {
    std::string prefix;
    std::vector<Entry> entries;
    int num_shards = 0;
};

std::string data_path(const std::string& prefix) { return prefix + ".data-00000-of-00001"; }

}  // namespace

extern "C" {

const char* psb_bundle_error() { return g_err.c_str(); }

uint32_t psb_crc32c(const void* data, uint64_t n) { return crc32c(data, n); }
uint32_t psb_crc32c_sw(const void* data, uint64_t n) { return crc32c_sw(0, static_cast<const uint8_t*>(data), n); }
uint32_t psb_crc32c_mask(uint32_t crc) { return mask_crc(crc); }

void* psb_bundle_writer_open(const char* prefix) {
    Writer* w = new Writer();
    w->prefix = prefix;
    w->data = fopen((data_path(w->prefix) + ".tmp").c_str(), "wb");
    if (!w->data) { g_err = "cannot create " + data_path(w->prefix) + ".tmp"; delete w; return nullptr; }
    w->table[""] = encode_header();
    return w;
}

// dtype uses TensorFlow's DataType numbering (DT_FLOAT = 1, DT_INT32 = 3, ...).
int psb_bundle_writer_add(void* handle, const char* name, int dtype, int ndim, const int64_t* dims,
                          const void* bytes, uint64_t nbytes) {
    Writer* w = static_cast<Writer*>(handle);
    if (!w || !name || !*name) return fail("psb_bundle_writer_add: bad argument");
    if (w->table.count(name)) return fail(std::string("psb_bundle_writer_add: duplicate key ") + name);
    Entry e;
    e.dtype = dtype;
    e.dims.assign(dims, dims + ndim);
    e.offset = w->offset;
    e.size = static_cast<int64_t>(nbytes);
    e.crc = mask_crc(crc32c(bytes, nbytes));
    if (nbytes && fwrite(bytes, 1, nbytes, w->data) != nbytes) return fail("psb_bundle_writer_add: short write");
    w->offset += static_cast<int64_t>(nbytes);
    w->table[name] = encode_entry(e);
    return 0;
}

// Closes the data shard, writes the index table, and renames both into place (data first, so an
// index never names bytes that are not there).  Frees the writer either way.
int psb_bundle_writer_finish(void* handle) {
    Writer* w = static_cast<Writer*>(handle);
    if (!w) return fail("psb_bundle_writer_finish: null writer");
    int rc = 0;
    if (fclose(w->data) != 0) rc = fail("psb_bundle_writer_finish: close failed");
    const std::string dpath = data_path(w->prefix), ipath = w->prefix + ".index";
    if (rc == 0) {
        const std::string file = build_table(w->table);
        FILE* f = fopen((ipath + ".tmp").c_str(), "wb");
        if (!f || fwrite(file.data(), 1, file.size(), f) != file.size()) rc = fail("cannot write " + ipath + ".tmp");
        if (f && fclose(f) != 0 && rc == 0) rc = fail("close failed on " + ipath + ".tmp");
    }
    if (rc == 0 && rename((dpath + ".tmp").c_str(), dpath.c_str()) != 0) rc = fail("rename failed for " + dpath);
    if (rc == 0 && rename((ipath + ".tmp").c_str(), ipath.c_str()) != 0) rc = fail("rename failed for " + ipath);
    delete w;
    return rc;
}

void* psb_bundle_reader_open(const char* prefix) {
    std::string file;
    if (!slurp(std::string(prefix) + ".index", &file)) return nullptr;
    std::vector<std::pair<std::string, std::string>> kv;
    if (!read_table(file, &kv)) return nullptr;
    Reader* r = new Reader();
    r->prefix = prefix;
    for (const auto& e : kv) {
        if (e.first.empty()) {
            const uint8_t* p = reinterpret_cast<const uint8_t*>(e.second.data());
            int endian = 0;
            pb_walk(p, p + e.second.size(), [&](int field, int wt, uint64_t v, const uint8_t*) {
                if (field == 1 && wt == 0) r->num_shards = static_cast<int>(v);
                if (field == 2 && wt == 0) endian = static_cast<int>(v);
            });
            if (r->num_shards != 1 || endian != 0) {
                g_err = "only single-shard little-endian bundles are supported";
                delete r;
                return nullptr;
            }
            continue;
        }
        Entry en;
        en.name = e.first;
        if (!decode_entry(e.second, &en)) { g_err = "unsupported or corrupt entry for " + e.first; delete r; return nullptr; }
        r->entries.push_back(en);
    }
    if (r->num_shards != 1) { g_err = "bundle header missing"; delete r; return nullptr; }
    return r;
}

int psb_bundle_reader_count(void* handle) { return handle ? static_cast<int>(static_cast<Reader*>(handle)->entries.size()) : -1; }

// name_out (capacity name_cap) receives the NUL-terminated key; dims_out must hold 8 values.
int psb_bundle_reader_entry(void* handle, int i, char* name_out, int name_cap, int* dtype, int* ndim,
                            int64_t* dims_out, int64_t* nbytes) {
    Reader* r = static_cast<Reader*>(handle);
    if (!r || i < 0 || i >= static_cast<int>(r->entries.size())) return fail("psb_bundle_reader_entry: bad index");
    const Entry& e = r->entries[i];
    if (static_cast<int>(e.name.size()) + 1 > name_cap || e.dims.size() > 8) return fail("psb_bundle_reader_entry: buffer too small");
    memcpy(name_out, e.name.c_str(), e.name.size() + 1);
    *dtype = e.dtype;
    *ndim = static_cast<int>(e.dims.size());
    for (size_t k = 0; k < e.dims.size(); ++k) dims_out[k] = e.dims[k];
    *nbytes = e.size;
    return 0;
}

// Copies entry i's bytes into dst (capacity cap) and verifies its checksum.
int psb_bundle_reader_read(void* handle, int i, void* dst, uint64_t cap) {
    Reader* r = static_cast<Reader*>(handle);
    if (!r || i < 0 || i >= static_cast<int>(r->entries.size())) return fail("psb_bundle_reader_read: bad index");
    const Entry& e = r->entries[i];
    if (static_cast<uint64_t>(e.size) > cap) return fail("psb_bundle_reader_read: destination too small");
    FILE* f = fopen(data_path(r->prefix).c_str(), "rb");
    if (!f) return fail("cannot open " + data_path(r->prefix));
    bool ok = fseek(f, static_cast<long>(e.offset), SEEK_SET) == 0 &&
              (e.size == 0 || fread(dst, 1, e.size, f) == static_cast<size_t>(e.size));
    fclose(f);
    if (!ok) return fail("short read for " + e.name);
    if (mask_crc(crc32c(dst, e.size)) != e.crc) return fail("tensor checksum mismatch for " + e.name);
    return 0;
}

void psb_bundle_reader_close(void* handle) { delete static_cast<Reader*>(handle); }

}  // extern "C"
