// GGUF tensor loader -> resident weights: reads the block_* arrays of a GGUF file straight from
// an mmap and hands them to tb200_weights_upload(), which does the one-time device repack. Host code only.
//
// File layout followed (llama.cpp/ggml/src/ggml.c:20896-21160 gguf_init_from_file, ggml.h:249-256):
//   magic "GGUF" | u32 version (2, 3) | u64 n_tensors | u64 n_kv
//   n_kv x { string key | u32 type | value }   string = u64 len + bytes; ARRAY = u32 elem type + u64 count + elements
//   n_tensors x { string name | u32 n_dims | u64 ne[n_dims] | u32 ggml_type | u64 offset }
//   padding to `general.alignment` (u32 kv, default 32) | tensor data; offsets are relative to the start of the data
// ne[0] is the row length K (elements), ne[1] the row count; a 3-D tensor (MoE experts) is ne[2] stacked matrices,
// which is exactly the row-stacked form tb200_mixmul_dev takes.
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "tb200_common.cuh"

size_t tb200_gguf_row_bytes(int type, long k);

struct tb200_gguf_tensor_info {
    std::string name;
    int n_dims;
    int64_t ne[4];
    int type;
    uint64_t offset;   // from the start of the data section
};

struct tb200_gguf {
    int fd;
    const uint8_t *map;
    size_t size;
    uint32_t version;
    size_t alignment;
    size_t data_start;
    std::vector<tb200_gguf_tensor_info> tensors;
};

namespace {

struct Cursor {
    const uint8_t *p, *end;
    bool ok;
    template <typename V> V get() {
        V v{};
        if (!ok || (size_t)(end - p) < sizeof(V)) { ok = false; return v; }
        memcpy(&v, p, sizeof(V));
        p += sizeof(V);
        return v;
    }
    bool skip(uint64_t n) {
        if (!ok || (uint64_t)(end - p) < n) { ok = false; return false; }
        p += n;
        return true;
    }
    bool str(std::string *out) {
        const uint64_t n = get<uint64_t>();
        if (!ok || (uint64_t)(end - p) < n) { ok = false; return false; }
        if (out) out->assign((const char *)p, (size_t)n);
        p += n;
        return true;
    }
};

// gguf_type sizes (ggml.c:20804-20822); STRING = 8 and ARRAY = 9 are variable
size_t scalar_size(uint32_t t) {
    switch (t) {
    case 0: case 1: case 7: return 1;            // UINT8, INT8, BOOL
    case 2: case 3: return 2;                    // UINT16, INT16
    case 4: case 5: case 6: return 4;            // UINT32, INT32, FLOAT32
    case 10: case 11: case 12: return 8;         // UINT64, INT64, FLOAT64
    default: return 0;
    }
}

bool skip_value(Cursor &c, uint32_t type, uint64_t *u32_value) {
    if (type == 8) return c.str(nullptr);
    if (type == 9) {
        const uint32_t et = c.get<uint32_t>();
        const uint64_t n = c.get<uint64_t>();
        if (!c.ok) return false;
        if (et == 8) {
            for (uint64_t i = 0; i < n; ++i)
                if (!c.str(nullptr)) return false;
            return true;
        }
        const size_t es = scalar_size(et);
        return es && n <= (uint64_t)(c.end - c.p) / es && c.skip(n * es);
    }
    const size_t es = scalar_size(type);
    if (!es) { c.ok = false; return false; }
    if (u32_value && type == 4) { *u32_value = c.get<uint32_t>(); return c.ok; }
    return c.skip(es);
}

}  // namespace

extern "C" {

tb200_gguf *tb200_gguf_open(const char *path) {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { tb200_set_error("gguf: cannot open %s", path); return nullptr; }
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size < 24) { close(fd); tb200_set_error("gguf: %s is too small", path); return nullptr; }
    void *m = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    if (m == MAP_FAILED) { close(fd); tb200_set_error("gguf: mmap failed"); return nullptr; }
    tb200_gguf *g = new tb200_gguf();
    g->fd = fd, g->map = (const uint8_t *)m, g->size = (size_t)st.st_size, g->alignment = 32;
    Cursor c = {g->map, g->map + g->size, true};
    auto fail = [&](const char *why) -> tb200_gguf * {
        tb200_set_error("gguf: %s", why);
        munmap(m, g->size); close(fd); delete g;
        return nullptr;
    };
    if (memcmp(c.p, "GGUF", 4) != 0) return fail("bad magic");
    c.skip(4);
    g->version = c.get<uint32_t>();
    if (g->version < 2 || g->version > 3) return fail("unsupported version (2 and 3 are read)");
    const uint64_t n_tensors = c.get<uint64_t>(), n_kv = c.get<uint64_t>();
    if (!c.ok || n_tensors > (1u << 24) || n_kv > (1u << 24)) return fail("implausible header counts");
    for (uint64_t i = 0; i < n_kv; ++i) {
        std::string key;
        if (!c.str(&key)) return fail("truncated key");
        const uint32_t type = c.get<uint32_t>();
        uint64_t v = 0;
        const bool is_align = key == "general.alignment";
        if (!skip_value(c, type, is_align ? &v : nullptr)) return fail("truncated or unknown metadata value");
        if (is_align && type == 4 && v && (v & (v - 1)) == 0) g->alignment = (size_t)v;
    }
    // a tensor info is at least 8 (name length) + 4 (rank) + 8 (one extent) + 4 (type) + 8 (offset) bytes: a header that announces
    // more of them than the rest of the file can hold is refused before anything is allocated for it
    if (n_tensors > (uint64_t)(c.end - c.p) / 32) return fail("more tensor infos announced than the file can hold");
    g->tensors.resize((size_t)n_tensors);
    for (auto &t : g->tensors) {
        if (!c.str(&t.name)) return fail("truncated tensor name");
        t.n_dims = (int)c.get<uint32_t>();
        if (!c.ok || t.n_dims < 1 || t.n_dims > 4) return fail("tensor rank outside 1..4");
        for (int d = 0; d < 4; ++d) t.ne[d] = 1;
        uint64_t elems = 1;
        for (int d = 0; d < t.n_dims; ++d) {
            const uint64_t e = c.get<uint64_t>();
            if (e > (uint64_t)INT64_MAX || __builtin_mul_overflow(elems, e, &elems) || elems > (uint64_t)INT64_MAX) return fail("tensor extents overflow");
            t.ne[d] = (int64_t)e;
        }
        t.type = (int)c.get<uint32_t>();
        t.offset = c.get<uint64_t>();
        if (!c.ok) return fail("truncated tensor info");
    }
    const size_t pos = (size_t)(c.p - g->map);
    g->data_start = (pos + g->alignment - 1) / g->alignment * g->alignment;
    if (g->data_start > g->size) return fail("data section beyond end of file");
    return g;
}

void tb200_gguf_close(tb200_gguf *g) {
    if (!g) return;
    munmap((void *)g->map, g->size);
    close(g->fd);
    delete g;
}

long tb200_gguf_tensor_count(const tb200_gguf *g) { return g ? (long)g->tensors.size() : 0; }

const char *tb200_gguf_tensor_name(const tb200_gguf *g, long i) {
    return (g && i >= 0 && (size_t)i < g->tensors.size()) ? g->tensors[(size_t)i].name.c_str() : nullptr;
}

// type: ggml_type code; ne[4]; data: host pointer into the mapping (valid until tb200_gguf_close); bytes: tensor size when the
// type is one this module knows (F32, F16, Q4_0, Q8_0, Q4_K, Q8_K, I32), else 0
int tb200_gguf_find(const tb200_gguf *g, const char *name, int *type, int64_t ne[4], const void **data, size_t *bytes) {
    if (!g || !name) return TB200_ERR_INVALID;
    for (const auto &t : g->tensors) {
        if (t.name != name) continue;
        if (type) *type = t.type;
        if (ne) memcpy(ne, t.ne, sizeof(t.ne));
        uint64_t row = 0;
        bool wraps = false;
        switch (t.type) {
        case TB200_TYPE_F32: case TB200_TYPE_I32: wraps = __builtin_mul_overflow((uint64_t)t.ne[0], (uint64_t)4, &row); break;
        default: row = tb200_gguf_row_bytes(t.type, (long)t.ne[0]); break;
        }
        // offset and extents come from the file: compare without letting either sum or product wrap
        const uint64_t avail = (uint64_t)(g->size - g->data_start);
        uint64_t total = row;
        for (int d = 1; d < 4; ++d) wraps |= __builtin_mul_overflow(total, (uint64_t)t.ne[d], &total);
        if (wraps || t.offset > avail || total > avail - t.offset) {
            tb200_set_error("gguf: tensor %s lies outside the file", name);
            return TB200_ERR_INVALID;
        }
        if (data) *data = g->map + g->data_start + t.offset;
        if (bytes) *bytes = (size_t)total;
        return TB200_OK;
    }
    tb200_set_error("gguf: no tensor named %s", name);
    return TB200_ERR_INVALID;
}

// weight matrix (or stack of ne[2] expert matrices) -> resident weights; rows [row_lo, row_hi) of every matrix when
// row_hi > row_lo (a rank's shard), else all rows. Only a 2-D tensor can be row-sliced.
tb200_weights *tb200_gguf_upload(const tb200_gguf *g, const char *name, long row_lo, long row_hi) {
    int type = 0;
    int64_t ne[4];
    const void *data = nullptr;
    size_t bytes = 0;
    if (tb200_gguf_find(g, name, &type, ne, &data, &bytes) != TB200_OK) return nullptr;
    if (type != TB200_TYPE_F16 && type != TB200_TYPE_Q4_0 && type != TB200_TYPE_Q8_0 && !tb200_is_kquant(type) && !tb200_is_g8(type)) {
        tb200_set_error("gguf: tensor %s has type %d, which this module does not hold", name, type);
        return nullptr;
    }
    if (ne[3] != 1) { tb200_set_error("gguf: tensor %s is 4-D", name); return nullptr; }
    const size_t row = tb200_gguf_row_bytes(type, (long)ne[0]);
    if (!row) { tb200_set_error("gguf: tensor %s: K=%ld is not a whole number of blocks", name, (long)ne[0]); return nullptr; }
    long rows = (long)(ne[1] * ne[2]);
    const uint8_t *base = (const uint8_t *)data;
    if (row_hi > row_lo) {
        if (ne[2] != 1 || row_lo < 0 || row_hi > ne[1]) { tb200_set_error("gguf: bad row slice of %s", name); return nullptr; }
        base += (size_t)row_lo * row, rows = row_hi - row_lo;
    }
    return tb200_weights_upload(base, type, rows, (long)ne[0], 0);
}

}  // extern "C"
