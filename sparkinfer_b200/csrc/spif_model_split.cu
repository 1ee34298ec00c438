// spif_model_split.cu -- the model-split file (`-ms <path>`): consumer and producer.  Host code only, no device work.
//
// Reference consumer: sparkinfer_cache_manager ctor, src/llama-sparkinfer.cpp:150-158 (KV `ffn_group_size` i32,
// `ffn_normalized_pattern` f32[n_layer]), :257 / :269-276 (tensors `blk.N.ffn_reorder_perms`, I32 {n_ff}: new_row -> old_row,
// applied by :283-352 == spif_reorder_rows).  The container is GGUF v3 (ggml/include/gguf.h:1-60, ggml/src/gguf.cpp:320-640
// read, :1355-1420 write).  The reference ships no producer for this file; spif_model_split_write is one: it
// emits exactly what the consumer above reads, so a file written here loads in the unmodified reference (tests/test_model_split.py
// reads it back with the reference's own gguf_init_from_file, and reads a file written by the reference's gguf writer with
// this parser).
//
// Deliberately minimal: the parser walks EVERY GGUF value type so that unknown keys are skipped correctly, but only the keys
// and tensors above are kept.  Errors are return codes (SPIF_E_SHAPE for a malformed / truncated file), never aborts.
#include "spif_common.cuh"

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

enum : uint32_t { GT_U8, GT_I8, GT_U16, GT_I16, GT_U32, GT_I32, GT_F32, GT_BOOL, GT_STR, GT_ARR, GT_U64, GT_I64, GT_F64, GT_COUNT };
const size_t k_gt_size[GT_COUNT] = { 1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8 };
constexpr uint32_t k_ggml_type_i32 = 26;  // GGML_TYPE_I32, ggml/include/ggml.h:385-415
constexpr uint32_t k_ggml_type_f32 = 0;

struct Reader {
    const std::vector<unsigned char> & b;
    size_t                             pos = 0;
    bool                               ok  = true;
    explicit Reader(const std::vector<unsigned char> & buf) : b(buf) {}
    bool take(void * dst, size_t n) {
        if (!ok || n > b.size() - pos) {
            ok = false;
            return false;
        }
        memcpy(dst, b.data() + pos, n);
        pos += n;
        return true;
    }
    template <typename T> T get() {
        T v = T();
        take(&v, sizeof(T));
        return v;
    }
    std::string str() {
        const uint64_t n = get<uint64_t>();
        if (!ok || n > b.size() - pos) {
            ok = false;
            return std::string();
        }
        std::string s((const char *) b.data() + pos, (size_t) n);
        pos += (size_t) n;
        return s;
    }
    void skip(size_t n) {
        if (!ok || n > b.size() - pos) {
            ok = false;
            return;
        }
        pos += n;
    }
};

struct TensorInfo {
    std::string name;
    uint32_t    type;
    uint64_t    ne[4];
    uint32_t    n_dims;
    uint64_t    offset;
};

}  // namespace

struct spif_model_split {
    int32_t                           group_size = 0;
    std::vector<float>                pattern;  // ffn_normalized_pattern
    std::vector<std::vector<int32_t>> perms;    // [layer][n_ff], empty where the file has no tensor for the layer
    int32_t                           n_ff = 0;
};

extern "C" {

int spif_model_split_open(const char * path, spif_model_split ** out) {
    SPIF_REQUIRE(path && out, SPIF_E_NULL);
    *out    = nullptr;
    FILE * f = fopen(path, "rb");
    SPIF_REQUIRE(f != nullptr, SPIF_E_NULL);
    std::vector<unsigned char> buf;
    {
        fseek(f, 0, SEEK_END);
        const long sz = ftell(f);
        fseek(f, 0, SEEK_SET);
        if (sz < 24) {
            fclose(f);
            return SPIF_E_SHAPE;
        }
        buf.resize((size_t) sz);
        const size_t got = fread(buf.data(), 1, (size_t) sz, f);
        fclose(f);
        SPIF_REQUIRE(got == (size_t) sz, SPIF_E_SHAPE);
    }
    Reader r(buf);
    char   magic[4];
    r.take(magic, 4);
    SPIF_REQUIRE(r.ok && memcmp(magic, "GGUF", 4) == 0, SPIF_E_SHAPE);
    const uint32_t version = r.get<uint32_t>();
    SPIF_REQUIRE(r.ok && version >= 2 && version <= 3, SPIF_E_UNSUPPORTED);  // v1 had 32-bit counts (gguf.cpp:368-374 rejects it too)
    const uint64_t n_tensors = r.get<uint64_t>();
    const uint64_t n_kv      = r.get<uint64_t>();
    SPIF_REQUIRE(r.ok && n_tensors < (1u << 20) && n_kv < (1u << 20), SPIF_E_SHAPE);

    auto ms = new spif_model_split();
    uint32_t alignment = 32;  // GGUF_DEFAULT_ALIGNMENT
    bool     have_gs   = false;
    for (uint64_t i = 0; i < n_kv && r.ok; ++i) {
        const std::string key  = r.str();
        const uint32_t    type = r.get<uint32_t>();
        if (!r.ok || type >= GT_COUNT) {
            r.ok = false;
            break;
        }
        if (type == GT_STR) {
            r.str();
        } else if (type == GT_ARR) {
            const uint32_t et = r.get<uint32_t>();
            const uint64_t n  = r.get<uint64_t>();
            if (!r.ok || et >= GT_COUNT || et == GT_ARR) {
                r.ok = false;
                break;
            }
            if (et == GT_STR) {
                for (uint64_t k = 0; k < n && r.ok; ++k) {
                    r.str();
                }
            } else if (key == "ffn_normalized_pattern" && et == GT_F32) {
                if (n > (buf.size() - r.pos) / 4) {
                    r.ok = false;
                    break;
                }
                ms->pattern.resize((size_t) n);
                r.take(ms->pattern.data(), (size_t) n * 4);
            } else {
                if (n > (buf.size() - r.pos) / k_gt_size[et]) {
                    r.ok = false;
                    break;
                }
                r.skip((size_t) n * k_gt_size[et]);
            }
        } else if (key == "ffn_group_size" && type == GT_I32) {
            ms->group_size = r.get<int32_t>();
            have_gs        = true;
        } else if (key == "general.alignment" && type == GT_U32) {
            alignment = r.get<uint32_t>();
        } else {
            r.skip(k_gt_size[type]);
        }
    }
    std::vector<TensorInfo> infos;
    for (uint64_t i = 0; i < n_tensors && r.ok; ++i) {
        TensorInfo t;
        t.name   = r.str();
        t.n_dims = r.get<uint32_t>();
        if (!r.ok || t.n_dims > 4) {
            r.ok = false;
            break;
        }
        for (int d = 0; d < 4; ++d) {
            t.ne[d] = 1;
        }
        for (uint32_t d = 0; d < t.n_dims; ++d) {
            t.ne[d] = r.get<uint64_t>();
        }
        t.type   = r.get<uint32_t>();
        t.offset = r.get<uint64_t>();
        infos.push_back(t);
    }
    if (!r.ok || !have_gs || ms->group_size <= 0 || ms->pattern.empty() || alignment == 0 || (alignment & (alignment - 1))) {
        delete ms;
        return SPIF_E_SHAPE;
    }
    const size_t data_start = (r.pos + alignment - 1) / alignment * alignment;
    ms->perms.resize(ms->pattern.size());
    for (const TensorInfo & t : infos) {
        int layer = -1, consumed = 0;
        if (sscanf(t.name.c_str(), "blk.%d.ffn_reorder_perms%n", &layer, &consumed) != 1 || consumed != (int) t.name.size()) {
            continue;  // not ours: the reference copies by name too (src/llama-sparkinfer.cpp:269-276)
        }
        const uint64_t n = t.ne[0];
        if (layer < 0 || layer >= (int) ms->perms.size() || t.type != k_ggml_type_i32 || t.ne[1] * t.ne[2] * t.ne[3] != 1 ||
            n == 0 || n > (1u << 30) || data_start + t.offset > buf.size() || n * 4 > buf.size() - data_start - t.offset) {
            delete ms;
            return SPIF_E_SHAPE;
        }
        if (ms->n_ff == 0) {
            ms->n_ff = (int32_t) n;
        }
        if ((int32_t) n != ms->n_ff) {
            delete ms;
            return SPIF_E_SHAPE;
        }
        ms->perms[layer].resize((size_t) n);
        memcpy(ms->perms[layer].data(), buf.data() + data_start + t.offset, (size_t) n * 4);
    }
    *out = ms;
    return 0;
}

void spif_model_split_close(spif_model_split * ms) {
    delete ms;
}

int spif_model_split_info(const spif_model_split * ms, int32_t * group_size, int32_t * n_layer, int32_t * n_ff) {
    SPIF_REQUIRE(ms, SPIF_E_NULL);
    if (group_size) {
        *group_size = ms->group_size;
    }
    if (n_layer) {
        *n_layer = (int32_t) ms->pattern.size();
    }
    if (n_ff) {
        *n_ff = ms->n_ff;
    }
    return 0;
}

int spif_model_split_pattern(const spif_model_split * ms, float * pattern_out, int n_layer) {
    SPIF_REQUIRE(ms && pattern_out, SPIF_E_NULL);
    SPIF_REQUIRE(n_layer == (int) ms->pattern.size(), SPIF_E_SHAPE);
    memcpy(pattern_out, ms->pattern.data(), sizeof(float) * ms->pattern.size());
    return 0;
}

// perm_out[new_row] = old_row; a layer without a tensor in the file, or a tensor that is not a permutation of [0, n_ff),
// is an error (the reference would index out of bounds, src/llama-sparkinfer.cpp:292-296)
int spif_model_split_perm(const spif_model_split * ms, int layer, int32_t * perm_out, int n_ff) {
    SPIF_REQUIRE(ms && perm_out, SPIF_E_NULL);
    SPIF_REQUIRE(layer >= 0 && layer < (int) ms->perms.size() && n_ff == ms->n_ff && !ms->perms[layer].empty(), SPIF_E_SHAPE);
    const std::vector<int32_t> & p = ms->perms[layer];
    std::vector<unsigned char>   seen((size_t) n_ff, 0);
    for (int i = 0; i < n_ff; ++i) {
        SPIF_REQUIRE(p[i] >= 0 && p[i] < n_ff && !seen[p[i]], SPIF_E_SHAPE);
        seen[p[i]] = 1;
    }
    memcpy(perm_out, p.data(), sizeof(int32_t) * (size_t) n_ff);
    return 0;
}

// producer: GGUF v3, little endian, default alignment; KV order and tensor order as listed in the header comment
int spif_model_split_write(const char * path, int group_size, const float * pattern, int n_layer, const int32_t * perms, int n_ff) {
    SPIF_REQUIRE(path && pattern && perms, SPIF_E_NULL);
    SPIF_REQUIRE(group_size > 0 && n_layer > 0 && n_ff > 0 && n_ff % group_size == 0, SPIF_E_SHAPE);
    std::vector<unsigned char> b;
    auto put = [&](const void * p, size_t n) { b.insert(b.end(), (const unsigned char *) p, (const unsigned char *) p + n); };
    auto u32 = [&](uint32_t v) { put(&v, 4); };
    auto u64 = [&](uint64_t v) { put(&v, 8); };
    auto str = [&](const std::string & s) {
        u64(s.size());
        put(s.data(), s.size());
    };
    put("GGUF", 4);
    u32(3);
    u64((uint64_t) n_layer);  // tensors
    u64(2);                   // KV pairs
    str("ffn_group_size");
    u32(GT_I32);
    {
        const int32_t v = group_size;
        put(&v, 4);
    }
    str("ffn_normalized_pattern");
    u32(GT_ARR);
    u32(GT_F32);
    u64((uint64_t) n_layer);
    put(pattern, sizeof(float) * (size_t) n_layer);
    const size_t   align  = 32;
    const uint64_t tbytes = ((uint64_t) n_ff * 4 + align - 1) / align * align;
    for (int l = 0; l < n_layer; ++l) {
        char name[64];
        snprintf(name, sizeof(name), "blk.%d.ffn_reorder_perms", l);
        str(name);
        u32(1);
        u64((uint64_t) n_ff);
        u32(k_ggml_type_i32);
        u64((uint64_t) l * tbytes);
    }
    b.resize((b.size() + align - 1) / align * align, 0);
    for (int l = 0; l < n_layer; ++l) {
        const size_t at = b.size();
        put(perms + (size_t) l * n_ff, (size_t) n_ff * 4);
        b.resize(at + (size_t) tbytes, 0);
    }
    FILE * f = fopen(path, "wb");
    SPIF_REQUIRE(f != nullptr, SPIF_E_NULL);
    const size_t w = fwrite(b.data(), 1, b.size(), f);
    const int    c = fclose(f);
    SPIF_REQUIRE(w == b.size() && c == 0, SPIF_E_SHAPE);
    return 0;
}

}  // extern "C"
