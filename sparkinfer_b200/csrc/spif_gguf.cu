// spif_gguf.cu -- the MODEL file of the path (GGUF): index of its hyper-parameters and tensors, row-range reads.  Host code only.
//
// What the reference reads from the model GGUF for the sparse-FFN path:
//   KV   general.architecture (prosparse-llama, bamboo, ...: src/llama-arch.cpp:20-110), {arch}.block_count, {arch}.embedding_length,
//        {arch}.feed_forward_length, {arch}.attention.layer_norm_rms_epsilon, {arch}.pred_lora = predictor rank per layer
//        (src/llama-arch.cpp:151, src/llama-model.cpp:661; writer gguf-py/gguf/gguf_writer.py:1168)
//   blk.N.ffn_up / ffn_gate .weight {n_embd, n_ff};  blk.N.ffn_down.weight stored TRANSPOSED {n_embd, n_ff} (one row per neuron,
//        convert_hf_to_gguf.py:4640-4648);  blk.N.ffn_pred_up.weight {n_embd, r}, blk.N.ffn_pred_down.weight {r, n_ff} and
//        their optional .bias (src/llama-model.cpp:2759-2768);  blk.N.ffn_norm.weight {n_embd}
// The container is GGUF v2/v3 (ggml/src/gguf.cpp:320-640).  The file is mapped read-only; nothing is copied until
// spif_gguf_read, which copies a ROW RANGE of one tensor into caller memory (pinned staging for the H2D copy): a GPU shard
// reads only the neuron rows it holds.  Errors are return codes (SPIF_E_SHAPE: malformed / truncated file), never aborts.
#include "spif_common.cuh"

#include <cstring>
#include <fcntl.h>
#include <map>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <vector>

namespace {

enum : uint32_t { KT_U8, KT_I8, KT_U16, KT_I16, KT_U32, KT_I32, KT_F32, KT_BOOL, KT_STR, KT_ARR, KT_U64, KT_I64, KT_F64, KT_COUNT };
const size_t k_kt_size[KT_COUNT] = { 1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8 };

struct Cursor {  // bounds-checked walk over the mapping
    const unsigned char * b;
    size_t                size, pos = 0;
    bool                  ok = true;
    Cursor(const unsigned char * p, size_t n) : b(p), size(n) {}
    bool has(size_t n) const { return ok && n <= size - pos; }
    template <typename T> T get() {
        T v = T();
        if (!has(sizeof(T))) {
            ok = false;
            return v;
        }
        memcpy(&v, b + pos, sizeof(T));
        pos += sizeof(T);
        return v;
    }
    std::string str() {
        const uint64_t n = get<uint64_t>();
        if (!has((size_t) n) || n > (1u << 24)) {
            ok = false;
            return std::string();
        }
        std::string s((const char *) b + pos, (size_t) n);
        pos += (size_t) n;
        return s;
    }
    void skip(size_t n) {
        if (!has(n)) {
            ok = false;
            return;
        }
        pos += n;
    }
};

struct KvRef {      // where a value sits in the mapping
    uint32_t type;  // KT_*; for arrays: the element type, with is_array set
    bool     is_array;
    uint64_t n;     // array length (1 for scalars)
    size_t   pos;   // first value byte (strings: the u64 length)
};

// bytes of one row of ne0 elements; 0 = a type the path does not read
size_t row_bytes_of(int32_t type, int64_t ne0) {
    if (type == 26 /* GGML_TYPE_I32 */) {
        return (size_t) ne0 * 4;
    }
    if ((type == SPIF_TYPE_Q8_0 || type == SPIF_TYPE_Q4_0) && ne0 % SPIF_QK != 0) {
        return 0;  // block formats hold whole blocks of 32 per row (ggml_row_size asserts it)
    }
    return spif_row_size(type, ne0);
}

}  // namespace

struct spif_gguf {
    int                              fd   = -1;
    const unsigned char *            map  = nullptr;
    size_t                           size = 0;
    std::map<std::string, KvRef>     kv;
    std::vector<std::string>         names;
    std::vector<spif_gguf_tensor>    tensors;
    std::map<std::string, int>       by_name;
    std::string                      arch;

    ~spif_gguf() {
        if (map) {
            munmap((void *) map, size);
        }
        if (fd >= 0) {
            close(fd);
        }
    }
    // scalar integer / float KV as double; false when absent or not numeric
    bool number(const std::string & key, double * out, uint64_t index = 0) const {
        auto it = kv.find(key);
        if (it == kv.end() || index >= it->second.n || it->second.type == KT_STR || it->second.type == KT_ARR) {
            return false;
        }
        const KvRef &         r = it->second;
        const unsigned char * p = map + r.pos + index * k_kt_size[r.type];
        switch (r.type) {
            case KT_U8: *out = *(const uint8_t *) p; break;
            case KT_I8: *out = *(const int8_t *) p; break;
            case KT_BOOL: *out = *(const uint8_t *) p ? 1 : 0; break;
            case KT_U16: { uint16_t v; memcpy(&v, p, 2); *out = v; } break;
            case KT_I16: { int16_t v; memcpy(&v, p, 2); *out = v; } break;
            case KT_U32: { uint32_t v; memcpy(&v, p, 4); *out = v; } break;
            case KT_I32: { int32_t v; memcpy(&v, p, 4); *out = v; } break;
            case KT_F32: { float v; memcpy(&v, p, 4); *out = v; } break;
            case KT_U64: { uint64_t v; memcpy(&v, p, 8); *out = (double) v; } break;
            case KT_I64: { int64_t v; memcpy(&v, p, 8); *out = (double) v; } break;
            case KT_F64: { double v; memcpy(&v, p, 8); *out = v; } break;
            default: return false;
        }
        return true;
    }
};

extern "C" {

int spif_gguf_open(const char * path, spif_gguf ** out) {
    SPIF_REQUIRE(path && out, SPIF_E_NULL);
    *out     = nullptr;
    auto * g = new spif_gguf();
    g->fd    = open(path, O_RDONLY);
    struct stat st;
    if (g->fd < 0 || fstat(g->fd, &st) != 0 || st.st_size < 24) {
        const int rc = g->fd < 0 ? SPIF_E_NULL : SPIF_E_SHAPE;
        delete g;
        return rc;
    }
    g->size = (size_t) st.st_size;
    void * m = mmap(nullptr, g->size, PROT_READ, MAP_PRIVATE, g->fd, 0);
    if (m == MAP_FAILED) {
        delete g;
        return SPIF_E_NULL;
    }
    g->map = (const unsigned char *) m;
    Cursor c(g->map, g->size);
    if (memcmp(g->map, "GGUF", 4) != 0) {
        delete g;
        return SPIF_E_SHAPE;
    }
    c.skip(4);
    const uint32_t version = c.get<uint32_t>();
    if (!c.ok || version < 2 || version > 3) {  // v1 had 32-bit counts; gguf.cpp:368-374 rejects it too
        delete g;
        return SPIF_E_UNSUPPORTED;
    }
    const uint64_t n_tensors = c.get<uint64_t>();
    const uint64_t n_kv      = c.get<uint64_t>();
    if (!c.ok || n_tensors > (1u << 22) || n_kv > (1u << 22)) {
        delete g;
        return SPIF_E_SHAPE;
    }
    for (uint64_t i = 0; i < n_kv && c.ok; ++i) {
        const std::string key  = c.str();
        const uint32_t    type = c.get<uint32_t>();
        if (!c.ok || type >= KT_COUNT) {
            c.ok = false;
            break;
        }
        KvRef r = { type, false, 1, c.pos };
        if (type == KT_STR) {
            c.str();
        } else if (type == KT_ARR) {
            const uint32_t et = c.get<uint32_t>();
            const uint64_t n  = c.get<uint64_t>();
            if (!c.ok || et >= KT_COUNT || et == KT_ARR) {
                c.ok = false;
                break;
            }
            r = { et, true, n, c.pos };
            if (et == KT_STR) {
                for (uint64_t k = 0; k < n && c.ok; ++k) {
                    c.str();
                }
            } else {
                if (n > (c.size - c.pos) / k_kt_size[et]) {
                    c.ok = false;
                    break;
                }
                c.skip((size_t) n * k_kt_size[et]);
            }
        } else {
            c.skip(k_kt_size[type]);
        }
        g->kv[key] = r;
    }
    g->tensors.reserve((size_t) n_tensors);
    for (uint64_t i = 0; i < n_tensors && c.ok; ++i) {
        std::string      name = c.str();
        spif_gguf_tensor t    = {};
        t.n_dims              = (int32_t) c.get<uint32_t>();
        if (!c.ok || t.n_dims < 0 || t.n_dims > 4) {
            c.ok = false;
            break;
        }
        for (int d = 0; d < 4; ++d) {
            t.ne[d] = 1;
        }
        for (int d = 0; d < t.n_dims; ++d) {
            const uint64_t v = c.get<uint64_t>();
            if (v == 0 || v > ((uint64_t) 1 << 40)) {
                c.ok = false;
            }
            t.ne[d] = (int64_t) v;
        }
        t.type   = (int32_t) c.get<uint32_t>();
        t.offset = c.get<uint64_t>();  // relative to the data section for now
        if (!c.ok) {
            break;
        }
        g->by_name[name] = (int) g->tensors.size();
        g->names.push_back(std::move(name));
        g->tensors.push_back(t);
    }
    double alignment = 32;  // GGUF_DEFAULT_ALIGNMENT
    g->number("general.alignment", &alignment);
    const uint64_t al = (uint64_t) alignment;
    if (!c.ok || al == 0 || (al & (al - 1))) {
        delete g;
        return SPIF_E_SHAPE;
    }
    const uint64_t data_start = (c.pos + al - 1) / al * al;
    for (spif_gguf_tensor & t : g->tensors) {
        const size_t rb   = row_bytes_of(t.type, t.ne[0]);
        uint64_t     rows = 1, nbytes = 0;
        bool         wraps = false;
        for (int d = 1; d < 4; ++d) {
            wraps |= __builtin_mul_overflow(rows, (uint64_t) t.ne[d], &rows);
        }
        wraps |= __builtin_mul_overflow(rows, (uint64_t) rb, &nbytes);
        if (wraps) {
            delete g;
            return SPIF_E_SHAPE;
        }
        t.nbytes = nbytes;  // 0: a type the path does not read
        t.offset += data_start;
        if (t.offset > g->size || t.nbytes > g->size - t.offset) {
            delete g;
            return SPIF_E_SHAPE;  // truncated file
        }
    }
    {
        auto it = g->kv.find("general.architecture");
        if (it != g->kv.end() && it->second.type == KT_STR && !it->second.is_array) {
            Cursor s(g->map, g->size);
            s.pos   = it->second.pos;
            g->arch = s.str();
        }
    }
    *out = g;
    return 0;
}

void spif_gguf_close(spif_gguf * g) {
    delete g;
}

// arch_out: general.architecture, NUL-terminated (truncated to arch_cap); n_ff: {arch}.feed_forward_length, 0 when it differs
// between layers (per-layer arrays); rms_eps: 0 when the key is absent.  Absent integer keys give SPIF_E_SHAPE.
int spif_gguf_hparams(const spif_gguf * g, char * arch_out, int arch_cap, int32_t * n_layer, int32_t * n_embd, int32_t * n_ff,
                      float * rms_eps) {
    SPIF_REQUIRE(g, SPIF_E_NULL);
    SPIF_REQUIRE(!g->arch.empty(), SPIF_E_SHAPE);
    if (arch_out && arch_cap > 0) {
        snprintf(arch_out, (size_t) arch_cap, "%s", g->arch.c_str());
    }
    double v = 0;
    if (n_layer) {
        SPIF_REQUIRE(g->number(g->arch + ".block_count", &v), SPIF_E_SHAPE);
        *n_layer = (int32_t) v;
    }
    if (n_embd) {
        SPIF_REQUIRE(g->number(g->arch + ".embedding_length", &v), SPIF_E_SHAPE);
        *n_embd = (int32_t) v;
    }
    if (n_ff) {
        const std::string key = g->arch + ".feed_forward_length";
        SPIF_REQUIRE(g->number(key, &v), SPIF_E_SHAPE);
        *n_ff          = (int32_t) v;
        const KvRef & r = g->kv.at(key);
        for (uint64_t i = 1; r.is_array && i < r.n; ++i) {
            double w = 0;
            g->number(key, &w, i);
            if (w != v) {
                *n_ff = 0;
            }
        }
    }
    if (rms_eps) {
        *rms_eps = g->number(g->arch + ".attention.layer_norm_rms_epsilon", &v) ? (float) v : 0.f;
    }
    return 0;
}

// ranks_out[l] = {arch}.pred_lora[l] (0: the layer has no predictor, it runs dense -- llama-graph.cpp:932-937); a file without
// the key (a stock model) gives all zeros, like the reference's std::fill default (src/llama-model.cpp:535)
int spif_gguf_pred_rank(const spif_gguf * g, int32_t * ranks_out, int n_layer) {
    SPIF_REQUIRE(g && ranks_out, SPIF_E_NULL);
    SPIF_REQUIRE(n_layer > 0, SPIF_E_SHAPE);
    const std::string key = g->arch + ".pred_lora";
    auto              it  = g->kv.find(key);
    for (int l = 0; l < n_layer; ++l) {
        ranks_out[l] = 0;
    }
    if (it == g->kv.end()) {
        return 0;
    }
    SPIF_REQUIRE(it->second.n == 1 || it->second.n >= (uint64_t) n_layer, SPIF_E_SHAPE);
    for (int l = 0; l < n_layer; ++l) {
        double v = 0;
        SPIF_REQUIRE(g->number(key, &v, it->second.n == 1 ? 0 : (uint64_t) l), SPIF_E_SHAPE);  // get_key_or_arr: scalar = all layers
        ranks_out[l] = (int32_t) v;
    }
    return 0;
}

int spif_gguf_n_tensors(const spif_gguf * g) {
    return g ? (int) g->tensors.size() : SPIF_E_NULL;
}

int spif_gguf_tensor_at(const spif_gguf * g, int i, char * name_out, int name_cap, spif_gguf_tensor * out) {
    SPIF_REQUIRE(g, SPIF_E_NULL);
    SPIF_REQUIRE(i >= 0 && i < (int) g->tensors.size(), SPIF_E_SHAPE);
    if (name_out && name_cap > 0) {
        snprintf(name_out, (size_t) name_cap, "%s", g->names[i].c_str());
    }
    if (out) {
        *out = g->tensors[i];
    }
    return 0;
}

// SPIF_E_SHAPE when the file has no tensor of that name (optional tensors: biases, ffn_gate of OPT / Falcon)
int spif_gguf_find(const spif_gguf * g, const char * name, spif_gguf_tensor * out) {
    SPIF_REQUIRE(g && name && out, SPIF_E_NULL);
    auto it = g->by_name.find(name);
    SPIF_REQUIRE(it != g->by_name.end(), SPIF_E_SHAPE);
    *out = g->tensors[it->second];
    return 0;
}

// rows [row0, row0 + n_rows) of tensor t (as returned by _find / _tensor_at of THIS file) -> dst_host.  A row is ne[0] elements;
// a 1-D tensor is ne[0] rows of one element (bias and norm vectors are sliced by neuron like the matrices).
int spif_gguf_read(const spif_gguf * g, const spif_gguf_tensor * t, int64_t row0, int64_t n_rows, void * dst, size_t dst_bytes) {
    SPIF_REQUIRE(g && t && dst, SPIF_E_NULL);
    const bool    vec   = t->n_dims <= 1;
    const int64_t rows  = vec ? t->ne[0] : t->ne[1] * t->ne[2] * t->ne[3];
    const size_t  rb    = row_bytes_of(t->type, vec ? 1 : t->ne[0]);
    SPIF_REQUIRE(rb != 0 && t->nbytes != 0, SPIF_E_TYPE);
    SPIF_REQUIRE(!(vec && t->type != 0 && t->type != 1 && t->type != 30 && t->type != 26), SPIF_E_TYPE);  // no block types by element
    SPIF_REQUIRE(row0 >= 0 && n_rows >= 0 && row0 <= rows && n_rows <= rows - row0, SPIF_E_SHAPE);
    SPIF_REQUIRE(dst_bytes >= (size_t) n_rows * rb, SPIF_E_WORKSPACE);
    SPIF_REQUIRE(t->offset <= g->size && t->nbytes <= g->size - t->offset && (uint64_t) rows * rb == t->nbytes, SPIF_E_SHAPE);
    memcpy(dst, g->map + t->offset + (size_t) row0 * rb, (size_t) n_rows * rb);
    return 0;
}

// rows[i] of tensor t -> row i of dst_host, for a row list in any order (a model-split permutation, a shard's neuron_idx):
// what spif_gguf_read does for a range, row by row in one call
int spif_gguf_gather(const spif_gguf * g, const spif_gguf_tensor * t, const int32_t * rows_idx, int64_t n_rows, void * dst,
                     size_t dst_bytes) {
    SPIF_REQUIRE(g && t && dst && rows_idx, SPIF_E_NULL);
    const bool    vec  = t->n_dims <= 1;
    const int64_t rows = vec ? t->ne[0] : t->ne[1] * t->ne[2] * t->ne[3];
    const size_t  rb   = row_bytes_of(t->type, vec ? 1 : t->ne[0]);
    SPIF_REQUIRE(rb != 0 && t->nbytes != 0, SPIF_E_TYPE);
    SPIF_REQUIRE(n_rows >= 0, SPIF_E_SHAPE);
    SPIF_REQUIRE(dst_bytes >= (size_t) n_rows * rb, SPIF_E_WORKSPACE);
    SPIF_REQUIRE(t->offset <= g->size && t->nbytes <= g->size - t->offset && (uint64_t) rows * rb == t->nbytes, SPIF_E_SHAPE);
    for (int64_t i = 0; i < n_rows; ++i) {
        SPIF_REQUIRE(rows_idx[i] >= 0 && rows_idx[i] < rows, SPIF_E_SHAPE);
    }
    const unsigned char * src = g->map + t->offset;
    unsigned char *       out = (unsigned char *) dst;
    for (int64_t i = 0; i < n_rows; ++i) {
        memcpy(out + (size_t) i * rb, src + (size_t) rows_idx[i] * rb, rb);
    }
    return 0;
}

}  // extern "C"
