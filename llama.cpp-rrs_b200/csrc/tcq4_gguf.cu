// tcq4_gguf.cu -- the caller side of the path: GGUF files carrying GGML_TYPE_TCQ4_K32 tensors and their
// channel permutations, read and written on the host.
//
// Replaces, for this path only:
//   * the loader's scan for "tcq4.reorder.enabled" / "tcq4.<tensor>.perm" and the registration of the permutations
//     with the GPU backend (/root/reference/src/llama-model.cpp:7129-7205) -- here per DEVICE, not "first GPU only";
//   * the upload of a tensor's bytes into a device buffer (ggml_backend_cuda_buffer_set_tensor,
//     /root/reference/ggml/src/ggml-cuda/ggml-cuda.cu:627-633: device bytes == file bytes);
//   * the writer side of llama_model_quantize for these keys (/root/reference/src/llama-quant.cpp:840-854).
// File format: GGUF v3 (/root/reference/ggml/src/gguf.cpp, ggml/include/gguf.h): header {magic "GGUF", u32 version,
// u64 n_tensors, u64 n_kv}, key/value pairs, tensor infos {name, n_dims, dims, type, offset}, padding to
// general.alignment (default 32), tensor data.  Little-endian hosts only, like the reference.
#include "../../include/tcq4_b200.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

enum GgufType : uint32_t {
    GT_U8 = 0, GT_I8, GT_U16, GT_I16, GT_U32, GT_I32, GT_F32, GT_BOOL, GT_STRING, GT_ARRAY, GT_U64, GT_I64, GT_F64, GT_COUNT
};
const size_t kScalarSize[GT_COUNT] = {1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8};

struct TensorInfo {
    std::string name;
    int         type = 0;
    int64_t     ne[4] = {1, 1, 1, 1};
    uint64_t    offset = 0;     // absolute file offset
    uint64_t    nbytes = 0;     // exact for F32 / F16 / TCQ4_K32, else the extent up to the next tensor
};
struct PermInfo {
    std::string          tensor;
    std::vector<int32_t> perm;
};

struct Reader {
    FILE  *f = nullptr;
    bool   ok = true;
    bool   rd(void *p, size_t n) {
        if (!ok) return false;
        if (n && fread(p, 1, n, f) != n) ok = false;
        return ok;
    }
    template <typename T> T get() {
        T v{};
        rd(&v, sizeof(T));
        return v;
    }
    std::string str() {
        const uint64_t n = get<uint64_t>();
        if (!ok || n > (1u << 24)) { ok = false; return {}; }
        std::string s(n, '\0');
        rd(&s[0], n);
        return s;
    }
    void skip(uint64_t n) {
        if (ok && fseeko(f, (off_t)n, SEEK_CUR) != 0) ok = false;
    }
};

size_t exact_bytes(int type, const int64_t ne[4]) {
    const uint64_t rows = (uint64_t)ne[1] * (uint64_t)ne[2] * (uint64_t)ne[3];
    switch (type) {
        case 0: return (size_t)(4 * (uint64_t)ne[0] * rows);                                   // F32
        case 1: return (size_t)(2 * (uint64_t)ne[0] * rows);                                   // F16
        case TCQ4_B200_GGML_TYPE: return (size_t)(((uint64_t)ne[0] / 256) * 148 * rows);       // ggml.c:750-757
        default: return 0;
    }
}

}  // namespace

struct tcq4_gguf {
    std::string             path;
    uint32_t                version = 0, alignment = 32;
    int                     reorder = -1;              // -1: key absent
    uint64_t                data_offset = 0, file_size = 0;
    std::vector<TensorInfo> tensors;
    std::vector<PermInfo>   perms;
};

extern "C" {

int tcq4_b200_gguf_open(const char *path, tcq4_gguf **out) {
    if (!path || !out) return TCQ4_ERR_ARG;
    *out = nullptr;
    Reader r;
    r.f = fopen(path, "rb");
    if (!r.f) return TCQ4_ERR_ARG;
    tcq4_gguf *g = new tcq4_gguf;
    g->path = path;
    char magic[4] = {0, 0, 0, 0};
    r.rd(magic, 4);
    g->version = r.get<uint32_t>();
    const uint64_t n_tensors = r.get<uint64_t>(), n_kv = r.get<uint64_t>();
    if (!r.ok || memcmp(magic, "GGUF", 4) != 0 || g->version < 2 || g->version > 3 || n_tensors > (1u << 24) ||
        n_kv > (1u << 24)) {
        fclose(r.f);
        delete g;
        return TCQ4_ERR_ARG;
    }
    for (uint64_t i = 0; i < n_kv && r.ok; i++) {
        const std::string key = r.str();
        const uint32_t type = r.get<uint32_t>();
        if (!r.ok || type >= GT_COUNT) { r.ok = false; break; }
        if (type == GT_STRING) {
            (void)r.str();
        } else if (type == GT_ARRAY) {
            const uint32_t et = r.get<uint32_t>();
            const uint64_t n = r.get<uint64_t>();
            if (!r.ok || et >= GT_COUNT || et == GT_ARRAY) { r.ok = false; break; }
            // "tcq4.<tensor>.perm", int32 array (llama-model.cpp:7149-7166: prefix, last '.', length > 10)
            const size_t dot = key.rfind('.');
            const bool is_perm = et == GT_I32 && key.size() > 10 && key.compare(0, 5, "tcq4.") == 0 &&
                                 dot != std::string::npos && key.compare(dot, std::string::npos, ".perm") == 0 && dot >= 5;
            if (is_perm && n <= (1u << 30)) {
                PermInfo p;
                p.tensor = key.substr(5, dot - 5);
                p.perm.resize(n);
                r.rd(p.perm.data(), n * 4);
                g->perms.push_back(std::move(p));
            } else if (et == GT_STRING) {
                for (uint64_t j = 0; j < n && r.ok; j++) (void)r.str();
            } else {
                r.skip(n * kScalarSize[et]);
            }
        } else {
            uint8_t buf[8] = {0};
            r.rd(buf, kScalarSize[type]);
            if (key == "tcq4.reorder.enabled" && type == GT_BOOL) g->reorder = buf[0] ? 1 : 0;
            if (key == "general.alignment" && type == GT_U32) memcpy(&g->alignment, buf, 4);
        }
    }
    if (g->alignment == 0 || (g->alignment & (g->alignment - 1))) r.ok = false;
    g->tensors.resize(r.ok ? n_tensors : 0);
    for (uint64_t i = 0; i < n_tensors && r.ok; i++) {
        TensorInfo &t = g->tensors[i];
        t.name = r.str();
        const uint32_t nd = r.get<uint32_t>();
        if (!r.ok || nd > 4) { r.ok = false; break; }
        for (uint32_t d = 0; d < nd; d++) t.ne[d] = (int64_t)r.get<uint64_t>();
        t.type = (int)r.get<uint32_t>();
        t.offset = r.get<uint64_t>();          // relative to the data section for now
    }
    if (r.ok) {
        const uint64_t pos = (uint64_t)ftello(r.f);
        g->data_offset = (pos + g->alignment - 1) / g->alignment * g->alignment;
        fseeko(r.f, 0, SEEK_END);
        g->file_size = (uint64_t)ftello(r.f);
        for (TensorInfo &t : g->tensors) t.offset += g->data_offset;
        for (TensorInfo &t : g->tensors) {
            uint64_t end = g->file_size;                                 // extent: up to the next tensor in the file
            for (const TensorInfo &o : g->tensors)
                if (o.offset > t.offset && o.offset < end) end = o.offset;
            const size_t ex = exact_bytes(t.type, t.ne);
            t.nbytes = ex ? ex : (end > t.offset ? end - t.offset : 0);
            if (t.offset + t.nbytes > g->file_size) r.ok = false;       // truncated file
            if (t.type == TCQ4_B200_GGML_TYPE && (t.ne[0] % 256) != 0) r.ok = false;
        }
    }
    fclose(r.f);
    if (!r.ok) {
        delete g;
        return TCQ4_ERR_ARG;
    }
    *out = g;
    return TCQ4_OK;
}

void tcq4_b200_gguf_close(tcq4_gguf *g) { delete g; }

int tcq4_b200_gguf_reorder_enabled(const tcq4_gguf *g) { return g ? g->reorder : -1; }
int64_t tcq4_b200_gguf_n_tensors(const tcq4_gguf *g) { return g ? (int64_t)g->tensors.size() : 0; }
int64_t tcq4_b200_gguf_n_perms(const tcq4_gguf *g) { return g ? (int64_t)g->perms.size() : 0; }

int tcq4_b200_gguf_tensor_info(const tcq4_gguf *g, int64_t i, const char **name, int *type, int64_t ne[4],
                               uint64_t *file_offset, uint64_t *nbytes) {
    if (!g || i < 0 || i >= (int64_t)g->tensors.size()) return TCQ4_ERR_ARG;
    const TensorInfo &t = g->tensors[(size_t)i];
    if (name) *name = t.name.c_str();
    if (type) *type = t.type;
    if (ne) memcpy(ne, t.ne, sizeof(t.ne));
    if (file_offset) *file_offset = t.offset;
    if (nbytes) *nbytes = t.nbytes;
    return TCQ4_OK;
}

int tcq4_b200_gguf_perm(const tcq4_gguf *g, int64_t i, const char **tensor_name, const int32_t **perm, int64_t *K) {
    if (!g || i < 0 || i >= (int64_t)g->perms.size()) return TCQ4_ERR_ARG;
    const PermInfo &p = g->perms[(size_t)i];
    if (tensor_name) *tensor_name = p.tensor.c_str();
    if (perm) *perm = p.perm.data();
    if (K) *K = (int64_t)p.perm.size();
    return TCQ4_OK;
}

// llama-model.cpp:7176-7197 for the CURRENT device: set the global enable flag (when the file has the key), then register
// every permutation under its tensor name.  Call once per device that will run the model.
int tcq4_b200_gguf_register_perms(const tcq4_gguf *g) {
    if (!g) return TCQ4_ERR_ARG;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return TCQ4_ERR_NO_DEVICE;
    if (g->reorder >= 0) ggml_backend_cuda_tcq4_set_reorder_enabled(g->reorder != 0);
    for (const PermInfo &p : g->perms)
        ggml_backend_cuda_tcq4_register_perm(p.tensor.c_str(), p.perm.data(), (int)p.perm.size());
    return TCQ4_OK;
}

// Read tensor i into HOST memory (dst_bytes >= nbytes).
int tcq4_b200_gguf_read_tensor(const tcq4_gguf *g, int64_t i, void *dst_host, size_t dst_bytes) {
    if (!g || !dst_host || i < 0 || i >= (int64_t)g->tensors.size()) return TCQ4_ERR_ARG;
    const TensorInfo &t = g->tensors[(size_t)i];
    if (dst_bytes < t.nbytes) return TCQ4_ERR_ARG;
    FILE *f = fopen(g->path.c_str(), "rb");
    if (!f) return TCQ4_ERR_ARG;
    const bool ok = fseeko(f, (off_t)t.offset, SEEK_SET) == 0 && fread(dst_host, 1, t.nbytes, f) == t.nbytes;
    fclose(f);
    return ok ? TCQ4_OK : TCQ4_ERR_ARG;
}

// Upload tensor i verbatim into device memory (device bytes == file bytes, ggml-cuda.cu:627-633) through a pinned
// staging buffer, 16 MB at a time, on `stream`; returns after the last chunk has been copied.
int tcq4_b200_gguf_load_tensor(const tcq4_gguf *g, int64_t i, void *dst_device, size_t dst_bytes, cudaStream_t stream) {
    if (!g || !dst_device || i < 0 || i >= (int64_t)g->tensors.size()) return TCQ4_ERR_ARG;
    const TensorInfo &t = g->tensors[(size_t)i];
    if (dst_bytes < t.nbytes) return TCQ4_ERR_ARG;
    FILE *f = fopen(g->path.c_str(), "rb");
    if (!f) return TCQ4_ERR_ARG;
    const size_t chunk = 16u << 20;
    void *stage[2] = {nullptr, nullptr};
    cudaEvent_t done[2] = {nullptr, nullptr};
    int rc = TCQ4_OK;
    for (int b = 0; b < 2 && rc == TCQ4_OK; b++)
        if (cudaMallocHost(&stage[b], chunk) != cudaSuccess || cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming) != cudaSuccess)
            rc = TCQ4_ERR_CUDA;
    if (rc == TCQ4_OK && fseeko(f, (off_t)t.offset, SEEK_SET) != 0) rc = TCQ4_ERR_ARG;
    size_t pos = 0;
    for (int b = 0; rc == TCQ4_OK && pos < t.nbytes; b ^= 1) {
        const size_t n = t.nbytes - pos < chunk ? t.nbytes - pos : chunk;
        cudaEventSynchronize(done[b]);                                   // the copy that last used this buffer has finished
        if (fread(stage[b], 1, n, f) != n) { rc = TCQ4_ERR_ARG; break; }
        if (cudaMemcpyAsync((char *)dst_device + pos, stage[b], n, cudaMemcpyHostToDevice, stream) != cudaSuccess ||
            cudaEventRecord(done[b], stream) != cudaSuccess)
            rc = TCQ4_ERR_CUDA;
        pos += n;
    }
    if (cudaStreamSynchronize(stream) != cudaSuccess && rc == TCQ4_OK) rc = TCQ4_ERR_CUDA;
    for (int b = 0; b < 2; b++) {
        if (done[b]) cudaEventDestroy(done[b]);
        if (stage[b]) cudaFreeHost(stage[b]);
    }
    fclose(f);
    return rc;
}

// ---- writer: what llama_model_quantize emits for this path (llama-quant.cpp:840-854 + gguf_write_to_file) ----------
}  // extern "C"
static void put(FILE *f, const void *p, size_t n) { fwrite(p, 1, n, f); }
template <typename T> static void putv(FILE *f, T v) { put(f, &v, sizeof(T)); }
static void puts_(FILE *f, const char *s) {
    const uint64_t n = strlen(s);
    putv<uint64_t>(f, n);
    put(f, s, n);
}
extern "C" {

int tcq4_b200_gguf_write(const char *path, int n_tensors, const char *const *names, const int *types, const int64_t *ne0,
                         const int64_t *ne1, const void *const *data_host, int n_perms, const char *const *perm_tensors,
                         const int32_t *const *perms, const int64_t *perm_K, int reorder_enabled) {
    if (!path || n_tensors < 0 || n_perms < 0 || (n_tensors && (!names || !types || !ne0 || !ne1 || !data_host)) ||
        (n_perms && (!perm_tensors || !perms || !perm_K)))
        return TCQ4_ERR_ARG;
    const uint32_t alignment = 32;
    std::vector<uint64_t> sizes((size_t)n_tensors), offs((size_t)n_tensors);
    uint64_t off = 0;
    for (int i = 0; i < n_tensors; i++) {
        const int64_t ne[4] = {ne0[i], ne1[i], 1, 1};
        sizes[i] = exact_bytes(types[i], ne);
        if (sizes[i] == 0 || !names[i] || !data_host[i]) return TCQ4_ERR_ARG;      // F32, F16 and TCQ4_K32 only
        if (types[i] == TCQ4_B200_GGML_TYPE && ((ne0[i] % 256) != 0 || (ne1[i] % 8) != 0)) return TCQ4_ERR_SHAPE;
        offs[i] = off;
        off = (off + sizes[i] + alignment - 1) / alignment * alignment;
    }
    FILE *f = fopen(path, "wb");
    if (!f) return TCQ4_ERR_ARG;
    put(f, "GGUF", 4);
    putv<uint32_t>(f, 3);
    putv<uint64_t>(f, (uint64_t)n_tensors);
    putv<uint64_t>(f, (uint64_t)n_perms + (reorder_enabled >= 0 ? 1 : 0) + 1);
    puts_(f, "general.alignment");
    putv<uint32_t>(f, GT_U32);
    putv<uint32_t>(f, alignment);
    if (reorder_enabled >= 0) {
        puts_(f, "tcq4.reorder.enabled");
        putv<uint32_t>(f, GT_BOOL);
        putv<uint8_t>(f, reorder_enabled ? 1 : 0);
    }
    for (int i = 0; i < n_perms; i++) {
        const std::string key = std::string("tcq4.") + perm_tensors[i] + ".perm";
        puts_(f, key.c_str());
        putv<uint32_t>(f, GT_ARRAY);
        putv<uint32_t>(f, GT_I32);
        putv<uint64_t>(f, (uint64_t)perm_K[i]);
        put(f, perms[i], (size_t)perm_K[i] * 4);
    }
    for (int i = 0; i < n_tensors; i++) {
        puts_(f, names[i]);
        putv<uint32_t>(f, 2);
        putv<uint64_t>(f, (uint64_t)ne0[i]);
        putv<uint64_t>(f, (uint64_t)ne1[i]);
        putv<uint32_t>(f, (uint32_t)types[i]);
        putv<uint64_t>(f, offs[i]);
    }
    const uint8_t zero[64] = {0};
    uint64_t pos = (uint64_t)ftello(f);
    put(f, zero, (size_t)((alignment - pos % alignment) % alignment));
    for (int i = 0; i < n_tensors; i++) {
        put(f, data_host[i], sizes[i]);
        put(f, zero, (size_t)((alignment - sizes[i] % alignment) % alignment));
    }
    const bool ok = ferror(f) == 0;
    fclose(f);
    return ok ? TCQ4_OK : TCQ4_ERR_ARG;
}

}  // extern "C"

// ---- which tensors of a model become TCQ4_K32 (quantiser side) ------------------------------------------------
// llama_tensor_get_type for ftype LLAMA_FTYPE_MOSTLY_TCQ4_K32 (src/llama-quant.cpp:204-526; default type :590) restated for the
// tensor classes it distinguishes.  Returns the ggml_type id the reference would store the tensor as.
extern "C" TCQ4_B200_API int tcq4_b200_quant_type_for(const char *name, int n_dims, int64_t K, int64_t N, int has_output) {
    enum { T_F16 = 1, T_Q5_0 = 6, T_Q8_0 = 8, T_Q4_K = 12, T_Q6_K = 14, T_TCQ4 = 42 };
    const std::string nm = name ? name : "";
    const bool is_weight = nm.size() >= 6 && nm.compare(nm.size() - 6, 6, "weight") == 0;
    constexpr int UNCHANGED = -2;                                // the tensor keeps the type it has in the source model
    if (n_dims < 2 || !is_weight) return UNCHANGED;              // only 2-D+ "weight" tensors are quantised (:885-888)
    if (nm.find("_norm.weight") != std::string::npos || nm.find("ffn_gate_inp.weight") != std::string::npos)
        return UNCHANGED;                                        // norms and the MoE router are not quantised (:891, :898)
    int t = T_TCQ4;
    const int64_t blck = 256;                                    // ggml_blck_size(TCQ4_K32) = TCQ4_TILE_K (ggml.c:750-753)
    if (nm == "output.weight" || (!has_output && nm == "token_embd.weight")) {
        // :232-253: the output matrix gets more bits; K is tested against the DEFAULT type's block size (256)
        t = (K % blck != 0) ? T_Q8_0 : T_Q6_K;
    } else if (nm == "token_embd.weight" || nm == "per_layer_token_embd.weight") {
        t = T_Q4_K;                                              // :280-285: the FWHT domain does not suit GET_ROWS consumers
    }
    // :468-497: shape constraints of the chosen type
    bool incompatible;
    if (t == T_TCQ4) incompatible = (K % 256 != 0) || (N % 8 != 0);
    else incompatible = K % ((t == T_Q8_0) ? 32 : 256) != 0;
    if (incompatible) {                                          // :499-523
        switch (t) {
            case T_TCQ4: t = T_Q4_K; break;
            case T_Q4_K: t = T_Q5_0; break;
            case T_Q6_K: t = T_Q8_0; break;
            default: return -1;                                  // "Unsupported tensor size encountered"
        }
        const int64_t b2 = (t == T_Q4_K) ? 256 : 32;
        if (K % b2 != 0) t = T_F16;
    }
    return t;
}
