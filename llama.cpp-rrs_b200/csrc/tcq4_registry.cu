// tcq4_registry.cu -- channel-permutation registry and the host-side mirror of the ggml operator
// entry, i.e. the rest of the fork's plug-in surface for GGML_TYPE_TCQ4_K32:
//   ggml_cuda_rrs_register_perm / get_perm / has_perm / clear_perms / set/get_reorder_enabled
//       /root/reference/ggml/src/ggml-cuda/rrs.cu:82-127
//   ggml_backend_cuda_tcq4_* wrappers resolved by name from libllama
//       /root/reference/ggml/src/ggml-cuda/ggml-cuda.cu:4148-4158, 4876-4885
//   ggml_cuda_rrs_mul_mat shape extraction and asserts
//       /root/reference/ggml/src/ggml-cuda/rrs.cu:241-270
//
// Differences from the reference, on purpose: the registry is keyed by (device, name)
// and mutex-guarded, CUDA errors are checked, and lookups on the launch path take a shared lock and
// do not allocate (the reference builds a std::string and hashes it per call, rrs.cu:104-106).
#include "../../include/tcq4_b200.h"

#include <cuda_runtime.h>

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <utility>

namespace {

struct PermEntry {
    int32_t *d_perm = nullptr;
    int      K = 0;
    std::vector<int32_t> h;      // host copy: the fan-out mirror compares permutations of different tensors
};
struct KeyLess {
    using is_transparent = void;
    using Key = std::pair<int, std::string>;
    using View = std::pair<int, const char *>;
    bool operator()(const Key &a, const Key &b) const { return a < b; }
    bool operator()(const Key &a, const View &b) const {
        return a.first != b.first ? a.first < b.first : std::strcmp(a.second.c_str(), b.second) < 0;
    }
    bool operator()(const View &a, const Key &b) const {
        return a.first != b.first ? a.first < b.first : std::strcmp(a.second, b.second.c_str()) < 0;
    }
};
std::map<std::pair<int, std::string>, PermEntry, KeyLess> g_perms;
std::shared_mutex g_perm_mu;
std::atomic<bool> g_reorder_enabled{true};      // rrs.cu:83

int current_dev() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        (void)cudaGetLastError();
        return -1;
    }
    return dev;
}

bool is_contiguous_2d(const tcq4_tensor_view *t, size_t elem_row_bytes) {
    return t->nb[1] == elem_row_bytes && t->ne[2] == 1 && t->ne[3] == 1;
}

}  // namespace

extern "C" {

void ggml_backend_cuda_tcq4_register_perm(const char *tensor_name, const int32_t *h_perm, int K) {
    if (!tensor_name || !h_perm || K <= 0) return;
    const int dev = current_dev();
    if (dev < 0) return;                                   // no device: nothing to register on
    int32_t *d_perm = nullptr;
    if (cudaMalloc(&d_perm, (size_t)K * sizeof(int32_t)) != cudaSuccess) {
        (void)cudaGetLastError();
        return;
    }
    if (cudaMemcpy(d_perm, h_perm, (size_t)K * sizeof(int32_t), cudaMemcpyHostToDevice) != cudaSuccess) {
        (void)cudaGetLastError();
        cudaFree(d_perm);
        return;
    }
    std::unique_lock<std::shared_mutex> lk(g_perm_mu);
    PermEntry &e = g_perms[std::make_pair(dev, std::string(tensor_name))];
    if (e.d_perm) cudaFree(e.d_perm);                      // rrs.cu:88-90: replace an existing entry
    e.d_perm = d_perm;
    e.K = K;
    e.h.assign(h_perm, h_perm + K);
}

void ggml_backend_cuda_tcq4_clear_perms(void) {
    std::unique_lock<std::shared_mutex> lk(g_perm_mu);
    for (auto &kv : g_perms)
        if (kv.second.d_perm) cudaFree(kv.second.d_perm);
    g_perms.clear();
}

void ggml_backend_cuda_tcq4_set_reorder_enabled(bool enabled) { g_reorder_enabled.store(enabled); }
int tcq4_b200_get_reorder_enabled(void) { return g_reorder_enabled.load() ? 1 : 0; }

const int32_t *tcq4_b200_get_perm(const char *tensor_name) {
    if (!tensor_name || tensor_name[0] == '\0' || !g_reorder_enabled.load()) return nullptr;   // rrs.cu:102,268
    const int dev = current_dev();
    if (dev < 0) return nullptr;
    std::shared_lock<std::shared_mutex> lk(g_perm_mu);
    auto it = g_perms.find(std::make_pair(dev, tensor_name));
    return it == g_perms.end() ? nullptr : it->second.d_perm;
}

// length of the permutation registered for this tensor on the current device (0: none)
static int64_t perm_len(const char *tensor_name) {
    if (!tensor_name || !tensor_name[0]) return 0;
    const int dev = current_dev();
    if (dev < 0) return 0;
    std::shared_lock<std::shared_mutex> lk(g_perm_mu);
    auto it = g_perms.find(std::make_pair(dev, tensor_name));
    return it == g_perms.end() ? 0 : (int64_t)it->second.K;
}

int tcq4_b200_has_perm(const char *tensor_name) { return tcq4_b200_get_perm(tensor_name) != nullptr ? 1 : 0; }

int tcq4_b200_ggml_supports_op(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1, const tcq4_tensor_view *dst,
                               int cc) {
    if (!src0 || !src1 || !dst) return 0;
    if (src0->type != TCQ4_B200_GGML_TYPE || src1->type != 0 || dst->type != 0) return 0;
    if (src0->ne[0] != src1->ne[0]) return 0;                                   // rrs.cu:256
    if (dst->ne[0] != src0->ne[1] || dst->ne[1] != src1->ne[1]) return 0;
    return tcq4_b200_supports(src1->ne[1], src0->ne[1], src0->ne[0], cc);
}

int tcq4_b200_ggml_mul_mat(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1, tcq4_tensor_view *dst,
                           void *workspace, size_t workspace_bytes, cudaStream_t stream) {
    if (!src0 || !src1 || !dst) return TCQ4_ERR_ARG;
    const int64_t K = src0->ne[0], N = src0->ne[1];          // rrs.cu:247-254
    const int64_t M = src1->ne[1];
    if (src0->type != TCQ4_B200_GGML_TYPE || src1->type != 0 || dst->type != 0) return TCQ4_ERR_ARG;
    if (src0->ne[0] != src1->ne[0]) return TCQ4_ERR_SHAPE;   // GGML_ASSERT(ne00 == ne10)
    if (dst->ne[0] != N || dst->ne[1] != M) return TCQ4_ERR_SHAPE;
    // the reference silently ignores ne[2], ne[3] and strides (rrs.cu:247-254); assert instead
    if (src0->ne[2] != 1 || src0->ne[3] != 1) return TCQ4_ERR_SHAPE;
    // dst may be a row-strided view (ggml_view_2d of a wider matrix: nb[1] >= N*4, a multiple of 16 bytes)
    if (!is_contiguous_2d(src1, (size_t)K * 4) || dst->nb[0] != 4 || dst->nb[1] < (size_t)N * 4 || (dst->nb[1] & 15)) return TCQ4_ERR_SHAPE;
    if (src0->nb[1] != (size_t)(K / 256) * 148) return TCQ4_ERR_SHAPE;          // ggml_row_size, ggml.c:750-757
    const int32_t *d_perm = src0->name ? tcq4_b200_get_perm(src0->name) : nullptr;   // rrs.cu:266-270
    if (d_perm && perm_len(src0->name) != K) return TCQ4_ERR_SHAPE;          // a registered permutation of another length would gather out of bounds
    if (dst->nb[1] != (size_t)N * 4) {
        if (workspace) return TCQ4_ERR_SHAPE;                                 // the strided store goes through the arena entry
        float *yp[1] = {(float *)dst->data};
        return tcq4_b200_mul_mat_peers(src0->data, (const float *)src1->data, yp, 1, (int64_t)(dst->nb[1] / 4), 0, M, N, K, d_perm, stream);
    }
    if (workspace)
        return tcq4_b200_mul_mat_ws(src0->data, (const float *)src1->data, (float *)dst->data, M, N, K, d_perm,
                                    workspace, workspace_bytes, TCQ4_KERNEL_AUTO, stream);
    return tcq4_b200_mul_mat(src0->data, (const float *)src1->data, (float *)dst->data, M, N, K, d_perm, stream);
}

// Fan-out over consecutive MUL_MAT nodes that share src1 (tcq4_b200_mul_mat_fanout).  Each node is checked like
// tcq4_b200_ggml_mul_mat; the permutations registered for the n weight tensors must be equal (same registry entry or same
// contents) -- otherwise the nodes are run one by one, which is always correct.
int tcq4_b200_ggml_mul_mat_fanout(int n, const tcq4_tensor_view *const *src0, const tcq4_tensor_view *src1,
                                  tcq4_tensor_view *const *dst, cudaStream_t stream) {
    if (n < 1 || n > 8 || !src0 || !src1 || !dst) return TCQ4_ERR_ARG;
    const int64_t K = src1->ne[0], M = src1->ne[1];
    const void *w[8];
    int64_t Ns[8], ld[8];
    float *y[8];
    for (int i = 0; i < n; i++) {
        if (!src0[i] || !dst[i]) return TCQ4_ERR_ARG;
        if (src0[i]->type != TCQ4_B200_GGML_TYPE || src1->type != 0 || dst[i]->type != 0) return TCQ4_ERR_ARG;
        const int64_t N = src0[i]->ne[1];
        if (src0[i]->ne[0] != K || dst[i]->ne[0] != N || dst[i]->ne[1] != M) return TCQ4_ERR_SHAPE;
        if (src0[i]->ne[2] != 1 || src0[i]->ne[3] != 1) return TCQ4_ERR_SHAPE;
        if (!is_contiguous_2d(src1, (size_t)K * 4) || dst[i]->nb[0] != 4 || dst[i]->nb[1] < (size_t)N * 4 || (dst[i]->nb[1] & 15)) return TCQ4_ERR_SHAPE;
        if (src0[i]->nb[1] != (size_t)(K / 256) * 148) return TCQ4_ERR_SHAPE;
        w[i] = src0[i]->data;
        Ns[i] = N;
        ld[i] = (int64_t)(dst[i]->nb[1] / 4);
        y[i] = (float *)dst[i]->data;
    }
    const int32_t *d_perm = src0[0]->name ? tcq4_b200_get_perm(src0[0]->name) : nullptr;
    if (d_perm && perm_len(src0[0]->name) != K) return TCQ4_ERR_SHAPE;
    bool same = true;
    if (g_reorder_enabled.load()) {
        const int dev = current_dev();
        std::shared_lock<std::shared_mutex> lk(g_perm_mu);
        const PermEntry *first = nullptr;
        for (int i = 0; i < n && same; i++) {
            const PermEntry *e = nullptr;
            if (dev >= 0 && src0[i]->name && src0[i]->name[0]) {
                auto it = g_perms.find(std::make_pair(dev, src0[i]->name));
                if (it != g_perms.end()) e = &it->second;
            }
            if (i == 0) first = e;
            else same = (e == first) || (e && first && e->h == first->h);
        }
    }
    if (same) return tcq4_b200_mul_mat_fanout_ld(n, w, Ns, y, ld, (const float *)src1->data, M, K, d_perm, stream);
    for (int i = 0; i < n; i++) {
        const int st = tcq4_b200_ggml_mul_mat(src0[i], src1, dst[i], nullptr, 0, stream);
        if (st != TCQ4_OK) return st;
    }
    return TCQ4_OK;
}

// The graph-level entry: the TCQ4 MUL_MAT nodes of a ggml graph in execution order (the loop of
// ggml_cuda_graph_evaluate_and_capture, ggml-cuda.cu:3481-3688, where the backend tries its fusions before it falls back
// to one kernel per node).  Consecutive nodes with ONE src1 (same data pointer, shape and strides: Q/K/V, gate/up) are
// fanned out -- one activation stage and one segmented GEMM / GEMV launch per run of up to 4 -- everything else goes
// node by node.  Always equal to calling tcq4_b200_ggml_mul_mat per node (bit-identical at prefill; at decode batch the
// fp32 sum over 256-blocks may be cut between warps at other places, see tcq4_b200_mul_mat_fanout).
int tcq4_b200_ggml_graph_compute(int n_nodes, const tcq4_tensor_view *const *src0, const tcq4_tensor_view *const *src1,
                                 tcq4_tensor_view *const *dst, cudaStream_t stream) {
    if (n_nodes < 0 || (n_nodes > 0 && (!src0 || !src1 || !dst))) return TCQ4_ERR_ARG;
    static const bool no_fuse = getenv("TCQ4_NO_GRAPH_FUSION") != nullptr;
    auto same_src1 = [](const tcq4_tensor_view *a, const tcq4_tensor_view *b) {
        if (!a || !b || a->data != b->data || a->type != b->type) return false;
        for (int d = 0; d < 4; d++)
            if (a->ne[d] != b->ne[d] || a->nb[d] != b->nb[d]) return false;
        return true;
    };
    int i = 0;
    while (i < n_nodes) {
        int j = i + 1;
        if (!no_fuse)
            while (j < n_nodes && j - i < 4 && same_src1(src1[i], src1[j]) && src0[j] && src0[i] && src0[j]->ne[0] == src0[i]->ne[0]) j++;
        // an output that aliases the shared input (or another output of the run) breaks the "one src1" premise
        for (int a = i; a < j && j - i > 1; a++)
            for (int b = i; b < j; b++)
                if (dst[a] && src1[i] && (dst[a]->data == src1[i]->data || (a != b && dst[b] && dst[a]->data == dst[b]->data))) j = i + 1;
        const int st = (j - i > 1) ? tcq4_b200_ggml_mul_mat_fanout(j - i, src0 + i, src1[i], dst + i, stream)
                                   : tcq4_b200_ggml_mul_mat(src0[i], src1[i], dst[i], nullptr, 0, stream);
        if (st != TCQ4_OK) return st;
        i = j;
    }
    return TCQ4_OK;
}

// ggml_cuda_mul_mat_id (ggml-cuda.cu:2287-2436) for a TCQ4 src0.  Shape rules as there (:2293-2295, 2326-2327) plus the
// contiguity this library needs; the reference's asserts become status codes.
int tcq4_b200_ggml_mul_mat_id(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1, const tcq4_tensor_view *ids,
                              tcq4_tensor_view *dst, int *err_flag, cudaStream_t stream) {
    if (!src0 || !src1 || !ids || !dst) return TCQ4_ERR_ARG;
    if (src0->type != TCQ4_B200_GGML_TYPE || src1->type != 0 || dst->type != 0 || ids->type != 26 /* GGML_TYPE_I32 */)
        return TCQ4_ERR_ARG;
    const int64_t K = src0->ne[0], N = src0->ne[1], n_expert = src0->ne[2];
    const int64_t ne11 = src1->ne[1], n_tokens = src1->ne[2];
    const int64_t n_used = ids->ne[0];
    if (src1->ne[0] != K || ids->ne[1] != n_tokens || src0->ne[3] != 1 || src1->ne[3] != 1) return TCQ4_ERR_SHAPE;
    if (dst->ne[0] != N || dst->ne[1] != n_used || dst->ne[2] != n_tokens) return TCQ4_ERR_SHAPE;
    if (ne11 != 1 && ne11 != n_used) return TCQ4_ERR_SHAPE;
    if (src0->nb[1] != (size_t)(K / 256) * 148) return TCQ4_ERR_SHAPE;
    if (src1->nb[0] != 4 || src1->nb[1] != (size_t)K * 4 || src1->nb[2] != (size_t)ne11 * K * 4) return TCQ4_ERR_SHAPE;
    if (dst->nb[0] != 4 || dst->nb[1] != (size_t)N * 4 || dst->nb[2] != (size_t)n_used * N * 4) return TCQ4_ERR_SHAPE;
    if (ids->nb[0] != 4 || (ids->nb[1] & 3)) return TCQ4_ERR_SHAPE;
    const int32_t *d_perm = src0->name ? tcq4_b200_get_perm(src0->name) : nullptr;
    if (d_perm && perm_len(src0->name) != K) return TCQ4_ERR_SHAPE;
    return tcq4_b200_mul_mat_id(src0->data, n_expert, src0->nb[2], (const float *)src1->data, ne11,
                                (const int32_t *)ids->data, (int64_t)(ids->nb[1] / 4), n_used, n_tokens, (float *)dst->data, N,
                                K, d_perm, err_flag, stream);
}

// ggml_cuda_op_get_rows for a TCQ4 src0 (ggml/src/ggml-cuda/getrows.cu:397-430 -> get_rows_cuda :348-380):
//   src0 [K, nrows, ne02, ne03] tiles, src1 I32 [ne10, ne11, ne12], dst F32 [K, ne10, ne11, ne12];
//   slice (i11, i12) of the indices selects rows of slice (i11, i12) of src0 (k_get_rows_tcq4_tile :126).
// One launch per (i11, i12) slice; the usual embedding lookup has one.
int tcq4_b200_ggml_get_rows(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1, tcq4_tensor_view *dst,
                            int *err_flag, cudaStream_t stream) {
    if (!src0 || !src1 || !dst) return TCQ4_ERR_ARG;
    if (src0->type != TCQ4_B200_GGML_TYPE || src1->type != 26 /* GGML_TYPE_I32 */ || dst->type != 0) return TCQ4_ERR_ARG;
    const int64_t K = src0->ne[0], nrows = src0->ne[1];
    const int64_t ne10 = src1->ne[0], ne11 = src1->ne[1], ne12 = src1->ne[2];
    if (dst->ne[0] != K || dst->ne[1] != ne10 || dst->ne[2] != ne11 || dst->ne[3] != ne12) return TCQ4_ERR_SHAPE;
    if (src1->nb[0] != 4 || dst->nb[0] != 4 || (dst->nb[1] & 15)) return TCQ4_ERR_SHAPE;
    if (src0->nb[1] != (size_t)(K / 256) * 148) return TCQ4_ERR_SHAPE;
    for (int64_t i12 = 0; i12 < ne12; i12++)
        for (int64_t i11 = 0; i11 < ne11; i11++) {
            const char *w = (const char *)src0->data + i11 * src0->nb[2] + i12 * src0->nb[3];
            const int32_t *idx = (const int32_t *)((const char *)src1->data + i11 * src1->nb[1] + i12 * src1->nb[2]);
            float *out = (float *)((char *)dst->data + i11 * dst->nb[2] + i12 * dst->nb[3]);
            const int st = tcq4_b200_get_rows(w, idx, out, ne10, nrows, K, (int64_t)(dst->nb[1] / 4), err_flag, stream);
            if (st != TCQ4_OK) return st;
        }
    return TCQ4_OK;
}

}  // extern "C"
