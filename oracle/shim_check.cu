// oracle/shim_check.cu -- TEST INFRASTRUCTURE ONLY: compile check of the integration shim of INTEGRATION.md sections 2 and 3
// against the FORK'S OWN headers (ggml.h, ggml-cuda/common.cuh), built by `make -C oracle shimcheck` where /root/reference
// exists.  Nothing here is linked into the product; the object file goes to oracle/_ref/.  It proves that the shim a
// maintainer pastes into ggml/src/ggml-cuda/rrs.cu (replacing ggml_cuda_rrs_mul_mat, rrs.cu:241-344) and the supports_op
// case (ggml-cuda.cu:4457-4459) compile against the fork's types: ggml_tensor's ne/nb/data/name/type, ggml_backend_cuda_context's
// pool() and stream(), ggml_cuda_pool_alloc, GGML_ASSERT, and the value of GGML_TYPE_TCQ4_K32.
#include "common.cuh"        // $(REF)/ggml/src/ggml-cuda/common.cuh (pulls ggml.h, ggml-cuda.h, the CUDA runtime)
#include "tcq4_b200.h"

static_assert((int) GGML_TYPE_TCQ4_K32 == TCQ4_B200_GGML_TYPE, "type id of the tile tensor");
static_assert((int) GGML_TYPE_F32 == 0, "type id of the activations");

static tcq4_tensor_view tcq4_view(const ggml_tensor * t) {
    tcq4_tensor_view v;
    v.type = (int) t->type;                       // 0 = F32, 42 = GGML_TYPE_TCQ4_K32
    for (int i = 0; i < 4; i++) { v.ne[i] = t->ne[i]; v.nb[i] = t->nb[i]; }
    v.data = t->data;                             // device pointer inside a ggml CUDA buffer
    v.name = t->name;                             // perm lookup key, as rrs.cu:268-270
    return v;
}

void ggml_cuda_rrs_mul_mat(ggml_backend_cuda_context & ctx,
                           const ggml_tensor * src0, const ggml_tensor * src1, ggml_tensor * dst) {
    const int64_t M = src1->ne[1], N = src0->ne[1], K = src0->ne[0];
    // scratch from the backend pool, exactly where the fork's #else path takes it (rrs.cu:326-328)
    const size_t ws_bytes = tcq4_b200_workspace_size(M, N, K);
    ggml_cuda_pool_alloc<uint8_t> ws(ctx.pool(), ws_bytes + 256);
    void * ws_ptr = (void *) (((uintptr_t) ws.get() + 255) & ~(uintptr_t) 255);
    GGML_ASSERT(tcq4_b200_workspace_init(ws_ptr, ws_bytes, ctx.stream()) == TCQ4_OK);   // zeroes split-K tickets
    tcq4_tensor_view s0 = tcq4_view(src0), s1 = tcq4_view(src1), d = tcq4_view(dst);
    const int st = tcq4_b200_ggml_mul_mat(&s0, &s1, &d, ws_ptr, ws_bytes, ctx.stream());
    GGML_ASSERT(st == TCQ4_OK && "tcq4_b200_ggml_mul_mat failed (see tcq4_b200_last_error())");
}

// the MUL_MAT case of ggml_backend_cuda_device_supports_op for a TCQ4 src0 (ggml-cuda.cu:4457-4459)
bool tcq4_shim_supports_op(const ggml_tensor * op, int device) {
    tcq4_tensor_view s0 = tcq4_view(op->src[0]), s1 = tcq4_view(op->src[1]), d = tcq4_view(op);
    return tcq4_b200_ggml_supports_op(&s0, &s1, &d, ggml_cuda_info().devices[device].cc / 10) != 0;
}

// the permutation registry keeps the fork's names and signatures (ggml-cuda.cu:4148-4158)
void tcq4_shim_registry(const char * name, const int32_t * perm, int K) {
    ggml_backend_cuda_tcq4_register_perm(name, perm, K);
    ggml_backend_cuda_tcq4_set_reorder_enabled(true);
    ggml_backend_cuda_tcq4_clear_perms();
}
