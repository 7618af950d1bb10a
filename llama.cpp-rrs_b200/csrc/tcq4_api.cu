// tcq4_api.cu -- C-ABI layer (include/tcq4_b200.h): dispatch, workspace carving, per-device scratch
// arena.  Host-side equivalent of ggml_cuda_rrs_mul_mat's launch logic
// (/root/reference/ggml/src/ggml-cuda/rrs.cu:241-344).  The channel-permutation registry and the
// ggml operator mirror live in tcq4_registry.cu.
//
// Nothing here allocates or synchronises on the launch path except the explicit convenience
// entries (tcq4_b200_mul_mat's arena growth outside capture, tcq4_b200_mul_mat_host).
#include "../../include/tcq4_b200.h"
#ifdef TCQ4_DIAG
#include "../../include/tcq4_b200_diag.h"
#endif

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "tcq4_kernels.h"

namespace {

thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_err(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
    set_err("%s: %s", what, cudaGetErrorString(e));
    return TCQ4_ERR_CUDA;
}

// ---- per-launch timing tap (bench.py's roofline leg): CUDA events around every launch of the main
// kernel (GEMV or GEMM) and around the activation kernel, on the launching stream.  Diagnostic only:
// single-threaded use, never active during stream capture.
struct ProfState {
    bool   on = false;
    int    cap = 0, n = 0;
    cudaEvent_t *ev = nullptr;      // 3 per launch: before act, before main, after main
    int   *kind = nullptr;          // TCQ4_KERNEL_GEMV / TCQ4_KERNEL_GEMM
};
ProfState g_prof;

constexpr int MAX_DEV = 16;
struct DevInfo {
    bool   ready = false;
    bool   ok = false;          // sm_100 device with attributes set
    int    cc = 0;
    int    num_sms = 0;
    size_t smem_optin = 0;
    // arena
    void  *arena = nullptr;
    size_t arena_bytes = 0, arena_high = 0;
    std::vector<void *> retired;      // earlier arena blocks (kept: captured graphs may still point into them)
    // host staging for tcq4_b200_mul_mat_host
    float *hx = nullptr, *hy = nullptr;
    size_t hx_bytes = 0, hy_bytes = 0;
};
DevInfo g_dev[MAX_DEV];
std::mutex g_dev_mu;

// Resolve (and lazily initialise) the current device.  Returns nullptr + status on failure.
DevInfo *current_device(int *status) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
        (void)cudaGetLastError();
        set_err("no CUDA device available (%s): this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
        *status = TCQ4_ERR_NO_DEVICE;
        return nullptr;
    }
    DevInfo &d = g_dev[dev];
    if (!d.ready) {
        std::lock_guard<std::mutex> lk(g_dev_mu);
        if (!d.ready) {
            int major = 0, minor = 0, sms = 0, optin = 0;
            cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
            cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
            d.cc = major * 10 + minor;
            d.num_sms = sms;
            d.smem_optin = (size_t)optin;
            d.ok = false;
            if (major == 10) {
                cudaError_t e1 = tcq4::gemv_init_attributes(d.smem_optin);
                cudaError_t e2 = tcq4::gemm_init_attributes(d.smem_optin);
                if (e2 == cudaSuccess) e2 = tcq4::gemm2_init_attributes(d.smem_optin);
                if (e2 == cudaSuccess) e2 = tcq4::gemm3_init_attributes(d.smem_optin);
                if (e2 == cudaSuccess) e2 = tcq4::gemm4_init_attributes(d.smem_optin);
                d.ok = (e1 == cudaSuccess && e2 == cudaSuccess);
                if (!d.ok) set_err("kernel attribute setup failed: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
            } else {
                set_err("device %d is sm_%d, this library is sm_100a only and has no fallback", dev, d.cc);
            }
            d.ready = true;
        }
    }
    if (!d.ok) {
        *status = (d.cc / 10 == 10) ? TCQ4_ERR_CUDA : TCQ4_ERR_NO_DEVICE;
        return nullptr;
    }
    *status = TCQ4_OK;
    return &d;
}

bool shape_ok(int64_t M, int64_t N, int64_t K) {
    return M >= 1 && N >= 8 && K >= 256 && (K % TCQ4_B200_TILE_K) == 0 && (N % TCQ4_B200_TILE_CHANNELS) == 0 &&
           M <= (1 << 24) && N <= (1 << 28) && K <= (1 << 24);
}

// Workspace layout (all sub-buffers 256-byte aligned):
//   [0, tickets)           int[NT]         split-K tickets of the GEMV (must be zero between launches)
//   [tickets, +partials)   float[...]      GEMV split-K partial sums
//   [.., +codes/a7/sumq)   GEMM activation staging
struct WsLayout {
    size_t ticket_off = 0, ticket_bytes = 0;
    size_t partial_off = 0, partial_bytes = 0;
    size_t codes_off = 0, a7_off = 0, sumq_off = 0, a7t_off = 0;
    size_t total = 0;
};

// worst-case sizes that do not depend on the device, so tcq4_b200_workspace_size is callable anywhere
WsLayout ws_layout(int64_t M, int64_t N, int64_t K) {
    WsLayout L;
    const int64_t NT = N / 8, KT = K / 256;
    L.ticket_off = 0;
    L.ticket_bytes = (size_t)tcq4::round_up(NT * 4, 256);
    L.partial_off = L.ticket_off + L.ticket_bytes;
    // GEMV split-K: a row of KT tiles is cut into at most KT CTA pieces whatever the grid is; a slot holds
    // 8 or 16 activation rows x 8 channels
    const int64_t maxslots = KT;
    const int64_t slot_rows = (M <= 8) ? 8 : 16;
    L.partial_bytes = (size_t)tcq4::round_up(NT * maxslots * slot_rows * 8 * 4, 256);   // kept for every M: size is monotone in M
    if (M > 8 && ((N + 127) / 128) * 2 <= tcq4::GEMM4_MAX_CTAS) {      // batched-decode GEMM, K parts: one [32|64 tokens][128 channels] slot per CTA
        const int64_t g4 = (int64_t)tcq4::GEMM4_MAX_CTAS * (M <= 32 ? 32 : 64) * 128 * 4;
        L.partial_bytes = std::max(L.partial_bytes, (size_t)tcq4::round_up(g4, 256));
    }
    L.codes_off = L.partial_off + L.partial_bytes;
    const int64_t MB = ((M + 255) / 256) * 2;       // whole CTA pairs: the pair kernel bulk-copies both units
    const size_t codes = (size_t)tcq4::round_up(MB * KT * (int64_t)tcq4::ACT_UNIT_BYTES, 256);
    L.a7_off = L.codes_off + codes;
    L.sumq_off = L.a7_off + (size_t)tcq4::round_up(M * KT * 4, 256);
    L.a7t_off = L.sumq_off + (size_t)tcq4::round_up(M * KT * 16, 256);
    L.total = L.a7t_off + (size_t)tcq4::round_up(MB * 128 * KT * 4, 256);       // scales per (128-row block, k-tile), tcq4_gemm3.cu
    return L;
}

int select_kernel(const DevInfo *d, int64_t M, int64_t N, int64_t K, tcq4::GemvPlan *gv) {
    const int sms = d ? d->num_sms : 148;
    const size_t optin = d ? d->smem_optin : (size_t)232448;
    tcq4::GemvPlan tmp;
    static const bool no_g4 = getenv("TCQ4_NO_GEMM4") != nullptr;
    if (tcq4::gemv_plan(M, N, K, sms, optin, gv ? gv : &tmp)) {
        // 9 <= M <= 16: the batched-decode GEMM is 10-25 % ahead of the GEMV on every shape of BASELINE's configs (measured,
        // tests/gpu_g4_vs_gemv.py, profiles/batched_decode_r02.txt); M <= 8 stays with the GEMV (1.2-1.9x ahead there)
        tcq4::Gemm4Plan g4;
        if (M >= 9 && !no_g4 && tcq4::gemm4_plan(M, &N, 1, K, sms, optin, &g4)) return TCQ4_KERNEL_GEMM4;
        return TCQ4_KERNEL_GEMV;
    }
    // 17 <= M <= 64 (batched decode): weights through TMEM, tokens as the MMA's N (tcq4_gemm4.cu)
    if (M <= 64 && !no_g4) return TCQ4_KERNEL_GEMM4;
    // 65 <= M <= 128: the single-CTA GEMM (128-row tiles, every SM busy) only wins on wide, shallow matrices;
    // elsewhere the persistent pair GEMM is up to 2x faster even with its second row block empty (measured:
    // profiles/, "M <= 128")
    if (M <= 128 && N >= 8192 && K <= 4096) return TCQ4_KERNEL_GEMM;
    // One matrix per call: the kind::i8 pair GEMM (measured per shape, tests/gpu_g3_shapes.py: 5-40 % ahead of the kind::f16
    // kernel on most single matrices).  Fan-out groups take the kind::f16 kernel (tcq4_b200_mul_mat_fanout: its segmented
    // tile list makes Q/K/V or gate/up ONE launch), as does an explicit TCQ4_KERNEL_GEMM3 / TCQ4_AUTO_GEMM3=1.
    static const bool auto_g3 = getenv("TCQ4_AUTO_GEMM3") != nullptr;
    return auto_g3 ? TCQ4_KERNEL_GEMM3 : TCQ4_KERNEL_GEMM2;
}

tcq4::YOut single_out(float *y, int64_t N) {
    tcq4::YOut o;
    for (int i = 0; i < tcq4::MAX_PEERS; i++) o.p[i] = y;
    o.n = 1;
    o.ldc = N;
    o.mc = nullptr;
    return o;
}

// K1 (unless the activations of x already sit in the workspace) + ONE batched-decode GEMM launch over up to four matrices
// that share x.  What must start at zero -- the outputs when blocks are cut in two (red.add), the arrival counters when
// they are cut in K parts -- is cleared by K1 itself (strided outputs row by row), or by a small kernel when K1 does not run.
int run_gemm4_group(DevInfo *d, const tcq4::Gemm4Seg *segs, int nseg, const float *x, const int32_t *perm, int64_t M, int64_t K,
                    uint8_t *wsb, const WsLayout &L, bool reuse_act, int32_t *dbg_isum, cudaStream_t stream, int *launches,
                    cudaEvent_t after_act = nullptr, const float *res = nullptr, const float *bias = nullptr) {
    int64_t Ns[tcq4::GEMM4_MAX_SEG];
    for (int i = 0; i < nseg; i++) Ns[i] = segs[i].N;
    tcq4::Gemm4Plan g4;
    if (!tcq4::gemm4_plan(M, Ns, nseg, K, d->num_sms, d->smem_optin, &g4, res != nullptr || bias != nullptr)) {
        set_err("batched-decode GEMM cannot be planned for this shape/device");
        return TCQ4_ERR_SHAPE;
    }
    if (g4.partial_bytes > L.partial_bytes || g4.counter_bytes > L.ticket_bytes || g4.grid > tcq4::GEMM4_MAX_CTAS) {
        set_err("internal: batched-decode GEMM workspace bound too small");
        return TCQ4_ERR_WORKSPACE;
    }
    tcq4::ActStage st;
    st.codes = reinterpret_cast<int8_t *>(wsb + L.codes_off);
    st.a7 = reinterpret_cast<float *>(wsb + L.a7_off);
    st.sumq = reinterpret_cast<int16_t *>(wsb + L.sumq_off);
    st.a7t = reinterpret_cast<float *>(wsb + L.a7t_off);
    tcq4::ZeroList zl{};
    if (g4.mode == 1) {
        zl.n = nseg;
        for (int i = 0; i < nseg; i++) {
            zl.p[i] = segs[i].y;
            zl.rows[i] = M;
            zl.cols[i] = segs[i].N;       // N % 8 == 0
            zl.ld[i] = segs[i].ldc;
        }
    } else if (g4.mode == 2) {
        zl.n = 1;
        zl.p[0] = reinterpret_cast<float *>(wsb + L.ticket_off);
        zl.rows[0] = 1;
        zl.cols[0] = tcq4::round_up((int64_t)g4.counter_bytes, 16) / 4;
        zl.ld[0] = zl.cols[0];
    }
    cudaError_t e = cudaSuccess;
    if (!reuse_act) {
        e = tcq4::launch_act_quant_nib_zl(x, perm, M, K, st, d->num_sms, stream, &zl);
        if (e != cudaSuccess) return cuda_fail(e, "activation kernel launch");
        (*launches)++;
    } else if (zl.n) {               // (the counters too: another shape's layout may have put other data where they live)
        e = tcq4::launch_zero_list(zl, d->num_sms, stream);
        if (e != cudaSuccess) return cuda_fail(e, "output clear launch");
        (*launches)++;
    }
    if (after_act) cudaEventRecord(after_act, stream);      // profile tap: activation stage | GEMM
    e = tcq4::launch_gemm4(segs, nseg, st, M, K, g4, reinterpret_cast<float *>(wsb + L.partial_off),
                           reinterpret_cast<int *>(wsb + L.ticket_off), dbg_isum, stream, res, bias);
    if (e != cudaSuccess) return cuda_fail(e, "gemm4 launch");
    (*launches)++;
    return TCQ4_OK;
}

// N_layout / reuse_act (fan-out, tcq4_b200_mul_mat_fanout): several weight matrices applied to the SAME activations share
// one workspace layout (that of the widest matrix) and the GEMMs after the first skip the activation kernel.
int run_mul_mat(const void *w, const float *x, tcq4::YOut yo, int64_t M, int64_t N, int64_t K, const int32_t *perm,
                void *ws, size_t ws_bytes, int kernel, int32_t *dbg_isum, cudaStream_t stream, int64_t N_layout = 0,
                bool reuse_act = false, bool *act_done = nullptr, const float *res = nullptr, const float *bias = nullptr) {
    float *y = yo.p[0];
    const bool plain = (yo.n == 1 && yo.ldc == N && !yo.mc);     // ggml's dst; otherwise strided / several destinations
    if (!shape_ok(M, N, K)) {
        set_err("unsupported shape M=%lld N=%lld K=%lld (need K%%256==0, N%%8==0)", (long long)M, (long long)N,
                (long long)K);
        return TCQ4_ERR_SHAPE;
    }
    if (!w || !x || !y || ((uintptr_t)w & 15) || ((uintptr_t)x & 15) || ((uintptr_t)y & 15) ||
        (perm && ((uintptr_t)perm & 15))) {
        set_err("null or misaligned pointer (w/x/y/perm must be 16-byte aligned device pointers)");
        return TCQ4_ERR_ARG;
    }
    if (yo.n < 1 || yo.n > tcq4::MAX_PEERS || yo.ldc < N || (yo.ldc & 3)) {
        set_err("bad output spec: 1..%d destinations, ldc >= N and a multiple of 4", tcq4::MAX_PEERS);
        return TCQ4_ERR_ARG;
    }
    for (int i = 0; i < yo.n; i++)
        if (!yo.p[i] || ((uintptr_t)yo.p[i] & 15)) {
            set_err("null or misaligned destination %d", i);
            return TCQ4_ERR_ARG;
        }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    const WsLayout L = ws_layout(M, N_layout > N ? N_layout : N, K);
    if (!ws || ((uintptr_t)ws & 255) || ws_bytes < L.total) {
        set_err("workspace too small or misaligned: need %zu bytes (256-aligned), got %zu", L.total, ws_bytes);
        return TCQ4_ERR_WORKSPACE;
    }
    uint8_t *wsb = reinterpret_cast<uint8_t *>(ws);

    tcq4::GemvPlan gv;
    int k = kernel;
    if (k == TCQ4_KERNEL_AUTO) k = select_kernel(d, M, N, K, &gv);
    const bool one_dest = (yo.n == 1 && !yo.mc);
    if (k == TCQ4_KERNEL_GEMM && !plain) k = TCQ4_KERNEL_GEMM2;      // only the pair GEMMs and the GEMV write strided / to peers
    if (k == TCQ4_KERNEL_GEMM && (res || bias)) k = TCQ4_KERNEL_GEMM2;            // ... and fuse bias / residual
    if (k == TCQ4_KERNEL_GEMM3 && !one_dest) k = TCQ4_KERNEL_GEMM2;               // several destinations / multicast: kind::i8 pair GEMM
    if (k == TCQ4_KERNEL_GEMM4 && (!one_dest || M > 64 || ((res || bias) && yo.ldc != N))) k = TCQ4_KERNEL_GEMM2;
    if (k == TCQ4_KERNEL_GEMV) {
        if (!tcq4::gemv_plan(M, N, K, d->num_sms, d->smem_optin, &gv)) {
            set_err("GEMV kernel does not take M=%lld K=%lld (M<=%d and M*K codes must fit shared memory)",
                    (long long)M, (long long)K, tcq4::GEMV_MAX_M);
            return TCQ4_ERR_SHAPE;
        }
        if (gv.partial_bytes > L.partial_bytes) {
            set_err("internal: GEMV partial buffer bound too small");
            return TCQ4_ERR_WORKSPACE;
        }
        const bool prof = g_prof.on && g_prof.n < g_prof.cap;
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 0], stream);
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 1], stream);
        }
        // the region once used for split-K partials now receives the pre-pass output when it is large enough
        float *blob = gv.blob_bytes <= L.partial_bytes ? reinterpret_cast<float *>(wsb + L.partial_off) : nullptr;
        tcq4::GvExpert ex{};
        ex.pdl = 1;
        ex.res = res;
        ex.bias = bias;
        cudaError_t e = tcq4::launch_gemv(w, x, perm, yo, M, N, K, gv, blob, reinterpret_cast<int *>(wsb + L.ticket_off),
                                          dbg_isum, stream, (res || bias) ? &ex : nullptr);
        if (e != cudaSuccess) return cuda_fail(e, "gemv launch");
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 2], stream);
            g_prof.kind[g_prof.n++] = TCQ4_KERNEL_GEMV;
        }
        g_launches.fetch_add((blob && tcq4::gemv_uses_prepass(gv, perm)) ? 2 : 1, std::memory_order_relaxed);
        return TCQ4_OK;
    }
    if (k == TCQ4_KERNEL_GEMM4) {
        // batched decode: nibble staging (K1 MODE_NIB), weights widened into TMEM, stream-K / K parts
        const bool prof = g_prof.on && g_prof.n < g_prof.cap;
        if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 0], stream);
        const tcq4::Gemm4Seg sg{w, y, N, yo.ldc};
        int launches = 0;
        const int rc = run_gemm4_group(d, &sg, 1, x, perm, M, K, wsb, L, reuse_act, dbg_isum, stream, &launches,
                                       prof ? g_prof.ev[3 * g_prof.n + 1] : nullptr, res, bias);
        if (rc != TCQ4_OK) return rc;
        if (act_done) *act_done = true;
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 2], stream);
            g_prof.kind[g_prof.n++] = k;
        }
        g_launches.fetch_add(launches, std::memory_order_relaxed);
        return TCQ4_OK;
    }
    if (k == TCQ4_KERNEL_GEMM3) {
        // kind::f16 pair GEMM: nibble staging (K1 MODE_NIB), optional K split with the output cleared by K1
        tcq4::Gemm3Plan g3;
        const int64_t Ns[1] = {N};
        const bool allow_split = (yo.ldc == N) && !(res != nullptr && res == y);
        if (!tcq4::gemm3_plan(M, Ns, 1, K, d->num_sms, d->smem_optin, allow_split, &g3)) {
            set_err("pair GEMM (kind::f16) cannot be planned for this shape/device");
            return TCQ4_ERR_SHAPE;
        }
        tcq4::ActStage st;
        st.codes = reinterpret_cast<int8_t *>(wsb + L.codes_off);
        st.a7 = reinterpret_cast<float *>(wsb + L.a7_off);
        st.sumq = reinterpret_cast<int16_t *>(wsb + L.sumq_off);
        st.a7t = reinterpret_cast<float *>(wsb + L.a7t_off);
        const bool prof = g_prof.on && g_prof.n < g_prof.cap;
        if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 0], stream);
        cudaError_t e = cudaSuccess;
        if (!reuse_act) {
            e = tcq4::launch_act_quant_nib(x, perm, M, K, st, d->num_sms, stream, g3.ksplit > 1 ? y : nullptr, M * N);
            if (e != cudaSuccess) return cuda_fail(e, "activation kernel launch");
        } else if (g3.ksplit > 1) {
            e = tcq4::launch_zero_floats(y, M * N, d->num_sms, stream);
            if (e != cudaSuccess) return cuda_fail(e, "output clear launch");
        }
        if (act_done) *act_done = true;
        if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 1], stream);
        const tcq4::Gemm3Seg sg{w, y, N, yo.ldc};
        e = tcq4::launch_gemm3(&sg, 1, st, M, K, g3, dbg_isum, stream, res, bias);
        if (e != cudaSuccess) return cuda_fail(e, "gemm3 launch");
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 2], stream);
            g_prof.kind[g_prof.n++] = k;
        }
        g_launches.fetch_add((!reuse_act || g3.ksplit > 1) ? 2 : 1, std::memory_order_relaxed);
        return TCQ4_OK;
    }
    if (k != TCQ4_KERNEL_GEMM && k != TCQ4_KERNEL_GEMM2) {
        set_err("unknown kernel selector %d", kernel);
        return TCQ4_ERR_ARG;
    }
    tcq4::GemmPlan gm;
    if (!tcq4::gemm_plan(M, N, K, d->num_sms, d->smem_optin, &gm)) {
        set_err("GEMM kernel cannot be planned for this shape/device");
        return TCQ4_ERR_SHAPE;
    }
    tcq4::ActStage st;
    st.codes = reinterpret_cast<int8_t *>(wsb + L.codes_off);
    st.a7 = reinterpret_cast<float *>(wsb + L.a7_off);
    st.sumq = reinterpret_cast<int16_t *>(wsb + L.sumq_off);
    st.a7t = nullptr;
    const bool prof = g_prof.on && g_prof.n < g_prof.cap;
    if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 0], stream);
    // TCQ4_G2_FUSE=1: the pair GEMM runs the activation stage itself on its idle warps (no separate launch).  Opt-in:
    // measured SLOWER (Qwen3-4B prefill chain 7.93 vs 7.04 ms; +7 us per MUL_MAT at K=2560, +15 us at K=9728) -- with 16
    // warps per SM the butterfly is a latency chain of ~1.8 us per round, the stand-alone kernel hides it with 64.
    static const bool want_fuse = getenv("TCQ4_G2_FUSE") != nullptr;
    const bool fuse = (k == TCQ4_KERNEL_GEMM2) && want_fuse && plain && L.ticket_bytes >= 8;
    cudaError_t e = cudaSuccess;
    if (!fuse && !reuse_act) {
        e = tcq4::launch_act_quant_stage(x, perm, M, K, st, d->num_sms, stream);
        if (e != cudaSuccess) return cuda_fail(e, "activation kernel launch");
    }
    if (act_done) *act_done = !fuse;         // the codes of x now sit in the workspace
    if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 1], stream);
    if (k == TCQ4_KERNEL_GEMM2) {
        tcq4::Gemm2Plan g2;
        if (!tcq4::gemm2_plan(M, N, K, d->num_sms, d->smem_optin, &g2)) {
            set_err("pair GEMM kernel cannot be planned for this shape/device");
            return TCQ4_ERR_SHAPE;
        }
        e = tcq4::launch_gemm2(w, st, yo, M, N, K, g2, dbg_isum, stream, fuse ? x : nullptr, perm,
                               reinterpret_cast<int *>(wsb + L.ticket_off), res, bias);
    } else {
        e = tcq4::launch_gemm(w, st, y, M, N, K, gm, dbg_isum, stream);
    }
    if (e != cudaSuccess) return cuda_fail(e, "gemm launch");
    if (prof) {
        cudaEventRecord(g_prof.ev[3 * g_prof.n + 2], stream);
        g_prof.kind[g_prof.n++] = k;
    }
    g_launches.fetch_add((!fuse && !reuse_act) ? 2 : 1, std::memory_order_relaxed);
    return TCQ4_OK;
}

}  // namespace

// =================================================================================================
// exported C ABI
// =================================================================================================
extern "C" {

const char *tcq4_b200_version(void) { return "tcq4-b200 0.1 (sm_100a)"; }
const char *tcq4_b200_last_error(void) { return g_err; }
uint64_t tcq4_b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int tcq4_b200_supports(int64_t M, int64_t N, int64_t K, int cc) {
    return (cc >= 100 && cc < 110 && shape_ok(M, N, K)) ? 1 : 0;
}

int tcq4_b200_select_kernel(int64_t M, int64_t N, int64_t K) {
    if (!shape_ok(M, N, K)) return TCQ4_KERNEL_AUTO;
    int dev = -1;
    const DevInfo *d = nullptr;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < MAX_DEV && g_dev[dev].ready && g_dev[dev].ok)
        d = &g_dev[dev];
    else
        (void)cudaGetLastError();
    return select_kernel(d, M, N, K, nullptr);
}

size_t tcq4_b200_workspace_size(int64_t M, int64_t N, int64_t K) {
    if (!shape_ok(M, N, K)) return 0;
    return ws_layout(M, N, K).total;
}

int tcq4_b200_workspace_init(void *workspace, size_t workspace_bytes, cudaStream_t stream) {
    if (!workspace || ((uintptr_t)workspace & 255)) {
        set_err("workspace must be a 256-byte aligned device pointer");
        return TCQ4_ERR_ARG;
    }
    int status;
    if (!current_device(&status)) return status;
    // only the ticket region has to be zero, but its size depends on N: clear everything once
    cudaError_t e = cudaMemsetAsync(workspace, 0, workspace_bytes, stream);
    if (e != cudaSuccess) return cuda_fail(e, "workspace memset");
    return TCQ4_OK;
}

int tcq4_b200_mul_mat_ws(const void *w_tiles, const float *x, float *y, int64_t M, int64_t N, int64_t K,
                         const int32_t *d_perm, void *workspace, size_t workspace_bytes, int kernel,
                         cudaStream_t stream) {
    return run_mul_mat(w_tiles, x, single_out(y, N), M, N, K, d_perm, workspace, workspace_bytes, kernel, nullptr, stream);
}

int tcq4_b200_mul_mat_debug(const void *w_tiles, const float *x, float *y, int64_t M, int64_t N, int64_t K,
                            const int32_t *d_perm, void *workspace, size_t workspace_bytes, int kernel,
                            int32_t *isum, cudaStream_t stream) {
    return run_mul_mat(w_tiles, x, single_out(y, N), M, N, K, d_perm, workspace, workspace_bytes, kernel, isum, stream);
}

// grow the per-device arena (never while the stream is capturing)
static int arena_ensure(DevInfo *d, size_t need, cudaStream_t stream) {      // `need` may be raised below (geometric growth)
    if (d->arena_bytes >= need) return TCQ4_OK;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (stream != nullptr && cudaStreamIsCapturing(stream, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone) {
        set_err("scratch arena too small (%zu < %zu) and the stream is capturing: call tcq4_b200_reserve first",
                d->arena_bytes, need);
        return TCQ4_ERR_WORKSPACE;
    }
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (d->arena_bytes >= need) return TCQ4_OK;
    cudaError_t e = cudaDeviceSynchronize();          // old arena may still be in use
    if (e != cudaSuccess) return cuda_fail(e, "arena sync");
    // The old block is RETIRED, never freed: a CUDA graph captured earlier holds its address (activation codes, scales,
    // pre-pass blob) and may be replayed after this growth.  Blocks grow geometrically, so the retired ones sum to less than
    // the live one.
    if (d->arena) d->retired.push_back(d->arena);
    d->arena = nullptr;
    d->arena_bytes = 0;
    if (need < 2 * d->arena_high) need = 2 * d->arena_high;
    d->arena_high = need;
    const size_t bytes = (size_t)tcq4::round_up((int64_t)need, 1 << 20);
    e = cudaMalloc(&d->arena, bytes);
    if (e != cudaSuccess) return cuda_fail(e, "arena cudaMalloc");
    e = cudaMemset(d->arena, 0, bytes);
    if (e != cudaSuccess) return cuda_fail(e, "arena memset");
    d->arena_bytes = bytes;
    return TCQ4_OK;
}

int tcq4_b200_reserve(int64_t max_M, int64_t max_N, int64_t max_K) {
    if (!shape_ok(max_M, max_N, max_K)) {
        set_err("reserve: unsupported shape");
        return TCQ4_ERR_SHAPE;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    size_t need = ws_layout(max_M, max_N, max_K).total;
    // the GEMV partial region exists only for M <= GEMV_MAX_M: make sure small-M calls of the same N,K fit too
    for (int64_t m : {(int64_t)8, (int64_t)16}) {
        const size_t n2 = ws_layout(max_M < m ? max_M : m, max_N, max_K).total;
        if (n2 > need) need = n2;
    }
    return arena_ensure(d, need, nullptr);
}

int tcq4_b200_mul_mat(const void *w_tiles, const float *x, float *y, int64_t M, int64_t N, int64_t K,
                      const int32_t *d_perm, cudaStream_t stream) {
    if (!shape_ok(M, N, K)) {
        set_err("unsupported shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
        return TCQ4_ERR_SHAPE;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    const size_t need = ws_layout(M, N, K).total;
    status = arena_ensure(d, need, stream);
    if (status != TCQ4_OK) return status;
    return run_mul_mat(w_tiles, x, single_out(y, N), M, N, K, d_perm, d->arena, d->arena_bytes, TCQ4_KERNEL_AUTO, nullptr,
                       stream);
}

// ---- allocation-free entry -----------------------------------------------------------------------------------------
// tcq4_b200_mul_mat_static: the same MUL_MAT with its scratch (activation codes, scales, group sums, GEMV pre-pass
// blob) in a file-scope device array instead of a caller workspace or the cudaMalloc'ed arena, for callers that may
// not allocate: code under stream capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in
// the reference.  Every GPU that loads the library holds the array until
// exit; every call on a device shares it (it keeps the previous call's contents), so two calls that overlap on the
// GPU -- different streams, or graphs replayed concurrently -- race on it.
constexpr size_t TCQ4_STATIC_SCRATCH_BYTES = (size_t)TCQ4_B200_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_static_scratch[TCQ4_STATIC_SCRATCH_BYTES];
static void *g_static_scratch_ptr[MAX_DEV] = {};      // device ADDRESS of the array, per device (the name above is only
static std::mutex g_static_mu;                         // a host-side stand-in and must never be handed to a kernel)

int tcq4_b200_mul_mat_static(const void *w_tiles, const float *x, float *y, int64_t M, int64_t N, int64_t K,
                             const int32_t *d_perm, cudaStream_t stream) {
    if (!shape_ok(M, N, K)) {
        set_err("unsupported shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
        return TCQ4_ERR_SHAPE;
    }
    const WsLayout L = ws_layout(M, N, K);
    if (L.total > TCQ4_STATIC_SCRATCH_BYTES) {
        set_err("mul_mat_static: M=%lld N=%lld K=%lld needs %zu bytes of scratch, the static array holds %zu "
                "(use tcq4_b200_mul_mat_ws with a caller workspace for larger shapes)",
                (long long)M, (long long)N, (long long)K, L.total, TCQ4_STATIC_SCRATCH_BYTES);
        return TCQ4_ERR_WORKSPACE;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    void *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_scratch_ptr[dev]) {      // once per device, with that device current
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e != cudaSuccess || !p) return cuda_fail(e, "cudaGetSymbolAddress(static scratch)");
            g_static_scratch_ptr[dev] = p;
        }
        scratch = g_static_scratch_ptr[dev];
    }
    // the only region that must start at zero (counters of the opt-in fused activation stage); everything else is
    // written before it is read within the call
    e = cudaMemsetAsync(scratch, 0, L.ticket_bytes < 1024 ? L.ticket_bytes : 1024, stream);
    if (e != cudaSuccess) return cuda_fail(e, "static scratch memset");
    return run_mul_mat(w_tiles, x, single_out(y, N), M, N, K, d_perm, scratch, TCQ4_STATIC_SCRATCH_BYTES, TCQ4_KERNEL_AUTO,
                       nullptr, stream);
}

// Epilogue fusion: y = W x + bias[n] + residual[m][n] in the MUL_MAT's own epilogue -- the ggml_add nodes
// that follow the attention-output and ffn-down projections (residual stream) and the bias add of models that have one.
// residual may alias y (in-place add).  The pair GEMM starts its fp32 accumulators from bias + residual, the GEMV adds
// them at its store; both differ from "MUL_MAT, then add" only in fp32 summation order.
int tcq4_b200_mul_mat_add(const void *w_tiles, const float *x, float *y, int64_t M, int64_t N, int64_t K,
                          const int32_t *d_perm, const float *residual, const float *bias, cudaStream_t stream) {
    if (!shape_ok(M, N, K)) {
        set_err("unsupported shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
        return TCQ4_ERR_SHAPE;
    }
    if (((uintptr_t)residual & 15) || ((uintptr_t)bias & 15)) {
        set_err("mul_mat_add: residual / bias must be 16-byte aligned");
        return TCQ4_ERR_ARG;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    status = arena_ensure(d, ws_layout(M, N, K).total, stream);
    if (status != TCQ4_OK) return status;
    return run_mul_mat(w_tiles, x, single_out(y, N), M, N, K, d_perm, d->arena, d->arena_bytes, TCQ4_KERNEL_AUTO, nullptr, stream,
                       0, false, nullptr, residual, bias);
}

// Fan-out (graph-level fusion): n weight matrices applied to the SAME activations -- Q/K/V and gate/up of a
// transformer layer are consecutive MUL_MAT nodes with one src1 (the hook is where ggml-cuda tries its other fusions,
// ggml-cuda.cu:3481-3688).  The activation stage (perm + FWHT + int4 quantisation, K1) runs ONCE; every GEMM after the
// first reads the codes already in the workspace: bit-identical to n separate tcq4_b200_mul_mat calls.  With a decode
// batch (M <= 16) all matrices run as ONE GEMV launch: same int32 accumulators per 256-block; the fp32 sum over a
// channel's blocks may be cut between warps at other places than in a separate launch (last-bit differences).
// All matrices must share the channel permutation (they do when it derives from the imatrix of their common input).
int tcq4_b200_mul_mat_fanout(int n, const void *const *w_tiles, const int64_t *Ns, float *const *ys, const float *x, int64_t M,
                             int64_t K, const int32_t *d_perm, cudaStream_t stream) {
    return tcq4_b200_mul_mat_fanout_ld(n, w_tiles, Ns, ys, nullptr, x, M, K, d_perm, stream);
}

// Same with a leading dimension per output (ldcs[i] >= Ns[i], a multiple of 4; null: contiguous): ys[i] may be a column window
// of a wider row-major matrix -- a ggml view, or this GPU's column block of an N-sharded C.
int tcq4_b200_mul_mat_fanout_ld(int n, const void *const *w_tiles, const int64_t *Ns, float *const *ys, const int64_t *ldcs,
                                const float *x, int64_t M, int64_t K, const int32_t *d_perm, cudaStream_t stream) {
    if (n < 1 || n > 8 || !w_tiles || !Ns || !ys) {
        set_err("mul_mat_fanout: 1..8 weight matrices");
        return TCQ4_ERR_ARG;
    }
    int64_t ld[8];
    for (int i = 0; i < n; i++) {
        ld[i] = ldcs ? ldcs[i] : Ns[i];
        if (ld[i] < Ns[i] || (ld[i] & 3)) {
            set_err("mul_mat_fanout: ldc of matrix %d must be >= N and a multiple of 4", i);
            return TCQ4_ERR_ARG;
        }
    }
    int64_t Nmax = 0;
    for (int i = 0; i < n; i++) {
        if (!shape_ok(M, Ns[i], K)) {
            set_err("unsupported shape M=%lld N=%lld K=%lld", (long long)M, (long long)Ns[i], (long long)K);
            return TCQ4_ERR_SHAPE;
        }
        Nmax = std::max(Nmax, Ns[i]);
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    // decode batch (M <= 16): ONE GEMV launch over the sum(N_i) channels of all matrices (SEG mode of tcq4_gemv_kernel:
    // a channel block takes its tiles from, and stores to, the matrix it belongs to) -- a decode MUL_MAT costs a fixed
    // ~4.5 us per launch, so Q/K/V in one launch is worth more than the shared activation stage
    int64_t Nsum = 0;
    for (int i = 0; i < n; i++) Nsum += Ns[i];
    tcq4::GemvPlan gv;
    static const bool no_seg = getenv("TCQ4_FANOUT_NOSEG") != nullptr;       // diagnostics / tests: one launch per matrix
    // (9 <= M <= 16: the batched-decode GEMM takes the group, as it takes single matrices -- select_kernel)
    const bool gemv_group = select_kernel(d, M, Nmax, K, nullptr) == TCQ4_KERNEL_GEMV;
    if (n >= 2 && n <= tcq4::GV_MAX_SEG && !no_seg && gemv_group && shape_ok(M, Nsum, K) &&
        tcq4::gemv_plan(M, Nsum, K, d->num_sms, d->smem_optin, &gv)) {
        tcq4::GvExpert ex{};
        ex.pdl = 1;
        ex.nseg = n;
        for (int i = 0; i < n; i++) {
            if (!w_tiles[i] || !ys[i] || ((uintptr_t)w_tiles[i] & 15) || ((uintptr_t)ys[i] & 15)) {
                set_err("mul_mat_fanout: null or misaligned pointer of matrix %d", i);
                return TCQ4_ERR_ARG;
            }
            ex.seg_w[i] = reinterpret_cast<const uint8_t *>(w_tiles[i]);
            ex.seg_y[i] = ys[i];
            ex.seg_ldc[i] = ld[i];
            ex.seg_nt0[i + 1] = ex.seg_nt0[i] + (int)(Ns[i] / 8);
        }
        if (!x || ((uintptr_t)x & 15) || (d_perm && ((uintptr_t)d_perm & 15))) {
            set_err("mul_mat_fanout: null or misaligned x / perm");
            return TCQ4_ERR_ARG;
        }
        const WsLayout L = ws_layout(M, Nsum, K);
        status = arena_ensure(d, L.total, stream);
        if (status != TCQ4_OK) return status;
        uint8_t *wsb = reinterpret_cast<uint8_t *>(d->arena);
        float *blob = gv.blob_bytes <= L.partial_bytes ? reinterpret_cast<float *>(wsb + L.partial_off) : nullptr;
        const bool prof = g_prof.on && g_prof.n < g_prof.cap;
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 0], stream);
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 1], stream);
        }
        cudaError_t e = tcq4::launch_gemv(nullptr, x, d_perm, single_out(ys[0], Nsum), M, Nsum, K, gv, blob,
                                          reinterpret_cast<int *>(wsb + L.ticket_off), nullptr, stream, &ex);
        if (e != cudaSuccess) return cuda_fail(e, "fan-out gemv launch");
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 2], stream);
            g_prof.kind[g_prof.n++] = TCQ4_KERNEL_GEMV;
        }
        g_launches.fetch_add((blob && tcq4::gemv_uses_prepass(gv, d_perm)) ? 2 : 1, std::memory_order_relaxed);
        return TCQ4_OK;
    }
    status = arena_ensure(d, ws_layout(M, Nmax, K).total, stream);
    if (status != TCQ4_OK) return status;
    static const bool no_g3 = getenv("TCQ4_NO_GEMM3") != nullptr;       // experiments: fan-out through the round-1 kernels
    const int ksel = select_kernel(d, M, Nmax, K, nullptr);
    if (ksel == TCQ4_KERNEL_GEMM4 && n >= 2) {
        // batched decode: ONE activation launch and ONE GEMM launch per group of up to four matrices (their 128-channel
        // blocks form one stream-K line)
        for (int i = 0; i < n; i++)
            if (!w_tiles[i] || !ys[i] || ((uintptr_t)w_tiles[i] & 15) || ((uintptr_t)ys[i] & 15)) {
                set_err("mul_mat_fanout: null or misaligned pointer of matrix %d", i);
                return TCQ4_ERR_ARG;
            }
        if (!x || ((uintptr_t)x & 15) || (d_perm && ((uintptr_t)d_perm & 15))) {
            set_err("mul_mat_fanout: null or misaligned x / perm");
            return TCQ4_ERR_ARG;
        }
        const WsLayout L = ws_layout(M, Nmax, K);
        uint8_t *wsb = reinterpret_cast<uint8_t *>(d->arena);
        const bool prof = g_prof.on && g_prof.n < g_prof.cap;
        if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 0], stream);
        int launches = 0;
        for (int i0 = 0; i0 < n; i0 += tcq4::GEMM4_MAX_SEG) {
            const int ns = std::min(tcq4::GEMM4_MAX_SEG, n - i0);
            tcq4::Gemm4Seg sg[tcq4::GEMM4_MAX_SEG];
            for (int i = 0; i < ns; i++) sg[i] = tcq4::Gemm4Seg{w_tiles[i0 + i], ys[i0 + i], Ns[i0 + i], ld[i0 + i]};
            const int rc = run_gemm4_group(d, sg, ns, x, d_perm, M, K, wsb, L, i0 > 0, nullptr, stream, &launches,
                                           (prof && i0 == 0) ? g_prof.ev[3 * g_prof.n + 1] : nullptr);
            if (rc != TCQ4_OK) return rc;
        }
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 2], stream);
            g_prof.kind[g_prof.n++] = TCQ4_KERNEL_GEMM4;
        }
        g_launches.fetch_add(launches, std::memory_order_relaxed);
        return TCQ4_OK;
    }
    if (!no_g3 && n >= 2 && (ksel == TCQ4_KERNEL_GEMM2 || ksel == TCQ4_KERNEL_GEMM3 || ksel == TCQ4_KERNEL_GEMM)) {
        // ONE activation launch and ONE GEMM launch per group of up to four matrices: their column blocks form one tile list
        for (int i = 0; i < n; i++)
            if (!w_tiles[i] || !ys[i] || ((uintptr_t)w_tiles[i] & 15) || ((uintptr_t)ys[i] & 15)) {
                set_err("mul_mat_fanout: null or misaligned pointer of matrix %d", i);
                return TCQ4_ERR_ARG;
            }
        if (!x || ((uintptr_t)x & 15) || (d_perm && ((uintptr_t)d_perm & 15))) {
            set_err("mul_mat_fanout: null or misaligned x / perm");
            return TCQ4_ERR_ARG;
        }
        const WsLayout L = ws_layout(M, Nmax, K);
        uint8_t *wsb = reinterpret_cast<uint8_t *>(d->arena);
        tcq4::ActStage st;
        st.codes = reinterpret_cast<int8_t *>(wsb + L.codes_off);
        st.a7 = reinterpret_cast<float *>(wsb + L.a7_off);
        st.sumq = reinterpret_cast<int16_t *>(wsb + L.sumq_off);
        st.a7t = reinterpret_cast<float *>(wsb + L.a7t_off);
        const bool prof = g_prof.on && g_prof.n < g_prof.cap;
        if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 0], stream);
        cudaError_t e = tcq4::launch_act_quant_nib(x, d_perm, M, K, st, d->num_sms, stream);
        if (e != cudaSuccess) return cuda_fail(e, "activation kernel launch");
        if (prof) cudaEventRecord(g_prof.ev[3 * g_prof.n + 1], stream);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        for (int i0 = 0; i0 < n; i0 += 4) {
            const int ns = std::min(4, n - i0);
            tcq4::Gemm3Seg sg[4];
            int64_t Nl[4];
            for (int i = 0; i < ns; i++) {
                sg[i] = tcq4::Gemm3Seg{w_tiles[i0 + i], ys[i0 + i], Ns[i0 + i], ld[i0 + i]};
                Nl[i] = Ns[i0 + i];
            }
            tcq4::Gemm3Plan g3;
            if (!tcq4::gemm3_plan(M, Nl, ns, K, d->num_sms, d->smem_optin, false, &g3)) {
                set_err("pair GEMM (kind::f16) cannot be planned for this fan-out");
                return TCQ4_ERR_SHAPE;
            }
            e = tcq4::launch_gemm3(sg, ns, st, M, K, g3, nullptr, stream);
            if (e != cudaSuccess) return cuda_fail(e, "gemm3 fan-out launch");
            g_launches.fetch_add(1, std::memory_order_relaxed);
        }
        if (prof) {
            cudaEventRecord(g_prof.ev[3 * g_prof.n + 2], stream);
            g_prof.kind[g_prof.n++] = TCQ4_KERNEL_GEMM3;
        }
        return TCQ4_OK;
    }
    bool have_codes = false;
    for (int i = 0; i < n; i++) {
        bool done = false;
        tcq4::YOut yo = single_out(ys[i], Ns[i]);
        yo.ldc = ld[i];
        status = run_mul_mat(w_tiles[i], x, yo, M, Ns[i], K, d_perm, d->arena, d->arena_bytes,
                             TCQ4_KERNEL_AUTO, nullptr, stream, Nmax, have_codes, &done);
        if (status != TCQ4_OK) return status;
        have_codes = have_codes || done;
    }
    return TCQ4_OK;
}

// MUL_MAT_ID (ggml_mul_mat_id: dst[tok][slot][:] = W[ids[tok][slot]] * x[tok][slot % x_slots][:]), the operator the
// reference admits for TCQ4 in supports_op (ggml-cuda.cu:4457-4459) and serves through its generic sorted path
// (ggml_cuda_mul_mat_id, :2287-2436: ids to the host, rows sorted by expert, get_rows, one ggml_cuda_mul_mat per expert,
// get_rows back).
//  * n_tokens == 1 (decode): NO host round trip and ONE launch -- a GEMV over the virtual [n_used * N][K] matrix whose
//    channel blocks take their tiles from the expert the ids name, read by the kernel on the device; x and y are used
//    in place.  Capturable in a graph.  (Fallback when x_slots activation rows do not fit the GEMV's shared memory: one
//    GEMV per slot; slot 0 in plain stream order -- the ids come from the router kernel before it -- slots 1.. with
//    programmatic dependent launch: they cannot start before slot 0 has, i.e. not before the router has completed.)
//    Up to 8 tokens (a decode batch) take the same route, one grouped launch per token.
//  * more tokens: the reference's scheme -- ids are read on the host (stream synchronise; not capturable), rows are
//    counting-sorted by expert, gathered by one copy kernel, multiplied per expert through the normal dispatch (GEMV
//    for <= 16 rows, tcgen05 GEMMs above), and scattered back by the same copy kernel.
int tcq4_b200_mul_mat_id(const void *w_experts, int64_t n_expert, size_t expert_stride_bytes, const float *x,
                         int64_t x_slots, const int32_t *ids, int64_t ids_stride, int64_t n_used, int64_t n_tokens, float *y,
                         int64_t N, int64_t K, const int32_t *d_perm, int *err_flag, cudaStream_t stream) {
    if (n_expert < 1 || n_used < 1 || n_tokens < 1 || (n_tokens > 1 && ids_stride < n_used) ||
        (x_slots != 1 && x_slots != n_used)) {
        set_err("mul_mat_id: need n_expert, n_used, n_tokens >= 1, ids_stride >= n_used and x_slots in {1, n_used}");
        return TCQ4_ERR_ARG;
    }
    if (!shape_ok(1, N, K)) {
        set_err("unsupported shape N=%lld K=%lld (need K%%256==0, N%%8==0)", (long long)N, (long long)K);
        return TCQ4_ERR_SHAPE;
    }
    if (!w_experts || !x || !ids || !y || ((uintptr_t)w_experts & 15) || (expert_stride_bytes & 15) ||
        expert_stride_bytes < (size_t)(N / 8) * (size_t)(K / 256) * 1184 || ((uintptr_t)ids & 3)) {
        set_err("mul_mat_id: null / misaligned pointer or expert stride smaller than one matrix");
        return TCQ4_ERR_ARG;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    const int64_t R = n_tokens * n_used;
    tcq4::GemvPlan gv;
    static const bool no_group = getenv("TCQ4_MOE_NOGROUP") != nullptr;      // diagnostics / tests: one launch per slot
    constexpr int64_t MOE_GROUPED_MAX_TOKENS = 8;      // a decode batch: one grouped launch per token beats the host round trip
    if (n_tokens <= MOE_GROUPED_MAX_TOKENS && !no_group &&
        tcq4::gemv_plan(x_slots, n_used * N, K, d->num_sms, d->smem_optin, &gv)) {
        // ONE launch for all slots of a token: a GEMV over the virtual [n_used * N][K] matrix (launch_gemv_grouped)
        for (int64_t t = 0; t < n_tokens; t++) {
            const tcq4::GvExpert ex{ids + t * ids_stride, (long long)expert_stride_bytes, (int)n_expert, err_flag, 0,
                                    (int)x_slots,         (int)(N / 8)};
            cudaError_t e = tcq4::launch_gemv_grouped(w_experts, x + t * x_slots * K, d_perm, y + t * n_used * N, n_used, N, K,
                                                      gv, ex, stream);
            if (e != cudaSuccess) return cuda_fail(e, "grouped gemv launch (mul_mat_id)");
            g_launches.fetch_add(1, std::memory_order_relaxed);
        }
        return TCQ4_OK;
    }
    if (n_tokens == 1 && tcq4::gemv_plan(1, N, K, d->num_sms, d->smem_optin, &gv)) {
        const WsLayout L = ws_layout(1, N, K);
        status = arena_ensure(d, L.total, stream);
        if (status != TCQ4_OK) return status;
        uint8_t *wsb = reinterpret_cast<uint8_t *>(d->arena);
        for (int64_t s = 0; s < n_used; s++) {
            const tcq4::GvExpert ex{ids + s, (long long)expert_stride_bytes, (int)n_expert, err_flag, s > 0 ? 1 : 0, 0, 0};
            cudaError_t e = tcq4::launch_gemv(w_experts, x + (s % x_slots) * K, d_perm, single_out(y + s * N, N), 1, N, K, gv,
                                              nullptr, reinterpret_cast<int *>(wsb + L.ticket_off), nullptr, stream, &ex);
            if (e != cudaSuccess) return cuda_fail(e, "gemv launch (mul_mat_id)");
            g_launches.fetch_add(1, std::memory_order_relaxed);
        }
        return TCQ4_OK;
    }

    // ---- more tokens: routing on the DEVICE, three launches, no host round trip, capturable (round 2) ----------------------
    //  (1) one CTA bins the rows by expert into chunks of <= 64 rows at 64-aligned virtual rows (tcq4_moe_route_kernel);
    //  (2) K1 quantises the virtual rows straight from x through the row map (no gathered copy of x);
    //  (3) the batched-decode GEMM runs over (chunk, 128-channel block) pairs -- a chunk's tokens are its N, the expert's tiles
    //      come through a 3-D tensor map -- and scatters every token's row to its place in y.
    static const bool host_sort = getenv("TCQ4_MOE_HOSTSORT") != nullptr;      // the reference's scheme (below), for comparison
    if (!host_sort && (size_t)(2 * n_expert + 1) * sizeof(int) <= 48 * 1024 && d->smem_optin >= 227 * 1024 - 4096) {
        const int64_t max_chunks = n_expert + (R + 63) / 64;
        const int64_t vmax = max_chunks * 64;
        const WsLayout L = ws_layout(vmax, N, K);
        const size_t map_off = (size_t)tcq4::round_up((int64_t)L.total, 256);
        const size_t src_off = map_off + 256;                                               // [n_chunks]
        const size_t dst_off = src_off + (size_t)tcq4::round_up(vmax * 4, 256);
        const size_t ce_off = dst_off + (size_t)tcq4::round_up(vmax * 4, 256);
        const size_t pos_off = ce_off + (size_t)tcq4::round_up(max_chunks * 4, 256);
        const size_t total = pos_off + (size_t)tcq4::round_up(R * 4, 256);
        status = arena_ensure(d, total, stream);
        if (status != TCQ4_OK) return status;
        uint8_t *base = reinterpret_cast<uint8_t *>(d->arena);
        int *n_chunks = reinterpret_cast<int *>(base + map_off);
        int32_t *src_map = reinterpret_cast<int32_t *>(base + src_off), *dst_map = reinterpret_cast<int32_t *>(base + dst_off);
        int *chunk_expert = reinterpret_cast<int *>(base + ce_off);
        int32_t *pos_tmp = reinterpret_cast<int32_t *>(base + pos_off);
        cudaError_t e = tcq4::launch_moe_route(ids, n_tokens == 1 ? n_used : ids_stride, n_used, n_tokens, n_expert, x_slots,
                                               (int)max_chunks, n_chunks, chunk_expert, src_map, dst_map, pos_tmp, err_flag, stream);
        if (e != cudaSuccess) return cuda_fail(e, "mul_mat_id: routing launch");
        tcq4::ActStage st;
        st.codes = reinterpret_cast<int8_t *>(base + L.codes_off);
        st.a7 = reinterpret_cast<float *>(base + L.a7_off);
        st.sumq = reinterpret_cast<int16_t *>(base + L.sumq_off);
        st.a7t = reinterpret_cast<float *>(base + L.a7t_off);
        e = tcq4::launch_act_quant_nib_rows(x, d_perm, vmax, K, st, d->num_sms, stream, src_map, n_chunks);
        if (e != cudaSuccess) return cuda_fail(e, "mul_mat_id: activation kernel launch");
        e = tcq4::launch_gemm4_id(w_experts, n_expert, expert_stride_bytes, st, y, N, K, n_chunks, chunk_expert, dst_map,
                                  (int)max_chunks, d->num_sms, stream);
        if (e != cudaSuccess) return cuda_fail(e, "mul_mat_id: grouped GEMM launch");
        g_launches.fetch_add(3, std::memory_order_relaxed);
        return TCQ4_OK;
    }

    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (stream != nullptr && cudaStreamIsCapturing(stream, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone) {
        set_err("mul_mat_id (host-sorted path) reads the ids on the host (as ggml_cuda_mul_mat_id does): not capturable");
        return TCQ4_ERR_ARG;
    }
    // ---- ids to the host, counting sort by expert (ggml-cuda.cu:2349-2368) -----------------------------------------
    if (n_tokens == 1) ids_stride = n_used;
    std::vector<int32_t> ids_host((size_t)((n_tokens - 1) * ids_stride + n_used));
    cudaError_t e = cudaMemcpyAsync(ids_host.data(), ids, ids_host.size() * 4, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return cuda_fail(e, "mul_mat_id: ids to host");
    std::vector<int64_t> count((size_t)n_expert + 1, 0);
    for (int64_t t = 0; t < n_tokens; t++)
        for (int64_t s = 0; s < n_used; s++) {
            const int32_t ex = ids_host[(size_t)(t * ids_stride + s)];
            if (ex < 0 || ex >= n_expert) {
                set_err("mul_mat_id: expert id %d of token %lld slot %lld outside [0, %lld)", ex, (long long)t, (long long)s,
                        (long long)n_expert);
                return TCQ4_ERR_ARG;
            }
            count[(size_t)ex + 1]++;
        }
    for (int64_t i = 0; i < n_expert; i++) count[(size_t)i + 1] += count[(size_t)i];     // count[e] = first sorted row of e
    std::vector<int32_t> idx((size_t)2 * R);            // [0, R): source row of x per sorted row; [R, 2R): row of dst
    {
        std::vector<int64_t> fill(count.begin(), count.end() - 1);
        for (int64_t t = 0; t < n_tokens; t++)
            for (int64_t s = 0; s < n_used; s++) {
                const int32_t ex = ids_host[(size_t)(t * ids_stride + s)];
                const int64_t r = fill[(size_t)ex]++;
                idx[(size_t)r] = (int32_t)(t * x_slots + s % x_slots);
                idx[(size_t)(R + r)] = (int32_t)(t * n_used + s);
            }
    }
    // ---- scratch: [MUL_MAT workspace][x sorted R x K][y sorted R x N][2R indices] ---------------------------------
    size_t ws_need = ws_layout(R, N, K).total;
    for (int64_t m : {(int64_t)8, (int64_t)16}) ws_need = std::max(ws_need, ws_layout(std::min(R, m), N, K).total);
    const size_t xs_off = (size_t)tcq4::round_up((int64_t)ws_need, 256);
    const size_t ys_off = xs_off + (size_t)tcq4::round_up(R * K * 4, 256);
    const size_t ix_off = ys_off + (size_t)tcq4::round_up(R * N * 4, 256);
    const size_t total = ix_off + (size_t)tcq4::round_up(2 * R * 4, 256);
    status = arena_ensure(d, total, stream);
    if (status != TCQ4_OK) return status;
    uint8_t *base = reinterpret_cast<uint8_t *>(d->arena);
    float *xs = reinterpret_cast<float *>(base + xs_off), *ys = reinterpret_cast<float *>(base + ys_off);
    int32_t *ix = reinterpret_cast<int32_t *>(base + ix_off);
    e = cudaMemcpyAsync(ix, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);          // idx is a local: the copy must have left it
    if (e != cudaSuccess) return cuda_fail(e, "mul_mat_id: indices to device");
    e = tcq4::launch_copy_rows(x, ix, xs, nullptr, R, K, d->num_sms, stream);
    if (e != cudaSuccess) return cuda_fail(e, "mul_mat_id: gather");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    for (int64_t ex = 0; ex < n_expert; ex++) {
        const int64_t r0 = count[(size_t)ex], m = count[(size_t)ex + 1] - r0;
        if (m == 0) continue;
        status = run_mul_mat(reinterpret_cast<const uint8_t *>(w_experts) + (size_t)ex * expert_stride_bytes, xs + r0 * K,
                             single_out(ys + r0 * N, N), m, N, K, d_perm, base, ws_need, TCQ4_KERNEL_AUTO, nullptr, stream);
        if (status != TCQ4_OK) return status;
    }
    e = tcq4::launch_copy_rows(ys, nullptr, y, ix + R, R, N, d->num_sms, stream);
    if (e != cudaSuccess) return cuda_fail(e, "mul_mat_id: scatter");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return TCQ4_OK;
}

// N-sharded MUL_MAT with the all-gather fused away: the N columns this GPU owns are written, by the kernel's own
// epilogue, at column col0 of EVERY destination matrix (row-major [M][ldc] fp32).  With peer-mapped pointers (one per
// GPU of the node) every GPU ends up with the full C without a collective on the path; the caller orders consumers
// behind all producers (one cross-GPU barrier per step or per matmul).  n_peers = 1 is a strided single-GPU store.
int tcq4_b200_mul_mat_peers(const void *w_tiles, const float *x, float *const *y_peers, int n_peers, int64_t ldc,
                            int64_t col0, int64_t M, int64_t N, int64_t K, const int32_t *d_perm, cudaStream_t stream) {
    if (!shape_ok(M, N, K) || !y_peers || n_peers < 1 || n_peers > tcq4::MAX_PEERS || col0 < 0 || (col0 & 3) ||
        col0 + N > ldc) {
        set_err("mul_mat_peers: bad shape, destination count (1..%d) or column window", tcq4::MAX_PEERS);
        return TCQ4_ERR_SHAPE;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    status = arena_ensure(d, ws_layout(M, N, K).total, stream);
    if (status != TCQ4_OK) return status;
    tcq4::YOut o;
    for (int i = 0; i < tcq4::MAX_PEERS; i++) o.p[i] = y_peers[i < n_peers ? i : 0] ? y_peers[i < n_peers ? i : 0] + col0 : nullptr;
    o.n = n_peers;
    o.ldc = ldc;
    o.mc = nullptr;
    return run_mul_mat(w_tiles, x, o, M, N, K, d_perm, d->arena, d->arena_bytes, TCQ4_KERNEL_AUTO, nullptr, stream);
}

// Same exchange through the switch: y_multicast is the multicast (NVLS) address of the symmetric [M][ldc] C; every
// store of the epilogue is a multimem.st that the NVSwitch replicates into all GPUs' copies, this GPU's included.
int tcq4_b200_mul_mat_multicast(const void *w_tiles, const float *x, float *y_multicast, int64_t ldc, int64_t col0,
                                int64_t M, int64_t N, int64_t K, const int32_t *d_perm, cudaStream_t stream) {
    if (!shape_ok(M, N, K) || !y_multicast || col0 < 0 || (col0 & 3) || col0 + N > ldc) {
        set_err("mul_mat_multicast: bad shape or column window");
        return TCQ4_ERR_SHAPE;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    status = arena_ensure(d, ws_layout(M, N, K).total, stream);
    if (status != TCQ4_OK) return status;
    tcq4::YOut o = single_out(y_multicast + col0, N);
    o.ldc = ldc;
    o.mc = y_multicast + col0;
    return run_mul_mat(w_tiles, x, o, M, N, K, d_perm, d->arena, d->arena_bytes, TCQ4_KERNEL_AUTO, nullptr, stream);
}

// Cross-GPU barrier of the exchange: every rank calls it with the same peer-mapped flag arrays (int32[TCQ4_B200_MAX_PEERS + 1]
// per rank, zeroed once) after the stores / pushes of the outputs its consumers need; returns (on the stream) when every rank
// has.  One 32-thread kernel, capturable; see tcq4_peer.cu.
int tcq4_b200_peer_barrier(int *const *flags_peers, int n_peers, int rank, cudaStream_t stream) {
    int status;
    if (!current_device(&status)) return status;
    cudaError_t e = tcq4::launch_peer_barrier(flags_peers, n_peers, rank, stream);
    if (e != cudaSuccess) return cuda_fail(e, "peer barrier launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return TCQ4_OK;
}

// Copy-engine half of the same exchange: push `ncols` columns starting at col0 of a local row-major [M][ldc] fp32
// matrix into the same window of a (peer-mapped) destination matrix.  One 2-D async copy, no SM involved; with the
// plain MUL_MAT writing the local C and this call per peer on side streams, the transfer of matmul i overlaps the
// kernels of matmul i+1.
int tcq4_b200_push_columns(float *dst_base, const float *src_base, int64_t M, int64_t ldc, int64_t col0, int64_t ncols,
                           cudaStream_t stream) {
    if (!dst_base || !src_base || M < 1 || ncols < 1 || col0 < 0 || col0 + ncols > ldc) {
        set_err("push_columns: bad window");
        return TCQ4_ERR_ARG;
    }
    cudaError_t e = cudaMemcpy2DAsync(dst_base + col0, (size_t)ldc * 4, src_base + col0, (size_t)ldc * 4, (size_t)ncols * 4,
                                      (size_t)M, cudaMemcpyDefault, stream);
    if (e != cudaSuccess) return cuda_fail(e, "push_columns");
    return TCQ4_OK;
}

int tcq4_b200_mul_mat_host(const void *w_tiles_dev, const float *x_host, float *y_host, int64_t M, int64_t N,
                           int64_t K, const int32_t *d_perm, cudaStream_t stream) {
    if (!shape_ok(M, N, K) || !x_host || !y_host) {
        set_err("mul_mat_host: bad shape or null host pointer");
        return TCQ4_ERR_SHAPE;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    const size_t xb = (size_t)M * K * 4, yb = (size_t)M * N * 4;
    if (d->hx_bytes < xb) {
        cudaDeviceSynchronize();
        if (d->hx) cudaFree(d->hx);
        d->hx = nullptr; d->hx_bytes = 0;
        cudaError_t e = cudaMalloc(&d->hx, xb);
        if (e != cudaSuccess) return cuda_fail(e, "x staging cudaMalloc");
        d->hx_bytes = xb;
    }
    if (d->hy_bytes < yb) {
        cudaDeviceSynchronize();
        if (d->hy) cudaFree(d->hy);
        d->hy = nullptr; d->hy_bytes = 0;
        cudaError_t e = cudaMalloc(&d->hy, yb);
        if (e != cudaSuccess) return cuda_fail(e, "y staging cudaMalloc");
        d->hy_bytes = yb;
    }
    cudaError_t e = cudaMemcpyAsync(d->hx, x_host, xb, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return cuda_fail(e, "H2D copy");
    status = tcq4_b200_mul_mat(w_tiles_dev, d->hx, d->hy, M, N, K, d_perm, stream);
    if (status != TCQ4_OK) return status;
    e = cudaMemcpyAsync(y_host, d->hy, yb, cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) return cuda_fail(e, "D2H copy");
    e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return cuda_fail(e, "stream sync");
    return TCQ4_OK;
}

// ---- pipelined host-buffer path ---------------------------------------------------------------------------------
// A caller with HOST activations (the end-to-end leg of bench.py; a CPU-resident ggml graph handing single
// MUL_MATs to this device) queues matmuls back to back: the H2D copy of call i+1, the kernels of call i and the D2H
// copy of call i-1 run on three library-owned streams over double-buffered staging, so PCIe moves in both
// directions while the SMs compute.  Results are valid after tcq4_b200_host_sync().
namespace {
struct HostPipe {
    bool ready = false;
    cudaStream_t s_in = nullptr, s_cmp = nullptr, s_out = nullptr;
    cudaEvent_t in_done[2] = {nullptr, nullptr}, cmp_done[2] = {nullptr, nullptr}, out_done[2] = {nullptr, nullptr};
    float *hx[2] = {nullptr, nullptr}, *hy[2] = {nullptr, nullptr};
    size_t hx_bytes[2] = {0, 0}, hy_bytes[2] = {0, 0};
    float *ws = nullptr;              // the pipeline's OWN workspace: its compute stream must not share the device arena
    size_t ws_bytes = 0;              // with whatever stream the caller runs tcq4_b200_mul_mat on
    uint64_t n = 0;
};
HostPipe g_pipe[MAX_DEV];

cudaError_t pipe_init(HostPipe &p) {
    if (p.ready) return cudaSuccess;
    cudaError_t e;
    if ((e = cudaStreamCreateWithFlags(&p.s_in, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if ((e = cudaStreamCreateWithFlags(&p.s_cmp, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if ((e = cudaStreamCreateWithFlags(&p.s_out, cudaStreamNonBlocking)) != cudaSuccess) return e;
    for (int i = 0; i < 2; i++) {
        if ((e = cudaEventCreateWithFlags(&p.in_done[i], cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&p.cmp_done[i], cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&p.out_done[i], cudaEventDisableTiming)) != cudaSuccess) return e;
    }
    p.ready = true;
    return cudaSuccess;
}
cudaError_t pipe_grow(HostPipe &p, float **buf, size_t *have, size_t need) {
    if (*have >= need) return cudaSuccess;
    cudaError_t e = cudaDeviceSynchronize();      // rare: only when a larger shape than any before arrives
    if (e != cudaSuccess) return e;
    if (*buf) cudaFree(*buf);
    *buf = nullptr;
    *have = 0;
    e = cudaMalloc(buf, need);
    if (e == cudaSuccess) *have = need;
    (void)p;
    return e;
}
}  // namespace

int tcq4_b200_mul_mat_host_async(const void *w_tiles_dev, const float *x_host, float *y_host, int64_t M, int64_t N,
                                 int64_t K, const int32_t *d_perm) {
    if (!shape_ok(M, N, K) || !x_host || !y_host) {
        set_err("mul_mat_host_async: bad shape or null host pointer");
        return TCQ4_ERR_SHAPE;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    int dev = 0;
    cudaGetDevice(&dev);
    HostPipe &p = g_pipe[dev];
    cudaError_t e = pipe_init(p);
    if (e != cudaSuccess) return cuda_fail(e, "host pipeline setup");
    const size_t xb = (size_t)M * K * 4, yb = (size_t)M * N * 4;
    const int slot = (int)(p.n & 1);
    if ((e = pipe_grow(p, &p.hx[slot], &p.hx_bytes[slot], xb)) != cudaSuccess) return cuda_fail(e, "x staging cudaMalloc");
    if ((e = pipe_grow(p, &p.hy[slot], &p.hy_bytes[slot], yb)) != cudaSuccess) return cuda_fail(e, "y staging cudaMalloc");
    if (p.n >= 2) {
        cudaStreamWaitEvent(p.s_in, p.cmp_done[slot], 0);      // the kernels of call n-2 have read this x slot
        cudaStreamWaitEvent(p.s_cmp, p.out_done[slot], 0);     // the D2H copy of call n-2 has drained this y slot
    }
    if ((e = cudaMemcpyAsync(p.hx[slot], x_host, xb, cudaMemcpyHostToDevice, p.s_in)) != cudaSuccess)
        return cuda_fail(e, "H2D copy");
    cudaEventRecord(p.in_done[slot], p.s_in);
    cudaStreamWaitEvent(p.s_cmp, p.in_done[slot], 0);
    {
        size_t need = ws_layout(M, N, K).total;
        for (int64_t m : {(int64_t)8, (int64_t)16}) need = std::max(need, ws_layout(std::min(M, m), N, K).total);
        if (p.ws_bytes < need) {
            if ((e = pipe_grow(p, &p.ws, &p.ws_bytes, need)) != cudaSuccess) return cuda_fail(e, "pipeline workspace cudaMalloc");
            if ((e = cudaMemset(p.ws, 0, p.ws_bytes)) != cudaSuccess) return cuda_fail(e, "pipeline workspace memset");
        }
    }
    status = tcq4_b200_mul_mat_ws(w_tiles_dev, p.hx[slot], p.hy[slot], M, N, K, d_perm, p.ws, p.ws_bytes, TCQ4_KERNEL_AUTO, p.s_cmp);
    if (status != TCQ4_OK) return status;
    cudaEventRecord(p.cmp_done[slot], p.s_cmp);
    cudaStreamWaitEvent(p.s_out, p.cmp_done[slot], 0);
    if ((e = cudaMemcpyAsync(y_host, p.hy[slot], yb, cudaMemcpyDeviceToHost, p.s_out)) != cudaSuccess)
        return cuda_fail(e, "D2H copy");
    cudaEventRecord(p.out_done[slot], p.s_out);
    p.n++;
    return TCQ4_OK;
}

// The same pipeline for a fan-out group (consecutive MUL_MAT nodes that share src1: Q/K/V, gate/up): x crosses PCIe ONCE for
// the group, the outputs come back one copy per matrix.
int tcq4_b200_mul_mat_fanout_host_async(int n, const void *const *w_tiles_dev, const int64_t *Ns, float *const *y_host,
                                        const float *x_host, int64_t M, int64_t K, const int32_t *d_perm) {
    if (n < 1 || n > 8 || !w_tiles_dev || !Ns || !y_host || !x_host) {
        set_err("mul_mat_fanout_host_async: 1..8 matrices, non-null tables and host pointers");
        return TCQ4_ERR_ARG;
    }
    int64_t Nsum = 0, Nmax = 0;
    for (int i = 0; i < n; i++) {
        if (!shape_ok(M, Ns[i], K) || !y_host[i]) {
            set_err("mul_mat_fanout_host_async: bad shape or null host pointer of matrix %d", i);
            return TCQ4_ERR_SHAPE;
        }
        Nsum += Ns[i];
        Nmax = std::max(Nmax, Ns[i]);
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    int dev = 0;
    cudaGetDevice(&dev);
    HostPipe &p = g_pipe[dev];
    cudaError_t e = pipe_init(p);
    if (e != cudaSuccess) return cuda_fail(e, "host pipeline setup");
    const size_t xb = (size_t)M * K * 4, yb = (size_t)M * Nsum * 4;
    const int slot = (int)(p.n & 1);
    if ((e = pipe_grow(p, &p.hx[slot], &p.hx_bytes[slot], xb)) != cudaSuccess) return cuda_fail(e, "x staging cudaMalloc");
    if ((e = pipe_grow(p, &p.hy[slot], &p.hy_bytes[slot], yb)) != cudaSuccess) return cuda_fail(e, "y staging cudaMalloc");
    status = tcq4_b200_reserve(M, Nmax, K);      // the fan-out's scratch is the device arena (grown here, outside the streams' work)
    if (status != TCQ4_OK) return status;
    if (p.n >= 2) {
        cudaStreamWaitEvent(p.s_in, p.cmp_done[slot], 0);      // the kernels of call n-2 have read this x slot
        cudaStreamWaitEvent(p.s_cmp, p.out_done[slot], 0);     // the D2H copies of call n-2 have drained this y slot
    }
    if ((e = cudaMemcpyAsync(p.hx[slot], x_host, xb, cudaMemcpyHostToDevice, p.s_in)) != cudaSuccess)
        return cuda_fail(e, "H2D copy");
    cudaEventRecord(p.in_done[slot], p.s_in);
    cudaStreamWaitEvent(p.s_cmp, p.in_done[slot], 0);
    float *ys[8];
    {
        float *yd = reinterpret_cast<float *>(p.hy[slot]);
        for (int i = 0; i < n; i++) {
            ys[i] = yd;
            yd += (size_t)M * Ns[i];
        }
    }
    status = tcq4_b200_mul_mat_fanout(n, w_tiles_dev, Ns, ys, reinterpret_cast<const float *>(p.hx[slot]), M, K, d_perm, p.s_cmp);
    if (status != TCQ4_OK) return status;
    cudaEventRecord(p.cmp_done[slot], p.s_cmp);
    cudaStreamWaitEvent(p.s_out, p.cmp_done[slot], 0);
    for (int i = 0; i < n; i++)
        if ((e = cudaMemcpyAsync(y_host[i], ys[i], (size_t)M * Ns[i] * 4, cudaMemcpyDeviceToHost, p.s_out)) != cudaSuccess)
            return cuda_fail(e, "D2H copy");
    cudaEventRecord(p.out_done[slot], p.s_out);
    p.n++;
    return TCQ4_OK;
}

// ---- input-side fusion: the op that produced the activations, applied while K1 loads them -----------------------
// RMSNorm * weight in front of Q/K/V and gate/up (build_norm, llama-graph.cpp; kernel rms_norm_f32<.., do_multiply>, norm.cu:75-150)
// and SwiGLU in front of down (build_ffn, llama-graph.cpp:731-755; unary_gated_op_kernel<op_silu>, unary.cu:255-267).  One fp32
// round trip of the [M][K] activations less; results are bit-identical to applying tcq4_b200_apply_input and then the plain
// entry, because the fused load evaluates the same expression in the same order.  M <= 8 (decode GEMV, whose activation stage
// lives inside the kernel) takes exactly that unfused route.
int tcq4_b200_rms_row_scale(const float *x, float *row_scale, int64_t M, int64_t K, float eps, cudaStream_t stream) {
    if (!x || !row_scale || M < 1 || K < 1) {
        set_err("rms_row_scale: null pointer or empty shape");
        return TCQ4_ERR_ARG;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    cudaError_t e = tcq4::launch_rms_row_scale(x, row_scale, M, K, eps, stream);
    if (e != cudaSuccess) return cuda_fail(e, "rms_row_scale launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return TCQ4_OK;
}

static int make_act_input(int mode, const float *x2, const float *row_scale, const float *col_weight, tcq4::ActInput *in) {
    in->mode = mode;
    in->x2 = x2;
    in->row_scale = row_scale;
    in->col_weight = col_weight;
    if (mode == TCQ4_INPUT_RMS_NORM_MUL && row_scale && col_weight) return TCQ4_OK;
    if (mode == TCQ4_INPUT_SWIGLU && x2) return TCQ4_OK;
    set_err("input transform: mode %d needs %s", mode, mode == TCQ4_INPUT_SWIGLU ? "x2" : "row_scale and col_weight");
    return TCQ4_ERR_ARG;
}

int tcq4_b200_apply_input(int mode, const float *x, const float *x2, const float *row_scale, const float *col_weight, float *out,
                          int64_t M, int64_t K, cudaStream_t stream) {
    tcq4::ActInput in;
    int status = make_act_input(mode, x2, row_scale, col_weight, &in);
    if (status != TCQ4_OK) return status;
    if (!x || !out || M < 1 || K < 1) {
        set_err("apply_input: null pointer or empty shape");
        return TCQ4_ERR_ARG;
    }
    DevInfo *d = current_device(&status);
    if (!d) return status;
    cudaError_t e = tcq4::launch_apply_input(in, x, out, M, K, d->num_sms, stream);
    if (e != cudaSuccess) return cuda_fail(e, "apply_input launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return TCQ4_OK;
}

int tcq4_b200_mul_mat_fanout_input(int n, const void *const *w_tiles, const int64_t *Ns, float *const *ys, const int64_t *ldcs,
                                   const float *x, int mode, const float *x2, const float *row_scale, const float *col_weight,
                                   int64_t M, int64_t K, const int32_t *d_perm, cudaStream_t stream) {
    tcq4::ActInput in;
    int status = make_act_input(mode, x2, row_scale, col_weight, &in);
    if (status != TCQ4_OK) return status;
    if (n < 1 || n > 8 || !Ns) {
        set_err("mul_mat_fanout_input: 1..8 weight matrices");
        return TCQ4_ERR_ARG;
    }
    int64_t Nmax = 0, Nsum = 0;
    for (int i = 0; i < n; i++) {
        if (!shape_ok(M, Ns[i], K)) {
            set_err("unsupported shape M=%lld N=%lld K=%lld", (long long)M, (long long)Ns[i], (long long)K);
            return TCQ4_ERR_SHAPE;
        }
        Nmax = std::max(Nmax, Ns[i]);
        Nsum += Ns[i];
    }
    DevInfo *d = current_device(&status);
    if (!d) return status;
    static const bool g2_fuse = getenv("TCQ4_G2_FUSE") != nullptr;      // (that opt-in runs the activation stage inside the GEMM)
    if (select_kernel(d, M, Nmax, K, nullptr) == TCQ4_KERNEL_GEMV || g2_fuse) {
        // decode: transform into scratch behind the MUL_MAT's own workspace, then the plain fan-out
        size_t tail = std::max(ws_layout(M, Nmax, K).total, shape_ok(M, Nsum, K) ? ws_layout(M, Nsum, K).total : (size_t)0);
        tail = (size_t)tcq4::round_up((int64_t)tail, 256);
        status = arena_ensure(d, tail + (size_t)M * K * 4, stream);
        if (status != TCQ4_OK) return status;
        float *xt = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(d->arena) + tail);
        cudaError_t e = tcq4::launch_apply_input(in, x, xt, M, K, d->num_sms, stream);
        if (e != cudaSuccess) return cuda_fail(e, "apply_input launch");
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return tcq4_b200_mul_mat_fanout_ld(n, w_tiles, Ns, ys, ldcs, xt, M, K, d_perm, stream);
    }
    tcq4::set_pending_act_input(&in);      // picked up by the one activation-stage launch of the call below
    status = tcq4_b200_mul_mat_fanout_ld(n, w_tiles, Ns, ys, ldcs, x, M, K, d_perm, stream);
    tcq4::set_pending_act_input(nullptr);
    return status;
}

int tcq4_b200_host_sync(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) return TCQ4_ERR_NO_DEVICE;
    HostPipe &p = g_pipe[dev];
    if (!p.ready) return TCQ4_OK;
    cudaError_t e = cudaStreamSynchronize(p.s_out);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p.s_cmp);
    if (e != cudaSuccess) return cuda_fail(e, "host pipeline sync");
    return TCQ4_OK;
}

int tcq4_b200_fwht_quantize_tc(const float *x, void *y_blocks, const int32_t *d_perm, int64_t K, int64_t M,
                               cudaStream_t stream) {
    if (M < 1 || K < 256 || (K % 256) != 0) {
        set_err("fwht_quantize_tc: need M>=1 and K%%256==0");
        return TCQ4_ERR_SHAPE;
    }
    if (!x || !y_blocks || ((uintptr_t)x & 15) || ((uintptr_t)y_blocks & 15) || (d_perm && ((uintptr_t)d_perm & 15))) {
        set_err("fwht_quantize_tc: null or misaligned pointer");
        return TCQ4_ERR_ARG;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    cudaError_t e = tcq4::launch_act_quant_ref(x, d_perm, M, K, y_blocks, d->num_sms, stream);
    if (e != cudaSuccess) return cuda_fail(e, "activation kernel launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return TCQ4_OK;
}

size_t tcq4_b200_quantize_size(int64_t nrow, int64_t K) {
    if (nrow < 0 || K < 0 || (K % 256) != 0) return 0;
    return (size_t)((nrow + 7) / 8) * (size_t)(K / 256) * TCQ4_B200_TILE_BYTES;
}

int tcq4_b200_quantize_tiles(const float *w, void *tiles, int64_t nrow, int64_t K, const int32_t *d_perm,
                             cudaStream_t stream) {
    if (nrow < 1 || K < 256 || (K % 256) != 0 || K > (int64_t)1 << 30) {
        set_err("quantize_tiles: need nrow >= 1 and K %% 256 == 0 (got nrow=%lld K=%lld)", (long long)nrow, (long long)K);
        return TCQ4_ERR_SHAPE;
    }
    if (!w || !tiles || ((uintptr_t)w & 15) || ((uintptr_t)tiles & 15) || (d_perm && ((uintptr_t)d_perm & 15))) {
        set_err("quantize_tiles: null or misaligned pointer (16-byte alignment)");
        return TCQ4_ERR_ARG;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    cudaError_t e = tcq4::launch_quantize_tiles(w, d_perm, tiles, nrow, K, d->num_sms, stream);
    if (e != cudaSuccess) return cuda_fail(e, "quantiser launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return TCQ4_OK;
}

int tcq4_b200_get_rows(const void *w_tiles, const int32_t *rows, float *dst, int64_t n_idx, int64_t nrows_src, int64_t K,
                       int64_t dst_stride, int *err_flag, cudaStream_t stream) {
    if (n_idx < 0 || nrows_src < 1 || K < 256 || (K % 256) != 0 || K > (int64_t)1 << 30 || dst_stride < K || (dst_stride & 3)) {
        set_err("get_rows: need K %% 256 == 0, nrows_src >= 1, dst_stride >= K and a multiple of 4");
        return TCQ4_ERR_SHAPE;
    }
    if (n_idx == 0) return TCQ4_OK;
    if (!w_tiles || !rows || !dst || ((uintptr_t)w_tiles & 15) || ((uintptr_t)dst & 15) || ((uintptr_t)rows & 3)) {
        set_err("get_rows: null or misaligned pointer");
        return TCQ4_ERR_ARG;
    }
    int status;
    DevInfo *d = current_device(&status);
    if (!d) return status;
    cudaError_t e = tcq4::launch_get_rows(w_tiles, rows, dst, n_idx, nrows_src, K, dst_stride, err_flag, d->num_sms, stream);
    if (e != cudaSuccess) return cuda_fail(e, "get_rows launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return TCQ4_OK;
}

int tcq4_b200_perm_from_imatrix(const float *imatrix_host, int64_t K, int32_t *perm_host) {
    if (!imatrix_host || !perm_host || K < 1 || K > (int64_t)1 << 30) {
        set_err("perm_from_imatrix: null pointer or bad K");
        return TCQ4_ERR_ARG;
    }
    for (int64_t i = 0; i < K; i++) perm_host[i] = (int32_t)i;
    for (int64_t b0 = 0; b0 < K; b0 += 256) {
        const int64_t b1 = b0 + 256 < K ? b0 + 256 : K;
        std::stable_sort(perm_host + b0, perm_host + b1,
                         [imatrix_host](int32_t a, int32_t b) { return imatrix_host[a] > imatrix_host[b]; });
    }
    return TCQ4_OK;
}

#ifdef TCQ4_DIAG
int tcq4_b200_debug_gemm2_timeline(long long *out, size_t bytes) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = tcq4::gemm2_read_prof(out, bytes);
    return e == cudaSuccess ? TCQ4_OK : TCQ4_ERR_CUDA;
}

int tcq4_b200_debug_gemm3_timeline(long long *out, size_t bytes) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = tcq4::gemm3_read_prof(out, bytes);
    return e == cudaSuccess ? TCQ4_OK : TCQ4_ERR_CUDA;
}

int tcq4_b200_debug_gemm4_timeline(long long *out, size_t bytes) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = tcq4::gemm4_read_prof(out, bytes);
    return e == cudaSuccess ? TCQ4_OK : TCQ4_ERR_CUDA;
}

int tcq4_b200_debug_gemv_phases(unsigned long long *out, size_t bytes) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = tcq4::gemv_read_prof(out, bytes);
    return e == cudaSuccess ? TCQ4_OK : cuda_fail(e, "gemv phase read");
}
#endif

int tcq4_b200_profile_begin(int max_launches) {
    if (max_launches < 1 || max_launches > (1 << 20)) {
        set_err("profile_begin: max_launches out of range");
        return TCQ4_ERR_ARG;
    }
    int status;
    if (!current_device(&status)) return status;
    if (g_prof.cap < max_launches) {
        for (int i = 0; i < 3 * g_prof.cap; i++) cudaEventDestroy(g_prof.ev[i]);
        delete[] g_prof.ev;
        delete[] g_prof.kind;
        g_prof.ev = new cudaEvent_t[3 * (size_t)max_launches];
        g_prof.kind = new int[max_launches];
        for (int i = 0; i < 3 * max_launches; i++) {
            cudaError_t e = cudaEventCreate(&g_prof.ev[i]);
            if (e != cudaSuccess) return cuda_fail(e, "profile event create");
        }
        g_prof.cap = max_launches;
    }
    g_prof.n = 0;
    g_prof.on = true;
    return TCQ4_OK;
}

int tcq4_b200_profile_end(int kernel, double *main_ms_sum, double *pre_ms_sum, int *launches) {
    g_prof.on = false;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return cuda_fail(e, "profile sync");
    double a = 0.0, b = 0.0;
    int n = 0;
    for (int i = 0; i < g_prof.n; i++) {
        if (kernel != TCQ4_KERNEL_AUTO && g_prof.kind[i] != kernel) continue;
        float pre = 0.0f, mn = 0.0f;
        cudaEventElapsedTime(&pre, g_prof.ev[3 * i + 0], g_prof.ev[3 * i + 1]);
        cudaEventElapsedTime(&mn, g_prof.ev[3 * i + 1], g_prof.ev[3 * i + 2]);
        a += mn;
        b += pre;
        n++;
    }
    if (main_ms_sum) *main_ms_sum = a;
    if (pre_ms_sum) *pre_ms_sum = b;
    if (launches) *launches = n;
    return TCQ4_OK;
}

int tcq4_b200_shard_range(int64_t N, int64_t K, int world, int rank, int64_t align, int64_t *n0, int64_t *n1,
                          size_t *byte_offset, size_t *byte_len) {
    if (world < 1 || rank < 0 || rank >= world || align < 8 || (align % 8) != 0 || N < 8 || (N % 8) != 0 ||
        K < 256 || (K % 256) != 0) {
        set_err("shard_range: bad arguments");
        return TCQ4_ERR_ARG;
    }
    const int64_t units = (N + align - 1) / align;          // last unit may be short
    const int64_t base = units / world, rem = units % world;
    const int64_t u0 = rank * base + (rank < rem ? rank : rem);
    const int64_t u1 = u0 + base + (rank < rem ? 1 : 0);
    int64_t a = u0 * align, b = u1 * align;
    if (a > N) a = N;
    if (b > N) b = N;
    if (n0) *n0 = a;
    if (n1) *n1 = b;
    const size_t row_bytes = (size_t)(K / 256) * 148;       // ggml_row_size (ggml.c:750-757)
    if (byte_offset) *byte_offset = (size_t)a * row_bytes;
    if (byte_len) *byte_len = (size_t)(b - a) * row_bytes;
    return TCQ4_OK;
}

}  // extern "C"
