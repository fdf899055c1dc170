// extern "C" surface of libvq_b200.so (declared in include/vq_b200.h).  Argument checking, workspace carve-up
// and dispatch only; the kernels live in vq_assign_tc.cu / vq_assign_simt.cu / vq_pass.cu / vq_ema.cu.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "vq_common.cuh"

namespace vq {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static std::atomic<int> g_pdl_mask{-1};
int pdl_mask() {
    int m = g_pdl_mask.load(std::memory_order_relaxed);
    if (m < 0) {
#ifdef VQ_TUNING
        const char* ev = getenv("VQ_PDL");           // tuning build only; the product takes vq_set_pdl_mask()
        m = ev == nullptr ? 31 : (atoi(ev) & 31);
#else
        m = 31;
#endif
        g_pdl_mask.store(m, std::memory_order_relaxed);
    }
    return m;
}

// Cross-rank rendezvous bound of the fused peer exchange: 30 minutes unless the host says otherwise (a torch process
// group's own default is 10 minutes for NCCL; VectorQuantizer passes the group's timeout through vq_set_peer_timeout_ms).
static std::atomic<long long> g_peer_timeout_ns{1800ll * 1000000000ll};
long long peer_timeout_ns() { return g_peer_timeout_ns.load(std::memory_order_relaxed); }

int current_device() {
    int dev = -1;
    return cudaGetDevice(&dev) == cudaSuccess ? dev : -1;
}

int sm_count() {
    static int cached_dev = -1, cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached = n;
        cached_dev = dev;
    }
    return cached;
}

int launch_ema_update(float* E_f32, float* ema_cnt, float* ema_sum, const float* code_cnt, const float* code_sum,
                      int K, int D, float decay, float eps, int do_ema, int search, void* E_bf16, float* e_norm2,
                      int64_t* usage_total, int32_t* steps_since_used, float* perplexity, int32_t* n_dead,
                      int dead_after, int update_usage, void* ws, cudaStream_t st, const void* const* peer_stats,
                      void* const* peer_flags, int world, int rank, int epoch, const void* mc_stats, void* mc_reduced,
                      const void* reduced_local);
int launch_backward(const void* z, const void* q, const void* gq, int64_t N, int D, int64_t ldz,
                    const float* grad_loss, float coef, void* gz, int dtype, cudaStream_t st);
int launch_codebook_grad(const float* resid, int K, int D, const float* grad_loss, float coef, float* gE,
                         cudaStream_t st);

}  // namespace vq

using namespace vq;

extern "C" {

int vq_abi_version(void) { return VQ_ABI_VERSION; }
const char* vq_last_error(void) { return g_err; }
int64_t vq_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
int vq_set_pdl_mask(int mask) {
    const int prev = pdl_mask();
    g_pdl_mask.store(mask & 31, std::memory_order_relaxed);
    return prev;
}

int64_t vq_set_peer_timeout_ms(int64_t ms) {
    const long long prev = g_peer_timeout_ns.exchange(ms <= 0 ? 0 : (long long)ms * 1000000ll, std::memory_order_relaxed);
    return prev / 1000000ll;
}

// K_pad fp32 norms followed by the K_pad x 32 B "aug" image (vq_ema.cu: write_aug)
int vq_norm_elems(int K) { return K <= 0 ? 0 : round_up(K, 256) * 9; }

static bool dtype_ok(int dt) { return dt == VQ_DT_BF16 || dt == VQ_DT_F32 || dt == VQ_DT_F16; }
static bool search_ok(int s) { return s == VQ_SEARCH_BF16 || s == VQ_SEARCH_F32 || s == VQ_SEARCH_F16; }

int vq_workspace_bytes(int64_t N, int K, int D, int in_dtype, int search, size_t* out_bytes) {
    VQ_REQUIRE(out_bytes != nullptr, "vq_workspace_bytes: out_bytes is NULL");
    VQ_REQUIRE(N >= 0 && K > 0 && D > 0 && dtype_ok(in_dtype) && search_ok(search),
               "vq_workspace_bytes: bad arguments N=%lld K=%d D=%d in_dtype=%d search=%d", (long long)N, K, D, in_dtype,
               search);
    *out_bytes = ws_layout(N, K, D, in_dtype, search).total;
    return VQ_OK;
}

// 16-bit elements the search shadow must hold: K * D rows of bf16 (VQ_SEARCH_BF16) or fp16 (VQ_SEARCH_F16); for
// VQ_SEARCH_F32 the hi / lo bf16 planes of the split tensor-core tier, [2][K_pad][D], K_pad = K rounded up to 256
int64_t vq_shadow_elems(int K, int D, int search) {
    if (K <= 0 || D <= 0) return 0;
    return search == VQ_SEARCH_F32 ? (int64_t)2 * round_up(K, 256) * D : (int64_t)K * D;
}

int vq_prepare_codebook(const float* E_f32, int K, int D, int search, void* E_bf16, float* e_norm2,
                        vq_stream_t stream) {
    VQ_REQUIRE(search_ok(search), "vq_prepare_codebook: bad search mode %d", search);
    VQ_REQUIRE(search == VQ_SEARCH_F32 || E_bf16 != nullptr, "vq_prepare_codebook: the 16-bit shadow is required for "
               "VQ_SEARCH_BF16 / VQ_SEARCH_F16");
    return launch_ema_update(const_cast<float*>(E_f32), nullptr, nullptr, nullptr, nullptr, K, D, 0.f, 0.f, 0, search,
                             E_bf16, e_norm2, nullptr, nullptr, nullptr, nullptr, 0, 0, nullptr, stream, nullptr, nullptr,
                             0, 0, 0, nullptr, nullptr, nullptr);
}

int vq_assign(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, const float* E_f32, const void* E_bf16,
              const float* e_norm2, int K, int search, int32_t* idx, float* smin, void* ws, vq_stream_t stream) {
    VQ_REQUIRE(dtype_ok(in_dtype), "vq_assign: bad in_dtype %d", in_dtype);
    VQ_REQUIRE(N >= 0 && D > 0 && K > 0 && ldz >= D, "vq_assign: bad shape N=%lld D=%d K=%d ldz=%lld", (long long)N, D,
               K, (long long)ldz);
    if (N == 0) return VQ_OK;
    VQ_REQUIRE(idx != nullptr && e_norm2 != nullptr, "vq_assign: idx and e_norm2 are required");
    VQ_REQUIRE(z != nullptr, "vq_assign: z is NULL");
    Presort ps{nullptr, nullptr, nullptr, nullptr};
    const bool ws_clean = (search & VQ_SEARCH_WS_CLEAN) != 0;
    const bool want_presort = (search & VQ_SEARCH_PRESORT) != 0;
    const bool force_simt = (search & VQ_SEARCH_SIMT) != 0;
    search &= ~(VQ_SEARCH_WS_CLEAN | VQ_SEARCH_PRESORT | VQ_SEARCH_SIMT);
    VQ_REQUIRE(search_ok(search), "vq_assign: bad search mode %d", search);
    if (want_presort) {
        VQ_REQUIRE(ws != nullptr, "vq_assign: VQ_SEARCH_PRESORT needs the workspace");
        VQ_REQUIRE(N < (int64_t(1) << 31), "vq_assign: VQ_SEARCH_PRESORT needs N < 2^31");
        const WsLayout L = ws_layout(N, K, D, in_dtype, search);
        uint8_t* w = static_cast<uint8_t*>(ws);
        ps.hist = reinterpret_cast<int*>(w + L.cnt);
        ps.rank = reinterpret_cast<int*>(w + L.rank);
        ps.start = reinterpret_cast<int*>(w + L.start);
        ps.done = ps.hist + 2 * K + 1;               // next to the pass kernel's own completion counter
        // [histogram K | cursor K | pass done | search done]: one memset, unless the caller vouches for the region
        // (the search kernel's last block re-zeroes the histogram, the completion counters reset themselves)
        if (!ws_clean) VQ_CHECK_CUDA(cudaMemsetAsync(ps.hist, 0, (size_t)(2 * K + 2) * 4, stream));
    }
    if (search == VQ_SEARCH_F32) {
        // the 1e-5 tier.  With the split shadow ([2][K_pad][D] hi / lo bf16 planes from vq_prepare_codebook /
        // vq_ema_update) and a shape the tensor-core kernel takes: three bf16 MMAs per product on tcgen05; otherwise
        // (odd D, D > 256, no shadow, VQ_SEARCH_SIMT) the fp32 FMA kernel, which has no shape limits
        if (!force_simt && E_bf16 != nullptr && ws != nullptr && assign_tc_split_supported(N, D, K)) {
            void* stage = static_cast<uint8_t*>(ws) + ws_layout(N, K, D, in_dtype, search).zbf16;
            int rc = launch_split_bf16(z, in_dtype, N, D, ldz, stage, stream);
            if (rc != VQ_OK) return rc;
            return launch_assign_tc(stage, N, D, D, E_bf16, e_norm2, K, VQ_SEARCH_F32, idx, smin, ps, stream);
        }
        VQ_REQUIRE(E_f32 != nullptr, "vq_assign: E_f32 required for VQ_SEARCH_F32");
        return launch_assign_simt(z, in_dtype, N, D, ldz, E_f32, e_norm2, K, idx, smin, ps, stream);
    }
    VQ_REQUIRE(E_bf16 != nullptr, "vq_assign: the 16-bit codebook shadow is required for VQ_SEARCH_BF16 / VQ_SEARCH_F16");
    const void* zb = z;
    int64_t ldb = ldz;
    if (search == VQ_SEARCH_F16) {
        if (in_dtype != VQ_DT_F16) {
            set_error("vq_assign(VQ_SEARCH_F16) takes fp16 latents; use VQ_SEARCH_BF16 or VQ_SEARCH_F32 for in_dtype %d", in_dtype);
            return VQ_ERR_UNSUPPORTED;
        }
    } else if (in_dtype != VQ_DT_BF16) {             // fp32 / fp16 latents under the bf16 search: staged as bf16
        VQ_REQUIRE(ws != nullptr, "vq_assign: workspace required to stage the latents as bf16");
        void* stage = static_cast<uint8_t*>(ws) + ws_layout(N, K, D, in_dtype, search).zbf16;
        int rc = launch_cast_bf16(z, in_dtype, N, D, ldz, stage, stream);
        if (rc != VQ_OK) return rc;
        zb = stage;
        ldb = D;
    }
    if (!assign_tc_supported(N, D, ldb, K)) {
        set_error("vq_assign(tensor core): shape N=%lld D=%d ld=%lld K=%d not supported by the tensor-core kernel "
                  "(needs D %% 8 == 0, 8 <= D <= 512, ld %% 8 == 0); use VQ_SEARCH_F32",
                  (long long)N, D, (long long)ldb, K);
        return VQ_ERR_UNSUPPORTED;
    }
    return launch_assign_tc(zb, N, D, ldb, E_bf16, e_norm2, K, search, idx, smin, ps, stream);
}

int vq_gather_st_loss_scatter(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, int32_t* idx,
                              const float* E_f32, int K, void* q_out, int out_dtype, int64_t ldq, float* sq_err_sum,
                              float* code_cnt, float* code_sum, int stat_kind, int deterministic,
                              int64_t* idx64_out, float* loss_out, float loss_scale, void* ws, vq_stream_t stream) {
    VQ_REQUIRE(dtype_ok(in_dtype), "vq_gather_st_loss_scatter: bad in_dtype %d", in_dtype);
    VQ_REQUIRE(out_dtype == in_dtype, "vq_gather_st_loss_scatter: out_dtype must equal in_dtype");
    {
        const int base_kind = stat_kind & ~(VQ_STAT_ACCUMULATE | VQ_STAT_PRESORTED);
        VQ_REQUIRE(base_kind == VQ_STAT_SUM_Z || base_kind == VQ_STAT_SUM_RESID, "bad stat_kind %d", stat_kind);
    }
    VQ_REQUIRE(E_f32 != nullptr && sq_err_sum != nullptr && (N == 0 || (idx != nullptr && z != nullptr)),
               "vq_gather_st_loss_scatter: NULL argument");
    if (q_out == nullptr) ldq = ldq < D ? D : ldq;     // q_out NULL: loss and statistics only (vq_gather_rows_cf writes q)
    const PassExtra extra{idx64_out, loss_out, loss_scale, 0};
    return launch_gather_pass(z, in_dtype, N, D, ldz, idx, E_f32, K, q_out, ldq, sq_err_sum, code_cnt, code_sum,
                              stat_kind, deterministic, extra, ws, stream);
}

int vq_ema_update(float* E_f32, float* ema_cnt, float* ema_sum, const float* code_cnt, const float* code_sum, int K,
                  int D, float decay, float eps, int do_ema, int search, void* E_bf16, float* e_norm2,
                  int64_t* usage_total, int32_t* steps_since_used, float* perplexity, int32_t* n_dead, int dead_after,
                  int update_usage, void* ws, vq_stream_t stream) {
    VQ_REQUIRE(code_cnt != nullptr, "vq_ema_update: code_cnt is NULL");
    VQ_REQUIRE(search_ok(search), "vq_ema_update: bad search mode %d", search);
    return launch_ema_update(E_f32, ema_cnt, ema_sum, code_cnt, code_sum, K, D, decay, eps, do_ema, search, E_bf16,
                             e_norm2, usage_total, steps_since_used, perplexity, n_dead, dead_after, update_usage, ws,
                             stream, nullptr, nullptr, 0, 0, 0, nullptr, nullptr, nullptr);
}

int vq_ema_update_dp(float* E_f32, float* ema_cnt, float* ema_sum, int K, int D, float decay, float eps, int do_ema,
                     int search, void* E_bf16, float* e_norm2, int64_t* usage_total, int32_t* steps_since_used,
                     float* perplexity, int32_t* n_dead, int dead_after, int update_usage, void* ws,
                     const void* const* peer_stats, void* const* peer_flags, int world, int rank, int epoch,
                     const void* mc_stats, void* mc_reduced, const void* reduced_local, vq_stream_t stream) {
    VQ_REQUIRE(peer_stats != nullptr && peer_flags != nullptr, "vq_ema_update_dp: peer arrays are NULL");
    VQ_REQUIRE((mc_reduced == nullptr) == (reduced_local == nullptr) && (mc_reduced == nullptr || mc_stats != nullptr),
               "vq_ema_update_dp: mc_reduced and reduced_local go together and need mc_stats");
    VQ_REQUIRE(search_ok(search), "vq_ema_update_dp: bad search mode %d", search);
    return launch_ema_update(E_f32, ema_cnt, ema_sum, nullptr, nullptr, K, D, decay, eps, do_ema, search, E_bf16, e_norm2,
                             usage_total, steps_since_used, perplexity, n_dead, dead_after, update_usage, ws, stream,
                             peer_stats, peer_flags, world, rank, epoch, mc_stats, mc_reduced, reduced_local);
}

int vq_knn_workspace_bytes(int64_t N, int C, size_t* out_bytes) {
    VQ_REQUIRE(out_bytes != nullptr && N >= 0 && C > 0, "vq_knn_workspace_bytes: bad arguments N=%lld C=%d", (long long)N, C);
    *out_bytes = knn_workspace_bytes(N, C);
    return VQ_OK;
}
int vq_knn_topk(const float* X, int64_t N, int C, const float* coords, const int32_t* patch, float pos_min_spatial, int k,
                int64_t* knn_idx, float* knn_d2, void* ws, vq_stream_t stream) {
    return launch_knn_topk(X, N, C, coords, patch, pos_min_spatial, k, knn_idx, knn_d2, ws, stream);
}
int vq_knn_mutual_pairs(const int64_t* knn_idx, int64_t N, int k, int64_t* pairs, int64_t* n_pairs, void* ws,
                        vq_stream_t stream) {
    return launch_knn_mutual(knn_idx, N, k, pairs, n_pairs, ws, stream);
}

// ---- CUDA IPC plumbing for the peer exchange (one allocation per rank: [stats buffer 0 | stats buffer 1 | flags])
int vq_ipc_alloc(size_t bytes, void** dev_ptr, void* handle_out) {
    VQ_REQUIRE(dev_ptr != nullptr && handle_out != nullptr && bytes > 0, "vq_ipc_alloc: bad arguments");
    void* p = nullptr;
    VQ_CHECK_CUDA(cudaMalloc(&p, bytes));
    VQ_CHECK_CUDA(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        set_error("cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
        return VQ_ERR_CUDA;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == VQ_IPC_HANDLE_BYTES, "IPC handle size");
    memcpy(handle_out, &h, sizeof(h));
    *dev_ptr = p;
    return VQ_OK;
}
int vq_ipc_free(void* dev_ptr) {
    if (dev_ptr != nullptr) VQ_CHECK_CUDA(cudaFree(dev_ptr));
    return VQ_OK;
}
int vq_ipc_open(const void* handle, void** dev_ptr) {
    VQ_REQUIRE(handle != nullptr && dev_ptr != nullptr, "vq_ipc_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    VQ_CHECK_CUDA(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return VQ_OK;
}
int vq_ipc_close(void* dev_ptr) {
    if (dev_ptr != nullptr) VQ_CHECK_CUDA(cudaIpcCloseMemHandle(dev_ptr));
    return VQ_OK;
}

int vq_backward(const void* z, const void* q, const void* grad_q, int64_t N, int D, int64_t ldz,
                const float* grad_loss, float coef, void* grad_z, int dtype, vq_stream_t stream) {
    VQ_REQUIRE(dtype_ok(dtype), "vq_backward: bad dtype %d", dtype);
    VQ_REQUIRE(N == 0 || (z != nullptr && q != nullptr && grad_z != nullptr && grad_loss != nullptr),
               "vq_backward: NULL argument");
    return launch_backward(z, q, grad_q, N, D, ldz, grad_loss, coef, grad_z, dtype, stream);
}

int vq_codebook_grad(const float* resid_sum, int K, int D, const float* grad_loss, float coef, float* grad_E,
                     vq_stream_t stream) {
    VQ_REQUIRE(resid_sum != nullptr && grad_loss != nullptr && grad_E != nullptr, "vq_codebook_grad: NULL argument");
    return launch_codebook_grad(resid_sum, K, D, grad_loss, coef, grad_E, stream);
}

int vq_cast_bf16(const float* src, int64_t N, int D, int64_t ld, void* dst, vq_stream_t stream) {
    VQ_REQUIRE(N == 0 || (src != nullptr && dst != nullptr), "vq_cast_bf16: NULL argument");
    return launch_cast_bf16(src, VQ_DT_F32, N, D, ld, dst, stream);
}

int vq_gather_rows_cf(const int32_t* idx, const float* E_f32, int K, int D, int64_t B, int64_t S, void* q_cf, int dtype,
                      vq_stream_t stream) {
    VQ_REQUIRE(dtype_ok(dtype), "vq_gather_rows_cf: bad dtype %d", dtype);
    VQ_REQUIRE(B >= 0 && S >= 0 && D > 0 && K > 0, "vq_gather_rows_cf: bad shape B=%lld S=%lld D=%d K=%d", (long long)B,
               (long long)S, D, K);
    VQ_REQUIRE(B == 0 || S == 0 || (idx != nullptr && E_f32 != nullptr && q_cf != nullptr), "vq_gather_rows_cf: NULL argument");
    return launch_gather_rows_cf(idx, E_f32, K, D, B, S, q_cf, dtype, stream);
}

int vq_layout_transpose(const void* src, void* dst, int64_t B, int64_t R, int64_t C, int dtype, vq_stream_t stream) {
    VQ_REQUIRE(dtype_ok(dtype), "vq_layout_transpose: bad dtype %d", dtype);
    VQ_REQUIRE((B == 0 || R == 0 || C == 0) || (src != nullptr && dst != nullptr), "vq_layout_transpose: NULL argument");
    return launch_layout_transpose(src, dst, B, R, C, dtype, stream);
}

}  // extern "C"
