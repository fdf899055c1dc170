// vq_gather_st_loss_scatter_noalloc: the fused gather / straight-through / loss / code-statistics pass (a-2 .. a-5) for
// callers that cannot hand the library a workspace they allocated -- code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  No counterpart in the reference.
//
// The scratch the pass needs (counting sort of idx: histogram, cursor, scan, permutation; loss partials) lives in ONE
// file-scope __device__ array of VQ_NOALLOC_SCRATCH_BYTES (4 MiB) that every GPU loading the library holds until the
// process exits.  In host code the array's NAME is a host-side stand-in, not a device address: its device address is
// looked up once per device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels get.
// The entry point calls no cudaMalloc / cudaFree / allocating routine.  All calls share the array and it keeps the last
// call's contents: the launcher clears the counters that must start at zero (one cudaMemsetAsync on the caller's stream),
// and two calls that overlap on the GPU (different streams, no ordering between them) RACE -- order them.
#include <mutex>

#include "vq_common.cuh"

namespace vq {
namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[VQ_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
void* g_scratch_ptr[MAX_DEVICES] = {};

// device address of the scratch array on the CURRENT device (cached per device ordinal)
int scratch_for_current_device(void** out) {
    int dev = -1;
    VQ_CHECK_CUDA(cudaGetDevice(&dev));
    VQ_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "vq_gather_st_loss_scatter_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_mu);
    if (g_scratch_ptr[dev] == nullptr) {
        void* p = nullptr;
        VQ_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        VQ_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0,
                   "vq_gather_st_loss_scatter_noalloc: bad scratch address %p", p);
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return VQ_OK;
}

}  // namespace
}  // namespace vq

using namespace vq;

extern "C" int vq_gather_st_loss_scatter_noalloc(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, int32_t* idx,
                                                 const float* E_f32, int K, void* q_out, int64_t ldq, float* sq_err_sum,
                                                 float* code_cnt, float* code_sum, int stat_kind, int64_t* idx64_out,
                                                 float* loss_out, float loss_scale, vq_stream_t stream) {
    VQ_REQUIRE(in_dtype == VQ_DT_BF16 || in_dtype == VQ_DT_F32 || in_dtype == VQ_DT_F16,
               "vq_gather_st_loss_scatter_noalloc: bad in_dtype %d", in_dtype);
    VQ_REQUIRE(N >= 0 && D > 0 && K > 0, "vq_gather_st_loss_scatter_noalloc: bad shape N=%lld D=%d K=%d", (long long)N, D, K);
    VQ_REQUIRE((stat_kind & VQ_STAT_PRESORTED) == 0,
               "vq_gather_st_loss_scatter_noalloc: VQ_STAT_PRESORTED refers to a caller workspace; this entry point has none");
    {
        const int base_kind = stat_kind & ~VQ_STAT_ACCUMULATE;
        VQ_REQUIRE(base_kind == VQ_STAT_SUM_Z || base_kind == VQ_STAT_SUM_RESID, "bad stat_kind %d", stat_kind);
    }
    VQ_REQUIRE(E_f32 != nullptr && sq_err_sum != nullptr && (N == 0 || (idx != nullptr && z != nullptr && q_out != nullptr)),
               "vq_gather_st_loss_scatter_noalloc: NULL argument");
    // Non-deterministic pass only: it touches the regions in front of the deterministic-mode slots (counters, scan,
    // permutation, ranks, loss partials) -- that prefix of the workspace layout is what must fit the static array.
    const WsLayout L = ws_layout(N, K, D, in_dtype, -1);
    if (L.chunk_cnt > (size_t)VQ_NOALLOC_SCRATCH_BYTES) {
        set_error("vq_gather_st_loss_scatter_noalloc: N=%lld K=%d needs %zu bytes of scratch, the built-in array holds %d "
                  "(about 12 bytes per row + 12 bytes per code); use vq_gather_st_loss_scatter with a caller workspace",
                  (long long)N, K, L.chunk_cnt, (int)VQ_NOALLOC_SCRATCH_BYTES);
        return VQ_ERR_UNSUPPORTED;
    }
    void* scratch = nullptr;
    const int rc = scratch_for_current_device(&scratch);
    if (rc != VQ_OK) return rc;
    const PassExtra extra{idx64_out, loss_out, loss_scale, 0};
    // launch_gather_pass clears the counter region it uses ([histogram | cursor | done], one cudaMemsetAsync on `stream`)
    // before its first kernel, so nothing the previous call left in the array is read
    return launch_gather_pass(z, in_dtype, N, D, ldz, idx, E_f32, K, q_out, ldq, sq_err_sum, code_cnt, code_sum, stat_kind,
                              /*deterministic=*/0, extra, scratch, stream);
}
