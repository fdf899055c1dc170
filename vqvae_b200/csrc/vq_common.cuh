// Shared device/host helpers for the vq_b200 kernels (sm_100a only).
// Inline-PTX wrappers for mbarrier, TMA (cp.async.bulk.tensor) and tcgen05 (UMMA + TMEM).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/vq_b200.h"

// ----------------------------------------------------------------------------- host side
namespace vq {

void set_error(const char* fmt, ...);          // thread-local message behind vq_last_error()
int  sm_count();                               // cached cudaDevAttrMultiProcessorCount of the current device
int  current_device();                         // cudaGetDevice, -1 on failure (function attributes are per device)

#define VQ_CHECK_CUDA(expr)                                                              \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            vq::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),        \
                          __FILE__, __LINE__);                                           \
            return VQ_ERR_CUDA;                                                          \
        }                                                                                \
    } while (0)

#define VQ_REQUIRE(cond, ...)                                                            \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            vq::set_error(__VA_ARGS__);                                                  \
            return VQ_ERR_ARG;                                                           \
        }                                                                                \
    } while (0)

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int     round_up(int a, int b) { return (a + b - 1) / b * b; }

// kernels implemented in the other translation units
// VQ_SEARCH_PRESORT: the search kernels also leave, in the workspace, what the gather pass needs to visit rows in
// code-sorted order -- per-code histogram, every row's rank inside its code, and (written by the block that finishes
// last) the exclusive scan of the histogram.  All NULL = off.  hist / done must be zero when the kernel starts.
struct Presort {
    int* hist;    // [K]   rows per code
    int* rank;    // [N]   rank of row n among the rows of its code (arrival order, not row order)
    int* start;   // [K+1] exclusive scan of hist
    int* done;    // [1]   blocks finished
};
int launch_assign_tc(const void* z16, int64_t N, int D, int64_t ldz, const void* E16, const float* e_norm2,
                     int K, int kind, int32_t* idx, float* smin, Presort ps, cudaStream_t st);
bool assign_tc_split_supported(int64_t N, int D, int K);
int launch_split_bf16(const void* src, int in_dtype, int64_t N, int D, int64_t ld, void* dst_planes, cudaStream_t st);
int launch_assign_simt(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, const float* E_f32,
                       const float* e_norm2, int K, int32_t* idx, float* smin, Presort ps, cudaStream_t st);
int launch_cast_bf16(const void* src, int in_dtype, int64_t N, int D, int64_t ld, void* dst, cudaStream_t st);
int launch_layout_transpose(const void* src, void* dst, int64_t B, int64_t R, int64_t C, int dtype, cudaStream_t st);
int launch_gather_rows_cf(const int32_t* idx, const float* E_f32, int K, int D, int64_t B, int64_t S, void* q_cf, int dtype,
                          cudaStream_t st);
size_t knn_workspace_bytes(int64_t N, int C);
int launch_knn_topk(const float* X, int64_t N, int C, const float* coords, const int32_t* patch, float pos_min_spatial,
                    int k, int64_t* knn_idx, float* knn_d2, void* ws, cudaStream_t st);
int launch_knn_mutual(const int64_t* knn_idx, int64_t N, int k, int64_t* pairs, int64_t* n_pairs, void* ws,
                      cudaStream_t st);
bool assign_tc_supported(int64_t N, int D, int64_t ldz, int K);
void count_launch(int n = 1);
long long peer_timeout_ns();                   // bound of the cross-rank waits (vq_set_peer_timeout_ms), <= 0 = none

// Byte offsets of the regions inside the caller-provided workspace.
struct WsLayout {
    size_t scal, zbf16, cnt, start, cursor, perm, rank, partial, chunk_cnt, dpart, total;
    int rows_per_chunk, nchunks;
};
WsLayout ws_layout(int64_t N, int K, int D, int in_dtype, int search);     // search < 0: the VQ_SEARCH_BF16 sizing
// optional by-products of the gather pass: the indices widened to int64 (the torch-facing dtype) and the scaled loss
struct PassExtra {
    int64_t* idx64;      // [N] or NULL
    float* loss_out;     // [1] or NULL: sq_err_sum (after this call) * loss_scale
    float loss_scale;
    int idx64_pre;       // set by the launcher: idx64 was already written in natural row order (coalesced) by the scatter
                         // kernel; the pass only rewrites rows it canonicalises (non-finite latents)
};
int launch_gather_pass(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, int32_t* idx, const float* E_f32,
                       int K, void* q_out, int64_t ldq, float* sq_err_sum, float* code_cnt, float* code_sum,
                       int stat_kind, int deterministic, PassExtra extra, void* ws, cudaStream_t st);

}  // namespace vq

// ----------------------------------------------------------------------------- device side
#ifdef __CUDACC__
namespace vq {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .pred P;\n\t.reg .b32 R;\n\t"
        "elect.sync R|P, 0xffffffff;\n\t"
        "selp.b32 %0, 1, 0, P;\n\t}\n"
        : "=r"(pred));
    return pred != 0;
}

// ---- mbarrier ---------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(
                     smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
#ifdef VQ_MBAR_POLL          // experiment: non-suspending poll instead of the (possibly time-sliced) try_wait
        "mbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
#else
        "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
#endif
        "selp.b32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

// Bounded wait: a protocol bug must not hang a shared GPU box.  After ~4 s of spinning the
// kernel records where it was stuck and traps (the launch then fails with a CUDA error).
#ifndef VQ_HANG_CYCLES
#define VQ_HANG_CYCLES 8000000000ll
#endif
static __device__ __noinline__ void mbar_hang(int tag, uint32_t parity) {
    printf("vq_b200: mbarrier wait timed out (tag %d, block %d, thread %d, parity %u)\n", tag, (int)blockIdx.x,
           (int)threadIdx.x, parity);
    __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int tag = 0) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > VQ_HANG_CYCLES) mbar_hang(tag, parity);
    }
}

// ---- TMA --------------------------------------------------------------------------------
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load: coordinates are (c0 = element along the contiguous dim, c1 = row).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                                 uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, "
        "%4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
// 1-D bulk copy global -> shared, completion on an mbarrier (bytes and addresses multiples of 16)
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// L2 eviction policies (createpolicy encodings used by CUTLASS' TMA::CacheHintSm90)
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast  = 0x14F0000000000000ull;

// ---- tcgen05 / TMEM ---------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {   // whole warp, .sync.aligned
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
                 "n"(kCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {        // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]^T ; kind::f16 covers bf16 inputs with fp32 accumulate.
__device__ __forceinline__ void umma_bf16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// ---- CTA pair (cluster of 2, tcgen05 cta_group::2) --------------------------------------------------------------
// One tcgen05.mma.cta_group::2 drives the tensor cores of both SMs of a pair: CTA r holds rows [128 r, 128 r + 128) of A
// and columns [N/2 r, N/2 r + N/2) of B in its own shared memory (same offsets in both CTAs) and receives rows
// [128 r, +128) x N of D in its own TMEM.  Only the leader (rank 0) issues MMAs and waits on the "full" barriers; both CTAs
// issue TMA loads that complete on the LEADER's barrier, and commits are multicast to the same barrier offset in both.
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {         // every thread of both CTAs
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same variable in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar_addr) {
#ifdef VQ_ARRIVE_RELEASE_CLUSTER
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar_addr) : "memory");
#else       // the form CUTLASS uses for the accumulator hand-back of a CTA pair (umma_arrive_2x1SM_sm0)
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar_addr) : "memory");
#endif
}
// destination = this CTA's shared memory, completion bytes -> the barrier at cluster address bar_addr (the leader's)
__device__ __forceinline__ void tma_load_2d_hint_pair(void* smem_dst, const CUtensorMap* m, uint32_t bar_addr, int c0, int c1,
                                                      uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], "
        "[%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_addr), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_result) {   // the same warp of BOTH CTAs
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
                 "n"(kCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void umma_bf16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                  uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs once the issued MMAs have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    const uint16_t mask = 3;
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"(mask)
                 : "memory");
}

// Instruction descriptor for kind::f16, A/B = bf16 K-major, D = fp32 (cute::UMMA::InstrDescriptor layout):
//   [4,6) c_format=1(F32)  [7,10) a_format=1(BF16)  [10,13) b_format=1(BF16)  [15] a_major=0(K)  [16] b_major=0(K)
//   [17,23) N>>3   [24,29) M>>4
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) |
           (static_cast<uint32_t>(M >> 4) << 24);
}
// Shared-memory matrix descriptor, K-major operand whose rows are 128 B (64 bf16) with the 128-byte swizzle
// (what TMA SWIZZLE_128B writes): 8-row groups are 1024 B apart (SBO), LBO unused, version 1 (Blackwell).
//   [0,14) addr>>4   [16,30) LBO>>4   [32,46) SBO>>4   [46,48) version=1   [61,64) layout=2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_sdesc_sw128(uint32_t smem_addr) {
    return static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu) | (static_cast<uint64_t>(1024 >> 4) << 32) |
           (1ull << 46) | (2ull << 61);
}

// Shared-memory matrix descriptor, K-major operand WITHOUT swizzle (canonical "interleave" layout, cute
// ((8,m),(T,2)):((1T,SBO),(1,LBO))): a core matrix is 8 rows x 16 bytes stored contiguously (128 B); lbo = byte distance
// between the two 16-byte K halves of a UMMA K step, sbo = byte distance between consecutive groups of 8 rows.
__device__ __forceinline__ uint64_t make_sdesc_none(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    return static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu) | (static_cast<uint64_t>((lbo >> 4) & 0x3FFFu) << 16) |
           (static_cast<uint64_t>((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}

// TMEM -> registers: lane i of the warp reads TMEM lane (base_lane + i), 32 consecutive fp32 columns.
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- programmatic dependent launch ------------------------------------------------------
// The kernels of one step form a chain on one stream.  Launched with the programmatic-stream-serialization attribute
// (launch_pdl below), a kernel's grid may be set up while its predecessor is still draining; pdl_wait() -- the first
// statement of every such kernel -- blocks until the predecessor has completed and its writes are visible, so only the
// launch latency overlaps, never a data access.  pdl_trigger() lets the successor's launch begin.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- misc -------------------------------------------------------------------------------
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int nthreads) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int4 ld_stream_16(const void* p) {     // read-once data: do not pollute L1
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_16(void* p, const int4& v) {
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// ---- VQ_SEARCH_PRESORT helpers (see struct Presort) -----------------------------------------
// Warp-aggregated claim of one histogram slot per valid lane.  The atomic's result (`base`, meaningful on the
// group's leader lane only) is NOT consumed here, so a caller can let it land while it does other work and finish
// with presort_rank() later: rank = shfl(base, leader) + off.
__device__ __forceinline__ void presort_claim(int* hist, int code, bool valid, int lane, int& base, int& leader,
                                              int& off) {
    const unsigned m = __match_any_sync(0xffffffffu, valid ? code : -1 - lane);
    leader = __ffs(m) - 1;
    off = __popc(m & ((1u << lane) - 1u));
    base = 0;
    if (valid && lane == leader) base = atomicAdd(hist + code, __popc(m));
}
__device__ __forceinline__ int presort_rank(int base, int leader, int off) {
    return __shfl_sync(0xffffffffu, base, leader) + off;
}
// Called by EVERY thread of every block once the block's rows are claimed: the block that arrives last turns the
// histogram into its exclusive scan start[0..K] and leaves the histogram ZERO again (so a caller that keeps the
// workspace to itself never needs the memset: VQ_SEARCH_WS_CLEAN).  sh: >= 34 ints of shared memory.
// K % 4 == 0: every thread owns 24 consecutive counters per round (six int4 loads issued together: one L2 round trip
// and one block scan for K <= 24 * blockDim); otherwise a scalar walk over per-thread chunks.
__device__ __forceinline__ void presort_finish(const Presort& P, int K, int* sh) {
    // Every claim of this block has completed (its result was consumed for the rank store) and atomics are performed
    // at L2; the barrier orders them before thread 0's acq_rel fence + counter increment (cumulative), so one fence per
    // block is enough -- an all-thread __threadfence() (MEMBAR.SC) here costs microseconds per block.
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        const int last = atomicAdd(P.done, 1) == (int)gridDim.x - 1;
        if (last) asm volatile("fence.acq_rel.gpu;" ::: "memory");
        sh[33] = last;
    }
    __syncthreads();
    if (sh[33] == 0) return;
    const int nt = blockDim.x, t = threadIdx.x, lane = t & 31, w = t >> 5, nw = (nt + 31) >> 5;
    // exclusive prefix over the block of one value per thread (thread order); returns the prefix, *total = block sum
    auto block_excl = [&](int v, int* total) {
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        __syncthreads();                         // sh[] of the previous round has been read
        if (lane == 31) sh[w] = inc;
        __syncthreads();
        if (w == 0) {
            const int x = lane < nw ? sh[lane] : 0;
            int ix = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, ix, o);
                if (lane >= o) ix += u;
            }
            sh[lane] = ix - x;
            if (lane == 31) sh[32] = ix;
        }
        __syncthreads();
        *total = sh[32];
        return sh[w] + inc - v;
    };
    int carry = 0;
    if ((K & 3) == 0) {
        constexpr int T = 6;                     // tiles whose loads are in flight together
        const int K4 = K >> 2;
        const int4* h4 = reinterpret_cast<const int4*>(P.hist);
        int4* s4 = reinterpret_cast<int4*>(P.start);
        int4* z4 = reinterpret_cast<int4*>(P.hist);
        for (int base = 0; base < K4; base += T * nt) {      // thread t owns T consecutive int4 (96 bytes) per round
            const int i0 = base + t * T;
            int4 v[T];
            int s = 0;
#pragma unroll
            for (int j = 0; j < T; ++j) v[j] = i0 + j < K4 ? __ldcg(h4 + i0 + j) : make_int4(0, 0, 0, 0);
#pragma unroll
            for (int j = 0; j < T; ++j) s += v[j].x + v[j].y + v[j].z + v[j].w;
            int tot;
            int pre = carry + block_excl(s, &tot);
#pragma unroll
            for (int j = 0; j < T; ++j) {
                if (i0 + j < K4) {
                    s4[i0 + j] = make_int4(pre, pre + v[j].x, pre + v[j].x + v[j].y, pre + v[j].x + v[j].y + v[j].z);
                    z4[i0 + j] = make_int4(0, 0, 0, 0);
                }
                pre += v[j].x + v[j].y + v[j].z + v[j].w;
            }
            carry += tot;
        }
    } else {
        const int per = (K + nt - 1) / nt;
        const int k0 = min(K, t * per), k1 = min(K, k0 + per);
        int s = 0;
        for (int k = k0; k < k1; ++k) s += __ldcg(P.hist + k);
        int tot;
        int run = block_excl(s, &tot);
        for (int k = k0; k < k1; ++k) {
            P.start[k] = run;
            run += __ldcg(P.hist + k);
            P.hist[k] = 0;
        }
        carry = tot;
    }
    if (t == 0) {
        P.start[K] = carry;
        *P.done = 0;
    }
}
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}


// host side of the above.  The mask (vq_set_pdl_mask; VQ_PDL in the environment of the tuning build) is a bit mask over the kernels of the step (tuning knob):
// 1 search, 2 scatter, 4 pass, 8 EMA totals, 16 EMA apply; 0 = plain stream order everywhere.
enum { PDL_SEARCH = 1, PDL_SCATTER = 2, PDL_PASS = 4, PDL_TOTALS = 8, PDL_APPLY = 16, PDL_DEFAULT = 31 };
int pdl_mask();                                  // vq_abi.cu: VQ_PDL from the environment, or the last vq_set_pdl_mask()
inline bool pdl_enabled(int which) { return (pdl_mask() & which) != 0; }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int which, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled(which) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
}  // namespace vq
#endif  // __CUDACC__
