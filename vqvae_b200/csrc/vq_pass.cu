// a-2 .. a-5  One pass over z: codebook gather, straight-through output, squared-error reduction and the
// per-code counts / sums for the EMA update (or, in st mode, the residual sums the codebook gradient needs).
//
// HBM-bound: algorithmic traffic is D*s (read z) + 4 (idx) + D*s (write q) bytes per latent vector; the fp32
// master codebook (K*D*4 <= 8 MiB at K=8192, D=256) stays L2-resident.  To keep the scatter-add off the L2
// atomic units, rows are visited in CODE-SORTED order: a counting sort of idx yields `perm` -- with VQ_STAT_PRESORTED
// the histogram, its scan and every row's rank come out of the search kernel and one atomics-free scatter is left
// here; stand-alone calls run histogram -> scan + slot claim themselves -- and every lane group then walks 64
// consecutive sorted positions, keeping the running per-code sum in registers and issuing one vector `red`
// per (code, 64-row slice) -- about K + N/64 reds of D floats per step instead of N.
// Required by: scripts/export_codebook.py:104 (gather from [K, D]); frl/config/frl_model_v0.yaml:35 (straight
// through); frl/config/frl_bindings_v0.yaml:887-891 (commitment + codebook terms); scripts/train_vqvae.py:433-436
// (EMA statistics).  Definitions: SURVEY.md 8(a-2 .. a-5).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "vq_common.cuh"

namespace vq {
namespace {

constexpr int LG = 64;   // sorted positions walked by one lane group

__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ------------------------------------------------------------------------------ counting sort of idx
__global__ void hist_kernel(const int32_t* __restrict__ idx, int64_t N, int K, int* __restrict__ cnt,
                            int* __restrict__ chunk_cnt, int rows_per_chunk) {
    const int lane = threadIdx.x & 31;
    for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (int64_t)gridDim.x * blockDim.x) {
        int c = idx[n];
        if ((unsigned)c >= (unsigned)K) c = 0;
        if (chunk_cnt != nullptr) {                  // deterministic mode: per-chunk histograms as well
            atomicAdd(&chunk_cnt[(n / rows_per_chunk) * K + c], 1);
        }
        const unsigned m = __match_any_sync(__activemask(), c);
        if (lane == __ffs(m) - 1) atomicAdd(&cnt[c], __popc(m));
    }
}

// Shared-memory privatised variants (default path): block b owns rows [b*R, (b+1)*R); per-code counters live in
// shared memory, so the global atomics drop from one per row to one per (block, code present in the block) and
// hot codes no longer serialise on a single L2 address.
__global__ void __launch_bounds__(512) hist_smem_kernel(const int32_t* __restrict__ idx, int64_t N, int K, int R,
                                                        int* __restrict__ cnt) {
    extern __shared__ int sh_cnt[];
    for (int k = threadIdx.x; k < K; k += blockDim.x) sh_cnt[k] = 0;
    __syncthreads();
    const int64_t a = (int64_t)blockIdx.x * R, b = min(N, a + R);
    for (int64_t n = a + threadIdx.x; n < b; n += blockDim.x) {
        int c = idx[n];
        if ((unsigned)c >= (unsigned)K) c = 0;
        atomicAdd(&sh_cnt[c], 1);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const int v = sh_cnt[k];
        if (v != 0) atomicAdd(&cnt[k], v);
    }
}

// Fused variant used by the default path: every block first rebuilds the exclusive scan of the (complete) global
// histogram in its own shared memory -- K counters, a few microseconds, in parallel on all blocks -- which removes
// the single-block scan launch from the critical path; `cursor` starts at zero and holds the claimed count per code.
// The blocks also clear the statistics the pass will accumulate into (code_sum, code_cnt), replacing two memsets.
__global__ void __launch_bounds__(512)
scan_scatter_smem_kernel(const int32_t* __restrict__ idx, int64_t N, int K, int R, const int* __restrict__ cnt,
                         int* __restrict__ cursor, int32_t* __restrict__ perm, float4* __restrict__ zero_a, int64_t na4,
                         float* __restrict__ zero_b, int nb) {
    extern __shared__ int sh[];
    int* sh_cnt = sh;
    int* sh_base = sh + K;
    int* sh_start = sh + 2 * K;
    __shared__ int warp_tot[16];
    __shared__ int carry_sh;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;

    for (int64_t i = (int64_t)blockIdx.x * 512 + t; i < na4; i += (int64_t)gridDim.x * 512)
        zero_a[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = blockIdx.x * 512 + t; i < nb; i += gridDim.x * 512) zero_b[i] = 0.f;

    if (t == 0) carry_sh = 0;
    for (int k = t; k < K; k += 512) sh_cnt[k] = 0;
    __syncthreads();
    for (int base = 0; base < K; base += 2048) {          // exclusive scan, 4 counters per thread per tile
        const int k0 = base + t * 4;
        int v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (k0 + j < K) ? cnt[k0 + j] : 0;
        const int sum4 = v[0] + v[1] + v[2] + v[3];
        int inc = sum4;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) warp_tot[w] = inc;
        __syncthreads();
        if (w == 0) {
            const int x = lane < 16 ? warp_tot[lane] : 0;
            int ix = x;
#pragma unroll
            for (int o = 1; o < 16; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, ix, o);
                if (lane >= o) ix += u;
            }
            if (lane < 16) warp_tot[lane] = ix - x;
        }
        __syncthreads();
        int run = carry_sh + warp_tot[w] + inc - sum4;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (k0 + j < K) sh_start[k0 + j] = run;
            run += v[j];
        }
        __syncthreads();
        if (t == 511) carry_sh = run;
        __syncthreads();
    }

    const int64_t a = (int64_t)blockIdx.x * R, b = min(N, a + R);
    for (int64_t n = a + t; n < b; n += 512) {
        int c = idx[n];
        if ((unsigned)c >= (unsigned)K) c = 0;
        atomicAdd(&sh_cnt[c], 1);
    }
    __syncthreads();
    for (int k = t; k < K; k += 512) {
        const int v = sh_cnt[k];
        if (v != 0) sh_base[k] = sh_start[k] + atomicAdd(&cursor[k], v);   // contiguous range inside code k's segment
        sh_cnt[k] = 0;
    }
    __syncthreads();
    for (int64_t n = a + t; n < b; n += 512) {
        int c = idx[n];
        if ((unsigned)c >= (unsigned)K) c = 0;
        const int r = atomicAdd(&sh_cnt[c], 1);
        perm[sh_base[c] + r] = (int32_t)n;
    }
}

// VQ_STAT_PRESORTED: the search kernel already left the histogram's exclusive scan (`start`) and every row's rank
// inside its code (vq_assign with VQ_SEARCH_PRESORT), so the counting sort is this one atomics-free scatter; like the
// kernel above it clears the statistics buffers on the way.
// kVec: four consecutive rows per thread (idx and rank read as int4), one thread per group -- every load of the
// idx -> start -> perm chain is in flight at once instead of a latency-bound loop (12.9 -> ~5 us at C2).
// idx64 (nullable): the indices widened to int64 for the torch-facing module, written HERE because this kernel walks the
// rows in natural order -- 32 coalesced bytes per thread -- whereas the pass visits them in code-sorted order and paid
// 18 us at C2 (12 % of its time) for 1.3 M scattered 8-byte stores.
template <bool kVec>
__global__ void __launch_bounds__(256)
scatter_ranked_kernel(const int32_t* __restrict__ idx, const int* __restrict__ rank, int64_t N, int K,
                      const int* __restrict__ start, int32_t* __restrict__ perm, float4* __restrict__ zero_a,
                      int64_t na4, float* __restrict__ zero_b, int nb, int64_t* __restrict__ idx64) {
    pdl_wait();
    pdl_trigger();
    const int64_t tid = (int64_t)blockIdx.x * 256 + threadIdx.x, nth = (int64_t)gridDim.x * 256;
    if (kVec) {
        const int64_t n4 = N >> 2;
        for (int64_t g = tid; g < n4; g += nth) {
            const int4 c = __ldcs(reinterpret_cast<const int4*>(idx) + g);
            const int4 r = __ldcs(reinterpret_cast<const int4*>(rank) + g);
            int cc[4] = {c.x, c.y, c.z, c.w};
            const int rr[4] = {r.x, r.y, r.z, r.w};
            int st[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if ((unsigned)cc[j] >= (unsigned)K) cc[j] = 0;
                st[j] = __ldg(start + cc[j]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) perm[st[j] + rr[j]] = (int32_t)(g * 4 + j);
            if (idx64 != nullptr) {           // 16-byte aligned (checked by the launcher): two streaming 16-byte stores
                longlong2* o = reinterpret_cast<longlong2*>(idx64 + g * 4);
                __stcs(o, make_longlong2(cc[0], cc[1]));
                __stcs(o + 1, make_longlong2(cc[2], cc[3]));
            }
        }
        for (int64_t n = (n4 << 2) + tid; n < N; n += nth) {
            int c = idx[n];
            if ((unsigned)c >= (unsigned)K) c = 0;
            perm[__ldg(start + c) + rank[n]] = (int32_t)n;
            if (idx64 != nullptr) idx64[n] = c;
        }
    } else {
        for (int64_t n = tid; n < N; n += nth) {
            int c = idx[n];
            if ((unsigned)c >= (unsigned)K) c = 0;
            perm[__ldg(start + c) + rank[n]] = (int32_t)n;
            if (idx64 != nullptr) idx64[n] = c;
        }
    }
    for (int64_t i = tid; i < na4; i += nth) zero_a[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = tid; i < nb; i += nth) zero_b[i] = 0.f;
}

// one block; exclusive scan of cnt[0..K) -> start[0..K], cursor[k] = start[k].  Tiles of 4096 counters, int4 loads.
__global__ void __launch_bounds__(1024) scan_kernel(const int* __restrict__ cnt, int K, int* __restrict__ start,
                                                    int* __restrict__ cursor) {
    __shared__ int warp_tot[32];
    __shared__ int carry_sh;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    if (t == 0) carry_sh = 0;
    __syncthreads();
    for (int base = 0; base < K; base += 4096) {
        const int k0 = base + t * 4;
        int v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (k0 + j < K) ? cnt[k0 + j] : 0;
        const int s = v[0] + v[1] + v[2] + v[3];
        int inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) warp_tot[w] = inc;
        __syncthreads();
        if (w == 0) {
            const int x = warp_tot[lane];
            int ix = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, ix, o);
                if (lane >= o) ix += u;
            }
            warp_tot[lane] = ix - x;
        }
        __syncthreads();
        int run = carry_sh + warp_tot[w] + inc - s;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (k0 + j < K) {
                start[k0 + j] = run;
                cursor[k0 + j] = run;
            }
            run += v[j];
        }
        __syncthreads();
        if (t == 1023) carry_sh = run;
        __syncthreads();
    }
    if (t == 0) start[K] = carry_sh;
}

__global__ void scatter_kernel(const int32_t* __restrict__ idx, int64_t N, int K, int* __restrict__ cursor,
                               int32_t* __restrict__ perm) {
    const int lane = threadIdx.x & 31;
    for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (int64_t)gridDim.x * blockDim.x) {
        int c = idx[n];
        if ((unsigned)c >= (unsigned)K) c = 0;
        const unsigned m = __match_any_sync(__activemask(), c);
        const int leader = __ffs(m) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&cursor[c], __popc(m));
        base = __shfl_sync(m, base, leader);
        perm[base + __popc(m & lanemask_lt())] = (int32_t)n;
    }
}

// deterministic slot claim.  chunk_cnt[b][k] holds the number of rows of chunk b with code k; turn every column
// into its running offset (start[k] + rows of code k in earlier chunks) ...
__global__ void chunk_offsets_kernel(int* __restrict__ chunk_cnt, int nchunks, int K, const int* __restrict__ start) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    int run = start[k];
    for (int b = 0; b < nchunks; ++b) {
        const int t = chunk_cnt[(int64_t)b * K + k];
        chunk_cnt[(int64_t)b * K + k] = run;
        run += t;
    }
}
// ... then one warp per chunk walks its rows in order, 32 at a time, so that equal codes keep row order.
__global__ void scatter_stable_kernel(const int32_t* __restrict__ idx, int64_t N, int K, int* __restrict__ chunk_off,
                                      int rows_per_chunk, int nchunks, int32_t* __restrict__ perm) {
    const int lane = threadIdx.x & 31;
    const int chunk = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (chunk >= nchunks) return;
    const int64_t a = (int64_t)chunk * rows_per_chunk;
    const int64_t b = min(N, a + rows_per_chunk);
    volatile int* off = chunk_off + (int64_t)chunk * K;
    for (int64_t n0 = a; n0 < b; n0 += 32) {
        const int64_t n = n0 + lane;
        const bool valid = n < b;
        int c = valid ? idx[n] : -1 - lane;           // distinct dummies never match a real code
        if (valid && (unsigned)c >= (unsigned)K) c = 0;
        const unsigned m = __match_any_sync(0xffffffffu, c);
        const int leader = __ffs(m) - 1;
        int base = 0;
        if (valid && lane == leader) {
            base = off[c];
            off[c] = base + __popc(m);
        }
        __syncwarp();
        base = __shfl_sync(0xffffffffu, base, leader);
        if (valid) perm[base + __popc(m & lanemask_lt())] = (int32_t)n;
    }
}

// ------------------------------------------------------------------------------ the fused pass
template <typename T> struct VecIO;
template <> struct VecIO<__nv_bfloat16> {
    static constexpr int EPV = 8;
    __device__ static __forceinline__ void unpack(const int4& v, float (&f)[8]) {
        f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
        f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
    }
    __device__ static __forceinline__ int4 pack(const float (&f)[8]) {
        return make_int4((int)pack_bf16(f[0], f[1]), (int)pack_bf16(f[2], f[3]), (int)pack_bf16(f[4], f[5]),
                         (int)pack_bf16(f[6], f[7]));
    }
    __device__ static __forceinline__ bool nonfinite(const int4& v) {   // any bf16 with an all-ones exponent
        const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
        // exponent field + one carries into the half's top bit exactly when the exponent is all ones (3 ops per word
        // instead of two masked compares per half)
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) acc |= (w[i] & 0x7f807f80u) + 0x00800080u;
        return (acc & 0x80008000u) != 0;
    }
};
template <> struct VecIO<__half> {      // fp16 latents: the reference script's default autocast dtype
    static constexpr int EPV = 8;
    __device__ static __forceinline__ void unpack(const int4& v, float (&f)[8]) {
        const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            f[2 * i] = t.x; f[2 * i + 1] = t.y;
        }
    }
    __device__ static __forceinline__ int4 pack(const float (&f)[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<const uint32_t*>(&h);
        }
        return make_int4((int)w[0], (int)w[1], (int)w[2], (int)w[3]);
    }
    __device__ static __forceinline__ bool nonfinite(const int4& v) {   // any fp16 with an all-ones exponent (5 bits)
        const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) acc |= (w[i] & 0x7c007c00u) + 0x04000400u;
        return (acc & 0x80008000u) != 0;
    }
};
template <> struct VecIO<float> {
    static constexpr int EPV = 4;
    __device__ static __forceinline__ void unpack(const int4& v, float (&f)[4]) {
        f[0] = __int_as_float(v.x); f[1] = __int_as_float(v.y); f[2] = __int_as_float(v.z); f[3] = __int_as_float(v.w);
    }
    __device__ static __forceinline__ int4 pack(const float (&f)[4]) {
        return make_int4(__float_as_int(f[0]), __float_as_int(f[1]), __float_as_int(f[2]), __float_as_int(f[3]));
    }
    __device__ static __forceinline__ bool nonfinite(const int4& v) {
        const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
        bool bad = false;
#pragma unroll
        for (int i = 0; i < 4; ++i) bad |= (w[i] & 0x7f800000u) == 0x7f800000u;
        return bad;
    }
};

// T: storage type of z and q.  LPR lanes cooperate on one row (16 bytes per lane per vector), VPL vectors per lane.
// perm == nullptr visits rows in natural order (no statistics wanted: eval mode).
//
// Latent rows reach the SM through a per-thread cp.async ring in shared memory (STAGES batches of UB rows): every
// thread copies exactly the 16-byte pieces it will later consume, so cp.async.wait_group is the only
// synchronisation, no registers are tied up by loads in flight, and (STAGES - 1) * UB rows per lane group are
// outstanding at any time -- memory-level parallelism independent of occupancy.
#ifndef VQ_PASS_BULK
#define VQ_PASS_BULK 0      // 0: per-lane cp.async ring (default); 1: one cp.async.bulk (TMA unit) per row + mbarriers --
                            // measured slower (C4 pass group 146 -> 165 us, C1 46 -> 63 us): 256-512-byte bulk copies are
                            // too small for the TMA unit, see profiles/r01/tuning_experiments.log
#endif
#ifndef VQ_PASS_BLOCKSUM
#define VQ_PASS_BLOCKSUM 1   // 1: lane groups whose last run is the block's first code leave as ONE red burst per block
#endif
#ifndef VQ_PASS_STAGES
#define VQ_PASS_STAGES 4
#endif
#ifndef VQ_PASS_ROWS
#define VQ_PASS_ROWS 4
#endif
constexpr int PASS_STAGES = VQ_PASS_STAGES;
// rows per batch: VQ_PASS_ROWS 16-byte pieces per thread per batch
template <int VPL> struct PassCfg { static constexpr int U = VQ_PASS_ROWS / VPL > 1 ? VQ_PASS_ROWS / VPL : 1; };

__device__ __forceinline__ void cp_async_16(uint32_t smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }
__device__ __forceinline__ int4 lds_16(uint32_t addr) {
    int4 r;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}

#ifndef VQ_PASS_MINB
#define VQ_PASS_MINB 1
#endif
template <typename T, int VPL, int LPR>
__global__ void __launch_bounds__(256, VQ_PASS_MINB)
pass_kernel(const T* __restrict__ z, int64_t N, int D, int64_t ldz, int32_t* __restrict__ idx,
            const int32_t* __restrict__ perm, const float* __restrict__ E, int K, T* __restrict__ q, int64_t ldq,
            float* __restrict__ loss_partial, float* __restrict__ code_cnt, float* __restrict__ code_sum,
            float* __restrict__ dpart, int stat_kind, int* __restrict__ done, float* __restrict__ sq_err_out,
            const PassExtra X, const int lgr) {
    // lgr <= LG: sorted positions per lane group (64 for large batches; 16 / 32 when N is small, so that a 40 960-row
    // batch still spreads over every SM: 160 blocks instead of 40)
    pdl_wait();
    pdl_trigger();
    using IO = VecIO<T>;
    constexpr int EPV = IO::EPV;
    constexpr int G = 32 / LPR;
    constexpr int U = PassCfg<VPL>::U;
    constexpr int UB = U < LPR ? U : LPR;             // rows per batch
    constexpr int S = PASS_STAGES;
    static_assert(LG % UB == 0 && 16 % UB == 0 && 16 / UB >= S - 1, "batching");
    const int NB = lgr / UB;                          // batches per lane group

    extern __shared__ int4 pass_smem[];
    const int lane = threadIdx.x & 31;
    const int g = lane / LPR, lg = lane % LPR;
#if VQ_PASS_BULK
    // per lane group [S][UB] row slots of ROWB bytes (a latent row lands contiguously: one bulk copy per row), then
    // [8 * G][S] mbarriers, then per warp [G][LG] row numbers and [G][LG] codes
    constexpr int ROWB = VPL * LPR * 16;
    const int grp = (threadIdx.x >> 5) * G + g;
    const uint32_t ring_g = smem_u32(pass_smem) + (uint32_t)grp * (uint32_t)(S * UB * ROWB);
    const uint32_t ring = ring_g + lg * 16u;
    uint64_t* bars = reinterpret_cast<uint64_t*>(pass_smem + S * UB * VPL * 256) + grp * S;
    int* pos_base = reinterpret_cast<int*>(reinterpret_cast<uint64_t*>(pass_smem + S * UB * VPL * 256) + 8 * G * S) +
                    (threadIdx.x >> 5) * (2 * G * LG);
    if (lg == 0) {
#pragma unroll
        for (int i = 0; i < S; ++i) mbar_init(&bars[i], 1);
        fence_mbar_init();
    }
    const uint32_t rowbytes = (uint32_t)D * (uint32_t)sizeof(T);
#else
    // [S][UB][VPL][256] 16-byte ring slots, then per warp [G][LG] row numbers and [G][LG] codes
    const uint32_t ring = smem_u32(pass_smem) + threadIdx.x * 16u;
    int* pos_base = reinterpret_cast<int*>(pass_smem + S * UB * VPL * 256) + (threadIdx.x >> 5) * (2 * G * LG);
#endif
    const unsigned gmask = (LPR == 32 ? 0xffffffffu : ((1u << LPR) - 1u)) << (g * LPR);
    const int64_t warp_global = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t p0 = (warp_global * G + g) * lgr;
    const int V = D / EPV;
    int* prow = pos_base + g * LG;
    int* pcode = pos_base + (G + g) * LG;

#pragma unroll
    for (int i = 0; i < LG / LPR; ++i) {
        const int64_t pos = p0 + i * LPR + lg;
        int r = -1, c = -1;
        if (i * LPR + lg < lgr && pos < N) {
            r = perm != nullptr ? perm[pos] : (int)pos;
            c = idx[r];
            if ((unsigned)c >= (unsigned)K) c = 0;
        }
        prow[i * LPR + lg] = r;
        pcode[i * LPR + lg] = c;
    }
    __syncwarp();
#if VQ_PASS_BLOCKSUM
    // A code with many rows (trained codebooks: a fifth of all rows can share one code) spans whole blocks, and every
    // lane group of such a block would fire its own red burst at the same 4*D bytes of code_sum (C2, Zipf-like usage:
    // pass group 198 us against 163 us on uniform usage).  Lane groups whose LAST run is the code the block starts with
    // hand their sums over through shared memory (the ring, idle by then) and leave as one burst per block.
    __shared__ int blk_code;
    __shared__ int blk_cnts[8 * G];
    const bool blk_on = code_sum != nullptr && dpart == nullptr && perm != nullptr;     // block-uniform
    if (blk_on) {
        if (threadIdx.x == 0) blk_code = pcode[0];    // warp 0, lane group 0, position 0: the block's first sorted row
        __syncthreads();
    }
#endif

#if VQ_PASS_BULK
    // Latent rows reach shared memory as ONE bulk copy (TMA unit, cp.async.bulk) per row, issued by the group's first
    // lane and completing on the stage's mbarrier: no per-lane cp.async instructions, and the bytes in flight are
    // bounded by the ring, not by the L1 lines that LDGSTS requests occupy until their data returns.
    auto issue = [&](int b) {                         // batch b of this lane group -> ring stage b % S
        if (lg == 0) {
            const int stage = b % S;
            int nvalid = 0;
#pragma unroll
            for (int u = 0; u < UB; ++u) nvalid += prow[b * UB + u] >= 0;
            mbar_arrive_expect_tx(&bars[stage], (uint32_t)nvalid * rowbytes);
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const int r = prow[b * UB + u];
                if (r >= 0)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(ring_g + (uint32_t)((stage * UB + u) * ROWB)), "l"(z + (int64_t)r * ldz),
                                   "r"(rowbytes), "r"(smem_u32(&bars[stage])) : "memory");
            }
        }
    };
#pragma unroll
    for (int b = 0; b < S - 1; ++b) issue(b);
#else
    auto issue = [&](int b) {                         // batch b of this lane group -> ring stage b % S
        const int stage = b % S;
#pragma unroll
        for (int u = 0; u < UB; ++u) {
            const int r = prow[b * UB + u];
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int v = lg + j * LPR;
                if (r >= 0 && v < V)
                    cp_async_16(ring + (uint32_t)(((stage * UB + u) * VPL + j) * 256) * 16u, z + (int64_t)r * ldz + v * EPV);
            }
        }
    };
#pragma unroll
    for (int b = 0; b < S - 1; ++b) {
        issue(b);
        cp_async_commit();
    }
#endif

    float e[VPL][EPV], acc[VPL][EPV];
    int4 qv[VPL];
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
        qv[j] = make_int4(0, 0, 0, 0);
#pragma unroll
        for (int x = 0; x < EPV; ++x) { e[j][x] = 0.f; acc[j][x] = 0.f; }
    }
    float loss = 0.f;
    int cnt = 0, cur = -1;

    // Hand the finished run (cur, this 64-position slice) over.  Default: vector reds straight into code_sum.
    // Deterministic: plain stores into slot (slice + code) of dpart -- slice and code are both non-decreasing along
    // the sorted order and one of them changes between consecutive runs, so the slot is unique; a second kernel
    // adds the slots of each code in slice order.
    const int64_t slice = p0 / LG;
    auto flush = [&]() {
        if (cur >= 0 && code_sum != nullptr && (cnt > 0 || dpart != nullptr)) {
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int v = lg + j * LPR;
                if (v < V) {
                    if (dpart != nullptr) {
                        float4* dst = reinterpret_cast<float4*>(dpart + (slice + cur) * D + v * EPV);
#pragma unroll
                        for (int x = 0; x < EPV; x += 4)
                            dst[x / 4] = make_float4(acc[j][x], acc[j][x + 1], acc[j][x + 2], acc[j][x + 3]);
                    } else {
                        float* dst = code_sum + (int64_t)cur * D + v * EPV;
#pragma unroll
                        for (int x = 0; x < EPV; x += 4)
                            red_add_v4(dst + x, acc[j][x], acc[j][x + 1], acc[j][x + 2], acc[j][x + 3]);
                    }
                }
            }
            if (lg == 0 && cnt > 0) atomicAdd(code_cnt + cur, (float)cnt);
        }
#pragma unroll
        for (int j = 0; j < VPL; ++j)
#pragma unroll
            for (int x = 0; x < EPV; ++x) acc[j][x] = 0.f;
        cnt = 0;
    };

#pragma unroll 1
    for (int b = 0; b < NB; ++b) {
        const int stage = b % S;
#if VQ_PASS_BULK
        __syncwarp();                                 // every lane is done reading the stage the next copy overwrites
        if (b + S - 1 < NB) issue(b + S - 1);
        mbar_wait(&bars[stage], (uint32_t)(b / S) & 1u, 20);
#else
        if (b + S - 1 < NB) issue(b + S - 1);
        cp_async_commit();                            // (possibly empty) group keeps the wait distance constant
        cp_async_wait<S - 1>();                       // batch b has landed in this thread's slots
#endif
#pragma unroll
        for (int u = 0; u < UB; ++u) {
            const int r = prow[b * UB + u];
            const int c = pcode[b * UB + u];
            const bool valid = r >= 0;
            int4 zv[VPL];
            bool nf = false;
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int v = lg + j * LPR;
                zv[j] = make_int4(0, 0, 0, 0);
#if VQ_PASS_BULK
                if (valid && v < V) zv[j] = lds_16(ring + (uint32_t)((stage * UB + u) * ROWB + j * LPR * 16));
#else
                if (valid && v < V) zv[j] = lds_16(ring + (uint32_t)(((stage * UB + u) * VPL + j) * 256) * 16u);
#endif
                nf |= IO::nonfinite(zv[j]);
            }
            const bool bad = (__ballot_sync(0xffffffffu, valid && nf) & gmask) != 0;
            if (valid) {
                if (c != cur) {
                    flush();
                    cur = c;
#pragma unroll
                    for (int j = 0; j < VPL; ++j) {
                        const int v = lg + j * LPR;
                        if (v < V) {
                            const float4* src = reinterpret_cast<const float4*>(E + (int64_t)cur * D + v * EPV);
#pragma unroll
                            for (int x = 0; x < EPV; x += 4) {
                                const float4 t = __ldg(src + x / 4);
                                e[j][x] = t.x; e[j][x + 1] = t.y; e[j][x + 2] = t.z; e[j][x + 3] = t.w;
                            }
                            qv[j] = IO::pack(e[j]);
                        }
                    }
                }
                if (!bad) {
#pragma unroll
                    for (int j = 0; j < VPL; ++j) {
                        const int v = lg + j * LPR;
                        if (v < V) {
                            float zf[EPV];
                            IO::unpack(zv[j], zf);
#pragma unroll
                            for (int x = 0; x < EPV; ++x) {
                                const float d = zf[x] - e[j][x];
                                loss = fmaf(d, d, loss);
                                acc[j][x] += (stat_kind & 15) == VQ_STAT_SUM_RESID ? d : zf[x];
                            }
                            if (q != nullptr) st_stream_16(q + (int64_t)r * ldq + v * EPV, qv[j]);
                        }
                    }
                    ++cnt;
                    if (X.idx64 != nullptr && !X.idx64_pre && lg == 0) X.idx64[r] = c;
                } else {
                    // non-finite latent row: idx := 0, q := E[0], no contribution to loss or statistics
#pragma unroll
                    for (int j = 0; j < VPL; ++j) {
                        const int v = lg + j * LPR;
                        if (v < V) {
                            float e0[EPV];
#pragma unroll
                            for (int x = 0; x < EPV; ++x) e0[x] = __ldg(E + v * EPV + x);
                            if (q != nullptr) st_stream_16(q + (int64_t)r * ldq + v * EPV, IO::pack(e0));
                        }
                    }
                    if (lg == 0) {
                        idx[r] = 0;
                        if (X.idx64 != nullptr) X.idx64[r] = 0;
                    }
                }
            }
        }
    }
#if VQ_PASS_BLOCKSUM
    {
        const bool mine = blk_on && cur >= 0 && cur == blk_code && cnt > 0;
        if (!mine) flush();
        if (blk_on) {
            constexpr int ROWF = VPL * LPR * EPV;     // floats per lane-group slot
            const int grp_i = (threadIdx.x >> 5) * G + g;
            float* slots = reinterpret_cast<float*>(pass_smem);
            __syncthreads();                          // every lane group is done with the ring
            if (lg == 0) blk_cnts[grp_i] = mine ? cnt : 0;
            if (mine) {
#pragma unroll
                for (int j = 0; j < VPL; ++j) {
                    const int v = lg + j * LPR;
                    if (v < V) {
#pragma unroll
                        for (int x = 0; x < EPV; x += 4)
                            *reinterpret_cast<float4*>(slots + grp_i * ROWF + v * EPV + x) =
                                make_float4(acc[j][x], acc[j][x + 1], acc[j][x + 2], acc[j][x + 3]);
                    }
                }
            }
            __syncthreads();
            int total = 0;
#pragma unroll
            for (int i = 0; i < 8 * G; ++i) total += blk_cnts[i];
            if (total > 0) {
                float* dst = code_sum + (int64_t)blk_code * D;
                for (int i = threadIdx.x * 4; i < D; i += 256 * 4) {
                    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int gi = 0; gi < 8 * G; ++gi) {          // fixed order
                        if (blk_cnts[gi] > 0) {
                            const float4 t = *reinterpret_cast<const float4*>(slots + gi * ROWF + i);
                            sum.x += t.x; sum.y += t.y; sum.z += t.z; sum.w += t.w;
                        }
                    }
                    red_add_v4(dst + i, sum.x, sum.y, sum.z, sum.w);
                }
                if (threadIdx.x == 0) atomicAdd(code_cnt + blk_code, (float)total);
            }
        }
    }
#else
    flush();
#endif
    loss = warp_sum(loss);
    if (lane == 0) loss_partial[warp_global] = loss;

    // The block that finishes last adds the per-warp partial sums in a fixed order (double accumulate): no extra
    // launch, and the result does not depend on which block happens to be last.
    __shared__ bool is_last;
    __shared__ double red[8];
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(done, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (is_last) {
        __threadfence();
        const int64_t nparts = (int64_t)gridDim.x * 8;
        double acc_d = 0.0;
        for (int64_t i = threadIdx.x; i < nparts; i += 256) acc_d += (double)__ldcg(loss_partial + i);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc_d += __shfl_xor_sync(0xffffffffu, acc_d, o);
        if (lane == 0) red[threadIdx.x >> 5] = acc_d;
        __syncthreads();
        if (threadIdx.x == 0) {
            double tot = 0.0;
#pragma unroll
            for (int i = 0; i < 8; ++i) tot += red[i];
            const float total = (stat_kind & VQ_STAT_ACCUMULATE) ? sq_err_out[0] + (float)tot : (float)tot;
            sq_err_out[0] = total;
            if (X.loss_out != nullptr) X.loss_out[0] = total * X.loss_scale;
            *done = 0;
        }
    }
}

// any D / alignment: one warp per row, scalar accesses, plain atomics.  Correct but slow; only odd shapes land here.
template <typename T>
__global__ void __launch_bounds__(256)
pass_generic_kernel(const T* __restrict__ z, int64_t N, int D, int64_t ldz, int32_t* __restrict__ idx,
                    const float* __restrict__ E, int K, T* __restrict__ q, int64_t ldq,
                    float* __restrict__ loss_partial, float* __restrict__ code_cnt, float* __restrict__ code_sum,
                    int stat_kind, const PassExtra X) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    float loss = 0.f;
    if (row < N) {
        int c = idx[row];
        if ((unsigned)c >= (unsigned)K) c = 0;
        bool nf = false;
        for (int d = lane; d < D; d += 32) {
            const float v = (float)z[row * ldz + d];
            nf |= !isfinite(v);
        }
        const bool bad = __any_sync(0xffffffffu, nf);
        if (bad) {
            c = 0;
            if (lane == 0) idx[row] = 0;
        }
        if (X.idx64 != nullptr && lane == 0) X.idx64[row] = c;
        for (int d = lane; d < D; d += 32) {
            const float ev = E[(int64_t)c * D + d];
            if (q != nullptr) q[row * ldq + d] = (T)ev;
            if (!bad) {
                const float zf = (float)z[row * ldz + d];
                const float df = zf - ev;
                loss = fmaf(df, df, loss);
                if (code_sum != nullptr) atomicAdd(code_sum + (int64_t)c * D + d, (stat_kind & 15) == VQ_STAT_SUM_RESID ? df : zf);
            }
        }
        if (!bad && lane == 0 && code_cnt != nullptr) atomicAdd(code_cnt + c, 1.f);
    }
    loss = warp_sum(loss);
    if (lane == 0) loss_partial[row < N ? row : N] = loss;   // slot N is a discard slot
}

// deterministic mode: code_sum[k] = sum over the slices that hold rows of code k, in slice order (one warp per code)
__global__ void __launch_bounds__(256)
reduce_dpart_kernel(const float* __restrict__ dpart, const int* __restrict__ start, int K, int D,
                    float* __restrict__ code_sum, int accumulate) {
    const int lane = threadIdx.x & 31;
    const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (k >= K) return;
    const int a = start[k], b = start[k + 1];
    for (int d = lane; d < D; d += 32) {
        float acc = 0.f;
        if (b > a)
            for (int sl = a / LG; sl <= (b - 1) / LG; ++sl) acc += dpart[((int64_t)sl + k) * D + d];
        code_sum[(int64_t)k * D + d] = accumulate ? code_sum[(int64_t)k * D + d] + acc : acc;
    }
}

// fixed-order reduction of the per-warp partial sums (double accumulate) -> out[0]
__global__ void __launch_bounds__(1024) reduce_partials_kernel(const float* __restrict__ part, int64_t n,
                                                               float* __restrict__ out, int accumulate,
                                                               const PassExtra X) {
    __shared__ double sh[32];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += 1024) s += (double)part[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = sh[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) {
            const float total = accumulate ? out[0] + (float)s : (float)s;
            out[0] = total;
            if (X.loss_out != nullptr) X.loss_out[0] = total * X.loss_scale;
        }
    }
}

template <typename T, int VPL, int LPR>
int launch_pass_t(const void* z, int64_t N, int D, int64_t ldz, int32_t* idx, const int32_t* perm, const float* E,
                  int K, void* q, int64_t ldq, float* part, int64_t* nparts, float* code_cnt, float* code_sum,
                  float* dpart, int stat_kind, int* done, float* sq_err, PassExtra X, cudaStream_t st) {
    constexpr int G = 32 / LPR;
    // rows per lane group: 64, fewer while that leaves SMs without a block (deterministic mode sizes its slots for 64)
    int lgr = LG;
    if (dpart == nullptr) {
        const int64_t want_groups = (int64_t)sm_count() * 2 * 8 * G;          // two blocks per SM
        while (lgr > 16 && ceil_div64(N, lgr) < want_groups) lgr >>= 1;
#ifdef VQ_TUNING
        static int env_lgr = -1;                 // tuning knob: VQ_PASS_LGR=<rows per lane group, multiple of 4, 16..64>
        if (env_lgr < 0) {
            const char* ev = getenv("VQ_PASS_LGR");
            env_lgr = ev != nullptr ? atoi(ev) : 0;
        }
        if (env_lgr >= 16 && env_lgr <= LG && env_lgr % 4 == 0) lgr = env_lgr;
#endif
    }
    const int64_t groups = ceil_div64(N, lgr);
    const int64_t warps = ceil_div64(groups, G);
    const int64_t blocks = ceil_div64(warps, 8);
    *nparts = blocks * 8;
    constexpr int U = PassCfg<VPL>::U;
    constexpr int UB = U < LPR ? U : LPR;
    const size_t smem = (size_t)PASS_STAGES * UB * VPL * 256 * 16 + (size_t)8 * 2 * G * LG * sizeof(int) +
                        (VQ_PASS_BULK ? (size_t)8 * G * PASS_STAGES * 8 : 0);
    static int attr_dev = -1;                  // per template instance; function attributes are per device
    const int cur_dev = current_device();
    if (attr_dev != cur_dev) {
        cudaFuncSetAttribute(pass_kernel<T, VPL, LPR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        // (left to the driver: forcing the largest shared-memory carve-out for a third block per SM shrinks L1, which
        // the codebook-row and idx gathers live on -- C4 pass group 145 -> 160 us)
        attr_dev = cur_dev;
    }
    VQ_CHECK_CUDA(launch_pdl(PDL_PASS, pass_kernel<T, VPL, LPR>, dim3((unsigned)blocks), dim3(256), smem, st, static_cast<const T*>(z),
                             N, D, ldz, idx, perm, E, K, static_cast<T*>(q), ldq, part, code_cnt, code_sum, dpart, stat_kind,
                             done, sq_err, X, lgr));
    return VQ_OK;
}

template <typename T>
int launch_pass_dispatch(int V, const void* z, int64_t N, int D, int64_t ldz, int32_t* idx, const int32_t* perm,
                         const float* E, int K, void* q, int64_t ldq, float* part, int64_t* nparts, float* code_cnt,
                         float* code_sum, float* dpart, int stat_kind, int* done, float* sq_err, PassExtra X,
                         cudaStream_t st) {
#define VQ_PASS(VPL, LPR) \
    return launch_pass_t<T, VPL, LPR>(z, N, D, ldz, idx, perm, E, K, q, ldq, part, nparts, code_cnt, code_sum, dpart, stat_kind, done, sq_err, X, st)
    if (V <= 4) VQ_PASS(1, 4);
    if (V <= 8) VQ_PASS(1, 8);
    if (V <= 16) VQ_PASS(1, 16);
    if (V <= 32) VQ_PASS(2, 16);      // two 16-byte pieces per lane: less per-row control work (C4 group 158.6 -> 151.4 us; <1,32> before)
    if (V <= 64) VQ_PASS(2, 32);
    VQ_PASS(4, 32);
#undef VQ_PASS
}

}  // namespace

// Workspace carve-up shared by vq_workspace_bytes and the launchers.
WsLayout ws_layout(int64_t N, int K, int D, int in_dtype, int search) {
    auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
    WsLayout L;
    size_t o = 0;
    L.scal = o;      o += 256;
    o += al((size_t)K * 4);                               // offset 256: rank-reduced counts of vq_ema_update_dp
    // the counter region sits at an offset that depends on K only (the VQ_SEARCH_WS_CLEAN contract is about it)
    L.cnt = o;       o += al((size_t)(2 * K + 4) * 4);     // [histogram K | cursor K | done counter]: one memset
    L.cursor = L.cnt + (size_t)K * 4;
    L.start = o;     o += al((size_t)(K + 1) * 4);
    L.perm = o;      o += al((size_t)N * 4);
    L.rank = o;      o += al((size_t)N * 4);              // VQ_SEARCH_PRESORT: rank of every row inside its code
    L.partial = o;   o += al((size_t)(N + 64) * 4);       // generic kernel: one slot per row (+ discard slot)
    L.rows_per_chunk = (int)((ceil_div64(N, 256) + 31) / 32 * 32);
    if (L.rows_per_chunk < 1024) L.rows_per_chunk = 1024;
    L.nchunks = (int)ceil_div64(N, L.rows_per_chunk);
    L.chunk_cnt = o; o += al((size_t)L.nchunks * K * 4);
    L.dpart = o;     o += al((size_t)(ceil_div64(N, LG) + K + 1) * D * 4);   // deterministic mode: one slot per run
    // staging for the tensor-core searches of non-native latents (LAST region: nothing else moves with the search kind):
    // the split fp32 tier takes two bf16 planes of any latent dtype, fp32 / fp16 latents under VQ_SEARCH_BF16 take one
    const int planes = search == VQ_SEARCH_F32 ? (assign_tc_split_supported(N, D, K) ? 2 : 0)
                                               : ((search == VQ_SEARCH_BF16 || search < 0) && in_dtype != VQ_DT_BF16 ? 1 : 0);
    L.zbf16 = o;     o += al((size_t)N * D * 2 * planes);
    L.total = o;
    return L;
}

int launch_gather_pass(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, int32_t* idx, const float* E_f32,
                       int K, void* q_out, int64_t ldq, float* sq_err_sum, float* code_cnt, float* code_sum,
                       int stat_kind, int deterministic, PassExtra X, void* ws, cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && N < (int64_t(1) << 31) && D > 0 && K > 0 && ldz >= D && ldq >= D,
               "vq_gather_st_loss_scatter: bad shape N=%lld D=%d K=%d", (long long)N, D, K);
    VQ_REQUIRE((code_cnt == nullptr) == (code_sum == nullptr), "code_cnt and code_sum must both be given or both NULL");
    const bool stats = code_cnt != nullptr;
    const bool accum = (stat_kind & VQ_STAT_ACCUMULATE) != 0;    // add to code_cnt / code_sum / sq_err_sum instead of overwriting
    if (N == 0) {
        if (accum) return VQ_OK;
        if (X.loss_out != nullptr) VQ_CHECK_CUDA(cudaMemsetAsync(X.loss_out, 0, 4, st));
        if (stats) {
            VQ_CHECK_CUDA(cudaMemsetAsync(code_cnt, 0, (size_t)K * 4, st));
            VQ_CHECK_CUDA(cudaMemsetAsync(code_sum, 0, (size_t)K * D * 4, st));
        }
        VQ_CHECK_CUDA(cudaMemsetAsync(sq_err_sum, 0, 4, st));
        return VQ_OK;
    }
    VQ_REQUIRE(ws != nullptr, "vq_gather_st_loss_scatter: workspace required");
    const WsLayout L = ws_layout(N, K, D, in_dtype, -1);         // the pass uses none of the search's staging
    uint8_t* w = static_cast<uint8_t*>(ws);
    int* cnt_i = reinterpret_cast<int*>(w + L.cnt);
    int* start = reinterpret_cast<int*>(w + L.start);
    int* cursor = reinterpret_cast<int*>(w + L.cursor);
    int* done = cnt_i + 2 * K;
    int32_t* perm = reinterpret_cast<int32_t*>(w + L.perm);
    float* part = reinterpret_cast<float*>(w + L.partial);
    int* chunk_cnt = reinterpret_cast<int*>(w + L.chunk_cnt);
    bool stats_cleared = false;

    const int es = in_dtype == VQ_DT_F32 ? 4 : 2;                // bf16 and fp16 rows are both 2-byte
    const int epv = 16 / es;
    const bool vec_ok = D % epv == 0 && (ldz * es) % 16 == 0 && (ldq * es) % 16 == 0 &&
                        (reinterpret_cast<uintptr_t>(z) & 15) == 0 && (reinterpret_cast<uintptr_t>(q_out) & 15) == 0 &&
                        (reinterpret_cast<uintptr_t>(E_f32) & 15) == 0 && D / epv <= 128;
    // VQ_STAT_PRESORTED: histogram, scan and ranks are already in ws (left by the search kernel); the pass kernel's
    // completion counter is zero (the search's memset cleared it, and the pass resets it when it ends)
    const bool presort_flag = (stat_kind & VQ_STAT_PRESORTED) != 0;
    const bool presorted = presort_flag && vec_ok && stats && !deterministic;
    stat_kind &= ~VQ_STAT_PRESORTED;
    // a path that builds its own histogram in the counter region puts the zeros back afterwards when the caller works
    // under the VQ_SEARCH_PRESORT / VQ_SEARCH_WS_CLEAN contract
    struct Rezero {
        int* p; size_t n; cudaStream_t st; bool on;
        ~Rezero() { if (on) (void)cudaMemsetAsync(p, 0, n, st); }
    } rezero{cnt_i, (size_t)(2 * K + 2) * 4, st, presort_flag && !presorted};
    // otherwise histogram, cursor and the pass kernel's completion counter are cleared by ONE memset
    if (!presorted) VQ_CHECK_CUDA(cudaMemsetAsync(cnt_i, 0, (size_t)(2 * K + 1) * 4, st));
    int64_t nparts = 0;
    if (!vec_ok) {
        const int64_t blocks = ceil_div64(N, 8);
        nparts = N;
        if (stats && !accum) {
            VQ_CHECK_CUDA(cudaMemsetAsync(code_cnt, 0, (size_t)K * 4, st));
            VQ_CHECK_CUDA(cudaMemsetAsync(code_sum, 0, (size_t)K * D * 4, st));
        }
        if (in_dtype == VQ_DT_F32)
            pass_generic_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const float*>(z), N, D, ldz, idx,
                                                                         E_f32, K, static_cast<float*>(q_out), ldq,
                                                                         part, code_cnt, code_sum, stat_kind, X);
        else if (in_dtype == VQ_DT_F16)
            pass_generic_kernel<__half><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const __half*>(z), N, D, ldz, idx,
                                                                          E_f32, K, static_cast<__half*>(q_out), ldq,
                                                                          part, code_cnt, code_sum, stat_kind, X);
        else
            pass_generic_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(
                static_cast<const __nv_bfloat16*>(z), N, D, ldz, idx, E_f32, K, static_cast<__nv_bfloat16*>(q_out), ldq,
                part, code_cnt, code_sum, stat_kind, X);
        VQ_CHECK_CUDA(cudaGetLastError());
        count_launch();
    } else {
        const int32_t* perm_arg = nullptr;
        float* dpart = nullptr;
        if (stats) {
            if (deterministic) {
                const int gs = (int)(ceil_div64(N, 256) < sm_count() * 8 ? ceil_div64(N, 256) : sm_count() * 8);
                VQ_CHECK_CUDA(cudaMemsetAsync(chunk_cnt, 0, (size_t)L.nchunks * K * 4, st));
                hist_kernel<<<gs, 256, 0, st>>>(idx, N, K, cnt_i, chunk_cnt, L.rows_per_chunk);
                scan_kernel<<<1, 1024, 0, st>>>(cnt_i, K, start, cursor);
                chunk_offsets_kernel<<<(K + 255) / 256, 256, 0, st>>>(chunk_cnt, L.nchunks, K, start);
                scatter_stable_kernel<<<(L.nchunks + 3) / 4, 128, 0, st>>>(idx, N, K, chunk_cnt, L.rows_per_chunk,
                                                                           L.nchunks, perm);
                count_launch(4);
                dpart = reinterpret_cast<float*>(w + L.dpart);
            } else if (presorted) {
                const bool zero_here = !accum && (reinterpret_cast<uintptr_t>(code_sum) & 15) == 0 && ((size_t)K * D) % 4 == 0;
                const bool vec4 = (reinterpret_cast<uintptr_t>(idx) & 15) == 0 &&
                                  (reinterpret_cast<uintptr_t>(X.idx64) & 15) == 0;
                const int64_t work = vec4 ? ceil_div64(N, 4) : N;
                const int64_t want = ceil_div64(work > (int64_t)K * D / 8 ? work : (int64_t)K * D / 8, 256);
                const int gs = (int)(want < (int64_t)sm_count() * 32 ? (want < 1 ? 1 : want) : (int64_t)sm_count() * 32);
                const int* rank = reinterpret_cast<const int*>(w + L.rank);
                float4* za = reinterpret_cast<float4*>(code_sum);
                const int64_t na4 = zero_here ? (int64_t)K * D / 4 : 0;
                VQ_CHECK_CUDA(launch_pdl(PDL_SCATTER, vec4 ? scatter_ranked_kernel<true> : scatter_ranked_kernel<false>, dim3(gs), dim3(256),
                                         0, st, idx, rank, N, K, start, perm, za, na4, code_cnt, zero_here ? K : 0, X.idx64));
                count_launch(1);
                stats_cleared = zero_here;
                X.idx64_pre = X.idx64 != nullptr;
            } else if ((size_t)K * 12 <= 200 * 1024 && (reinterpret_cast<uintptr_t>(code_sum) & 15) == 0 &&
                       ((size_t)K * D) % 4 == 0) {
                // shared-memory privatised counting sort, 2 launches: block b owns R consecutive rows; the second
                // kernel rebuilds the scan per block and clears code_sum / code_cnt on the way
                int R = (int)ceil_div64(N, 2 * sm_count());
                R = round_up(R < 1024 ? 1024 : R, 1024);
                const int nb = (int)ceil_div64(N, R);
                static int attr_dev = -1;      // function attributes are per device
                const int cur_dev = current_device();
                if (attr_dev != cur_dev) {
                    VQ_CHECK_CUDA(cudaFuncSetAttribute(hist_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                    VQ_CHECK_CUDA(cudaFuncSetAttribute(scan_scatter_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                    attr_dev = cur_dev;
                }
                hist_smem_kernel<<<nb, 512, (size_t)K * 4, st>>>(idx, N, K, R, cnt_i);
                scan_scatter_smem_kernel<<<nb, 512, (size_t)K * 12, st>>>(idx, N, K, R, cnt_i, cursor, perm,
                                                                          reinterpret_cast<float4*>(code_sum),
                                                                          accum ? 0 : (int64_t)K * D / 4, code_cnt,
                                                                          accum ? 0 : K);
                count_launch(2);
                stats_cleared = true;
            } else {
                const int gs = (int)(ceil_div64(N, 256) < sm_count() * 8 ? ceil_div64(N, 256) : sm_count() * 8);
                hist_kernel<<<gs, 256, 0, st>>>(idx, N, K, cnt_i, nullptr, L.rows_per_chunk);
                scan_kernel<<<1, 1024, 0, st>>>(cnt_i, K, start, cursor);
                scatter_kernel<<<gs, 256, 0, st>>>(idx, N, K, cursor, perm);
                count_launch(3);
            }
            VQ_CHECK_CUDA(cudaGetLastError());
            perm_arg = perm;
            if (!stats_cleared && !accum) {
                VQ_CHECK_CUDA(cudaMemsetAsync(code_cnt, 0, (size_t)K * 4, st));
                VQ_CHECK_CUDA(cudaMemsetAsync(code_sum, 0, (size_t)K * D * 4, st));
            }
        }
        const int V = D / epv;
        if (in_dtype == VQ_DT_F32)
            launch_pass_dispatch<float>(V, z, N, D, ldz, idx, perm_arg, E_f32, K, q_out, ldq, part, &nparts, code_cnt,
                                        code_sum, dpart, stat_kind, done, sq_err_sum, X, st);
        else if (in_dtype == VQ_DT_F16)
            launch_pass_dispatch<__half>(V, z, N, D, ldz, idx, perm_arg, E_f32, K, q_out, ldq, part, &nparts, code_cnt,
                                         code_sum, dpart, stat_kind, done, sq_err_sum, X, st);
        else
            launch_pass_dispatch<__nv_bfloat16>(V, z, N, D, ldz, idx, perm_arg, E_f32, K, q_out, ldq, part, &nparts,
                                                code_cnt, code_sum, dpart, stat_kind, done, sq_err_sum, X, st);
        VQ_CHECK_CUDA(cudaGetLastError());
        count_launch();
        if (dpart != nullptr) {
            reduce_dpart_kernel<<<(K + 7) / 8, 256, 0, st>>>(dpart, start, K, D, code_sum, accum ? 1 : 0);
            VQ_CHECK_CUDA(cudaGetLastError());
            count_launch();
        }
        return VQ_OK;                       // sq_err_sum was written by the pass kernel's last block
    }
    reduce_partials_kernel<<<1, 1024, 0, st>>>(part, nparts, sq_err_sum, accum ? 1 : 0, X);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

}  // namespace vq
