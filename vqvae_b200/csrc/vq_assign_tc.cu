// a-1  Nearest-codebook search on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
//   idx[n] = argmin_k ( ||e_k||^2 - 2 z_n . e_k )        z: [N, D] bf16,  E: [K, D] bf16,  ||e||^2: fp32
//
// The N x K score matrix never exists in HBM: a CTA pair (two SMs of a TPC, tcgen05 cta_group::2) owns a 256-row slab
// of z, 128 rows per CTA, kept resident in shared memory for the whole sweep; the codebook streams through a TMA ring in
// 256-code x 64-feature stages of which each CTA holds 128 codes; S = z E^T accumulates for one 256 x 256 tile in the
// TMEM of both SMs (fp32, 128 rows each), and the epilogue warps fold each finished tile into a running (min, argmin)
// per row straight out of TMEM while the tensor cores work on the next tile (two accumulator buffers = all 512 TMEM
// columns).  kPair = false is the same kernel on single CTAs (cta_group::1, 128-row slabs, whole stages per CTA).
//
// Warp roles (128 + 32 * kEpiWarps threads, 1 CTA / SM, persistent over row slabs):
//   warp 0      TMA producer (one elected lane; in a pair both CTAs load, completion goes to the leader's barriers)
//   warp 1      tcgen05.mma issuer (one elected lane; in a pair the leader CTA only)
//   warp 2      TMEM allocator / deallocator
//   warp 3      VQ_SEARCH_PRESORT only: takes each slab's final indices from the epilogue (shared memory, named barriers)
//               and claims their slots in the per-code histogram (warp-aggregated atomics) -> rank of every row inside
//               its code; the block that finishes last scans the histogram (presort_finish).  The gather pass then
//               needs no histogram launches of its own.
//   warps 4..   epilogue (8 warps in pairs, 16 on single CTAs): warp w reads TMEM lanes 32*(w%4)..+31 (one row per
//               thread) and one column part of each tile; the parts are merged through shared memory per slab.
//
// Operand kinds (all tcgen05 kind::f16, fp32 accumulate): bf16 latents x bf16 codebook shadow, fp16 latents x fp16 shadow
// (the reference script's default autocast dtype, scripts/train_vqvae.py:284-286 -- no staging pass, no cast), and the
// SPLIT form of the fp32 (1e-5) tier: z = z_hi + z_lo, e = e_hi + e_lo (bf16 pieces, 16 significant bits each), and
// z.e ~ z_hi.e_hi + z_lo.e_hi + z_hi.e_lo accumulated in ONE TMEM accumulator -- three MMAs per product, the dropped
// terms are below 2^-16 |z||e|.  The slab then holds two planes of z and the ring alternates e_hi / e_lo stages.
//
// Required by: scripts/train_vqvae.py:11,287 (VectorQuantizer inside model(batch));
//              frl/config/frl_model_v0.yaml:29-33 (kind: vq, K, D).  Definition: SURVEY.md 8(a-1).
#include <math_constants.h>
#include <stdio.h>
#include <stdlib.h>

#include "vq_common.cuh"

namespace vq {
namespace {

constexpr int BM = 128;             // rows per slab   (UMMA M)
constexpr int BN = 256;             // codes per tile  (UMMA N)
constexpr int BK = 64;              // features per smem K-block = one 128-byte swizzle row of bf16
constexpr int UK = 16;              // UMMA K for 16-bit inputs
constexpr int A_KB_BYTES = BM * BK * 2;   // 16 KiB
constexpr int B_ST_BYTES = BN * BK * 2;   // 32 KiB
constexpr int MAX_STAGES = 8;
constexpr int MAX_EPI_PARTS = 4;   // epilogue warps = 4 TMEM lane quarters x column parts (2 or 4)
constexpr int TMEM_COLS = 512;
constexpr int SMEM_LIMIT = 232448;  // 227 KiB opt-in maximum per CTA
constexpr int NORM_SMEM_MAX = 8192; // codes whose ||e||^2 may be staged in shared memory (32 KiB)

// What the MMAs multiply (host-chosen, see launch_assign_tc): instruction descriptor with the operand formats, the
// A-side partner row (c1, c2, c3, 0, ...) of the mode-3 norm image as 16-bit patterns, and the split-plane geometry.
struct OperandKind {
    uint32_t idesc;          // kind::f16 instruction descriptor (bf16 or fp16 operands, fp32 accumulate, M, N)
    uint32_t ones01, ones2;  // c1 | c2 << 16, c3: sum_i c_i * piece_i reproduces -||e||^2 / 2
    int planes;              // 1, or 2 = split fp32 tier (hi / lo bf16 planes of z and of E)
    int a_plane_rows;        // row offset of z's second plane in the staged [2][N][D] image
    int b_plane_rows;        // row offset of E's second plane in the [2][K_pad][D] shadow
};

struct Ctl {
    uint64_t a_full[2], a_empty[2];
    uint64_t b_full[MAX_STAGES], b_empty[MAX_STAGES];
    uint64_t t_full[2], t_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
    float xbest[2][MAX_EPI_PARTS][BM];
    int xidx[2][MAX_EPI_PARTS][BM];
    int scan_sh[36];                   // VQ_SEARCH_PRESORT: scratch of presort_finish
    int fin_idx[2][BM];                // VQ_SEARCH_PRESORT: a slab's final indices on their way to warp 3
};

__device__ __forceinline__ float min3f(float a, float b, float c) {
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
// (d0, d1) = (-2, -2) * (a0, a1) + (c0, c1): one FFMA2, each half an IEEE fp32 fma (same bits as fmaf)
__device__ __forceinline__ void score2(float& d0, float& d1, uint32_t a0, uint32_t a1, float c0, float c1) {
    uint64_t a, c, d;
    const uint64_t m2 = 0xC0000000C0000000ull;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "r"(a0), "r"(a1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(c0), "f"(c1));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(m2), "l"(c));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

// Fold 32 accumulator columns of one row into the running (min, argmin); lowest column wins ties, earlier chunks
// win ties across chunks.  The comparison pipe (FMNMX / FSETP / SEL) and the power budget are the scarce resources
// next to a saturated tensor pipe, so the common case is kept to 16 FFMA2 (scores) + 18 FMNMX3 (minimum, two new
// values per instruction) + one compare; the index search runs only when this chunk holds a new strict minimum
// for some lane.  kMode selects what happens then:
//   0  straight-line compare/select search over the 32 columns (always exact, 62 ALU instructions)
//   1  the same search restricted to the group of 8 columns that holds the minimum (pays off when few lanes hit)
//   2  DEFERRED: the lane only parks the winning group of 8 scores in its private shared-memory slot (two predicated
//      STS.128) and remembers the group's first column in bidx; the within-group search happens once per slab in
//      finish_deferred().  With K/32 chunks against 32 rows per warp nearly every chunk is a hit for some lane when
//      K is small, so this takes the per-chunk hit cost from ~68 to ~18 issue slots and leaves the comparison pipe
//      to the minimum itself.
template <bool kNormSmem, int kMode>
__device__ __forceinline__ void fold_chunk(const uint32_t (&r)[32], const float* __restrict__ nrm_g, uint32_t nrm_s,
                                           int col0, float& best, int& bidx, uint32_t slot, uint32_t slot_stride, int dbg = 0) {
    float s[32];
#ifdef VQ_TUNING
    if (dbg == 3) {                    // timing experiment (results are WRONG): no norm loads, no FFMA
#pragma unroll
        for (int j = 0; j < 32; ++j) s[j] = __uint_as_float(r[j]);
    } else
#endif
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float4 n4;
        if (kNormSmem)      // same address in every lane: one broadcast LDS.128
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(n4.x), "=f"(n4.y), "=f"(n4.z), "=f"(n4.w)
                         : "r"(nrm_s + (uint32_t)(col0 + 4 * j) * 4u));
        else
            n4 = __ldg(reinterpret_cast<const float4*>(nrm_g + col0) + j);
        score2(s[4 * j + 0], s[4 * j + 1], r[4 * j + 0], r[4 * j + 1], n4.x, n4.y);
        score2(s[4 * j + 2], s[4 * j + 3], r[4 * j + 2], r[4 * j + 3], n4.z, n4.w);
    }
    float m[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        m[c] = min3f(s[8 * c], s[8 * c + 1], s[8 * c + 2]);
        m[c] = min3f(m[c], s[8 * c + 3], s[8 * c + 4]);
        m[c] = min3f(m[c], s[8 * c + 5], s[8 * c + 6]);
        m[c] = fminf(m[c], s[8 * c + 7]);
    }
    const float mm = fminf(min3f(m[0], m[1], m[2]), m[3]);
    if (kMode == 2) {
        const bool hit = mm < best;
        const bool p0 = hit && m[0] == mm;
        const bool p1 = hit && !p0 && m[1] == mm;
        const bool p2 = hit && !p0 && !p1 && m[2] == mm;
        const bool p3 = hit && !p0 && !p1 && !p2;
#define VQ_PARK(P, B)                                                                                            \
    if (P) {                                                                                                     \
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(slot), "f"(s[(B)]), "f"(s[(B) + 1]),       \
                     "f"(s[(B) + 2]), "f"(s[(B) + 3]) : "memory");                                               \
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(slot + slot_stride), "f"(s[(B) + 4]),      \
                     "f"(s[(B) + 5]), "f"(s[(B) + 6]), "f"(s[(B) + 7]) : "memory");                              \
        bidx = col0 + (B);                                                                                       \
    }
        VQ_PARK(p0, 0)
        VQ_PARK(p1, 8)
        VQ_PARK(p2, 16)
        VQ_PARK(p3, 24)
#undef VQ_PARK
        if (hit) best = mm;
        return;
    }
    if (mm < best) {
        best = mm;
        int jj;
#define VQ_SEARCH8(B)                                     \
    {                                                     \
        jj = (B) + 7;                                     \
        _Pragma("unroll") for (int i = 6; i >= 0; --i)    \
            if (s[(B) + i] == mm) jj = (B) + i;           \
    }
        if (kMode == 1) {   // few lanes hit per chunk (large K): look only inside the group of 8 that holds the minimum
            if (m[0] == mm) VQ_SEARCH8(0)
            else if (m[1] == mm) VQ_SEARCH8(8)
            else if (m[2] == mm) VQ_SEARCH8(16)
            else VQ_SEARCH8(24)
        } else {            // most lanes hit in most chunks (small K): one straight-line pass, no divergence
            jj = 31;
#pragma unroll
            for (int j = 30; j >= 0; --j)
                if (s[j] == mm) jj = j;
        }
#undef VQ_SEARCH8
        bidx = col0 + jj;
    }
}

__device__ __forceinline__ float max3f(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// kMode 3 ("aug"): the accumulator already holds t = z.e - ||e||^2/2 (the norm rode through the tensor core as one more
// K step), so the nearest code is the ARG-MAX of the raw accumulator: no norm loads, no FFMA -- 18 FMNMX3 plus the
// parking of the winning group of 8 (see kMode 2).  nbest is the running maximum.
__device__ __forceinline__ void fold_chunk_aug(const uint32_t (&r)[32], int col0, float& nbest, int& bidx, uint32_t slot,
                                               uint32_t slot_stride) {
    float m[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        m[c] = max3f(__uint_as_float(r[8 * c]), __uint_as_float(r[8 * c + 1]), __uint_as_float(r[8 * c + 2]));
        m[c] = max3f(m[c], __uint_as_float(r[8 * c + 3]), __uint_as_float(r[8 * c + 4]));
        m[c] = max3f(m[c], __uint_as_float(r[8 * c + 5]), __uint_as_float(r[8 * c + 6]));
        m[c] = fmaxf(m[c], __uint_as_float(r[8 * c + 7]));
    }
    const float mm = fmaxf(max3f(m[0], m[1], m[2]), m[3]);
    const bool hit = mm > nbest;
    const bool p0 = hit && m[0] == mm;
    const bool p1 = hit && !p0 && m[1] == mm;
    const bool p2 = hit && !p0 && !p1 && m[2] == mm;
    const bool p3 = hit && !p0 && !p1 && !p2;
#define VQ_PARK(P, B)                                                                                              \
    if (P) {                                                                                                       \
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(slot), "r"(r[(B)]), "r"(r[(B) + 1]),         \
                     "r"(r[(B) + 2]), "r"(r[(B) + 3]) : "memory");                                                 \
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(slot + slot_stride), "r"(r[(B) + 4]),        \
                     "r"(r[(B) + 5]), "r"(r[(B) + 6]), "r"(r[(B) + 7]) : "memory");                                \
        bidx = col0 + (B);                                                                                         \
    }
    VQ_PARK(p0, 0)
    VQ_PARK(p1, 8)
    VQ_PARK(p2, 16)
    VQ_PARK(p3, 24)
#undef VQ_PARK
    if (hit) nbest = mm;
}

// kMode 2 / 3: resolve the parked group of 8 scores into the column index (once per slab and row)
__device__ __forceinline__ int finish_deferred(float best, int bcol, uint32_t slot, uint32_t slot_stride, bool any_hit) {
    float v[8];
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(slot));
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "r"(slot + slot_stride));
    int i = 7;
#pragma unroll
    for (int j = 6; j >= 0; --j)
        if (v[j] == best) i = j;
    return any_hit ? bcol + i : 0;       // no hit at all (every score NaN or at the initial bound): index 0
}

// kPair: the kernel runs as clusters of two CTAs on the two SMs of a TPC and every MMA is a tcgen05 cta_group::2
// instruction over a 256-row slab (128 rows per CTA) x 256 codes (128 codes of each stage in each CTA's shared memory,
// see vq_common.cuh).  Per SM that halves the codebook bytes pulled from L2 (the binding resource of the single-CTA
// kernel: 148 SMs x 64 B/clk against an L2 that delivers about 43 B/clk/SM) and the operand bytes read from shared memory.
// kPlanes (1, or 2 = split fp32 tier) is a template parameter on purpose: as a run-time loop bound in the producer and the MMA
// issuer it cost the bf16 search 12-15 % at D = 256 (1.42 -> 1.60 ms at C4, same box, profiles/r02/search_planes_ab.log).
// kNKB: number of 64-feature K-blocks when D is 64, 128 or 256 (the shapes of BASELINE.json), 0 = read it from the argument.
// Same reason: with the K-block loops fully unrolled the search is another 4 % faster (C4 1.33 -> 1.28 ms, C2 261 -> 250 us).
template <bool kNormSmem, int kEpiWarps, int kMode, bool kPair, int kPlanes, int kNKB>
__global__ void __launch_bounds__(128 + 32 * kEpiWarps, 1)
assign_tc_kernel(const __grid_constant__ CUtensorMap tm_z, const __grid_constant__ CUtensorMap tm_e,
                 const __grid_constant__ CUtensorMap tm_aug, const float* __restrict__ norms, int64_t N, int K, int D,
                 int nkb_arg, int nstages, int nabuf, const OperandKind ok, int32_t* __restrict__ idx, float* __restrict__ smin,
#ifdef VQ_TUNING
                 int dbg,               // 1..4: timing experiments of the tuning build (results are WRONG); 0 = off
#endif
                 const Presort ps) {
#ifndef VQ_TUNING
    constexpr int dbg = 0;             // the product build has no result-changing switches
#endif
    const int nkb = kNKB != 0 ? kNKB : nkb_arg;          // compile-time for D = 64 * kNKB
    constexpr int B_ROWS = kPair ? BN / 2 : BN;          // codes of a stage held by this CTA
    constexpr int B_ST = B_ROWS * BK * 2;                // ring slot bytes (16 KiB / 32 KiB)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smA = smem;
    uint8_t* smB = smA + nabuf * kPlanes * nkb * A_KB_BYTES;
    const int num_nb = (K + BN - 1) / BN;
    float* smN = reinterpret_cast<float*>(smB + nstages * B_ST);
    uint8_t* smP = reinterpret_cast<uint8_t*>(smN) + (kNormSmem ? num_nb * BN * 4 : 0);   // kMode 2: parked groups
    uint8_t* smO = smP + (kMode >= 2 ? 2 * kEpiWarps * 32 * 16 : 0);     // kMode 3: constant [128 x 16] "ones" operand (4 KiB)
    Ctl* ctl = reinterpret_cast<Ctl*>(smO + (kMode == 3 ? 4096 : 0));
    const uint8_t* aug_img = reinterpret_cast<const uint8_t*>(norms + num_nb * BN);   // behind the K_pad norms

    pdl_wait();                   // the codebook shadow this kernel reads was written by the previous step's EMA kernel
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = kPair ? cluster_ctarank() : 0u;                 // 0 = leader (MMA issuer) of the pair
    const int64_t unit = kPair ? (blockIdx.x >> 1) : blockIdx.x;          // slab-owning unit: a CTA or a CTA pair
    const int64_t nunits = kPair ? (gridDim.x >> 1) : gridDim.x;
    constexpr int SLAB = kPair ? 2 * BM : BM;                             // rows per slab of the unit
    const int64_t ntiles = (N + SLAB - 1) / SLAB;
    constexpr int npl = kPlanes;                                            // 2: split fp32 tier
    const int a_bytes = npl * nkb * A_KB_BYTES;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_z);
        prefetch_tmap(&tm_e);
        if (kPair && kMode == 3) prefetch_tmap(&tm_aug);
    }
    if (warp == 1 && lane == 0) {
        for (int b = 0; b < 2; ++b) {
            mbar_init(&ctl->a_full[b], 1);
            mbar_init(&ctl->a_empty[b], 1);
            mbar_init(&ctl->t_full[b], 1);
            // pair: both CTAs' epilogues release it (dbg 4: timing experiment, the peer's epilogue stays silent)
            mbar_init(&ctl->t_empty[b], (kPair && dbg != 4) ? 2 * kEpiWarps : kEpiWarps);
        }
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&ctl->b_full[s], 1);
            mbar_init(&ctl->b_empty[s], 1);
        }
        fence_mbar_init();
    }
    if (warp == 2) {
        if (kPair) tmem_alloc_pair<TMEM_COLS>(&ctl->tmem_base);
        else tmem_alloc<TMEM_COLS>(&ctl->tmem_base);
    }
    if (kNormSmem && warp >= 4) {          // ||e||^2 (padded with +inf to a multiple of BN) -> shared memory, once
        const float4* src = reinterpret_cast<const float4*>(norms);
        float4* dst = reinterpret_cast<float4*>(smN);
        for (int i = threadIdx.x - 128; i < num_nb * (BN / 4); i += kEpiWarps * 32) dst[i] = __ldg(src + i);
    }
    if (kMode == 3 && threadIdx.x >= 128 && threadIdx.x < 256) {
        // A-side partner of the aug image: every row is (c1, c2, c3, 0, ..., 0), canonical no-swizzle K-major layout
        const int r = threadIdx.x - 128;
        uint8_t* p = smO + (r >> 3) * 256 + (r & 7) * 16;
        *reinterpret_cast<uint4*>(p) = make_uint4(ok.ones01, ok.ones2, 0u, 0u);
        *reinterpret_cast<uint4*>(p + 128) = make_uint4(0u, 0u, 0u, 0u);
        fence_proxy_async();               // generic-proxy writes -> visible to the tensor core's async proxy
    }
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync_all();         // the peer's barriers are initialised before anything remote touches them
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (elect_one()) {
            // slab number j of this CTA lives in A buffer (j % nabuf); it is that buffer's (j / nabuf)-th use
            // pair: loads land in THIS CTA's shared memory and complete on the LEADER's "full" barriers, which expect the
            // bytes of both CTAs; each producer throttles on its own copy of the "empty" barriers (multicast commits)
            const uint32_t lead_a_full = kPair ? map_to_cta(smem_u32(&ctl->a_full[0]), 0) : 0u;     // [buf]: + 8 * buf
            const uint32_t lead_b_full = kPair ? map_to_cta(smem_u32(&ctl->b_full[0]), 0) : 0u;     // [st]:  + 8 * st
            auto load_a = [&](uint32_t j, int64_t tile) {
                const uint32_t buf = nabuf == 2 ? (j & 1u) : 0u;
                const uint32_t use = nabuf == 2 ? (j >> 1) : j;
                mbar_wait(&ctl->a_empty[buf], (use & 1u) ^ 1u, 1);
                if (rank == 0) mbar_arrive_expect_tx(&ctl->a_full[buf], kPair ? 2 * a_bytes : a_bytes);
                const int row0 = (int)(tile * SLAB + rank * BM);
                for (int pa = 0; pa < npl; ++pa)
                    for (int kb = 0; kb < nkb; ++kb) {
                        uint8_t* dst = smA + buf * a_bytes + (pa * nkb + kb) * A_KB_BYTES;
                        const int r = row0 + pa * ok.a_plane_rows;
                        if (kPair) tma_load_2d_hint_pair(dst, &tm_z, lead_a_full + 8u * buf, kb * BK, r, kEvictFirst);
                        else tma_load_2d_hint(dst, &tm_z, &ctl->a_full[buf], kb * BK, r, kEvictFirst);
                    }
            };
            uint32_t st = 0, ph = 0, j = 0;
            const int loads_per_slab = num_nb * (nkb * npl + (kMode == 3 ? 1 : 0));
            const int trigger = loads_per_slab < nstages ? loads_per_slab : nstages;
            if (nabuf == 2 && unit < ntiles) load_a(0, unit);
            for (int64_t tile = unit; tile < ntiles; tile += nunits, ++j) {
                if (nabuf == 1) load_a(j, tile);
                const int64_t next = tile + nunits;
                int issued = 0;
                for (int nb = 0; nb < num_nb; ++nb) {
                    for (int kb = 0; kb < nkb; ++kb)
                      for (int pb = 0; pb < npl; ++pb) {       // split tier: the e_hi stage, then the e_lo stage
                        mbar_wait(&ctl->b_empty[st], ph ^ 1, 2);
                        if (rank == 0) mbar_arrive_expect_tx(&ctl->b_full[st], B_ST_BYTES);   // 32 KiB per stage either way
                        const int erow = nb * BN + pb * ok.b_plane_rows;
                        if (kPair)
                            tma_load_2d_hint_pair(smB + st * B_ST, &tm_e, lead_b_full + 8u * st, kb * BK,
                                                  erow + (int)rank * B_ROWS, kEvictLast);
                        else
                            tma_load_2d_hint(smB + st * B_ST, &tm_e, &ctl->b_full[st], kb * BK, erow, kEvictLast);
                        if (++st == (uint32_t)nstages) { st = 0; ph ^= 1; }
                        // next slab's rows: requested once this slab's first ring-full of codebook stages is in
                        // flight, i.e. about when the previous slab retires and frees the other A buffer
                        if (++issued == trigger && nabuf == 2 && next < ntiles) load_a(j + 1, next);
                      }
                    if (kMode == 3) {       // one more ring stage per tile: the 8 KiB aug block of these 256 codes
                        mbar_wait(&ctl->b_empty[st], ph ^ 1, 2);
                        if (rank == 0) mbar_arrive_expect_tx(&ctl->b_full[st], 8192);
                        if (kPair)          // u64 view [K_pad / 64][256]: this CTA's 128 codes are 2 rows of 2 KiB
                            tma_load_2d_hint_pair(smB + st * B_ST, &tm_aug, lead_b_full + 8u * st, 0,
                                                  nb * 4 + (int)rank * 2, kEvictLast);
                        else
                            bulk_load_1d(smB + st * B_ST, aug_img + (size_t)nb * 8192, 8192, &ctl->b_full[st]);
                        if (++st == (uint32_t)nstages) { st = 0; ph ^= 1; }
                        if (++issued == trigger && nabuf == 2 && next < ntiles) load_a(j + 1, next);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (pair: the leader CTA only)
        if (rank == 0 && elect_one()) {
            const uint32_t idesc = ok.idesc;
            const uint32_t a_addr = smem_u32(smA), b_addr = smem_u32(smB);
            auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t acc_flag) {
                if (kPair) umma_bf16_ss_pair(d, da, db, idesc, acc_flag);
                else umma_bf16_ss(d, da, db, idesc, acc_flag);
            };
            auto commit = [&](uint64_t* bar) {
                if (kPair) umma_commit_pair(bar);
                else umma_commit(bar);
            };
            uint32_t st = 0, ph = 0, it = 0, j = 0;
            for (int64_t tile = unit; tile < ntiles; tile += nunits, ++j) {
                const uint32_t buf = nabuf == 2 ? (j & 1u) : 0u;
                const uint32_t use = nabuf == 2 ? (j >> 1) : j;
                mbar_wait(&ctl->a_full[buf], use & 1u, 3);
                tc_fence_after();
                const uint32_t a_base = a_addr + buf * a_bytes;
                for (int nb = 0; nb < num_nb; ++nb, ++it) {
                    const uint32_t acc = it & 1, acc_ph = (it >> 1) & 1;
                    mbar_wait(&ctl->t_empty[acc], acc_ph ^ 1, 4);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc * BN;
                    for (int kb = 0; kb < nkb; ++kb) {
                        const int rem = D - kb * BK;
                        const int ksteps = (kNKB != 0 || rem >= BK) ? BK / UK : (rem + UK - 1) / UK;
                        for (int pb = 0; pb < npl; ++pb) {
                            mbar_wait(&ctl->b_full[st], ph, 5);
                            tc_fence_after();
                            // split tier: the e_hi stage meets z_hi and z_lo, the e_lo stage z_hi only (z_lo.e_lo is
                            // below the tier's 2^-16 error budget)
                            const int npa = pb == 0 ? npl : 1;
                            for (int pa = 0; pa < npa; ++pa)
                                for (int k = 0; k < ksteps; ++k) {
                                    const uint64_t da = make_sdesc_sw128(a_base + (pa * nkb + kb) * A_KB_BYTES + k * (UK * 2));
                                    const uint64_t db = make_sdesc_sw128(b_addr + st * B_ST + k * (UK * 2));
                                    mma(d_tmem, da, db, (kb | k | pb | pa) != 0 ? 1u : 0u);
                                }
                            commit(&ctl->b_empty[st]);         // stage reusable once these MMAs retire
                            if (++st == (uint32_t)nstages) { st = 0; ph ^= 1; }
                        }
                    }
                    if (kMode == 3) {       // + [1,1,1,0..] x [h1,h2,h3,0..]^T = -||e||^2/2 into the same accumulator
                        mbar_wait(&ctl->b_full[st], ph, 5);
                        tc_fence_after();
                        mma(d_tmem, make_sdesc_none(smem_u32(smO), 128, 256), make_sdesc_none(b_addr + st * B_ST, 128, 256), 1u);
                        commit(&ctl->b_empty[st]);
                        if (++st == (uint32_t)nstages) { st = 0; ph ^= 1; }
                    }
                    commit(&ctl->t_full[acc]);                 // accumulator tile complete
                }
                commit(&ctl->a_empty[buf]);                    // slab of z no longer read
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue: running argmin
        constexpr int PARTS = kEpiWarps / 4;          // column parts of a tile
        constexpr int CW = BN / PARTS;                // columns per warp per tile (128 or 64)
        const int ew = warp - 4;
        const int q = ew & 3;          // TMEM lane quarter == warp % 4
        const int part = ew >> 2;
        const uint32_t nsm = smem_u32(smN);
        const uint32_t slot = smem_u32(smP) + (threadIdx.x - 128) * 16u;     // [2][epilogue threads] x 16 B: conflict-free
        constexpr uint32_t slot_stride = kEpiWarps * 32 * 16;
        const uint32_t lead_t_empty = kPair ? map_to_cta(smem_u32(&ctl->t_empty[0]), 0) : 0u;       // [acc]: + 8 * acc
        auto release_acc = [&](uint32_t acc) {      // this warp has read everything it needs from accumulator `acc`
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (kPair) mbar_arrive_cluster(lead_t_empty + 8u * acc);
                else mbar_arrive(&ctl->t_empty[acc]);
            }
        };
        // VQ_SEARCH_PRESORT: the histogram claim of a slab's rows is issued when the slab ends and consumed one slab later,
        // so the atomic's round trip to L2 never stalls the fold of the next accumulators
        uint32_t it = 0, xb = 0, slab_no = 0;
        for (int64_t tile = unit; tile < ntiles; tile += nunits, xb ^= 1, ++slab_no) {
            float best = kMode == 3 ? -CUDART_INF_F : CUDART_INF_F;      // kMode 3 tracks a maximum
            int bi = 0;
            for (int nb = 0; nb < num_nb; ++nb, ++it) {
                const uint32_t acc = it & 1, acc_ph = (it >> 1) & 1;
                const int col0 = nb * BN + part * CW;
                const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + acc * BN + part * CW;
                uint32_t ra[32], rb[32];
                mbar_wait(&ctl->t_full[acc], acc_ph, 6);
                tc_fence_after();
                if (dbg == 2 || dbg == 4) {           // timing experiment: no TMEM read, no fold
                    if (dbg == 2 || rank == 0) release_acc(acc);
                    continue;
                }
                tmem_ld_32x32b_x32(taddr, ra);
                tmem_ld_32x32b_x32(taddr + 32, rb);
                tmem_wait_ld();
#pragma unroll
                for (int c = 0; c + 2 < CW / 32; c += 2) {        // CW = 128 / 256: more chunk pairs before the last one
                    if (kMode == 3) fold_chunk_aug(ra, col0 + 32 * c, best, bi, slot, slot_stride);
                    else fold_chunk<kNormSmem, kMode>(ra, norms, nsm, col0 + 32 * c, best, bi, slot, slot_stride, dbg);
                    tmem_ld_32x32b_x32(taddr + 32 * (c + 2), ra);    // in flight while the second chunk is folded
                    if (kMode == 3) fold_chunk_aug(rb, col0 + 32 * (c + 1), best, bi, slot, slot_stride);
                    else fold_chunk<kNormSmem, kMode>(rb, norms, nsm, col0 + 32 * (c + 1), best, bi, slot, slot_stride, dbg);
                    tmem_ld_32x32b_x32(taddr + 32 * (c + 3), rb);
                    tmem_wait_ld();
                }
                // every TMEM read of this accumulator has landed: hand it back before folding what is in registers
                release_acc(acc);
                if (dbg == 1) {                       // timing experiment: TMEM read but no fold
                    bi += (int)(ra[0] ^ rb[5] ^ ra[31] ^ rb[17]);
                    continue;
                }
                if (kMode == 3) fold_chunk_aug(ra, col0 + CW - 64, best, bi, slot, slot_stride);
                    else fold_chunk<kNormSmem, kMode>(ra, norms, nsm, col0 + CW - 64, best, bi, slot, slot_stride, dbg);
                if (kMode == 3) fold_chunk_aug(rb, col0 + CW - 32, best, bi, slot, slot_stride);
                    else fold_chunk<kNormSmem, kMode>(rb, norms, nsm, col0 + CW - 32, best, bi, slot, slot_stride, dbg);
            }
            if (kMode == 2) bi = finish_deferred(best, bi, slot, slot_stride, best < CUDART_INF_F);
            if (kMode == 3) {
                bi = finish_deferred(best, bi, slot, slot_stride, best > -CUDART_INF_F);
                best = -2.f * best;           // back to the score scale ||e||^2 - 2 z.e (exact, order-reversing)
            }
            const int rl = q * 32 + lane;
            if (part != 0) {
                ctl->xbest[xb][part][rl] = best;
                ctl->xidx[xb][part][rl] = bi;
            }
            named_bar_sync(1, kEpiWarps * 32);
            if (part == 0) {
#pragma unroll
                for (int p = 1; p < PARTS; ++p) {     // ascending column order: strict < keeps the lowest index
                    const float ob = ctl->xbest[xb][p][rl];
                    const int oi = ctl->xidx[xb][p][rl];
                    if (ob < best) { best = ob; bi = oi; }
                }
                const int64_t row = tile * SLAB + rank * BM + rl;
                if (row < N) {
                    idx[row] = bi;
                    if (smin != nullptr) smin[row] = best;
                }
                if (ps.hist != nullptr) {
                    // hand the slab's indices to warp 3 (histogram claims + ranks): issued from here, the 32-address
                    // atomics cost the search 5 % at D=128 although nothing waits for their results
                    if (slab_no >= 2) named_bar_sync(4 + xb, 160);        // warp 3 has read this buffer's previous slab
                    ctl->fin_idx[xb][rl] = row < N ? bi : -1;
                    __threadfence_block();
                    named_bar_arrive(2 + xb, 160);
                }
            }
        }
    } else if (warp == 3 && ps.hist != nullptr) {
        // ------------------------------------------------------------------ VQ_SEARCH_PRESORT: claims and ranks
        // Warp-aggregated claims: one atomic per distinct code among 32 rows.  With trained (Zipf-like) code usage a
        // fifth of all rows can share one code, and un-aggregated claims on its counter serialise in L2 for longer than
        // the whole search takes (C2, trained-like latents: 270 -> 405 us); MATCH costs this warp ~0.35 us per 32 rows,
        // which it has (nothing else waits for it; on the epilogue warps the same instruction cost the kernel 5 %).
        uint32_t xb = 0;
        for (int64_t tile = unit; tile < ntiles; tile += nunits, xb ^= 1) {
            named_bar_sync(2 + xb, 160);
            int c[4], base[4], leader[4], off[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) c[j] = ctl->fin_idx[xb][j * 32 + lane];
            __threadfence_block();
            named_bar_arrive(4 + xb, 160);
#pragma unroll
            for (int j = 0; j < 4; ++j) presort_claim(ps.hist, c[j], c[j] >= 0, lane, base[j], leader[j], off[j]);
            const int64_t row0 = tile * SLAB + rank * BM + lane;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int r = presort_rank(base[j], leader[j], off[j]);
                if (c[j] >= 0) ps.rank[row0 + j * 32] = r;
            }
        }
    }

#ifndef VQ_PRESORT_NO_FINISH
    pdl_trigger();                // this CTA's rows are done: the next kernel's launch may begin (it still waits for us)
    if (ps.hist != nullptr) presort_finish(ps, K, ctl->scan_sh);
#endif

    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync_all();         // neither CTA leaves (or frees TMEM) while the other can still signal it
    if (warp == 2) {
        if (kPair) tmem_dealloc_pair<TMEM_COLS>(tmem_base);
        else tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------- host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// [rows, D] 16-bit elements, row stride ld elements; box = 64 features x box_rows rows, 128-byte swizzle, zero OOB fill.
int make_tmap(CUtensorMap* m, const void* base, int64_t rows, int D, int64_t ld, int box_rows, bool fp16 = false) {
    EncodeTiledFn enc = get_encode_fn();
    if (enc == nullptr) {
        set_error("cuTensorMapEncodeTiled entry point not available");
        return VQ_ERR_CUDA;
    }
    cuuint64_t gdim[2] = {(cuuint64_t)D, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, fp16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                     const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d) rows=%lld D=%d ld=%lld box_rows=%d", (int)r,
                  (long long)rows, D, (long long)ld, box_rows);
        return VQ_ERR_CUDA;
    }
    return VQ_OK;
}

// The aug image behind the norms, viewed as [K_pad / 64 rows][256 x u64] (2 KiB rows, no swizzle): a box of 2 rows is the
// 4 KiB block of 128 consecutive codes (pair mode loads it through TMA so that completion can go to the leader's barrier).
int make_tmap_aug(CUtensorMap* m, const void* base, int K_pad) {
    EncodeTiledFn enc = get_encode_fn();
    if (enc == nullptr) {
        set_error("cuTensorMapEncodeTiled entry point not available");
        return VQ_ERR_CUDA;
    }
    cuuint64_t gdim[2] = {256, (cuuint64_t)(K_pad / 64)};
    cuuint64_t gstr[1] = {2048};
    cuuint32_t box[2] = {256, 2};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(aug) failed (CUresult %d) K_pad=%d", (int)r, K_pad);
        return VQ_ERR_CUDA;
    }
    return VQ_OK;
}

}  // namespace

bool assign_tc_supported(int64_t N, int D, int64_t ldz, int K) {
    return N > 0 && N < (int64_t(1) << 31) - BM && K > 0 && D >= 8 && D % 8 == 0 && D <= 512 && ldz % 8 == 0;
}
// split fp32 tier: two planes of the slab must leave room for the ring (D <= 256), plane offsets must fit a TMA coordinate
bool assign_tc_split_supported(int64_t N, int D, int K) {
    return N > 0 && N < (int64_t(1) << 30) - BM && K > 0 && K < (1 << 29) && D >= 8 && D % 8 == 0 && D <= 256;
}

// kind: VQ_SEARCH_BF16 (bf16 z, bf16 shadow), VQ_SEARCH_F16 (fp16 z, fp16 shadow), VQ_SEARCH_F32 (split tier: z16 is the
// staged [2][N][D] bf16 hi / lo image of z with ldz == D, E16 the [2][K_pad][D] hi / lo shadow, norms of the fp32 rows)
int launch_assign_tc(const void* z_bf16, int64_t N, int D, int64_t ldz, const void* E_bf16, const float* e_norm2,
                     int K, int kind, int32_t* idx, float* smin, Presort ps, cudaStream_t st) {
    const bool split = kind == VQ_SEARCH_F32, fp16 = kind == VQ_SEARCH_F16;
    const int planes = split ? 2 : 1;
    VQ_REQUIRE(assign_tc_supported(N, D, ldz, K) && (!split || (assign_tc_split_supported(N, D, K) && ldz == D)),
               "vq_assign(tensor core): unsupported shape N=%lld D=%d ldz=%lld K=%d", (long long)N, D, (long long)ldz, K);
    VQ_REQUIRE((reinterpret_cast<uintptr_t>(z_bf16) & 15) == 0 && (reinterpret_cast<uintptr_t>(E_bf16) & 15) == 0,
               "vq_assign(tensor core): z and the codebook shadow must be 16-byte aligned");
    const int nkb = (D + BK - 1) / BK;
    const int num_nb = (K + BN - 1) / BN;
    // The product build reads nothing from the environment: epilogue warps, epilogue mode and pairing follow the shape.
    // Only the tuning build (build.py --variant, -DVQ_TUNING) has the knobs below, including timing experiments whose
    // results are WRONG.
    int env_epi = 0, mode_override = -1, pair_override = -1;
#ifdef VQ_TUNING
    static int t_epi = -1, dbg = 0, t_mode = -1, t_pair = -1;
    if (t_epi < 0) {
        const char* ev = getenv("VQ_ASSIGN_EPI_WARPS");          // 4, 8 or 16 epilogue warps
        t_epi = (ev != nullptr && (atoi(ev) == 16 || atoi(ev) == 8 || atoi(ev) == 4)) ? atoi(ev) : 0;
        ev = getenv("VQ_ASSIGN_DBG");                            // 1..4: timing experiments, results are WRONG
        if (ev != nullptr) dbg = atoi(ev);
        ev = getenv("VQ_ASSIGN_MODE");                           // force an epilogue mode (0..3)
        if (ev != nullptr) t_mode = atoi(ev);
        ev = getenv("VQ_ASSIGN_PAIR");                           // 0: single-CTA kernel, 1: CTA-pair kernel
        if (ev != nullptr) t_pair = atoi(ev);
    }
    env_epi = t_epi; mode_override = t_mode; pair_override = t_pair;
#endif
    // CTA pairs (cta_group::2) whenever there is more than one 128-row slab: a pair retires 256 rows in the time one CTA
    // retires 128, so the load balance over the SMs is the same and each SM pulls half the codebook bytes.
    static bool pair_unavailable = false;            // set once if this device / context cannot co-schedule CTA pairs
    bool pair = N > BM && !pair_unavailable;
    if (pair_override == 0) pair = false;
    if (pair_override == 1) pair = true;
  for (int attempt = 0; attempt < 2; ++attempt) {    // second round only if the pair launch turned out to be impossible
    // epilogue warps (measured on B200): pairs 8 (C4 1409 us / C2 277 us; 16: 1427 / 292; 4: 1471 / 323), single CTA 16
    // ... except the D = 256, K >= 4096 shape (epilogue mode 1) on the final kernel: 16 warps 1284 us vs 1320 us at C4
    // (profiles/r02/search_split2.log); at D = 128 both read 258 us
    const bool wide_epi = pair && planes == 1 && D == 256 && K >= 4096 && mode_override < 0;
    const int epi_warps = env_epi != 0 ? env_epi : (pair ? (wide_epi ? 16 : 8) : 16);
    // Epilogue mode (see fold_chunk / fold_chunk_aug).  The epilogue, not the tensor pipe, bounds this kernel when
    // D is small (a 128 x 256 tile costs D/16 MMAs but always 32768 compare-selects), so:
    //   D <= 128          mode 3: ||e||^2 rides through the tensor core as one more K step, arg-max of the raw accumulator,
    //                     winning group parked in shared memory
    //   D > 128, K < 4096 mode 2: FFMA2 scores, winning group parked
    //   D > 128, K >=4096 mode 1: FFMA2 scores, immediate search inside the group of 8
    //   measured on B200 (C2: K=1024 D=128): mode 0 0.365 ms, 1 0.386, 2 0.347; (C4: K=8192 D=256): mode 0 1.53 ms, 1 1.52
    int mode = D <= 128 ? 3 : (K >= 4096 ? 1 : 2);
    if (mode_override >= 0 && mode_override <= 3) mode = mode_override;
    // shared-memory plan: [A slab x nabuf][codebook ring x nstages][||e||^2 (modes 0-2, optional)][parked groups
    // (modes 2, 3)][ones operand (mode 3)][Ctl]
    //   norms go to shared memory when K <= 8192 and that still leaves >= 3 ring stages;
    //   the slab is double-buffered when that still leaves >= 4 ring stages.
    const int park_bytes = 2 * epi_warps * 32 * 16;
    const int a_bytes = planes * nkb * A_KB_BYTES;
    const int norm_bytes = num_nb * BN * 4;
    const int b_st = pair ? B_ST_BYTES / 2 : B_ST_BYTES;           // a pair CTA holds half of every codebook stage
    const int max_stages = pair ? MAX_STAGES : 6;
    int avail = 0, nabuf = 1, nstages = 0, extra = 0;
    bool norm_smem = false;
    for (;;) {
        extra = (mode >= 2 ? park_bytes : 0) + (mode == 3 ? 4096 : 0);
        avail = SMEM_LIMIT - 1024 - (int)sizeof(Ctl) - extra;
        norm_smem = mode != 3 && K <= NORM_SMEM_MAX && (avail - a_bytes - norm_bytes) / b_st >= 3;
        const int after_norm = avail - (norm_smem ? norm_bytes : 0);
        nabuf = (after_norm - 2 * a_bytes) / b_st >= 4 ? 2 : 1;
        nstages = (after_norm - nabuf * a_bytes) / b_st;
        if (nstages > max_stages) nstages = max_stages;
        if (nstages >= 3 || mode == 0) break;
        mode = mode == 3 ? 2 : 0;                      // not enough room for the ring: fall back to a leaner mode
    }
    VQ_REQUIRE(nstages >= 2, "vq_assign(tensor core): D=%d leaves no room for the codebook ring", D);
    const size_t smem_total = 1024 + (size_t)nabuf * a_bytes + (size_t)nstages * b_st +
                              (norm_smem ? norm_bytes : 0) + extra + sizeof(Ctl);

    CUtensorMap tm_z, tm_e, tm_aug;
    const int K_pad = num_nb * BN;
    int rc = make_tmap(&tm_z, z_bf16, planes * N, D, ldz, BM, fp16);
    if (rc != VQ_OK) return rc;
    rc = make_tmap(&tm_e, E_bf16, split ? (int64_t)2 * K_pad : K, D, D, pair ? BN / 2 : BN, fp16);
    if (rc != VQ_OK) return rc;
    OperandKind ok;
    // [4,6) D = f32; [7,10) A format, [10,13) B format: 0 = f16, 1 = bf16; K-major both; [17,23) N >> 3; [24,29) M >> 4
    ok.idesc = (1u << 4) | (fp16 ? 0u : (1u << 7) | (1u << 10)) | (static_cast<uint32_t>(BN >> 3) << 17) |
               (static_cast<uint32_t>((pair ? 2 * BM : BM) >> 4) << 24);
    // A-side partner of the norm image (vq_ema.cu: write_aug): bf16 pieces sum with weights (1, 1, 1); fp16 pieces carry
    // the weights (2^8, 2^-4, 2^-14) so that -||e||^2/2 keeps 24 significant bits from 1e-11 up to 1.6e7
    ok.ones01 = fp16 ? 0x2C005C00u : 0x3F803F80u;
    ok.ones2 = fp16 ? 0x00000400u : 0x00003F80u;
    ok.planes = planes;
    ok.a_plane_rows = (int)N;
    ok.b_plane_rows = K_pad;
    if (pair && mode == 3) {
        rc = make_tmap_aug(&tm_aug, e_norm2 + num_nb * BN, num_nb * BN);
        if (rc != VQ_OK) return rc;
    } else {
        tm_aug = tm_e;                                 // never dereferenced
    }

#ifdef VQ_TUNING
    void (*kern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const float*, int64_t, int, int, int, int, int,
                 const OperandKind, int32_t*, float*, int, const Presort);
#else
    void (*kern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const float*, int64_t, int, int, int, int, int,
                 const OperandKind, int32_t*, float*, const Presort);
#endif
#define VQ_PICK4(NS, EW, P, PL)                                                                                  \
    (mode == 3 ? assign_tc_kernel<false, EW, 3, P, PL, 0> : mode == 2 ? assign_tc_kernel<NS, EW, 2, P, PL, 0>       \
                                                        : mode == 1 ? assign_tc_kernel<NS, EW, 1, P, PL, 0> : assign_tc_kernel<NS, EW, 0, P, PL, 0>)
#define VQ_PICK3(NS, EW, P) (planes == 2 ? VQ_PICK4(NS, EW, P, 2) : VQ_PICK4(NS, EW, P, 1))
#define VQ_PICK(NS, EW) (pair ? VQ_PICK3(NS, EW, true) : VQ_PICK3(NS, EW, false))
#ifdef VQ_TUNING                // the epilogue-warp knob exists for the one-plane kernels only
    if (planes == 2) {
        if (norm_smem) kern = pair ? VQ_PICK4(true, 8, true, 2) : VQ_PICK4(true, 16, false, 2);
        else kern = pair ? VQ_PICK4(false, 8, true, 2) : VQ_PICK4(false, 16, false, 2);
    } else if (norm_smem) {
        kern = epi_warps == 4 ? (pair ? VQ_PICK4(true, 4, true, 1) : VQ_PICK4(true, 4, false, 1))
             : epi_warps == 8 ? (pair ? VQ_PICK4(true, 8, true, 1) : VQ_PICK4(true, 8, false, 1))
                              : (pair ? VQ_PICK4(true, 16, true, 1) : VQ_PICK4(true, 16, false, 1));
    } else {
        kern = epi_warps == 4 ? (pair ? VQ_PICK4(false, 4, true, 1) : VQ_PICK4(false, 4, false, 1))
             : epi_warps == 8 ? (pair ? VQ_PICK4(false, 8, true, 1) : VQ_PICK4(false, 8, false, 1))
                              : (pair ? VQ_PICK4(false, 16, true, 1) : VQ_PICK4(false, 16, false, 1));
    }
#else       // product build: pairs run 8 epilogue warps, single CTAs 16 (28 instantiations: 14 per plane count)
    if (norm_smem) kern = pair ? VQ_PICK3(true, 8, true) : VQ_PICK3(true, 16, false);
    else kern = pair ? VQ_PICK3(false, 8, true) : VQ_PICK3(false, 16, false);
#endif
    // the BASELINE shapes (D = 64, 128, 256 on CTA pairs, one plane, the epilogue mode the shape picks by itself) run the
    // instantiations with a compile-time K-block count
    if (wide_epi && epi_warps == 16) {
        VQ_REQUIRE(mode == 1, "vq_assign(tensor core): internal: epilogue mode %d with the 16-warp plan", mode);
        kern = norm_smem ? assign_tc_kernel<true, 16, 1, true, 1, 4> : assign_tc_kernel<false, 16, 1, true, 1, 4>;
    } else if (pair && planes == 1 && epi_warps == 8 && D % BK == 0) {
        if (D == 64 && mode == 3) kern = assign_tc_kernel<false, 8, 3, true, 1, 1>;
        else if (D == 128 && mode == 3) kern = assign_tc_kernel<false, 8, 3, true, 1, 2>;
        else if (D == 256 && mode == 1) kern = norm_smem ? assign_tc_kernel<true, 8, 1, true, 1, 4> : assign_tc_kernel<false, 8, 1, true, 1, 4>;
        else if (D == 256 && mode == 2) kern = norm_smem ? assign_tc_kernel<true, 8, 2, true, 1, 4> : assign_tc_kernel<false, 8, 2, true, 1, 4>;
    }
#ifdef VQ_TUNING     // the epilogue-warp knob on the specialised kernels too (4 / 16 warps), so that it compares like with like
    if (pair && planes == 1 && epi_warps != 8 && D % BK == 0) {
        if (D == 128 && mode == 3)
            kern = epi_warps == 4 ? assign_tc_kernel<false, 4, 3, true, 1, 2> : assign_tc_kernel<false, 16, 3, true, 1, 2>;
        else if (D == 256 && mode == 1 && norm_smem)
            kern = epi_warps == 4 ? assign_tc_kernel<true, 4, 1, true, 1, 4> : assign_tc_kernel<true, 16, 1, true, 1, 4>;
    }
#endif
#undef VQ_PICK
#undef VQ_PICK3
#undef VQ_PICK4
    VQ_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));

    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[2];
    cfg.blockDim = dim3(128 + 32 * epi_warps);
    cfg.dynamicSmemBytes = smem_total;
    cfg.stream = st;
    if (pair) {
        const int64_t nslabs = (N + 2 * BM - 1) / (2 * BM);
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        // persistent: one pair per TPC that can host one (a TPC with a fused-off SM cannot)
        static int max_pairs = 0;
        static const void* cached_kern = nullptr;
        static size_t cached_smem = 0;
        static int cached_dev = -1;
        const int cur_dev = current_device();
        if (max_pairs == 0 || cached_kern != (const void*)kern || cached_smem != smem_total || cached_dev != cur_dev) {
            cached_kern = (const void*)kern;
            cached_smem = smem_total;
            cached_dev = cur_dev;
            cfg.gridDim = dim3((unsigned)(sm_count() & ~1));
            int nc = 0;
            if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) != cudaSuccess || nc <= 0) {
                (void)cudaGetLastError();            // no room for a single pair (MPS slice, odd partition, ...): single-CTA kernel
                max_pairs = 0;
                pair_unavailable = true;
                pair = false;
                continue;
            }
            max_pairs = nc < sm_count() / 2 ? nc : sm_count() / 2;
        }
        cfg.gridDim = dim3((unsigned)(2 * (nslabs < max_pairs ? nslabs : max_pairs)));
    } else {
        const int64_t ntiles = (N + BM - 1) / BM;
        cfg.gridDim = dim3((unsigned)(ntiles < sm_count() ? ntiles : sm_count()));
    }
#ifdef VQ_TUNING
    static const bool verbose = getenv("VQ_ASSIGN_VERBOSE") != nullptr;
#else
    constexpr bool verbose = false;
#endif
    if (verbose)
        fprintf(stderr, "vq_assign: N=%lld K=%d D=%d pair=%d mode=%d nstages=%d nabuf=%d norm_smem=%d grid=%u smem=%zu\n",
                (long long)N, K, D, (int)pair, mode, nstages, nabuf, (int)norm_smem, cfg.gridDim.x, smem_total);
    {   // programmatic dependent launch: this grid may be set up while the previous kernel of the stream drains
        int na = pair ? 1 : 0;
        if (pdl_enabled(PDL_SEARCH)) {
            attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        cfg.attrs = attr;
        cfg.numAttrs = na;
    }
#ifdef VQ_TUNING
    const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, tm_z, tm_e, tm_aug, e_norm2, N, K, D, nkb, nstages, nabuf, ok, idx, smin, dbg, ps);
#else
    const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, tm_z, tm_e, tm_aug, e_norm2, N, K, D, nkb, nstages, nabuf, ok, idx, smin, ps);
#endif
    if (le != cudaSuccess && pair && attempt == 0) {  // cluster launch refused: fall back to the single-CTA kernel for good
        (void)cudaGetLastError();
        pair_unavailable = true;
        pair = false;
        continue;
    }
    VQ_CHECK_CUDA(le);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
  }
    set_error("vq_assign(tensor core): kernel launch failed");
    return VQ_ERR_CUDA;
}

}  // namespace vq
