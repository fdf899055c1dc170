// a-1  Nearest-codebook search on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
//   idx[n] = argmin_k ( ||e_k||^2 - 2 z_n . e_k )        z: [N, D] bf16,  E: [K, D] bf16,  ||e||^2: fp32
//
// The N x K score matrix never exists in HBM: each CTA owns a 128-row slab of z (kept resident in shared
// memory for the whole sweep), streams the codebook through a TMA ring in 256-code x 64-feature stages,
// accumulates S = z E^T for one 128 x 256 tile in TMEM (fp32), and the epilogue warps fold each finished tile
// into a running (min, argmin) per row straight out of TMEM while the tensor core works on the next tile
// (two accumulator buffers = all 512 TMEM columns).
//
// Warp roles (384 threads, 1 CTA / SM, persistent over row slabs):
//   warp 0      TMA producer (one elected lane)
//   warp 1      tcgen05.mma issuer (one elected lane)
//   warp 2      TMEM allocator / deallocator
//   warps 4-11  epilogue: warp w reads TMEM lanes 32*(w%4)..+31 (one row per thread), columns
//               [128*((w-4)/4), +128) of each tile; the two halves are merged through smem per slab.
//
// Required by: scripts/train_vqvae.py:11,287 (VectorQuantizer inside model(batch));
//              frl/config/frl_model_v0.yaml:29-33 (kind: vq, K, D).  Definition: SURVEY.md 8(a-1).
#include <math_constants.h>
#include <stdio.h>

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
constexpr int NUM_THREADS = 384;
constexpr int NUM_EPI_WARPS = 8;
constexpr int TMEM_COLS = 512;
constexpr int SMEM_LIMIT = 232448;  // 227 KiB opt-in maximum per CTA

struct Ctl {
    uint64_t a_full, a_empty;
    uint64_t b_full[MAX_STAGES], b_empty[MAX_STAGES];
    uint64_t t_full[2], t_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
    float xbest[2][BM];
    int xidx[2][BM];
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
assign_tc_kernel(const __grid_constant__ CUtensorMap tm_z, const __grid_constant__ CUtensorMap tm_e,
                 const float* __restrict__ norms, int64_t N, int K, int D, int nkb, int nstages,
                 int32_t* __restrict__ idx, float* __restrict__ smin) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smA = smem;
    uint8_t* smB = smA + nkb * A_KB_BYTES;
    Ctl* ctl = reinterpret_cast<Ctl*>(smB + nstages * B_ST_BYTES);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t ntiles = (N + BM - 1) / BM;
    const int num_nb = (K + BN - 1) / BN;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_z);
        prefetch_tmap(&tm_e);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(&ctl->a_full, 1);
        mbar_init(&ctl->a_empty, 1);
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&ctl->b_full[s], 1);
            mbar_init(&ctl->b_empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&ctl->t_full[b], 1);
            mbar_init(&ctl->t_empty[b], NUM_EPI_WARPS);
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc<TMEM_COLS>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (elect_one()) {
            uint32_t st = 0, ph = 0, a_ph = 0;
            for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                mbar_wait(&ctl->a_empty, a_ph ^ 1, 1);
                mbar_arrive_expect_tx(&ctl->a_full, nkb * A_KB_BYTES);
                for (int kb = 0; kb < nkb; ++kb)
                    tma_load_2d_hint(smA + kb * A_KB_BYTES, &tm_z, &ctl->a_full, kb * BK, (int)(tile * BM),
                                     kEvictFirst);
                a_ph ^= 1;
                for (int nb = 0; nb < num_nb; ++nb) {
                    for (int kb = 0; kb < nkb; ++kb) {
                        mbar_wait(&ctl->b_empty[st], ph ^ 1, 2);
                        mbar_arrive_expect_tx(&ctl->b_full[st], B_ST_BYTES);
                        tma_load_2d_hint(smB + st * B_ST_BYTES, &tm_e, &ctl->b_full[st], kb * BK, nb * BN,
                                         kEvictLast);
                        if (++st == (uint32_t)nstages) { st = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (elect_one()) {
            constexpr uint32_t idesc = make_idesc_bf16(BM, BN);
            const uint32_t a_addr = smem_u32(smA), b_addr = smem_u32(smB);
            uint32_t st = 0, ph = 0, a_ph = 0, it = 0;
            for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                mbar_wait(&ctl->a_full, a_ph, 3);
                a_ph ^= 1;
                tc_fence_after();
                for (int nb = 0; nb < num_nb; ++nb, ++it) {
                    const uint32_t acc = it & 1, acc_ph = (it >> 1) & 1;
                    mbar_wait(&ctl->t_empty[acc], acc_ph ^ 1, 4);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc * BN;
                    for (int kb = 0; kb < nkb; ++kb) {
                        mbar_wait(&ctl->b_full[st], ph, 5);
                        tc_fence_after();
                        const int rem = D - kb * BK;
                        const int ksteps = rem >= BK ? BK / UK : (rem + UK - 1) / UK;
                        for (int k = 0; k < ksteps; ++k) {
                            const uint64_t da = make_sdesc_sw128(a_addr + kb * A_KB_BYTES + k * (UK * 2));
                            const uint64_t db = make_sdesc_sw128(b_addr + st * B_ST_BYTES + k * (UK * 2));
                            umma_bf16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                        }
                        umma_commit(&ctl->b_empty[st]);        // stage reusable once these MMAs retire
                        if (++st == (uint32_t)nstages) { st = 0; ph ^= 1; }
                    }
                    umma_commit(&ctl->t_full[acc]);            // accumulator tile complete
                }
                umma_commit(&ctl->a_empty);                    // slab of z no longer read
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue: running argmin
        const int ew = warp - 4;
        const int q = ew & 3;          // TMEM lane quarter == warp % 4
        const int half = ew >> 2;      // which 128 columns of each 256-column tile
        uint32_t it = 0, xb = 0;
        for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, xb ^= 1) {
            float best[4];
            int bi[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) { best[c] = CUDART_INF_F; bi[c] = 0; }
            for (int nb = 0; nb < num_nb; ++nb, ++it) {
                const uint32_t acc = it & 1, acc_ph = (it >> 1) & 1;
                mbar_wait(&ctl->t_full[acc], acc_ph, 6);
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < 4; ++c) {
                    const int col0 = nb * BN + half * 128 + c * 32;
                    const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + acc * BN + half * 128 + c * 32;
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(taddr, r);
                    const float4* np = reinterpret_cast<const float4*>(norms + col0);
                    float4 nv[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) nv[j] = __ldg(np + j);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float s0 = fmaf(-2.f, __uint_as_float(r[4 * j + 0]), nv[j].x);
                        const float s1 = fmaf(-2.f, __uint_as_float(r[4 * j + 1]), nv[j].y);
                        const float s2 = fmaf(-2.f, __uint_as_float(r[4 * j + 2]), nv[j].z);
                        const float s3 = fmaf(-2.f, __uint_as_float(r[4 * j + 3]), nv[j].w);
                        const int cj = col0 + 4 * j;
                        if (s0 < best[0]) { best[0] = s0; bi[0] = cj; }
                        if (s1 < best[1]) { best[1] = s1; bi[1] = cj + 1; }
                        if (s2 < best[2]) { best[2] = s2; bi[2] = cj + 2; }
                        if (s3 < best[3]) { best[3] = s3; bi[3] = cj + 3; }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&ctl->t_empty[acc]);
            }
            float b = best[0];
            int i = bi[0];
#pragma unroll
            for (int c = 1; c < 4; ++c)
                if (best[c] < b || (best[c] == b && bi[c] < i)) { b = best[c]; i = bi[c]; }
            const int rl = q * 32 + lane;
            if (half == 1) {
                ctl->xbest[xb][rl] = b;
                ctl->xidx[xb][rl] = i;
            }
            named_bar_sync(1, NUM_EPI_WARPS * 32);
            if (half == 0) {
                const float ob = ctl->xbest[xb][rl];
                const int oi = ctl->xidx[xb][rl];
                if (ob < b || (ob == b && oi < i)) { b = ob; i = oi; }
                const int64_t row = tile * BM + rl;
                if (row < N) {
                    idx[row] = i;
                    if (smin != nullptr) smin[row] = b;
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
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

// [rows, D] bf16, row stride ld elements; box = 64 features x box_rows rows, 128-byte swizzle, zero OOB fill.
int make_tmap(CUtensorMap* m, const void* base, int64_t rows, int D, int64_t ld, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    if (enc == nullptr) {
        set_error("cuTensorMapEncodeTiled entry point not available");
        return VQ_ERR_CUDA;
    }
    cuuint64_t gdim[2] = {(cuuint64_t)D, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d) rows=%lld D=%d ld=%lld box_rows=%d", (int)r,
                  (long long)rows, D, (long long)ld, box_rows);
        return VQ_ERR_CUDA;
    }
    return VQ_OK;
}

}  // namespace

bool assign_tc_supported(int64_t N, int D, int64_t ldz, int K) {
    return N > 0 && N < (int64_t(1) << 31) - BM && K > 0 && D >= 8 && D % 8 == 0 && D <= 512 && ldz % 8 == 0;
}

int launch_assign_tc(const void* z_bf16, int64_t N, int D, int64_t ldz, const void* E_bf16, const float* e_norm2,
                     int K, int32_t* idx, float* smin, cudaStream_t st) {
    VQ_REQUIRE(assign_tc_supported(N, D, ldz, K), "vq_assign(bf16): unsupported shape N=%lld D=%d ldz=%lld K=%d",
               (long long)N, D, (long long)ldz, K);
    VQ_REQUIRE((reinterpret_cast<uintptr_t>(z_bf16) & 15) == 0 && (reinterpret_cast<uintptr_t>(E_bf16) & 15) == 0,
               "vq_assign(bf16): z and E_bf16 must be 16-byte aligned");
    const int nkb = (D + BK - 1) / BK;
    int nstages = (SMEM_LIMIT - 1024 - (int)sizeof(Ctl) - nkb * A_KB_BYTES) / B_ST_BYTES;
    if (nstages > 6) nstages = 6;
    VQ_REQUIRE(nstages >= 2, "vq_assign(bf16): D=%d leaves no room for the codebook ring", D);
    const size_t smem_bytes = 1024 + (size_t)nkb * A_KB_BYTES + (size_t)nstages * B_ST_BYTES + sizeof(Ctl);

    CUtensorMap tm_z, tm_e;
    int rc = make_tmap(&tm_z, z_bf16, N, D, ldz, BM);
    if (rc != VQ_OK) return rc;
    rc = make_tmap(&tm_e, E_bf16, K, D, D, BN);
    if (rc != VQ_OK) return rc;

    static bool attr_set = false;
    if (!attr_set) {
        VQ_CHECK_CUDA(cudaFuncSetAttribute(assign_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
        attr_set = true;
    }
    const int64_t ntiles = (N + BM - 1) / BM;
    const int grid = (int)(ntiles < sm_count() ? ntiles : sm_count());
    assign_tc_kernel<<<grid, NUM_THREADS, smem_bytes, st>>>(tm_z, tm_e, e_norm2, N, K, D, nkb, nstages, idx, smin);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

}  // namespace vq
