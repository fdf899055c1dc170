// a-1, fp32 tier: nearest-codebook search with fp32 FMA on the CUDA cores (VQ_SEARCH_F32).
// Same contract as the tensor-core kernel (lowest index wins ties, score = ||e||^2 - 2 z.e), any D and K,
// fp32 or bf16 latents, fp32 master codebook.  Classic 64x64 register-tiled SGEMM shape with the score
// tile folded into a per-row running argmin, so the N x K matrix is never stored either.
// Definition: SURVEY.md 8(a-1); required by scripts/train_vqvae.py:11,287.
#include <cuda_fp16.h>
#include <math_constants.h>

#include "vq_common.cuh"

namespace vq {
namespace {

constexpr int TM = 64, TN = 64, TK = 32, LDS_PAD = 68;

template <typename T>
__device__ __forceinline__ float load_as_float(const T* p);
template <>
__device__ __forceinline__ float load_as_float<float>(const float* p) { return __ldg(p); }
template <>
__device__ __forceinline__ float load_as_float<__nv_bfloat16>(const __nv_bfloat16* p) {
    return __bfloat162float(*p);
}

template <>
__device__ __forceinline__ float load_as_float<__half>(const __half* p) { return __half2float(*p); }

template <typename T>
__global__ void __launch_bounds__(256)
assign_simt_kernel(const T* __restrict__ z, int64_t N, int D, int64_t ldz, const float* __restrict__ E,
                   const float* __restrict__ norms, int K, int32_t* __restrict__ idx, float* __restrict__ smin,
                   const Presort ps) {
    __shared__ __align__(16) float zs[TK][LDS_PAD];
    __shared__ int scan_sh[36];
    __shared__ __align__(16) float es[TK][LDS_PAD];
    const int tid = threadIdx.x;
    const int ty = tid >> 4, tx = tid & 15;
    const int64_t row0 = (int64_t)blockIdx.x * TM;

    float best[4];
    int bi[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { best[i] = CUDART_INF_F; bi[i] = 0; }

    for (int c0 = 0; c0 < K; c0 += TN) {
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

        for (int k0 = 0; k0 < D; k0 += TK) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int lin = tid + i * 256;
                const int r = lin >> 5, k = lin & 31;
                const int64_t gr = row0 + r;
                const int gk = k0 + k;
                zs[k][r] = (gr < N && gk < D) ? load_as_float<T>(z + gr * ldz + gk) : 0.f;
                const int gc = c0 + r;
                es[k][r] = (gc < K && gk < D) ? __ldg(E + (int64_t)gc * D + gk) : 0.f;
            }
            __syncthreads();
#pragma unroll 8
            for (int k = 0; k < TK; ++k) {
                const float4 a = *reinterpret_cast<const float4*>(&zs[k][ty * 4]);
                const float4 b = *reinterpret_cast<const float4*>(&es[k][tx * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w};
                const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
            __syncthreads();
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int col = c0 + tx * 4 + j;
            const float nrm = norms[col];            // padded with +inf beyond K
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float s = fmaf(-2.f, acc[i][j], nrm);
                if (s < best[i]) { best[i] = s; bi[i] = col; }
            }
        }
    }
    // reduce over the 16 threads (half-warp) that share these rows: lexicographic (score, index) minimum
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float b = best[i];
        int bidx = bi[i];
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, b, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
            if (ob < b || (ob == b && oi < bidx)) { b = ob; bidx = oi; }
        }
        const int64_t gr = row0 + ty * 4 + i;
        if (tx == 0 && gr < N) {
            idx[gr] = bidx;
            if (smin != nullptr) smin[gr] = b;
            if (ps.hist != nullptr) ps.rank[gr] = atomicAdd(ps.hist + bidx, 1);
        }
    }
    if (ps.hist != nullptr) presort_finish(ps, K, scan_sh);
}

template <typename T>
__global__ void cast_bf16_kernel(const T* __restrict__ src, int64_t N, int D, int64_t ld,
                                 __nv_bfloat16* __restrict__ dst) {
    const int64_t total = N * (int64_t)D;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / D;
        const int c = (int)(i - r * D);
        dst[i] = __float2bfloat16_rn(load_as_float<T>(src + r * ld + c));
    }
}

__global__ void cast_bf16_vec_kernel(const float4* __restrict__ src, int64_t n4, uint2* __restrict__ dst) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = src[i];
        dst[i] = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
    }
}

// Split fp32 tier of the tensor-core search: x = hi + lo with hi = bf16(x), lo = bf16(x - hi) (16 significant bits; the
// subtraction is exact in fp32), written as two dense [N, D] bf16 planes, plane 1 behind plane 0.
template <typename T>
__global__ void split_bf16_kernel(const T* __restrict__ src, int64_t N, int D, int64_t ld,
                                  __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo) {
    const int64_t total = N * (int64_t)D;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / D;
        const int c = (int)(i - r * D);
        const float x = load_as_float<T>(src + r * ld + c);
        const __nv_bfloat16 h = __float2bfloat16_rn(x);
        hi[i] = h;
        lo[i] = __float2bfloat16_rn(x - __bfloat162float(h));
    }
}
__global__ void split_bf16_vec_kernel(const float4* __restrict__ src, int64_t n4, uint2* __restrict__ hi,
                                      uint2* __restrict__ lo) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = __ldcs(src + i);
        const uint2 h = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
        hi[i] = h;
        lo[i] = make_uint2(pack_bf16(v.x - bf16_lo(h.x), v.y - bf16_hi(h.x)), pack_bf16(v.z - bf16_lo(h.y), v.w - bf16_hi(h.y)));
    }
}

}  // namespace

int launch_split_bf16(const void* src, int in_dtype, int64_t N, int D, int64_t ld, void* dst_planes, cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && D > 0 && ld >= D, "vq_assign(split): bad shape");
    if (N == 0) return VQ_OK;
    const int64_t total = N * (int64_t)D;
    __nv_bfloat16* hi = static_cast<__nv_bfloat16*>(dst_planes);
    __nv_bfloat16* lo = hi + total;
    const int blocks = sm_count() * 8;
    if (in_dtype == VQ_DT_F32 && ld == D && total % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 &&
        (reinterpret_cast<uintptr_t>(dst_planes) & 7) == 0 && (total * 2) % 8 == 0)
        split_bf16_vec_kernel<<<blocks, 256, 0, st>>>(static_cast<const float4*>(src), total / 4,
                                                      reinterpret_cast<uint2*>(hi), reinterpret_cast<uint2*>(lo));
    else if (in_dtype == VQ_DT_F32)
        split_bf16_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(src), N, D, ld, hi, lo);
    else if (in_dtype == VQ_DT_F16)
        split_bf16_kernel<__half><<<blocks, 256, 0, st>>>(static_cast<const __half*>(src), N, D, ld, hi, lo);
    else
        split_bf16_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(src), N, D, ld, hi, lo);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

int launch_assign_simt(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, const float* E_f32,
                       const float* e_norm2, int K, int32_t* idx, float* smin, Presort ps, cudaStream_t st) {
    VQ_REQUIRE(N > 0 && D > 0 && K > 0 && ldz >= D, "vq_assign(f32): bad shape N=%lld D=%d K=%d ldz=%lld",
               (long long)N, D, K, (long long)ldz);
    const int64_t grid = (N + TM - 1) / TM;
    VQ_REQUIRE(grid < (int64_t(1) << 31), "vq_assign(f32): N too large");
    if (in_dtype == VQ_DT_F32)
        assign_simt_kernel<float><<<(unsigned)grid, 256, 0, st>>>(static_cast<const float*>(z), N, D, ldz, E_f32,
                                                                  e_norm2, K, idx, smin, ps);
    else if (in_dtype == VQ_DT_F16)
        assign_simt_kernel<__half><<<(unsigned)grid, 256, 0, st>>>(static_cast<const __half*>(z), N, D, ldz, E_f32,
                                                                   e_norm2, K, idx, smin, ps);
    else
        assign_simt_kernel<__nv_bfloat16><<<(unsigned)grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(z), N, D,
                                                                          ldz, E_f32, e_norm2, K, idx, smin, ps);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

int launch_cast_bf16(const void* src, int in_dtype, int64_t N, int D, int64_t ld, void* dst, cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && D > 0 && ld >= D, "vq_cast_bf16: bad shape");
    VQ_REQUIRE(in_dtype == VQ_DT_F32 || in_dtype == VQ_DT_F16, "vq_cast_bf16: source must be fp32 or fp16");
    if (N == 0) return VQ_OK;
    const int64_t total = N * (int64_t)D;
    const int blocks = sm_count() * 8;
    if (in_dtype == VQ_DT_F16)
        cast_bf16_kernel<__half><<<blocks, 256, 0, st>>>(static_cast<const __half*>(src), N, D, ld,
                                                         static_cast<__nv_bfloat16*>(dst));
    else if (ld == D && total % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 &&
        (reinterpret_cast<uintptr_t>(dst) & 7) == 0)
        cast_bf16_vec_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const float4*>(src), total / 4,
                                                     static_cast<uint2*>(dst));
    else
        cast_bf16_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(src), N, D, ld,
                                                        static_cast<__nv_bfloat16*>(dst));
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

}  // namespace vq
