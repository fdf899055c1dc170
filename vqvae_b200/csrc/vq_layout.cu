// a-0  Layout ingest: batched transpose [B, R, C] -> [B, C, R] so that channel-first latents [B, D, *spatial]
// (the live model's layout, frl/models/representation.py:317-334) become the [N, D] rows the quantiser takes, and
// the quantised rows go back to channel-first.  One pass, 64 x 64 tiles through shared memory, 4- or 8-byte
// accesses coalesced on both sides.  HBM-bound: 2 * B*R*C*s bytes.   Definition: SURVEY.md 8(a-0), 8f row 3.
#include <cuda_fp16.h>

#include "vq_common.cuh"

namespace vq {
namespace {

constexpr int TT = 64;

template <typename T>   // T = 2- or 4-byte element handled as an opaque integer
struct Pair;
template <> struct Pair<uint16_t> { using type = uint32_t; };
template <> struct Pair<uint32_t> { using type = uint2; };

template <typename T, bool kVec>
__global__ void __launch_bounds__(256)
transpose_kernel(const T* __restrict__ src, T* __restrict__ dst, int64_t R, int64_t C, int tiles_c, int tiles_r) {
    __shared__ T tile[TT][TT + 2];
    const int64_t b = blockIdx.x / ((int64_t)tiles_c * tiles_r);
    const int t = (int)(blockIdx.x % ((int64_t)tiles_c * tiles_r));
    const int64_t r0 = (int64_t)(t / tiles_c) * TT, c0 = (int64_t)(t % tiles_c) * TT;
    const T* s = src + b * R * C;
    T* d = dst + b * R * C;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    using P = typename Pair<T>::type;
    if (kVec) {
#pragma unroll
        for (int k = 0; k < TT / 8; ++k) {
            const int r = ty + 8 * k;
            if (r0 + r < R && c0 + 2 * tx < C) {
                const P v = *reinterpret_cast<const P*>(s + (r0 + r) * C + c0 + 2 * tx);
                const T* e = reinterpret_cast<const T*>(&v);
                tile[r][2 * tx] = e[0];
                tile[r][2 * tx + 1] = e[1];
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < TT / 8; ++k) {
            const int c = ty + 8 * k;
            if (c0 + c < C && r0 + 2 * tx < R) {
                P v;
                T* e = reinterpret_cast<T*>(&v);
                e[0] = tile[2 * tx][c];
                e[1] = tile[2 * tx + 1][c];
                *reinterpret_cast<P*>(d + (c0 + c) * R + r0 + 2 * tx) = v;
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < TT / 8; ++k)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int r = ty + 8 * k, c = tx + 32 * h;
                if (r0 + r < R && c0 + c < C) tile[r][c] = s[(r0 + r) * C + c0 + c];
            }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < TT / 8; ++k)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = ty + 8 * k, r = tx + 32 * h;
                if (c0 + c < C && r0 + r < R) d[(c0 + c) * R + r0 + r] = tile[r][c];
            }
    }
}

template <typename T>
int launch_t(const void* src, void* dst, int64_t B, int64_t R, int64_t C, cudaStream_t st) {
    const int tiles_c = (int)ceil_div64(C, TT), tiles_r = (int)ceil_div64(R, TT);
    const int64_t blocks = B * tiles_c * tiles_r;
    VQ_REQUIRE(blocks < (int64_t(1) << 31), "vq_layout_transpose: too many tiles");
    const bool vec = R % 2 == 0 && C % 2 == 0 && (reinterpret_cast<uintptr_t>(src) % (2 * sizeof(T))) == 0 &&
                     (reinterpret_cast<uintptr_t>(dst) % (2 * sizeof(T))) == 0;
    if (vec)
        transpose_kernel<T, true><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const T*>(src), static_cast<T*>(dst), R, C,
                                                                   tiles_c, tiles_r);
    else
        transpose_kernel<T, false><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const T*>(src), static_cast<T*>(dst), R,
                                                                    C, tiles_c, tiles_r);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

// q_cf[b, d, s] = E[idx[b * S + s]][d] cast to the latent dtype: the quantised output written STRAIGHT in the live model's
// channel-first layout [B, D, *spatial] (frl/models/representation.py:317-334) from the indices alone -- no row-major q,
// no transpose-out pass.  One block per (b, 64 consecutive s): the 64 code rows come from the L2-resident fp32 master
// (coalesced along d), are converted and parked in shared memory [d][s], and leave as 128-byte (64 x 2 B) or 256-byte
// segments along s.  HBM traffic: N*D*s bytes written + 4 N read.
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

constexpr int GS = 64;        // s positions per block
template <typename T>
__global__ void __launch_bounds__(256)
gather_rows_cf_kernel(const int32_t* __restrict__ idx, const float* __restrict__ E, int K, int D, int64_t S, int tiles_s,
                      T* __restrict__ q) {
    extern __shared__ unsigned char gsm[];
    T* tile = reinterpret_cast<T*>(gsm);                   // [D][GS + 2]
    __shared__ int codes[GS];
    const int64_t b = blockIdx.x / tiles_s;
    const int64_t s0 = (int64_t)(blockIdx.x % tiles_s) * GS;
    const int ns = (int)(S - s0 < GS ? S - s0 : GS);
    if (threadIdx.x < GS) {
        int c = threadIdx.x < ns ? idx[b * S + s0 + threadIdx.x] : 0;
        if ((unsigned)c >= (unsigned)K) c = 0;
        codes[threadIdx.x] = c;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int LD = GS + 2;
    for (int r = warp; r < ns; r += 8) {                   // a warp copies one code row, coalesced along d
        const float* e = E + (int64_t)codes[r] * D;
        for (int d = lane; d < D; d += 32) tile[d * LD + r] = from_f32<T>(__ldg(e + d));
    }
    __syncthreads();
    T* out = q + b * D * S + s0;
    for (int i = threadIdx.x; i < D * GS; i += 256) {      // consecutive threads -> consecutive s of one d
        const int d = i / GS, r = i % GS;
        if (r < ns) out[(int64_t)d * S + r] = tile[d * LD + r];
    }
}

template <typename T>
int launch_gather_cf_t(const int32_t* idx, const float* E, int K, int D, int64_t B, int64_t S, void* q, cudaStream_t st) {
    const int tiles_s = (int)ceil_div64(S, GS);
    const int64_t blocks = B * tiles_s;
    VQ_REQUIRE(blocks < (int64_t(1) << 31), "vq_gather_rows_cf: too many tiles");
    const size_t smem = (size_t)D * (GS + 2) * sizeof(T);
    VQ_REQUIRE(smem <= 200 * 1024, "vq_gather_rows_cf: D=%d too large", D);
    static int attr_dev = -1;
    const int cur = current_device();
    if (attr_dev != cur) {
        VQ_CHECK_CUDA(cudaFuncSetAttribute(gather_rows_cf_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_dev = cur;
    }
    gather_rows_cf_kernel<T><<<(unsigned)blocks, 256, smem, st>>>(idx, E, K, D, S, tiles_s, static_cast<T*>(q));
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

}  // namespace

int launch_gather_rows_cf(const int32_t* idx, const float* E_f32, int K, int D, int64_t B, int64_t S, void* q_cf, int dtype,
                          cudaStream_t st) {
    if (B == 0 || S == 0) return VQ_OK;
    if (dtype == VQ_DT_F32) return launch_gather_cf_t<float>(idx, E_f32, K, D, B, S, q_cf, st);
    if (dtype == VQ_DT_F16) return launch_gather_cf_t<__half>(idx, E_f32, K, D, B, S, q_cf, st);
    return launch_gather_cf_t<__nv_bfloat16>(idx, E_f32, K, D, B, S, q_cf, st);
}

int launch_layout_transpose(const void* src, void* dst, int64_t B, int64_t R, int64_t C, int dtype, cudaStream_t st) {
    VQ_REQUIRE(B >= 0 && R >= 0 && C >= 0, "vq_layout_transpose: bad shape");
    if (B == 0 || R == 0 || C == 0) return VQ_OK;
    return dtype == VQ_DT_F32 ? launch_t<uint32_t>(src, dst, B, R, C, st) : launch_t<uint16_t>(src, dst, B, R, C, st);
}

}  // namespace vq
