// a-0  Layout ingest: batched transpose [B, R, C] -> [B, C, R] so that channel-first latents [B, D, *spatial]
// (the live model's layout, frl/models/representation.py:317-334) become the [N, D] rows the quantiser takes, and
// the quantised rows go back to channel-first.  One pass, 64 x 64 tiles through shared memory, 4- or 8-byte
// accesses coalesced on both sides.  HBM-bound: 2 * B*R*C*s bytes.   Definition: SURVEY.md 8(a-0), 8f row 3.
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

}  // namespace

int launch_layout_transpose(const void* src, void* dst, int64_t B, int64_t R, int64_t C, int dtype, cudaStream_t st) {
    VQ_REQUIRE(B >= 0 && R >= 0 && C >= 0, "vq_layout_transpose: bad shape");
    if (B == 0 || R == 0 || C == 0) return VQ_OK;
    return dtype == VQ_DT_F32 ? launch_t<uint32_t>(src, dst, B, R, C, st) : launch_t<uint16_t>(src, dst, B, R, C, st);
}

}  // namespace vq
