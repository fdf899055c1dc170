// chunked mutual-kNN pair mining, the one dense-distance hot spot of the reference's LIVE training
// step (frl/losses/pairs.py:531-610 `pairs_mutual_knn_chunked`, called at frl/training/representation/step.py:734-741).
// It is the top-k generalisation of the fused distance + argmin search: the N x N distance matrix never exists, each
// query keeps its k best candidates in registers while candidate tiles stream through shared memory.
//
//   d2[i, j] = |x_i|^2 + |x_j|^2 - 2 x_i.x_j   (fp32; the reference's torch.cdist takes the same matmul form for N > 25
//                                               and orders by sqrt(d2), a monotone map)
//   masked to +inf:  j == i;  same patch and |coord_i - coord_j| < pos_min_spatial        (pairs.py:573-588)
//   knn[i, :] = the k smallest (ascending; ties: lower index first), -1 where the distance is +inf   (pairs.py:590-594)
//   pairs     = every (i, j) with j in knn[i] and i in knn[j], ordered by i, then by rank of j     (pairs.py:596-610)
//
// N is a few thousand anchors (batch x ~100-300 per patch), C = 64..128: latency / launch bound in the reference
// (N / 128 iterations of cdist + masking + topk); here 1 + 1 + 3 launches.
#include <math_constants.h>

#include "vq_common.cuh"

namespace vq {
namespace {

constexpr int KMAX = 32;          // largest k
constexpr int QPB = 8;            // queries per block (one warp each)
constexpr int TJ = 32;            // candidates per shared-memory tile (one per lane)

__global__ void knn_norms_kernel(const float* __restrict__ X, int64_t N, int C, float* __restrict__ norms) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= N) return;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) {
        const float v = X[i * C + c];
        s = fmaf(v, v, s);
    }
    s = warp_sum(s);
    if (lane == 0) norms[i] = s;
}

__global__ void __launch_bounds__(QPB * 32)
knn_topk_kernel(const float* __restrict__ X, const float* __restrict__ norms, const float* __restrict__ coords,
                const int32_t* __restrict__ patch, int64_t N, int C, float min_sp, int k, int64_t* __restrict__ knn_idx,
                float* __restrict__ knn_d2) {
    extern __shared__ float sm[];
    float* qs = sm;                               // [QPB][C]
    float* ct = qs + QPB * C;                     // [TJ][C + 1]
    float* cn = ct + TJ * (C + 1);                // [TJ] norms
    float* cc = cn + TJ;                          // [TJ][2] coords
    int* cp = reinterpret_cast<int*>(cc + 2 * TJ);   // [TJ] patch ids
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * QPB + w;
    const bool live = i < N;
    for (int t = threadIdx.x; t < QPB * C; t += QPB * 32) {
        const int64_t qi = (int64_t)blockIdx.x * QPB + t / C;
        qs[t] = qi < N ? X[qi * C + t % C] : 0.f;
    }
    const float qn = live ? norms[i] : 0.f;
    const int qp = (live && patch != nullptr) ? patch[i] : -1;
    const float qr = (live && coords != nullptr) ? coords[2 * i] : 0.f;
    const float qc = (live && coords != nullptr) ? coords[2 * i + 1] : 0.f;

    float dk[KMAX];
    int ik[KMAX];
#pragma unroll
    for (int p = 0; p < KMAX; ++p) { dk[p] = CUDART_INF_F; ik[p] = -1; }

    for (int64_t j0 = 0; j0 < N; j0 += TJ) {
        __syncthreads();
        for (int t = threadIdx.x; t < TJ * C; t += QPB * 32) {
            const int r = t / C, c = t % C;
            ct[r * (C + 1) + c] = (j0 + r < N) ? X[(j0 + r) * C + c] : 0.f;
        }
        if (threadIdx.x < TJ) {
            const int64_t j = j0 + threadIdx.x;
            cn[threadIdx.x] = j < N ? norms[j] : CUDART_INF_F;
            cp[threadIdx.x] = (j < N && patch != nullptr) ? patch[j] : -2;
            cc[2 * threadIdx.x] = (j < N && coords != nullptr) ? coords[2 * j] : 0.f;
            cc[2 * threadIdx.x + 1] = (j < N && coords != nullptr) ? coords[2 * j + 1] : 0.f;
        }
        __syncthreads();
        const int64_t j = j0 + lane;
        float dot = 0.f;
        const float* q = qs + w * C;
        const float* cr = ct + lane * (C + 1);
#pragma unroll 4
        for (int c = 0; c < C; ++c) dot = fmaf(q[c], cr[c], dot);
        float d = fmaxf(fmaf(-2.f, dot, qn + cn[lane]), 0.f);
        bool masked = !live || j >= N || j == i;
        if (cp[lane] == qp && coords != nullptr) {
            const float dr = qr - cc[2 * lane], dc = qc - cc[2 * lane + 1];
            masked |= sqrtf(fmaf(dr, dr, dc * dc)) < min_sp;
        }
        if (masked) d = CUDART_INF_F;
        if (d < dk[KMAX - 1]) {                    // strict: an equal distance at a higher index never displaces
            dk[KMAX - 1] = d;
            ik[KMAX - 1] = (int)j;
#pragma unroll
            for (int p = KMAX - 1; p > 0; --p) {
                if (dk[p] < dk[p - 1]) {
                    const float td = dk[p]; dk[p] = dk[p - 1]; dk[p - 1] = td;
                    const int ti = ik[p]; ik[p] = ik[p - 1]; ik[p - 1] = ti;
                }
            }
        }
    }
    // merge the 32 sorted lane lists: k rounds of (distance, index)-lexicographic warp minimum
    for (int r = 0; r < k; ++r) {
        float bd = dk[0];
        int bi = ik[0] < 0 ? 0x7fffffff : ik[0];
        int bl = lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, bd, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            const int ol = __shfl_xor_sync(0xffffffffu, bl, o);
            if (od < bd || (od == bd && oi < bi)) { bd = od; bi = oi; bl = ol; }
        }
        if (lane == bl) {                          // pop the winner's head
#pragma unroll
            for (int p = 0; p < KMAX - 1; ++p) { dk[p] = dk[p + 1]; ik[p] = ik[p + 1]; }
            dk[KMAX - 1] = CUDART_INF_F;
            ik[KMAX - 1] = -1;
        }
        if (live && lane == 0) {
            const bool ok = bd < CUDART_INF_F;
            knn_idx[i * k + r] = ok ? (int64_t)bi : -1;
            if (knn_d2 != nullptr) knn_d2[i * k + r] = bd;
        }
    }
}

// count[i] = number of slots of row i whose neighbour j lists i as well
__global__ void knn_mutual_count_kernel(const int64_t* __restrict__ knn, int64_t N, int k, int* __restrict__ count) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int c = 0;
    for (int s = 0; s < k; ++s) {
        const int64_t j = knn[i * k + s];
        if (j < 0) continue;
        bool m = false;
        for (int t = 0; t < k; ++t) m |= knn[j * k + t] == i;
        c += m;
    }
    count[i] = c;
}

// one block: exclusive scan of count[0..N) in place, total -> n_pairs
__global__ void __launch_bounds__(1024) knn_scan_kernel(int* __restrict__ count, int64_t N, int64_t* __restrict__ n_pairs) {
    __shared__ int wt[32];
    __shared__ int carry;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    if (t == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < N; base += 1024) {
        const int v = base + t < N ? count[base + t] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) wt[w] = inc;
        __syncthreads();
        if (w == 0) {
            const int x = wt[lane];
            int ix = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, ix, o);
                if (lane >= o) ix += u;
            }
            wt[lane] = ix - x;
        }
        __syncthreads();
        const int excl = carry + wt[w] + inc - v;
        if (base + t < N) count[base + t] = excl;
        __syncthreads();
        if (t == 1023) carry = excl + v;
        __syncthreads();
    }
    if (t == 0) n_pairs[0] = carry;
}

__global__ void knn_mutual_write_kernel(const int64_t* __restrict__ knn, int64_t N, int k, const int* __restrict__ start,
                                        int64_t* __restrict__ pairs) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    int64_t o = start[i];
    for (int s = 0; s < k; ++s) {
        const int64_t j = knn[i * k + s];
        if (j < 0) continue;
        bool m = false;
        for (int t = 0; t < k; ++t) m |= knn[j * k + t] == i;
        if (m) {
            pairs[2 * o] = i;
            pairs[2 * o + 1] = j;
            ++o;
        }
    }
}

}  // namespace

int launch_knn_topk(const float* X, int64_t N, int C, const float* coords, const int32_t* patch, float pos_min_spatial,
                    int k, int64_t* knn_idx, float* knn_d2, void* ws, cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && N < (int64_t(1) << 31) && C > 0 && C <= 1024 && k > 0 && k <= KMAX,
               "vq_knn_topk: bad shape N=%lld C=%d k=%d (k <= %d, C <= 1024)", (long long)N, C, k, KMAX);
    VQ_REQUIRE((coords == nullptr) == (patch == nullptr), "vq_knn_topk: coords and patch go together");
    if (N == 0) return VQ_OK;
    VQ_REQUIRE(X != nullptr && knn_idx != nullptr && ws != nullptr, "vq_knn_topk: NULL argument");
    float* norms = static_cast<float*>(ws);
    knn_norms_kernel<<<(unsigned)ceil_div64(N, 8), 256, 0, st>>>(X, N, C, norms);
    const size_t smem = ((size_t)QPB * C + (size_t)TJ * (C + 1) + TJ + 2 * TJ) * 4 + TJ * 4;
    static bool attr_set = false;
    if (!attr_set) {
        VQ_CHECK_CUDA(cudaFuncSetAttribute(knn_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    knn_topk_kernel<<<(unsigned)ceil_div64(N, QPB), QPB * 32, smem, st>>>(X, norms, coords, patch, N, C, pos_min_spatial, k,
                                                                        knn_idx, knn_d2);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch(2);
    return VQ_OK;
}

int launch_knn_mutual(const int64_t* knn_idx, int64_t N, int k, int64_t* pairs, int64_t* n_pairs, void* ws,
                      cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && k > 0 && n_pairs != nullptr, "vq_knn_mutual_pairs: bad arguments");
    if (N == 0) {
        VQ_CHECK_CUDA(cudaMemsetAsync(n_pairs, 0, 8, st));
        return VQ_OK;
    }
    VQ_REQUIRE(knn_idx != nullptr && pairs != nullptr && ws != nullptr, "vq_knn_mutual_pairs: NULL argument");
    int* count = static_cast<int*>(ws);
    knn_mutual_count_kernel<<<(unsigned)ceil_div64(N, 128), 128, 0, st>>>(knn_idx, N, k, count);
    knn_scan_kernel<<<1, 1024, 0, st>>>(count, N, n_pairs);
    knn_mutual_write_kernel<<<(unsigned)ceil_div64(N, 128), 128, 0, st>>>(knn_idx, N, k, count, pairs);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch(3);
    return VQ_OK;
}

}  // namespace vq
