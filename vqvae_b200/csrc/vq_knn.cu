// chunked mutual-kNN pair mining, the one dense-distance hot spot of the reference's LIVE training
// step (frl/losses/pairs.py:531-610 `pairs_mutual_knn_chunked`, called at frl/training/representation/step.py:734-741).
// It is the top-k generalisation of the fused distance + argmin search: the N x N distance matrix never exists.
//
//   d2[i, j] = |x_i|^2 + |x_j|^2 - 2 x_i.x_j   (fp32; the reference's torch.cdist takes the same matmul form for N > 25
//                                               and orders by sqrt(d2), a monotone map)
//   masked to +inf:  j == i;  same patch and |coord_i - coord_j| < pos_min_spatial        (pairs.py:573-588)
//   knn[i, :] = the k smallest (ascending; ties: lower index first), -1 where the distance is +inf   (pairs.py:590-594)
//   pairs     = every (i, j) with j in knn[i] and i in knn[j], ordered by i, then by rank of j     (pairs.py:596-610)
//
// The arithmetic has to stay fp32 (neighbour ranks are compared, bf16 operands would reorder them), so this is a
// register-tiled SIMT distance GEMM, not a tensor-core one.  Exact selection without per-row sorted lists in registers:
//
//   prep    |x|^2 and X^T ([C, N_pad], so tiles load as coalesced 16-byte rows)
//   pass 1  64 x 64 distance tiles (4 x 4 per thread); every thread keeps the minimum of each of its 16 (row, column slot)
//           buckets over all candidate tiles of its split -> 64 x splits bucket minima per row
//   tau     per row, the k-th smallest bucket minimum: k distinct candidates lie at or below it, so it bounds the k-th
//           nearest distance from above (in expectation it is about the 1.15 k-th nearest)
//   pass 2  the same tiles again (bit-identical arithmetic); the few distances <= tau go to a per-row buffer
//   select  per row, the k lexicographically smallest (distance, index) of the buffer
//   redo    rows whose buffer overflowed (masses of exact ties) -- and every row when the shape does not fit the tiled
//           path -- go through the streaming kernel with per-lane sorted lists, which handles anything
//
// N is a few thousand anchors (batch x ~100-300 per patch), C = 64..128: the reference runs N / 128 rounds of
// cdist + masking + topk (about 10 launches each).
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
                float* __restrict__ knn_d2, const int* __restrict__ redo) {
    extern __shared__ float sm[];
    float* qs = sm;                               // [QPB][C]
    float* ct = qs + QPB * C;                     // [TJ][C + 1]
    float* cn = ct + TJ * (C + 1);                // [TJ] norms
    float* cc = cn + TJ;                          // [TJ][2] coords
    int* cp = reinterpret_cast<int*>(cc + 2 * TJ);   // [TJ] patch ids
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * QPB + w;
    bool live = i < N;
    if (redo != nullptr) {                         // second-chance launch: only rows the tiled path flagged
        live = live && redo[i] != 0;
        if (__syncthreads_or(live ? 1 : 0) == 0) return;
    }
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

// ------------------------------------------------------------------------------------------ tiled path
constexpr int GT = 64;            // tile edge: 64 queries x 64 candidates per block iteration, 4 x 4 per thread
constexpr int GS = GT + 4;        // shared-memory row stride in floats (16-byte aligned rows, staggered banks)
constexpr int CAP = 128;          // per-row buffer of pass-2 survivors
constexpr int MAX_SPLITS = 8;     // candidate-range splits (grid.y): 64 bucket minima per row and split

// |x_i|^2 (same summation order as knn_norms_kernel) and X^T [C, N_pad], zero beyond N
__global__ void __launch_bounds__(256)
knn_prep_kernel(const float* __restrict__ X, int64_t N, int64_t Npad, int C, float* __restrict__ XT,
                float* __restrict__ norms) {
    __shared__ float t[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t i0 = (int64_t)blockIdx.x * 32;
    float part[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c0 = 0; c0 < C; c0 += 32) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const int r = ty + 8 * s;
            const float v = (i0 + r < N && c0 + tx < C) ? X[(i0 + r) * C + c0 + tx] : 0.f;
            t[r][tx] = v;
            part[s] = fmaf(v, v, part[s]);
        }
        __syncthreads();
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const int cc = ty + 8 * s;
            if (c0 + cc < C) XT[(int64_t)(c0 + cc) * Npad + i0 + tx] = t[tx][cc];
        }
        __syncthreads();
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const float v = warp_sum(part[s]);
        const int64_t i = i0 + ty + 8 * s;
        if (tx == 0 && i < N) norms[i] = v;
    }
}

// kPass 1: bucket minima.  kPass 2: append every distance <= tau[row] to the row's buffer.
template <int kPass>
__global__ void __launch_bounds__(256)
knn_tile_kernel(const float* __restrict__ XT, const float* __restrict__ norms, const float* __restrict__ coords,
                const int32_t* __restrict__ patch, int64_t N, int64_t Npad, int C, float min_sp, int tiles_per_split,
                float* __restrict__ bmin, int nb, const float* __restrict__ tau, int* __restrict__ cnt,
                int2* __restrict__ buf) {
    extern __shared__ float sm[];
    float* qs = sm;                                 // [C][GS]  queries, feature-major
    float* cs = qs + (size_t)C * GS;                // [C][GS]  candidates, feature-major
    float* cn = cs + (size_t)C * GS;                // [GT] candidate norms (+inf beyond N)
    float* ccr = cn + GT;                           // [GT] candidate row coordinate
    float* ccc = ccr + GT;                          // [GT] candidate column coordinate
    int* cpt = reinterpret_cast<int*>(ccc + GT);    // [GT] candidate patch id
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int64_t i0 = (int64_t)blockIdx.x * GT;
    const int64_t ntiles = Npad / GT;
    const int64_t t0 = (int64_t)blockIdx.y * tiles_per_split;
    const int64_t t1 = t0 + tiles_per_split < ntiles ? t0 + tiles_per_split : ntiles;
    const bool has_mask = coords != nullptr;

    for (int t = tid; t < C * (GT / 4); t += 256) {
        const int c = t >> 4, v = t & 15;
        *reinterpret_cast<float4*>(qs + c * GS + 4 * v) =
            __ldg(reinterpret_cast<const float4*>(XT + (int64_t)c * Npad + i0) + v);
    }
    float qn[4], qr[4], qc[4], tv[4];
    int qp[4];
    bool live[4];
#pragma unroll
    for (int v = 0; v < 4; ++v) {
        const int64_t i = i0 + 4 * ty + v;
        live[v] = i < N;
        qn[v] = live[v] ? norms[i] : 0.f;
        qp[v] = (live[v] && has_mask) ? patch[i] : -1;
        qr[v] = (live[v] && has_mask) ? coords[2 * i] : 0.f;
        qc[v] = (live[v] && has_mask) ? coords[2 * i + 1] : 0.f;
        tv[v] = (kPass == 2 && live[v]) ? tau[i] : -1.f;          // d >= 0: nothing passes for dead rows
    }
    float bm[4][4];
#pragma unroll
    for (int v = 0; v < 4; ++v)
#pragma unroll
        for (int u = 0; u < 4; ++u) bm[v][u] = CUDART_INF_F;

    for (int64_t tile = t0; tile < t1; ++tile) {
        const int64_t j0 = tile * GT;
        __syncthreads();                            // previous tile fully consumed (and, first time, qs visible below)
        for (int t = tid; t < C * (GT / 4); t += 256) {
            const int c = t >> 4, v = t & 15;
            *reinterpret_cast<float4*>(cs + c * GS + 4 * v) =
                __ldg(reinterpret_cast<const float4*>(XT + (int64_t)c * Npad + j0) + v);
        }
        if (tid < GT) {
            const int64_t j = j0 + tid;
            cn[tid] = j < N ? norms[j] : CUDART_INF_F;
            cpt[tid] = (j < N && has_mask) ? patch[j] : -2;
            ccr[tid] = (j < N && has_mask) ? coords[2 * j] : 0.f;
            ccc[tid] = (j < N && has_mask) ? coords[2 * j + 1] : 0.f;
        }
        __syncthreads();
        float acc[4][4];
#pragma unroll
        for (int v = 0; v < 4; ++v)
#pragma unroll
            for (int u = 0; u < 4; ++u) acc[v][u] = 0.f;
#pragma unroll 8
        for (int c = 0; c < C; ++c) {               // ascending c, one fma per term: the streaming kernel's order
            const float4 a = *reinterpret_cast<const float4*>(qs + c * GS + 4 * ty);
            const float4 b = *reinterpret_cast<const float4*>(cs + c * GS + 4 * tx);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int v = 0; v < 4; ++v)
#pragma unroll
                for (int u = 0; u < 4; ++u) acc[v][u] = fmaf(av[v], bv[u], acc[v][u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int jl = 4 * tx + u;
            const int64_t j = j0 + jl;
            const float cnj = cn[jl];
            const int cpj = cpt[jl];
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                float d = fmaxf(fmaf(-2.f, acc[v][u], qn[v] + cnj), 0.f);
                bool masked = j >= N || j == i0 + 4 * ty + v;
                if (cpj == qp[v]) {                 // same patch (never true without a mask: -1 vs -2)
                    const float dr = qr[v] - ccr[jl], dc = qc[v] - ccc[jl];
                    masked |= sqrtf(fmaf(dr, dr, dc * dc)) < min_sp;
                }
                if (masked) d = CUDART_INF_F;
                if (kPass == 1) {
                    bm[v][u] = fminf(bm[v][u], d);
                } else if (d <= tv[v] && d < CUDART_INF_F) {
                    const int64_t i = i0 + 4 * ty + v;
                    const int pos = atomicAdd(cnt + i, 1);
                    if (pos < CAP) buf[i * CAP + pos] = make_int2(__float_as_int(d), (int)j);
                }
            }
        }
    }
    if (kPass == 1) {
#pragma unroll
        for (int v = 0; v < 4; ++v)
            if (live[v]) {
                float4* dst = reinterpret_cast<float4*>(bmin + (i0 + 4 * ty + v) * nb + blockIdx.y * GT + 4 * tx);
                *dst = make_float4(bm[v][0], bm[v][1], bm[v][2], bm[v][3]);
            }
    }
}

// lexicographic (distance, index) minimum across the warp; returns the owning lane
__device__ __forceinline__ int warp_lexmin(float& d, int& i, int lane) {
    int l = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float od = __shfl_xor_sync(0xffffffffu, d, o);
        const int oi = __shfl_xor_sync(0xffffffffu, i, o);
        const int ol = __shfl_xor_sync(0xffffffffu, l, o);
        if (od < d || (od == d && (oi < i || (oi == i && ol < l)))) { d = od; i = oi; l = ol; }
    }
    return l;
}

// tau[i] = k-th smallest of the row's nb bucket minima (duplicates count); also clears the row's pass-2 counter
__global__ void __launch_bounds__(256)
knn_tau_kernel(const float* __restrict__ bmin, int64_t N, int nb, int k, float* __restrict__ tau, int* __restrict__ cnt) {
    constexpr int PER = MAX_SPLITS * GT / 32;       // 16 values per lane at most
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= N) return;
    float v[PER];
#pragma unroll
    for (int p = 0; p < PER; ++p) v[p] = (lane + 32 * p < nb) ? bmin[i * nb + lane + 32 * p] : CUDART_INF_F;
    float kth = CUDART_INF_F;
    for (int r = 0; r < k; ++r) {
        float m = v[0];
        int mp = 0;
#pragma unroll
        for (int p = 1; p < PER; ++p)
            if (v[p] < m) { m = v[p]; mp = p; }
        float d = m;
        int tag = lane;
        const int owner = warp_lexmin(d, tag, lane);
        kth = d;
        if (lane == owner) {
#pragma unroll
            for (int p = 0; p < PER; ++p)
                if (p == mp) v[p] = CUDART_INF_F;
        }
        if (!(d < CUDART_INF_F)) break;             // fewer than k finite buckets: tau = +inf, pass 2 takes every finite distance
    }
    if (lane == 0) {
        tau[i] = kth;
        cnt[i] = 0;
    }
}

// the k lexicographically smallest (distance, index) entries of the row's buffer; overflowed rows are flagged for the
// streaming kernel instead
__global__ void __launch_bounds__(256)
knn_select_kernel(const int2* __restrict__ buf, const int* __restrict__ cnt, int64_t N, int k,
                  int64_t* __restrict__ knn_idx, float* __restrict__ knn_d2, int* __restrict__ redo) {
    constexpr int PER = CAP / 32;
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= N) return;
    const int n = cnt[i];
    if (n > CAP) {
        if (lane == 0) redo[i] = 1;
        return;
    }
    if (lane == 0) redo[i] = 0;
    float d[PER];
    int ix[PER];
#pragma unroll
    for (int p = 0; p < PER; ++p) {
        const int e = lane + 32 * p;
        const int2 w = e < n ? buf[i * CAP + e] : make_int2(__float_as_int(CUDART_INF_F), 0x7fffffff);
        d[p] = __int_as_float(w.x);
        ix[p] = w.y;
    }
    for (int r = 0; r < k; ++r) {
        float m = d[0];
        int mi = ix[0], mp = 0;
#pragma unroll
        for (int p = 1; p < PER; ++p)
            if (d[p] < m || (d[p] == m && ix[p] < mi)) { m = d[p]; mi = ix[p]; mp = p; }
        const int owner = warp_lexmin(m, mi, lane);
        if (lane == owner) {
#pragma unroll
            for (int p = 0; p < PER; ++p)
                if (p == mp) { d[p] = CUDART_INF_F; ix[p] = 0x7fffffff; }
        }
        if (lane == 0) {
            const bool ok = m < CUDART_INF_F;
            knn_idx[i * k + r] = ok ? (int64_t)mi : -1;
            if (knn_d2 != nullptr) knn_d2[i * k + r] = ok ? m : CUDART_INF_F;
        }
    }
}

// ------------------------------------------------------------------------------------------ mutual filter
// one warp per row, one lane per slot: mask[i] bit s = (i is in the list of knn[i, s]); count[i] = popc
__global__ void __launch_bounds__(256)
knn_mutual_mask_kernel(const int64_t* __restrict__ knn, int64_t N, int k, int* __restrict__ count,
                       uint32_t* __restrict__ mask) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= N) return;
    bool m = false;
    if (lane < k) {
        const int64_t j = knn[i * k + lane];
        if (j >= 0)
            for (int t = 0; t < k; ++t) m |= knn[j * k + t] == i;
    }
    const uint32_t b = __ballot_sync(0xffffffffu, m);
    if (lane == 0) {
        mask[i] = b;
        count[i] = __popc(b);
    }
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

__global__ void __launch_bounds__(256)
knn_mutual_write_kernel(const int64_t* __restrict__ knn, int64_t N, int k, const int* __restrict__ start,
                        const uint32_t* __restrict__ mask, int64_t* __restrict__ pairs) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= N) return;
    const uint32_t b = mask[i];
    if ((b >> lane) & 1u) {
        const int64_t o = start[i] + __popc(b & ((1u << lane) - 1u));        // rank order within the row is kept
        reinterpret_cast<longlong2*>(pairs)[o] = make_longlong2(i, knn[i * k + lane]);
    }
}

struct KnnWs {
    size_t norms, xt, bmin, tau, cnt, redo, buf, total;
    int64_t Npad;
    int splits, tiles_per_split, nb;
    bool tiled;
};

// Offsets are sized for MAX_SPLITS so that they depend on (N, C) alone; how many splits a launch really uses is decided
// from the SM count at launch time (sms > 0).
KnnWs knn_ws_layout(int64_t N, int C, int sms) {
    KnnWs w{};
    auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
    w.Npad = (N + GT - 1) / GT * GT;
    // tiled path: both tiles resident in shared memory, 16-byte rows; tiny problems stay on the streaming kernel
    w.tiled = N >= 4 * GT && C % 4 == 0 && (size_t)2 * C * GS * 4 + 4 * GT * 4 <= 200 * 1024;
    const int64_t ntiles = w.Npad / GT > 0 ? w.Npad / GT : 1;
    int64_t want = sms > 0 ? (2 * (int64_t)sms + ntiles - 1) / ntiles : MAX_SPLITS;      // about two blocks per SM
    if (want < 1) want = 1;
    if (want > MAX_SPLITS) want = MAX_SPLITS;
    if (want > ntiles) want = ntiles;
    w.tiles_per_split = (int)((ntiles + want - 1) / want);
    w.splits = (int)((ntiles + w.tiles_per_split - 1) / w.tiles_per_split);
    w.nb = w.splits * GT;
    size_t off = 0;
    w.norms = off; off += al((size_t)(N > 0 ? N : 1) * 4);
    if (w.tiled) {
        w.xt = off; off += al((size_t)C * w.Npad * 4);
        w.bmin = off; off += al((size_t)N * MAX_SPLITS * GT * 4);
        w.tau = off; off += al((size_t)N * 4);
        w.cnt = off; off += al((size_t)N * 4);
        w.redo = off; off += al((size_t)N * 4);
        w.buf = off; off += al((size_t)N * CAP * 8);
    }
    w.total = off;
    return w;
}

}  // namespace

size_t knn_workspace_bytes(int64_t N, int C) {
    const size_t topk = knn_ws_layout(N, C, 0).total;
    const size_t mutual = 2 * (((size_t)(N > 0 ? N : 1) * 4 + 255) & ~size_t(255));     // count + mask
    return topk > mutual ? topk : mutual;
}

int launch_knn_topk(const float* X, int64_t N, int C, const float* coords, const int32_t* patch, float pos_min_spatial,
                    int k, int64_t* knn_idx, float* knn_d2, void* ws, cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && N < (int64_t(1) << 31) - GT && C > 0 && C <= 1024 && k > 0 && k <= KMAX,
               "vq_knn_topk: bad shape N=%lld C=%d k=%d (k <= %d, C <= 1024)", (long long)N, C, k, KMAX);
    VQ_REQUIRE((coords == nullptr) == (patch == nullptr), "vq_knn_topk: coords and patch go together");
    if (N == 0) return VQ_OK;
    VQ_REQUIRE(X != nullptr && knn_idx != nullptr && ws != nullptr, "vq_knn_topk: NULL argument");
    VQ_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "vq_knn_topk: ws must be 256-byte aligned");
    static int attr_dev = -1;                  // function attributes are per device
    const int cur_dev = current_device();
    if (attr_dev != cur_dev) {
        VQ_CHECK_CUDA(cudaFuncSetAttribute(knn_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        VQ_CHECK_CUDA(cudaFuncSetAttribute(knn_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        VQ_CHECK_CUDA(cudaFuncSetAttribute(knn_tile_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_dev = cur_dev;
    }
    const KnnWs w = knn_ws_layout(N, C, sm_count());
    uint8_t* base = static_cast<uint8_t*>(ws);
    float* norms = reinterpret_cast<float*>(base + w.norms);
    const size_t smem_stream = ((size_t)QPB * C + (size_t)TJ * (C + 1) + TJ + 2 * TJ) * 4 + TJ * 4;
    const char* force = getenv("VQ_KNN_PATH");                               // "stream": tuning / test knob
    const bool tiled = w.tiled && !(force != nullptr && force[0] == 's');
    if (!tiled) {
        knn_norms_kernel<<<(unsigned)ceil_div64(N, 8), 256, 0, st>>>(X, N, C, norms);
        knn_topk_kernel<<<(unsigned)ceil_div64(N, QPB), QPB * 32, smem_stream, st>>>(X, norms, coords, patch, N, C,
                                                                               pos_min_spatial, k, knn_idx, knn_d2, nullptr);
        VQ_CHECK_CUDA(cudaGetLastError());
        count_launch(2);
        return VQ_OK;
    }
    float* XT = reinterpret_cast<float*>(base + w.xt);
    float* bmin = reinterpret_cast<float*>(base + w.bmin);
    float* tau = reinterpret_cast<float*>(base + w.tau);
    int* cnt = reinterpret_cast<int*>(base + w.cnt);
    int* redo = reinterpret_cast<int*>(base + w.redo);
    int2* buf = reinterpret_cast<int2*>(base + w.buf);
    const size_t smem_tile = (size_t)2 * C * GS * 4 + 4 * GT * 4;
    const dim3 grid((unsigned)(w.Npad / GT), (unsigned)w.splits);
    knn_prep_kernel<<<(unsigned)(w.Npad / 32), 256, 0, st>>>(X, N, w.Npad, C, XT, norms);
    knn_tile_kernel<1><<<grid, 256, smem_tile, st>>>(XT, norms, coords, patch, N, w.Npad, C, pos_min_spatial,
                                                    w.tiles_per_split, bmin, w.nb, nullptr, nullptr, nullptr);
    knn_tau_kernel<<<(unsigned)ceil_div64(N, 8), 256, 0, st>>>(bmin, N, w.nb, k, tau, cnt);
    knn_tile_kernel<2><<<grid, 256, smem_tile, st>>>(XT, norms, coords, patch, N, w.Npad, C, pos_min_spatial,
                                                    w.tiles_per_split, nullptr, w.nb, tau, cnt, buf);
    knn_select_kernel<<<(unsigned)ceil_div64(N, 8), 256, 0, st>>>(buf, cnt, N, k, knn_idx, knn_d2, redo);
    knn_topk_kernel<<<(unsigned)ceil_div64(N, QPB), QPB * 32, smem_stream, st>>>(X, norms, coords, patch, N, C,
                                                                           pos_min_spatial, k, knn_idx, knn_d2, redo);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch(6);
    return VQ_OK;
}

int launch_knn_mutual(const int64_t* knn_idx, int64_t N, int k, int64_t* pairs, int64_t* n_pairs, void* ws,
                      cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && k > 0 && k <= 32 && n_pairs != nullptr, "vq_knn_mutual_pairs: bad arguments (k <= 32)");
    if (N == 0) {
        VQ_CHECK_CUDA(cudaMemsetAsync(n_pairs, 0, 8, st));
        return VQ_OK;
    }
    VQ_REQUIRE(knn_idx != nullptr && pairs != nullptr && ws != nullptr, "vq_knn_mutual_pairs: NULL argument");
    VQ_REQUIRE((reinterpret_cast<uintptr_t>(pairs) & 15) == 0, "vq_knn_mutual_pairs: pairs must be 16-byte aligned");
    int* count = static_cast<int*>(ws);
    uint32_t* mask = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(ws) + (((size_t)N * 4 + 255) & ~size_t(255)));
    knn_mutual_mask_kernel<<<(unsigned)ceil_div64(N, 8), 256, 0, st>>>(knn_idx, N, k, count, mask);
    knn_scan_kernel<<<1, 1024, 0, st>>>(count, N, n_pairs);
    knn_mutual_write_kernel<<<(unsigned)ceil_div64(N, 8), 256, 0, st>>>(knn_idx, N, k, count, mask, pairs);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch(3);
    return VQ_OK;
}

}  // namespace vq
