// a-7 .. a-9  EMA codebook update, usage / dead-code accounting, perplexity; plus the search-shadow refresh
// (bf16 copy + ||e||^2) and the two small backward kernels (a-3 / a-4).
// K*D elementwise work (2 M elements at K=8192, D=256): launch-latency bound, reported in microseconds.
// Required by: scripts/train_vqvae.py:191-194,433-436 (quantizer="ema", ema_decay, ema_eps);
// scripts/train_vqvae.py:92,197-198 and scripts/export_codebook.py:124-128 (usage); scripts/train_vqvae.py:357
// (perplexity).  Definitions: SURVEY.md 8(a-7 .. a-9).
#include <cuda_fp16.h>
#include <math_constants.h>

#include "vq_common.cuh"

namespace vq {
namespace {

struct Scal {          // lives at offset 0 of the workspace
    float tot_new;     // sum_k n_k after the decay step
    float wait_us;     // peer exchange, SUM over steps: time between this rank's own signal and the last peer's signal
    float read_us;     //   ... and reading + adding the W count vectors over NVLink (diagnostics: tools/kernel_times.py)
    float n_steps;     //   ... steps summed (the host zeroes the three to start a measurement)
};

// ---- a-6 fused into a-7: peer-memory exchange ---------------------------------------------------------------
// With one process per GPU every rank exposes its step statistics [code_sum (K*D) | code_cnt (K)] through CUDA IPC.
// Instead of an NCCL all-reduce followed by the update, the update kernels READ every rank's buffer over
// NVLink / NVSwitch and add them in rank order 0..W-1, so all ranks compute bit-identical codebooks and the small
// message (0.5 MiB at K=1024, D=128) costs one flag exchange instead of a collective's launch + ring latency.
struct Peers {
    const float* stats[VQ_MAX_PEERS];   // rank r's statistics buffer of this step (own buffer included)
    int32_t* flags[VQ_MAX_PEERS];       // rank r's flag array [VQ_MAX_PEERS]; slot s = last epoch signalled by rank s
    const float* mc;                    // multicast (NVLS) address of this step's statistics buffer over all ranks, or NULL:
                                        // one multimem.ld_reduce returns the SUM over ranks, added inside the NVSwitch --
                                        // one NVLink read per element instead of W
    float* mc_red;                      // two-phase exchange (large K*D): multicast address of the [K*D] "reduced" region
    const float* red;                   //   ... and this rank's own copy of it; NULL: every rank sums every element itself
    int world, rank, epoch;
    long long timeout_ns;               // cross-rank wait bound (vq_set_peer_timeout_ms); <= 0: wait for ever
};
// flag array of a rank (int32, zero-initialised by the caller): [0, MAX_PEERS) phase 1 "statistics of epoch e complete",
// [MAX_PEERS, 2 MAX_PEERS) phase 2 "my slice of the reduced sums of epoch e is written everywhere", [2 MAX_PEERS] block
// counter of ema_reduce_totals_kernel (returns to 0 at the end of every launch)
constexpr int FLAG_PHASE2 = VQ_MAX_PEERS, FLAG_COUNTER = 2 * VQ_MAX_PEERS;
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ float ld_peer(const float* p) {
    float v;
    asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
// sum over all ranks of 4 consecutive floats / one float of the multicast-mapped statistics buffer (reduced in the switch)
__device__ __forceinline__ float4 ld_reduce_v4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ float ld_reduce_f32(const float* mc) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
    return v;
}
// one store that lands in EVERY rank's copy of the multicast-mapped region (replicated by the switch)
__device__ __forceinline__ void st_multicast_v4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_release_sys(int32_t* p, int32_t v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_acquire_sys(const int32_t* p) {
    int32_t v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Tell every rank that this rank's statistics of `epoch` are complete (they were written by earlier kernels of this
// stream), then wait until every rank has said the same.  The wait is a rendezvous of PROCESSES, not a protocol step
// inside one kernel: a rank may arrive seconds or minutes late (rank-0-only validation or checkpointing, a data-loader
// stall), so its bound is its own -- wall-clock (globaltimer), minutes by default, set by the host with
// vq_set_peer_timeout_ms to match the process group's timeout -- and unrelated to the mbarrier hang bound.  The poll
// backs off to one system-scope load per ~2 us so that a long wait does not hammer NVLink.  A rank that is still
// missing when the bound expires traps (the launch fails with a CUDA error, like a collective's watchdog would).
__device__ __forceinline__ void peers_wait(const Peers& P, int phase_off = 0) {
    const int t = threadIdx.x;
    if (t < P.world) {
        const int32_t* mine = P.flags[P.rank] + phase_off + t;
        if (ld_acquire_sys(mine) < P.epoch) {
            const unsigned long long t0 = globaltimer_ns();
            unsigned ns = 32;
            while (ld_acquire_sys(mine) < P.epoch) {
                __nanosleep(ns);
                if (ns < 2048) ns <<= 1;
                if (P.timeout_ns > 0 && globaltimer_ns() - t0 > (unsigned long long)P.timeout_ns) {
                    printf("vq_b200: rank %d waited %lld ms for rank %d (epoch %d): giving up\n", P.rank,
                           P.timeout_ns / 1000000ll, t, P.epoch);
                    __trap();
                }
            }
        }
    }
    __syncthreads();
}
__device__ __forceinline__ void peers_signal(const Peers& P, int phase_off = 0) {      // threads 0 .. world-1
    const int t = threadIdx.x;
    if (t < P.world) {
        __threadfence_system();
        st_release_sys(P.flags[t] + phase_off + P.rank, P.epoch);
    }
}
__device__ __forceinline__ void peers_signal_and_wait(const Peers& P) {
    peers_signal(P);
    peers_wait(P);
}

// one block.  tot_new = sum_k (g n_k + (1-g) cnt_k) in a fixed order (double); usage / dead / perplexity.
// kPeer: code_cnt is the sum over ranks of the peers' counts (read over NVLink after the flag exchange); the reduced
// counts are also written to cnt_red (local) for the second pass below and for ema_apply_kernel.
struct TotalsArgs {
    const float* ema_cnt;
    const float* code_cnt;
    int K;
    float decay;
    int do_ema;
    Scal* scal;
    int64_t* usage_total;
    int32_t* steps_since_used;
    float* perplexity;
    int32_t* n_dead;
    int dead_after, update_usage;
    int64_t KD;
    float* cnt_red;
};
// body of the totals step for ONE block of 1024 threads (sh: 3 x 32 doubles of shared memory)
template <bool kPeer>
__device__ __forceinline__ void totals_body(const TotalsArgs& A, const Peers& P, double (*sh)[32]) {
    const float* __restrict__ ema_cnt = A.ema_cnt;
    const float* __restrict__ code_cnt = A.code_cnt;
    const int K = A.K, do_ema = A.do_ema, dead_after = A.dead_after, update_usage = A.update_usage;
    const float decay = A.decay;
    Scal* __restrict__ scal = A.scal;
    int64_t* __restrict__ usage_total = A.usage_total;
    int32_t* __restrict__ steps_since_used = A.steps_since_used;
    float* __restrict__ perplexity = A.perplexity;
    int32_t* __restrict__ n_dead = A.n_dead;
    const int64_t KD = A.KD;
    float* __restrict__ cnt_red = A.cnt_red;
    const int t = threadIdx.x;
    const float omg = 1.0f - decay;
    double s_n = 0.0, s_c = 0.0;
    int dead = 0;
    unsigned long long tp0 = 0, tp1 = 0;
    if (kPeer) {
        const bool waited = P.red != nullptr;     // two-phase: this block of ema_reduce_totals_kernel has already waited for phase 1
        if (t == 0) tp0 = globaltimer_ns();
        if (!waited) peers_signal_and_wait(P);
        if (t == 0) tp1 = globaltimer_ns();
        // all ranks' values of a batch are requested before the first is used (one NVLink round trip per batch, not one
        // per rank); the adds then run in rank order, so every rank forms bit-identical sums
        constexpr int KB = 4;
        if (P.mc != nullptr) {                    // in-switch sum: one request per 4 codes (K_pad % 4 == 0 by layout)
            for (int k4 = t * 4; k4 < K; k4 += 4096) {
                if (k4 + 4 <= K) {
                    const float4 c = ld_reduce_v4(P.mc + KD + k4);
                    cnt_red[k4] = c.x; cnt_red[k4 + 1] = c.y; cnt_red[k4 + 2] = c.z; cnt_red[k4 + 3] = c.w;
                } else {
                    for (int k = k4; k < K; ++k) cnt_red[k] = ld_reduce_f32(P.mc + KD + k);
                }
            }
        } else
        for (int k0 = t; k0 < K; k0 += 1024 * KB) {
            float v[KB][VQ_MAX_PEERS];
#pragma unroll
            for (int b = 0; b < KB; ++b)
#pragma unroll
                for (int r = 0; r < VQ_MAX_PEERS; ++r)
                    v[b][r] = (r < P.world && k0 + 1024 * b < K) ? ld_peer(P.stats[r] + KD + k0 + 1024 * b) : 0.f;
#pragma unroll
            for (int b = 0; b < KB; ++b) {
                float c = 0.f;
#pragma unroll
                for (int r = 0; r < VQ_MAX_PEERS; ++r)
                    if (r < P.world) c = __fadd_rn(c, v[b][r]);
                if (k0 + 1024 * b < K) cnt_red[k0 + 1024 * b] = c;
            }
        }
        __syncthreads();
        code_cnt = cnt_red;
        if (t == 0 && !waited) {
            scal->wait_us += (float)(tp1 - tp0) * 1e-3f;
            scal->read_us += (float)(globaltimer_ns() - tp1) * 1e-3f;
            scal->n_steps += 1.f;
        }
    }
    // one block walks K counters: the loads of TB codes per thread are issued together (the loop is latency-, not
    // bandwidth-bound: 15.8 -> us at K = 8192 with one load chain per iteration); sums keep the order k = t, t+1024, ...
    constexpr int TB = 8;
    for (int k0 = t; k0 < K; k0 += 1024 * TB) {
        float c[TB], e[TB];
        int64_t u[TB];
        int32_t ssu[TB];
#pragma unroll
        for (int b = 0; b < TB; ++b) {
            const int k = k0 + 1024 * b;
            const bool in = k < K;
            c[b] = in ? code_cnt[k] : 0.f;
            e[b] = (in && do_ema) ? ema_cnt[k] : 0.f;
            u[b] = (in && update_usage) ? usage_total[k] : 0;
            ssu[b] = in ? steps_since_used[k] : 0;
        }
#pragma unroll
        for (int b = 0; b < TB; ++b) {
            const int k = k0 + 1024 * b;
            if (k >= K) break;
            if (do_ema) s_n += (double)__fadd_rn(__fmul_rn(decay, e[b]), __fmul_rn(omg, c[b]));
            s_c += (double)c[b];
            if (update_usage) {
                usage_total[k] = u[b] + (int64_t)llrintf(c[b]);
                const int s = c[b] > 0.f ? 0 : ssu[b] + 1;
                steps_since_used[k] = s;
                dead += s >= dead_after;
            } else {
                dead += ssu[b] >= dead_after;
            }
        }
    }
    double v[3] = {s_n, s_c, (double)dead};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
        if ((t & 31) == 0) sh[i][t >> 5] = v[i];
    }
    __syncthreads();
    if (t < 32) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            v[i] = sh[i][t];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
        }
        if (t == 0) {
            scal->tot_new = (float)v[0];
            n_dead[0] = (int32_t)v[2];
            sh[1][0] = v[1];
        }
    }
    __syncthreads();
    const double tot_c = sh[1][0];
    __syncthreads();
    // perplexity = exp(-sum p log(p + 1e-10))
    double h = 0.0;
    if (tot_c > 0.0)
        for (int k0 = t; k0 < K; k0 += 1024 * TB) {
            float c[TB];
#pragma unroll
            for (int b = 0; b < TB; ++b) c[b] = k0 + 1024 * b < K ? code_cnt[k0 + 1024 * b] : 0.f;
#pragma unroll
            for (int b = 0; b < TB; ++b) {
                if (k0 + 1024 * b >= K) break;
                const float p = (float)((double)c[b] / tot_c);
                h += (double)(p * logf(p + 1e-10f));
            }
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((t & 31) == 0) sh[0][t >> 5] = h;
    __syncthreads();
    if (t < 32) {
        h = sh[0][t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
        if (t == 0) perplexity[0] = tot_c > 0.0 ? (float)exp(-h) : 1.0f;
    }
}

template <bool kPeer>
__global__ void __launch_bounds__(1024) ema_totals_kernel(const TotalsArgs A, const Peers P) {
    __shared__ double sh[3][32];
    pdl_wait();
    pdl_trigger();
    totals_body<kPeer>(A, P, sh);
}

// Two-phase exchange for large codebooks (reduce-scatter -> all-gather inside the switch).  With one-phase reads every
// rank pulls the whole [K*D] sum through the switch, so each GPU's NVLink carries W x K*D x 4 bytes outbound per step
// (measured: ema_apply_kernel<true> 35 us at 2 GPUs, 55 us at 4, K=8192 D=256).  Here rank r reduces only the r-th 1/W of
// the flat [K*D] array (multimem.ld_reduce: summed in the switch) and writes the result to EVERY rank's "reduced" region
// with one multimem.st per 16 bytes (replicated by the switch): K*D x 4 bytes out and in per GPU whatever W is.  Every
// rank then applies the EMA step from its own copy, which holds the same bits on all ranks.
// First kernel of the update, blocks of 1024 threads: block 0 signals phase 1, all blocks wait for it (polling this
// rank's own flag array); blocks 0 .. gridDim.x-2 move the slice and the last of them to finish signals phase 2; the
// LAST block does the totals step (counts summed in the switch, n totals, usage, dead codes, perplexity) at the same
// time, so the K-length reductions cost no time of their own and the phase-2 flags fly while they run.
__global__ void __launch_bounds__(1024)
ema_reduce_totals_kernel(int64_t KD4, const TotalsArgs A, const Peers P) {
    __shared__ double sh[3][32];
    __shared__ int s_last;
    pdl_wait();
    pdl_trigger();
    const int t = threadIdx.x;
    const int nslice = (int)gridDim.x - 1;                // blocks that move the slice
    unsigned long long tp0 = 0, tp1 = 0;
    if (blockIdx.x == 0) {
        if (t == 0) tp0 = globaltimer_ns();
        peers_signal(P);
    }
    peers_wait(P);
    if ((int)blockIdx.x == nslice) {                      // totals block (P.red != NULL: it does not wait again)
        totals_body<true>(A, P, sh);
        return;
    }
    if (blockIdx.x == 0 && t == 0) tp1 = globaltimer_ns();
    const int64_t per = (KD4 + P.world - 1) / P.world;
    const int64_t i0 = per * P.rank, i1 = i0 + per < KD4 ? i0 + per : KD4;
    // one 16-byte request per thread per round, spread over as many SMs as the slice has 16 KiB pieces: 16 blocks x 4
    // requests per thread moved the 1 MiB slice of an 8-GPU step in 43 us, 64 blocks x 256 threads x 4 in 26 us
    constexpr int U = 1;
    for (int64_t b = i0 + ((int64_t)blockIdx.x * U) * 1024 + t; b < i1; b += (int64_t)nslice * U * 1024) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (b + u * 1024 < i1) v[u] = ld_reduce_v4(P.mc + 4 * (b + u * 1024));
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (b + u * 1024 < i1) st_multicast_v4(P.mc_red + 4 * (b + u * 1024), v[u]);
    }
    __threadfence_system();
    __syncthreads();
    int32_t* ctr = P.flags[P.rank] + FLAG_COUNTER;
    if (t == 0) s_last = atomicAdd(ctr, 1) == nslice - 1;
    __syncthreads();
    if (s_last) {
        peers_signal(P, FLAG_PHASE2);
        if (t == 0) *ctr = 0;
    }
    if (blockIdx.x == 0 && t == 0) {
        A.scal->wait_us += (float)(tp1 - tp0) * 1e-3f;
        A.scal->read_us += (float)(globaltimer_ns() - tp1) * 1e-3f;
        A.scal->n_steps += 1.f;
    }
}

// "aug" image behind the K_pad norms (float offset K_pad): per 256-code tile an 8 KiB block holding -||e_k||^2 / 2 split
// into three bf16 pieces (h1 + h2 + h3 reproduces the fp32 value), laid out as the canonical no-swizzle K-major UMMA
// operand [256 codes x 16] (core matrix = 8 rows x 16 B; group of 8 rows g at g*256, second K half at +128).  The
// tensor-core search multiplies it by a constant [1,1,1,0..] block so the accumulator already holds z.e - ||e||^2/2.
// fp16 shadow (VQ_SEARCH_F16): the pieces are fp16 and carry the weights (2^8, 2^-4, 2^-14) of the A-side partner row
// (vq_assign_tc.cu: OperandKind::ones*): p1 = fp16(h 2^-8), p2 = fp16((h - 2^8 p1) 2^4), p3 = fp16((.. - 2^-4 p2) 2^14);
// every residual is exact in fp32, so 2^8 p1 + 2^-4 p2 + 2^-14 p3 keeps 24 significant bits of h for 1e-11 <= |h| <= 1.6e7
// (fp16 alone would overflow at ||e||^2 = 131008 and lose bits below 1e-4).
__device__ __forceinline__ void write_aug(float* e_norm2, int norm_elems, int k, float nrm, bool pad, bool fp16) {
    uint8_t* img = reinterpret_cast<uint8_t*>(e_norm2 + norm_elems) + (size_t)(k >> 8) * 8192 + ((k & 255) >> 3) * 256 +
                   (k & 7) * 16;
    uint4 lo = make_uint4(0u, 0u, 0u, 0u);
    if (pad) {
        lo.x = fp16 ? 0xFC00u : 0xFF80u;                   // -inf: a padding code never wins the arg-max
    } else if (fp16) {
        const float h = -0.5f * nrm;
        const __half p1 = __float2half_rn(h * 0.00390625f);
        const float r1 = h - __half2float(p1) * 256.0f;
        const __half p2 = __float2half_rn(r1 * 16.0f);
        const float r2 = r1 - __half2float(p2) * 0.0625f;
        const __half p3 = __float2half_rn(r2 * 16384.0f);
        lo.x = (uint32_t)__half_as_ushort(p1) | ((uint32_t)__half_as_ushort(p2) << 16);
        lo.y = (uint32_t)__half_as_ushort(p3);
    } else {
        const float h = -0.5f * nrm;
        const __nv_bfloat16 h1 = __float2bfloat16_rn(h);
        const float r1 = h - __bfloat162float(h1);
        const __nv_bfloat16 h2 = __float2bfloat16_rn(r1);
        const __nv_bfloat16 h3 = __float2bfloat16_rn(r1 - __bfloat162float(h2));
        lo.x = (uint32_t)__bfloat16_as_ushort(h1) | ((uint32_t)__bfloat16_as_ushort(h2) << 16);
        lo.y = (uint32_t)__bfloat16_as_ushort(h3);
    }
    *reinterpret_cast<uint4*>(img) = lo;
    *reinterpret_cast<uint4*>(img + 128) = make_uint4(0u, 0u, 0u, 0u);
}

// one warp per code row.  do_ema: decay step + Laplace-normalised codebook row; always: refresh the search shadow + norm.
// The shadow follows the search kind: bf16 rows (VQ_SEARCH_BF16), fp16 rows (VQ_SEARCH_F16), or -- VQ_SEARCH_F32 with a
// shadow -- the hi / lo bf16 planes [2][norm_elems][D] of the split tensor-core tier (pad rows zero).
// kPeer: the per-code sums are added over the ranks' buffers (peer loads) in rank order.
template <bool kPeer>
__global__ void __launch_bounds__(256)
ema_apply_kernel(float* __restrict__ E, float* __restrict__ ema_cnt, float* __restrict__ ema_sum,
                 const float* __restrict__ code_cnt, const float* __restrict__ code_sum, int K, int D, float decay,
                 float eps, float k_eps, int do_ema, int search, const Scal* __restrict__ scal,
                 __nv_bfloat16* __restrict__ E_bf16, float* __restrict__ e_norm2, int norm_elems, const Peers P) {
    pdl_wait();
    pdl_trigger();
    const bool use_red = kPeer && do_ema && P.red != nullptr;
    if (use_red) peers_wait(P, FLAG_PHASE2);      // every rank's slice of the reduced sums has landed in this rank's copy
    const int lane = threadIdx.x & 31;
    const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (k >= norm_elems) return;
    const bool f16 = search == VQ_SEARCH_F16, split = search == VQ_SEARCH_F32 && E_bf16 != nullptr;
    if (k >= K) {
        if (lane == 0) {
            e_norm2[k] = CUDART_INF_F;
            write_aug(e_norm2, norm_elems, k, 0.f, true, f16);
        }
        if (split)
            for (int d = lane; d < D; d += 32) {
                E_bf16[(int64_t)k * D + d] = __float2bfloat16_rn(0.f);
                E_bf16[((int64_t)norm_elems + k) * D + d] = __float2bfloat16_rn(0.f);
            }
        return;
    }
    const float omg = 1.0f - decay;
    float nhat = 1.f;
    if (do_ema) {
        const float n = __fadd_rn(__fmul_rn(decay, ema_cnt[k]), __fmul_rn(omg, code_cnt[k]));
        const float tot = scal->tot_new;
        nhat = __fmul_rn(__fdiv_rn(__fadd_rn(n, eps), __fadd_rn(tot, k_eps)), tot);
        __syncwarp();
        if (lane == 0) ema_cnt[k] = n;
    }
    float nrm = 0.f;
    constexpr int DB = 4;                     // elements per lane whose peer loads are in flight together
    extern __shared__ float mc_row[];         // multicast path: [8 warps][D] rank-summed rows
    const bool use_mc = kPeer && do_ema && P.mc != nullptr && !use_red;
    float* my_row = mc_row + (threadIdx.x >> 5) * D;
    if (use_mc) {
        // in-switch reduction: every lane asks for 16 bytes of the row's sum over ranks (all requests of the row are in
        // flight together: one NVLink round trip per row, one read per element instead of W)
        const float* src = P.mc + (int64_t)k * D;
        if ((D & 3) == 0) {
            for (int d4 = lane * 4; d4 < D; d4 += 128) *reinterpret_cast<float4*>(my_row + d4) = ld_reduce_v4(src + d4);
        } else {
            for (int d = lane; d < D; d += 32) my_row[d] = ld_reduce_f32(src + d);
        }
        __syncwarp();
    }
    for (int d0 = lane; d0 < D; d0 += 32 * DB) {
        float cs[DB];
        if (use_red) {                        // written by the peers through the switch: read from L2, never a stale L1 line
#pragma unroll
            for (int b = 0; b < DB; ++b) cs[b] = d0 + 32 * b < D ? __ldcg(P.red + (int64_t)k * D + d0 + 32 * b) : 0.f;
        } else if (use_mc) {
#pragma unroll
            for (int b = 0; b < DB; ++b) cs[b] = d0 + 32 * b < D ? my_row[d0 + 32 * b] : 0.f;
        } else if (do_ema && kPeer) {         // one NVLink round trip for DB x world values, adds in rank order
            float v[DB][VQ_MAX_PEERS];
#pragma unroll
            for (int b = 0; b < DB; ++b)
#pragma unroll
                for (int r = 0; r < VQ_MAX_PEERS; ++r)
                    v[b][r] = (r < P.world && d0 + 32 * b < D) ? ld_peer(P.stats[r] + (int64_t)k * D + d0 + 32 * b) : 0.f;
#pragma unroll
            for (int b = 0; b < DB; ++b) {
                cs[b] = 0.f;
#pragma unroll
                for (int r = 0; r < VQ_MAX_PEERS; ++r)
                    if (r < P.world) cs[b] = __fadd_rn(cs[b], v[b][r]);
            }
        }
#pragma unroll
        for (int b = 0; b < DB; ++b) {
            const int d = d0 + 32 * b;
            if (d >= D) break;
            const int64_t o = (int64_t)k * D + d;
            float ev;
            if (do_ema) {
                const float c = kPeer ? cs[b] : code_sum[o];
                const float m = __fadd_rn(__fmul_rn(decay, ema_sum[o]), __fmul_rn(omg, c));
                ema_sum[o] = m;
                ev = __fdiv_rn(m, nhat);
                E[o] = ev;
            } else {
                ev = E[o];
            }
            float en = ev;                        // the value whose norm the search adds: what the MMA operand holds
            if (f16) {
                const __half eh = __float2half_rn(ev);
                if (E_bf16 != nullptr) reinterpret_cast<__half*>(E_bf16)[o] = eh;
                en = __half2float(eh);
            } else {
                const __nv_bfloat16 eb = __float2bfloat16_rn(ev);
                if (E_bf16 != nullptr) E_bf16[o] = eb;
                if (search == VQ_SEARCH_BF16) en = __bfloat162float(eb);
                else if (split) E_bf16[(int64_t)norm_elems * D + o] = __float2bfloat16_rn(ev - __bfloat162float(eb));
            }
            nrm = fmaf(en, en, nrm);
        }
    }
    nrm = warp_sum(nrm);
    if (lane == 0) {
        e_norm2[k] = nrm;
        write_aug(e_norm2, norm_elems, k, nrm, false, f16);
    }
}

template <typename T>
__global__ void backward_kernel(const T* __restrict__ z, const T* __restrict__ q, const T* __restrict__ gq, int64_t N,
                                int D, int64_t ldz, const float* __restrict__ grad_loss, float coef,
                                T* __restrict__ gz) {
    const float c = coef * grad_loss[0];
    const int64_t total = N * (int64_t)D;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / D;
        const int d = (int)(i - r * D);
        const float zf = (float)z[r * ldz + d];
        const float qf = (float)q[i];
        float g = c * (zf - qf);
        if (gq != nullptr) g += (float)gq[i];
        gz[i] = (T)g;
    }
}

__global__ void codebook_grad_kernel(const float* __restrict__ resid, int64_t n, const float* __restrict__ grad_loss,
                                     float coef, float* __restrict__ gE) {
    const float c = -coef * grad_loss[0];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        gE[i] = c * resid[i];
}

}  // namespace

int launch_ema_update(float* E_f32, float* ema_cnt, float* ema_sum, const float* code_cnt, const float* code_sum,
                      int K, int D, float decay, float eps, int do_ema, int search, void* E_bf16, float* e_norm2,
                      int64_t* usage_total, int32_t* steps_since_used, float* perplexity, int32_t* n_dead,
                      int dead_after, int update_usage, void* ws, cudaStream_t st, const void* const* peer_stats,
                      void* const* peer_flags, int world, int rank, int epoch, const void* mc_stats, void* mc_reduced,
                      const void* reduced_local) {
    Peers P;
    P.world = 0; P.rank = 0; P.epoch = 0; P.timeout_ns = peer_timeout_ns();
    P.mc = static_cast<const float*>(mc_stats);
    // two-phase form: needs the in-switch sum, 16-byte granules, and something to apply (st mode exchanges counts only)
    const bool two_phase = mc_stats != nullptr && mc_reduced != nullptr && reduced_local != nullptr && do_ema &&
                           ((int64_t)K * D) % 4 == 0 && peer_stats != nullptr;
    P.mc_red = two_phase ? static_cast<float*>(mc_reduced) : nullptr;
    P.red = two_phase ? static_cast<const float*>(reduced_local) : nullptr;
    const bool peer = peer_stats != nullptr;
    if (peer) {
        VQ_REQUIRE(world >= 1 && world <= VQ_MAX_PEERS && rank >= 0 && rank < world && peer_flags != nullptr && epoch > 0,
                   "vq_ema_update_dp: bad world=%d rank=%d epoch=%d (at most %d ranks)", world, rank, epoch, VQ_MAX_PEERS);
        for (int r = 0; r < world; ++r) {
            VQ_REQUIRE(peer_stats[r] != nullptr && peer_flags[r] != nullptr, "vq_ema_update_dp: NULL peer pointer %d", r);
            P.stats[r] = static_cast<const float*>(peer_stats[r]);
            P.flags[r] = static_cast<int32_t*>(peer_flags[r]);
        }
        P.world = world; P.rank = rank; P.epoch = epoch;
        code_cnt = P.stats[rank] + (int64_t)K * D;          // non-NULL marker; the kernels read the peers' buffers
        code_sum = P.stats[rank];
    }
    VQ_REQUIRE(K > 0 && D > 0, "vq_ema_update: bad shape K=%d D=%d", K, D);
    VQ_REQUIRE(E_f32 != nullptr && e_norm2 != nullptr, "vq_ema_update: E_f32 and e_norm2 are required");
    Scal* scal = static_cast<Scal*>(ws);
    if (code_cnt != nullptr) {
        VQ_REQUIRE(ws != nullptr && usage_total != nullptr && steps_since_used != nullptr && perplexity != nullptr &&
                       n_dead != nullptr,
                   "vq_ema_update: statistics buffers and workspace are required when code_cnt is given");
        VQ_REQUIRE(!do_ema || (ema_cnt != nullptr && ema_sum != nullptr && code_sum != nullptr),
                   "vq_ema_update: do_ema needs ema_cnt, ema_sum and code_sum");
        float* cnt_red = reinterpret_cast<float*>(static_cast<uint8_t*>(ws) + 256);    // reserved by ws_layout
        TotalsArgs TA{ema_cnt, code_cnt, K, decay, do_ema, scal, usage_total, steps_since_used, perplexity, n_dead, dead_after,
                      update_usage, peer ? (int64_t)K * D : (int64_t)0, peer ? cnt_red : nullptr};
        if (two_phase) {
            const int64_t KD4 = (int64_t)K * D / 4, per = (KD4 + world - 1) / world;
            int64_t blocks = (per + 1023) / 1024;                       // 16 bytes per thread per round
            blocks = blocks < 1 ? 1 : (blocks > sm_count() - 1 ? sm_count() - 1 : blocks);
            VQ_CHECK_CUDA(launch_pdl(PDL_TOTALS, ema_reduce_totals_kernel, dim3((unsigned)blocks + 1), dim3(1024), 0, st, KD4, TA, P));
        } else if (peer) {
            VQ_CHECK_CUDA(launch_pdl(PDL_TOTALS, ema_totals_kernel<true>, dim3(1), dim3(1024), 0, st, TA, P));
        } else {
            VQ_CHECK_CUDA(launch_pdl(PDL_TOTALS, ema_totals_kernel<false>, dim3(1), dim3(1024), 0, st, TA, P));
        }
        if (peer) code_cnt = cnt_red;
        VQ_CHECK_CUDA(cudaGetLastError());
        count_launch();
    } else {
        VQ_REQUIRE(!do_ema, "vq_ema_update: do_ema without code_cnt");
    }
    const int ne = round_up(K, 256);                 // codes incl. padding; vq_norm_elems(K) floats = ne norms + aug image
    const float k_eps = (float)((double)K * (double)eps);
    const size_t apply_smem = (peer && do_ema && P.mc != nullptr && !two_phase) ? (size_t)8 * D * sizeof(float) : 0;
    VQ_REQUIRE(apply_smem <= 48 * 1024, "vq_ema_update_dp: D=%d too large for the multicast path (D <= 1536)", D);
    VQ_CHECK_CUDA(launch_pdl(PDL_APPLY, (peer && do_ema) ? ema_apply_kernel<true> : ema_apply_kernel<false>, dim3((ne + 7) / 8),
                             dim3(256), apply_smem, st, E_f32, ema_cnt, ema_sum, code_cnt, code_sum, K, D, decay, eps, k_eps, do_ema,
                             search, static_cast<const Scal*>(scal), static_cast<__nv_bfloat16*>(E_bf16), e_norm2, ne, P));
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

int launch_backward(const void* z, const void* q, const void* gq, int64_t N, int D, int64_t ldz,
                    const float* grad_loss, float coef, void* gz, int dtype, cudaStream_t st) {
    VQ_REQUIRE(N >= 0 && D > 0 && ldz >= D, "vq_backward: bad shape");
    if (N == 0) return VQ_OK;
    const int blocks = sm_count() * 8;
    if (dtype == VQ_DT_F32)
        backward_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(z), static_cast<const float*>(q),
                                                       static_cast<const float*>(gq), N, D, ldz, grad_loss, coef,
                                                       static_cast<float*>(gz));
    else if (dtype == VQ_DT_F16)
        backward_kernel<__half><<<blocks, 256, 0, st>>>(static_cast<const __half*>(z), static_cast<const __half*>(q),
                                                        static_cast<const __half*>(gq), N, D, ldz, grad_loss, coef,
                                                        static_cast<__half*>(gz));
    else
        backward_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(
            static_cast<const __nv_bfloat16*>(z), static_cast<const __nv_bfloat16*>(q),
            static_cast<const __nv_bfloat16*>(gq), N, D, ldz, grad_loss, coef, static_cast<__nv_bfloat16*>(gz));
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

int launch_codebook_grad(const float* resid, int K, int D, const float* grad_loss, float coef, float* gE,
                         cudaStream_t st) {
    VQ_REQUIRE(K > 0 && D > 0, "vq_codebook_grad: bad shape");
    codebook_grad_kernel<<<sm_count() * 4, 256, 0, st>>>(resid, (int64_t)K * D, grad_loss, coef, gE);
    VQ_CHECK_CUDA(cudaGetLastError());
    count_launch();
    return VQ_OK;
}

}  // namespace vq
