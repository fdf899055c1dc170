/* vq_b200.h -- C ABI of libvq_b200.so: the B200 (sm_100a) vector-quantisation hot path.
 *
 * What this boundary replaces.  The reference (nnnagle/vq-vae) is pure in-process Python and its
 * quantiser source is absent from the tree; there is no reference FFI to mirror.
 * The entry points below are the device-side halves of the steps the reference's call sites require
 * of `model.quant` (row labels = SURVEY.md section 8a):
 *
 *   a-1  nearest-codebook search        scripts/train_vqvae.py:11,287 ; frl/config/frl_model_v0.yaml:29-33
 *   a-2  gather q = E[idx]              scripts/export_codebook.py:104 (codebook is [K, D])
 *   a-3  straight-through output        frl/config/frl_model_v0.yaml:35 ; scripts/train_vqvae.py:433-434
 *   a-4  commitment / codebook loss     frl/config/frl_bindings_v0.yaml:887-891 ; scripts/train_vqvae.py:189,321,328
 *   a-5  per-code counts and sums       scripts/train_vqvae.py:191-194,433-436 ; scripts/config.yaml:61
 *   a-7  EMA codebook update            scripts/train_vqvae.py:435-436 (ema_decay 0.99, ema_eps 1e-5)
 *   a-8  usage / dead-code accounting   scripts/train_vqvae.py:92,197-198 ; scripts/export_codebook.py:124-128
 *   a-9  perplexity                     scripts/train_vqvae.py:32,287,357
 *
 * Conventions.  Every pointer is a DEVICE pointer owned by the caller (torch tensors, kept alive until the
 * stream has drained) unless it says "host".  Every function enqueues work on `stream` and returns without
 * synchronising.  Return value 0 = ok; otherwise one of VQ_ERR_* and vq_last_error() holds a message
 * (thread-local).  No torch / ATen types appear here; the Python side binds these with ctypes
 * (vqvae_b200/_lib.py) -- see INTEGRATION.md.  There is no CPU fallback behind any of these.
 */
#ifndef VQ_B200_H_
#define VQ_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef VQ_STREAM_T_DEFINED
#define VQ_STREAM_T_DEFINED
typedef struct CUstream_st* vq_stream_t; /* == cudaStream_t */
#endif

#define VQ_ABI_VERSION 5

enum { VQ_OK = 0, VQ_ERR_ARG = 1, VQ_ERR_CUDA = 2, VQ_ERR_UNSUPPORTED = 3 };
/* storage type of latents and quantised rows.  fp16 is the reference script's DEFAULT autocast dtype
 * (scripts/train_vqvae.py:284-286); bf16 with --bf16, fp32 with --no_amp. */
enum { VQ_DT_BF16 = 0, VQ_DT_F32 = 1, VQ_DT_F16 = 2 };
/* arithmetic of the distance search */
enum {
    VQ_SEARCH_BF16 = 0, /* z, E rounded to bf16, fp32 accumulate: tcgen05 tensor-core kernel (bf16 latents native;
                           fp32 / fp16 latents are staged as bf16 in ws)                                   */
    VQ_SEARCH_F32 = 1,  /* the 1e-5 tier: fp32 products.  On the tensor cores as a split-operand search (z and E each
                           as hi + lo bf16 pieces, three MMAs per product, fp32 accumulate in TMEM) when the caller
                           passes the split shadow and D % 8 == 0, D <= 256; fp32 FMA on the CUDA cores otherwise   */
    VQ_SEARCH_F16 = 2,  /* fp16 latents x fp16-rounded codebook, fp32 accumulate: the same tcgen05 kernel with fp16
                           operands -- no staging, no cast (in_dtype must be VQ_DT_F16)                     */
    VQ_SEARCH_SIMT = 1024, /* flag, with VQ_SEARCH_F32: force the CUDA-core kernel (comparison / tests)     */
    VQ_SEARCH_PRESORT = 256, /* flag, OR-ed in (vq_assign only, needs ws): the search also leaves in ws the per-code
                               histogram of idx, each row's rank inside its code and the exclusive scan, so that
                               vq_gather_st_loss_scatter(... | VQ_STAT_PRESORTED) can skip its own histogram passes */
    VQ_SEARCH_WS_CLEAN = 512 /* flag, with VQ_SEARCH_PRESORT: the caller guarantees that the counter region of ws is
                               zero -- ws was zero-filled when it was allocated and, since then, only calls carrying
                               VQ_SEARCH_PRESORT / VQ_STAT_PRESORTED have used it (those leave it zero again) -- so the
                               search skips its memset (5-8 us in front of the kernel) */
};
/* what vq_gather_st_loss_scatter accumulates per code */
enum {
    VQ_STAT_SUM_Z = 0,  /* sum_k = sum z_n         (EMA statistics, a-5)                                 */
    VQ_STAT_SUM_RESID = 1, /* sum_k = sum (z_n - e_k) (st mode: codebook gradient = -2g/(ND) * sum_k)     */
    VQ_STAT_ACCUMULATE = 16, /* flag, OR-ed in: add to code_cnt / code_sum / sq_err_sum instead of overwriting them
                               (a batch processed in row chunks, e.g. while the next chunk is still in flight) */
    VQ_STAT_PRESORTED = 32  /* flag, OR-ed in: idx was produced by vq_assign(... | VQ_SEARCH_PRESORT) with the same
                               N, K, D, in_dtype and ws, and neither idx nor ws has been written since: the counting
                               sort shrinks to one scatter launch.  Ignored where the pass cannot use it. */
};

int vq_abi_version(void);
const char* vq_last_error(void);

/* Number of floats the e_norm2 array must hold: K_pad = K rounded up to the search kernel's 256-code tile norms
 * (padding = +inf) followed by K_pad * 8 floats of search-private data (-||e||^2/2 as a tensor-core operand image). */
int vq_norm_elems(int K);

/* Scratch bytes for one step at these sizes and this search kind (covers vq_assign and vq_gather_st_loss_scatter;
 * the staging region the search kind needs comes last, everything else sits at the same offsets for every kind). */
int vq_workspace_bytes(int64_t N, int K, int D, int in_dtype, int search, size_t* out_bytes /* host */);

/* Number of 16-bit elements the search shadow ("E_bf16" below) must hold for this search kind: K * D (bf16 rows for
 * VQ_SEARCH_BF16, fp16 rows for VQ_SEARCH_F16) or, for VQ_SEARCH_F32, 2 * K_pad * D: the hi and lo bf16 planes of the
 * split tensor-core tier (K_pad = K rounded up to 256; pad rows zero).  16-byte aligned. */
int64_t vq_shadow_elems(int K, int D, int search);

/* a-7 tail: refresh what the search reads.  E_bf16: the 16-bit shadow of vq_shadow_elems(K, D, search) elements (for
 * VQ_SEARCH_F32 it may be NULL: the search then runs on the CUDA cores); e_norm2 [vq_norm_elems(K)] fp32 = ||e_k||^2
 * of the rounded rows (VQ_SEARCH_BF16 / VQ_SEARCH_F16) or of the fp32 rows (VQ_SEARCH_F32); padding entries are +inf. */
int vq_prepare_codebook(const float* E_f32, int K, int D, int search, void* E_bf16, float* e_norm2,
                        vq_stream_t stream);

/* a-1: idx[n] = argmin_k ( ||e_k||^2 - 2 z_n.e_k ), lowest index wins ties.  z is [N, D] with row stride
 * ldz elements.  smin (nullable) receives the minimal score (add ||z_n||^2 for the squared distance).
 * Rows containing non-finite values get an unspecified in-range index here; vq_gather_st_loss_scatter
 * canonicalises them to 0. */
int vq_assign(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, const float* E_f32, const void* E_bf16,
              const float* e_norm2, int K, int search, int32_t* idx, float* smin /* nullable */, void* ws,
              vq_stream_t stream);

/* a-2..a-5 in ONE pass over z: q_out[n] = E_f32[idx[n]] (cast to out_dtype == in_dtype), sq_err_sum[0] =
 * sum (z - q)^2 (fp32, over finite rows; q_out may be NULL: loss and statistics only, see vq_gather_rows_cf), and -- when code_cnt/code_sum are non-NULL -- cnt_k and the per-code
 * sums selected by stat_kind.  Rows are visited in code-sorted order (a counting sort of idx built in ws),
 * so the sums are segmented reductions in registers with one vector red per (code, 64-row slice).
 * deterministic != 0 makes the visiting order, hence every fp32 sum, run-to-run reproducible.
 * idx is in/out: rows with non-finite z are rewritten to 0, get q = E[0] and are left out of all sums.
 * By-products for the torch-facing module (all nullable): idx64_out [N] = the final indices widened to int64 (the
 * `indices` the forward returns), loss_out [1] = sq_err_sum * loss_scale (loss_scale = (1+beta)/(N*D) or beta/(N*D)). */
int vq_gather_st_loss_scatter(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, int32_t* idx,
                              const float* E_f32, int K, void* q_out, int out_dtype, int64_t ldq, float* sq_err_sum,
                              float* code_cnt /* [K] or NULL */, float* code_sum /* [K*D] or NULL */, int stat_kind,
                              int deterministic, int64_t* idx64_out, float* loss_out, float loss_scale, void* ws,
                              vq_stream_t stream);

/* The same pass for callers that cannot give the library a workspace of their own (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scratch lives in a file-scope __device__ array of VQ_NOALLOC_SCRATCH_BYTES inside the library (every
 * GPU that loads the library holds it until the process exits); the call allocates nothing (no cudaMalloc / cudaFree /
 * allocating routine).  Limits: non-deterministic mode only, no VQ_STAT_PRESORTED, and the counting-sort scratch of the
 * shape -- about 12 bytes per row + 12 bytes per code -- must fit: N up to ~340 000 rows at K = 8192; larger shapes
 * return VQ_ERR_UNSUPPORTED with a message.  All calls on one GPU share the array: calls that overlap on the GPU (two
 * streams without ordering) race; calls on one stream, or ordered by events, are fine (each call clears what it needs).
 * Results equal vq_gather_st_loss_scatter(..., deterministic = 0, ...) on the same inputs. */
#define VQ_NOALLOC_SCRATCH_BYTES (4 << 20)
int vq_gather_st_loss_scatter_noalloc(const void* z, int in_dtype, int64_t N, int D, int64_t ldz, int32_t* idx,
                                      const float* E_f32, int K, void* q_out, int64_t ldq, float* sq_err_sum,
                                      float* code_cnt /* [K] or NULL */, float* code_sum /* [K*D] or NULL */, int stat_kind,
                                      int64_t* idx64_out, float* loss_out, float loss_scale, vq_stream_t stream);

/* a-7..a-9 (after the cross-GPU reduce of code_cnt/code_sum, a-6):
 *   n_k <- g n_k + (1-g) cnt_k ; m_k <- g m_k + (1-g) sum_k ; n^_k = (n_k+eps)/(sum n + K eps) * sum n ;
 *   e_k <- m_k / n^_k ; then the search shadow (E_bf16, e_norm2) is refreshed in the same launch.
 * do_ema == 0 skips the codebook update (st mode) but still does usage / dead / perplexity:
 *   usage_total += cnt ; steps_since_used = cnt>0 ? 0 : +1 ; n_dead = #{steps_since_used >= dead_after} ;
 *   perplexity = exp(-sum p log(p + 1e-10)), p = cnt / sum cnt. */
int vq_ema_update(float* E_f32, float* ema_cnt, float* ema_sum, const float* code_cnt, const float* code_sum, int K,
                  int D, float decay, float eps, int do_ema, int search, void* E_bf16, float* e_norm2,
                  int64_t* usage_total, int32_t* steps_since_used, float* perplexity /* [1] */,
                  int32_t* n_dead /* [1] */, int dead_after, int update_usage, void* ws, vq_stream_t stream);

/* a-6 fused into a-7..a-9 (one process per GPU on one node): instead of all-reducing code_cnt / code_sum first, the
 * update kernels read every rank's statistics buffer [code_sum (K*D) | code_cnt (K)] through peer memory (NVLink /
 * NVSwitch) and add them in rank order, so every rank computes the bit-identical update.  peer_stats / peer_flags are
 * HOST arrays of `world` DEVICE pointers (entry `rank` = this rank's own buffers, the others opened with vq_ipc_open);
 * each flag array holds VQ_MAX_PEERS int32, zero-initialised; `epoch` must increase by one per call on every rank
 * (start at 1).  The caller alternates between two statistics buffers from step to step (a rank may still be reading
 * the previous step's buffer of a faster peer).  mc_stats (nullable): the MULTICAST (NVLS) address that maps this step's
 * statistics buffer of all `world` ranks (cuMulticast* / torch symmetric memory): the kernels then fetch every element
 * with one multimem.ld_reduce -- the sum over ranks is formed inside the NVSwitch, one NVLink read per element instead of
 * `world` -- and peer_stats is used for nothing but the flags' companions.
 * mc_reduced + reduced_local (nullable, together, with mc_stats; used when do_ema and K*D % 4 == 0): a second region of
 * K*D floats at the same offset of every rank's allocation, given by its MULTICAST address and by this rank's own
 * address.  The exchange then runs in two phases, whose NVLink traffic per GPU does not grow with `world`: rank r sums the
 * r-th 1/world of the flat [K*D] array in the switch (multimem.ld_reduce) and writes it to every rank's region with
 * multimem.st, a second flag round follows, and every rank applies the update from its own copy.  With it each flag array
 * holds 2 * VQ_MAX_PEERS + 1 int32 (phase 1, phase 2, a block counter), zero-initialised.  Worth it for large codebooks
 * (K*D of a few MB); for small ones the one-phase form has one kernel and one flag round less.
 * The reference has no distributed code: SURVEY.md 8(e). */
#define VQ_MAX_PEERS 16
#define VQ_IPC_HANDLE_BYTES 64
int vq_ema_update_dp(float* E_f32, float* ema_cnt, float* ema_sum, int K, int D, float decay, float eps, int do_ema,
                     int search, void* E_bf16, float* e_norm2, int64_t* usage_total, int32_t* steps_since_used,
                     float* perplexity /* [1] */, int32_t* n_dead /* [1] */, int dead_after, int update_usage, void* ws,
                     const void* const* peer_stats /* host [world] */, void* const* peer_flags /* host [world] */,
                     int world, int rank, int epoch, const void* mc_stats /* nullable */,
                     void* mc_reduced /* nullable */, const void* reduced_local /* nullable */, vq_stream_t stream);
/* Bound of the cross-rank wait inside vq_ema_update_dp, in milliseconds of wall-clock time (default 1 800 000 = 30 min;
 * <= 0 waits for ever); returns the previous value.  It is a rendezvous of processes -- a rank may be late by seconds or
 * minutes (rank-0-only validation, checkpointing, a data-loader stall) -- so it is NOT the kernels' internal hang bound.
 * When it expires the waiting kernel traps and the launch fails with a CUDA error, like a collective's watchdog. */
int64_t vq_set_peer_timeout_ms(int64_t ms);
/* CUDA IPC plumbing for the above: cudaMalloc + zero + export a 64-byte handle / open a peer's handle / release. */
int vq_ipc_alloc(size_t bytes, void** dev_ptr /* host out */, void* handle_out /* host, 64 bytes */);
int vq_ipc_free(void* dev_ptr);
int vq_ipc_open(const void* handle /* host, 64 bytes */, void** dev_ptr /* host out */);
int vq_ipc_close(void* dev_ptr);

/* a-3/a-4 backward: grad_z = grad_q + (*grad_loss) * coef * (z - q), coef = 2*beta/(N*D) passed by the host.
 * grad_q may be NULL (no gradient arrived through the quantised output). */
int vq_backward(const void* z, const void* q, const void* grad_q /* nullable */, int64_t N, int D, int64_t ldz,
                const float* grad_loss /* device [1] */, float coef, void* grad_z, int dtype, vq_stream_t stream);

/* st mode: grad_E[k] = -(*grad_loss) * coef * resid_sum[k]  (resid_sum from VQ_STAT_SUM_RESID, coef = 2/(N*D)). */
int vq_codebook_grad(const float* resid_sum, int K, int D, const float* grad_loss /* device [1] */, float coef,
                     float* grad_E, vq_stream_t stream);

/* fp32 -> bf16 row cast (what vq_assign does internally when fp32 latents go through VQ_SEARCH_BF16; dst [N, D] dense). */
int vq_cast_bf16(const float* src, int64_t N, int D, int64_t ld, void* dst, vq_stream_t stream);

/* a-0 layout ingest: batched transpose src [B, R, C] -> dst [B, C, R] (dense, dtype VQ_DT_*).  Channel-first latents
 * [B, D, S] (frl/models/representation.py:317-334) -> rows [B*S, D]: R = D, C = S; quantised rows back: R = S, C = D. */
int vq_layout_transpose(const void* src, void* dst, int64_t B, int64_t R, int64_t C, int dtype, vq_stream_t stream);

/* a-2 for channel-first callers: q_cf[b, d, s] = E_f32[idx[b*S + s]][d] cast to `dtype`, i.e. the
 * quantised output written straight in the live model's [B, D, *spatial] layout (frl/models/representation.py:317-334)
 * from the indices alone -- no row-major q and no transpose-out pass.  Pair it with vq_gather_st_loss_scatter(q_out = NULL):
 * the pass then only reads the latent rows (loss + statistics).  idx must be final (after the pass canonicalised
 * non-finite rows to 0). */
int vq_gather_rows_cf(const int32_t* idx, const float* E_f32, int K, int D, int64_t B, int64_t S, void* q_cf, int dtype,
                      vq_stream_t stream);

/* SURVEY.md 8(f) row 4 -- mutual-kNN pair mining of the live training step (frl/losses/pairs.py:531-610
 * `pairs_mutual_knn_chunked`; call site frl/training/representation/step.py:734-741), the top-k generalisation of the
 * fused distance + argmin search.  X [N, C] fp32 dense; coords [N, 2] fp32 and patch [N] int32 (both or neither) give the
 * within-patch spatial mask |coord_i - coord_j| < pos_min_spatial; k <= 32.  knn_idx [N, k] int64: the k nearest
 * (ascending squared distance, ties lower index first), -1 where fewer than k candidates survive the masks; knn_d2
 * (nullable) the squared distances.  ws: vq_knn_workspace_bytes(N, C) bytes, 256-byte aligned (scratch for both calls). */
int vq_knn_workspace_bytes(int64_t N, int C, size_t* out_bytes);
int vq_knn_topk(const float* X, int64_t N, int C, const float* coords, const int32_t* patch, float pos_min_spatial, int k,
                int64_t* knn_idx, float* knn_d2 /* nullable */, void* ws, vq_stream_t stream);
/* pairs [N*k, 2] int64 receives every (i, j) with j in knn[i] and i in knn[j], ordered by i then by the rank of j;
 * n_pairs (device, [1] int64) the number written.  pairs 16-byte aligned; ws as above (any C). */
int vq_knn_mutual_pairs(const int64_t* knn_idx, int64_t N, int k, int64_t* pairs, int64_t* n_pairs, void* ws,
                        vq_stream_t stream);

/* Programmatic dependent launch along the step's kernel chain (search -> scatter -> pass -> EMA totals -> EMA apply):
 * every kernel starts with griddepcontrol.wait and is launched with the stream-serialization attribute, so the next
 * grid is set up while the current one drains.  mask bits: 1 search, 2 scatter, 4 pass, 8 EMA totals, 16 EMA apply
 * (default 31, or VQ_PDL from the environment); returns the previous mask.  A caller that runs several chains
 * concurrently on different streams should pass 0: early-resident grids of one chain can keep the other chain's
 * CTA pairs off the SMs (measured: two-stream C3 step 0.89 -> 1.2 ms in one run out of three). */
int vq_set_pdl_mask(int mask);

/* Number of kernels this library has launched in this process (for bench.py's gpu_launches). */
int64_t vq_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* VQ_B200_H_ */
