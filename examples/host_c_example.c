/* One quantiser step through the C ABI from plain C: no Python, no torch.
 *
 *   gcc -std=c99 -O2 examples/host_c_example.c -Iinclude -I/usr/local/cuda/include \
 *       -Lvqvae_b200 -lvq_b200 -L/usr/local/cuda/lib64 -lcudart -lm \
 *       -Wl,-rpath,$PWD/vqvae_b200 -o host_c_example && ./host_c_example          (needs a B200: sm_100a code only)
 *
 * BASELINE config 1's shape (40 960 fp32 latents of width 64, 512 codes), bf16 tensor-core search, the stand-alone call
 * order of INTEGRATION.md: vq_prepare_codebook -> vq_assign -> vq_gather_st_loss_scatter -> vq_ema_update.
 * tests/test_host_cpu.py compiles and links this file on every CPU run; running it needs the GPU. */
#include <cuda_runtime_api.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "vq_b200.h"

#define CK(call)                                                                                    \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));             \
            return 2;                                                                               \
        }                                                                                           \
    } while (0)
#define VQ(call)                                                                                    \
    do {                                                                                            \
        int rc_ = (call);                                                                           \
        if (rc_ != VQ_OK) {                                                                         \
            fprintf(stderr, "%s:%d: rc %d: %s\n", __FILE__, __LINE__, rc_, vq_last_error());        \
            return 3;                                                                               \
        }                                                                                           \
    } while (0)

static unsigned long long rng_state = 0x9E3779B97F4A7C15ull;
static float gauss(void) {                      /* sum of 12 uniforms - 6: close enough to N(0, 1) for a demo */
    double s = 0.0;
    for (int i = 0; i < 12; ++i) {
        rng_state = rng_state * 6364136223846793005ull + 1442695040888963407ull;
        s += (double)(rng_state >> 11) / 9007199254740992.0;
    }
    return (float)(s - 6.0);
}

int main(void) {
    const int64_t N = 40960;
    const int K = 512, D = 64;
    const int search = VQ_SEARCH_BF16, dt = VQ_DT_F32;
    if (vq_abi_version() != VQ_ABI_VERSION) {
        fprintf(stderr, "library ABI %d, header %d\n", vq_abi_version(), VQ_ABI_VERSION);
        return 1;
    }
    float* hz = (float*)malloc((size_t)N * D * sizeof(float));
    float* hE = (float*)malloc((size_t)K * D * sizeof(float));
    float* hq = (float*)malloc((size_t)N * D * sizeof(float));
    int32_t* hidx = (int32_t*)malloc((size_t)N * sizeof(int32_t));
    float* hcnt = (float*)malloc((size_t)K * sizeof(float));
    if (!hz || !hE || !hq || !hidx || !hcnt) return 1;
    for (int64_t i = 0; i < N * D; ++i) hz[i] = gauss();
    for (int i = 0; i < K * D; ++i) hE[i] = gauss();

    size_t ws_bytes = 0;
    VQ(vq_workspace_bytes(N, K, D, dt, search, &ws_bytes));
    const int norm_elems = vq_norm_elems(K);
    const int64_t shadow_elems = vq_shadow_elems(K, D, search);

    float *z, *E, *norms, *q, *sq, *cnt, *sums, *ema_cnt, *ema_sum, *perp;
    void *shadow, *ws;
    int32_t *idx, *ssu, *n_dead;
    int64_t* usage;
    CK(cudaMalloc((void**)&z, (size_t)N * D * 4));
    CK(cudaMalloc((void**)&E, (size_t)K * D * 4));
    CK(cudaMalloc(&shadow, (size_t)shadow_elems * 2));
    CK(cudaMalloc((void**)&norms, (size_t)norm_elems * 4));
    CK(cudaMalloc(&ws, ws_bytes));
    CK(cudaMalloc((void**)&idx, (size_t)N * 4));
    CK(cudaMalloc((void**)&q, (size_t)N * D * 4));
    CK(cudaMalloc((void**)&sq, 4));
    CK(cudaMalloc((void**)&cnt, (size_t)K * 4));
    CK(cudaMalloc((void**)&sums, (size_t)K * D * 4));
    CK(cudaMalloc((void**)&ema_cnt, (size_t)K * 4));
    CK(cudaMalloc((void**)&ema_sum, (size_t)K * D * 4));
    CK(cudaMalloc((void**)&usage, (size_t)K * 8));
    CK(cudaMalloc((void**)&ssu, (size_t)K * 4));
    CK(cudaMalloc((void**)&perp, 4));
    CK(cudaMalloc((void**)&n_dead, 4));
    CK(cudaMemcpy(z, hz, (size_t)N * D * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(E, hE, (size_t)K * D * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ema_sum, hE, (size_t)K * D * 4, cudaMemcpyHostToDevice));       /* EMA state: n = 1, m = E */
    for (int k = 0; k < K; ++k) hcnt[k] = 1.0f;
    CK(cudaMemcpy(ema_cnt, hcnt, (size_t)K * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(usage, 0, (size_t)K * 8));
    CK(cudaMemset(ssu, 0, (size_t)K * 4));

    const vq_stream_t st = (vq_stream_t)0;                                          /* the default stream */
    VQ(vq_prepare_codebook(E, K, D, search, shadow, norms, st));
    VQ(vq_assign(z, dt, N, D, D, E, shadow, norms, K, search, idx, NULL, ws, st));
    VQ(vq_gather_st_loss_scatter(z, dt, N, D, D, idx, E, K, q, dt, D, sq, cnt, sums, VQ_STAT_SUM_Z, 0, NULL, NULL, 0.f,
                                 ws, st));
    CK(cudaMemcpy(hidx, idx, (size_t)N * 4, cudaMemcpyDeviceToHost));               /* before the codebook moves */
    CK(cudaMemcpy(hq, q, (size_t)N * D * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hcnt, cnt, (size_t)K * 4, cudaMemcpyDeviceToHost));
    VQ(vq_ema_update(E, ema_cnt, ema_sum, cnt, sums, K, D, 0.99f, 1e-5f, 1, search, shadow, norms, usage, ssu, perp,
                     n_dead, 100, 1, ws, st));
    CK(cudaDeviceSynchronize());

    float h_sq = 0.f, h_perp = 0.f;
    int32_t h_dead = -1;
    CK(cudaMemcpy(&h_sq, sq, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&h_perp, perp, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&h_dead, n_dead, 4, cudaMemcpyDeviceToHost));

    /* host checks: q is the gathered fp32 row, every row is counted once, the chosen code is (nearly) the nearest one */
    double total = 0.0, sq_host = 0.0;
    int bad_q = 0, far = 0;
    for (int k = 0; k < K; ++k) total += hcnt[k];
    for (int64_t n = 0; n < N; ++n) {
        const int c = hidx[n];
        if (c < 0 || c >= K) { ++bad_q; continue; }
        if (memcmp(hq + n * D, hE + (size_t)c * D, (size_t)D * 4) != 0) ++bad_q;
        double dc = 0.0;
        for (int d = 0; d < D; ++d) dc += ((double)hz[n * D + d] - hE[c * D + d]) * ((double)hz[n * D + d] - hE[c * D + d]);
        sq_host += dc;
        if (n % 64 == 0) {                      /* brute force on a sample; bf16 operands may flip near ties only */
            double best = 1e300;
            for (int k = 0; k < K; ++k) {
                double dk = 0.0;
                for (int d = 0; d < D; ++d) dk += ((double)hz[n * D + d] - hE[k * D + d]) * ((double)hz[n * D + d] - hE[k * D + d]);
                if (dk < best) best = dk;
            }
            if (dc > best * 1.02 + 1e-6) ++far;
        }
    }
    printf("N=%lld K=%d D=%d: sum (z-q)^2 = %.4f (host %.4f), vq loss (beta 0.25) = %.6f, perplexity = %.2f, dead codes = %d\n",
           (long long)N, K, D, h_sq, sq_host, 0.25 * h_sq / ((double)N * D), h_perp, h_dead);
    printf("rows counted %.0f of %lld, rows whose q is not the gathered code row: %d, sampled rows far from the nearest code: %d\n",
           total, (long long)N, bad_q, far);
    const int ok = total == (double)N && bad_q == 0 && far == 0 && fabs(h_sq - sq_host) <= 1e-4 * sq_host;
    printf(ok ? "OK\n" : "MISMATCH\n");
    return ok ? 0 : 4;
}
