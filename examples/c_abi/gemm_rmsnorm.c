/* Using libscaletorch_b200.so from plain C: one bf16 GEMM (y = x W^T) followed by a fused RMSNorm, checked on the host.
 *
 *   gcc -std=c11 -O2 -I include -I /usr/local/cuda/include examples/c_abi/gemm_rmsnorm.c \
 *       -L scaletorch_b200/lib -lscaletorch_b200 -Wl,-rpath,$PWD/scaletorch_b200/lib \
 *       -L /usr/local/cuda/lib64 -lcudart -lm -o gemm_rmsnorm && ./gemm_rmsnorm          (needs a B200)
 *
 * Conventions shown here: every call returns 0 or leaves a message in stb_last_error(); pointers are
 * device pointers, bf16 travels as void*; the library allocates nothing and synchronises nothing -- the caller owns the
 * buffers and the stream.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "scaletorch_b200.h"

#define CHECK_CUDA(call)                                                                              \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));               \
            return 1;                                                                                 \
        }                                                                                             \
    } while (0)

#define CHECK_STB(call)                                                                               \
    do {                                                                                              \
        if ((call) != 0) {                                                                            \
            fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, stb_last_error());                     \
            return 1;                                                                                 \
        }                                                                                             \
    } while (0)

/* round-to-nearest-even float -> bf16 and back, for host-side reference data */
static uint16_t to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    u += 0x7FFFu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

static float from_bf16(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

int main(void) {
    const int M = 256, N = 512, K = 1024;       /* x [M,K], W [N,K] (both K-major), y [M,N] */
    const float eps = 1e-6f;
    size_t nx = (size_t)M * K, nw = (size_t)N * K, ny = (size_t)M * N;
    uint16_t *hx = malloc(nx * 2), *hw = malloc(nw * 2), *hy = malloc(ny * 2), *hn = malloc(ny * 2), *hg = malloc((size_t)N * 2);
    if (!hx || !hw || !hy || !hn || !hg) return 1;
    srand(1);
    for (size_t i = 0; i < nx; i++) hx[i] = to_bf16((float)rand() / RAND_MAX - 0.5f);
    for (size_t i = 0; i < nw; i++) hw[i] = to_bf16(((float)rand() / RAND_MAX - 0.5f) * 0.1f);
    for (int j = 0; j < N; j++) hg[j] = to_bf16(1.0f + 0.01f * (float)(j % 7));

    printf("libscaletorch_b200 ABI version %d\n", stb_abi_version());
    void *dx, *dw, *dy, *dn, *dg;
    float* drstd;
    cudaStream_t stream;
    CHECK_CUDA(cudaStreamCreate(&stream));
    CHECK_CUDA(cudaMalloc(&dx, nx * 2));
    CHECK_CUDA(cudaMalloc(&dw, nw * 2));
    CHECK_CUDA(cudaMalloc(&dy, ny * 2));
    CHECK_CUDA(cudaMalloc(&dn, ny * 2));
    CHECK_CUDA(cudaMalloc(&dg, (size_t)N * 2));
    CHECK_CUDA(cudaMalloc((void**)&drstd, (size_t)M * sizeof(float)));
    CHECK_CUDA(cudaMemcpyAsync(dx, hx, nx * 2, cudaMemcpyHostToDevice, stream));
    CHECK_CUDA(cudaMemcpyAsync(dw, hw, nw * 2, cudaMemcpyHostToDevice, stream));
    CHECK_CUDA(cudaMemcpyAsync(dg, hg, (size_t)N * 2, cudaMemcpyHostToDevice, stream));

    /* y = x W^T : both operands K-major (a_mn = b_mn = 0), bf16 output, overwrite, automatic tile choice */
    CHECK_STB(stb_gemm_bf16(dx, dw, dy, M, N, K, /*lda*/ K, /*ldb*/ K, /*ldc*/ N, 0, 0, /*out_fp32*/ 0, /*accumulate*/ 0,
                            /*force_bn*/ 0, stream));
    /* n = g * (y * rsqrt(mean(y^2) + eps)) : no residual input, no residual output, keep rstd */
    CHECK_STB(stb_rmsnorm_fwd(dy, NULL, dg, NULL, dn, drstd, M, N, eps, stream));
    CHECK_CUDA(cudaMemcpyAsync(hy, dy, ny * 2, cudaMemcpyDeviceToHost, stream));
    CHECK_CUDA(cudaMemcpyAsync(hn, dn, ny * 2, cudaMemcpyDeviceToHost, stream));
    CHECK_CUDA(cudaStreamSynchronize(stream));

    /* host reference in double precision on a few rows */
    double worst_y = 0.0, worst_n = 0.0;
    for (int r = 0; r < M; r += 37) {
        double ss = 0.0;
        for (int j = 0; j < N; j++) {
            double acc = 0.0;
            for (int k = 0; k < K; k++) acc += (double)from_bf16(hx[(size_t)r * K + k]) * (double)from_bf16(hw[(size_t)j * K + k]);
            double got = from_bf16(hy[(size_t)r * N + j]);
            double err = fabs(got - acc) / (fabs(acc) + 1e-2);
            if (err > worst_y) worst_y = err;
            ss += got * got;                      /* the norm sees the bf16 GEMM output */
        }
        double rstd = 1.0 / sqrt(ss / N + eps);
        for (int j = 0; j < N; j++) {
            double want = (double)from_bf16(hg[j]) * (double)from_bf16(to_bf16((float)(from_bf16(hy[(size_t)r * N + j]) * rstd)));
            double err = fabs((double)from_bf16(hn[(size_t)r * N + j]) - want) / (fabs(want) + 1e-2);
            if (err > worst_n) worst_n = err;
        }
    }
    printf("GEMM     %d x %d x %d : worst relative error %.4f\n", M, N, K, worst_y);
    printf("RMSNorm  %d x %d      : worst relative error %.4f\n", M, N, worst_n);
    int ok = worst_y < 2e-2 && worst_n < 2e-2;
    printf(ok ? "OK\n" : "MISMATCH\n");
    cudaFree(dx); cudaFree(dw); cudaFree(dy); cudaFree(dn); cudaFree(dg); cudaFree(drstd);
    cudaStreamDestroy(stream);
    free(hx); free(hw); free(hy); free(hn); free(hg);
    return ok ? 0 : 2;
}
