/* A plain C99 consumer of include/pe_rotation.h: proves that the header is valid C (not only C++), that the shared
 * library links from C, and - with the argument "gpu" - that the path runs from a host program that uses nothing but
 * the CUDA runtime (no Python, no torch): what a C-level binding of the reference's functions would do.
 *
 *   gcc -std=c99 -Wall -Wextra -Werror -pedantic -I include tests/c_abi/consumer.c \
 *       -L poseestimation_b200 -lpe_rotation -isystem /usr/local/cuda/include -L /usr/local/cuda/lib64 -lcudart -lm
 *
 * Known answers: identity -> identity; a rotation -> itself; diag(1, 1, -1) -> a proper rotation with
 * det +1 (the det correction of rotation_representation.py:199-205); any input -> R^T R = I. */
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "pe_rotation.h"

#define CHECK(cond, msg)                                             \
    do {                                                             \
        if (!(cond)) {                                               \
            fprintf(stderr, "FAIL %s (line %d): %s\n", msg, __LINE__, pe_last_error()); \
            return 1;                                                \
        }                                                            \
    } while (0)

static float det3(const float* m) {
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}

static int cpu_checks(void) {
    CHECK(pe_version() == PE_VERSION, "pe_version");
    CHECK(pe_svd_orth_fwd(NULL, 9, NULL, NULL, -1, NULL) == PE_ERR_BAD_ARG, "negative count");
    CHECK(strstr(pe_last_error(), "negative") != NULL, "error message");
    CHECK(pe_frobenius_fwd((const float*)1024, (const float*)2048, (double*)(4096 + 4), 1, NULL) == PE_ERR_BAD_ARG,
          "misaligned accumulator");
    CHECK(pe_kabsch_fwd(NULL, NULL, 4, 0, NULL, NULL, NULL) == PE_ERR_BAD_ARG, "empty clouds");
    CHECK(pe_frobenius_fwd_static((const float*)1024, (const float*)2048, (double*)4096, (int64_t)PE_STATIC_MAX_RECORDS + 1,
                                  NULL) == PE_ERR_BAD_ARG, "static scratch limit");
    CHECK(strstr(pe_last_error(), "PE_STATIC_MAX_RECORDS") != NULL, "static scratch limit message");
    printf("consumer: cpu checks ok (version %d)\n", pe_version());
    return 0;
}

static int gpu_checks(void) {
    enum { N = 5 };
    const float c = 0.8f, s = 0.6f;
    float h_m[N * 9] = {
        1, 0, 0, 0, 1, 0, 0, 0, 1,          /* identity */
        c, -s, 0, s, c, 0, 0, 0, 1,         /* a rotation about z */
        1, 0, 0, 0, 1, 0, 0, 0, -1,         /* a reflection: the det correction must act */
        2, 0, 0, 0, 3, 0, 0, 0, 0.5f,       /* positive diagonal -> identity */
        0.3f, -1.2f, 0.7f, 2.0f, 0.1f, -0.4f, -0.9f, 0.8f, 1.5f};
    float h_r[N * 9], h_t[N * 9], h_d[N];
    double h_sum = -1.0;
    float *d_m = NULL, *d_r = NULL, *d_t = NULL, *d_d = NULL;
    double* d_sum = NULL;
    int i, j, k;
    CHECK(cudaMalloc((void**)&d_m, sizeof h_m) == cudaSuccess, "cudaMalloc");
    CHECK(cudaMalloc((void**)&d_r, sizeof h_r) == cudaSuccess, "cudaMalloc");
    CHECK(cudaMalloc((void**)&d_t, sizeof h_t) == cudaSuccess, "cudaMalloc");
    CHECK(cudaMalloc((void**)&d_d, sizeof h_d) == cudaSuccess, "cudaMalloc");
    CHECK(cudaMalloc((void**)&d_sum, sizeof(double)) == cudaSuccess, "cudaMalloc");
    CHECK(cudaMemcpy(d_m, h_m, sizeof h_m, cudaMemcpyHostToDevice) == cudaSuccess, "H2D");
    CHECK(pe_svd_orth_fwd(d_m, 9, d_r, d_d, N, NULL) == 0, "pe_svd_orth_fwd");
    CHECK(cudaMemcpy(h_r, d_r, sizeof h_r, cudaMemcpyDeviceToHost) == cudaSuccess, "D2H");
    CHECK(cudaMemcpy(h_d, d_d, sizeof h_d, cudaMemcpyDeviceToHost) == cudaSuccess, "D2H");
    for (i = 0; i < N; ++i) {
        const float* r = h_r + 9 * i;
        for (j = 0; j < 3; ++j)
            for (k = 0; k < 3; ++k) {
                float dot = r[j] * r[k] + r[3 + j] * r[3 + k] + r[6 + j] * r[6 + k];
                CHECK(fabsf(dot - (j == k ? 1.0f : 0.0f)) < 1e-5f, "R^T R = I");
            }
        CHECK(fabsf(det3(r) - 1.0f) < 1e-5f, "det R = +1");
    }
    for (j = 0; j < 9; ++j) {
        CHECK(fabsf(h_r[j] - h_m[j]) < 1e-6f, "identity -> identity");
        CHECK(fabsf(h_r[9 + j] - h_m[9 + j]) < 1e-6f, "rotation -> itself");
        CHECK(fabsf(h_r[27 + j] - h_m[j]) < 1e-6f, "positive diagonal -> identity");
    }
    CHECK(h_d[0] == 1.0f && h_d[1] == 1.0f && h_d[2] == -1.0f && h_d[3] == 1.0f, "det signs");
    /* the fused loss of the outputs against themselves is zero; against the inputs' rotations likewise */
    CHECK(pe_svd_orth_frobenius_fwd(d_m, 9, d_r, d_t, d_sum, N, NULL) == 0, "pe_svd_orth_frobenius_fwd");
    CHECK(cudaMemcpy(&h_sum, d_sum, sizeof h_sum, cudaMemcpyDeviceToHost) == cudaSuccess, "D2H");
    CHECK(fabs(h_sum) < 1e-5, "loss(R, R) = 0");
    /* round 2: explicit set-up, the step with its metric ([sum |.|_F, sum angle, flag]), the allocation-free loss */
    {
        double h3[3] = {-1.0, -1.0, -1.0}, h_a = -1.0, h_b = -2.0;
        double* d3 = NULL;
        CHECK(pe_init(-1) == 0 && pe_init(-1) == 0, "pe_init (idempotent)");
        CHECK(cudaMalloc((void**)&d3, sizeof h3) == cudaSuccess, "cudaMalloc");
        CHECK(pe_svd_orth_frobenius_step_metrics(d_m, 9, d_r, 1.0f / N, d_t, d3, NULL, NULL, NULL, NULL, N, NULL) == 0,
              "pe_svd_orth_frobenius_step_metrics");
        CHECK(cudaMemcpy(h3, d3, sizeof h3, cudaMemcpyDeviceToHost) == cudaSuccess, "D2H");
        CHECK(fabs(h3[0]) < 1e-5 && h3[1] >= 0.0 && h3[1] < 0.05 * N && h3[2] == 0.0, "step metrics of (M, svd_orth(M))");
        CHECK(pe_frobenius_fwd(d_m, d_r, d_sum, N, NULL) == 0, "pe_frobenius_fwd");
        CHECK(cudaMemcpy(&h_a, d_sum, sizeof h_a, cudaMemcpyDeviceToHost) == cudaSuccess, "D2H");
        CHECK(pe_frobenius_fwd_static(d_m, d_r, d_sum, N, NULL) == 0, "pe_frobenius_fwd_static");
        CHECK(cudaMemcpy(&h_b, d_sum, sizeof h_b, cudaMemcpyDeviceToHost) == cudaSuccess, "D2H");
        CHECK(h_a > 1.0 && fabs(h_a - h_b) < 1e-9 * N, "static-scratch loss equals the slot-based one");
        cudaFree(d3);
    }
    CHECK(pe_launch_count() >= 2, "launch counter");
    cudaFree(d_m); cudaFree(d_r); cudaFree(d_t); cudaFree(d_d); cudaFree(d_sum);
    printf("consumer: gpu checks ok (%lld launches)\n", pe_launch_count());
    return 0;
}

int main(int argc, char** argv) {
    if (cpu_checks()) return 1;
    if (argc > 1 && strcmp(argv[1], "gpu") == 0) return gpu_checks();
    return 0;
}
