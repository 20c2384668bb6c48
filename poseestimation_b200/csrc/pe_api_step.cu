// extern "C" entry points of the single-pass training step (loss + dL/dx from one SVD; with the metric sums).
#include "pe_api_run.cuh"

using namespace pe;

extern "C" {

int pe_svd_orth_frobenius_step(const float* m, int64_t m_stride, const float* r_true, float inv_n, float* dm,
                               double* sum, int64_t n, void* stream) {
    return run<SvdFrobStep>(m, m_stride, r_true, nullptr, dm, nullptr, n, sum, nullptr, {inv_n}, stream);
}
int pe_svd_orth_frobenius_step_mean(const float* m, int64_t m_stride, const float* r_true, float inv_n, float* dm,
                                    double* sum, float* mean, int64_t n, void* stream) {
    if (!mean) return fail(PE_ERR_BAD_ARG, "null mean pointer");
    return run<SvdFrobStep>(m, m_stride, r_true, nullptr, dm, nullptr, n, sum, nullptr, {inv_n}, stream, mean,
                            n > 0 ? 1.0 / (double)n : 0.0);
}
int pe_svd_orth_frobenius_step_metrics(const float* m, int64_t m_stride, const float* r_true, float inv_n, float* dm,
                                       double* sums, float* loss_mean, double* angle_mean, int* flag, void* scratch,
                                       int64_t n, void* stream) {
    return run<SvdFrobStepMetrics>(m, m_stride, r_true, nullptr, dm, nullptr, n, sums, flag, {inv_n}, stream, loss_mean,
                                   n > 0 ? 1.0 / (double)n : 0.0, angle_mean, scratch, true);
}

}  // extern "C"
