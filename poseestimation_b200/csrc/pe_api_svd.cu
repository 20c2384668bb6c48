// extern "C" entry points of the SVD orthogonalisation: forward, backward, and the two-pass SVD + Frobenius pair.
#include "pe_api_run.cuh"

using namespace pe;

extern "C" {

int pe_svd_orth_fwd(const float* m, int64_t m_stride, float* r, float* det_sign, int64_t n, void* stream) {
    if (det_sign) return run<SvdFwd<true>>(m, m_stride, nullptr, nullptr, r, det_sign, n, nullptr, nullptr, {}, stream);
    return run<SvdFwd<false>>(m, m_stride, nullptr, nullptr, r, nullptr, n, nullptr, nullptr, {}, stream);
}
int pe_svd_orth_bwd(const float* m, int64_t m_stride, const float* r, const float* g, float* dm, int64_t n,
                    void* stream) {
    if (r) return run<SvdBwdSaved>(m, m_stride, r, g, dm, nullptr, n, nullptr, nullptr, {}, stream);
    return run<SvdBwdRecompute>(m, m_stride, g, nullptr, dm, nullptr, n, nullptr, nullptr, {}, stream);
}

int pe_svd_orth_frobenius_fwd(const float* m, int64_t m_stride, const float* r_true, float* r, double* sum,
                              int64_t n, void* stream) {
    if (r) return run<SvdFrobFwd<true>>(m, m_stride, r_true, nullptr, r, nullptr, n, sum, nullptr, {}, stream);
    return run<SvdFrobFwd<false>>(m, m_stride, r_true, nullptr, nullptr, nullptr, n, sum, nullptr, {}, stream);
}
int pe_svd_orth_frobenius_bwd(const float* m, int64_t m_stride, const float* r_true, const float* gout,
                              float inv_n, float* dm, int64_t n, void* stream) {
    if (!gout) return fail(PE_ERR_BAD_ARG, "null upstream-gradient pointer");
    return run<SvdFrobBwd>(m, m_stride, r_true, nullptr, dm, nullptr, n, nullptr, nullptr, {gout, inv_n}, stream);
}

}  // extern "C"
