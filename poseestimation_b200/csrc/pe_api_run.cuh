// make_params / run<Op>: how every extern "C" entry point of the per-record ops reaches stream_launch.  Shared by the
// pe_api*.cu translation units (the entry points are split over several files only so that nvcc --threads compiles the
// SVD kernels, the single-pass step, the fused map+metric kernels and the light ops side by side).
#pragma once
#include "pe_ops.cuh"

namespace pe {

template <class Op>
Params<Op> make_params(const float* i0, long long stride0, const float* i1, const float* i2, float* o0, float* o1,
                       long long n, double* sums, int* flag, typename Op::Extra ex) {
    Params<Op> p;
    p.in[0] = i0; p.in[1] = i1; p.in[2] = i2;
    p.out[0] = o0; p.out[1] = o1;
    p.n = n;
    p.in0_stride = stride0;
    p.sums = sums;
    p.mean32 = nullptr;
    p.mean_scale = 0.0;
    p.mean64 = nullptr;
    p.flag = flag;
    p.flag_to_sums = 0;
    p.sched = nullptr;  // assigned by stream_launch
    p.ex = ex;
    return p;
}

template <class Op>
int run(const float* i0, long long stride0, const float* i1, const float* i2, float* o0, float* o1, long long n,
        double* sums, int* flag, typename Op::Extra ex, void* stream, float* mean32 = nullptr, double mean_scale = 0.0,
        double* mean64 = nullptr, void* scratch = nullptr, bool flag_to_sums = false) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n < 0) return fail(PE_ERR_BAD_ARG, "negative record count");
    // natural alignment of every array (a misaligned access would be a sticky device fault, not an error code)
    if ((reinterpret_cast<uintptr_t>(i0) | reinterpret_cast<uintptr_t>(i1) | reinterpret_cast<uintptr_t>(i2) |
         reinterpret_cast<uintptr_t>(o0) | reinterpret_cast<uintptr_t>(o1) | reinterpret_cast<uintptr_t>(flag)) & 3u)
        return fail(PE_ERR_BAD_ARG, "array not aligned to 4 bytes");
    if (reinterpret_cast<uintptr_t>(sums) & 7u) return fail(PE_ERR_BAD_ARG, "float64 accumulator not aligned to 8 bytes");
    if (Op::NRED > 0) {
        if (!sums) return fail(PE_ERR_BAD_ARG, "null accumulator pointer");
        if (n == 0) {  // an empty batch still defines its sums; otherwise the kernel's last CTA writes them
            cudaError_t e = cudaMemsetAsync(sums, 0, sizeof(double) * Op::NRED, st);
            if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(sum)");
        }
    }
    if (Op::FLAG) {
        if (!flag && !flag_to_sums) return fail(PE_ERR_BAD_ARG, "null flag pointer");
        if (n == 0) {  // otherwise the kernel's last CTA writes the flag (its bits meet in the launch slot)
            cudaError_t e = flag ? cudaMemsetAsync(flag, 0, sizeof(int), st) : cudaSuccess;
            if (e == cudaSuccess && flag_to_sums) e = cudaMemsetAsync(sums + Op::NRED, 0, sizeof(double), st);
            if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(flag)");
        }
    }
    Params<Op> p = make_params<Op>(i0, stride0, i1, i2, o0, o1, n, sums, flag, ex);
    p.flag_to_sums = flag_to_sums ? 1 : 0;
    if (mean32) {
        if (reinterpret_cast<uintptr_t>(mean32) & 3u) return fail(PE_ERR_BAD_ARG, "array not aligned to 4 bytes");
        if (n == 0) {  // the mean of an empty batch is NaN, as torch's (all-ones bits are a quiet NaN)
            cudaError_t e = cudaMemsetAsync(mean32, 0xFF, sizeof(float), st);
            if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(mean)");
        }
        p.mean32 = mean32;
        p.mean_scale = mean_scale;
    }
    if (mean64) {
        if (reinterpret_cast<uintptr_t>(mean64) & 7u) return fail(PE_ERR_BAD_ARG, "float64 mean not aligned to 8 bytes");
        if (n == 0) {
            cudaError_t e = cudaMemsetAsync(mean64, 0xFF, sizeof(double), st);
            if (e != cudaSuccess) return fail_cuda(e, "cudaMemsetAsync(mean)");
        }
        p.mean64 = mean64;
        p.mean_scale = mean_scale;
    }
    return stream_launch<Op>(p, st, scratch);
}

}  // namespace pe
