// extern "C" pe_map_metric_fwd: representation -> rotation -> metric sum, one kernel per (map, metric) pair.
#include "pe_api_run.cuh"

using namespace pe;

namespace {
template <template <int> class Op>
int map_metric(int metric, const float* x, int64_t stride, const float* rt, Clamp c, double* sum, int* flag, int64_t n,
               void* stream) {
    switch (metric) {
        case PE_METRIC_FROBENIUS:
            return run<Op<0>>(x, stride, rt, nullptr, nullptr, nullptr, n, sum, nullptr, c, stream);
        case PE_METRIC_GEODESIC:
            return run<Op<1>>(x, stride, rt, nullptr, nullptr, nullptr, n, sum, nullptr, c, stream);
        case PE_METRIC_ANGLE_DEG:
            return run<Op<2>>(x, stride, rt, nullptr, nullptr, nullptr, n, sum, flag, c, stream);
    }
    return fail(PE_ERR_BAD_ARG, "unknown metric");
}

}  // namespace

extern "C" {

int pe_map_metric_fwd(int map, int metric, const float* x, int64_t x_stride, const float* r_true, float lo, float hi,
                      double* sum, int* flag, int64_t n, void* stream) {
    const Clamp c{lo, hi};
    switch (map) {
        case PE_MAP_SVD: return map_metric<SvdMetric>(metric, x, x_stride, r_true, c, sum, flag, n, stream);
        case PE_MAP_ORTHO6D: return map_metric<Ortho6dMetric>(metric, x, x_stride, r_true, c, sum, flag, n, stream);
        case PE_MAP_QUAT: return map_metric<QuatMetric>(metric, x, x_stride, r_true, c, sum, flag, n, stream);
        case PE_MAP_EULER: return map_metric<EulerMetric>(metric, x, x_stride, r_true, c, sum, flag, n, stream);
        case PE_MAP_ORTHO5D: return map_metric<Ortho5dMetric>(metric, x, x_stride, r_true, c, sum, flag, n, stream);
        case PE_MAP_EXPMAP: return map_metric<ExpmapMetric>(metric, x, x_stride, r_true, c, sum, flag, n, stream);
    }
    return fail(PE_ERR_BAD_ARG, "unknown map");
}

}  // extern "C"
