// Error convention, launch counter and test hooks shared by every translation unit of libpe_rotation.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>

#include "../../include/pe_rotation.h"

namespace pe {

inline char* last_error_buf() {
    static thread_local char buf[256] = "";
    return buf;
}
inline int fail(int code, const char* what) {
    snprintf(last_error_buf(), 256, "%s", what);
    return code;
}
inline int fail_cuda(cudaError_t e, const char* where) {
    snprintf(last_error_buf(), 256, "%s: %s", where, cudaGetErrorString(e));
    return (int)e;  // positive = cudaError_t
}
inline std::atomic<long long>& launch_counter() {
    static std::atomic<long long> c{0};
    return c;
}
inline void count_launch() { launch_counter().fetch_add(1, std::memory_order_relaxed); }
// PE_FORCE_DIRECT=1 routes aligned, dense inputs through the non-TMA kernel as well (tests only; still the GPU).
inline bool force_direct() {
    const char* v = getenv("PE_FORCE_DIRECT");
    return v && v[0] == '1';
}

// PE_NO_WIDE=1 sends 12-word-row inputs (the out[:, :9] view) to the direct kernel instead of the wide TMA stage (A/B runs).
inline bool no_wide() {
    static const bool off = [] { const char* v = getenv("PE_NO_WIDE"); return v && v[0] == '1'; }();
    return off;
}

// Programmatic dependent launch.  The streaming kernels and the scaling kernels are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization: their CTAs may be scheduled while the previous kernel of the
// stream is still draining, and block in `griddepcontrol.wait` (first statement of the kernel, before any global
// access) until that kernel has completed and its writes are visible - so the launch latency and the CTA start-up hide
// behind the predecessor's tail, whatever the predecessor is.  PE_NO_PDL=1 switches the attribute off (A/B runs).
inline bool pdl_enabled() {
    static const bool on = [] { const char* v = getenv("PE_NO_PDL"); return !(v && v[0] == '1'); }();
    return on;
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st,
                              Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}
#endif

}  // namespace pe
