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

}  // namespace pe
