// Staging primitives for the rotation-mapping kernels (sm_100a): TMA bulk copies between HBM and shared memory,
// the mbarrier that tracks them, and conflict-free moves of one W-word record between shared memory and the
// registers of one thread.
//
// Why this exists: a record is 3, 4, 6 or 9 floats (12..36 bytes).  A thread that loads "its" record straight
// from HBM produces 36-byte-strided scalar traffic.  Instead one elected thread asks the TMA unit to move a whole
// contiguous tile of T records (T*W*4 bytes, a multiple of 16) into shared memory as ONE bulk transaction
// (SASS: UBLKCP), and every thread then picks its record out of shared memory.  With W odd (3, 9) the word stride
// is coprime with the 32 banks, W = 6 is read as 3 x 64-bit and W = 4 as 1 x 128-bit: all conflict-free.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pe {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
    // make the initialised barriers visible to the async proxy (the TMA unit) before the first bulk copy
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// ---- TMA bulk copies (1-D, no tensor map) --------------------------------------------------------------------
// L2 policy for data that is touched exactly once.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// global -> shared, completion counted in bytes on `bar`.  dst/src 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar,
                                         uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
// shared -> global, tracked by the issuing thread's bulk async-group.
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
                 "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until at most N of this thread's bulk groups still READ their shared-memory source
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// order this thread's generic-proxy shared-memory writes before later async-proxy (TMA) reads of them
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- one record: shared memory <-> registers -------------------------------------------------------------------
// `s` points at word 0 of the tile, `t` is the record index inside the tile.
template <int W>
__device__ __forceinline__ void lds_record(const float* s, int t, float* r) {
    if constexpr (W % 4 == 0) {
        const float4* p = reinterpret_cast<const float4*>(s + W * t);
#pragma unroll
        for (int i = 0; i < W / 4; ++i) {
            float4 v = p[i];
            r[4 * i] = v.x; r[4 * i + 1] = v.y; r[4 * i + 2] = v.z; r[4 * i + 3] = v.w;
        }
    } else if constexpr (W % 2 == 0) {
        const float2* p = reinterpret_cast<const float2*>(s + W * t);
#pragma unroll
        for (int i = 0; i < W / 2; ++i) {
            float2 v = p[i];
            r[2 * i] = v.x; r[2 * i + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < W; ++i) r[i] = s[W * t + i];
    }
}
template <int W>
__device__ __forceinline__ void sts_record(float* s, int t, const float* r) {
    if constexpr (W % 4 == 0) {
        float4* p = reinterpret_cast<float4*>(s + W * t);
#pragma unroll
        for (int i = 0; i < W / 4; ++i) p[i] = make_float4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
    } else if constexpr (W % 2 == 0) {
        float2* p = reinterpret_cast<float2*>(s + W * t);
#pragma unroll
        for (int i = 0; i < W / 2; ++i) p[i] = make_float2(r[2 * i], r[2 * i + 1]);
    } else {
#pragma unroll
        for (int i = 0; i < W; ++i) s[W * t + i] = r[i];
    }
}

// ---- reductions ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace pe
