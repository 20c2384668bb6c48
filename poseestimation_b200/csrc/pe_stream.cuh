// The record-streaming engine: one persistent kernel template that every per-record op of the path runs in.
//
// An Op maps up to three input records (WI0, WI1, WI2 words) to up to two output records (WO0, WO1 words) and
// optionally adds into NRED double accumulators and an int flag, all for ONE record held in registers
// (pe_math.cuh).  The engine owns everything else:
//
//   HBM --TMA bulk copy--> smem ring (STAGES deep) --LDS--> registers --Op::run--> registers --STS--> smem
//       --TMA bulk store--> HBM
//
// * grid = (#SMs x resident CTAs/SM) persistent CTAs that draw tiles from ONE global counter (atomicAdd), in order:
//   at any moment the whole grid streams through one contiguous window of HBM, a CTA that starts late or shares its SM
//   with another kernel (the NCCL all-reduce of the loss, overlapped with the next step) simply draws fewer tiles, and
//   there is no last-wave quantisation (2^24 matrices = 32768 tiles over 444 CTAs would be 74 vs 73 tiles per CTA).
//   The index for the refill of a stage is drawn one iteration ahead, so the atomic's latency hides behind the math;
//   the last CTA to finish resets the counter pair for the next launch that uses the slot.
// * a tile is T consecutive records; one elected thread issues one bulk copy per input array per tile and the
//   copies complete on the stage's mbarrier (byte-counted).  STAGES-1 tiles are always in flight per CTA.
// * outputs are double-buffered in shared memory and leave with bulk stores tracked by bulk async-groups.
// * the last, partial tile (n % T records) is moved with plain loads/stores inside the same kernel.
// * inputs that are not 16-byte aligned, or whose row stride is not the record width (the `out[:, :9]` view of
//   a (B, 12) tensor, /root/reference/3D-Pose/main2.py:94), run the `direct` kernel: still on the GPU, no TMA.
#pragma once
#include <mutex>
#include <thread>
#include <unordered_map>

#include "pe_common.h"
#include "pe_math.cuh"
#include "pe_stage.cuh"

namespace pe {

constexpr int kTile = 256;  // records per tile = threads per CTA
// Output tiles of the ops that do not store in place wait in a ring of kOutBufs shared-memory buffers for their bulk
// stores.  With THREE buffers a tile needs ONE __syncthreads: the barrier that says "everyone is done with the input
// stage" also publishes the output tile, because the buffer written in iteration i+1 was last read by the store of
// iteration i-2, which thread 0 has waited for before it arrived at the barrier of iteration i.  (With two buffers the
// write has to sit between two barriers; ncu showed the write-heavy maps stalled 5-11 cycles per issue on them.)
#ifndef PE_OUT_BUFS
#define PE_OUT_BUFS 3
#endif
constexpr int kOutBufs = PE_OUT_BUFS;

template <class Op>
struct Params {
    const float* in[3];
    float* out[2];
    long long n;           // records
    long long in0_stride;  // words between consecutive records of in[0] (direct kernel only)
    double* sums;          // NRED results (device): written once, by the last CTA to finish
    float* mean32;         // optional: (float)(sums[0] * mean_scale), written by the same CTA - the float32 batch mean a
    double mean_scale;     // loss function returns, without a follow-up launch that divides and casts
    double* mean64;        // optional: sums[NRED - 1] * mean_scale in float64 (the mean angle error of the metrics step)
    int* flag;             // WRITTEN by the last CTA: the OR of the per-record flags (they meet in the slot)
    int flag_to_sums;      // also store that OR as a double behind the NRED sums: [sum |.|_F, sum theta, flag] is the
                           // vector the ranks all-reduce
    unsigned int* sched;   // this launch's slot (set by stream_launch): {next tile to hand out, CTAs finished, flag bits,
                           // -, two float64 accumulators}; all zero when idle, re-armed by the last CTA
    typename Op::Extra ex;
};

// ring depth so that a CTA stays inside `budget` bytes of shared memory
constexpr int stages_raw(int witot, int wotot, int budget, int nbuf) {
    return (budget - nbuf * kTile * 4 * wotot - 512) / (kTile * 4 * witot);
}
constexpr int stages_for(int witot, int wotot, int budget = 54 * 1024, int nbuf = 2) {
    int s = stages_raw(witot, wotot, budget, nbuf);
    return s < 2 ? 2 : (s > 6 ? 6 : s);
}

// WIDE: in[0] is the `out[:, :9]` view of a (B, 12) network output (/root/reference/3D-Pose/main2.py:94,
// Iterative/utility.py:105).  Its rows are contiguous in memory, so whole 48-byte rows travel by the same bulk copies
// (TR x 48 B per tile) and a thread reads its row as 3 x LDS.128 (12-word stride: conflict-free for 128-bit accesses,
// a quarter-warp covers all 32 banks) and uses the first nine words.  33 % more read bytes on that one array instead of
// the direct kernel's per-thread 36-byte-strided loads.
constexpr int kWideRow = 12;
// PE_WIDE_SPLIT=1 (default): a wide op that has a second input (R_true, the upstream gradient) does not keep it in the
// stages.  That input is only read AFTER the decomposition, two thirds into a tile's arithmetic, so it lives in ONE
// buffer of its own: its bulk copy is issued at the top of the iteration that needs it and awaited by the lazy source
// itself (PairSrcLate); the in-place output is written over it and stored from there.  The ring then holds 12-word rows
// only: 2 x 24.5 KB + 18.4 KB = 67 KB per CTA instead of 2 x 43 KB, i.e. three CTAs per SM like the dense kernels.
#ifndef PE_WIDE_SPLIT
#define PE_WIDE_SPLIT 1
#endif
template <class Op, bool WIDE = false>
struct Geometry {
    static constexpr bool SPLIT = WIDE && (PE_WIDE_SPLIT != 0) && Op::WI1 > 0 && Op::WI2 == 0;
    __host__ __device__ static constexpr int wi(int k) { return k == 0 ? Op::WI0 : (k == 1 ? Op::WI1 : Op::WI2); }
    __host__ __device__ static constexpr int wo(int k) { return k == 0 ? Op::WO0 : Op::WO1; }
    // words per record INSIDE A STAGE (differs from wi only for the wide in[0])
    __host__ __device__ static constexpr int sw(int k) { return (k == 0 && WIDE) ? kWideRow : wi(k); }
    static constexpr int WITOT = sw(0) + sw(1) + sw(2);
    static constexpr int SWTOT = SPLIT ? sw(0) : WITOT;        // words per record inside a stage of the ring
    static constexpr int LATEW = SPLIT ? sw(1) : 0;             // words per record in the one-deep buffer of in[1]
    static constexpr int WOTOT = Op::WO0 + Op::WO1;
    // in-place output: the op's single output record replaces, inside the stage, the thread's own record of in[0] -
    // or, when in[0] is wide, of in[1] (R_true or the upstream gradient: same width as the output, read once, by
    // the thread that owns the record)
    static constexpr bool IN_PLACE0 = !WIDE && Op::RPT == 2 && Op::WO0 == Op::WI0 && Op::WO1 == 0;
    static constexpr bool IN_PLACE1 = WIDE && Op::RPT == 2 && Op::WO0 == Op::WI1 && Op::WO1 == 0;
    // a wide op WITHOUT a second input (the forward): the dense output tile is written over the head of the stage the
    // rows came in - other threads' rows, hence one more CTA barrier between the last read and the first write - and
    // stored from there: no output ring, three stages of rows, three CTAs per SM
    static constexpr bool IN_PLACE0W = WIDE && (PE_WIDE_SPLIT != 0) && Op::RPT == 2 && Op::WI1 == 0 && Op::WI2 == 0 &&
                                       Op::WO0 > 0 && Op::WO0 <= kWideRow && Op::WO1 == 0;
    static constexpr bool IN_PLACE = IN_PLACE0 || IN_PLACE1 || IN_PLACE0W;
    static constexpr int OUTBUFS = WIDE ? 3 : Op::OUTBUFS;
    static constexpr int NOBUF = (WOTOT > 0 && !IN_PLACE) ? OUTBUFS : 0;
    // wide stages are 21 words per record pair-half: two CTAs per SM (110 KB each)
    static constexpr int STAGES =
        SPLIT  ? stages_for(Op::RPT * SWTOT, 0, 75 * 1024 - kTile * Op::RPT * 4 * LATEW, 2)
        : IN_PLACE0W ? stages_for(Op::RPT * WITOT, 0, 75 * 1024, 2)
        : WIDE ? stages_for(Op::RPT * WITOT, IN_PLACE ? 0 : Op::RPT * WOTOT, 110 * 1024, NOBUF >= 3 ? NOBUF : 2)
               : Op::STAGES;
    static_assert(!SPLIT || IN_PLACE1 || WOTOT == 0, "a split wide op stores in place or not at all");
    // How an output tile leaves shared memory (one-barrier path).  TMA bulk store, except for read-heavy ops whose output
    // is less than half of their input: there every thread drains the tile with plain 128-bit streaming stores.
    // Measured on one box, two repetitions each (r1m): angle_error_fwd with per-record float64 output 0.218 -> 0.203 ms,
    // the map pullbacks unchanged (+-0.7 %); with plain stores for EVERY op the write-heavy forward maps are 2-6 % slower.
#ifndef PE_NO_PLAIN_DRAIN
    static constexpr bool PLAIN_DRAIN = 2 * WOTOT < WITOT;
#else
    static constexpr bool PLAIN_DRAIN = false;
#endif
    static constexpr int TR = kTile * Op::RPT;  // records per tile (RPT records per thread)
    static constexpr bool DYNAMIC = Op::RPT == 2;  // tiles drawn from a global counter (see stream_tma_kernel)
    static constexpr bool SLOT = DYNAMIC || Op::NRED > 0 || Op::FLAG;  // the launch needs a slot (scheduler / accumulators / flag)
    static constexpr int kBarBytes = 128;
    static constexpr int kStageBytes = TR * 4 * SWTOT;
    static constexpr int kLateBytes = TR * 4 * LATEW;
    static constexpr int kOutBytes = TR * 4 * WOTOT;
    static constexpr int kSmemBytes =
        kBarBytes + STAGES * kStageBytes + kLateBytes + NOBUF * kOutBytes + 32 * 8 * (Op::NRED > 0 ? Op::NRED : 1);
};

template <int N>
struct RegArr {
    float v[N > 0 ? N : 1];
};

template <int N>
struct PairArr {
    F2 v[N > 0 ? N : 1];
};
// records t and t + stride of a tile -> the two halves of F2 registers
template <int W>
__device__ __forceinline__ void lds_pair(const float* s, int t, int stride, F2* r) {
    float a[W], b[W];
    lds_record<W>(s, t, a);
    lds_record<W>(s, t + stride, b);
#pragma unroll
    for (int i = 0; i < W; ++i) r[i] = F2{a[i], b[i]};
}

template <int W>
__device__ __forceinline__ void sts_pair(float* s, int t, int stride, const F2* r) {
    float a[W], b[W];
#pragma unroll
    for (int i = 0; i < W; ++i) { a[i] = r[i].x; b[i] = r[i].y; }
    sts_record<W>(s, t, a);
    sts_record<W>(s, t + stride, b);
}

// Sums: one atomic per CTA into the slot's accumulators (zero when the launch starts), and the last CTA to finish moves
// them to the caller's array and re-arms the slot - so a reducing launch is ONE stream operation (no memset of the
// caller's accumulators in front of it).  Also re-arms the tile counter of the dynamic scheduler: every CTA has drawn
// its last index before it gets here.
template <class Op>
__device__ __forceinline__ void block_reduce_and_publish(double* red, int flag, double* scratch, const Params<Op>& p) {
    static_assert(Op::NRED <= 2, "a slot holds two accumulators");
    unsigned int* slot = p.sched;
    double* acc = reinterpret_cast<double*>(slot + 4);
    if constexpr (Op::NRED > 0) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
        for (int k = 0; k < Op::NRED; ++k) {
            double v = warp_sum(red[k]);
            if (lane == 0) scratch[warp * Op::NRED + k] = v;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < Op::NRED; ++k) {
                double v = 0.0;
                for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += scratch[w * Op::NRED + k];
                atomicAdd(acc + k, v);
            }
        }
    }
    if constexpr (Op::FLAG) {
        int any = __syncthreads_or(flag);
        if (threadIdx.x == 0 && any) atomicOr(slot + 2, (unsigned int)any);
    }
    if constexpr (Geometry<Op>::SLOT) {
        if (threadIdx.x == 0) {
            __threadfence();  // this CTA's adds are visible before its ticket
            if (atomicAdd(slot + 1, 1u) == gridDim.x - 1) {
                __threadfence();
#pragma unroll
                for (int k = 0; k < Op::NRED; ++k) {
                    const double v = __ldcg(acc + k);  // L2: where the other CTAs' atomics landed
                    p.sums[k] = v;
                    if (p.mean32 && k == 0) p.mean32[0] = (float)(v * p.mean_scale);
                    if (p.mean64 && k == Op::NRED - 1) p.mean64[0] = v * p.mean_scale;
                    acc[k] = 0.0;
                }
                if constexpr (Op::FLAG) {
                    const int f = (int)__ldcg(slot + 2);
                    if (p.flag) *p.flag = f;
                    if (p.flag_to_sums) p.sums[Op::NRED] = (double)f;
                    slot[2] = 0u;
                }
                slot[0] = 0u;
                slot[1] = 0u;
                __threadfence();
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Lazy source of a record pair (records t and t + stride of one input tile): the op decides WHEN the registers
// are filled, so inputs that are only needed after the SVD (R_true, the second read of M) do not occupy registers
// while the Jacobi rotations run.  This is what lets the two-matrices-per-thread kernels stay under 85 registers.
template <int W>
struct PairSrc {
    const float* s;
    int t, stride;
    __device__ __forceinline__ void load(F2* dst) const { lds_pair<W>(s, t, stride, dst); }
};

// the same for a wide in[0]: 12-word rows in the stage, the first nine are the matrix.  Each row is read as three
// 128-bit loads (conflict-free at a 12-word stride).  The third one is spelled in PTX: only its first word is used, and
// the compiler otherwise narrows it to a 32-bit LDS, which at this stride is a 4-way bank conflict (ncu source page of
// r2c: 786 432 excessive wavefronts per launch on exactly those two instructions).
__device__ __forceinline__ float lds128_first(const float* p) {
    float x, y, z, w;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(smem_u32(p)) : "memory");
    return x;
}
struct PairSrcWide {
    const float* s;
    int t, stride;
    __device__ __forceinline__ void load(F2* dst) const {
        float a[8], b[8];
        lds_record<8>(s + 4 * t, t, a);               // words 0..7 of row t (row stride 12 = 8 + 4)
        lds_record<8>(s + 4 * (t + stride), t + stride, b);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = F2{a[i], b[i]};
        dst[8] = F2{lds128_first(s + kWideRow * t + 8), lds128_first(s + kWideRow * (t + stride) + 8)};
    }
};

// in[1] of a split wide op: the tile arrives in a buffer of its own while the decomposition runs; the first load waits
// for it (every thread by itself - an mbarrier wait is a shared-memory poll, no CTA barrier)
template <int W>
struct PairSrcLate {
    const float* s;
    int t, stride;
    uint64_t* bar;
    uint32_t parity;
    bool wait;  // false for the ragged last tile, which was written with plain loads
    __device__ __forceinline__ void load(F2* dst) const {
        if (wait) mbar_wait(bar, parity);
        lds_pair<W>(s, t, stride, dst);
    }
};

// A wide stage WITH in[1] inside (PE_WIDE_SPLIT=0, or no second input) is up to 21 words per record: two CTAs fit an SM
// whatever the register count, so those instantiations are compiled for two (up to 128 registers: no spills) instead
// of the dense kernels' three (80 registers).
#ifndef PE_WIDE_MINB
#define PE_WIDE_MINB 2
#endif
template <class Op, bool WIDE>
constexpr int tma_min_blocks() {
    return WIDE && !Geometry<Op, WIDE>::SPLIT && !Geometry<Op, WIDE>::IN_PLACE0W && Op::MINB > PE_WIDE_MINB ? PE_WIDE_MINB
                                                                                                              : Op::MINB;
}
template <class Op, bool WIDE = false>
__global__ void __launch_bounds__(kTile, (tma_min_blocks<Op, WIDE>())) stream_tma_kernel(const Params<Op> p) {
    using G = Geometry<Op, WIDE>;
    constexpr int T = kTile, TR = G::TR, S = G::STAGES;
    constexpr bool IP = G::IN_PLACE;  // the output record overwrites the thread's own in[0] (wide: in[1]) record in the stage
    pdl_wait();               // programmatic dependent launch (pe_common.h): the predecessor is complete from here on
    pdl_launch_dependents();  // and the successor may be scheduled as soon as CTAs of this grid retire
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);
    float* s_in = reinterpret_cast<float*>(smem_raw + G::kBarBytes);
    float* s_late = s_in + S * TR * G::SWTOT;                 // SPLIT: the one buffer of in[1] (and of the in-place output)
    float* s_out = s_late + TR * G::LATEW;
    uint64_t* bar_late = &bars[S];
    double* s_red = reinterpret_cast<double*>(s_out + G::NOBUF * TR * G::WOTOT);

    const int tid = threadIdx.x;
    const long long n = p.n;
    const long long ntiles = (n + TR - 1) / TR;
    // tiles [0, nfull) are complete and move by TMA.  Wide rows: the last tile always takes the plain-load path, which
    // stops after the ninth word of the last row - the three words behind it need not exist
    const long long nfull = WIDE ? (n - 1) / TR : n / TR;
    unsigned int* s_tile = reinterpret_cast<unsigned int*>(smem_raw + 64);  // tile held (or awaited) by each stage
    uint64_t policy = 0;
    // Tile hand-out (thread 0 only).  DYN: one global counter.  Otherwise the static round-robin c, c + grid, ...:
    // same code path, the "counter" is a register.  The heavy (two-records-per-thread SVD) ops are dynamic; the light
    // elementwise ops finish a 256-record tile in well under a microsecond and 65536 same-address atomics per launch
    // would serialise (measured: 0.144 -> 0.204 ms for euler_fwd), so they keep the static order.
    constexpr bool DYN = G::DYNAMIC;
    unsigned int static_next = blockIdx.x;
    auto draw = [&]() -> unsigned int {
        if constexpr (DYN) {
            return atomicAdd(p.sched, 1u);
        } else {
            const unsigned int t = static_next;
            static_next += gridDim.x;
            return t;
        }
    };

    auto issue = [&](int s, long long tile) {
        // one elected thread: arm the barrier with the byte count, then one bulk copy per input array
        mbar_arrive_expect_tx(&bars[s], (uint32_t)G::kStageBytes);
        float* dst = s_in + s * TR * G::SWTOT;
        int off = 0;
#pragma unroll
        for (int k = 0; k < (G::SPLIT ? 1 : 3); ++k) {
            if (G::sw(k) > 0) {
                bulk_g2s(dst + off, p.in[k] + tile * TR * G::sw(k), (uint32_t)(TR * 4 * G::sw(k)), &bars[s], policy);
                off += TR * G::sw(k);
            }
        }
    };

    auto issue_late = [&](long long tile) {  // SPLIT: in[1] of the tile that is being worked on
        mbar_arrive_expect_tx(bar_late, (uint32_t)G::kLateBytes);
        bulk_g2s(s_late, p.in[1] + tile * TR * G::sw(1), (uint32_t)G::kLateBytes, bar_late, policy);
    };

    if (tid == 0) {
        policy = l2_evict_first_policy();
#pragma unroll
        for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
        if constexpr (G::SPLIT) mbar_init(bar_late, 1);
        mbar_init_fence();
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const unsigned int t = draw();
            if constexpr (DYN) s_tile[s] = t;
            if (t < nfull) issue(s, t);
        }
    }
    // the index that will refill the stage of the current iteration is drawn ONE ITERATION AHEAD (tid 0 only), so the
    // round trip of the atomic is over long before the index is needed, also for ops with next to no arithmetic
    unsigned int ahead = 0;
    if (tid == 0) ahead = draw();
    __syncthreads();

    double red[Op::NRED > 0 ? Op::NRED : 1];
#pragma unroll
    for (int k = 0; k < (Op::NRED > 0 ? Op::NRED : 1); ++k) red[k] = 0.0;
    int flag = 0;

    int s = 0;
    uint32_t parity = 0;
    uint32_t late_parity = 0;
    int obuf = 0;
    bool refill = false;         // IP mode: the stage vacated by the previous iteration awaits its next tile
    unsigned int refill_tile = 0;
    int refill_stage = 0;
    long long static_tile = blockIdx.x;  // !DYN: the processing order is known without the shared-memory hand-over
    for (;;) {
        // DYN: the same for every thread, written by thread 0 before a __syncthreads
        const long long tile = DYN ? (long long)s_tile[s] : static_tile;
        if (tile >= ntiles) break;
        static_tile += gridDim.x;
        const unsigned int drawn = ahead;               // for the refill of THIS stage
        if (tid == 0) ahead = draw();                   // for the next iteration's
        const long long base = tile * TR;
        const bool full = tile < nfull;
        const int cnt = full ? TR : (int)(n - base);
        float* si = s_in + s * TR * G::SWTOT;
        float* so = (G::IN_PLACE0 || G::IN_PLACE0W) ? si
                                 : (G::IN_PLACE1 ? (G::SPLIT ? s_late : si + TR * G::sw(0)) : s_out + obuf * TR * G::WOTOT);

        if constexpr (G::SPLIT) {
            // in[1] of THIS tile: its buffer is free once the store of the previous tile has read it (in-place ops; the
            // others left it behind barrier (A)).  It is needed after the decomposition, microseconds from now.
            if (tid == 0) {
                if constexpr (IP) {
                    if (refill) bulk_wait_read<0>();
                }
                if (full) issue_late(tile);
            }
        } else if constexpr (IP) {
            // the stage used one iteration ago was stored from in place: once the TMA store has finished reading
            // it, refill it (prefetch distance: S - 1 tiles)
            if (tid == 0 && refill) {
                bulk_wait_read<0>();
                if constexpr (DYN) s_tile[refill_stage] = refill_tile;
                if (refill_tile < nfull) issue(refill_stage, refill_tile);
            }
        }
        if (full) {
            mbar_wait(&bars[s], parity);
        } else {
            // ragged tail: plain coalesced word copies into the same shared-memory layout
            if constexpr (G::SPLIT && IP) __syncthreads();  // thread 0 has waited for the store that was reading s_late
            int off = 0;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (G::sw(k) > 0) {
                    const float* src = p.in[k] + base * G::sw(k);
                    float* dst = (G::SPLIT && k == 1) ? s_late : si + off;
                    const int words = cnt * G::sw(k) - (G::sw(k) - G::wi(k));  // wide: not past the last matrix
                    for (int w = tid; w < words; w += T) dst[w] = src[w];
                    off += TR * G::sw(k);
                }
            }
            __syncthreads();
        }

        const float* si1 = G::SPLIT ? s_late : si + TR * G::sw(0);
        const float* si2 = si + TR * (G::sw(0) + G::sw(1));
        RegArr<Op::WO0 * Op::RPT> o0;
        RegArr<Op::WO1 * Op::RPT> o1;
        if constexpr (Op::RPT == 1) {
            RegArr<Op::WI0> i0;
            RegArr<Op::WI1> i1;
            RegArr<Op::WI2> i2;
            if constexpr (Op::WI0 > 0) lds_record<Op::WI0>(si, tid, i0.v);
            if constexpr (Op::WI1 > 0) lds_record<Op::WI1>(si1, tid, i1.v);
            if constexpr (Op::WI2 > 0) lds_record<Op::WI2>(si2, tid, i2.v);
            if (tid < cnt) Op::run(i0.v, i1.v, i2.v, o0.v, o1.v, red, flag, p.ex);
        } else {
            // two records per thread (tid and tid + T), one in each half of a packed fp32x2 register pair
            PairArr<Op::WO0> q0;
            PairArr<Op::WO1> q1;
            if constexpr (G::SPLIT)
                Op::run2(PairSrcWide{si, tid, T}, PairSrcLate<Op::WI1>{si1, tid, T, bar_late, late_parity, full},
                         PairSrc<Op::WI2>{si2, tid, T}, q0.v, q1.v, red, flag, tid < cnt, tid + T < cnt, p.ex);
            else if constexpr (WIDE)
                Op::run2(PairSrcWide{si, tid, T}, PairSrc<Op::WI1>{si1, tid, T}, PairSrc<Op::WI2>{si2, tid, T}, q0.v, q1.v,
                         red, flag, tid < cnt, tid + T < cnt, p.ex);
            else
                Op::run2(PairSrc<Op::WI0>{si, tid, T}, PairSrc<Op::WI1>{si1, tid, T}, PairSrc<Op::WI2>{si2, tid, T}, q0.v,
                         q1.v, red, flag, tid < cnt, tid + T < cnt, p.ex);
#pragma unroll
            for (int w = 0; w < Op::WO0; ++w) { o0.v[w] = q0.v[w].x; o0.v[Op::WO0 + w] = q0.v[w].y; }
#pragma unroll
            for (int w = 0; w < Op::WO1; ++w) { o1.v[w] = q1.v[w].x; o1.v[Op::WO1 + w] = q1.v[w].y; }
        }

        if constexpr (!IP && G::NOBUF >= 3) {
            // ONE barrier per tile (see kOutBufs): write the output tile, then the barrier, then thread 0 refills the
            // input stage and sends the output tile on its way
#pragma unroll
            for (int rr = 0; rr < Op::RPT; ++rr) {
                if constexpr (Op::WO0 > 0) sts_record<Op::WO0>(so, tid + rr * T, o0.v + rr * Op::WO0);
                if constexpr (Op::WO1 > 0) sts_record<Op::WO1>(so + TR * Op::WO0, tid + rr * T, o1.v + rr * Op::WO1);
            }
            if constexpr (G::PLAIN_DRAIN) {
                // read-heavy op (small output record): every thread drains the tile with 128-bit streaming stores; the
                // buffer is rewritten NOBUF tiles later, behind two more barriers, so there is no asynchronous wait
                __syncthreads();
                if (tid == 0) {
                    if constexpr (DYN) s_tile[s] = drawn;
                    if (drawn < nfull) issue(s, drawn);
                }
                if (full) {
                    if constexpr (Op::WO0 > 0) {
                        const float4* s4 = reinterpret_cast<const float4*>(so);
                        float4* g4 = reinterpret_cast<float4*>(p.out[0] + base * Op::WO0);
#pragma unroll
                        for (int w = tid; w < TR * Op::WO0 / 4; w += T) __stcs(g4 + w, s4[w]);
                    }
                    if constexpr (Op::WO1 > 0) {
                        const float4* s4 = reinterpret_cast<const float4*>(so + TR * Op::WO0);
                        float4* g4 = reinterpret_cast<float4*>(p.out[1] + base * Op::WO1);
#pragma unroll
                        for (int w = tid; w < TR * Op::WO1 / 4; w += T) __stcs(g4 + w, s4[w]);
                    }
                }
            } else {
                fence_async_smem();
                if (tid == 0) bulk_wait_read<G::NOBUF - 2>();  // all stores but the latest NOBUF-2 have finished reading their buffers
                __syncthreads();  // every thread is done with stage s (and has read s_tile[s]); `so` is complete
                if (tid == 0) {
                    if constexpr (DYN) s_tile[s] = drawn;
                    if (drawn < nfull) issue(s, drawn);
                    if (full) {
                        if constexpr (Op::WO0 > 0) bulk_s2g(p.out[0] + base * Op::WO0, so, (uint32_t)(TR * 4 * Op::WO0));
                        if constexpr (Op::WO1 > 0)
                            bulk_s2g(p.out[1] + base * Op::WO1, so + TR * Op::WO0, (uint32_t)(TR * 4 * Op::WO1));
                        bulk_commit();
                    }
                }
            }
            if (!full) {  // ragged last tile: plain coalesced word copies (no later tile reuses `so`)
                if constexpr (Op::WO0 > 0) {
                    float* dst = p.out[0] + base * Op::WO0;
                    for (int w = tid; w < cnt * Op::WO0; w += T) dst[w] = so[w];
                }
                if constexpr (Op::WO1 > 0) {
                    float* dst = p.out[1] + base * Op::WO1;
                    for (int w = tid; w < cnt * Op::WO1; w += T) dst[w] = so[TR * Op::WO0 + w];
                }
            }
            if (++obuf == G::NOBUF) obuf = 0;
        } else {
        if constexpr (!IP) {
            if constexpr (G::NOBUF > 0) {
                if (tid == 0) bulk_wait_read<1>();  // the store issued two tiles ago has finished reading `so`
            }
            __syncthreads();  // (A) every thread is done with stage s (and has read s_tile[s]); `so` is free
            if (tid == 0) {
                if constexpr (DYN) s_tile[s] = drawn;
                if (drawn < nfull) issue(s, drawn);
            }
        }
        if constexpr (G::WOTOT > 0) {
            // IP: a thread only ever reads its OWN records of in[0], and it has finished reading them, so it may
            // overwrite them without a barrier in between (wide rows: the dense output lands on other threads' rows)
            if constexpr (G::IN_PLACE0W) __syncthreads();
#pragma unroll
            for (int rr = 0; rr < Op::RPT; ++rr) {
                if constexpr (Op::WO0 > 0) sts_record<Op::WO0>(so, tid + rr * T, o0.v + rr * Op::WO0);
                if constexpr (Op::WO1 > 0) sts_record<Op::WO1>(so + TR * Op::WO0, tid + rr * T, o1.v + rr * Op::WO1);
            }
            if (full) {
                fence_async_smem();
                __syncthreads();  // (B)
                if (tid == 0) {
                    if constexpr (Op::WO0 > 0) bulk_s2g(p.out[0] + base * Op::WO0, so, (uint32_t)(TR * 4 * Op::WO0));
                    if constexpr (Op::WO1 > 0)
                        bulk_s2g(p.out[1] + base * Op::WO1, so + TR * Op::WO0, (uint32_t)(TR * 4 * Op::WO1));
                    bulk_commit();
                }
            } else {
                __syncthreads();
                if constexpr (Op::WO0 > 0) {
                    float* dst = p.out[0] + base * Op::WO0;
                    for (int w = tid; w < cnt * Op::WO0; w += T) dst[w] = so[w];
                }
                if constexpr (Op::WO1 > 0) {
                    float* dst = p.out[1] + base * Op::WO1;
                    for (int w = tid; w < cnt * Op::WO1; w += T) dst[w] = so[TR * Op::WO0 + w];
                }
            }
            obuf ^= 1;
        }
        }
        if constexpr (G::SPLIT) {
            if (full) late_parity ^= 1;
            if constexpr (IP) {
                // the rows of this stage have been read by everybody (barrier (B) / the tail's barrier is behind us) and the
                // store reads s_late, not the stage: refill it now, a whole tile ahead
                if (tid == 0) {
                    if constexpr (DYN) s_tile[s] = drawn;
                    if (drawn < nfull) issue(s, drawn);
                }
            }
        }
        if constexpr (IP) {
            refill = true;
            refill_tile = drawn;
            refill_stage = s;
        }
        if (++s == S) { s = 0; parity ^= 1; }
    }
    if constexpr (G::WOTOT > 0) {
        if (tid == 0) bulk_wait<0>();  // shared memory must outlive the stores that read it
    }
    block_reduce_and_publish<Op>(red, flag, s_red, p);
}

// ---------------------------------------------------------------------------------------------------------------
// Non-TMA path: arbitrary alignment and arbitrary row stride on in[0] (the `out[:, :9]` view of a (B, 12) network
// output, /root/reference/3D-Pose/main2.py:94).  Grid-stride, plain loads through L1.  The two-records-per-thread ops
// (the SVD family) keep their packed fp32x2 arithmetic here too: a thread takes records i and i + ceil(n/2), holds them
// in registers and hands them to the op's `run2` through the same lazy-source interface the TMA kernel uses (a PairSrc
// over a thread-local array; after inlining its loads are register moves).  Measured at 2^24 matrices: svd_orth on a
// stride-12 view 0.516 ms with one record per thread (scalar SVD), 0.347 ms packed; unaligned dense input 0.440 -> 0.334.
// Occupancy of the packed variant (measured on a stride-12 view of 2^24 matrices): the forward is fastest at 3 CTAs/SM
// (80 registers, a few spills: 0.346 ms against 0.400 at 2 CTAs), the ops that also hold a second input at 2 CTAs/SM
// (128 registers: the step 0.457 ms against 0.508 at 3 CTAs and 0.613 at 1).
template <class Op>
constexpr int direct_min_blocks() { return Op::RPT == 2 ? (Op::WI1 == 0 ? 3 : 2) : 1; }
template <class Op>
__global__ void __launch_bounds__(kTile, direct_min_blocks<Op>()) stream_direct_kernel(const Params<Op> p) {
    __shared__ double s_red[32 * (Op::NRED > 0 ? Op::NRED : 1)];
    double red[Op::NRED > 0 ? Op::NRED : 1];
#pragma unroll
    for (int k = 0; k < (Op::NRED > 0 ? Op::NRED : 1); ++k) red[k] = 0.0;
    int flag = 0;
    const long long step = (long long)gridDim.x * blockDim.x;
    if constexpr (Op::RPT == 2) {
        constexpr int W0 = Op::WI0, W1 = Op::WI1, W2 = Op::WI2;
        const long long half = (p.n + 1) / 2;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < half; i += step) {
            const bool v1 = i + half < p.n;
            const long long j = v1 ? i + half : i;  // an odd batch: the last thread's second lane repeats its first
            __align__(16) float l0[2 * (W0 > 0 ? W0 : 1)], l1[2 * (W1 > 0 ? W1 : 1)], l2[2 * (W2 > 0 ? W2 : 1)];
#pragma unroll
            for (int w = 0; w < W0; ++w) {
                l0[w] = __ldg(p.in[0] + i * p.in0_stride + w);
                l0[W0 + w] = __ldg(p.in[0] + j * p.in0_stride + w);
            }
#pragma unroll
            for (int w = 0; w < W1; ++w) { l1[w] = __ldg(p.in[1] + i * W1 + w); l1[W1 + w] = __ldg(p.in[1] + j * W1 + w); }
#pragma unroll
            for (int w = 0; w < W2; ++w) { l2[w] = __ldg(p.in[2] + i * W2 + w); l2[W2 + w] = __ldg(p.in[2] + j * W2 + w); }
            PairArr<Op::WO0> q0;
            PairArr<Op::WO1> q1;
            Op::run2(PairSrc<W0>{l0, 0, 1}, PairSrc<W1>{l1, 0, 1}, PairSrc<W2>{l2, 0, 1}, q0.v, q1.v, red, flag, true, v1,
                     p.ex);
#pragma unroll
            for (int w = 0; w < Op::WO0; ++w) p.out[0][i * Op::WO0 + w] = q0.v[w].x;
#pragma unroll
            for (int w = 0; w < Op::WO1; ++w) p.out[1][i * Op::WO1 + w] = q1.v[w].x;
            if (v1) {
#pragma unroll
                for (int w = 0; w < Op::WO0; ++w) p.out[0][j * Op::WO0 + w] = q0.v[w].y;
#pragma unroll
                for (int w = 0; w < Op::WO1; ++w) p.out[1][j * Op::WO1 + w] = q1.v[w].y;
            }
        }
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += step) {
            RegArr<Op::WI0> i0;
            RegArr<Op::WI1> i1;
            RegArr<Op::WI2> i2;
            RegArr<Op::WO0> o0;
            RegArr<Op::WO1> o1;
#pragma unroll
            for (int w = 0; w < Op::WI0; ++w) i0.v[w] = __ldg(p.in[0] + i * p.in0_stride + w);
#pragma unroll
            for (int w = 0; w < Op::WI1; ++w) i1.v[w] = __ldg(p.in[1] + i * Op::WI1 + w);
#pragma unroll
            for (int w = 0; w < Op::WI2; ++w) i2.v[w] = __ldg(p.in[2] + i * Op::WI2 + w);
            Op::run(i0.v, i1.v, i2.v, o0.v, o1.v, red, flag, p.ex);
#pragma unroll
            for (int w = 0; w < Op::WO0; ++w) p.out[0][i * Op::WO0 + w] = o0.v[w];
#pragma unroll
            for (int w = 0; w < Op::WO1; ++w) p.out[1][i * Op::WO1 + w] = o1.v[w];
        }
    }
    block_reduce_and_publish<Op>(red, flag, s_red, p);
}

// ---------------------------------------------------------------------------------------------------------------
struct LaunchCfg {
    int sms = 0;
    int ctas_per_sm = 0;
};

template <class Op, bool WIDE = false>
int stream_config(LaunchCfg* out) {
    static std::mutex mu;
    static LaunchCfg cache[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(PE_ERR_BAD_ARG, "device index out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (cache[dev].sms == 0) {
        using G = Geometry<Op, WIDE>;
        cudaDeviceProp prop;
        e = cudaGetDeviceProperties(&prop, dev);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetDeviceProperties");
        if (prop.major < 10) return fail(PE_ERR_NO_DEVICE, "this library is built for sm_100a (B200) only");
        e = cudaFuncSetAttribute(stream_tma_kernel<Op, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::kSmemBytes);
        if (e != cudaSuccess) return fail_cuda(e, "cudaFuncSetAttribute(smem)");
        int per_sm = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, stream_tma_kernel<Op, WIDE>, kTile, G::kSmemBytes);
        if (e != cudaSuccess) return fail_cuda(e, "cudaOccupancyMaxActiveBlocksPerMultiprocessor");
        if (per_sm < 1) return fail(PE_ERR_INTERNAL, "kernel does not fit on an SM");
        cache[dev].ctas_per_sm = per_sm;
        cache[dev].sms = prop.multiProcessorCount;
    }
    *out = cache[dev];
    return 0;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---------------------------------------------------------------------------------------------------------------
// Launch slots: {next tile, finished CTAs, -, -, two float64 accumulators} in device memory, 32 bytes each, all zero
// when idle (the last CTA of a launch re-arms its slot).  The one piece of device memory the library owns: kSlots x 32 B
// = 128 KB per device, allocated and zeroed by pe_init(device) (include/pe_rotation.h) - explicitly, once, synchronously,
// so that no launch ever allocates and a launch can be captured into a CUDA graph.  Who gets which slot:
//   * eager launches: ONE slot per (device, stream).  Launches on one stream never overlap - every kernel that uses a
//     slot starts with griddepcontrol.wait, which returns only when the preceding kernel of the stream has completed
//     and its writes are visible, and the non-PDL launches are serialised by the stream itself - so a stream can re-use
//     its slot for ever, and launches on different streams can never meet in one.  (Round 1 handed slots out of a ring
//     at enqueue time: a stalled stream plus 1024 launches elsewhere, or a replaying graph, could share a slot.)
//   * launches recorded by a stream capture: a slot of their own each, from a region eager launches never use and
//     that is never recycled, so a graph may replay while any stream runs eager launches.  (Two executable graphs
//     instantiated from the SAME captured graph share their slots: launch them on one stream, or capture twice.)
//   * a caller can bypass all of this by passing its own 32-byte zeroed scratch (pe_svd_orth_frobenius_step_metrics).
// Both regions are finite; running out is an error code (PE_ERR_INTERNAL), never a silently shared slot.
constexpr int kStreamSlots = 1024;
constexpr int kGraphSlots = 3072;
constexpr int kSlots = kStreamSlots + kGraphSlots;
struct SlotPool {
    unsigned int* base = nullptr;
    std::unordered_map<unsigned long long, int> by_stream;
    int graph_next = 0;
};
inline std::mutex& slot_mutex() {
    static std::mutex mu;
    return mu;
}
inline SlotPool* slot_pools() {
    static SlotPool pools[64];
    return pools;
}
// allocate + zero the pool of the given device (idempotent).  Synchronises the device once: the zeroes are in place
// before any stream, blocking or not, can run a kernel that reads them.
inline int slots_init(int dev) {
    if (dev < 0 || dev >= 64) return fail(PE_ERR_BAD_ARG, "device index out of range");
    std::lock_guard<std::mutex> lk(slot_mutex());
    SlotPool& sp = slot_pools()[dev];
    if (sp.base) return 0;
    int prev = 0;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    if (prev != dev) {
        e = cudaSetDevice(dev);
        if (e != cudaSuccess) return fail_cuda(e, "cudaSetDevice");
    }
    void* ptr = nullptr;
    e = cudaMalloc(&ptr, (size_t)kSlots * 32);
    if (e == cudaSuccess) e = cudaMemset(ptr, 0, (size_t)kSlots * 32);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e == cudaSuccess) sp.base = static_cast<unsigned int*>(ptr);
    if (prev != dev) cudaSetDevice(prev);
    if (e != cudaSuccess) return fail_cuda(e, "pe_init: slot pool");
    return 0;
}
inline int sched_slot(cudaStream_t st, unsigned int** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(PE_ERR_BAD_ARG, "device index out of range");
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    e = cudaStreamIsCapturing(st, &cap);
    if (e != cudaSuccess) return fail_cuda(e, "cudaStreamIsCapturing");
    const bool capturing = cap != cudaStreamCaptureStatusNone;
    if (!slot_pools()[dev].base) {
        if (capturing)
            return fail(PE_ERR_BAD_ARG, "pe_init(device) must be called before a launch is captured into a CUDA graph");
        if (int rc = slots_init(dev)) return rc;
    }
    std::lock_guard<std::mutex> lk(slot_mutex());
    SlotPool& sp = slot_pools()[dev];
    int idx;
    if (capturing) {
        if (sp.graph_next >= kGraphSlots) return fail(PE_ERR_INTERNAL, "out of launch slots for captured launches");
        idx = kStreamSlots + sp.graph_next++;
    } else {
        unsigned long long key = reinterpret_cast<unsigned long long>(st);
        if (st == cudaStreamPerThread)  // one handle, a different stream in every host thread
            key ^= (unsigned long long)std::hash<std::thread::id>()(std::this_thread::get_id()) << 8;
        auto it = sp.by_stream.find(key);
        if (it == sp.by_stream.end()) {
            if ((int)sp.by_stream.size() >= kStreamSlots) return fail(PE_ERR_INTERNAL, "out of launch slots (more than 1024 streams)");
            it = sp.by_stream.emplace(key, (int)sp.by_stream.size()).first;
        }
        idx = it->second;
    }
    *out = sp.base + (size_t)idx * 8;
    return 0;
}

// force_direct: test hook (PE_FORCE_DIRECT=1) so that the non-TMA kernel is covered on aligned inputs too
template <class Op, bool WIDE>
int stream_launch_tma(Params<Op>& p, cudaStream_t stream, long long ntiles) {
    using G = Geometry<Op, WIDE>;
    LaunchCfg cfg;
    if (int rc = stream_config<Op, WIDE>(&cfg)) return rc;
    long long grid = (long long)cfg.sms * cfg.ctas_per_sm;
    if (grid > ntiles) grid = ntiles;
    cudaError_t le = launch_pdl(stream_tma_kernel<Op, WIDE>, (unsigned)grid, (unsigned)kTile, (size_t)G::kSmemBytes, stream, p);
    if (le != cudaSuccess) return fail_cuda(le, Op::name());
    return 0;
}

template <class Op>
constexpr bool wide_ok() { return Op::RPT == 2 && Op::WI0 == 9 && Op::WIDE_OK; }

template <class Op>
int stream_launch(Params<Op> p, cudaStream_t stream, void* scratch = nullptr) {
    using G = Geometry<Op>;
    if (p.n < 0) return fail(PE_ERR_BAD_ARG, "negative record count");
    if (p.n == 0) return 0;
    for (int k = 0; k < 3; ++k)
        if (G::wi(k) > 0 && p.in[k] == nullptr) return fail(PE_ERR_BAD_ARG, "null input pointer");
    for (int k = 0; k < 2; ++k)
        if (G::wo(k) > 0 && p.out[k] == nullptr) return fail(PE_ERR_BAD_ARG, "null output pointer");
    if (Op::NRED > 0 && p.sums == nullptr) return fail(PE_ERR_BAD_ARG, "null accumulator pointer");
    if (Op::FLAG && p.flag == nullptr && !p.flag_to_sums) return fail(PE_ERR_BAD_ARG, "null flag pointer");
    if (p.in0_stride < G::wi(0)) return fail(PE_ERR_BAD_ARG, "row stride smaller than the record");
    // Dense, 16-byte aligned arrays take the TMA kernel; so does the SVD family on 12-word rows (wide stage).  Anything
    // else (other strides, unaligned bases) takes the direct kernel.  (Tried: keeping the dense TMA kernel and gathering
    // a strided in[0] tile by tile with plain coalesced loads - slower than the direct kernel, 0.49 vs 0.40 ms for
    // svd_orth on a stride-12 view of 2^24 matrices: the gather is not prefetched and costs a barrier per tile.)
    bool aligned = !force_direct();
    for (int k = 0; k < 3; ++k) aligned = aligned && (G::wi(k) == 0 || aligned16(p.in[k]));
    for (int k = 0; k < 2; ++k) aligned = aligned && (G::wo(k) == 0 || aligned16(p.out[k]));
    const bool dense = aligned && p.in0_stride == G::wi(0);
    const bool wide = aligned && wide_ok<Op>() && p.in0_stride == kWideRow && !no_wide();
    if constexpr (G::SLOT) {
        if (scratch) {
            if (reinterpret_cast<uintptr_t>(scratch) & 7u) return fail(PE_ERR_BAD_ARG, "scratch not aligned to 8 bytes");
            p.sched = static_cast<unsigned int*>(scratch);
        } else if (int rc = sched_slot(stream, &p.sched)) {
            return rc;
        }
    }
    const long long ntiles = (p.n + G::TR - 1) / G::TR;
    if (dense) {
        if (int rc = stream_launch_tma<Op, false>(p, stream, ntiles)) return rc;
    } else if (wide) {
        if constexpr (wide_ok<Op>()) {
            if (int rc = stream_launch_tma<Op, true>(p, stream, ntiles)) return rc;
        }
    } else {
        LaunchCfg cfg;
        if (int rc = stream_config<Op>(&cfg)) return rc;
        long long grid = (long long)cfg.sms * 8;
        if (grid > ntiles) grid = ntiles;
        stream_direct_kernel<Op><<<(unsigned)grid, kTile, 0, stream>>>(p);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, Op::name());
    count_launch();
    return 0;
}

}  // namespace pe
