// Shared device/host helpers for the gmat_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <atomic>
#include "../../include/gmat_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "gmat_b200 kernels are written for sm_100a (B200) only"
#endif

namespace gb {

// ---------------------------------------------------------------- host: errors, launch count
void set_error(const char *fmt, ...);
void set_engine_note(const char *fmt, ...);   // which engine a drop-in symbol ran (gmat_b200_last_engine_note)
extern std::atomic<long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define GB_CUDA(call)                                                                          \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            gb::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,                 \
                          cudaGetErrorString(e_));                                             \
            return GMAT_B200_ERR_CUDA;                                                         \
        }                                                                                      \
    } while (0)

#define GB_CHECK_ARG(cond, msg)                                                                \
    do {                                                                                       \
        if (!(cond)) {                                                                         \
            gb::set_error("bad argument: %s (%s)", msg, #cond);                                \
            return GMAT_B200_ERR_ARG;                                                          \
        }                                                                                      \
    } while (0)

#define GB_LAUNCH_CHECK()                                                                      \
    do {                                                                                       \
        cudaError_t e_ = cudaGetLastError();                                                   \
        if (e_ != cudaSuccess) {                                                               \
            gb::set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__,             \
                          cudaGetErrorString(e_));                                             \
            return GMAT_B200_ERR_CUDA;                                                         \
        }                                                                                      \
        gb::count_launch();                                                                    \
    } while (0)

int sm_count();   // cached multiprocessor count of the current device
void keep_scratch_pool();   // make the default stream-ordered pool of the current device retain freed scratch

// ---------------------------------------------------------------- device helpers
constexpr int BLOCK_ROWS = 128;   // row-block of Z'PZ in the packed layout / exact kernel M tile

__host__ __device__ inline long long round_up(long long x, long long m) { return (x + m - 1) / m * m; }

#ifdef __CUDACC__

// 2-bit code of individual k in a SNP row
__device__ __forceinline__ int code_at(const unsigned char *row, int k) {
    return (row[k >> 2] >> ((k & 3) * 2)) & 3;
}

// FP64 tensor-core MMA, the native Blackwell shape (SASS: DMMA.8x8x4).
//   A (8x4, row):  lane holds A[lane>>2][lane&3]
//   B (4x8, col):  lane holds B[lane&3][lane>>2]
//   C (8x8):       lane holds C[lane>>2][2*(lane&3) + {0,1}]
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// ---- mbarrier + TMA bulk copy (cp.async.bulk, SASS: UBLKCP) -------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// blocks until the phase with the given parity has completed (try_wait is a HW-suspended wait, not a
// busy poll: SASS SYNCS.PHASECHK.TRANS64.TRYWAIT)
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// warp-collective wait: lane 0 polls, the rest of the warp joins at __syncwarp (which orders memory among the
// lanes, so the acquire of lane 0 covers the warp).  32 lanes polling the same barrier word would cost 32
// serialized shared-memory transactions per poll.
__device__ __forceinline__ void mbar_wait_warp(uint64_t *bar, unsigned parity) {
    if ((threadIdx.x & 31) == 0) {
        while (!mbar_try_wait(bar, parity)) {
        }
    }
    __syncwarp();
}
// plain arrive (release semantics at CTA scope): one pending count of the current phase
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
// named barrier for a subset of the CTA's warps (id 1..15; id 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}
// global -> shared bulk copy through the TMA engine; completion is signalled on `bar` (tx bytes).
// bytes must be a multiple of 16, both addresses 16-byte aligned.
__device__ __forceinline__ void tma_bulk_g2s(void *smem_dst, const void *gmem_src, unsigned bytes,
                                             uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::
            "r"(smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// make generic-proxy smem reads (LDS) that precede this point visible/ordered before a later
// async-proxy write (TMA refill) to the same buffer
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}

__device__ __forceinline__ double shfl_xor_f64(double v, int mask) {
    return __shfl_xor_sync(0xffffffffu, v, mask);
}

#endif  // __CUDACC__
}  // namespace gb
