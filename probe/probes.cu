// Measurement scaffolding, kept OUT of the product library: probe/libgmat_b200_probe.so holds the issue-rate
// microbenchmarks bench.py and scripts/ use as roofline denominators (int8 tcgen05 MMA; FP64 DMMA / DFMA; the DMMA
// issue-model probes behind profiles/r01_dmma_probe.md).  Nothing under gmat_b200/ loads it; it links the product's
// runtime object (error string, SM count) and nothing else.  Declarations: probe/gmat_b200_probe.h.
#include "../gmat_b200/csrc/tc5.cuh"

namespace gb {

// Issue-rate microbenchmark of the int8 MMA (the roofline denominator of the engine): one thread per CTA issues
// back-to-back 128x256x32 MMAs on a resident (pseudo-random) smem tile pair, alternating two TMEM accumulators.
__global__ void __launch_bounds__(128, 1) i8_peak_kernel(int iters) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    for (int k = threadIdx.x; k < I8_STAGE_BYTES / 4; k += blockDim.x)
        reinterpret_cast<unsigned *>(smem)[k] = (unsigned)(k * 2654435761u) ^ (blockIdx.x * 40503u);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tbase_slot)),
                     "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tbase_slot;
    if (threadIdx.x == 0) {
        const uint32_t sa = smem_u32(smem);
        const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + I8_A_BYTES);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_i8(tbase + (uint32_t)((it & 1) * I8_TK), da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), I8_IDESC,
                        (uint32_t)(it > 1));
        }
        tc_commit(&bar);
        mbar_wait(&bar, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tbase), "r"(512) : "memory");
}


// The same for the FP4 MMA of the GRM (gram_i8.cu: tcgen05.mma kind::mxf4.block_scale.block32, E2M1 operands, unit
// UE8M0 scale factors in TMEM columns [256, 320)): back-to-back 128x256x64 MMAs on a resident tile pair into one
// 256-column accumulator, the way the GRM kernel's K loop issues them.  Every nibble pattern is a finite E2M1 value.
constexpr uint32_t F4_PROBE_IDESC = (1u << 7) | (1u << 10) | ((uint32_t)(I8_TK >> 3) << 17) | (1u << 23) |
                                    ((uint32_t)(I8_TJ >> 4) << 24);
__device__ __forceinline__ void umma_f4(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                        uint32_t tsfa, uint32_t tsfb, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tsfa), "r"(tsfb)
        : "memory");
}
__global__ void __launch_bounds__(128, 1) f4_peak_kernel(int iters) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    for (int k = threadIdx.x; k < I8_STAGE_BYTES / 4; k += blockDim.x)
        reinterpret_cast<unsigned *>(smem)[k] = (unsigned)(k * 2654435761u) ^ (blockIdx.x * 40503u);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tbase_slot)),
                     "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tbase_slot;
    {   // unit scale factors (UE8M0 2^0 = 0x7F) in columns [256, 320) of this warp's 32 lanes
        const uint32_t tq = tbase + ((uint32_t)((threadIdx.x >> 5) * 32) << 16) + (uint32_t)I8_TK;
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
            "{%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, "
            "%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};\n" ::"r"(tq), "r"(0x7F7F7F7Fu) : "memory");
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
            "{%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, "
            "%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};\n" ::"r"(tq + 32u), "r"(0x7F7F7F7Fu) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x == 0) {
        const uint32_t sa = smem_u32(smem);
        const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + I8_A_BYTES);
        const uint32_t tsf = tbase + (uint32_t)I8_TK;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k)                                // K = 64 nibbles = 32 bytes per MMA
                umma_f4(tbase, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), F4_PROBE_IDESC, tsf, tsf + 32u,
                        (uint32_t)(it > 0));
        }
        tc_commit(&bar);
        mbar_wait(&bar, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tbase), "r"(512) : "memory");
}

// ---- FP64 issue-rate microbenchmarks ---------------------------------------------------------
__global__ void __launch_bounds__(512) peak_dmma_kernel(double *out, int iters) {
    double c[8][2];
#pragma unroll
    for (int q = 0; q < 8; ++q) { c[q][0] = 0.0; c[q][1] = 0.0; }
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int q = 0; q < 8; ++q) dmma884(c[q][0], c[q][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += c[q][0] + c[q][1];
    if (s == 12345.678) out[0] = s;                 // keep the chains alive
}

__global__ void __launch_bounds__(512) peak_dfma_kernel(double *out, int iters) {
    double c[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) c[q] = threadIdx.x * 1e-3 + q;
    const double a = 1.0000001, b = 1e-7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int q = 0; q < 8; ++q) c[q] = fma(c[q], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += c[q];
    if (s == 12345.678) out[0] = s;
}

// DMMA issue model probe: one CTA per SM, `warps` warps, ILP independent accumulator tiles per warp arranged
// as an (IA x IB) register tile with distinct A / B fragments (like the real kernels), no memory traffic.
template <int IA, int IB, int MAXT>
__global__ void __launch_bounds__(MAXT) probe_dmma_kernel(double *out, int iters) {
    double c[IA][IB][2], a[IA], b[IB];
#pragma unroll
    for (int x = 0; x < IA; ++x) {
        a[x] = 1.0 + (threadIdx.x + x) * 1e-9;
#pragma unroll
        for (int y = 0; y < IB; ++y) { c[x][y][0] = 0.0; c[x][y][1] = 0.0; }
    }
#pragma unroll
    for (int y = 0; y < IB; ++y) b[y] = 1.0 - (threadIdx.x + y) * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int x = 0; x < IA; ++x)
#pragma unroll
            for (int y = 0; y < IB; ++y) dmma884(c[x][y][0], c[x][y][1], a[x], b[y]);
    }
    double s = 0.0;
#pragma unroll
    for (int x = 0; x < IA; ++x)
#pragma unroll
        for (int y = 0; y < IB; ++y) s += c[x][y][0] + c[x][y][1];
    if (s == 12345.678) out[0] = s;
}

// Same register tile, but the fragments of every k-step are (re)loaded from shared memory the way the
// GEMM-shaped kernels do it: A as LDS.128 (two m-tiles per load), B as LDS.64 (BWIDE = 0) or LDS.128
// (BWIDE = 1, two n-tiles per load).  No barriers, no producers: isolates the LDS + DMMA issue mix.
template <int IA, int IB, int BWIDE, int MAXT>
__global__ void __launch_bounds__(MAXT) probe_dmma_lds_kernel(double *out, int iters) {
    extern __shared__ __align__(16) double sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int x = threadIdx.x; x < 4096; x += blockDim.x) sm[x] = 1.0 + x * 1e-9;     // 32 KB
    __syncthreads();
    double c[IA][IB][2];
#pragma unroll
    for (int x = 0; x < IA; ++x)
#pragma unroll
        for (int y = 0; y < IB; ++y) { c[x][y][0] = 0.0; c[x][y][1] = 0.0; }
    const double2 *A2 = reinterpret_cast<const double2 *>(sm);
    for (int it = 0; it < iters; ++it) {
        const int base = ((it + warp) & 7) * 256;                 // moves every iteration: loads cannot be hoisted
        double a[IA], b[IB];
#pragma unroll
        for (int x = 0; x < IA / 2; ++x) {
            const double2 v = A2[(base + x * 32 + lane) & 2047];
            a[2 * x] = v.x; a[2 * x + 1] = v.y;
        }
        if (BWIDE) {
#pragma unroll
            for (int y = 0; y < IB / 2; ++y) {
                const double2 v = A2[(base + 1024 + y * 32 + lane) & 2047];
                b[2 * y] = v.x; b[2 * y + 1] = v.y;
            }
        } else {
#pragma unroll
            for (int y = 0; y < IB; ++y) b[y] = sm[(2 * base + 2048 + y * 32 + lane) & 4095];
        }
#pragma unroll
        for (int x = 0; x < IA; ++x)
#pragma unroll
            for (int y = 0; y < IB; ++y) dmma884(c[x][y][0], c[x][y][1], a[x], b[y]);
    }
    double s = 0.0;
#pragma unroll
    for (int x = 0; x < IA; ++x)
#pragma unroll
        for (int y = 0; y < IB; ++y) s += c[x][y][0] + c[x][y][1];
    if (s == 12345.678) out[0] = s;
}

// DMMA + DMUL mix probe: an 8 x 4 accumulator tile (32 DMMAs per step) plus NMUL independent DMULs per step,
// MODE 0: the DMULs grouped before the DMMAs and independent of them; MODE 1: spread between the DMMAs, independent;
// MODE 2: grouped, and the B fragments of the step ARE the DMUL results (dependent, like py folding in eff.cu).
template <int NMUL, int MODE>
__global__ void __launch_bounds__(256) probe_dmma_dmul_kernel(double *out, int iters) {
    double c[8][4][2], a[8], b[4], m[4];
#pragma unroll
    for (int x = 0; x < 8; ++x) {
        a[x] = 1.0 + (threadIdx.x + x) * 1e-9;
#pragma unroll
        for (int y = 0; y < 4; ++y) { c[x][y][0] = 0.0; c[x][y][1] = 0.0; }
    }
#pragma unroll
    for (int y = 0; y < 4; ++y) { b[y] = 1.0 - (threadIdx.x + y) * 1e-9; m[y] = 1.0 + y * 1e-12; }
    const double f = 1.0 + 1e-15 * (threadIdx.x & 1);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
            for (int q = 0; q < NMUL; ++q) asm volatile("mul.f64 %0, %0, %1;" : "+d"(m[q & 3]) : "d"(f));
        }
#pragma unroll
        for (int x = 0; x < 8; ++x)
#pragma unroll
            for (int y = 0; y < 4; ++y) {
                if (MODE == 1 && y == 0 && x < NMUL) asm volatile("mul.f64 %0, %0, %1;" : "+d"(m[x & 3]) : "d"(f));
                dmma884(c[x][y][0], c[x][y][1], a[x], MODE == 2 ? m[y] : b[y]);
            }
    }
    double s = m[0] + m[1] + m[2] + m[3];
#pragma unroll
    for (int x = 0; x < 8; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) s += c[x][y][0] + c[x][y][1];
    if (s == 12345.678) out[0] = s;
}

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_i8_peak(int iters, double *tops) {
    GB_CHECK_ARG(iters > 0 && tops, "iters/tops");
    const int smem_bytes = I8_STAGE_BYTES + 1024;
    GB_CUDA(cudaFuncSetAttribute(i8_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    cudaEvent_t e0, e1;
    GB_CUDA(cudaEventCreate(&e0));
    GB_CUDA(cudaEventCreate(&e1));
    const int grid = sm_count();
    i8_peak_kernel<<<grid, 128, smem_bytes>>>(iters / 8 + 1);          // warm-up
    GB_CUDA(cudaEventRecord(e0));
    i8_peak_kernel<<<grid, 128, smem_bytes>>>(iters);
    GB_CUDA(cudaEventRecord(e1));
    GB_CUDA(cudaEventSynchronize(e1));
    GB_LAUNCH_CHECK();
    float ms = 0.f;
    GB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tops = 2.0 * grid * (double)iters * 4.0 * I8_TJ * I8_TK * 32.0 / (ms * 1e-3) / 1e12;
    return GMAT_B200_OK;
}


extern "C" int gmat_b200_f4_peak(int iters, double *tops) {
    GB_CHECK_ARG(iters > 0 && tops, "iters/tops");
    const int smem_bytes = I8_STAGE_BYTES + 1024;
    GB_CUDA(cudaFuncSetAttribute(f4_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    cudaEvent_t e0, e1;
    GB_CUDA(cudaEventCreate(&e0));
    GB_CUDA(cudaEventCreate(&e1));
    const int grid = sm_count();
    f4_peak_kernel<<<grid, 128, smem_bytes>>>(iters / 8 + 1);          // warm-up
    GB_CUDA(cudaEventRecord(e0));
    f4_peak_kernel<<<grid, 128, smem_bytes>>>(iters);
    GB_CUDA(cudaEventRecord(e1));
    GB_CUDA(cudaEventSynchronize(e1));
    GB_LAUNCH_CHECK();
    float ms = 0.f;
    GB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tops = 2.0 * grid * (double)iters * 4.0 * I8_TJ * I8_TK * 64.0 / (ms * 1e-3) / 1e12;
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_dmma_dmul_probe(int warps, int nmul, int mode, int iters, double *tflops) {
    GB_CHECK_ARG(tflops != nullptr && iters > 0 && warps >= 1 && warps <= 8, "args");
    void (*kern)(double *, int) = nullptr;
    if (nmul == 0) kern = probe_dmma_dmul_kernel<0, 0>;
    else if (nmul == 4 && mode == 0) kern = probe_dmma_dmul_kernel<4, 0>;
    else if (nmul == 4 && mode == 1) kern = probe_dmma_dmul_kernel<4, 1>;
    else if (nmul == 4 && mode == 2) kern = probe_dmma_dmul_kernel<4, 2>;
    else if (nmul == 8 && mode == 0) kern = probe_dmma_dmul_kernel<8, 0>;
    else if (nmul == 8 && mode == 1) kern = probe_dmma_dmul_kernel<8, 1>;
    else if (nmul == 1 && mode == 0) kern = probe_dmma_dmul_kernel<1, 0>;
    GB_CHECK_ARG(kern != nullptr, "unsupported (nmul, mode)");
    double *d = nullptr;
    GB_CUDA(cudaMalloc(&d, sizeof(double)));
    cudaEvent_t e0, e1;
    GB_CUDA(cudaEventCreate(&e0));
    GB_CUDA(cudaEventCreate(&e1));
    const int grid = sm_count(), block = warps * 32;
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        GB_CUDA(cudaEventRecord(e0));
        kern<<<grid, block>>>(d, iters);
        GB_LAUNCH_CHECK();
        GB_CUDA(cudaEventRecord(e1));
        GB_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        GB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    *tflops = (double)grid * warps * (double)iters * 32 * 512.0 / (best * 1e-3) / 1e12;      // DMMA flop only
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    return GMAT_B200_OK;
}


extern "C" int gmat_b200_dmma_lds_probe(int warps, int shape, int ctas_per_sm, int iters, double *tflops) {
    GB_CHECK_ARG(tflops != nullptr && iters > 0 && warps >= 1 && warps <= 16 && ctas_per_sm >= 1, "args");
    void (*kern)(double *, int) = nullptr;
    int ia = 4, ib = 4;
    if (shape == 0) kern = probe_dmma_lds_kernel<4, 4, 0, 512>;
    else if (shape == 1) kern = probe_dmma_lds_kernel<4, 4, 1, 512>;
    else if (shape == 2) { kern = probe_dmma_lds_kernel<8, 4, 0, 256>; ia = 8; GB_CHECK_ARG(warps <= 8, "warps"); }
    else if (shape == 3) { kern = probe_dmma_lds_kernel<8, 4, 1, 256>; ia = 8; GB_CHECK_ARG(warps <= 8, "warps"); }
    else if (shape == 4) { kern = probe_dmma_lds_kernel<4, 2, 1, 512>; ib = 2; }
    GB_CHECK_ARG(kern != nullptr, "shape");
    double *d = nullptr;
    GB_CUDA(cudaMalloc(&d, sizeof(double)));
    cudaEvent_t e0, e1;
    GB_CUDA(cudaEventCreate(&e0));
    GB_CUDA(cudaEventCreate(&e1));
    const int grid = sm_count() * ctas_per_sm, block = warps * 32;
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        GB_CUDA(cudaEventRecord(e0));
        kern<<<grid, block, 32768>>>(d, iters);
        GB_LAUNCH_CHECK();
        GB_CUDA(cudaEventRecord(e1));
        GB_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        GB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    const double flop = (double)grid * warps * (double)iters * ia * ib * 512.0;
    *tflops = flop / (best * 1e-3) / 1e12;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_dmma_probe(int warps, int ia, int ib, int ctas_per_sm, int iters, double *tflops) {
    GB_CHECK_ARG(tflops != nullptr && iters > 0 && warps >= 1 && warps <= 32 && ctas_per_sm >= 1, "args");
    void (*kern)(double *, int) = nullptr;
    int maxt = 1024;
    if (ia == 1 && ib == 1) kern = probe_dmma_kernel<1, 1, 1024>;
    else if (ia == 2 && ib == 1) kern = probe_dmma_kernel<2, 1, 1024>;
    else if (ia == 2 && ib == 2) kern = probe_dmma_kernel<2, 2, 1024>;
    else if (ia == 4 && ib == 2) kern = probe_dmma_kernel<4, 2, 1024>;
    else if (ia == 4 && ib == 4) { kern = probe_dmma_kernel<4, 4, 512>; maxt = 512; }
    else if (ia == 8 && ib == 4) { kern = probe_dmma_kernel<8, 4, 256>; maxt = 256; }
    GB_CHECK_ARG(kern != nullptr, "unsupported (ia, ib)");
    GB_CHECK_ARG(warps * 32 <= maxt, "too many warps for this register tile");
    double *d = nullptr;
    GB_CUDA(cudaMalloc(&d, sizeof(double)));
    cudaEvent_t e0, e1;
    GB_CUDA(cudaEventCreate(&e0));
    GB_CUDA(cudaEventCreate(&e1));
    const int grid = sm_count() * ctas_per_sm, block = warps * 32;
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        GB_CUDA(cudaEventRecord(e0));
        kern<<<grid, block>>>(d, iters);
        GB_LAUNCH_CHECK();
        GB_CUDA(cudaEventRecord(e1));
        GB_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        GB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    const double flop = (double)grid * warps * (double)iters * ia * ib * 512.0;
    *tflops = flop / (best * 1e-3) / 1e12;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_fp64_peak(int kind, int iters, double *tflops) {
    GB_CHECK_ARG(tflops != nullptr && iters > 0, "tflops/iters");
    GB_CHECK_ARG(kind == 0 || kind == 1, "kind");
    double *d = nullptr;
    GB_CUDA(cudaMalloc(&d, sizeof(double)));
    cudaEvent_t e0, e1;
    GB_CUDA(cudaEventCreate(&e0));
    GB_CUDA(cudaEventCreate(&e1));
    const int grid = sm_count() * 2, block = 512;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {              // rep 0 is warm-up
        GB_CUDA(cudaEventRecord(e0));
        if (kind == 0) peak_dmma_kernel<<<grid, block>>>(d, iters);
        else peak_dfma_kernel<<<grid, block>>>(d, iters);
        GB_LAUNCH_CHECK();
        GB_CUDA(cudaEventRecord(e1));
        GB_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        GB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    const double per_thread_flop = kind == 0 ? 8.0 * 512.0 / 32.0 : 8.0 * 2.0;
    const double flop = (double)grid * block * (double)iters * per_thread_flop;
    *tflops = flop / (best * 1e-3) / 1e12;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    return GMAT_B200_OK;
}

