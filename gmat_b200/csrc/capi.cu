// Host side of the C-ABI: runtime/errors, .bed file I/O, FP64 peak microbenchmarks, and the seven
// DROP-IN symbols of the reference's cffi modules (host pointers + file side effects) implemented on
// top of the device-pointer API.  See include/gmat_b200.h for the contract of every symbol.
#include "common.cuh"
#include <algorithm>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace gb {

static thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count() {
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cached[dev] = v;
    }
    return cached[dev];
}

// Stream-ordered scratch (cudaMallocAsync) comes from the device's default memory pool, whose default is to hand
// freed memory back to the driver at the next synchronisation: the screen's 0.3 - 1 GiB operand scratch would then
// be re-mapped on every call (measured: 4.5 ms of a 28 ms scan).  Keep it cached in the pool instead.
void keep_scratch_pool() {
    static bool done[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done[dev]) return;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    done[dev] = true;
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

extern "C" int gmat_b200_version(void) { return 100; }
extern "C" const char *gmat_b200_last_error(void) { return g_err; }
extern "C" long long gmat_b200_kernel_launches(void) { return g_launches.load(); }

extern "C" int gmat_b200_device_count(void) {
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) {
        set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e));
        return GMAT_B200_ERR_CUDA;
    }
    return c;
}

extern "C" int gmat_b200_set_device(int device) {
    GB_CUDA(cudaSetDevice(device));
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_device_info(int *sm, int *cc_major, int *cc_minor, long long *hbm_bytes) {
    int dev = 0;
    GB_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    GB_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (sm) *sm = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (hbm_bytes) *hbm_bytes = (long long)prop.totalGlobalMem;
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

// ---- .bed file I/O ----------------------------------------------------------------------------
extern "C" int gmat_b200_bed_read_host(const char *bed_file, long long num_id, long long num_snp,
                                       long long row_bytes, unsigned char *host_buf) {
    GB_CHECK_ARG(bed_file != nullptr && num_id > 0 && num_snp > 0, "bed_file/num_id/num_snp");
    const long long src_row = (num_id + 3) / 4;
    GB_CHECK_ARG(host_buf == nullptr || row_bytes >= src_row, "row_bytes < ceil(num_id/4)");
    std::string path = std::string(bed_file) + ".bed";     // the reference appends ".bed" too
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) {
        set_error("Fail to open the plink bed file: %s (%s)", path.c_str(), strerror(errno));
        return GMAT_B200_ERR_IO;
    }
    unsigned char magic[3];
    if (fread(magic, 1, 3, f) != 3 || magic[0] != 0x6c || magic[1] != 0x1b) {
        fclose(f);
        set_error("%s: not a PLINK1 .bed file (bad magic)", path.c_str());
        return GMAT_B200_ERR_IO;
    }
    if (magic[2] != 0x01) {
        fclose(f);
        set_error("%s: individual-major .bed files are not supported (mode byte %d)", path.c_str(), magic[2]);
        return GMAT_B200_ERR_IO;
    }
    if (host_buf) {
        if (row_bytes == src_row) {
            if ((long long)fread(host_buf, 1, (size_t)(src_row * num_snp), f) != src_row * num_snp) {
                fclose(f);
                set_error("%s: file shorter than num_snp*ceil(num_id/4) bytes", path.c_str());
                return GMAT_B200_ERR_IO;
            }
        } else {
            memset(host_buf, 0, (size_t)(row_bytes * num_snp));
            for (long long s = 0; s < num_snp; ++s) {
                if ((long long)fread(host_buf + s * row_bytes, 1, (size_t)src_row, f) != src_row) {
                    fclose(f);
                    set_error("%s: file shorter than num_snp*ceil(num_id/4) bytes", path.c_str());
                    return GMAT_B200_ERR_IO;
                }
            }
        }
    }
    fclose(f);
    return GMAT_B200_OK;
}
