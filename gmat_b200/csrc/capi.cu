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

}  // namespace gb

using namespace gb;

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
