// Host side of the C-ABI: runtime/errors, .bed file I/O, and the seven
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

static thread_local char g_note[256] = "";
void set_engine_note(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_note, sizeof(g_note), fmt, ap);
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

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_version(void) { return 100; }
extern "C" const char *gmat_b200_last_error(void) { return g_err; }
extern "C" const char *gmat_b200_last_engine_note(void) { return g_note; }
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
