// PLINK .bed 2-bit genotype store: per-SNP counts, per-SNP coding tables (LUTs), decode, pack.
//
// Reference behaviour restated here (never copied):
//   decode map and orientation ......... gmat/process_plink/_read_plink_bed.c:19-46
//   additive coding x = g - 2p ........... gmat/remma/remma_epiAA/remma_epiAA.py:59-61
//   dominance coding d = h - 2p(1-p) ..... gmat/remma/remma_epiDD/remma_epiDD.py:59-65
//   C-side p accumulated term by term .... gmat/remma/_remma_epi_eff_cpu.c:103-112, 272-287, 454-466
//
// All of these are HBM-bound byte kernels (0.25 B per genotype in); grids are multiples of the SM
// count, loads are 16-byte vectors where the layout allows.
#include "common.cuh"

namespace gb {

// ------------------------------------------------------------------------------------------
// counts[snp][c]: one warp per SNP, 32-bit words, popcount on the even/odd bit planes.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bed_stats_kernel(const unsigned char *__restrict__ bed,
                                                        long long n, long long m, long long row_bytes,
                                                        int *__restrict__ counts) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const long long nwarp = (long long)gridDim.x * (blockDim.x >> 5);
    const int nwords = (int)((n + 15) / 16);          // 16 individuals per 32-bit word
    for (long long snp = warp0; snp < m; snp += nwarp) {
        const uint32_t *row = reinterpret_cast<const uint32_t *>(bed + snp * row_bytes);
        int c1 = 0, c2 = 0, c3 = 0;
        for (int w = lane; w < nwords; w += 32) {
            uint32_t x = __ldg(row + w);
            const long long first = (long long)w * 16;
            if (first + 16 > n) {                     // mask individuals >= n in the last word
                const int valid = (int)(n - first);   // 1..15
                x &= (valid >= 16) ? 0xffffffffu : ((1u << (2 * valid)) - 1u);
            }
            const uint32_t lo = x & 0x55555555u, hi = (x >> 1) & 0x55555555u;
            c3 += __popc(lo & hi);
            c2 += __popc(hi & ~lo);
            c1 += __popc(lo & ~hi);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            c1 += __shfl_xor_sync(0xffffffffu, c1, o);
            c2 += __shfl_xor_sync(0xffffffffu, c2, o);
            c3 += __shfl_xor_sync(0xffffffffu, c3, o);
        }
        if (lane == 0) {
            int4 v;
            v.y = c1; v.z = c2; v.w = c3;
            v.x = (int)n - c1 - c2 - c3;
            reinterpret_cast<int4 *>(counts)[snp] = v;
        }
    }
}

// ------------------------------------------------------------------------------------------
// LUT build.  One thread per SNP.  FREQ_NUMPY uses the exact integer allele sum; FREQ_CLOOP
// replays the reference C accumulation order so p is bit-identical to the C scan's pFreq.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) build_lut_kernel(const unsigned char *__restrict__ bed,
                                                        long long n, long long m, long long row_bytes,
                                                        const int *__restrict__ counts, int coding,
                                                        int freq_mode, double *__restrict__ lut,
                                                        double *__restrict__ freq_out) {
    const long long snp = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (snp >= m) return;
    const double third = 2.0 / 6.0;                   // (1*1+1)/6.0, _read_plink_bed.c:37
    double p;
    if (freq_mode == GMAT_B200_FREQ_NUMPY) {
        const int4 c = reinterpret_cast<const int4 *>(counts)[snp];
        // np.sum(snp_mat, axis=0) / (2 * num_id): the sum of exact small integers is exact
        const double s = (double)((long long)c.z + 2LL * (long long)c.w);
        p = s / (double)(2 * n);
    } else {
        // pFreq += marker_mat[i*num_id+j]/(2*num_id), j ascending (_remma_epi_eff_cpu.c:105-107)
        const double den = (double)(2 * n);
        const double q[4] = {0.0 / den, third / den, 1.0 / den, 2.0 / den};
        const unsigned char *row = bed + snp * row_bytes;
        p = 0.0;
        const long long nfull = n / 4;
        for (long long b = 0; b < nfull; ++b) {
            const unsigned x = row[b];
            p = __dadd_rn(p, q[x & 3]);
            p = __dadd_rn(p, q[(x >> 2) & 3]);
            p = __dadd_rn(p, q[(x >> 4) & 3]);
            p = __dadd_rn(p, q[(x >> 6) & 3]);
        }
        const int rem = (int)(n - nfull * 4);
        if (rem) {
            const unsigned x = row[nfull];
            for (int t = 0; t < rem; ++t) p = __dadd_rn(p, q[(x >> (2 * t)) & 3]);
        }
    }
    if (freq_out) freq_out[snp] = p;
    double v0, v1, v2, v3;
    const double miss = (freq_mode == GMAT_B200_FREQ_NUMPY) ? __longlong_as_double(0x7ff8000000000000LL)
                                                            : third;
    if (coding == GMAT_B200_CODING_RAW) {
        v0 = 0.0; v1 = miss; v2 = 1.0; v3 = 2.0;
    } else if (coding == GMAT_B200_CODING_ADD) {
        const double c = __dmul_rn(2.0, p);           // 2*pFreq
        v0 = __dsub_rn(0.0, c); v1 = __dsub_rn(miss, c); v2 = __dsub_rn(1.0, c); v3 = __dsub_rn(2.0, c);
    } else {
        // 2*p*(1-p) evaluated left to right like numpy and C: (2*p)*(1-p)
        const double c = __dmul_rn(__dmul_rn(2.0, p), __dsub_rn(1.0, p));
        v0 = __dsub_rn(0.0, c); v1 = __dsub_rn(miss, c); v2 = __dsub_rn(1.0, c); v3 = __dsub_rn(0.0, c);
    }
    double4 o;                                        // 32-byte aligned: lut is double[m][4]
    o.x = v0; o.y = v1; o.z = v2; o.w = v3;
    reinterpret_cast<double4 *>(lut)[snp] = o;
}

// ------------------------------------------------------------------------------------------
// decode: out[snp][k] = lut[snp][code] (or the raw (c*c+c)/6 map).  HBM-bound, writes dominate (8 B per
// genotype).  One warp per 512-genotype chunk of a SNP row: lane l loads ONE 32-bit word of codes (16
// genotypes; a coalesced 128 B per warp), then for j = 0..7 the warp writes 64 consecutive genotypes as one
// fully coalesced 512 B streaming store (lane l: genotypes 64j+2l, 64j+2l+1, codes fetched from the owning
// lane with a shuffle).  No integer division or FP64 arithmetic per genotype: values come from 4 registers.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double sel4(int c, double t0, double t1, double t2, double t3) {
    return (c & 2) ? ((c & 1) ? t3 : t2) : ((c & 1) ? t1 : t0);
}

template <bool HAS_LUT>
__global__ void __launch_bounds__(256) decode_kernel(const unsigned char *__restrict__ bed,
                                                     long long n, long long m, long long row_bytes,
                                                     const double *__restrict__ lut,
                                                     double *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const long long nwarp = (long long)gridDim.x * (blockDim.x >> 5);
    const long long cpr = (n + 511) / 512;                       // chunks per SNP row
    const long long total = cpr * m;
    for (long long chunk = warp0; chunk < total; chunk += nwarp) {
        const long long snp = chunk / cpr;
        const long long ch = chunk - snp * cpr;
        const long long woff = (ch * 32 + lane) * 4;                 // byte offset of this lane's code word
        unsigned word = 0;
        if (woff < row_bytes) word = __ldg(reinterpret_cast<const unsigned *>(bed + snp * row_bytes + woff));
        double t0, t1, t2, t3;
        if (HAS_LUT) {
            const double2 a = __ldg(reinterpret_cast<const double2 *>(lut) + 2 * snp);
            const double2 b = __ldg(reinterpret_cast<const double2 *>(lut) + 2 * snp + 1);
            t0 = a.x; t1 = a.y; t2 = b.x; t3 = b.y;
        } else {
            t0 = 0.0; t1 = 2.0 / 6.0; t2 = 1.0; t3 = 2.0;            // (c*c+c)/6.0, _read_plink_bed.c:37
        }
        double *row = out + snp * n;
        const bool vec_ok = ((snp * n) & 1) == 0;                    // 16-byte aligned double2 stores
        const int sh = (2 * lane & 15) * 2;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const unsigned w = __shfl_sync(0xffffffffu, word, 4 * j + (lane >> 3));
            const long long k = ch * 512 + j * 64 + 2 * lane;
            const double v0 = sel4((w >> sh) & 3, t0, t1, t2, t3);
            const double v1 = sel4((w >> (sh + 2)) & 3, t0, t1, t2, t3);
            if (k + 1 < n) {
                if (vec_ok) {
                    __stcs(reinterpret_cast<double2 *>(row + k), make_double2(v0, v1));
                } else {
                    __stcs(row + k, v0);
                    __stcs(row + k + 1, v1);
                }
            } else if (k < n) {
                __stcs(row + k, v0);
            }
        }
    }
}

// pack: geno[snp][k] in {0,1,2} -> 2-bit codes (anything else -> 01 missing)
__global__ void __launch_bounds__(256) pack_kernel(const double *__restrict__ geno, long long n, long long m,
                                                   long long row_bytes, unsigned char *__restrict__ bed) {
    const long long total = row_bytes * m;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long snp = e / row_bytes;
        const long long b = e - snp * row_bytes;
        unsigned x = 0;
        for (int q = 0; q < 4; ++q) {
            const long long k = b * 4 + q;
            if (k < n) {
                const double g = geno[snp * n + k];
                const unsigned c = (g == 0.0) ? 0u : (g == 1.0) ? 2u : (g == 2.0) ? 3u : 1u;
                x |= c << (2 * q);
            }
        }
        bed[e] = (unsigned char)x;
    }
}

}  // namespace gb

using namespace gb;

static int check_store(const void *bed_dev, long long n, long long m, long long row_bytes) {
    GB_CHECK_ARG(bed_dev != nullptr, "bed_dev is NULL");
    GB_CHECK_ARG(n > 0 && m > 0, "n and m must be positive");
    GB_CHECK_ARG(n < (1LL << 30) && m < (1LL << 31), "n or m too large");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4, "row_bytes < ceil(n/4)");
    GB_CHECK_ARG(row_bytes % 16 == 0, "row_bytes must be a multiple of 16");
    GB_CHECK_ARG((reinterpret_cast<uintptr_t>(bed_dev) & 15) == 0, "bed_dev must be 16-byte aligned");
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_bed_stats(const unsigned char *bed_dev, long long n, long long m,
                                   long long row_bytes, int *counts_dev, void *stream) {
    int rc = check_store(bed_dev, n, m, row_bytes);
    if (rc != GMAT_B200_OK) return rc;
    GB_CHECK_ARG(counts_dev != nullptr, "counts_dev is NULL");
    const int wpb = 8;
    long long blocks = (m + wpb - 1) / wpb;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    bed_stats_kernel<<<(unsigned)blocks, wpb * 32, 0, (cudaStream_t)stream>>>(bed_dev, n, m, row_bytes,
                                                                              counts_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_build_lut(const unsigned char *bed_dev, long long n, long long m,
                                   long long row_bytes, const int *counts_dev, int coding, int freq_mode,
                                   double *lut_dev, double *freq_dev, void *stream) {
    int rc = check_store(bed_dev, n, m, row_bytes);
    if (rc != GMAT_B200_OK) return rc;
    GB_CHECK_ARG(lut_dev != nullptr, "lut_dev is NULL");
    GB_CHECK_ARG(coding >= 0 && coding <= 2, "coding");
    GB_CHECK_ARG(freq_mode == GMAT_B200_FREQ_NUMPY || freq_mode == GMAT_B200_FREQ_CLOOP, "freq_mode");
    GB_CHECK_ARG(freq_mode == GMAT_B200_FREQ_CLOOP || counts_dev != nullptr, "counts_dev is NULL");
    build_lut_kernel<<<(unsigned)((m + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        bed_dev, n, m, row_bytes, counts_dev, coding, freq_mode, lut_dev, freq_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_decode(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                const double *lut_dev, double *out_dev, void *stream) {
    int rc = check_store(bed_dev, n, m, row_bytes);
    if (rc != GMAT_B200_OK) return rc;
    GB_CHECK_ARG(out_dev != nullptr, "out_dev is NULL");
    const long long total = (n + 511) / 512 * m;              // one warp per 512-genotype chunk
    long long blocks = (total + 7) / 8;
    const long long cap = (long long)sm_count() * 32;
    if (blocks > cap) blocks = cap;
    if (lut_dev)
        decode_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(bed_dev, n, m, row_bytes,
                                                                               lut_dev, out_dev);
    else
        decode_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(bed_dev, n, m, row_bytes,
                                                                                nullptr, out_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_pack(const double *geno_dev, long long n, long long m, long long row_bytes,
                              unsigned char *bed_dev_out, void *stream) {
    int rc = check_store(bed_dev_out, n, m, row_bytes);
    if (rc != GMAT_B200_OK) return rc;
    GB_CHECK_ARG(geno_dev != nullptr, "geno_dev is NULL");
    const long long total = row_bytes * m;
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)sm_count() * 32;
    if (blocks > cap) blocks = cap;
    pack_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(geno_dev, n, m, row_bytes, bed_dev_out);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
