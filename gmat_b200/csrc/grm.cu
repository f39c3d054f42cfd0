// Genomic relationship matrix  K = X X' / scale  with X[ind][snp] = lut[snp][code] generated on the fly
// from the 2-bit genotype store; FP64 tensor cores (DMMA.8x8x4); SYRK-style: only lower-triangle
// 128x128 tiles are computed, then mirrored; fused epilogue: / scale, diag += diag*small_val.
//
// Reference behaviour restated (never copied):
//   agmat:    centring, kin = dot/scale, diag inflation ... gmat/gmatrix/gmatrix.py:53-66
//   dgmat_as: dominance coding and scale .................. gmat/gmatrix/gmatrix.py:116-130
//
// K dimension = SNPs, in slabs of 16: for a tile (mi, ni) a slab needs 16 x 32 B of codes per operand
// (SNP-major rows, 128 individuals = 32 bytes), decoded into shared memory [snp][ind] (+4 padding =>
// conflict-free fragment loads for both operands), double buffered under the MMAs.
#include "common.cuh"

namespace gb {

constexpr int GR_T = 128;
constexpr int GR_BK = 16;
constexpr int GR_LD = GR_T + 4;

struct GrmParams {
    const unsigned char *bed;
    long long n, m, row_bytes;
    const double *lut;
    long long snp_begin, snp_end;
    double scale, small_val;
    int accumulate;
    double *K;
    long long nblk;
};

struct GrmSmem {
    double X[2][2][GR_BK][GR_LD];      // [buffer][operand: rows(mi) / cols(ni)][snp][individual]
    double lut[3][GR_BK][4];           // LUT slabs, written two slabs ahead
};

template <int WARPS_M, int WARPS_N>
__global__ void __launch_bounds__(WARPS_M *WARPS_N * 32, 1) grm_kernel(const GrmParams p) {
    constexpr int NT = WARPS_M * WARPS_N * 32;
    static_assert(NT == 512, "producer mapping assumes 16 warps");
    constexpr int WM = GR_T / WARPS_M, WN = GR_T / WARPS_N;
    constexpr int MT = WM / 8, NTL = WN / 8;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GrmSmem &S = *reinterpret_cast<GrmSmem *>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / WARPS_N, wn = warp % WARPS_N;
    // producer role: operand = warp/8, two SNPs of the slab per warp, lane -> individuals lane + 32q
    const int pop = warp >> 3;
    const int ps0 = (warp & 7) * 2;
    const long long nslab = (p.snp_end - p.snp_begin + GR_BK - 1) / GR_BK;
    const long long ntiles = p.nblk * (p.nblk + 1) / 2;

    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        // tile -> (mi, ni), ni <= mi
        long long mi = (long long)((sqrt(8.0 * (double)tile + 1.0) - 1.0) * 0.5);
        while (mi * (mi + 1) / 2 > tile) --mi;
        while ((mi + 1) * (mi + 2) / 2 <= tile) ++mi;
        const long long ni = tile - mi * (mi + 1) / 2;
        const long long blk = pop == 0 ? mi : ni;
        __syncthreads();

        uint32_t words[2][4];
        auto load_codes = [&](long long slab) {
#pragma unroll
            for (int s2 = 0; s2 < 2; ++s2) {
                const long long snp = p.snp_begin + slab * GR_BK + ps0 + s2;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const long long off = blk * (GR_T / 4) + (long long)((lane >> 4) + 2 * q) * 4;
                    uint32_t w = 0;
                    if (snp < p.snp_end && off < p.row_bytes)
                        w = __ldg(reinterpret_cast<const uint32_t *>(p.bed + snp * p.row_bytes + off));
                    words[s2][q] = w;
                }
            }
        };
        auto load_lut = [&](long long slab) -> double {     // threads 0..63: one LUT entry each
            const long long snp = p.snp_begin + slab * GR_BK + (tid >> 2);
            return (tid < 64 && snp < p.snp_end) ? __ldg(p.lut + snp * 4 + (tid & 3)) : 0.0;
        };
        auto store_lut = [&](long long slab, double v) {
            if (tid < 64) S.lut[slab % 3][tid >> 2][tid & 3] = v;
        };
        auto store_slab = [&](int buf, long long slab) {
            const int sh = 2 * (lane & 15);
#pragma unroll
            for (int s2 = 0; s2 < 2; ++s2) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = (words[s2][q] >> sh) & 3;
                    S.X[buf][pop][ps0 + s2][lane + 32 * q] = S.lut[slab % 3][ps0 + s2][c];
                }
            }
        };

        double acc[MT][NTL][2];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

        // prologue: LUT slabs 0 and 1, operand slab 0
        store_lut(0, load_lut(0));
        if (nslab > 1) store_lut(1, load_lut(1));
        load_codes(0);
        __syncthreads();
        store_slab(0, 0);
        __syncthreads();
        for (long long s = 0; s < nslab; ++s) {
            const int buf = (int)(s & 1);
            const bool more = s + 1 < nslab;
            double lutv = 0.0;
            if (more) load_codes(s + 1);
            if (s + 2 < nslab) lutv = load_lut(s + 2);
#pragma unroll
            for (int ks = 0; ks < GR_BK / 4; ++ks) {
                double af[MT], bf[NTL];
#pragma unroll
                for (int a = 0; a < MT; ++a) af[a] = S.X[buf][0][ks * 4 + t][wm * WM + a * 8 + g];
#pragma unroll
                for (int b = 0; b < NTL; ++b) bf[b] = S.X[buf][1][ks * 4 + t][wn * WN + b * 8 + g];
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
            }
            if (more) store_slab(buf ^ 1, s + 1);       // reads lut slab s+1 (stored one iteration ago)
            if (s + 2 < nslab) store_lut(s + 2, lutv);
            __syncthreads();
        }

        // ---- epilogue ---------------------------------------------------------------------------
#pragma unroll
        for (int a = 0; a < MT; ++a) {
            const long long r = mi * GR_T + wm * WM + a * 8 + g;
#pragma unroll
            for (int b = 0; b < NTL; ++b) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const long long col = ni * GR_T + wn * WN + b * 8 + 2 * t + c;
                    if (r < p.n && col < p.n && (mi != ni || col <= r)) {
                        if (p.accumulate) {
                            p.K[r * p.n + col] += acc[a][b][c];
                            if (col != r) p.K[col * p.n + r] += acc[a][b][c];
                        } else {
                            double v = acc[a][b][c] / p.scale;
                            if (col == r) v = v + v * p.small_val;       // gmatrix.py:64-66
                            p.K[r * p.n + col] = v;
                            if (col != r) p.K[col * p.n + r] = v;
                        }
                    }
                }
            }
        }
    }
}

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_grm(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                             const double *lut_dev, long long snp_begin, long long snp_end, double scale,
                             double small_val, int accumulate, double *K_dev, void *stream) {
    GB_CHECK_ARG(bed_dev && lut_dev && K_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && m > 0 && n < (1LL << 30), "sizes");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4 && row_bytes % 16 == 0, "row_bytes");
    GB_CHECK_ARG((reinterpret_cast<uintptr_t>(bed_dev) & 15) == 0, "bed_dev alignment");
    GB_CHECK_ARG(0 <= snp_begin && snp_begin <= snp_end && snp_end <= m, "snp range");
    GB_CHECK_ARG(accumulate || scale != 0.0, "scale");
    GrmParams p;
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes; p.lut = lut_dev;
    p.snp_begin = snp_begin; p.snp_end = snp_end; p.scale = scale; p.small_val = small_val;
    p.accumulate = accumulate; p.K = K_dev;
    p.nblk = (n + GR_T - 1) / GR_T;
    auto kern = grm_kernel<4, 4>;
    const size_t smem = sizeof(GrmSmem);
    GB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = p.nblk * (p.nblk + 1) / 2;
    if (grid > sm_count()) grid = sm_count();
    kern<<<(unsigned)grid, 512, smem, (cudaStream_t)stream>>>(p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
