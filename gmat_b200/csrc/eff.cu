// Approximate epistasis screen: eff_ij = sum_k u_i[k] * v_j[k] * py[k] for i in ilist, j > i,
// as tiled FP64 tensor-core contractions  Eff(I x J) = (U_I . diag(py))' V_J  with the operands
// generated on the fly from the 2-bit genotype store (never materialised as doubles in HBM).
//
// Reference behaviour restated (never copied):
//   scalar form of the contraction + strict threshold ... gmat/remma/_remma_epi_eff_cpu.c:69-77
//   MAF-binned threshold eff_cut[freq[i]*10+freq[j]] ...... gmat/remma/_remma_epi_eff_cpu.c:154
//   AD: two directions, '>=' then '>' ..................... gmat/remma/_remma_epi_eff_cpu.c:233-247
//
// Tile: 128 first-SNPs x 128 second-SNPs per CTA, K (individuals) in slabs of 16; operands are
// decoded into shared memory as doubles [k][snp] (+4 padding => conflict-free DMMA fragment loads),
// double buffered; DMMA.8x8x4 accumulates in registers; the epilogue thresholds and compacts hits
// with a warp ballot + one atomic per warp.
#include "common.cuh"

namespace gb {

constexpr int EFF_T = 128;          // tile edge (SNPs)
constexpr int EFF_BK = 16;          // individuals per slab (= 4 bytes of codes per SNP)
constexpr int EFF_LD = EFF_T + 4;   // padded leading dimension of the slab arrays

struct EffParams {
    const unsigned char *bed;
    long long n, m, row_bytes;
    const double *lutU, *lutV, *py;
    const long long *ilist;
    long long n_i;
    int flags;
    double eff_cut;
    const double *cut_table;
    const int *binI, *binJ;
    gmat_b200_hit *hits;
    long long cap;
    unsigned long long *count;
    long long n_iblk, n_jblk;
};

template <int WARPS_M, int WARPS_N>
struct EffSmem {
    double U[2][EFF_BK][EFF_LD];
    double V[2][EFF_BK][EFF_LD];
    double lutU[4][EFF_T];
    double lutV[4][EFF_T];
    double cut[112];
    long long isnp[EFF_T];          // SNP index of each tile row (-1: past the end of ilist)
    int binI[EFF_T];
    int binJ[EFF_T];
    long long min_i;
};

template <int WARPS_M, int WARPS_N>
__global__ void __launch_bounds__(WARPS_M *WARPS_N * 32, 1) epi_eff_kernel(const EffParams p) {
    constexpr int NT = WARPS_M * WARPS_N * 32;
    constexpr int WM = EFF_T / WARPS_M, WN = EFF_T / WARPS_N;
    constexpr int MT = WM / 8, NTL = WN / 8;
    constexpr int PARTS = NT / EFF_T;             // threads per SNP in the producer
    constexpr int KPER = EFF_BK / PARTS;          // individuals decoded per thread per operand
    static_assert(KPER == 4 || KPER == 8, "producer split");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EffSmem<WARPS_M, WARPS_N> &S = *reinterpret_cast<EffSmem<WARPS_M, WARPS_N> *>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / WARPS_N, wn = warp % WARPS_N;
    const int psnp = tid % EFF_T, ppart = tid / EFF_T;
    const bool upper = true;                      // the C screen only ever tests j > i
    const long long nslab = (p.n + EFF_BK - 1) / EFF_BK;
    const bool use_table = p.cut_table != nullptr;

    if (use_table) {
        for (int x = tid; x < 111; x += NT) S.cut[x] = p.cut_table[x];
    }

    const long long ntiles = p.n_iblk * p.n_jblk;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long ib = tile / p.n_jblk, jb = tile % p.n_jblk;
        const long long jbase = jb * EFF_T;
        __syncthreads();                          // previous tile fully consumed
        // ---- tile setup: row SNP indices, LUTs, bins -----------------------------------------
        if (tid < EFF_T) {
            const long long li = ib * EFF_T + tid;
            const long long si = (li < p.n_i) ? p.ilist[li] : -1;
            S.isnp[tid] = si;
            const long long sj = jbase + tid;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                S.lutU[c][tid] = (si >= 0) ? __ldg(p.lutU + si * 4 + c) : 0.0;
                S.lutV[c][tid] = (sj < p.m) ? __ldg(p.lutV + sj * 4 + c) : 0.0;
            }
            if (use_table) {
                S.binI[tid] = (si >= 0) ? p.binI[si] : 0;
                S.binJ[tid] = (sj < p.m) ? p.binJ[sj] : 0;
            }
        }
        __syncthreads();
        if (warp == 0) {       // min first-SNP of the block: lets tiles wholly at/below the diagonal be skipped
            long long v = (1LL << 62);
#pragma unroll
            for (int x = 0; x < EFF_T / 32; ++x) {
                const long long si = S.isnp[lane + 32 * x];
                if (si >= 0 && si < v) v = si;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const long long w = __shfl_xor_sync(0xffffffffu, v, o);
                v = w < v ? w : v;
            }
            if (lane == 0) S.min_i = v;
        }
        __syncthreads();
        if (upper && jbase + EFF_T <= S.min_i + 1) continue;      // no pair with j > i in this tile
        if (S.min_i == (1LL << 62)) continue;

        // ---- producer state ---------------------------------------------------------------------
        const long long usnp = S.isnp[psnp];
        const long long vsnp = jbase + psnp;
        const unsigned char *urow = p.bed + (usnp >= 0 ? usnp : 0) * p.row_bytes + ppart * (KPER / 4);
        const unsigned char *vrow = p.bed + (vsnp < p.m ? vsnp : 0) * p.row_bytes + ppart * (KPER / 4);

        auto load_codes = [&](long long slab, unsigned &cu, unsigned &cv) {
            if (KPER == 8) {
                cu = __ldg(reinterpret_cast<const unsigned short *>(urow + slab * 4));
                cv = __ldg(reinterpret_cast<const unsigned short *>(vrow + slab * 4));
            } else {
                cu = __ldg(urow + slab * 4);
                cv = __ldg(vrow + slab * 4);
            }
        };
        auto store_slab = [&](int buf, long long slab, unsigned cu, unsigned cv) {
            const long long k0 = slab * EFF_BK + ppart * KPER;
#pragma unroll
            for (int q = 0; q < KPER; ++q) {
                const long long k = k0 + q;
                const double pyk = (k < p.n) ? __ldg(p.py + k) : 0.0;
                const int c_u = (cu >> (2 * q)) & 3, c_v = (cv >> (2 * q)) & 3;
                // padded individuals (k >= n) contribute exactly 0 through pyk == 0 and a zeroed V
                S.U[buf][ppart * KPER + q][psnp] = S.lutU[c_u][psnp] * pyk;
                S.V[buf][ppart * KPER + q][psnp] = (k < p.n) ? S.lutV[c_v][psnp] : 0.0;
            }
        };

        double acc[MT][NTL][2];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

        unsigned cu, cv;
        load_codes(0, cu, cv);
        store_slab(0, 0, cu, cv);
        __syncthreads();
        for (long long s = 0; s < nslab; ++s) {
            const int buf = (int)(s & 1);
            const bool more = (s + 1 < nslab);
            if (more) load_codes(s + 1, cu, cv);
#pragma unroll
            for (int ks = 0; ks < EFF_BK / 4; ++ks) {
                double af[MT], bf[NTL];
#pragma unroll
                for (int a = 0; a < MT; ++a) af[a] = S.U[buf][ks * 4 + t][wm * WM + a * 8 + g];
#pragma unroll
                for (int b = 0; b < NTL; ++b) bf[b] = S.V[buf][ks * 4 + t][wn * WN + b * 8 + g];
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
            }
            if (more) store_slab(buf ^ 1, s + 1, cu, cv);
            __syncthreads();
        }

        // ---- epilogue: threshold + warp-ballot compaction ------------------------------------------
        const bool ge = (p.flags & GMAT_B200_EFF_GE) != 0, swap = (p.flags & GMAT_B200_EFF_SWAP) != 0;
#pragma unroll
        for (int a = 0; a < MT; ++a) {
            const int r = wm * WM + a * 8 + g;
            const long long si = S.isnp[r];
#pragma unroll
            for (int b = 0; b < NTL; ++b) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const int col = wn * WN + b * 8 + 2 * t + c;
                    const long long sj = jbase + col;
                    const double e = acc[a][b][c];
                    bool hit = (si >= 0) && (sj < p.m) && (sj > si);
                    if (hit) {
                        const double cut = use_table ? S.cut[S.binI[r] * 10 + S.binJ[col]] : p.eff_cut;
                        hit = ge ? (fabs(e) >= cut) : (fabs(e) > cut);
                    }
                    const unsigned ballot = __ballot_sync(0xffffffffu, hit);
                    if (ballot) {
                        unsigned long long base = 0;
                        if (lane == 0) base = atomicAdd(p.count, (unsigned long long)__popc(ballot));
                        base = __shfl_sync(0xffffffffu, base, 0);
                        if (hit) {
                            const unsigned long long pos = base + __popc(ballot & ((1u << lane) - 1u));
                            if ((long long)pos < p.cap) {
                                gmat_b200_hit h;
                                h.snp_0 = (int)(swap ? sj : si);
                                h.snp_1 = (int)(swap ? si : sj);
                                h.eff = e; h.var = 0.0; h.chi = 0.0;
                                p.hits[pos] = h;
                            }
                        }
                    }
                }
            }
        }
    }
}

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_epi_eff(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                 const double *lutU_dev, const double *lutV_dev, const double *py_dev,
                                 const long long *ilist_dev, long long n_i, int flags, double eff_cut,
                                 const double *cut_table_dev, const int *binI_dev, const int *binJ_dev,
                                 gmat_b200_hit *hits_dev, long long cap, unsigned long long *count_dev,
                                 void *stream) {
    GB_CHECK_ARG(bed_dev && lutU_dev && lutV_dev && py_dev && ilist_dev && count_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && m > 0 && n_i >= 0, "sizes");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4 && row_bytes % 16 == 0, "row_bytes");
    GB_CHECK_ARG(row_bytes * 4 >= round_up(n, EFF_BK), "row_bytes must cover n rounded up to 16 individuals");
    GB_CHECK_ARG(cap == 0 || hits_dev != nullptr, "hits_dev is NULL");
    GB_CHECK_ARG(cut_table_dev == nullptr || (binI_dev && binJ_dev), "bins required with cut table");
    if (n_i == 0) return GMAT_B200_OK;
    EffParams p;
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes;
    p.lutU = lutU_dev; p.lutV = lutV_dev; p.py = py_dev;
    p.ilist = ilist_dev; p.n_i = n_i; p.flags = flags; p.eff_cut = eff_cut;
    p.cut_table = cut_table_dev; p.binI = binI_dev; p.binJ = binJ_dev;
    p.hits = hits_dev; p.cap = cap; p.count = count_dev;
    p.n_iblk = (n_i + EFF_T - 1) / EFF_T;
    p.n_jblk = (m + EFF_T - 1) / EFF_T;
    constexpr int WMm = 4, WNn = 4;
    auto kern = epi_eff_kernel<WMm, WNn>;
    const size_t smem = sizeof(EffSmem<WMm, WNn>);
    GB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = p.n_iblk * p.n_jblk;
    const long long maxgrid = (long long)sm_count();
    if (grid > maxgrid) grid = maxgrid;
    kern<<<(unsigned)grid, WMm * WNn * 32, smem, (cudaStream_t)stream>>>(p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
