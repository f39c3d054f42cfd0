// Approximate epistasis screen: eff_ij = sum_k u_i[k] * v_j[k] * py[k] for i in ilist, j > i,
// as tiled FP64 tensor-core contractions  Eff(I x J) = (U_I . diag(py))' V_J  with the operands
// generated on the fly from the 2-bit genotype store (never materialised as doubles in HBM).
//
// Reference behaviour restated (never copied):
//   scalar form of the contraction + strict threshold ... gmat/remma/_remma_epi_eff_cpu.c:69-77
//   MAF-binned threshold eff_cut[freq[i]*10+freq[j]] ...... gmat/remma/_remma_epi_eff_cpu.c:154
//   AD: two directions, '>=' then '>' ..................... gmat/remma/_remma_epi_eff_cpu.c:233-247
//
// B200 design (the recipe of exact.cu, simpler because nothing is shared between warps in the K loop):
//  * tile = 128 first-SNPs (list entries) x 128 second-SNPs per persistent CTA; 8 warps (2 x 4), warp tile
//    64 x 32 SNPs = 8 x 4 DMMA.8x8x4 accumulator tiles (128 registers), K = individuals in slabs of 16.
//  * operands are built in registers: lane (g, t) keeps the 16-individual code word of each of its 8 + 4 SNP
//    rows (one LDG.32 per row per slab through L1, prefetched a slab ahead) and fetches every fragment element
//    with one LDS.64 from the per-tile value tables lutU[row][code], lutV[col][code] (32 B per SNP: the four rows
//    of a half warp cover one 128 B line, conflict-free); py[k] is folded into the 4 B-side values (4 DMUL per
//    32 DMMA).  No decoded slab in shared memory, no barrier of any kind inside the K loop.
//  * epilogue thresholds the 64 effects of a lane and compacts hits with a warp ballot + one atomic per warp.
#include "common.cuh"

namespace gb {

constexpr int EFF_T = 128;          // tile edge (SNPs)
constexpr int EFF_BK = 16;          // individuals per slab (= 4 bytes of codes per SNP)

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
    const long long *prefix;        // [n_iblk + 1] first work-list position of each I-block
    const long long *jb0;           // [n_iblk] first J-block holding a j > min(i of the block)
    unsigned long long *next_tile;  // dynamic tile counter (zeroed before the launch)
};

struct EffSmem {
    double lutU[EFF_T][4];          // value table of each tile row (first SNP), zero for masked rows
    double lutV[EFF_T][4];          // value table of each tile column (second SNP)
    double cut[112];
    long long isnp[EFF_T];          // SNP index of each tile row (-1: past the end of ilist)
    int binI[EFF_T];
    int binJ[EFF_T];
    long long tile_next;
};

// Work list of the screen: I-blocks of 128 list entries; block ib only needs the J-blocks from jb0[ib] =
// (min first-SNP of the block + 1) / 128 on (j > i).  prefix[] turns a position of the compacted list into
// (ib, jb); the tiles are then handed out dynamically (atomic counter), because a static round-robin over the
// triangular tile set leaves the CTAs up to ~10 % apart.
__global__ void eff_tiles_kernel(const long long *__restrict__ ilist, long long n_i, long long n_iblk,
                                 long long n_jblk, long long *__restrict__ prefix, long long *__restrict__ jb0) {
    for (long long ib = threadIdx.x; ib < n_iblk; ib += blockDim.x) {
        long long lo = (1LL << 62);
        for (int q = 0; q < EFF_T; ++q) {
            const long long li = ib * EFF_T + q;
            if (li < n_i) { const long long s = ilist[li]; if (s >= 0 && s < lo) lo = s; }
        }
        long long j0 = (lo == (1LL << 62)) ? n_jblk : (lo + 1) / EFF_T;
        if (j0 > n_jblk) j0 = n_jblk;
        jb0[ib] = j0;
        prefix[ib + 1] = n_jblk - j0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        prefix[0] = 0;
        for (long long ib = 0; ib < n_iblk; ++ib) prefix[ib + 1] += prefix[ib];
    }
}

__global__ void __launch_bounds__(256, 1) epi_eff_kernel(const EffParams p) {
    constexpr int NT = 256;
    constexpr int MT = 8, NTL = 4;                // warp tile 64 x 32
    extern __shared__ __align__(128) unsigned char smem_raw[];
    EffSmem &S = *reinterpret_cast<EffSmem *>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp >> 2, wn = warp & 3;
    const long long nslab = (p.n + EFF_BK - 1) / EFF_BK;
    const bool use_table = p.cut_table != nullptr;

    if (use_table) {
        for (int x = tid; x < 111; x += NT) S.cut[x] = p.cut_table[x];
    }

    const long long ntiles = p.prefix[p.n_iblk];
    if (tid == 0) S.tile_next = (long long)atomicAdd(p.next_tile, 1ull);
    __syncthreads();
    while (true) {
        const long long tile = S.tile_next;
        __syncthreads();                          // everyone holds `tile`; previous tile fully consumed
        if (tile >= ntiles) break;
        if (tid == 0) S.tile_next = (long long)atomicAdd(p.next_tile, 1ull);   // read after the next barrier pair
        long long ib;
        {
            long long lo = 0, hi = p.n_iblk;      // largest ib with prefix[ib] <= tile (uniform)
            while (hi - lo > 1) {
                const long long mid = (lo + hi) >> 1;
                if (p.prefix[mid] <= tile) lo = mid; else hi = mid;
            }
            ib = lo;
        }
        const long long jb = p.jb0[ib] + (tile - p.prefix[ib]);
        const long long jbase = jb * EFF_T;
        // ---- tile setup: row SNP indices, value tables, bins ---------------------------------
        if (tid < EFF_T) {
            const long long li = ib * EFF_T + tid;
            const long long si = (li < p.n_i) ? p.ilist[li] : -1;
            S.isnp[tid] = si;
            const long long sj = jbase + tid;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                S.lutU[tid][c] = (si >= 0) ? __ldg(p.lutU + si * 4 + c) : 0.0;
                S.lutV[tid][c] = (sj < p.m) ? __ldg(p.lutV + sj * 4 + c) : 0.0;
            }
            if (use_table) {
                S.binI[tid] = (si >= 0) ? p.binI[si] : 0;
                S.binJ[tid] = (sj < p.m) ? p.binJ[sj] : 0;
            }
        }
        __syncthreads();

        // ---- this lane's operand sources -------------------------------------------------------
        // A rows: wm*64 + 8a + g (list entries, arbitrary SNPs);  B columns: wn*32 + 8b + g (consecutive SNPs).
        // Row offsets are 32-bit (the host checks m * row_bytes < 2^32): registers are the scarce resource here.
        unsigned ua[MT], vb[NTL];
        unsigned ta[MT];
#pragma unroll
        for (int a = 0; a < MT; ++a) {
            const int r = wm * 64 + a * 8 + g;
            const long long si = S.isnp[r];
            ua[a] = (unsigned)((si >= 0 ? si : 0) * p.row_bytes);
            ta[a] = smem_u32(&S.lutU[r][0]);
        }
#pragma unroll
        for (int b = 0; b < NTL; ++b) {
            const long long sj = jbase + wn * 32 + 8 * b + g;      // columns past m re-read a valid row (table zero)
            vb[b] = (unsigned)((sj < p.m ? sj : p.m - 1) * p.row_bytes);
        }
        const unsigned tb = smem_u32(&S.lutV[wn * 32 + g][0]);

        unsigned nwa[MT], nwb[NTL];                  // code words of the next slab (prefetched through L1)
        double npy[EFF_BK / 4];
        auto load_slab = [&](long long slab) {
            const long long off = slab * (EFF_BK / 4);
            const bool ok = off + 4 <= p.row_bytes;
#pragma unroll
            for (int a = 0; a < MT; ++a) nwa[a] = ok ? __ldg(reinterpret_cast<const unsigned *>(p.bed + ua[a] + off)) : 0u;
#pragma unroll
            for (int b = 0; b < NTL; ++b) nwb[b] = ok ? __ldg(reinterpret_cast<const unsigned *>(p.bed + vb[b] + off)) : 0u;
#pragma unroll
            for (int k4 = 0; k4 < EFF_BK / 4; ++k4) {
                const long long k = slab * EFF_BK + k4 * 4 + t;
                npy[k4] = (k < p.n) ? __ldg(p.py + k) : 0.0;      // padded individuals contribute exactly 0
            }
        };
        auto lds64 = [](unsigned addr) -> double {
            double v;
            asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(v) : "r"(addr));
            return v;
        };

        double acc[MT][NTL][2];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

        load_slab(0);
        for (long long s = 0; s < nslab; ++s) {
            // byte k4 of xa / xb = 8 * code of individual 16 s + 4 k4 + t: a byte offset into the SNP's value table
            unsigned xa[MT], xb[NTL];
            double pyv[EFF_BK / 4];
#pragma unroll
            for (int a = 0; a < MT; ++a) xa[a] = ((nwa[a] >> (2 * t)) & 0x03030303u) << 3;
#pragma unroll
            for (int b = 0; b < NTL; ++b) xb[b] = ((nwb[b] >> (2 * t)) & 0x03030303u) << 3;
#pragma unroll
            for (int k4 = 0; k4 < EFF_BK / 4; ++k4) pyv[k4] = npy[k4];
            if (s + 1 < nslab) load_slab(s + 1);
            // (hand-pipelining the fragments a step ahead, or pinning the four DMULs together, was measured slower
            // than leaving the schedule to ptxas: 28.4 vs 30.3 TFLOP/s at n=2000)
#pragma unroll
            for (int k4 = 0; k4 < EFF_BK / 4; ++k4) {
                double af[MT], bf[NTL];
#pragma unroll
                for (int a = 0; a < MT; ++a) af[a] = lds64(ta[a] + ((xa[a] >> (8 * k4)) & 0xffu));
#pragma unroll
                for (int b = 0; b < NTL; ++b) bf[b] = lds64(tb + b * (8 * 32) + ((xb[b] >> (8 * k4)) & 0xffu)) * pyv[k4];
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
            }
        }

        // ---- epilogue: threshold + warp-ballot compaction ------------------------------------------
        const bool ge = (p.flags & GMAT_B200_EFF_GE) != 0, swap = (p.flags & GMAT_B200_EFF_SWAP) != 0;
#pragma unroll
        for (int a = 0; a < MT; ++a) {
            const int r = wm * 64 + a * 8 + g;
            const long long si = S.isnp[r];
#pragma unroll
            for (int b = 0; b < NTL; ++b) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const int col = wn * 32 + b * 8 + 2 * t + c;
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
    GB_CHECK_ARG(m * row_bytes < (1LL << 32), "genotype store larger than 4 GiB");
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
    cudaStream_t st = (cudaStream_t)stream;
    long long *scratch = nullptr;                 // prefix[n_iblk + 1], jb0[n_iblk], tile counter
    const size_t sbytes = sizeof(long long) * (2 * p.n_iblk + 2);
    GB_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&scratch), sbytes, st));
    GB_CUDA(cudaMemsetAsync(scratch, 0, sbytes, st));
    p.prefix = scratch;
    p.jb0 = scratch + p.n_iblk + 1;
    p.next_tile = reinterpret_cast<unsigned long long *>(scratch + 2 * p.n_iblk + 1);
    eff_tiles_kernel<<<1, 1024, 0, st>>>(ilist_dev, n_i, p.n_iblk, p.n_jblk, scratch, scratch + p.n_iblk + 1);
    GB_LAUNCH_CHECK();
    auto kern = epi_eff_kernel;
    const size_t smem = sizeof(EffSmem);
    GB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = p.n_iblk * p.n_jblk;         // upper bound of the tile count (exact count is on the device)
    const long long maxgrid = (long long)sm_count();
    if (grid > maxgrid) grid = maxgrid;
    kern<<<(unsigned)grid, 256, smem, st>>>(p);
    GB_LAUNCH_CHECK();
    GB_CUDA(cudaFreeAsync(scratch, st));
    return GMAT_B200_OK;
}
