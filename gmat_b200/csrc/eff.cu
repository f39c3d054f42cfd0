// Approximate epistasis screen: eff_ij = sum_k u_i[k] * v_j[k] * py[k] for i in ilist, j > i,
// as tiled FP64 tensor-core contractions  Eff(I x J) = (U_I . diag(py))' V_J  with the operands
// generated on the fly from the 2-bit genotype store (never materialised as doubles in HBM).
//
// Reference behaviour restated (never copied):
//   scalar form of the contraction + strict threshold ... gmat/remma/_remma_epi_eff_cpu.c:69-77
//   MAF-binned threshold eff_cut[freq[i]*10+freq[j]] ...... gmat/remma/_remma_epi_eff_cpu.c:154
//   AD: two directions, '>=' then '>' ..................... gmat/remma/_remma_epi_eff_cpu.c:233-247
//
// B200 design (the recipe of exact.cu):
//  * tile = 128 first-SNPs (list entries) x 128 second-SNPs per persistent CTA; 8 consumer warps (2 x 4), warp
//    tile 64 x 32 SNPs = 8 x 4 DMMA.8x8x4 accumulator tiles, K = individuals in slabs of 16; 4 producer warps
//    whose registers go to the consumers (setmaxnreg).
//  * A operand (first SNPs, py folded in): eff_pack_kernel writes  U_I[row][k] * py[k]  for a chunk of I-blocks
//    ONCE, as doubles in DMMA A-fragment order (128 x 16 slab = one contiguous 16 KB piece).  This is the only
//    place decoded doubles exist outside registers: a bounded scratch (<= 1 GiB per chunk), written once and
//    re-read from L2 by every J-tile of the I-block.  One producer lane streams the slabs through an 8-deep TMA
//    ring (cp.async.bulk + mbarrier expect_tx); fragment loads are conflict-free LDS.128.
//  * B operand (second SNPs) is built in registers: lane (g, t) keeps the 16-individual code word of its 4 SNP
//    rows (LDG.32 through L1, prefetched a slab ahead) and fetches each element with one LDS.64 from the per-tile
//    value table lutV[col][code].  No DMUL and no cross-warp hand-over in the K loop.
//    (v2 built both operands in registers with py multiplied in by DMUL: 82 % of the DMMA peak; the probe in
//    profiles/r01_dmma_probe.md shows what DMULs inside a DMMA stream cost.)
//  * epilogue thresholds the 64 effects of a lane and compacts hits with a warp ballot + one atomic per warp.
#include "common.cuh"
#include <cstdlib>

namespace gb {

constexpr int EFF_T = 128;          // tile edge (SNPs)
constexpr int EFF_BK = 16;          // individuals per slab (= 4 bytes of codes per SNP)
constexpr int EFF_NS = 8;           // ring depth (16 KB slabs of packed A)
constexpr int EFF_CW = 8;           // consumer warps (2 x 4)
constexpr int EFF_PW = 4;           // producer warps (one per SM sub-partition; only one lane works)
constexpr int EFF_SLAB = EFF_T * EFF_BK;      // doubles per slab

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
    const double *packedU;          // [chunk I-blocks][nslab][EFF_SLAB] A operand, fragment order
    long long ib_begin, ib_end;     // I-blocks of this launch (the chunk packedU holds)
    long long nslab;
};

struct EffSmem {
    double A[EFF_NS][EFF_SLAB];     // TMA destination
    double lutV[EFF_T][4];          // value table of each tile column (second SNP), zero past m
    double cut[112];
    long long isnp[EFF_T];          // SNP index of each tile row (-1: past the end of ilist)
    int binI[EFF_T];
    int binJ[EFF_T];
    uint64_t full[EFF_NS], empty[EFF_NS];
};

// Work list of the screen: I-blocks of 128 list entries; block ib only needs the J-blocks from jb0[ib] =
// (min first-SNP of the block + 1) / 128 on (j > i).  prefix[] turns a position of the compacted list into (ib, jb).
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

// packedU[blk][slab][k4][mtp][lane][h] = lutU[i][code(i, k)] * py[k],  i = ilist[(ib_begin + blk)*128 + (2*mtp+h)*8
// + (lane>>2)],  k = slab*16 + k4*4 + (lane&3);  zero for list positions past n_i and individuals past n.
__global__ void __launch_bounds__(256) eff_pack_kernel(const EffParams p, double *__restrict__ out) {
    // one CTA per (slab = blockIdx.x, I-block = blockIdx.y): 2048 doubles, 8 per thread, no 64-bit divisions
    const long long slab = blockIdx.x, blk = blockIdx.y;
    double *dst = out + (blk * p.nslab + slab) * EFF_SLAB;
#pragma unroll
    for (int x = 0; x < EFF_SLAB / 256; ++x) {
        const int q = x * 256 + threadIdx.x;
        const int h = q & 1, lane = (q >> 1) & 31, mtp = (q >> 6) & 7, k4 = q >> 9;
        const long long li = (p.ib_begin + blk) * EFF_T + (2 * mtp + h) * 8 + (lane >> 2);
        const long long k = slab * EFF_BK + k4 * 4 + (lane & 3);
        double v = 0.0;
        if (li < p.n_i && k < p.n) {
            const long long si = p.ilist[li];
            if (si >= 0) {
                const int c = (p.bed[si * p.row_bytes + (k >> 2)] >> ((k & 3) * 2)) & 3;
                v = __ldg(p.lutU + si * 4 + c) * __ldg(p.py + k);
            }
        }
        dst[q] = v;
    }
}

__global__ void __launch_bounds__((EFF_CW + EFF_PW) * 32, 1) epi_eff_kernel(const EffParams p) {
    constexpr int NCT = EFF_CW * 32;
    constexpr int MT = 8, NTL = 4;                // warp tile 64 x 32
    constexpr unsigned SLAB_BYTES = EFF_SLAB * sizeof(double);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    EffSmem &S = *reinterpret_cast<EffSmem *>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool use_table = p.cut_table != nullptr;
    // tiles of this launch = positions [prefix[ib_begin], prefix[ib_end]) of the compacted work list
    const long long t_begin = p.prefix[p.ib_begin], t_end = p.prefix[p.ib_end];

    if (tid == 0) {
        for (int s = 0; s < EFF_NS; ++s) {
            mbar_init(&S.full[s], 1);
            mbar_init(&S.empty[s], EFF_CW);
        }
        mbar_fence_init();
    }
    if (use_table) {
        for (int x = tid; x < 111; x += blockDim.x) S.cut[x] = p.cut_table[x];
    }
    __syncthreads();

    auto tile_iblock = [&](long long tile) {
        long long lo = p.ib_begin, hi = p.ib_end;     // largest ib with prefix[ib] <= tile (uniform)
        while (hi - lo > 1) {
            const long long mid = (lo + hi) >> 1;
            if (p.prefix[mid] <= tile) lo = mid; else hi = mid;
        }
        return lo;
    };

    if (warp >= EFF_CW) {
        // =============================== A-stream producer ==============================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
        if (warp == EFF_CW && lane == 0) {
            int stage = 0;
            unsigned use_odd = 0;
            bool first_pass = true;
            for (long long tile = t_begin + blockIdx.x; tile < t_end; tile += gridDim.x) {
                const long long ib = tile_iblock(tile);
                const double *src = p.packedU + (ib - p.ib_begin) * p.nslab * EFF_SLAB;
                for (long long s = 0; s < p.nslab; ++s) {
                    if (!first_pass) mbar_wait(&S.empty[stage], use_odd ^ 1u);
                    fence_proxy_async();             // consumers' LDS of this stage precede the async refill
                    mbar_expect_tx(&S.full[stage], SLAB_BYTES);
                    tma_bulk_g2s(S.A[stage], src + s * EFF_SLAB, SLAB_BYTES, &S.full[stage]);
                    if (++stage == EFF_NS) { stage = 0; use_odd ^= 1u; first_pass = false; }
                }
            }
        }
        return;
    }

    // =================================== consumer warps =================================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n");
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp >> 2, wn = warp & 3;
    int stage = 0;
    unsigned fphase = 0;
    for (long long tile = t_begin + blockIdx.x; tile < t_end; tile += gridDim.x) {
        const long long ib = tile_iblock(tile);
        const long long jb = p.jb0[ib] + (tile - p.prefix[ib]);
        const long long jbase = jb * EFF_T;
        named_bar_sync(1, NCT);                   // previous tile fully consumed
        // ---- tile setup: row SNP indices, column value tables, bins -----------------------------
        if (tid < EFF_T) {
            const long long li = ib * EFF_T + tid;
            const long long si = (li < p.n_i) ? p.ilist[li] : -1;
            S.isnp[tid] = si;
            const long long sj = jbase + tid;
#pragma unroll
            for (int c = 0; c < 4; ++c) S.lutV[tid][c] = (sj < p.m) ? __ldg(p.lutV + sj * 4 + c) : 0.0;
            if (use_table) {
                S.binI[tid] = (si >= 0) ? p.binI[si] : 0;
                S.binJ[tid] = (sj < p.m) ? p.binJ[sj] : 0;
            }
        }
        named_bar_sync(1, NCT);

        // ---- this lane's B sources: columns wn*32 + 8b + g (consecutive SNPs) ---------------------
        const unsigned char *vb[NTL];
#pragma unroll
        for (int b = 0; b < NTL; ++b) {
            const long long sj = jbase + wn * 32 + 8 * b + g;      // columns past m re-read a valid row (table zero)
            vb[b] = p.bed + (sj < p.m ? sj : p.m - 1) * p.row_bytes;
        }
        const char *tb = reinterpret_cast<const char *>(&S.lutV[wn * 32 + g][0]);
        unsigned nwb[NTL];                        // code words of the next slab (prefetched through L1)
        auto load_codes = [&](long long slab) {
            const long long off = slab * (EFF_BK / 4);
            const bool ok = off + 4 <= p.row_bytes;
#pragma unroll
            for (int b = 0; b < NTL; ++b) nwb[b] = ok ? __ldg(reinterpret_cast<const unsigned *>(vb[b] + off)) : 0u;
        };
        load_codes(0);

        double acc[MT][NTL][2];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

        for (long long s = 0; s < p.nslab; ++s) {
            // byte k4 of xb = 8 * code of individual 16 s + 4 k4 + t: a byte offset into the column's value table
            unsigned xb[NTL];
#pragma unroll
            for (int b = 0; b < NTL; ++b) xb[b] = ((nwb[b] >> (2 * t)) & 0x03030303u) << 3;
            load_codes(s + 1 < p.nslab ? s + 1 : 0);
            mbar_wait(&S.full[stage], fphase);
            const double2 *Af = reinterpret_cast<const double2 *>(S.A[stage]);
#pragma unroll
            for (int k4 = 0; k4 < EFF_BK / 4; ++k4) {
                double af[MT], bf[NTL];
#pragma unroll
                for (int a2 = 0; a2 < MT / 2; ++a2) {
                    const double2 v = Af[(k4 * 8 + wm * (MT / 2) + a2) * 32 + lane];
                    af[2 * a2] = v.x; af[2 * a2 + 1] = v.y;
                }
#pragma unroll
                for (int b = 0; b < NTL; ++b)
                    bf[b] = *reinterpret_cast<const double *>(tb + b * (8 * 32) + ((xb[b] >> (8 * k4)) & 0xffu));
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.empty[stage]);
            if (++stage == EFF_NS) { stage = 0; fphase ^= 1u; }
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
    keep_scratch_pool();
    p.nslab = (n + EFF_BK - 1) / EFF_BK;
    long long *scratch = nullptr;                 // prefix[n_iblk + 1], jb0[n_iblk]
    const size_t sbytes = sizeof(long long) * (2 * p.n_iblk + 2);
    GB_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&scratch), sbytes, st));
    p.prefix = scratch;
    p.jb0 = scratch + p.n_iblk + 1;
    eff_tiles_kernel<<<1, 1024, 0, st>>>(ilist_dev, n_i, p.n_iblk, p.n_jblk, scratch, scratch + p.n_iblk + 1);
    GB_LAUNCH_CHECK();
    // packed A operand for a chunk of I-blocks at a time: bounded scratch (<= 1 GiB), written once per chunk
    const long long blk_bytes = p.nslab * EFF_SLAB * (long long)sizeof(double);
    long long budget = 1LL << 30;
    if (const char *e = getenv("GMAT_B200_EFF_SCRATCH_BYTES")) {        // tests force several chunks with this
        const long long v = atoll(e);
        if (v > 0) budget = v;
    }
    long long chunk = budget / blk_bytes;
    if (chunk < 1) chunk = 1;
    if (chunk > p.n_iblk) chunk = p.n_iblk;
    double *packed = nullptr;
    GB_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&packed), (size_t)(chunk * blk_bytes), st));
    p.packedU = packed;
    auto kern = epi_eff_kernel;
    const size_t smem = sizeof(EffSmem);
    GB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (long long ib0 = 0; ib0 < p.n_iblk; ib0 += chunk) {
        p.ib_begin = ib0;
        p.ib_end = (ib0 + chunk < p.n_iblk) ? ib0 + chunk : p.n_iblk;
        const dim3 pgrid((unsigned)p.nslab, (unsigned)(p.ib_end - p.ib_begin));
        eff_pack_kernel<<<pgrid, 256, 0, st>>>(p, packed);
        GB_LAUNCH_CHECK();
        long long grid = (p.ib_end - p.ib_begin) * p.n_jblk;      // upper bound of the chunk's tile count
        if (grid > sm_count()) grid = sm_count();
        kern<<<(unsigned)grid, (EFF_CW + EFF_PW) * 32, smem, st>>>(p);
        GB_LAUNCH_CHECK();
    }
    GB_CUDA(cudaFreeAsync(packed, st));
    GB_CUDA(cudaFreeAsync(scratch, st));
    return GMAT_B200_OK;
}
