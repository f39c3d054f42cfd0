// Exact epistasis test: for a tile of 128 SNP pairs, w = u_i o v_j is generated on the fly from the
// 2-bit genotype store and the quadratic forms var = w' A w (A = Z'PZ) are evaluated as an FP64
// tensor-core GEMM  Y = A W  fused with the Hadamard/column-sum epilogue  var = sum_k W o Y,
// eff = W' py,  chi = eff^2 / var, thresholding and warp-ballot hit compaction.
//
// Reference behaviour restated (never copied):
//   statistics per pair ........................ gmat/remma/remma_epiAA/remma_epiAA.py:72-77
//   AD pair space (all j, incl. j == i) ........ gmat/remma/remma_epiAD/remma_epiAD.py:76-77
//   pair-list variant (dense eff/var/chi) ...... gmat/remma/remma_epiAA/remma_epiAA_pair.py:78-84
//
// B200 design:
//  * A is re-tiled once per run (pack_pvp) into the exact order the kernel consumes it: lower
//    block-triangle of 128-row blocks, each 128x16 slab stored in DMMA.8x8x4 A-fragment order, so a
//    slab is ONE contiguous 16 KB TMA bulk copy (cp.async.bulk + mbarrier) and every fragment load is
//    a conflict-free LDS.128.  Symmetry: w'Aw = sum_mb w_mb'( A[mb,mb] w_mb + 2 sum_{kb<mb} A[mb,kb] w_kb )
//    -- only the lower block-triangle is streamed and multiplied (~half the flops of a full GEMM);
//    the packed entries are (A_kl + A_lk)/2 so the doubled term equals the reference's full sum.
//  * W never exists in HBM: each K-slab (16 individuals x 128 pairs) is produced into shared memory
//    from 2 code bytes per pair and two 4-entry LUTs, double buffered under the MMAs.
//  * one persistent CTA per SM streams all of packed A per pair tile; all CTAs walk A in the same
//    order, so A is served from the 126 MB L2 (105 MB at n=5000), not HBM.
#include "common.cuh"

namespace gb {

constexpr int EX_T = 128;            // pairs per tile
constexpr int EX_BI = 8;             // first-SNPs per scan tile
constexpr int EX_BJ = 16;            // second-SNPs per scan tile
constexpr int EX_BK = 16;            // K-slab (columns of A / individuals)
constexpr int EX_LD = EX_T + 4;      // padded leading dimension of the W slab
constexpr int EX_STAGES = 5;         // ring of {A slab, W slab} stages
constexpr int EX_AHEAD = 2;          // slabs produced ahead of the one being consumed
constexpr int EX_SLAB_ELEMS = BLOCK_ROWS * EX_BK;          // 2048 doubles = 16 KB
constexpr int EX_SLABS_PER_BLOCK = BLOCK_ROWS / EX_BK;     // 8

__host__ __device__ inline long long ex_total_slabs(long long nb) {
    return (long long)EX_SLABS_PER_BLOCK * nb * (nb + 1) / 2;
}

// ------------------------------------------------------------------------------------------
// pack_pvp: packed[slab][k4][mtp][lane][h] = sym(A)[mb*128 + (2*mtp+h)*8 + (lane>>2)][ks*16 + k4*4 + (lane&3)]
// with slab = 8*mb(mb+1)/2 + ks, ks in [0, 8*(mb+1)); zero outside n x n.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_pvp_kernel(const double *__restrict__ A, long long n, long long nb,
                                                       double *__restrict__ packed) {
    const long long total = ex_total_slabs(nb) * EX_SLAB_ELEMS;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long slab = e / EX_SLAB_ELEMS;
        const int r = (int)(e - slab * EX_SLAB_ELEMS);
        // invert slab -> (mb, ks): slabs before block-row mb = 8*mb(mb+1)/2
        long long mb = (long long)((sqrt(8.0 * (double)(slab / EX_SLABS_PER_BLOCK) + 1.0) - 1.0) * 0.5);
        while (EX_SLABS_PER_BLOCK * mb * (mb + 1) / 2 > slab) --mb;
        while (EX_SLABS_PER_BLOCK * (mb + 1) * (mb + 2) / 2 <= slab) ++mb;
        const long long ks = slab - EX_SLABS_PER_BLOCK * mb * (mb + 1) / 2;
        const int h = r & 1, lane = (r >> 1) & 31, mtp = (r >> 6) & 7, k4 = r >> 9;
        const long long row = mb * BLOCK_ROWS + (2 * mtp + h) * 8 + (lane >> 2);
        const long long col = ks * EX_BK + k4 * 4 + (lane & 3);
        double v = 0.0;
        if (row < n && col < n) v = 0.5 * (A[row * n + col] + A[col * n + row]);
        packed[e] = v;
    }
}

// ------------------------------------------------------------------------------------------
// scan tiling: I-blocks of 8 list entries; block ib tests j in [jstart, m) in J-blocks of 16.
// prefix[ib] = first tile id of I-block ib, prefix[n_iblk] = total tiles.
// ------------------------------------------------------------------------------------------
__global__ void exact_tiles_kernel(const long long *__restrict__ ilist, long long n_i, long long m,
                                   int pair_space, long long n_iblk, long long *__restrict__ prefix,
                                   long long *__restrict__ jstart) {
    for (long long ib = threadIdx.x; ib < n_iblk; ib += blockDim.x) {
        long long lo = (1LL << 62);
        for (int q = 0; q < EX_BI; ++q) {
            const long long li = ib * EX_BI + q;
            if (li < n_i) { const long long s = ilist[li]; lo = s < lo ? s : lo; }
        }
        const long long js = (pair_space == GMAT_B200_PAIRS_ALL) ? 0 : lo + 1;
        jstart[ib] = js;
        prefix[ib + 1] = (js < m) ? (m - js + EX_BJ - 1) / EX_BJ : 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        prefix[0] = 0;
        for (long long ib = 0; ib < n_iblk; ++ib) prefix[ib + 1] += prefix[ib];
    }
}

struct ExactParams {
    const unsigned char *bed;
    long long n, m, row_bytes, nb;
    const double *lutU, *lutV, *py, *packed;
    // scan mode
    const long long *ilist;
    long long n_i, n_iblk;
    const long long *prefix, *jstart;
    int pair_space;
    double chi_cut;
    gmat_b200_hit *hits;
    long long cap;
    unsigned long long *count;
    // pair mode
    const int *pairs;
    long long n_pairs;
    gmat_b200_hit *out;
};

// Shared memory of one CTA (~205 KB): a ring of EX_STAGES x {A slab (TMA destination), W slab}, the per-tile pair
// metadata, and the mbarriers that let the 16 warps run the K loop without any CTA-wide barrier.
template <int WARPS_M>
struct ExactSmem {
    double A[EX_STAGES][EX_SLAB_ELEMS];           // 16 KB each, TMA destination (16-byte aligned)
    double W[EX_STAGES][EX_BK][EX_LD];            // 16.5 KB each: w = u_i o v_j, 16 individuals x 128 pairs
    double prod[16][EX_T];                        // per pair: lutU[i][a] * lutV[j][b] at [a*4+b]
    double red_var[WARPS_M][EX_T];
    double red_eff[WARPS_M][EX_T];
    long long uoff[EX_T];                         // byte offset of the pair's U / V genotype rows
    long long voff[EX_T];
    int pi[EX_T];                                 // SNP indices (-1: masked pair)
    int pj[EX_T];
    uint64_t full[EX_STAGES];                     // A landed (TMA tx bytes) + W written (one arrival per warp)
    uint64_t empty[EX_STAGES];                    // slab consumed (one arrival per warp)
    long long tile_ib;
};

}  // namespace gb

namespace gb {

// Persistent kernel, 16 warps (4 x 4, warp tile 32 rows x 32 pairs).  Every warp is both consumer (DMMA on
// slab s) and a 1/16 producer of the W slab EX_AHEAD slabs ahead (table lookups on 2-bit codes: LDS + STS, no
// FP64 work); thread 0 also issues the TMA bulk copy of that slab of packed A.  Hand-off is by mbarriers only
// (full: tx bytes + 16 warp arrivals, empty: 16 warp arrivals), so inside a pair tile there is no CTA-wide
// barrier and warps may drift min(EX_AHEAD, EX_STAGES-EX_AHEAD-1) slabs apart: when one warp waits, the DMMA
// pipe of its SM sub-partition issues from the other three.
template <int WARPS_M, int WARPS_N, bool PAIR_MODE>
__global__ void __launch_bounds__(WARPS_M *WARPS_N * 32, 1) epi_exact_kernel(const ExactParams p) {
    constexpr int NW = WARPS_M * WARPS_N;
    constexpr int NT = NW * 32;
    constexpr int WM = BLOCK_ROWS / WARPS_M, WN = EX_T / WARPS_N;
    constexpr int MT = WM / 8, NTL = WN / 8;
    constexpr int PARTS = NT / EX_T;                 // producer threads per pair
    constexpr int KPER = EX_BK / PARTS;              // individuals per producer thread per slab
    static_assert(MT % 2 == 0, "A fragments are loaded as pairs of m-tiles (LDS.128)");
    static_assert(KPER == 4, "a producer thread expands one code byte per operand per slab");
    static_assert(EX_AHEAD >= 1 && EX_AHEAD < EX_STAGES - 1, "ring must leave room for warp skew");
    constexpr unsigned SLAB_BYTES = EX_SLAB_ELEMS * sizeof(double);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ExactSmem<WARPS_M> &S = *reinterpret_cast<ExactSmem<WARPS_M> *>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / WARPS_N, wn = warp % WARPS_N;
    const int ppair = tid % EX_T, ppart = tid / EX_T;

    const long long ntiles = PAIR_MODE ? (p.n_pairs + EX_T - 1) / EX_T : p.prefix[p.n_iblk];
    const long long my_tiles =
        ((long long)blockIdx.x < ntiles) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long s_total = ex_total_slabs(p.nb);

    if (tid == 0) {
        for (int s = 0; s < EX_STAGES; ++s) {
            mbar_init(&S.full[s], NW + 1);           // 16 warp arrivals + the expect_tx arrival of thread 0
            mbar_init(&S.empty[s], NW);
        }
        mbar_fence_init();
    }
    long long c = 0;                                 // slabs consumed so far by this CTA (all tiles)
    long long cp = 0;                                // slabs produced so far

    for (long long tl = 0; tl < my_tiles; ++tl) {
        const long long tile = blockIdx.x + tl * gridDim.x;
        __syncthreads();                             // previous tile's results consumed (also orders mbar_init)
        // ---------------- tile setup: which 128 pairs, their rows and product tables -----------------
        if (!PAIR_MODE) {
            if (tid == 0) {
                long long lo = 0, hi = p.n_iblk;     // largest ib with prefix[ib] <= tile
                while (hi - lo > 1) {
                    const long long mid = (lo + hi) >> 1;
                    if (p.prefix[mid] <= tile) lo = mid; else hi = mid;
                }
                S.tile_ib = lo;
            }
            __syncthreads();
        }
        if (tid < EX_T) {
            long long si = -1, sj = -1;
            if (PAIR_MODE) {
                const long long idx = tile * EX_T + tid;
                if (idx < p.n_pairs) { si = p.pairs[2 * idx]; sj = p.pairs[2 * idx + 1]; }
            } else {
                const long long ib = S.tile_ib;
                const long long li = ib * EX_BI + tid / EX_BJ;
                si = (li < p.n_i) ? p.ilist[li] : -1;
                sj = p.jstart[ib] + (tile - p.prefix[ib]) * EX_BJ + (tid % EX_BJ);
                if (sj >= p.m) sj = -1;
                if (p.pair_space == GMAT_B200_PAIRS_UPPER && sj <= si) sj = -1;
            }
            const bool valid = (si >= 0) && (sj >= 0) && (si < p.m) && (sj < p.m);
            S.pi[tid] = valid ? (int)si : -1;
            S.pj[tid] = valid ? (int)sj : -1;
            S.uoff[tid] = valid ? si * p.row_bytes : 0;
            S.voff[tid] = valid ? sj * p.row_bytes : 0;
            double lu[4], lv[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                lu[a] = valid ? __ldg(p.lutU + si * 4 + a) : 0.0;       // masked pair: w == 0
                lv[a] = valid ? __ldg(p.lutV + sj * 4 + a) : 0.0;
            }
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) S.prod[a * 4 + b][tid] = lu[a] * lv[b];
#pragma unroll
            for (int q = 0; q < WARPS_M; ++q) { S.red_var[q][tid] = 0.0; S.red_eff[q][tid] = 0.0; }
        }
        __syncthreads();

        // ---------------- producer side: this thread's pair, 4 individuals per slab ---------------------
        const unsigned char *urow = p.bed + S.uoff[ppair] + ppart;
        const unsigned char *vrow = p.bed + S.voff[ppair] + ppart;
        long long mbp = 0, ksp = 0;                  // (row block, K slab) of the next slab to produce
        long long sp = 0;                            // its index inside the tile, 0 .. s_total-1
        unsigned cu, cv;                             // its code bytes (already loaded)
        auto load_codes = [&](long long ks) {
            const long long off = ks * (EX_BK / 4);
            const bool ok = off + ppart < p.row_bytes;
            cu = ok ? __ldg(urow + off) : 0u;
            cv = ok ? __ldg(vrow + off) : 0u;
        };
        auto produce = [&]() {                       // slab sp of this tile -> ring stage cp % EX_STAGES
            const int stage = (int)(cp % EX_STAGES);
            const long long use = cp / EX_STAGES;
            if (use >= 1) mbar_wait(&S.empty[stage], (unsigned)((use - 1) & 1));
            if (tid == 0) {
                fence_proxy_async();                 // the consumers' LDS of this stage precede the async refill
                mbar_expect_tx(&S.full[stage], SLAB_BYTES);
                tma_bulk_g2s(S.A[stage], p.packed + sp * EX_SLAB_ELEMS, SLAB_BYTES, &S.full[stage]);
            }
            const long long k0 = ksp * EX_BK + ppart * KPER;
#pragma unroll
            for (int q = 0; q < KPER; ++q) {
                const unsigned idx = (((cu >> (2 * q)) & 3u) << 2) | ((cv >> (2 * q)) & 3u);
                // individuals past n (row padding, stray tail bits of the .bed row) must contribute exactly 0
                S.W[stage][ppart * KPER + q][ppair] = (k0 + q < p.n) ? S.prod[idx][ppair] : 0.0;
            }
            ++cp; ++sp;
            if (++ksp == (mbp + 1) * EX_SLABS_PER_BLOCK) { ksp = 0; ++mbp; }
            if (sp < s_total) load_codes(ksp);       // next slab's codes travel under this slab's MMAs
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.full[stage]);     // release: the warp's W stores are visible
        };
        load_codes(0);
#pragma unroll 1
        for (int d = 0; d < EX_AHEAD; ++d)
            if (sp < s_total) produce();

        // ---------------- K loop over the lower block-triangle of A -------------------------------------
        for (long long mb = 0; mb < p.nb; ++mb) {
            double acc[MT][NTL][2];
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
            const long long nks = (mb + 1) * EX_SLABS_PER_BLOCK;
            const long long diag0 = mb * EX_SLABS_PER_BLOCK;
            for (long long ks = 0; ks < nks; ++ks, ++c) {
                if (sp < s_total) produce();
                if (ks == diag0 && mb > 0) {         // off-diagonal blocks count twice (symmetry)
#pragma unroll
                    for (int a = 0; a < MT; ++a)
#pragma unroll
                        for (int b = 0; b < NTL; ++b) { acc[a][b][0] *= 2.0; acc[a][b][1] *= 2.0; }
                }
                const int stage = (int)(c % EX_STAGES);
                mbar_wait(&S.full[stage], (unsigned)((c / EX_STAGES) & 1));
                const double2 *Af = reinterpret_cast<const double2 *>(S.A[stage]);
#pragma unroll
                for (int k4 = 0; k4 < EX_BK / 4; ++k4) {
                    double af[MT], bf[NTL];
#pragma unroll
                    for (int a2 = 0; a2 < MT / 2; ++a2) {
                        const double2 v = Af[(k4 * 8 + wm * (MT / 2) + a2) * 32 + lane];
                        af[2 * a2] = v.x; af[2 * a2 + 1] = v.y;
                    }
#pragma unroll
                    for (int b = 0; b < NTL; ++b) bf[b] = S.W[stage][k4 * 4 + t][wn * WN + b * 8 + g];
#pragma unroll
                    for (int a = 0; a < MT; ++a)
#pragma unroll
                        for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&S.empty[stage]);
            }
            // ---------- epilogue of row block mb: var += sum_r w_r * y_r ; eff += sum_r w_r * py_r ----
#pragma unroll
            for (int b = 0; b < NTL; ++b) {
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int pair = wn * WN + b * 8 + 2 * t + cc;
                    const unsigned char *ur = p.bed + S.uoff[pair];
                    const unsigned char *vr = p.bed + S.voff[pair];
                    double v = 0.0, e = 0.0;
#pragma unroll
                    for (int a = 0; a < MT; ++a) {
                        const long long r = mb * BLOCK_ROWS + wm * WM + a * 8 + g;
                        if (r < p.n) {
                            const int idx = (code_at(ur, (int)r) << 2) | code_at(vr, (int)r);
                            const double w = S.prod[idx][pair];
                            v += w * acc[a][b][cc];
                            e += w * __ldg(p.py + r);
                        }
                    }
#pragma unroll
                    for (int o = 4; o < 32; o <<= 1) { v += shfl_xor_f64(v, o); e += shfl_xor_f64(e, o); }
                    if (g == 0) { S.red_var[wm][pair] += v; S.red_eff[wm][pair] += e; }
                }
            }
        }

        // ---------------- tile result: chi-square, threshold, compaction ----------------------------
        __syncthreads();
        if (tid < EX_T) {
            double var = 0.0, eff = 0.0;
#pragma unroll
            for (int q = 0; q < WARPS_M; ++q) { var += S.red_var[q][tid]; eff += S.red_eff[q][tid]; }
            const int si = S.pi[tid], sj = S.pj[tid];
            const bool valid = si >= 0;
            const double chi = eff * eff / var;
            gmat_b200_hit h;
            h.snp_0 = si; h.snp_1 = sj; h.eff = eff; h.var = var; h.chi = chi;
            if (PAIR_MODE) {
                if (valid) p.out[tile * EX_T + tid] = h;
            } else {
                const bool hit = valid && (p.chi_cut < 0.0 || chi > p.chi_cut);
                const unsigned ballot = __ballot_sync(0xffffffffu, hit);
                if (ballot) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(p.count, (unsigned long long)__popc(ballot));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (hit) {
                        const unsigned long long pos = base + __popc(ballot & ((1u << lane) - 1u));
                        if ((long long)pos < p.cap) p.hits[pos] = h;
                    }
                }
            }
        }
    }
}

}  // namespace gb

using namespace gb;

// warp layout of the exact kernel: 4 (rows) x 4 (pairs) warps = 512 threads, warp tile 32 x 32
constexpr int EXW_M = 4, EXW_N = 4;

extern "C" long long gmat_b200_pvp_packed_elems(long long n) {
    if (n <= 0) return 0;
    const long long nb = round_up(n, BLOCK_ROWS) / BLOCK_ROWS;
    return ex_total_slabs(nb) * EX_SLAB_ELEMS;
}

extern "C" int gmat_b200_pack_pvp(const double *pvp_dev, long long n, double *packed_dev, void *stream) {
    GB_CHECK_ARG(pvp_dev && packed_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && n < (1LL << 30), "n");
    GB_CHECK_ARG((reinterpret_cast<uintptr_t>(packed_dev) & 15) == 0, "packed_dev must be 16-byte aligned");
    const long long nb = round_up(n, BLOCK_ROWS) / BLOCK_ROWS;
    const long long total = ex_total_slabs(nb) * EX_SLAB_ELEMS;
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)sm_count() * 32;
    if (blocks > cap) blocks = cap;
    pack_pvp_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(pvp_dev, n, nb, packed_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

static int exact_common_check(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                              const double *lutU_dev, const double *lutV_dev, const double *py_dev,
                              const double *pvp_packed_dev) {
    GB_CHECK_ARG(bed_dev && lutU_dev && lutV_dev && py_dev && pvp_packed_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && m > 0 && n < (1LL << 30) && m < (1LL << 31), "sizes");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4 && row_bytes % 16 == 0, "row_bytes");
    GB_CHECK_ARG((reinterpret_cast<uintptr_t>(pvp_packed_dev) & 15) == 0, "pvp_packed_dev alignment");
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_epi_exact_scan(const unsigned char *bed_dev, long long n, long long m,
                                        long long row_bytes, const double *lutU_dev, const double *lutV_dev,
                                        const double *py_dev, const double *pvp_packed_dev,
                                        const long long *ilist_dev, long long n_i, int pair_space,
                                        double chi_cut, gmat_b200_hit *hits_dev, long long cap,
                                        unsigned long long *count_dev, void *stream) {
    int rc = exact_common_check(bed_dev, n, m, row_bytes, lutU_dev, lutV_dev, py_dev, pvp_packed_dev);
    if (rc != GMAT_B200_OK) return rc;
    GB_CHECK_ARG(ilist_dev && count_dev, "NULL pointer");
    GB_CHECK_ARG(n_i >= 0, "n_i");
    GB_CHECK_ARG(cap == 0 || hits_dev, "hits_dev is NULL");
    GB_CHECK_ARG(pair_space == GMAT_B200_PAIRS_UPPER || pair_space == GMAT_B200_PAIRS_ALL, "pair_space");
    if (n_i == 0) return GMAT_B200_OK;
    cudaStream_t st = (cudaStream_t)stream;
    ExactParams p{};
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes;
    p.nb = round_up(n, BLOCK_ROWS) / BLOCK_ROWS;
    p.lutU = lutU_dev; p.lutV = lutV_dev; p.py = py_dev; p.packed = pvp_packed_dev;
    p.ilist = ilist_dev; p.n_i = n_i; p.n_iblk = (n_i + EX_BI - 1) / EX_BI;
    p.pair_space = pair_space; p.chi_cut = chi_cut;
    p.hits = hits_dev; p.cap = cap; p.count = count_dev;
    long long *scratch = nullptr;
    GB_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&scratch), sizeof(long long) * (2 * p.n_iblk + 1), st));
    p.prefix = scratch;
    p.jstart = scratch + p.n_iblk + 1;
    exact_tiles_kernel<<<1, 1024, 0, st>>>(ilist_dev, n_i, m, pair_space, p.n_iblk, scratch,
                                           scratch + p.n_iblk + 1);
    GB_LAUNCH_CHECK();
    auto kern = epi_exact_kernel<EXW_M, EXW_N, false>;
    const size_t smem = sizeof(ExactSmem<EXW_M>) + 128;
    GB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)sm_count(), EXW_M * EXW_N * 32, smem, st>>>(p);
    GB_LAUNCH_CHECK();
    GB_CUDA(cudaFreeAsync(scratch, st));
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_epi_exact_pairs(const unsigned char *bed_dev, long long n, long long m,
                                         long long row_bytes, const double *lutU_dev, const double *lutV_dev,
                                         const double *py_dev, const double *pvp_packed_dev,
                                         const int *pairs_dev, long long n_pairs, gmat_b200_hit *out_dev,
                                         void *stream) {
    int rc = exact_common_check(bed_dev, n, m, row_bytes, lutU_dev, lutV_dev, py_dev, pvp_packed_dev);
    if (rc != GMAT_B200_OK) return rc;
    GB_CHECK_ARG(n_pairs >= 0, "n_pairs");
    if (n_pairs == 0) return GMAT_B200_OK;
    GB_CHECK_ARG(pairs_dev && out_dev, "NULL pointer");
    ExactParams p{};
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes;
    p.nb = round_up(n, BLOCK_ROWS) / BLOCK_ROWS;
    p.lutU = lutU_dev; p.lutV = lutV_dev; p.py = py_dev; p.packed = pvp_packed_dev;
    p.pairs = pairs_dev; p.n_pairs = n_pairs; p.out = out_dev;
    auto kern = epi_exact_kernel<EXW_M, EXW_N, true>;
    const size_t smem = sizeof(ExactSmem<EXW_M>) + 128;
    GB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = (n_pairs + EX_T - 1) / EX_T;
    if (grid > sm_count()) grid = sm_count();
    kern<<<(unsigned)grid, EXW_M * EXW_N * 32, smem, (cudaStream_t)stream>>>(p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
