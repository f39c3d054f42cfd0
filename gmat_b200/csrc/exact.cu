// Exact epistasis test: for a tile of BN SNP pairs, w = u_i o v_j is generated on the fly from the
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
//  * FP64 has no tcgen05 kind; the FP64 tensor path of sm_100a is the warp-level DMMA.8x8x4 (cuBLAS DGEMM
//    on this part is cutlass_80_tensorop_d884gemm).  Measured issue-rate ceiling: 37.2 TFLOP/s.
//  * A is re-tiled once per run (pack_pvp) into the exact order the kernel consumes it: lower
//    block-triangle of BM-row blocks, each BM x 16 slab stored in DMMA A-fragment order, so a slab is ONE
//    contiguous TMA bulk copy (cp.async.bulk + mbarrier complete_tx) and every fragment load is a
//    conflict-free LDS.128.  Symmetry: w'Aw = sum_mb w_mb'( A[mb,mb] w_mb + 2 sum_{kb<mb} A[mb,kb] w_kb )
//    -- only the lower block-triangle is streamed and multiplied (~half the flops of a full GEMM);
//    the packed entries are (A_kl + A_lk)/2 so the doubled term equals the reference's full sum.
//  * W never exists in HBM: each K-slab (16 individuals x BN pairs) is produced into shared memory
//    from code bytes and two 4-entry LUTs per pair, into an XOR-swizzled (padding-free) layout.
//  * the K loop has no CTA-wide barrier: a ring of {A slab, W slab} stages is handed between the warps with
//    mbarriers only (full: TMA tx bytes + one arrival per warp; empty: one arrival per warp).
//  * several small independent CTAs per SM (EXACT_CFG 0: 4 CTAs of 4 warps, 64 x 64 tiles) instead of one
//    large one: a warp that waits at a slab boundary never idles the DMMA pipe of its SM sub-partition,
//    because the other resident warps there belong to other CTAs at unrelated phases (the same reason
//    cuBLAS' d884gemm uses 64x32 tiles).  Every CTA walks packed A in the same order, so A is served from
//    the 126 MB L2 (52 MB packed at n=5000, BM=64), not HBM.
#include "common.cuh"
#include <cstdlib>

namespace gb {

constexpr int EX_BK = 16;            // K-slab (columns of A / individuals)
constexpr int EX_BJ = 16;            // second-SNPs per scan tile (first-SNPs per tile = BN / 16)

__host__ __device__ inline long long ex_total_slabs(long long nb, int bm) {
    return (long long)(bm / EX_BK) * nb * (nb + 1) / 2;
}

// ------------------------------------------------------------------------------------------
// pack_pvp (row-block size bm = 64 or 128):
//   packed[slab][k4][mtp][lane][h] = sym(A)[mb*bm + (2*mtp+h)*8 + (lane>>2)][ks*16 + k4*4 + (lane&3)]
// with slab = (bm/16)*mb(mb+1)/2 + ks, ks in [0, (bm/16)*(mb+1)), mtp in [0, bm/16); zero outside n x n.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_pvp_kernel(const double *__restrict__ A, long long n, long long nb,
                                                       int bm, double *__restrict__ packed) {
    const int spb = bm / EX_BK;                       // slabs per block column, also m-tile pairs per slab
    const int slab_elems = bm * EX_BK;
    const long long total = ex_total_slabs(nb, bm) * slab_elems;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long slab = e / slab_elems;
        const int r = (int)(e - slab * slab_elems);
        // invert slab -> (mb, ks): slabs before block-row mb = spb*mb(mb+1)/2
        long long mb = (long long)((sqrt(8.0 * (double)(slab / spb) + 1.0) - 1.0) * 0.5);
        while (spb * mb * (mb + 1) / 2 > slab) --mb;
        while (spb * (mb + 1) * (mb + 2) / 2 <= slab) ++mb;
        const long long ks = slab - spb * mb * (mb + 1) / 2;
        const int h = r & 1, lane = (r >> 1) & 31, mtp = (r >> 6) % spb, k4 = (r >> 6) / spb;
        const long long row = mb * bm + (2 * mtp + h) * 8 + (lane >> 2);
        const long long col = ks * EX_BK + k4 * 4 + (lane & 3);
        double v = 0.0;
        if (row < n && col < n) v = 0.5 * (A[row * n + col] + A[col * n + row]);
        packed[e] = v;
    }
}

// ------------------------------------------------------------------------------------------
// scan tiling: I-blocks of `bi` list entries; block ib tests j in [jstart, m) in J-blocks of 16.
// prefix[ib] = first tile id of I-block ib, prefix[n_iblk] = total tiles.
// ------------------------------------------------------------------------------------------
__global__ void exact_tiles_kernel(const long long *__restrict__ ilist, long long n_i, long long m,
                                   int pair_space, int bi, long long n_iblk, long long *__restrict__ prefix,
                                   long long *__restrict__ jstart) {
    for (long long ib = threadIdx.x; ib < n_iblk; ib += blockDim.x) {
        long long lo = (1LL << 62);
        for (int q = 0; q < bi; ++q) {
            const long long li = ib * bi + q;
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

// Row-block epilogue shared by the exact kernels: var += sum_r w_r * y_r, eff += sum_r w_r * py_r for the rows of
// this warp (8 * MT consecutive individuals starting at r0) and its 8 * NTL pairs (pair0 ..).  w is re-derived
// from the 2-bit codes: ONE 8-byte word per 32 rows per pair and operand, all loads issued before any use (the
// first version fetched single code bytes inside the accumulate chains and cost as much as the MMAs).
// lutU / lutV point at [4][BN] tables in shared memory; red_var / red_eff at this warp-row's [BN] accumulators.
template <int MT, int NTL, int BN>
__device__ __forceinline__ void exact_row_epilogue(const double (&acc)[MT][NTL][2], const unsigned char *bed,
                                                   long long row_bytes, long long n, long long r0, int pair0,
                                                   int g, int t, const double *__restrict__ py,
                                                   const long long *uoff, const long long *voff,
                                                   const double *lutU, const double *lutV, double *red_var,
                                                   double *red_eff) {
    constexpr int NWORD = MT / 4;                    // 64-bit code words per pair and operand (32 rows each)
    static_assert(MT % 4 == 0, "rows per warp must be a multiple of 32");
    const long long boff = r0 >> 2;                  // r0 is a multiple of 32: 8-byte aligned
    double pyv[MT];
#pragma unroll
    for (int a = 0; a < MT; ++a) {
        const long long r = r0 + a * 8 + g;
        pyv[a] = (r < n) ? __ldg(py + r) : 0.0;
    }
#pragma unroll
    for (int b0 = 0; b0 < NTL; b0 += 2) {            // two n-tiles (4 pairs per thread) per batch of loads
        unsigned long long cwu[2][2][NWORD], cwv[2][2][NWORD];
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int pair = pair0 + (b0 + b) * 8 + 2 * t + cc;
                const unsigned char *ur = bed + uoff[pair] + boff, *vr = bed + voff[pair] + boff;
#pragma unroll
                for (int x = 0; x < NWORD; ++x) {
                    const bool ok = boff + 8 * (x + 1) <= row_bytes;
                    cwu[b][cc][x] = ok ? __ldg(reinterpret_cast<const unsigned long long *>(ur) + x) : 0ull;
                    cwv[b][cc][x] = ok ? __ldg(reinterpret_cast<const unsigned long long *>(vr) + x) : 0ull;
                }
            }
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int pair = pair0 + (b0 + b) * 8 + 2 * t + cc;
                double v = 0.0, e = 0.0;
#pragma unroll
                for (int a = 0; a < MT; ++a) {
                    const int k = (a & 3) * 8 + g;   // individual inside the 32-row word
                    const int c_u = (int)(cwu[b][cc][a >> 2] >> (2 * k)) & 3;
                    const int c_v = (int)(cwv[b][cc][a >> 2] >> (2 * k)) & 3;
                    double w = lutU[c_u * BN + pair] * lutV[c_v * BN + pair];
                    if (r0 + a * 8 + g >= n) w = 0.0;        // padding rows
                    v += w * acc[a][b0 + b][cc];
                    e += w * pyv[a];
                }
#pragma unroll
                for (int o = 4; o < 32; o <<= 1) { v += shfl_xor_f64(v, o); e += shfl_xor_f64(e, o); }
                if (g == 0) { red_var[pair] += v; red_eff[pair] += e; }
            }
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
    int dbg;                                       // timing experiments only (GMAT_B200_EXACT_DBG), 0 in production
};

// Shared memory of one CTA: a ring of NS x {A slab (TMA destination), W slab}, the per-tile pair metadata,
// and the mbarriers that sequence the warps.
template <int BM, int BN, int WARPS_M, int NS, int SPS>
struct ExactSmem {
    double A[NS][SPS * BM * EX_BK];                  // TMA destination (16-byte aligned), fragment order
    double W[NS][SPS * EX_BK][BN];                   // w = u_i o v_j; column index XOR-swizzled by (k & 3) << 2
    double lutU[4][BN];
    double lutV[4][BN];
    double red_var[WARPS_M][BN];
    double red_eff[WARPS_M][BN];
    long long uoff[BN];                           // byte offset of the pair's U / V genotype rows
    long long voff[BN];
    int pi[BN];                                   // SNP indices (-1: masked pair)
    int pj[BN];
    uint64_t full[NS];                            // A landed (TMA tx bytes) + W written (one arrival per warp)
    uint64_t empty[NS];                           // slab consumed (one arrival per warp)
    long long tile_ib;
};

// Persistent kernel; WARPS_M x WARPS_N warps, warp tile 32 rows x 32 pairs.  Every warp is both consumer
// (DMMA on slab s) and a 1/NW producer of the W slab AHEAD slabs ahead (table lookups on 2-bit codes + one
// DMUL per element); thread 0 also issues the TMA bulk copy of that slab of packed A.
template <int BM, int BN, int WARPS_M, int WARPS_N, int NS, int AHEAD, int SPS, int CTAS_PER_SM, bool PAIR_MODE>
__global__ void __launch_bounds__(WARPS_M *WARPS_N * 32, CTAS_PER_SM) epi_exact_kernel(const ExactParams p) {
    constexpr int NW = WARPS_M * WARPS_N;
    constexpr int NT = NW * 32;
    constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;
    constexpr int MT = WM / 8, NTL = WN / 8;
    constexpr int PARTS = NT / BN;                   // producer threads per pair
    constexpr int KPER = EX_BK / PARTS;              // individuals per producer thread per slab
    constexpr int SPB = BM / EX_BK;                  // slabs per block column
    constexpr int BI = BN / EX_BJ;                   // first-SNPs per scan tile
    static_assert(MT % 2 == 0, "A fragments are loaded as pairs of m-tiles (LDS.128)");
    static_assert(KPER == 4 || KPER == 8, "a producer thread expands one or two code bytes per operand per slab");
    static_assert(AHEAD >= 1 && AHEAD < NS, "ring");
    static_assert(BN <= NT, "tile setup uses one thread per pair");
    constexpr unsigned SLAB_BYTES = SPS * BM * EX_BK * sizeof(double);      // one ring stage = SPS slabs
    static_assert((BM / EX_BK) % SPS == 0, "a stage must not straddle row blocks");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ExactSmem<BM, BN, WARPS_M, NS, SPS> &S = *reinterpret_cast<ExactSmem<BM, BN, WARPS_M, NS, SPS> *>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / WARPS_N, wn = warp % WARPS_N;
    const int ppair = tid % BN, ppart = tid / BN;

    const long long ntiles = PAIR_MODE ? (p.n_pairs + BN - 1) / BN : p.prefix[p.n_iblk];
    const long long my_tiles =
        ((long long)blockIdx.x < ntiles) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long s_total = ex_total_slabs(p.nb, BM);

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&S.full[s], NW + 1);           // NW warp arrivals + the expect_tx arrival of thread 0
            mbar_init(&S.empty[s], NW);
        }
        mbar_fence_init();
    }
    long long c = 0;                                 // slabs consumed so far by this CTA (all tiles)
    long long cp = 0;                                // slabs produced so far

    for (long long tl = 0; tl < my_tiles; ++tl) {
        const long long tile = blockIdx.x + tl * gridDim.x;
        __syncthreads();                             // previous tile's results consumed (also orders mbar_init)
        // ---------------- tile setup: which BN pairs, their rows and LUTs -----------------------------
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
        if (tid < BN) {
            long long si = -1, sj = -1;
            if (PAIR_MODE) {
                const long long idx = tile * BN + tid;
                if (idx < p.n_pairs) { si = p.pairs[2 * idx]; sj = p.pairs[2 * idx + 1]; }
            } else {
                const long long ib = S.tile_ib;
                const long long li = ib * BI + tid / EX_BJ;
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
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                S.lutU[a][tid] = valid ? __ldg(p.lutU + si * 4 + a) : 0.0;       // masked pair: w == 0
                S.lutV[a][tid] = valid ? __ldg(p.lutV + sj * 4 + a) : 0.0;
            }
#pragma unroll
            for (int q = 0; q < WARPS_M; ++q) { S.red_var[q][tid] = 0.0; S.red_eff[q][tid] = 0.0; }
        }
        __syncthreads();

        // ---------------- producer side: this thread's pair, KPER individuals per slab -------------------
        const unsigned char *urow = p.bed + S.uoff[ppair] + ppart * (KPER / 4);
        const unsigned char *vrow = p.bed + S.voff[ppair] + ppart * (KPER / 4);
        double lu[4], lv[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { lu[a] = S.lutU[a][ppair]; lv[a] = S.lutV[a][ppair]; }
        long long mbp = 0, ksp = 0;                  // (row block, K slab) of the next slab to produce
        long long sp = 0;                            // its index inside the tile, 0 .. s_total-1
        unsigned cu[SPS], cv[SPS];                   // its code bytes (already loaded)
        auto load_codes = [&](long long ks) {
#pragma unroll
            for (int x = 0; x < SPS; ++x) {
                const long long off = (ks + x) * (EX_BK / 4);
                const bool ok = off + (ppart + 1) * (KPER / 4) <= p.row_bytes;
                if (KPER == 8) {
                    cu[x] = ok ? __ldg(reinterpret_cast<const unsigned short *>(urow + off)) : 0u;
                    cv[x] = ok ? __ldg(reinterpret_cast<const unsigned short *>(vrow + off)) : 0u;
                } else {
                    cu[x] = ok ? __ldg(urow + off) : 0u;
                    cv[x] = ok ? __ldg(vrow + off) : 0u;
                }
            }
        };
        auto produce = [&]() {                       // slab sp of this tile -> ring stage cp % NS
            const int stage = (int)(cp % NS);
            const long long use = cp / NS;
            if (use >= 1 && !(p.dbg & 64)) mbar_wait(&S.empty[stage], (unsigned)((use - 1) & 1));
            if (tid == 0 && !(p.dbg & 64)) {
                if (p.dbg & 4) {
                    mbar_arrive(&S.full[stage]);
                } else {
                    fence_proxy_async();             // the consumers' LDS of this stage precede the async refill
                    mbar_expect_tx(&S.full[stage], SLAB_BYTES);
                    tma_bulk_g2s(S.A[stage], p.packed + sp * (BM * EX_BK), SLAB_BYTES, &S.full[stage]);
                }
            }
            if (!(p.dbg & 2))
#pragma unroll
            for (int x = 0; x < SPS; ++x) {
                const long long k0 = (ksp + x) * EX_BK + ppart * KPER;
#pragma unroll
                for (int q = 0; q < KPER; ++q) {
                    const int c_u = (cu[x] >> (2 * q)) & 3, c_v = (cv[x] >> (2 * q)) & 3;
                    const double u = (c_u & 2) ? ((c_u & 1) ? lu[3] : lu[2]) : ((c_u & 1) ? lu[1] : lu[0]);
                    const double v = (c_v & 2) ? ((c_v & 1) ? lv[3] : lv[2]) : ((c_v & 1) ? lv[1] : lv[0]);
                    const int k = x * EX_BK + ppart * KPER + q;
                    // individuals past n (row padding, stray tail bits of the .bed row) must contribute exactly 0
                    S.W[stage][k][ppair ^ ((k & 3) << 2)] = (k0 + q < p.n) ? u * v : 0.0;
                }
            }
            ++cp; sp += SPS;
            ksp += SPS;
            if (ksp == (mbp + 1) * SPB) { ksp = 0; ++mbp; }
            if (sp < s_total) load_codes(ksp);       // next slab's codes travel under this slab's MMAs
            __syncwarp();
            if (lane == 0 && !(p.dbg & 64)) mbar_arrive(&S.full[stage]);     // release: the warp's W stores are visible
        };
        load_codes(0);
#pragma unroll 1
        for (int d = 0; d < AHEAD; ++d)
            if (sp < s_total) produce();
        if (p.dbg & 32) {                            // stagger the warps of one SM sub-partition (warp % 4)
            const long long t0 = clock64(), wait = (long long)(warp >> 2) * ((p.dbg >> 8) ? (p.dbg >> 8) : 600);
            while (clock64() - t0 < wait) {
            }
        }

        // ---------------- K loop over the lower block-triangle of A -------------------------------------
        for (long long mb = 0; mb < p.nb; ++mb) {
            double acc[MT][NTL][2];
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
            const long long nks = (mb + 1) * SPB;
            const long long diag0 = mb * SPB;
            for (long long ks = 0; ks < nks; ks += SPS, ++c) {
                if (sp < s_total) produce();
                if (ks == diag0 && mb > 0) {         // off-diagonal blocks count twice (symmetry)
#pragma unroll
                    for (int a = 0; a < MT; ++a)
#pragma unroll
                        for (int b = 0; b < NTL; ++b) { acc[a][b][0] *= 2.0; acc[a][b][1] *= 2.0; }
                }
                const int stage = (int)(c % NS);
                if (!(p.dbg & 64)) mbar_wait(&S.full[stage], (unsigned)((c / NS) & 1));
                const double2 *Af = reinterpret_cast<const double2 *>(S.A[(p.dbg & 8) ? 0 : stage]);
                if (!(p.dbg & 16))
#pragma unroll
                for (int k4 = 0; k4 < SPS * EX_BK / 4; ++k4) {
                    double af[MT], bf[NTL];
#pragma unroll
                    for (int a2 = 0; a2 < MT / 2; ++a2) {
                        const double2 v = Af[(k4 >> 2) * (BM * EX_BK / 2) + ((k4 & 3) * SPB + wm * (MT / 2) + a2) * 32 + lane];
                        af[2 * a2] = v.x; af[2 * a2 + 1] = v.y;
                    }
#pragma unroll
                    for (int b = 0; b < NTL; ++b)
                        bf[b] = S.W[stage][k4 * 4 + t][(wn * WN + b * 8 + g) ^ (t << 2)];
#pragma unroll
                    for (int a = 0; a < MT; ++a)
#pragma unroll
                        for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
                }
                __syncwarp();
                if (lane == 0 && !(p.dbg & 64)) mbar_arrive(&S.empty[stage]);
            }
            // ---------- epilogue of row block mb: var += sum_r w_r * y_r ; eff += sum_r w_r * py_r ----
            if (p.dbg & 1) {
                double z = 0.0;
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b) z += acc[a][b][0] + acc[a][b][1];
                if (z == 12345.678) S.red_var[0][0] = z;
            } else {
                exact_row_epilogue<MT, NTL, BN>(acc, p.bed, p.row_bytes, p.n, mb * BM + wm * WM, wn * WN, g, t, p.py,
                                                S.uoff, S.voff, &S.lutU[0][0], &S.lutV[0][0], S.red_var[wm],
                                                S.red_eff[wm]);
            }
        }

        // ---------------- tile result: chi-square, threshold, compaction ----------------------------
        __syncthreads();
        if (tid < BN) {
            double var = 0.0, eff = 0.0;
#pragma unroll
            for (int q = 0; q < WARPS_M; ++q) { var += S.red_var[q][tid]; eff += S.red_eff[q][tid]; }
            const int si = S.pi[tid], sj = S.pj[tid];
            const bool valid = si >= 0;
            const double chi = eff * eff / var;
            gmat_b200_hit h;
            h.snp_0 = si; h.snp_1 = sj; h.eff = eff; h.var = var; h.chi = chi;
            if (PAIR_MODE) {
                if (valid) p.out[tile * BN + tid] = h;
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

// ==========================================================================================
// Warp-specialised variant (EXACT_CFG 3): one CTA per SM, 128 x 128 tile.
//   warps 0..7  consumers, 2 (rows) x 4 (pairs), warp tile 64 x 32 = 8 x 4 DMMA accumulator tiles: they only
//               wait on `full`, load fragments, issue DMMA and release `empty`; two consumer warps per SM
//               sub-partition, 128 DMMAs each per slab;
//   warps 8..11 producers (one pair per thread): tile metadata, the TMA bulk copy of the next packed A slab
//               and the next W slab, up to WS_NS slabs ahead of the slowest consumer.
// Register split by setmaxnreg (producers 40, consumers 232).
// ==========================================================================================
constexpr int WS_BM = 128, WS_BN = 128, WS_NS = 5;
constexpr int WS_CW = 8;                          // consumer warps
constexpr int WS_PW = 4;                          // producer warps
constexpr int WS_WARPS_M = 2, WS_WARPS_N = 4;

struct ExactWsSmem {
    double A[WS_NS][WS_BM * EX_BK];
    double W[WS_NS][EX_BK][WS_BN];
    double lutU[2][4][WS_BN];
    double lutV[2][4][WS_BN];
    double red_var[2][WS_WARPS_M][WS_BN];
    double red_eff[2][WS_WARPS_M][WS_BN];
    long long uoff[2][WS_BN];
    long long voff[2][WS_BN];
    int pi[2][WS_BN];
    int pj[2][WS_BN];
    uint64_t full[WS_NS], empty[WS_NS];
    uint64_t meta_full[2], meta_empty[2];
};

template <bool PAIR_MODE>
__global__ void __launch_bounds__((WS_CW + WS_PW) * 32, 1) epi_exact_ws_kernel(const ExactParams p) {
    constexpr int BM = WS_BM, BN = WS_BN, NS = WS_NS;
    constexpr int SPB = BM / EX_BK;
    constexpr int BI = BN / EX_BJ;
    constexpr int MT = 8, NTL = 4;                   // warp tile 64 x 32
    constexpr unsigned SLAB_BYTES = BM * EX_BK * sizeof(double);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ExactWsSmem &S = *reinterpret_cast<ExactWsSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const long long ntiles = PAIR_MODE ? (p.n_pairs + BN - 1) / BN : p.prefix[p.n_iblk];
    const long long my_tiles =
        ((long long)blockIdx.x < ntiles) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long s_total = ex_total_slabs(p.nb, BM);

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&S.full[s], WS_PW + 1);        // one arrival per producer warp + the expect_tx arrival
            mbar_init(&S.empty[s], WS_CW);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&S.meta_full[s], WS_PW);
            mbar_init(&S.meta_empty[s], WS_CW);
        }
        mbar_fence_init();
    }
    for (int x = tid; x < 2 * WS_WARPS_M * BN; x += blockDim.x) {
        (&S.red_var[0][0][0])[x] = 0.0;
        (&S.red_eff[0][0][0])[x] = 0.0;
    }
    __syncthreads();

    if (warp >= WS_CW) {
        // =============================== producer warps ================================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
        const int pr = tid - WS_CW * 32;             // this thread's pair, 0..127
        long long cp = 0;                            // slabs produced so far (all tiles)
        for (long long tl = 0; tl < my_tiles; ++tl) {
            const long long tile = blockIdx.x + tl * gridDim.x;
            const int tp = (int)(tl & 1);
            const long long tuse = tl >> 1;
            if (tuse >= 1) mbar_wait(&S.meta_empty[tp], (unsigned)((tuse - 1) & 1));
            // ---------------- tile setup: pair pr of this tile -----------------------------------------
            long long si = -1, sj = -1;
            if (PAIR_MODE) {
                const long long idx = tile * BN + pr;
                if (idx < p.n_pairs) { si = p.pairs[2 * idx]; sj = p.pairs[2 * idx + 1]; }
            } else {
                long long lo = 0, hi = p.n_iblk;     // largest ib with prefix[ib] <= tile (uniform)
                while (hi - lo > 1) {
                    const long long mid = (lo + hi) >> 1;
                    if (p.prefix[mid] <= tile) lo = mid; else hi = mid;
                }
                const long long ib = lo;
                const long long li = ib * BI + pr / EX_BJ;
                si = (li < p.n_i) ? p.ilist[li] : -1;
                sj = p.jstart[ib] + (tile - p.prefix[ib]) * EX_BJ + (pr % EX_BJ);
                if (sj >= p.m) sj = -1;
                if (p.pair_space == GMAT_B200_PAIRS_UPPER && sj <= si) sj = -1;
            }
            const bool valid = (si >= 0) && (sj >= 0) && (si < p.m) && (sj < p.m);
            S.pi[tp][pr] = valid ? (int)si : -1;
            S.pj[tp][pr] = valid ? (int)sj : -1;
            const long long uo = valid ? si * p.row_bytes : 0, vo = valid ? sj * p.row_bytes : 0;
            S.uoff[tp][pr] = uo;
            S.voff[tp][pr] = vo;
            double lu[4], lv[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                lu[a] = valid ? __ldg(p.lutU + si * 4 + a) : 0.0;         // masked pair: w == 0
                lv[a] = valid ? __ldg(p.lutV + sj * 4 + a) : 0.0;
                S.lutU[tp][a][pr] = lu[a];
                S.lutV[tp][a][pr] = lv[a];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.meta_full[tp]);
            // ---------------- slabs of this tile, in the order the consumers walk them ------------------
            const unsigned char *urow = p.bed + uo, *vrow = p.bed + vo;
            unsigned cu = 0, cv = 0;
            auto load_codes = [&](long long ks) {
                const long long off = ks * (EX_BK / 4);
                const bool ok = off + 4 <= p.row_bytes;
                cu = ok ? __ldg(reinterpret_cast<const unsigned *>(urow + off)) : 0u;
                cv = ok ? __ldg(reinterpret_cast<const unsigned *>(vrow + off)) : 0u;
            };
            load_codes(0);
            long long sp = 0;
            for (long long mb = 0; mb < p.nb; ++mb) {
                const long long nks = (mb + 1) * SPB;
                for (long long ks = 0; ks < nks; ++ks, ++cp, ++sp) {
                    const int stage = (int)(cp % NS);
                    const long long use = cp / NS;
                    if (use >= 1) mbar_wait(&S.empty[stage], (unsigned)((use - 1) & 1));
                    if (pr == 0) {
                        fence_proxy_async();         // the consumers' LDS of this stage precede the async refill
                        mbar_expect_tx(&S.full[stage], SLAB_BYTES);
                        tma_bulk_g2s(S.A[stage], p.packed + sp * (BM * EX_BK), SLAB_BYTES, &S.full[stage]);
                    }
                    const unsigned wu = cu, wv = cv;
                    if (sp + 1 < s_total) load_codes((ks + 1 < nks) ? ks + 1 : 0);    // prefetch next slab's codes
                    const long long k0 = ks * EX_BK;
#pragma unroll
                    for (int k = 0; k < EX_BK; ++k) {
                        const int c_u = (wu >> (2 * k)) & 3, c_v = (wv >> (2 * k)) & 3;
                        const double u = (c_u & 2) ? ((c_u & 1) ? lu[3] : lu[2]) : ((c_u & 1) ? lu[1] : lu[0]);
                        const double v = (c_v & 2) ? ((c_v & 1) ? lv[3] : lv[2]) : ((c_v & 1) ? lv[1] : lv[0]);
                        // individuals past n (row padding, stray tail bits) must contribute exactly 0
                        S.W[stage][k][pr ^ ((k & 3) << 2)] = (k0 + k < p.n) ? u * v : 0.0;
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&S.full[stage]);      // release: the warp's W stores are visible
                }
            }
        }
    } else {
        // =============================== consumer warps ================================================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n");
        const int g = lane >> 2, t = lane & 3;
        const int wm = warp / WS_WARPS_N, wn = warp % WS_WARPS_N;
        long long c = 0;
        for (long long tl = 0; tl < my_tiles; ++tl) {
            const long long tile = blockIdx.x + tl * gridDim.x;
            const int tp = (int)(tl & 1);
            const unsigned tpar = (unsigned)((tl >> 1) & 1);
            for (long long mb = 0; mb < p.nb; ++mb) {
                double acc[MT][NTL][2];
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
                const long long nks = (mb + 1) * SPB;
                const long long diag0 = mb * SPB;
                for (long long ks = 0; ks < nks; ++ks, ++c) {
                    if (ks == diag0 && mb > 0) {     // off-diagonal blocks count twice (symmetry)
#pragma unroll
                        for (int a = 0; a < MT; ++a)
#pragma unroll
                            for (int b = 0; b < NTL; ++b) { acc[a][b][0] *= 2.0; acc[a][b][1] *= 2.0; }
                    }
                    const int stage = (int)(c % NS);
                    mbar_wait(&S.full[stage], (unsigned)((c / NS) & 1));
                    const double2 *Af = reinterpret_cast<const double2 *>(S.A[stage]);
#pragma unroll
                    for (int k4 = 0; k4 < EX_BK / 4; ++k4) {
                        double af[MT], bf[NTL];
#pragma unroll
                        for (int a2 = 0; a2 < MT / 2; ++a2) {
                            const double2 v = Af[(k4 * SPB + wm * (MT / 2) + a2) * 32 + lane];
                            af[2 * a2] = v.x; af[2 * a2 + 1] = v.y;
                        }
#pragma unroll
                        for (int b = 0; b < NTL; ++b)
                            bf[b] = S.W[stage][k4 * 4 + t][(wn * 32 + b * 8 + g) ^ (t << 2)];
#pragma unroll
                        for (int a = 0; a < MT; ++a)
#pragma unroll
                            for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&S.empty[stage]);
                }
                // ---------- epilogue of row block mb: var += sum_r w_r * y_r ; eff += sum_r w_r * py_r ----
                if (mb == 0) mbar_wait(&S.meta_full[tp], tpar);
                exact_row_epilogue<MT, NTL, BN>(acc, p.bed, p.row_bytes, p.n, mb * BM + wm * 64, wn * 32, g, t, p.py,
                                                S.uoff[tp], S.voff[tp], &S.lutU[tp][0][0], &S.lutV[tp][0][0],
                                                S.red_var[tp][wm], S.red_eff[tp][wm]);
            }
            // ---------------- tile result: chi-square, threshold, compaction ----------------------------
            named_bar_sync(1, WS_CW * 32);           // consumers only; the producers are already slabs ahead
            if (tid < BN) {
                double var = 0.0, eff = 0.0;
#pragma unroll
                for (int q = 0; q < WS_WARPS_M; ++q) {
                    var += S.red_var[tp][q][tid]; eff += S.red_eff[tp][q][tid];
                    S.red_var[tp][q][tid] = 0.0; S.red_eff[tp][q][tid] = 0.0;       // ready for tile tl + 2
                }
                const int si = S.pi[tp][tid], sj = S.pj[tp][tid];
                const bool valid = si >= 0;
                const double chi = eff * eff / var;
                gmat_b200_hit h;
                h.snp_0 = si; h.snp_1 = sj; h.eff = eff; h.var = var; h.chi = chi;
                if (PAIR_MODE) {
                    if (valid) p.out[tile * BN + tid] = h;
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
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.meta_empty[tp]);
        }
    }
}

// ------------------------------------------------------------------------------------------
// configurations (GMAT_B200_EXACT_CFG selects one at library load; the packed layout depends on BM)
// ------------------------------------------------------------------------------------------
struct ExactCfg {
    int bm, bn, threads, ctas_per_sm;
    size_t smem;
    void (*scan)(const ExactParams);
    void (*pairs)(const ExactParams);
};

template <int BM, int BN, int WMm, int WNn, int NS, int AHEAD, int SPS, int CPS>
static ExactCfg make_cfg() {
    ExactCfg c;
    c.bm = BM; c.bn = BN; c.threads = WMm * WNn * 32; c.ctas_per_sm = CPS;
    c.smem = sizeof(ExactSmem<BM, BN, WMm, NS, SPS>);
    c.scan = epi_exact_kernel<BM, BN, WMm, WNn, NS, AHEAD, SPS, CPS, false>;
    c.pairs = epi_exact_kernel<BM, BN, WMm, WNn, NS, AHEAD, SPS, CPS, true>;
    return c;
}

static const ExactCfg &exact_cfg() {
    static const ExactCfg cfgs[6] = {
        make_cfg<64, 64, 2, 2, 3, 1, 1, 4>(),        // 0: 4 CTAs/SM x 4 warps, 64 x 64 tiles
        make_cfg<128, 64, 4, 2, 4, 2, 1, 2>(),       // 1: 2 CTAs/SM x 8 warps, 128 x 64 tiles  (default: fastest measured)
        make_cfg<128, 128, 4, 4, 5, 2, 1, 1>(),      // 2: 1 CTA/SM x 16 warps, 128 x 128 tiles
        {WS_BM, WS_BN, (WS_CW + WS_PW) * 32, 1, sizeof(ExactWsSmem),       // 3: warp-specialised 8 + 4 warps
         epi_exact_ws_kernel<false>, epi_exact_ws_kernel<true>},
        make_cfg<128, 128, 4, 4, 3, 1, 2, 1>(),      // 4: as 2, but 32-individual stages (2 slabs), 3-deep ring
        make_cfg<128, 64, 4, 2, 2, 1, 2, 2>(),       // 5: as 1, but 32-individual stages, double buffered
    };
    static const int sel = [] {
        const char *e = getenv("GMAT_B200_EXACT_CFG");
        const int v = e ? atoi(e) : 1;
        return (v >= 0 && v < 6) ? v : 1;
    }();
    return cfgs[sel];
}

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_exact_config(int *bm, int *bn, int *ctas_per_sm) {
    const ExactCfg &cfg = exact_cfg();
    if (bm) *bm = cfg.bm;
    if (bn) *bn = cfg.bn;
    if (ctas_per_sm) *ctas_per_sm = cfg.ctas_per_sm;
    return GMAT_B200_OK;
}

extern "C" long long gmat_b200_pvp_packed_elems(long long n) {
    if (n <= 0) return 0;
    const int bm = exact_cfg().bm;
    const long long nb = round_up(n, bm) / bm;
    return ex_total_slabs(nb, bm) * bm * EX_BK;
}

extern "C" int gmat_b200_pack_pvp(const double *pvp_dev, long long n, double *packed_dev, void *stream) {
    GB_CHECK_ARG(pvp_dev && packed_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && n < (1LL << 30), "n");
    GB_CHECK_ARG((reinterpret_cast<uintptr_t>(packed_dev) & 15) == 0, "packed_dev must be 16-byte aligned");
    const int bm = exact_cfg().bm;
    const long long nb = round_up(n, bm) / bm;
    const long long total = ex_total_slabs(nb, bm) * bm * EX_BK;
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)sm_count() * 32;
    if (blocks > cap) blocks = cap;
    pack_pvp_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(pvp_dev, n, nb, bm, packed_dev);
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
    const ExactCfg &cfg = exact_cfg();
    const int bi = cfg.bn / EX_BJ;
    cudaStream_t st = (cudaStream_t)stream;
    ExactParams p{};
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes;
    p.nb = round_up(n, cfg.bm) / cfg.bm;
    p.lutU = lutU_dev; p.lutV = lutV_dev; p.py = py_dev; p.packed = pvp_packed_dev;
    p.ilist = ilist_dev; p.n_i = n_i; p.n_iblk = (n_i + bi - 1) / bi;
    p.pair_space = pair_space; p.chi_cut = chi_cut;
    p.hits = hits_dev; p.cap = cap; p.count = count_dev;
    { const char *e = getenv("GMAT_B200_EXACT_DBG"); p.dbg = e ? atoi(e) : 0; }
    long long *scratch = nullptr;
    GB_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&scratch), sizeof(long long) * (2 * p.n_iblk + 1), st));
    p.prefix = scratch;
    p.jstart = scratch + p.n_iblk + 1;
    exact_tiles_kernel<<<1, 1024, 0, st>>>(ilist_dev, n_i, m, pair_space, bi, p.n_iblk, scratch,
                                           scratch + p.n_iblk + 1);
    GB_LAUNCH_CHECK();
    GB_CUDA(cudaFuncSetAttribute(cfg.scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    // upper bound of the tile count (exact count lives on the device): every CTA strides over the tile ids
    const unsigned grid = (unsigned)(sm_count() * cfg.ctas_per_sm);
    cfg.scan<<<grid, cfg.threads, cfg.smem, st>>>(p);
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
    const ExactCfg &cfg = exact_cfg();
    ExactParams p{};
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes;
    p.nb = round_up(n, cfg.bm) / cfg.bm;
    p.lutU = lutU_dev; p.lutV = lutV_dev; p.py = py_dev; p.packed = pvp_packed_dev;
    p.pairs = pairs_dev; p.n_pairs = n_pairs; p.out = out_dev;
    GB_CUDA(cudaFuncSetAttribute(cfg.pairs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    long long grid = (n_pairs + cfg.bn - 1) / cfg.bn;
    const long long maxgrid = (long long)sm_count() * cfg.ctas_per_sm;
    if (grid > maxgrid) grid = maxgrid;
    cfg.pairs<<<(unsigned)grid, cfg.threads, cfg.smem, (cudaStream_t)stream>>>(p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
