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
//    on this part is cutlass_80_tensorop_d884gemm).  Measured issue-rate ceiling: 36.9 TFLOP/s (one DMMA per
//    16 clocks per SM sub-partition), reached by a single warp per sub-partition with an 8 x 4 accumulator tile.
//  * A is re-tiled once per run (pack_pvp) into the exact order the kernel consumes it: lower
//    block-triangle of 128-row blocks, each 128 x 16 slab stored in DMMA A-fragment order, so a slab is ONE
//    contiguous 16 KB TMA bulk copy (cp.async.bulk + mbarrier complete_tx) and every fragment load is a
//    conflict-free LDS.128.  Symmetry: w'Aw = sum_mb w_mb'( A[mb,mb] w_mb + 2 sum_{kb<mb} A[mb,kb] w_kb )
//    -- only the lower block-triangle is streamed and multiplied (~half the flops of a full GEMM; inside a
//    diagonal block the same identity is applied once more at the 64-row granularity of a warp);
//    the packed entries are (A_kl + A_lk)/2 so the doubled term equals the reference's full sum.
//  * W never exists as a matrix, not even in shared memory: every consumer warp builds its DMMA B fragments
//    in registers from the 2-bit codes and a per-pair 4 x 4 product table (see epi_exact_rb_kernel).  The
//    first versions staged W slabs in shared memory, produced cooperatively by all warps and handed over with
//    mbarriers; that all-to-all dependency between the warps of a CTA cost 15-20 % of the DMMA issue rate
//    (75 % of peak issued; profiles/r01_kernels_v1.md) -- the register-operand form issues 94 %.
//  * one persistent CTA per SM; every CTA walks packed A in the same order, so A is served from the 126 MB L2
//    (17 MB packed at n=2000, 105 MB at n=5000), not HBM.
#include "common.cuh"
#include <cstdlib>
#include <mutex>

namespace gb {

constexpr int EX_BK = 16;            // K-slab (columns of A / individuals)
constexpr int EX_BJ = 16;            // second-SNPs per scan tile (first-SNPs per tile = BN / 16)

__host__ __device__ inline long long ex_total_slabs(long long nb, int bm) {
    return (long long)(bm / EX_BK) * nb * (nb + 1) / 2;
}

// ------------------------------------------------------------------------------------------
// pack_pvp (row-block size bm; the kernel uses 128):
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

// ==========================================================================================
// The exact-test kernel (register-operand B): W is never a matrix anywhere, not even in shared memory.
//   * consumer warps (2 x BN/32, warp tile 64 rows x 32 pairs = 8 x 4 DMMA accumulator tiles) build their own
//     DMMA B fragments: lane (g, t) needs w[k = 4*k4 + t][pair = 8*b + g]; it keeps the 16-individual code
//     words of its u / v rows in registers (prefetched one slab ahead through L1), and one LDS.64 into the
//     per-pair 4 x 4 product table prod[pair][c_u*4 + c_v] = lutU[i][c_u] * lutV[j][c_v] yields the element.
//     No W slab, no DMUL in the K loop, and above all no warp ever waits for another warp's production.
//   * the only shared stream is packed A: one lane of an extra (producer) warp keeps an NS-deep ring of 16 KB slabs full
//     with TMA bulk copies; full[s] completes on the TMA byte count alone, empty[s] on one arrival per
//     consumer warp, so consumers may drift up to NS slabs apart before anything blocks.
//   * consumers synchronise among themselves only at tile boundaries (named barrier 1).
// ==========================================================================================
constexpr int RB_BM = 128;
constexpr int RB_PS = 17;                          // padded row of the product table (16 entries + 1)

template <int BN, int NS>
struct ExactRbSmem {
    double A[NS][RB_BM * EX_BK];                   // TMA destination, DMMA A-fragment order
    double prod[BN][RB_PS];
    double red_var[2][BN];
    double red_eff[2][BN];
    long long uoff[BN];
    long long voff[BN];
    int pi[BN];
    int pj[BN];
    uint64_t full[NS], empty[NS];
};

// epilogue of one row block with the product table: var += sum_r w_r y_r, eff += sum_r w_r py_r.
// Rows past n need no masking: packed A has zero rows/columns there (y_r == 0) and py_r is loaded as 0.
template <int MT, int NTL>
__device__ __forceinline__ void exact_row_epilogue_prod(const double (&acc)[MT][NTL][2], const unsigned char *bed,
                                                        long long row_bytes, long long n, long long r0, int pair0,
                                                        int g, int t, const double *__restrict__ py,
                                                        const long long *uoff, const long long *voff,
                                                        const double *prod, double *red_var, double *red_eff) {
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
                const double *pt = prod + pair * RB_PS;
                double v = 0.0, e = 0.0;
#pragma unroll
                for (int a = 0; a < MT; ++a) {
                    const int k = (a & 3) * 8 + g;   // individual inside the 32-row word
                    const int c_u = (int)(cwu[b][cc][a >> 2] >> (2 * k)) & 3;
                    const int c_v = (int)(cwv[b][cc][a >> 2] >> (2 * k)) & 3;
                    const double w = pt[c_u * 4 + c_v];
                    v += w * acc[a][b0 + b][cc];
                    e += w * pyv[a];
                }
#pragma unroll
                for (int o = 4; o < 32; o <<= 1) { v += shfl_xor_f64(v, o); e += shfl_xor_f64(e, o); }
                if (g == 0) { red_var[pair] += v; red_eff[pair] += e; }
            }
    }
}

// One K slab (16 individuals) of DMMAs for the first MTV of the warp's 8 m-tiles.  MTV < 8 is used in the last row
// block, where the rows past n are padding: with n = 5000 only 8 of its 128 rows exist, and that block has the
// longest K loop of all (it is 5 % of the kernel).
template <int MTV, int NTL>
__device__ __forceinline__ void exact_slab_mma(double (&acc)[8][NTL][2], const double2 *Af, int frag0, int lane,
                                               const char *const (&pb)[NTL], const unsigned (&z)[NTL]) {
    constexpr int SPB = RB_BM / EX_BK;
#pragma unroll
    for (int k4 = 0; k4 < EX_BK / 4; ++k4) {
        double af[MTV], bf[NTL];
#pragma unroll
        for (int a2 = 0; a2 < MTV / 2; ++a2) {
            const double2 v = Af[(k4 * SPB + frag0 + a2) * 32 + lane];
            af[2 * a2] = v.x; af[2 * a2 + 1] = v.y;
        }
#pragma unroll
        for (int b = 0; b < NTL; ++b)
            bf[b] = *reinterpret_cast<const double *>(pb[b] + ((z[b] >> (8 * k4)) & 0xffu));
#pragma unroll
        for (int a = 0; a < MTV; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
}

template <int BN, int NS, int CTAS_PER_SM, bool PAIR_MODE>
// Register budget: the file is 16 K registers per SM sub-partition, so what matters is warps per sub-partition.
// Four producer warps (one per sub-partition, three of them idle) keep the CTA symmetric: the kernel starts at
// 65536 / threads registers per thread, the producers hand theirs back (setmaxnreg.dec) and every consumer
// warp grows to 232 (setmaxnreg.inc) -- the 8 x 4 accumulator tile alone is 128 registers.
__global__ void __launch_bounds__((2 * (BN / 32) + 4) * 32, CTAS_PER_SM)
epi_exact_rb_kernel(const ExactParams p) {
    constexpr int BM = RB_BM, WARPS_N = BN / 32, NCW = 2 * WARPS_N, NCT = NCW * 32;
    constexpr int MT = 8, NTL = 4;                   // warp tile 64 rows x 32 pairs
    constexpr int SPB = BM / EX_BK;
    constexpr int BI = BN / EX_BJ;
    constexpr int NUV = PAIR_MODE ? 4 : 2;           // distinct u (and v) genotype rows among a lane's 4 pairs
    constexpr unsigned SLAB_BYTES = BM * EX_BK * sizeof(double);
    static_assert(BN % 32 == 0 && BN <= NCT, "tile setup uses one consumer thread per pair");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ExactRbSmem<BN, NS> &S = *reinterpret_cast<ExactRbSmem<BN, NS> *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const long long ntiles = PAIR_MODE ? (p.n_pairs + BN - 1) / BN : p.prefix[p.n_iblk];
    const long long my_tiles =
        ((long long)blockIdx.x < ntiles) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long s_total = ex_total_slabs(p.nb, BM);

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&S.full[s], 1);                // the expect_tx arrival; completes when the bytes landed
            mbar_init(&S.empty[s], NCW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp >= NCW) {
        // =============================== A-stream producer ==============================================
        if (BN == 128) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 24;\n");
        if (warp == NCW && lane == 0) {
            int stage = 0;
            unsigned use_odd = 0;                    // parity of (ring pass index)
            bool first_pass = true;
            for (long long tl = 0; tl < my_tiles; ++tl) {
                for (long long s = 0; s < s_total; ++s) {
                    if (!first_pass) mbar_wait(&S.empty[stage], use_odd ^ 1u);
                    fence_proxy_async();             // consumers' LDS of this stage precede the async refill
                    mbar_expect_tx(&S.full[stage], SLAB_BYTES);
                    tma_bulk_g2s(S.A[stage], p.packed + s * (BM * EX_BK), SLAB_BYTES, &S.full[stage]);
                    if (++stage == NS) { stage = 0; use_odd ^= 1u; first_pass = false; }
                }
            }
        }
        return;
    }

    // =================================== consumer warps =================================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n");
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / WARPS_N, wn = warp % WARPS_N;
    int stage = 0;
    unsigned fphase = 0;
    for (long long tl = 0; tl < my_tiles; ++tl) {
        const long long tile = blockIdx.x + tl * gridDim.x;
        named_bar_sync(1, NCT);                      // previous tile's results consumed
        // ---------------- tile setup: which BN pairs, their rows and product tables -------------------
        if (tid < BN) {
            long long si = -1, sj = -1;
            if (PAIR_MODE) {
                const long long idx = tile * BN + tid;
                if (idx < p.n_pairs) { si = p.pairs[2 * idx]; sj = p.pairs[2 * idx + 1]; }
            } else {
                long long lo = 0, hi = p.n_iblk;     // largest ib with prefix[ib] <= tile (uniform)
                while (hi - lo > 1) {
                    const long long mid = (lo + hi) >> 1;
                    if (p.prefix[mid] <= tile) lo = mid; else hi = mid;
                }
                const long long ib = lo;
                const long long li = ib * BI + tid / EX_BJ;
                si = (li < p.n_i) ? p.ilist[li] : -1;
                sj = p.jstart[ib] + (tile - p.prefix[ib]) * EX_BJ + (tid % EX_BJ);
            }
            const bool iok = (si >= 0) && (si < p.m), jok = (sj >= 0) && (sj < p.m);
            const bool valid = iok && jok && (PAIR_MODE || p.pair_space != GMAT_B200_PAIRS_UPPER || sj > si);
            S.pi[tid] = valid ? (int)si : -1;
            S.pj[tid] = valid ? (int)sj : -1;
            // row offsets follow the SNPs, not the pair mask: in a scan tile the lanes share u / v rows across
            // pairs, and a masked pair (j <= i) sits next to valid ones that use the same rows
            S.uoff[tid] = iok ? si * p.row_bytes : 0;
            S.voff[tid] = jok ? sj * p.row_bytes : 0;
            double lu[4], lv[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                lu[a] = valid ? __ldg(p.lutU + si * 4 + a) : 0.0;         // masked pair: w == 0
                lv[a] = valid ? __ldg(p.lutV + sj * 4 + a) : 0.0;
            }
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) S.prod[tid][a * 4 + b] = lu[a] * lv[b];
            S.red_var[0][tid] = 0.0; S.red_var[1][tid] = 0.0;
            S.red_eff[0][tid] = 0.0; S.red_eff[1][tid] = 0.0;
        }
        named_bar_sync(1, NCT);

        // ---------------- this lane's B-fragment sources ---------------------------------------------------
        // pair of n-tile b:  wn*32 + 8*b + g.  Scan tiles are (BN/16 first-SNPs) x (16 second-SNPs), so the four
        // pairs of a lane share two u rows (b >> 1) and two v rows (b & 1); pair lists have four of each.
        const char *pb[NTL];
#pragma unroll
        for (int b = 0; b < NTL; ++b) pb[b] = reinterpret_cast<const char *>(&S.prod[wn * 32 + b * 8 + g][0]);
        const unsigned char *ur[NUV], *vr[NUV];
#pragma unroll
        for (int x = 0; x < NUV; ++x) {
            ur[x] = p.bed + S.uoff[wn * 32 + (PAIR_MODE ? x * 8 : x * 16) + g];
            vr[x] = p.bed + S.voff[wn * 32 + x * 8 + g];
        }
        unsigned nu[NUV], nv[NUV];                   // code words (16 individuals) of the next slab
        auto load_codes = [&](long long ks) {
            const long long off = ks * (EX_BK / 4);
            const bool ok = off + 4 <= p.row_bytes;
#pragma unroll
            for (int x = 0; x < NUV; ++x) {
                nu[x] = ok ? __ldg(reinterpret_cast<const unsigned *>(ur[x] + off)) : 0u;
                nv[x] = ok ? __ldg(reinterpret_cast<const unsigned *>(vr[x] + off)) : 0u;
            }
        };
        load_codes(0);

        // ---------------- K loop over the lower block-triangle of A -------------------------------------
        for (long long mb = 0; mb < p.nb; ++mb) {
            double acc[MT][NTL][2];
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
            const long long nks = (mb + 1) * SPB;
            const long long diag0 = mb * SPB;
            // Symmetry, at the granularity of this warp's 64 rows: everything left of the warp's own 64 x 64
            // diagonal sub-block counts twice, the sub-block itself once, and what lies right of it (the upper
            // right quarter of the 128 x 128 diagonal block, warp row 0 only) is not multiplied at all.
            const long long dbl_at = diag0 + (wm ? SPB / 2 : 0);
            long long mma_end = wm ? nks : diag0 + SPB / 2;
            // padding: columns past n are zero in packed A (skip their slabs), and of this warp's 8 m-tiles only
            // those with a row below n need multiplying (pairs of m-tiles: fragments are loaded two at a time)
            const long long kslabs = (p.n + EX_BK - 1) / EX_BK;
            if (mma_end > kslabs) mma_end = kslabs;
            const long long rows_left = p.n - (mb * BM + wm * 64);
            const int mtv = rows_left >= 64 ? 8 : (rows_left <= 0 ? 0 : (int)((rows_left + 15) / 16) * 2);
            for (long long ks = 0; ks < nks; ++ks) {
                if (ks == dbl_at && ks > 0) {
#pragma unroll
                    for (int a = 0; a < MT; ++a)
#pragma unroll
                        for (int b = 0; b < NTL; ++b) { acc[a][b][0] *= 2.0; acc[a][b][1] *= 2.0; }
                }
                // byte k4 of z[b] = 8 * (c_u*4 + c_v) of individual 16*ks + 4*k4 + t: a byte offset into prod[pair]
                unsigned z[NTL];
                {
                    unsigned zu[NUV], zv[NUV];
#pragma unroll
                    for (int x = 0; x < NUV; ++x) {
                        zu[x] = ((nu[x] >> (2 * t)) & 0x03030303u) << 5;
                        zv[x] = ((nv[x] >> (2 * t)) & 0x03030303u) << 3;
                    }
#pragma unroll
                    for (int b = 0; b < NTL; ++b) z[b] = zu[PAIR_MODE ? b : (b >> 1)] | zv[PAIR_MODE ? b : (b & 1)];
                }
                load_codes((ks + 1 < nks) ? ks + 1 : 0);                 // next slab's codes travel under the MMAs
                mbar_wait(&S.full[stage], fphase);
                const double2 *Af = reinterpret_cast<const double2 *>(S.A[stage]);
                if (ks < mma_end) {
                    const int frag0 = wm * (MT / 2);
                    if (mtv == 8) exact_slab_mma<8, NTL>(acc, Af, frag0, lane, pb, z);
                    else if (mtv == 6) exact_slab_mma<6, NTL>(acc, Af, frag0, lane, pb, z);
                    else if (mtv == 4) exact_slab_mma<4, NTL>(acc, Af, frag0, lane, pb, z);
                    else if (mtv == 2) exact_slab_mma<2, NTL>(acc, Af, frag0, lane, pb, z);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&S.empty[stage]);
                if (++stage == NS) { stage = 0; fphase ^= 1u; }
            }
            exact_row_epilogue_prod<MT, NTL>(acc, p.bed, p.row_bytes, p.n, mb * BM + wm * 64, wn * 32, g, t, p.py,
                                             S.uoff, S.voff, &S.prod[0][0], S.red_var[wm], S.red_eff[wm]);
        }

        // ---------------- tile result: chi-square, threshold, compaction ----------------------------
        named_bar_sync(1, NCT);
        if (tid < BN) {
            const double var = S.red_var[0][tid] + S.red_var[1][tid];
            const double eff = S.red_eff[0][tid] + S.red_eff[1][tid];
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

// ------------------------------------------------------------------------------------------
// configurations (GMAT_B200_EXACT_CFG selects one at library load; the packed layout depends on BM)
// ------------------------------------------------------------------------------------------
struct ExactCfg {
    int bm, bn, threads, ctas_per_sm;
    size_t smem;
    void (*scan)(const ExactParams);
    void (*pairs)(const ExactParams);
};

template <int BN, int NS, int CPS>
static ExactCfg make_rb_cfg() {
    ExactCfg c;
    c.bm = RB_BM; c.bn = BN; c.threads = (2 * (BN / 32) + 4) * 32; c.ctas_per_sm = CPS;
    c.smem = sizeof(ExactRbSmem<BN, NS>);
    c.scan = epi_exact_rb_kernel<BN, NS, CPS, false>;
    c.pairs = epi_exact_rb_kernel<BN, NS, CPS, true>;
    return c;
}

static const ExactCfg &exact_cfg() {
    static const ExactCfg cfgs[2] = {
        make_rb_cfg<128, 8, 1>(),                    // 0: 8 + 4 warps, 128 x 128 tiles, 1 CTA/SM (default)
        make_rb_cfg<64, 5, 2>(),                     // 1: 4 + 4 warps, 128 x 64 tiles, 2 CTAs/SM (same speed, 2x L2 reads)
    };
    static const int sel = [] {
        const char *e = getenv("GMAT_B200_EXACT_CFG");
        const int v = e ? atoi(e) : 0;
        return (v == 1) ? 1 : 0;
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

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of gmat_b200_epi_exact_scan.  No counterpart in the reference: added
// for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  The tile table of the scan (prefix[n_iblk + 1] and jstart[n_iblk], int64) -- the only scratch the
// scan needs -- lives in a file-scope device array instead of a cudaMallocAsync block; every GPU that loads the
// library holds its GMAT_B200_NOALLOC_MAX_FIRST-sized copy (1 MiB + 8 bytes) until exit.
//
// In host code `exact_noalloc_scratch` is only the host-side stand-in of the symbol, NOT a device address: the
// address of each device's copy is taken once per device with cudaGetSymbolAddress, with that device current, and
// that pointer is what the kernels get.
//
// Every call on a device shares the one array, and it keeps the last call's contents.  Nothing in it has to start at
// zero: exact_tiles_kernel writes all 2 n_iblk + 1 entries this call's scan reads before the scan starts (same
// stream), so no clearing is needed.  Two calls whose kernels overlap on one GPU (different streams, or two host
// threads) RACE on the table: serialise them, or use gmat_b200_epi_exact_scan.
// ---------------------------------------------------------------------------------------------------------------
namespace gb {
constexpr long long NOALLOC_MAX_FIRST = GMAT_B200_NOALLOC_MAX_FIRST;
static __device__ __align__(256) long long exact_noalloc_scratch[2 * NOALLOC_MAX_FIRST + 1];

// device address of the current device's copy of exact_noalloc_scratch (nullptr + last_error on failure)
static long long *exact_noalloc_ptr() {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static long long *ptrs[MAX_DEV];
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
        set_error("noalloc scratch: cudaGetDevice failed or device index %d out of range (%s)", dev,
                  cudaGetErrorString(e));
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        long long *ptr = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), exact_noalloc_scratch);
        if (e != cudaSuccess || !ptr) {
            set_error("noalloc scratch: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        ptrs[dev] = ptr;
    }
    return ptrs[dev];
}
}  // namespace gb

extern "C" long long gmat_b200_noalloc_max_first(void) { return NOALLOC_MAX_FIRST; }

extern "C" int gmat_b200_epi_exact_scan_noalloc(const unsigned char *bed_dev, long long n, long long m,
                                                long long row_bytes, const double *lutU_dev,
                                                const double *lutV_dev, const double *py_dev,
                                                const double *pvp_packed_dev, const long long *ilist_dev,
                                                long long n_i, int pair_space, double chi_cut,
                                                gmat_b200_hit *hits_dev, long long cap,
                                                unsigned long long *count_dev, void *stream) {
    int rc = exact_common_check(bed_dev, n, m, row_bytes, lutU_dev, lutV_dev, py_dev, pvp_packed_dev);
    if (rc != GMAT_B200_OK) return rc;
    GB_CHECK_ARG(ilist_dev && count_dev, "NULL pointer");
    GB_CHECK_ARG(n_i >= 0, "n_i");
    if (n_i > NOALLOC_MAX_FIRST) {
        set_error("gmat_b200_epi_exact_scan_noalloc: n_i = %lld first SNPs exceeds the static scratch "
                  "(GMAT_B200_NOALLOC_MAX_FIRST = %lld); split the list or call gmat_b200_epi_exact_scan",
                  n_i, NOALLOC_MAX_FIRST);
        return GMAT_B200_ERR_ARG;
    }
    GB_CHECK_ARG(cap == 0 || hits_dev, "hits_dev is NULL");
    GB_CHECK_ARG(pair_space == GMAT_B200_PAIRS_UPPER || pair_space == GMAT_B200_PAIRS_ALL, "pair_space");
    if (n_i == 0) return GMAT_B200_OK;
    long long *scratch = exact_noalloc_ptr();
    if (!scratch) return GMAT_B200_ERR_CUDA;
    const ExactCfg &cfg = exact_cfg();
    const int bi = cfg.bn / EX_BJ;
    cudaStream_t st = (cudaStream_t)stream;
    ExactParams p{};
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes;
    p.nb = round_up(n, cfg.bm) / cfg.bm;
    p.lutU = lutU_dev; p.lutV = lutV_dev; p.py = py_dev; p.packed = pvp_packed_dev;
    p.ilist = ilist_dev; p.n_i = n_i; p.n_iblk = (n_i + bi - 1) / bi;      // n_iblk <= n_i <= NOALLOC_MAX_FIRST
    p.pair_space = pair_space; p.chi_cut = chi_cut;
    p.hits = hits_dev; p.cap = cap; p.count = count_dev;
    p.prefix = scratch;
    p.jstart = scratch + p.n_iblk + 1;
    exact_tiles_kernel<<<1, 1024, 0, st>>>(ilist_dev, n_i, m, pair_space, bi, p.n_iblk, scratch,
                                           scratch + p.n_iblk + 1);
    GB_LAUNCH_CHECK();
    GB_CUDA(cudaFuncSetAttribute(cfg.scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    const unsigned grid = (unsigned)(sm_count() * cfg.ctas_per_sm);
    cfg.scan<<<grid, cfg.threads, cfg.smem, st>>>(p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
