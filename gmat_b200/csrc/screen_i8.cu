// Approximate screen on the INT8 tensor cores (sm_100a: tcgen05.mma cta_group::2 kind::i8, TMEM, TMA) with an exact
// FP64 re-test of the survivors.
//
// The reference's screen (gmat/remma/_remma_epi_eff_cpu.c:61-86, 226-250) reports the pairs with |eff_ij| > cut,
// eff_ij = sum_k x_ik x_jk py_k.  With x_ik = a_ik - o_i (a the integer dosage / heterozygote indicator, o_i = 2p_i or
// 2p_i q_i = -lut[i][code 0]):
//
//     eff_ij = Q_ij - o_j t_i - o_i (t_j - o_j P),   Q_ij = sum_k a_ik a_jk py_k,  t_i = a_i . py,  P = sum_k py_k
//
// and only Q costs n per pair.  py / scale (scale = max|py| / 0.49) is cut into S = 2 signed base-128 digits d_s
// (|d| <= 64), so that (a_ik d_sk) is an int8 and  Q~ = scale sum_s 128^-(s+1) sum_k (a_ik d_sk) a_jk  is S exact
// int8 x int8 -> int32 GEMMs; |Q - Q~| <= scale 128^-S 2n.  Stage 1 (i8_screen_pair_kernel) evaluates
// e~ = Q~ - o_j t_i - o_i c_j in FP32 for EVERY pair and keeps the pairs with |e~| >= cut - margin, margin = a
// rigorous bound of everything stage 1 drops (digit truncation + FP32 rounding); stage 2 (screen_exact_kernel)
// recomputes eff of those few pairs in FP64 from the LUT values, exactly like the FP64 engine (eff.cu), and applies
// the reference's own threshold.  The reported set is therefore the set of pairs whose FP64 eff passes the cut, and
// the reported values are FP64 values; the tensor cores only decide which pairs are worth the FP64 arithmetic.
// 2 S n int8 ops per pair at ~3 POP/s against 2 n FP64 flop per pair at 35 TFLOP/s.
//
// Tiles: M = 256 second SNPs (two CTAs x 128 TMEM lanes) x N = 128 first SNPs x K = all ids in slabs of 128; the S
// plane accumulators of a tile sit side by side in TMEM (S x 128 columns), double-buffered when they fit twice.
//
// Missing calls.  The reference C screen does not impute: code 01 is the value 1/3 (_remma_epi_eff_cpu.c:42), so
// a in {0, 1/3, 1, 2} (dominance: {0, 1/3, 1, 0}).  Both operands then carry 3a in {0, 1, 3, 6} -- still small
// integers -- and Q = Q'/9.  (3 a_ik)(d_sk) must stay an int8, so py is cut into S = 3 signed base-32 digits
// (|d| <= 16, 15 bits) instead of 2 base-128 digits (14 bits): 1.5 x the MMAs of the clean case, same margin formula
// (2 n scale B^-S), same FP64 re-test.  Real PLINK files almost always have missing calls; before this the screen
// dropped to the 14-42 x slower FP64 DMMA kernel for them.
//
// No host synchronisation: the scale of py, the margin, the FP32 cuts and the "is the int8 path worth it" decision
// are formed ON THE DEVICE (sc_setup_kernel) and read by the kernels from a small scalar block; stage 2 is a
// grid-stride kernel over the device-side candidate count.  The caller reads one status word with the hit count.
#include "tc5.cuh"
#include <cstdlib>
#include <vector>

namespace gb {

constexpr int SC_SMAX = 3;                                        // digit planes of py: 2 (base 128) or 3 (base 32)
constexpr int SC_TI = 128;                                        // first SNPs per tile (UMMA N)
constexpr int SC_NS_MAX = 6;                                      // smem ring stages (6 for S = 2, 5 for S = 3)
constexpr int SC_THREADS = 192;                                   // producer, issuer, 4 epilogue warps
constexpr int SC_B_BYTES = (SC_TI / 2) * I8_SLAB;                 // 8 KB: this CTA's half of one plane tile
template <int S> struct ScreenCfg {
    static constexpr int NS = S == 2 ? 6 : 5;
    static constexpr int STAGE_BYTES = I8_A_BYTES + S * SC_B_BYTES;   // 32 / 40 KB per CTA
    static constexpr int NBUF = 2 * S * SC_TI <= 512 ? 2 : 1;         // accumulator buffers that fit the 512 TMEM columns
};

struct ScreenScalars {                                            // written by sc_setup_kernel, read by every kernel
    float cut;                                                    // eff_cut - margin, rounded down
    float qscale;                                                 // scale * B^-S (/ 9 with tripled codes)
    float base;                                                   // digit base B
    int fallback;                                                 // 1: margin not small against the cut -> nothing runs
    double margin;
};

struct ScreenBars {
    uint64_t full[SC_NS_MAX], empty[SC_NS_MAX], tfull[2], tempty[2];
    uint32_t tmem_base;
    int col_snp[2][SC_TI];                                        // per accumulator buffer: first SNP of a column (-1: none)
    float col_t[2][SC_TI], col_o[2][SC_TI];
    int col_bin[2][SC_TI];
    float table[112];
};

struct ScreenParams {
    int n_i, n_ib, m, n_slab, pair_space;
    int plane_rows;                                               // rows of one plane in the PU buffer (n_ib * 128)
    const int *item_prefix;                                       // [n_ib + 1]
    const int *jb0;                                               // [n_ib]
    const long long *ilist;                                       // [n_i]
    const float *tU, *oU;                                         // [n_i]  a_i . py,  o_i
    const float *oV, *cV;                                         // [m]    o_j,  t_j - o_j P
    const int *binI, *binJ;                                       // [m] each or NULL
    const float *cut_table;                                       // 111 floats (cut - margin, rounded down) or NULL
    const ScreenScalars *scal;
    int2 *cand; long long cand_cap; unsigned long long *cand_count;
};

__device__ __forceinline__ int sc_find(const int *prefix, int n, int item) {
    int lo = 0, hi = n - 1;                                       // largest ib with prefix[ib] <= item
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (prefix[mid] <= item) lo = mid; else hi = mid - 1;
    }
    return lo;
}

template <int S>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(SC_THREADS, 1)
i8_screen_pair_kernel(const __grid_constant__ CUtensorMap map_v, const __grid_constant__ CUtensorMap map_p,
                      const ScreenParams p) {
    constexpr int SC_S = S, SC_NS = ScreenCfg<S>::NS, SC_STAGE_BYTES = ScreenCfg<S>::STAGE_BYTES;
    constexpr int NBUF = ScreenCfg<S>::NBUF;
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    ScreenBars *bars = reinterpret_cast<ScreenBars *>(smem + SC_NS * SC_STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    const int total_items = p.scal->fallback ? 0 : p.item_prefix[p.n_ib];

    if (threadIdx.x == 0) {
        for (int s = 0; s < SC_NS; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&bars->tfull[b], 1); mbar_init(&bars->tempty[b], 8); }
        mbar_fence_init();
    }
    if (p.cut_table != nullptr)
        for (int q = threadIdx.x; q < 111; q += SC_THREADS) bars->table[q] = p.cut_table[q];
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                         smem_u32(&bars->tmem_base)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    // producer and issuer warps run converged; one elected lane issues the single-thread instructions (tc5.cuh)
    if (warp == 0) {
        // ================================ TMA producer (both CTAs) ===============================
        int stage = 0; unsigned phase = 0;
        for (int item = cluster_id; item < total_items; item += n_clusters) {
            const int ib = sc_find(p.item_prefix, p.n_ib, item);
            const int jb = p.jb0[ib] + (item - p.item_prefix[ib]);
            for (int l = 0; l < p.n_slab; ++l) {
                mbar_wait(&bars->empty[stage], phase ^ 1u);
                if (elect_one()) {
                    if (rank == 0) mbar_expect_tx(&bars->full[stage], 2 * SC_STAGE_BYTES);
                    unsigned char *sa = smem + stage * SC_STAGE_BYTES;
                    tma_load_2d_pair(sa, &map_v, l * I8_SLAB, (2 * jb + rank) * I8_TJ, &bars->full[stage]);
#pragma unroll
                    for (int s = 0; s < SC_S; ++s)
                        tma_load_2d_pair(sa + I8_A_BYTES + s * SC_B_BYTES, &map_p, l * I8_SLAB,
                                         s * p.plane_rows + ib * SC_TI + rank * (SC_TI / 2), &bars->full[stage]);
                }
                if (++stage == SC_NS) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader CTA) ================================
        if (rank == 0) {
            int stage = 0; unsigned phase = 0;
            int abuf = 0; unsigned aphase = 0;
            const uint32_t idesc = i8p_idesc(SC_TI);
            for (int item = cluster_id; item < total_items; item += n_clusters) {
                mbar_wait(&bars->tempty[abuf], aphase ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(abuf * SC_S * SC_TI);
                for (int l = 0; l < p.n_slab; ++l) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + (uint32_t)(stage * SC_STAGE_BYTES);
                    const uint64_t da = umma_desc_sw128(sa);
                    if (elect_one()) {
#pragma unroll
                        for (int s = 0; s < SC_S; ++s) {
                            const uint64_t db = umma_desc_sw128(sa + I8_A_BYTES + s * SC_B_BYTES);
#pragma unroll
                            for (int k = 0; k < I8_SLAB / 32; ++k)
                                umma_i8_pair(tmem_d + (uint32_t)(s * SC_TI), da + (uint64_t)(2 * k),
                                             db + (uint64_t)(2 * k), idesc, (uint32_t)((l | k) != 0));
                        }
                        tc_commit_pair(&bars->empty[stage]);
                        if (l == p.n_slab - 1) tc_commit_pair(&bars->tfull[abuf]);
                    }
                    if (++stage == SC_NS) { stage = 0; phase ^= 1u; }
                }
                if (++abuf == NBUF) { abuf = 0; aphase ^= 1u; }
            }
        }
    } else {
        // ================================ epilogue (both CTAs): TMEM lane = second SNP =============
        const int q = warp & 3;
        const int e = (warp - 2) * 32 + lane;                     // the column whose constants this thread stages
        const float qscale = p.scal->qscale, cut0 = p.scal->cut, base = p.scal->base;
        int abuf = 0; unsigned aphase = 0;
        for (int item = cluster_id; item < total_items; item += n_clusters) {
            const int ib = sc_find(p.item_prefix, p.n_ib, item);
            const int jb = p.jb0[ib] + (item - p.item_prefix[ib]);
            if (NBUF == 1) named_bar_sync(2, 128);                // single buffer: nobody still reads the last item's constants
            {
                const int ii = ib * SC_TI + e;
                const bool has = ii < p.n_i;
                const int snp = has ? (int)p.ilist[ii] : -1;
                bars->col_snp[abuf][e] = snp;
                bars->col_t[abuf][e] = has ? p.tU[ii] : 0.f;
                bars->col_o[abuf][e] = has ? p.oU[ii] : 0.f;
                bars->col_bin[abuf][e] = (has && p.binI != nullptr) ? p.binI[snp] * 10 : 0;
            }
            const int j = (2 * jb + rank) * I8_TJ + q * 32 + lane;
            const bool jok = j < p.m;
            const float oj = jok ? p.oV[j] : 0.f, cj = jok ? p.cV[j] : 0.f;
            const int bj = (jok && p.binJ != nullptr) ? p.binJ[j] : 0;
            named_bar_sync(1, 128);                               // the four epilogue warps: constants are staged
            mbar_wait_warp(&bars->tfull[abuf], aphase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(abuf * SC_S * SC_TI);
#pragma unroll 1
            for (int ch = 0; ch < SC_TI / 32; ++ch) {
                int v[SC_S][32];
#pragma unroll
                for (int sp = 0; sp < SC_S; ++sp) tmem_ld32(taddr + sp * SC_TI + ch * 32, v[sp]);
                tmem_ld_wait();
                unsigned mask = 0;
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const int col = ch * 32 + c;
                    float t32 = (float)v[0][c];                   // Horner over the planes, most significant first
#pragma unroll
                    for (int sp = 1; sp < SC_S; ++sp) t32 = fmaf(t32, base, (float)v[sp][c]);
                    const float r = fmaf(-oj, bars->col_t[abuf][col], -bars->col_o[abuf][col] * cj);
                    const float eff = fmaf(t32, qscale, r);
                    const float cut = p.cut_table != nullptr ? bars->table[bars->col_bin[abuf][col] + bj] : cut0;
                    const int snp = bars->col_snp[abuf][col];
                    const bool ok = jok && snp >= 0 && (p.pair_space == GMAT_B200_PAIRS_ALL || j > snp);
                    if (ok && fabsf(eff) >= cut) mask |= 1u << c;
                }
                if (__any_sync(0xffffffffu, mask != 0)) {         // rare: a handful of candidates per launch
                    while (mask) {
                        const int c = __ffs(mask) - 1;
                        mask &= mask - 1;
                        const unsigned long long pos = atomicAdd(p.cand_count, 1ULL);
                        if ((long long)pos < p.cand_cap)
                            p.cand[pos] = make_int2(bars->col_snp[abuf][ch * 32 + c], j);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&bars->tempty[abuf]);
            if (++abuf == NBUF) { abuf = 0; aphase ^= 1u; }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// ------------------------------------------------------------------------------------------ operand builders
// out[0] = max |py|, out[1] = sum py (one block, fixed order)
__global__ void __launch_bounds__(1024) sc_py_stats_kernel(const double *__restrict__ py, int n, double *__restrict__ out) {
    __shared__ double smax[32], ssum[32];
    double mx = 0.0, sm = 0.0;
    for (int k = threadIdx.x; k < n; k += 1024) { const double v = py[k]; mx = fmax(mx, fabs(v)); sm += v; }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, d));
        sm += __shfl_xor_sync(0xffffffffu, sm, d);
    }
    if ((threadIdx.x & 31) == 0) { smax[threadIdx.x >> 5] = mx; ssum[threadIdx.x >> 5] = sm; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < 32; ++w) { a = fmax(a, smax[w]); b += ssum[w]; }
        out[0] = a; out[1] = b;
    }
}
// balanced base-2^bits digits of py[k] / scale, scale = max|py| / 0.49 read from stats[0] on the device:
// dig[s][k] in [-B/2, B/2 - 1] (most significant plane first), zero for k >= n
__global__ void sc_digits_kernel(const double *__restrict__ py, int n, int ld, const double *__restrict__ stats,
                                 int n_planes, int bits, signed char *__restrict__ dig) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ld) return;
    const double maxabs = stats[0];
    const double inv_scale = maxabs > 0.0 ? 0.49 / maxabs : 0.0;
    const long long B = 1LL << bits, half = B >> 1;
    long long t = k < n ? __double2ll_rn(py[k] * inv_scale * (double)(1LL << (bits * n_planes))) : 0;
    for (int s = n_planes - 1; s >= 1; --s) {
        const long long d = ((t + half) & (B - 1)) - half;
        t = (t - d) >> bits;
        dig[(size_t)s * ld + k] = (signed char)d;
    }
    dig[k] = (signed char)t;
}
// Everything the kernels need that depends on the cut: scale, margin, FP32 cut(s), and whether the int8 path is worth
// it at all.  One block.  status[0] = 1 when it is not (the caller runs the FP64 kernel instead).
__global__ void sc_setup_kernel(const double *__restrict__ stats, int n, double eff_cut, const double *__restrict__ table,
                                int n_planes, int bits, int tripled, ScreenScalars *__restrict__ scal,
                                float *__restrict__ tab32, int *__restrict__ status) {
    __shared__ double smin[128];
    const double maxabs = stats[0];
    const double scale = maxabs / 0.49;
    const double bpow = 1.0 / (double)(1LL << (bits * n_planes));                  // B^-S
    // everything stage 1 drops: digit truncation scale B^-S 2n, FP32 rounding of Q~ and of the rank-one terms
    const double margin = scale * (double)n * (2.0 * bpow + 4.0e-6);
    double mn = 1e300;
    if (table != nullptr) {
        if (threadIdx.x < 111) { mn = table[threadIdx.x]; tab32[threadIdx.x] = __double2float_rd(table[threadIdx.x] - margin); }
        if (mn != mn) mn = -1.0;                                                   // a NaN cut: never the int8 path
    } else {
        mn = eff_cut == eff_cut ? eff_cut : -1.0;
    }
    smin[threadIdx.x] = mn;
    __syncthreads();
    for (int d = 64; d >= 1; d >>= 1) {
        if (threadIdx.x < d) smin[threadIdx.x] = fmin(smin[threadIdx.x], smin[threadIdx.x + d]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double cut_min = smin[0];
        const bool ok = maxabs > 0.0 && maxabs < 1e300 && margin < 0.25 * cut_min;  // a cut near zero keeps every pair
        scal->cut = __double2float_rd(eff_cut - margin);
        scal->qscale = (float)(scale * bpow * (tripled ? 1.0 / 9.0 : 1.0));
        scal->base = (float)(1 << bits);
        scal->fallback = ok ? 0 : 1;
        scal->margin = margin;
        if (!ok) status[0] = 1;
    }
}
// t[snp] = sum_k a_k py_k over the int8 codes of one SNP row (one warp per SNP, fixed order)
__global__ void __launch_bounds__(256) sc_dot_kernel(const signed char *__restrict__ g8, int ld, int n, long long m,
                                                     const double *__restrict__ py, double mult,
                                                     double *__restrict__ t) {
    const long long snp = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (snp >= m) return;
    const int lane = threadIdx.x & 31;
    const unsigned *row = reinterpret_cast<const unsigned *>(g8 + (size_t)snp * ld);
    double acc = 0.0;
    for (int w = lane; w * 4 < n; w += 32) {
        const unsigned x = __ldg(row + w);
        const int k = w * 4;
        acc += (double)(int)(x & 0xffu) * py[k];
        if (k + 1 < n) acc += (double)(int)((x >> 8) & 0xffu) * py[k + 1];
        if (k + 2 < n) acc += (double)(int)((x >> 16) & 0xffu) * py[k + 2];
        if (k + 3 < n) acc += (double)(int)(x >> 24) * py[k + 3];
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if (lane == 0) t[snp] = acc * mult;
}
// per second SNP: o_j = -lutV[j][0], c_j = t_j - o_j P  (FP32 copies for stage 1)
__global__ void sc_second_kernel(const double *__restrict__ lutV, const double *__restrict__ tV,
                                 const double *__restrict__ stats, long long m, float *__restrict__ oV,
                                 float *__restrict__ cV) {
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m) return;
    const double P = stats[1];
    const double o = -lutV[j * 4];
    oV[j] = (float)o;
    cV[j] = (float)(tV[j] - o * P);
}
// planes of the first SNPs of the launch: PU[s][row][k] = a_ik * dig[s][k] (row = position in ilist, zero past n_i)
// plus the FP32 per-first-SNP constants
__global__ void __launch_bounds__(256) sc_planes_kernel(const signed char *__restrict__ g8U, int ld,
                                                        const long long *__restrict__ ilist, int n_i, int plane_rows,
                                                        int n_planes, const signed char *__restrict__ dig,
                                                        const double *__restrict__ lutU, const double *__restrict__ tUall,
                                                        signed char *__restrict__ pu, float *__restrict__ tU,
                                                        float *__restrict__ oU) {
    const int row = blockIdx.x;
    const long long snp = row < n_i ? ilist[row] : -1;
    if (threadIdx.x == 0 && row < n_i) { tU[row] = (float)tUall[snp]; oU[row] = (float)(-lutU[snp * 4]); }
    for (int k0 = threadIdx.x * 16; k0 < ld; k0 += 256 * 16) {
        uint4 a = make_uint4(0u, 0u, 0u, 0u);
        if (snp >= 0) a = *reinterpret_cast<const uint4 *>(g8U + (size_t)snp * ld + k0);
        const unsigned aw[4] = {a.x, a.y, a.z, a.w};
        for (int s = 0; s < n_planes; ++s) {
            const uint4 d = *reinterpret_cast<const uint4 *>(dig + (size_t)s * ld + k0);
            const unsigned dw[4] = {d.x, d.y, d.z, d.w};
            unsigned o[4];
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                unsigned r = 0;
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int av = (int)((aw[w] >> (8 * b)) & 0xffu);
                    const int dv = (int)(signed char)((dw[w] >> (8 * b)) & 0xffu);
                    r |= ((unsigned)(av * dv) & 0xffu) << (8 * b);
                }
                o[w] = r;
            }
            *reinterpret_cast<uint4 *>(pu + ((size_t)s * plane_rows + row) * ld + k0) = make_uint4(o[0], o[1], o[2], o[3]);
        }
    }
}
// work list: I-blocks of 128 list entries x J-blocks of 256 second SNPs; UPPER: J-blocks from (min first SNP + 1) / 256
__global__ void sc_tiles_kernel(const long long *__restrict__ ilist, int n_i, int n_ib, int n_jb, int pair_space,
                                int *__restrict__ prefix, int *__restrict__ jb0) {
    for (int ib = threadIdx.x; ib < n_ib; ib += blockDim.x) {
        long long lo = 1LL << 62;
        for (int q = 0; q < SC_TI; ++q) {
            const int li = ib * SC_TI + q;
            if (li < n_i) { const long long s = ilist[li]; if (s < lo) lo = s; }
        }
        long long j0 = pair_space == GMAT_B200_PAIRS_ALL ? 0 : (lo + 1) / (2 * I8_TJ);
        if (j0 > n_jb) j0 = n_jb;
        jb0[ib] = (int)j0;
        prefix[ib + 1] = n_jb - (int)j0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        prefix[0] = 0;
        for (int ib = 0; ib < n_ib; ++ib) prefix[ib + 1] += prefix[ib];
    }
}
// Stage 2: eff of the candidates in FP64 from the LUT values (the arithmetic of the FP64 engine and of the
// reference: lutU[i][c_ik] * lutV[j][c_jk] * py[k]), the reference's threshold, the hit record.  One warp per
// candidate, lanes stride over the bytes of the two .bed rows, fixed reduction order.
__global__ void __launch_bounds__(256) screen_exact_kernel(
    const unsigned char *__restrict__ bed, int n, long long row_bytes, const double *__restrict__ lutU,
    const double *__restrict__ lutV, const double *__restrict__ py, const int2 *__restrict__ cand,
    const unsigned long long *__restrict__ n_cand, int flags, double eff_cut, const double *__restrict__ cut_table,
    const int *__restrict__ binI, const int *__restrict__ binJ, gmat_b200_hit *__restrict__ hits, long long cap,
    unsigned long long *__restrict__ count, long long cand_cap, int *__restrict__ status) {
    __shared__ double prod[8][16];                                // per warp: lutU[i][cu] * lutV[j][cv]
    __shared__ int s_flag[8], s_off[8];
    __shared__ unsigned long long s_base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long total = (long long)*n_cand;
    if (total > cand_cap) {                                       // more survivors than the buffer holds: the caller must
        if (blockIdx.x == 0 && threadIdx.x == 0) status[0] = 2;   // fall back (or retry with a larger buffer)
        return;
    }
    const bool ge = (flags & GMAT_B200_EFF_GE) != 0, swap = (flags & GMAT_B200_EFF_SWAP) != 0;
    // every warp of the block runs the same number of iterations (the hit records of an iteration are appended with
    // ONE atomic per block: a per-hit atomic on the single counter serialised 1e5 hits into ~0.4 ms)
    for (long long base = (long long)blockIdx.x * 8; base < total; base += (long long)gridDim.x * 8) {
        const long long c = base + warp;
        const bool active = c < total;
        int i = 0, j = 0;
        double acc = 0.0;
        bool hit = false;
        if (active) {
            i = cand[c].x; j = cand[c].y;
            if (lane < 16) prod[warp][lane] = lutU[(size_t)i * 4 + (lane >> 2)] * lutV[(size_t)j * 4 + (lane & 3)];
            __syncwarp();
            const unsigned *ri = reinterpret_cast<const unsigned *>(bed + (size_t)i * row_bytes);
            const unsigned *rj = reinterpret_cast<const unsigned *>(bed + (size_t)j * row_bytes);
            // lane <-> individual: py is read in full lines, the two code words of a step are warp broadcasts
            for (int k = lane; k < n; k += 32) {
                const unsigned ci = __ldg(ri + (k >> 4)), cj = __ldg(rj + (k >> 4));
                const int sh = 2 * (k & 15);
                acc += prod[warp][(((ci >> sh) & 3u) << 2) | ((cj >> sh) & 3u)] * py[k];
            }
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
            const double cut = cut_table != nullptr ? cut_table[binI[i] * 10 + binJ[j]] : eff_cut;
            const double a = fabs(acc);
            hit = ge ? a >= cut : a > cut;
        }
        if (lane == 0) s_flag[warp] = hit ? 1 : 0;
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) { s_off[w] = tot; tot += s_flag[w]; }
            s_base = tot ? atomicAdd(count, (unsigned long long)tot) : 0ULL;
        }
        __syncthreads();
        if (lane == 0 && hit) {
            const unsigned long long pos = s_base + (unsigned long long)s_off[warp];
            if ((long long)pos < cap) {
                gmat_b200_hit h;
                h.snp_0 = swap ? j : i; h.snp_1 = swap ? i : j; h.eff = acc; h.var = 0.0; h.chi = 0.0;
                hits[pos] = h;
            }
        }
    }
}


// Stage 2, word-per-lane form (n <= SC_ROWS_MAX_N): the lane-per-individual loop above walks a .bed row 64 bytes per
// step and waits for memory at every line it enters -- 13.7 us per candidate and warp at n = 2000, 0.37 ms for the 1.3e5
// candidates of the C3 screen, as long as the whole prefilter.  Here every lane loads ITS words of both rows up front
// (32 lanes x 4 bytes = one line per instruction, up to eight loads in flight) and then works from registers: word w
// holds individuals 16w .. 16w+15, whose Z'Py values sit transposed in shared memory (pyT[x][w] = py[16w + x], zero
// beyond n: staged once per block), so a warp reads 32 consecutive doubles per step.  Same products, same threshold,
// same hit records; the sum runs per lane over its words, then the same shuffle tree.
constexpr int SC_ROWS_MAX_N = 20480;                              // 16 x (n/16 rounded up to 32) doubles of shared memory
// append a warp's buffered hit records to the global list: one atomic, the records copied by the lanes
__device__ __forceinline__ void sc_flush_hits(const gmat_b200_hit *buf, int nbuf, int lane, gmat_b200_hit *hits,
                                              long long cap, unsigned long long *count) {
    __syncwarp();
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(count, (unsigned long long)nbuf);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (lane < nbuf && (long long)(base + lane) < cap) hits[base + lane] = buf[lane];
    __syncwarp();
}
__global__ void __launch_bounds__(256) screen_exact_rows_kernel(
    const unsigned char *__restrict__ bed, int n, long long row_bytes, const double *__restrict__ lutU,
    const double *__restrict__ lutV, const double *__restrict__ py, const int2 *__restrict__ cand,
    const unsigned long long *__restrict__ n_cand, int flags, double eff_cut, const double *__restrict__ cut_table,
    const int *__restrict__ binI, const int *__restrict__ binJ, gmat_b200_hit *__restrict__ hits, long long cap,
    unsigned long long *__restrict__ count, long long cand_cap, int *__restrict__ status) {
    extern __shared__ double pyT[];                               // [16][wpad]
    __shared__ double prod[8][16];                                // per warp: lutU[i][cu] * lutV[j][cv]
    __shared__ gmat_b200_hit hitbuf[8][32];                       // per warp: hit records waiting for their append
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long total = (long long)*n_cand;
    if (total > cand_cap) {
        if (blockIdx.x == 0 && threadIdx.x == 0) status[0] = 2;
        return;
    }
    if ((long long)blockIdx.x * 8 >= total) return;               // nothing for this block: skip the staging too
    const int nwords = (n + 15) >> 4, wpad = (nwords + 31) & ~31;
    for (int idx = threadIdx.x; idx < 16 * wpad; idx += 256) {
        const int x = idx / wpad, w = idx - x * wpad, k = 16 * w + x;
        pyT[idx] = k < n ? py[k] : 0.0;
    }
    __syncthreads();
    const bool ge = (flags & GMAT_B200_EFF_GE) != 0, swap = (flags & GMAT_B200_EFF_SWAP) != 0;
    // Every warp is on its own: it keeps its hit records in a private shared-memory buffer and appends them to the
    // global list 32 at a time with ONE atomic (the per-iteration block-wide append of the kernel above put an atomic
    // round trip and two block barriers on the path of every candidate).
    gmat_b200_hit *mybuf = hitbuf[warp];
    int nbuf = 0;                                                 // warp-uniform
    for (long long c = (long long)blockIdx.x * 8 + warp; c < total; c += (long long)gridDim.x * 8) {
        const int i = cand[c].x, j = cand[c].y;
        if (lane < 16) prod[warp][lane] = lutU[(size_t)i * 4 + (lane >> 2)] * lutV[(size_t)j * 4 + (lane & 3)];
        __syncwarp();
        const unsigned *ri = reinterpret_cast<const unsigned *>(bed + (size_t)i * row_bytes);
        const unsigned *rj = reinterpret_cast<const unsigned *>(bed + (size_t)j * row_bytes);
        double acc = 0.0;
        for (int w0 = 0; w0 < wpad; w0 += 128) {                  // passes of four words per row and lane
            unsigned ci[4], cj[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int w = w0 + 32 * t + lane;
                const bool in = w < nwords;
                ci[t] = in ? __ldg(ri + w) : 0u;
                cj[t] = in ? __ldg(rj + w) : 0u;
            }
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int w = w0 + 32 * t + lane;
                if (w0 + 32 * t < wpad) {                         // warp-uniform
#pragma unroll
                    for (int x = 0; x < 16; ++x)
                        acc += prod[warp][(((ci[t] >> (2 * x)) & 3u) << 2) | ((cj[t] >> (2 * x)) & 3u)] *
                               pyT[x * wpad + w];
                }
            }
        }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        const double cut = cut_table != nullptr ? cut_table[binI[i] * 10 + binJ[j]] : eff_cut;
        const double a = fabs(acc);
        const bool hit = ge ? a >= cut : a > cut;                 // warp-uniform: acc is the same in every lane
        __syncwarp();                                             // prod[warp] is rewritten by the next candidate
        if (hit) {
            if (lane == 0) {
                gmat_b200_hit h;
                h.snp_0 = swap ? j : i; h.snp_1 = swap ? i : j; h.eff = acc; h.var = 0.0; h.chi = 0.0;
                mybuf[nbuf] = h;
            }
            ++nbuf;
        }
        if (nbuf == 32) { sc_flush_hits(mybuf, nbuf, lane, hits, cap, count); nbuf = 0; }
    }
    if (nbuf > 0) sc_flush_hits(mybuf, nbuf, lane, hits, cap, count);
}

}  // namespace gb

using namespace gb;

namespace {
struct Scratch {                                                  // stream-ordered scratch, freed on scope exit
    cudaStream_t st;
    std::vector<void *> ptrs;
    explicit Scratch(cudaStream_t s) : st(s) {}
    ~Scratch() { for (void *q : ptrs) cudaFreeAsync(q, st); }
    template <class T> T *get(size_t count) {
        void *q = nullptr;
        if (cudaMallocAsync(&q, count * sizeof(T) + 16, st) != cudaSuccess) return nullptr;
        ptrs.push_back(q);
        return reinterpret_cast<T *>(q);
    }
};
inline int sc_planes_of(int tripled) { return tripled ? 3 : 2; }
inline int sc_bits_of(int tripled) { return tripled ? 5 : 7; }
}  // namespace

extern "C" long long gmat_b200_screen_ld(long long n) { return round_up(n, I8_TK); }

extern "C" int gmat_b200_screen_operand(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                        int coding, int tripled, const double *py_dev, signed char *g8_dev,
                                        double *t_dev, void *stream) {
    GB_CHECK_ARG(bed_dev && py_dev && g8_dev && t_dev && n > 0 && m > 0, "null buffer or empty shape");
    GB_CHECK_ARG(coding == GMAT_B200_CODING_ADD || coding == GMAT_B200_CODING_DOM, "coding");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4 && row_bytes % 16 == 0, "row_bytes");
    cudaStream_t st = (cudaStream_t)stream;
    const int ld = (int)round_up(n, I8_TK);
    // byte[code]: plain {0, 0, 1, 2} / {0, 0, 1, 0}; tripled (code 01 = 1/3, x 3): {0, 1, 3, 6} / {0, 1, 3, 0}
    const unsigned tab = tripled ? (coding == GMAT_B200_CODING_DOM ? 0x00030100u : 0x06030100u)
                                 : (coding == GMAT_B200_CODING_DOM ? 0x00010000u : 0x02010000u);
    int rc = launch_i8_dosage(bed_dev, (int)n, m, row_bytes, tab, nullptr, g8_dev, nullptr, ld, st);
    if (rc != GMAT_B200_OK) return rc;
    sc_dot_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(g8_dev, ld, (int)n, m, py_dev, tripled ? 1.0 / 3.0 : 1.0,
                                                          t_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_screen_py(const double *py_dev, long long n, int tripled, signed char *dig_dev,
                                   double *stats_dev, void *stream) {
    GB_CHECK_ARG(py_dev && dig_dev && stats_dev && n > 0, "null buffer or empty shape");
    cudaStream_t st = (cudaStream_t)stream;
    const int ld = (int)round_up(n, I8_TK);
    sc_py_stats_kernel<<<1, 1024, 0, st>>>(py_dev, (int)n, stats_dev);
    GB_LAUNCH_CHECK();
    sc_digits_kernel<<<(unsigned)((ld + 255) / 256), 256, 0, st>>>(py_dev, (int)n, ld, stats_dev, sc_planes_of(tripled),
                                                                   sc_bits_of(tripled), dig_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

template <int S>
static int sc_launch_stage1(const CUtensorMap &map_v, const CUtensorMap &map_p, const ScreenParams &p, int n_ib, int n_jb,
                            cudaStream_t st) {
    const int smem_bytes = ScreenCfg<S>::NS * ScreenCfg<S>::STAGE_BYTES + 1024 + (int)sizeof(ScreenBars);
    static bool attr_set[64];                                     // per device: one driver call, not one per launch
    int dev = 0;
    GB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        GB_CUDA(cudaFuncSetAttribute(i8_screen_pair_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    long long pairs_of_sms = sm_count() / 2;
    const long long max_items = (long long)n_ib * n_jb;
    if (pairs_of_sms > max_items) pairs_of_sms = max_items;
    i8_screen_pair_kernel<S><<<(unsigned)(2 * pairs_of_sms), SC_THREADS, smem_bytes, st>>>(map_v, map_p, p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_epi_eff_i8p(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                     const double *lutU_dev, const double *lutV_dev, const signed char *g8U_dev,
                                     const signed char *g8V_dev, const double *tU_dev, const double *tV_dev,
                                     const signed char *dig_dev, const double *stats_dev, int tripled,
                                     const double *py_dev, const long long *ilist_dev, long long n_i, int flags,
                                     double eff_cut, const double *cut_table_dev, const int *binI_dev,
                                     const int *binJ_dev, gmat_b200_hit *hits_dev, long long cap,
                                     unsigned long long *count_dev, int *cand_dev, long long cand_cap,
                                     unsigned long long *cand_count_dev, int *status_dev, void *stream) {
    GB_CHECK_ARG(bed_dev && lutU_dev && lutV_dev && g8U_dev && g8V_dev && tU_dev && tV_dev && dig_dev && stats_dev &&
                 py_dev && ilist_dev && count_dev && cand_dev && cand_count_dev && status_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && m > 0 && n_i >= 0 && cand_cap > 0, "sizes");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4 && row_bytes % 16 == 0, "row_bytes");
    GB_CHECK_ARG(cap == 0 || hits_dev != nullptr, "hits_dev is NULL");
    GB_CHECK_ARG(cut_table_dev == nullptr || (binI_dev && binJ_dev), "bins required with cut table");
    // int32 plane sums must convert to FP32 without losing the margin's assumptions (|Y| <= 576 n < 2^24) and SNP
    // indices are ints in the kernels
    GB_CHECK_ARG(n < (tripled ? 29000 : 60000) && m < (1LL << 31) && n_i < (1LL << 31),
                 "shape outside the int8 screen's integer ranges (use gmat_b200_epi_eff)");
    cudaStream_t st = (cudaStream_t)stream;
    GB_CUDA(cudaMemsetAsync(status_dev, 0, sizeof(int), st));
    GB_CUDA(cudaMemsetAsync(cand_count_dev, 0, sizeof(unsigned long long), st));
    if (n_i == 0) return GMAT_B200_OK;
    keep_scratch_pool();
    const int S = sc_planes_of(tripled), bits = sc_bits_of(tripled);
    const int ld = (int)round_up(n, I8_TK);
    const int n_slab = (int)((n + I8_SLAB - 1) / I8_SLAB);
    const int n_jb = (int)((m + 2 * I8_TJ - 1) / (2 * I8_TJ));
    Scratch sc(st);
    // one stream-ordered allocation for the call's small buffers: [scalars | tab32 (112) | oV (m) | cV (m)]
    unsigned char *blob = sc.get<unsigned char>(256 + 512 + 8 * (size_t)m);
    GB_CHECK_ARG(blob != nullptr, "out of device memory");
    ScreenScalars *scal = reinterpret_cast<ScreenScalars *>(blob);
    float *tab32 = reinterpret_cast<float *>(blob + 256);
    float *oV = reinterpret_cast<float *>(blob + 768), *cV = oV + m;
    sc_setup_kernel<<<1, 128, 0, st>>>(stats_dev, (int)n, eff_cut, cut_table_dev, S, bits, tripled, scal, tab32,
                                       status_dev);
    GB_LAUNCH_CHECK();
    sc_second_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(lutV_dev, tV_dev, stats_dev, m, oV, cV);
    GB_LAUNCH_CHECK();
    CUtensorMap map_v;
    int rc = make_map(&map_v, g8V_dev, ld, m, ld, I8_TJ);
    if (rc != GMAT_B200_OK) return rc;

    // ---- first SNPs in chunks (plane buffer <= 1 GiB) ----
    long long budget = 1LL << 30;
    if (const char *e = getenv("GMAT_B200_SCREEN_SCRATCH_BYTES")) {     // tests force several chunks with this
        const long long v = atoll(e);
        if (v > 0) budget = v;
    }
    long long chunk = budget / ((long long)S * ld);
    chunk = chunk / SC_TI * SC_TI;
    if (chunk < SC_TI) chunk = SC_TI;
    for (long long c0 = 0; c0 < n_i; c0 += chunk) {
        const int ni = (int)((n_i - c0 < chunk) ? n_i - c0 : chunk);
        const int n_ib = (ni + SC_TI - 1) / SC_TI, plane_rows = n_ib * SC_TI;
        Scratch csc(st);                                          // this chunk's buffers: freed (stream-ordered) per chunk
        // one allocation per chunk: [planes | tU | oU | tiles]
        const size_t plane_bytes = (size_t)S * plane_rows * ld;
        signed char *pu = csc.get<signed char>(plane_bytes + 8 * (size_t)plane_rows + 4 * ((size_t)2 * n_ib + 2));
        GB_CHECK_ARG(pu != nullptr, "out of device memory");
        float *tU = reinterpret_cast<float *>(pu + plane_bytes), *oU = tU + plane_rows;
        int *tiles = reinterpret_cast<int *>(oU + plane_rows);
        sc_planes_kernel<<<(unsigned)plane_rows, 256, 0, st>>>(g8U_dev, ld, ilist_dev + c0, ni, plane_rows, S, dig_dev,
                                                               lutU_dev, tU_dev, pu, tU, oU);
        GB_LAUNCH_CHECK();
        const int pair_space = GMAT_B200_PAIRS_UPPER;            // the screen only ever tests j > i (:69-77, :233-247)
        sc_tiles_kernel<<<1, 1024, 0, st>>>(ilist_dev + c0, ni, n_ib, n_jb, pair_space, tiles, tiles + n_ib + 1);
        GB_LAUNCH_CHECK();
        ScreenParams p;
        p.n_i = ni; p.n_ib = n_ib; p.m = (int)m; p.n_slab = n_slab; p.pair_space = pair_space;
        p.plane_rows = plane_rows; p.item_prefix = tiles; p.jb0 = tiles + n_ib + 1; p.ilist = ilist_dev + c0;
        p.tU = tU; p.oU = oU; p.oV = oV; p.cV = cV; p.binI = binI_dev; p.binJ = binJ_dev;
        p.cut_table = cut_table_dev ? tab32 : nullptr;
        p.scal = scal;
        p.cand = reinterpret_cast<int2 *>(cand_dev); p.cand_cap = cand_cap; p.cand_count = cand_count_dev;
        CUtensorMap map_p;
        rc = make_map(&map_p, pu, ld, (long long)S * plane_rows, ld, SC_TI / 2);
        if (rc != GMAT_B200_OK) return rc;
        rc = S == 2 ? sc_launch_stage1<2>(map_v, map_p, p, n_ib, n_jb, st) : sc_launch_stage1<3>(map_v, map_p, p, n_ib, n_jb, st);
        if (rc != GMAT_B200_OK) return rc;
    }
    // stage 2 over the device-side candidate count (a grid-stride kernel: no host read of the count)
    const bool lanewise = n > SC_ROWS_MAX_N || getenv("GMAT_B200_SCREEN_STAGE2_LANEWISE") != nullptr;
    if (lanewise) {
        screen_exact_kernel<<<(unsigned)(4 * sm_count()), 256, 0, st>>>(
            bed_dev, (int)n, row_bytes, lutU_dev, lutV_dev, py_dev, reinterpret_cast<const int2 *>(cand_dev),
            cand_count_dev, flags, eff_cut, cut_table_dev, binI_dev, binJ_dev, hits_dev, cap, count_dev, cand_cap,
            status_dev);
    } else {
        const int wpad = (int)((((n + 15) >> 4) + 31) & ~31LL);
        const int smem_bytes = 16 * wpad * (int)sizeof(double);
        static bool attr_set[64];
        int dev = 0;
        GB_CUDA(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && !attr_set[dev]) {
            GB_CUDA(cudaFuncSetAttribute(screen_exact_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         16 * ((SC_ROWS_MAX_N / 16 + 31) & ~31) * (int)sizeof(double)));
            attr_set[dev] = true;
        }
        // as many blocks per SM as shared memory and the 64 warp slots allow: the kernel is a chain of dependent loads
        // per candidate (pair -> tables -> rows), latency is hidden by warps
        int per_sm = (200 * 1024) / (smem_bytes + 10 * 1024);
        per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
        screen_exact_rows_kernel<<<(unsigned)(per_sm * sm_count()), 256, smem_bytes, st>>>(
            bed_dev, (int)n, row_bytes, lutU_dev, lutV_dev, py_dev, reinterpret_cast<const int2 *>(cand_dev),
            cand_count_dev, flags, eff_cut, cut_table_dev, binI_dev, binJ_dev, hits_dev, cap, count_dev, cand_cap,
            status_dev);
    }
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

// One-shot form: builds the operands, runs the prepared screen, reads the status back (one synchronisation) and runs
// the FP64 kernel itself when the int8 path declined.  Callers that screen more than once with the same Z'Py (the two
// directions of AD, the stages of the _approx pipelines, repeated calls) keep the operands and use the prepared form.
extern "C" int gmat_b200_epi_eff_i8(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                    const double *lutU_dev, const double *lutV_dev, int codingU, int codingV,
                                    int tripled, const double *py_dev, const long long *ilist_dev, long long n_i,
                                    int flags, double eff_cut, const double *cut_table_dev, const int *binI_dev,
                                    const int *binJ_dev, gmat_b200_hit *hits_dev, long long cap,
                                    unsigned long long *count_dev, int *engine_used, void *stream) {
    GB_CHECK_ARG(bed_dev && lutU_dev && lutV_dev && py_dev && ilist_dev && count_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && m > 0 && n_i >= 0, "sizes");
    GB_CHECK_ARG((codingU == GMAT_B200_CODING_ADD || codingU == GMAT_B200_CODING_DOM) &&
                 (codingV == GMAT_B200_CODING_ADD || codingV == GMAT_B200_CODING_DOM), "coding");
    if (engine_used) *engine_used = 0;
    if (n_i == 0) return GMAT_B200_OK;
    cudaStream_t st = (cudaStream_t)stream;
    keep_scratch_pool();
#define SC_FALLBACK()                                                                                            \
    return gmat_b200_epi_eff(bed_dev, n, m, row_bytes, lutU_dev, lutV_dev, py_dev, ilist_dev, n_i, flags, eff_cut,   \
                             cut_table_dev, binI_dev, binJ_dev, hits_dev, cap, count_dev, stream)
    if (n >= (tripled ? 29000 : 60000) || m >= (1LL << 31) || n_i >= (1LL << 31)) SC_FALLBACK();
    const int ld = (int)round_up(n, I8_TK);
    const long long cand_cap = 1LL << 22;
    Scratch sc(st);
    signed char *g8V = sc.get<signed char>((size_t)m * ld), *g8U = g8V;
    double *tV = sc.get<double>((size_t)m), *tU = tV;
    signed char *dig = sc.get<signed char>((size_t)SC_SMAX * ld);
    double *stats = sc.get<double>(4);
    int *cand = sc.get<int>((size_t)2 * cand_cap), *status = sc.get<int>(2);
    unsigned long long *cand_count = sc.get<unsigned long long>(1);
    GB_CHECK_ARG(g8V && tV && dig && stats && cand && status && cand_count, "out of device memory");
    int rc = gmat_b200_screen_operand(bed_dev, n, m, row_bytes, codingV, tripled, py_dev, g8V, tV, stream);
    if (rc != GMAT_B200_OK) return rc;
    if (codingU != codingV) {
        g8U = sc.get<signed char>((size_t)m * ld);
        tU = sc.get<double>((size_t)m);
        GB_CHECK_ARG(g8U && tU, "out of device memory");
        rc = gmat_b200_screen_operand(bed_dev, n, m, row_bytes, codingU, tripled, py_dev, g8U, tU, stream);
        if (rc != GMAT_B200_OK) return rc;
    }
    rc = gmat_b200_screen_py(py_dev, n, tripled, dig, stats, stream);
    if (rc != GMAT_B200_OK) return rc;
    rc = gmat_b200_epi_eff_i8p(bed_dev, n, m, row_bytes, lutU_dev, lutV_dev, g8U, g8V, tU, tV, dig, stats, tripled, py_dev,
                               ilist_dev, n_i, flags, eff_cut, cut_table_dev, binI_dev, binJ_dev, hits_dev, cap, count_dev,
                               cand, cand_cap, cand_count, status, stream);
    if (rc != GMAT_B200_OK) return rc;
    int hstatus = 0;
    GB_CUDA(cudaMemcpyAsync(&hstatus, status, sizeof(int), cudaMemcpyDeviceToHost, st));
    GB_CUDA(cudaStreamSynchronize(st));
    if (hstatus != 0) {                                           // margin not small against the cut, or too many survivors
        GB_CUDA(cudaMemsetAsync(count_dev, 0, sizeof(unsigned long long), st));
        SC_FALLBACK();
    }
#undef SC_FALLBACK
    if (engine_used) *engine_used = 1;
    return GMAT_B200_OK;
}
