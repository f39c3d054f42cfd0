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
// plane accumulators of a tile sit side by side in TMEM (S x 128 columns), double-buffered (2 x 256 columns).
#include "tc5.cuh"
#include <cstdlib>
#include <vector>

namespace gb {

constexpr int SC_S = 2;                                           // digit planes of py
constexpr int SC_TI = 128;                                        // first SNPs per tile (UMMA N)
constexpr int SC_NS = 6;                                          // smem ring stages
constexpr int SC_THREADS = 192;                                   // producer, issuer, 4 epilogue warps
constexpr int SC_B_BYTES = (SC_TI / 2) * I8_SLAB;                 // 8 KB: this CTA's half of one plane tile
constexpr int SC_STAGE_BYTES = I8_A_BYTES + SC_S * SC_B_BYTES;    // 32 KB per CTA

struct ScreenBars {
    uint64_t full[SC_NS], empty[SC_NS], tfull[2], tempty[2];
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
    float cut;                                                    // eff_cut - margin, rounded down
    float qscale;                                                 // scale * 128^-S
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

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(SC_THREADS, 1)
i8_screen_pair_kernel(const __grid_constant__ CUtensorMap map_v, const __grid_constant__ CUtensorMap map_p,
                      const ScreenParams p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    ScreenBars *bars = reinterpret_cast<ScreenBars *>(smem + SC_NS * SC_STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    const int total_items = p.item_prefix[p.n_ib];

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
                if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
            }
        }
    } else {
        // ================================ epilogue (both CTAs): TMEM lane = second SNP =============
        const int q = warp & 3;
        const int e = (warp - 2) * 32 + lane;                     // the column whose constants this thread stages
        int abuf = 0; unsigned aphase = 0;
        for (int item = cluster_id; item < total_items; item += n_clusters) {
            const int ib = sc_find(p.item_prefix, p.n_ib, item);
            const int jb = p.jb0[ib] + (item - p.item_prefix[ib]);
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
                int v0[32], v1[32];
                tmem_ld32(taddr + ch * 32, v0);
                tmem_ld32(taddr + SC_TI + ch * 32, v1);
                tmem_ld_wait();
                unsigned mask = 0;
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const int col = ch * 32 + c;
                    const float t32 = fmaf((float)v0[c], 128.f, (float)v1[c]);
                    const float r = fmaf(-oj, bars->col_t[abuf][col], -bars->col_o[abuf][col] * cj);
                    const float eff = fmaf(t32, p.qscale, r);
                    const float cut = p.cut_table != nullptr ? bars->table[bars->col_bin[abuf][col] + bj] : p.cut;
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
            if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
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
// balanced base-128 digits of py[k] / scale, |py / scale| <= 0.49: dig[s][k] in [-64, 63], zero for k >= n
__global__ void sc_digits_kernel(const double *__restrict__ py, int n, int ld, double inv_scale,
                                 signed char *__restrict__ dig) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ld) return;
    long long t = k < n ? __double2ll_rn(py[k] * inv_scale * 16384.0) : 0;    // 128^SC_S
#pragma unroll
    for (int s = SC_S - 1; s >= 1; --s) {
        const long long d = ((t + 64) & 127) - 64;
        t = (t - d) >> 7;
        dig[(size_t)s * ld + k] = (signed char)d;
    }
    dig[k] = (signed char)t;
}
// t[snp] = sum_k a_k py_k over the int8 codes of one SNP row (one warp per SNP, fixed order)
__global__ void __launch_bounds__(256) sc_dot_kernel(const signed char *__restrict__ g8, int ld, int n, long long m,
                                                     const double *__restrict__ py, double *__restrict__ t) {
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
    if (lane == 0) t[snp] = acc;
}
// per second SNP: o_j = -lutV[j][0], c_j = t_j - o_j P  (FP32 copies for stage 1)
__global__ void sc_second_kernel(const double *__restrict__ lutV, const double *__restrict__ tV, double P, long long m,
                                 float *__restrict__ oV, float *__restrict__ cV) {
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m) return;
    const double o = -lutV[j * 4];
    oV[j] = (float)o;
    cV[j] = (float)(tV[j] - o * P);
}
// planes of the first SNPs of the launch: PU[s][row][k] = a_ik * dig[s][k] (row = position in ilist, zero past n_i)
// plus the FP32 per-first-SNP constants
__global__ void __launch_bounds__(256) sc_planes_kernel(const signed char *__restrict__ g8U, int ld,
                                                        const long long *__restrict__ ilist, int n_i, int plane_rows,
                                                        const signed char *__restrict__ dig,
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
#pragma unroll
        for (int s = 0; s < SC_S; ++s) {
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
__global__ void sc_table_kernel(const double *__restrict__ table, double margin, float *__restrict__ out) {
    if (threadIdx.x < 111) out[threadIdx.x] = __double2float_rd(table[threadIdx.x] - margin);
}

// Stage 2: eff of the candidates in FP64 from the LUT values (the arithmetic of the FP64 engine and of the
// reference: lutU[i][c_ik] * lutV[j][c_jk] * py[k]), the reference's threshold, the hit record.  One warp per
// candidate, lanes stride over the bytes of the two .bed rows, fixed reduction order.
__global__ void __launch_bounds__(256) screen_exact_kernel(
    const unsigned char *__restrict__ bed, int n, long long row_bytes, const double *__restrict__ lutU,
    const double *__restrict__ lutV, const double *__restrict__ py, const int2 *__restrict__ cand,
    const unsigned long long *__restrict__ n_cand, int flags, double eff_cut, const double *__restrict__ cut_table,
    const int *__restrict__ binI, const int *__restrict__ binJ, gmat_b200_hit *__restrict__ hits, long long cap,
    unsigned long long *__restrict__ count) {
    const int lane = threadIdx.x & 31;
    const long long total = (long long)*n_cand;
    const bool ge = (flags & GMAT_B200_EFF_GE) != 0, swap = (flags & GMAT_B200_EFF_SWAP) != 0;
    for (long long c = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); c < total; c += (long long)gridDim.x * 8) {
        const int i = cand[c].x, j = cand[c].y;
        const double lu[4] = {lutU[(size_t)i * 4], lutU[(size_t)i * 4 + 1], lutU[(size_t)i * 4 + 2], lutU[(size_t)i * 4 + 3]};
        const double lv[4] = {lutV[(size_t)j * 4], lutV[(size_t)j * 4 + 1], lutV[(size_t)j * 4 + 2], lutV[(size_t)j * 4 + 3]};
        const unsigned char *ri = bed + (size_t)i * row_bytes, *rj = bed + (size_t)j * row_bytes;
        double acc = 0.0;
        for (int b = lane; b * 4 < n; b += 32) {
            const unsigned ci = ri[b], cj = rj[b];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int k = b * 4 + r;
                if (k < n) acc += lu[(ci >> (2 * r)) & 3u] * lv[(cj >> (2 * r)) & 3u] * py[k];
            }
        }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        if (lane == 0) {
            const double cut = cut_table != nullptr ? cut_table[binI[i] * 10 + binJ[j]] : eff_cut;
            const double a = fabs(acc);
            if (ge ? a >= cut : a > cut) {
                const unsigned long long pos = atomicAdd(count, 1ULL);
                if ((long long)pos < cap) {
                    gmat_b200_hit h;
                    h.snp_0 = swap ? j : i; h.snp_1 = swap ? i : j; h.eff = acc; h.var = 0.0; h.chi = 0.0;
                    hits[pos] = h;
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
                                 void *stream);

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
}  // namespace

extern "C" int gmat_b200_epi_eff_i8(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                    const double *lutU_dev, const double *lutV_dev, int codingU, int codingV,
                                    const double *py_dev, const long long *ilist_dev, long long n_i, int flags,
                                    double eff_cut, const double *cut_table_dev, const int *binI_dev,
                                    const int *binJ_dev, gmat_b200_hit *hits_dev, long long cap,
                                    unsigned long long *count_dev, int *engine_used, void *stream) {
    GB_CHECK_ARG(bed_dev && lutU_dev && lutV_dev && py_dev && ilist_dev && count_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && m > 0 && n_i >= 0, "sizes");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4 && row_bytes % 16 == 0, "row_bytes");
    GB_CHECK_ARG((codingU == GMAT_B200_CODING_ADD || codingU == GMAT_B200_CODING_DOM) &&
                 (codingV == GMAT_B200_CODING_ADD || codingV == GMAT_B200_CODING_DOM), "coding");
    GB_CHECK_ARG(cap == 0 || hits_dev != nullptr, "hits_dev is NULL");
    GB_CHECK_ARG(cut_table_dev == nullptr || (binI_dev && binJ_dev), "bins required with cut table");
    if (engine_used) *engine_used = 0;
    if (n_i == 0) return GMAT_B200_OK;
    cudaStream_t st = (cudaStream_t)stream;
    keep_scratch_pool();
#define SC_FALLBACK()                                                                                            \
    return gmat_b200_epi_eff(bed_dev, n, m, row_bytes, lutU_dev, lutV_dev, py_dev, ilist_dev, n_i, flags, eff_cut,   \
                             cut_table_dev, binI_dev, binJ_dev, hits_dev, cap, count_dev, stream)
    // int32 plane sums must convert to FP32 exactly (|Y| <= 256 n < 2^24) and SNP indices are ints in the kernels
    if (n >= 60000 || m >= (1LL << 31) || n_i >= (1LL << 31)) SC_FALLBACK();
    Scratch sc(st);
    // ---- scale of py, margin of stage 1, is the int8 path worth it? ----
    double *stats = sc.get<double>(2);
    GB_CHECK_ARG(stats != nullptr, "out of device memory");
    sc_py_stats_kernel<<<1, 1024, 0, st>>>(py_dev, (int)n, stats);
    GB_LAUNCH_CHECK();
    double hstats[2];
    double htable[111];
    GB_CUDA(cudaMemcpyAsync(hstats, stats, sizeof(hstats), cudaMemcpyDeviceToHost, st));
    if (cut_table_dev) GB_CUDA(cudaMemcpyAsync(htable, cut_table_dev, sizeof(htable), cudaMemcpyDeviceToHost, st));
    GB_CUDA(cudaStreamSynchronize(st));
    const double maxabs = hstats[0], P = hstats[1];
    double cut_min = eff_cut;
    if (cut_table_dev) { cut_min = htable[0]; for (int q = 1; q < 111; ++q) cut_min = htable[q] < cut_min ? htable[q] : cut_min; }
    if (!(maxabs > 0.0) || !(maxabs < 1e300) || !(cut_min == cut_min)) SC_FALLBACK();
    const double scale = maxabs / 0.49;
    // everything stage 1 drops: digit truncation scale 128^-S 2n, FP32 rounding of Q~ and of the rank-one terms
    const double margin = scale * (double)n * (2.0 / 16384.0 + 4.0e-6);
    if (!(margin < 0.25 * cut_min)) SC_FALLBACK();            // a cut near zero keeps (almost) every pair: FP64 kernel
    const int ld = (int)round_up(n, I8_TK);
    const int n_slab = (int)((n + I8_SLAB - 1) / I8_SLAB);
    const int n_jb = (int)((m + 2 * I8_TJ - 1) / (2 * I8_TJ));

    // ---- operands that do not depend on the first-SNP list ----
    signed char *g8V = sc.get<signed char>((size_t)m * ld), *g8U = g8V;
    signed char *dig = sc.get<signed char>((size_t)SC_S * ld);
    double *tV = sc.get<double>((size_t)m), *tUall = tV;
    float *oV = sc.get<float>((size_t)m), *cV = sc.get<float>((size_t)m), *tab32 = sc.get<float>(112);
    GB_CHECK_ARG(g8V && dig && tV && oV && cV && tab32, "out of device memory");
    int rc = gmat_b200_i8_dosage(bed_dev, n, m, row_bytes, codingV, g8V, nullptr, ld, stream);
    if (rc != GMAT_B200_OK) return rc;
    sc_dot_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(g8V, ld, (int)n, m, py_dev, tV);
    GB_LAUNCH_CHECK();
    if (codingU != codingV) {
        g8U = sc.get<signed char>((size_t)m * ld);
        tUall = sc.get<double>((size_t)m);
        GB_CHECK_ARG(g8U && tUall, "out of device memory");
        rc = gmat_b200_i8_dosage(bed_dev, n, m, row_bytes, codingU, g8U, nullptr, ld, stream);
        if (rc != GMAT_B200_OK) return rc;
        sc_dot_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(g8U, ld, (int)n, m, py_dev, tUall);
        GB_LAUNCH_CHECK();
    }
    sc_digits_kernel<<<(unsigned)((ld + 255) / 256), 256, 0, st>>>(py_dev, (int)n, ld, 1.0 / scale, dig);
    GB_LAUNCH_CHECK();
    sc_second_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(lutV_dev, tV, P, m, oV, cV);
    GB_LAUNCH_CHECK();
    if (cut_table_dev) {
        sc_table_kernel<<<1, 128, 0, st>>>(cut_table_dev, margin, tab32);
        GB_LAUNCH_CHECK();
    }

    // ---- candidates of the whole call (stage 2 runs once, after the last chunk) ----
    const long long cand_cap = 1LL << 22;
    int2 *cand = sc.get<int2>((size_t)cand_cap);
    unsigned long long *cand_count = sc.get<unsigned long long>(1);
    GB_CHECK_ARG(cand && cand_count, "out of device memory");
    GB_CUDA(cudaMemsetAsync(cand_count, 0, sizeof(unsigned long long), st));

    CUtensorMap map_v;
    rc = make_map(&map_v, g8V, ld, m, ld, I8_TJ);
    if (rc != GMAT_B200_OK) return rc;
    const int smem_bytes = SC_NS * SC_STAGE_BYTES + 1024 + (int)sizeof(ScreenBars);
    GB_CUDA(cudaFuncSetAttribute(i8_screen_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));

    // ---- first SNPs in chunks (plane buffer <= 1 GiB) ----
    long long budget = 1LL << 30;
    if (const char *e = getenv("GMAT_B200_SCREEN_SCRATCH_BYTES")) {     // tests force several chunks with this
        const long long v = atoll(e);
        if (v > 0) budget = v;
    }
    long long chunk = budget / ((long long)SC_S * ld);
    chunk = chunk / SC_TI * SC_TI;
    if (chunk < SC_TI) chunk = SC_TI;
    for (long long c0 = 0; c0 < n_i; c0 += chunk) {
        const int ni = (int)((n_i - c0 < chunk) ? n_i - c0 : chunk);
        const int n_ib = (ni + SC_TI - 1) / SC_TI, plane_rows = n_ib * SC_TI;
        Scratch csc(st);                                          // this chunk's buffers: freed (stream-ordered) per chunk
        signed char *pu = csc.get<signed char>((size_t)SC_S * plane_rows * ld);
        float *tU = csc.get<float>((size_t)plane_rows), *oU = csc.get<float>((size_t)plane_rows);
        int *tiles = csc.get<int>((size_t)2 * n_ib + 2);
        GB_CHECK_ARG(pu && tU && oU && tiles, "out of device memory");
        sc_planes_kernel<<<(unsigned)plane_rows, 256, 0, st>>>(g8U, ld, ilist_dev + c0, ni, plane_rows, dig, lutU_dev,
                                                               tUall, pu, tU, oU);
        GB_LAUNCH_CHECK();
        const int pair_space = GMAT_B200_PAIRS_UPPER;            // the screen only ever tests j > i (:69-77, :233-247)
        sc_tiles_kernel<<<1, 1024, 0, st>>>(ilist_dev + c0, ni, n_ib, n_jb, pair_space, tiles, tiles + n_ib + 1);
        GB_LAUNCH_CHECK();
        ScreenParams p;
        p.n_i = ni; p.n_ib = n_ib; p.m = (int)m; p.n_slab = n_slab; p.pair_space = pair_space;
        p.plane_rows = plane_rows; p.item_prefix = tiles; p.jb0 = tiles + n_ib + 1; p.ilist = ilist_dev + c0;
        p.tU = tU; p.oU = oU; p.oV = oV; p.cV = cV; p.binI = binI_dev; p.binJ = binJ_dev;
        p.cut_table = cut_table_dev ? tab32 : nullptr;
        p.cut = __builtin_nextafterf((float)(eff_cut - margin), -1.0e38f);
        p.qscale = (float)(scale / 16384.0);
        p.cand = cand; p.cand_cap = cand_cap; p.cand_count = cand_count;
        CUtensorMap map_p;
        rc = make_map(&map_p, pu, ld, (long long)SC_S * plane_rows, ld, SC_TI / 2);
        if (rc != GMAT_B200_OK) return rc;
        long long pairs_of_sms = sm_count() / 2;
        const long long max_items = (long long)n_ib * n_jb;
        if (pairs_of_sms > max_items) pairs_of_sms = max_items;
        i8_screen_pair_kernel<<<(unsigned)(2 * pairs_of_sms), SC_THREADS, smem_bytes, st>>>(map_v, map_p, p);
        GB_LAUNCH_CHECK();
    }
    unsigned long long hcount = 0;
    GB_CUDA(cudaMemcpyAsync(&hcount, cand_count, sizeof(hcount), cudaMemcpyDeviceToHost, st));
    GB_CUDA(cudaStreamSynchronize(st));
    if ((long long)hcount > cand_cap) SC_FALLBACK();              // too many survivors for the buffer: FP64 kernel
#undef SC_FALLBACK
    if (hcount > 0) {
        long long blocks = ((long long)hcount + 7) / 8;
        if (blocks > 8LL * sm_count()) blocks = 8LL * sm_count();
        screen_exact_kernel<<<(unsigned)blocks, 256, 0, st>>>(bed_dev, (int)n, row_bytes, lutU_dev, lutV_dev, py_dev,
                                                              cand, cand_count, flags, eff_cut, cut_table_dev, binI_dev,
                                                              binJ_dev, hits_dev, cap, count_dev);
        GB_LAUNCH_CHECK();
    }
    if (engine_used) *engine_used = 1;
    return GMAT_B200_OK;
}
