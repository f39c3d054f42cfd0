// INT8 tensor-core engine of the exact epistasis test (sm_100a: tcgen05.mma kind::i8, TMEM, TMA).
//
// The exact test of the reference (gmat/remma/remma_epiAA.py:72-77) needs, per SNP pair, var = w' A w with
// A = Z'PZ (n x n, FP64) and w = u_i o v_j.  With v_j = a_j - o_j (a_j the integer dosage / heterozygote
// indicator, o_j = 2p_j or 2p_j q_j) and B_i = diag(u_i) A diag(u_i):
//
//     var_ij = a_j' B_i a_j  -  2 o_j a_j.r_i  +  o_j^2 s_i,      r_i = u_i o (A u_i),  s_i = sum(r_i)
//     eff_ij = a_j.(u_i o Z'Py) - o_j (u_i . Z'Py)
//
// Only a_j' B_i a_j costs n^2 per pair, and a_j is an exact small-integer vector.  B_i is split error-free into
// S signed base-128 digit planes (int8); "plane x genotype" products are exact int8 x int8 -> int32 tensor-core
// GEMMs, recombined in FP64 (S = 8 carries 56 bits: the split is finer than the FP64 rounding of B_i itself).
// The FP64 DMMA kernel (exact.cu) issues 2 n^2 FP64 flop per pair at 37 TFLOP/s; this path issues S n^2 / 2
// int8 MACs per pair (block-lower-triangle of the symmetric B_i, diagonal blocks halved, result doubled) on a
// pipe that is two orders of magnitude faster.
//
// Kernel roles (one CTA per SM, 192 threads): warp 0 = TMA producer, warp 1 = tcgen05.mma issuer (one lane),
// warps 2..5 = epilogue (TMEM -> registers -> a_jk-weighted row sums, kept per thread: the TMEM lane IS the
// pair, so no cross-lane reduction exists anywhere).  MMA shape M=128 (second SNPs j) x N=256 (rows k of a
// plane) x K=32 ids; accumulators double-buffered in TMEM (2 x 256 columns).
#include "tc5.cuh"

namespace gb {

constexpr int I8_THREADS = 192;
constexpr int I8_MAX_PLANES = 8;

struct I8Params {
    int n, m, ld, np, n_planes, n_kt, n_slab;
    int n_i, total_items, pair_space;
    const long long *ilist;
    const int *item_prefix;      // [n_i+1] first work item of every first SNP
    const int *jt0;              // [n_i]   first j tile of every first SNP
    const signed char *g8;       // [m][ld] integer coded second SNPs
    const double *scale, *s_i, *upy;      // [n_i]
    const double *t1, *e;        // [n_i][m]  a_j.r_i  and  a_j.(u_i o py)
    const double *offs;          // [m]   o_j
    double chi_cut;
    gmat_b200_hit *hits;
    long long cap;
    unsigned long long *count;
    double *q_debug;             // optional [n_i][m]: a_j' B_i a_j
    int trim;                    // CTA-pair kernel: issue only the plane rows that can be non-zero (i8p_shape)
};

// Epilogue of one accumulator tile for one TMEM lane (= one second SNP): sum_k a_jk * Y[j][k] over the 256 plane
// rows of the tile, all in INTEGER registers.  No FP64 instruction may sit in this loop: DADD / I2F.F64 issue into
// the pipe the tensor MMAs occupy (ncu: 37 % of all stall samples sat on the one DADD of the first version and the
// tensor pipe idled 37 % of the time).  |Y| <= 2*127*n, so a 32-column partial fits int32 for n < 1.3e5.
__device__ __forceinline__ int i8_dot_chunk(const int (&v)[32], const int4 &lo, const int4 &hi) {
    const int aw[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;                           // four independent chains
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        c0 += (aw[w] & 0xff) * v[4 * w];
        c1 += ((aw[w] >> 8) & 0xff) * v[4 * w + 1];
        c2 += ((aw[w] >> 16) & 0xff) * v[4 * w + 2];
        c3 += ((aw[w] >> 24) & 0xff) * v[4 * w + 3];
    }
    return (c0 + c1) + (c2 + c3);                                 // |.| <= 32 * 2 * 2*127*n: int32 for n < 1.3e5
}
// The TMEM load of chunk ch+1 is in flight while chunk ch is multiplied (two register buffers).
__device__ __forceinline__ long long i8_drain_tile(uint32_t taddr, const int4 (&a)[I8_TK / 16]) {
    long long total = 0;
    int v0[32], v1[32];
    tmem_ld32(taddr, v0);
#pragma unroll
    for (int ch = 0; ch < I8_TK / 32; ch += 2) {
        tmem_ld_wait();
        tmem_ld32(taddr + (ch + 1) * 32, v1);
        total += i8_dot_chunk(v0, a[2 * ch], a[2 * ch + 1]);
        tmem_ld_wait();
        if (ch + 2 < I8_TK / 32) tmem_ld32(taddr + (ch + 2) * 32, v0);
        total += i8_dot_chunk(v1, a[2 * ch + 2], a[2 * ch + 3]);
    }
    return total;
}
// the 256 dosage bytes of this lane's SNP for one row tile: issued BEFORE the wait on the accumulator, so the L2
// latency (the first version prefetched one 32-byte chunk ahead and stalled on it: 23 % of all samples) is hidden
__device__ __forceinline__ void i8_load_dosage(const signed char *row, int4 (&a)[I8_TK / 16]) {
    const int4 *ap = reinterpret_cast<const int4 *>(row);
#pragma unroll
    for (int q = 0; q < I8_TK / 16; ++q) a[q] = __ldg(ap + q);
}

// Finish one pair (first SNP ii of the launch, second SNP j) from qacc = sum_s 128^-(s+1) sum_k a_jk Y_s[j][k]: the
// FP64 side terms, chi, threshold, warp-ballot compaction, 32-byte hit record (remma_epiAA.py:75-82).  Called by
// all 32 lanes of an epilogue warp.
__device__ __forceinline__ void i8_finish_pair(const I8Params &p, int ii, int j, int lane, double qacc) {
    const int i = (int)p.ilist[ii];
    const bool valid = j < p.m && (p.pair_space == GMAT_B200_PAIRS_ALL || j > i);
    const int jc = j < p.m ? j : p.m - 1;
    const double quad = 2.0 * p.scale[ii] * qacc;                   // a_j' B_i a_j
    const double o = p.offs[jc];
    const size_t at = (size_t)ii * p.m + jc;
    double var = quad - 2.0 * o * p.t1[at] + o * o * p.s_i[ii];
    double eff = p.e[at] - o * p.upy[ii];
    if (o != o) { var = 0.0; eff = 0.0; }                          // monomorphic second SNP: w == 0 exactly
    else if (o > 1.0e300) { var = o - o; eff = var; }             // missing genotype in the second SNP: NaN, as numpy
    const double chi = eff * eff / var;
    if (p.q_debug != nullptr && j < p.m) p.q_debug[at] = quad;
    gmat_b200_hit h;
    h.snp_0 = i; h.snp_1 = j; h.eff = eff; h.var = var; h.chi = chi;
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

struct I8Bars {
    uint64_t full[I8_NS], empty[I8_NS], tfull[2], tempty[2];
    uint32_t tmem_base;
};

__device__ __forceinline__ int i8_slabs_of(int kt, int n_slab) {
    const int want = 2 * kt + 2;              // block-lower-triangle: id slabs 0 .. (last 128-row block of the tile)
    return want < n_slab ? want : n_slab;
}
__device__ __forceinline__ int i8_find_first(const int *prefix, int n_i, int item) {
    int lo = 0, hi = n_i - 1;                 // largest ii with prefix[ii] <= item
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (prefix[mid] <= item) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(I8_THREADS, 1)
i8_scan_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_p,
               const I8Params p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    I8Bars *bars = reinterpret_cast<I8Bars *>(smem + I8_NS * I8_STAGE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < I8_NS; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&bars->tfull[b], 1); mbar_init(&bars->tempty[b], 4); }
        mbar_fence_init();
    }
    if (warp == 1) {                          // TMEM: all 512 columns (two 256-column accumulators)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                         smem_u32(&bars->tmem_base)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================================ TMA producer ==========================================
        if (lane == 0) {
            int stage = 0; unsigned phase = 0;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                const int ii = i8_find_first(p.item_prefix, p.n_i, item);
                const int jt = p.jt0[ii] + (item - p.item_prefix[ii]);
                for (int s = 0; s < p.n_planes; ++s) {
                    const int prow = (ii * p.n_planes + s) * p.np;
                    for (int kt = 0; kt < p.n_kt; ++kt) {
                        const int nl = i8_slabs_of(kt, p.n_slab);
                        for (int l = 0; l < nl; ++l) {
                            mbar_wait(&bars->empty[stage], phase ^ 1u);
                            mbar_expect_tx(&bars->full[stage], I8_STAGE_BYTES);
                            unsigned char *sa = smem + stage * I8_STAGE_BYTES;
                            tma_load_2d(sa, &map_g, l * I8_SLAB, jt * I8_TJ, &bars->full[stage]);
                            tma_load_2d(sa + I8_A_BYTES, &map_p, l * I8_SLAB, prow + kt * I8_TK, &bars->full[stage]);
                            if (++stage == I8_NS) { stage = 0; phase ^= 1u; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (one thread) ================================
        if (lane == 0) {
            int stage = 0; unsigned phase = 0;
            int abuf = 0; unsigned aphase = 0;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                for (int s = 0; s < p.n_planes; ++s) {
                    for (int kt = 0; kt < p.n_kt; ++kt) {
                        const int nl = i8_slabs_of(kt, p.n_slab);
                        mbar_wait(&bars->tempty[abuf], aphase ^ 1u);       // epilogue drained this accumulator
                        tc_fence_after();
                        const uint32_t tmem_d = tmem_base + (uint32_t)(abuf * I8_TK);
                        for (int l = 0; l < nl; ++l) {
                            mbar_wait(&bars->full[stage], phase);           // TMA bytes have landed
                            tc_fence_after();
                            const uint32_t sa = smem_u32(smem + stage * I8_STAGE_BYTES);
                            const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + I8_A_BYTES);
#pragma unroll
                            for (int k = 0; k < I8_SLAB / 32; ++k)          // K = 32 ids per MMA: +32 B in the atom
                                umma_i8(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), I8_IDESC,
                                        (uint32_t)((l | k) != 0));
                            tc_commit(&bars->empty[stage]);                 // frees the smem slot when the MMAs retire
                            if (++stage == I8_NS) { stage = 0; phase ^= 1u; }
                        }
                        tc_commit(&bars->tfull[abuf]);                      // accumulator complete
                        if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
                    }
                }
            }
        }
    } else {
        // ================================ epilogue warps =========================================
        const int q = warp & 3;                                             // TMEM lane quarter of this warp
        int abuf = 0; unsigned aphase = 0;
        for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
            const int ii = i8_find_first(p.item_prefix, p.n_i, item);
            const int jt = p.jt0[ii] + (item - p.item_prefix[ii]);
            const int j = jt * I8_TJ + q * 32 + lane;
            const signed char *arow = p.g8 + (size_t)(j < p.m ? j : p.m - 1) * p.ld;
            double qacc = 0.0, wgt = 1.0;
            for (int s = 0; s < p.n_planes; ++s) {
                wgt *= 0.0078125;                                           // 128^-(s+1)
                long long plane_sum = 0;                                    // exact
                for (int kt = 0; kt < p.n_kt; ++kt) {
                    int4 abytes[I8_TK / 16];
                    i8_load_dosage(arow + kt * I8_TK, abytes);
                    mbar_wait_warp(&bars->tfull[abuf], aphase);
                    tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(abuf * I8_TK);
                    plane_sum += i8_drain_tile(taddr, abytes);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bars->tempty[abuf]);
                    if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
                }
                qacc += (double)plane_sum * wgt;                         // the only FP64 work per plane
            }
            i8_finish_pair(p, ii, j, lane, qacc);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// ---- CTA-pair variant: tcgen05.mma.cta_group::2 ---------------------------------------------------------------
// Two CTAs of a cluster (one TPC) own the two 128-SNP tiles of a work item and HALF of every plane tile each
// (128 of its 256 rows); one MMA of the leader CTA computes M = 256 SNPs x N = 256 plane rows across both SMs,
// each tensor core reading its own genotype tile and the two plane halves.  Per SM and per 512 MMA cycles the
// shared memory sees 32 KB of TMA writes and 32 KB of operand reads instead of 48 + 48, which is what holds the
// single-CTA kernel at ~60 % of the issue rate.  Accumulators stay double-buffered (2 x 256 TMEM columns per CTA).
constexpr int I8P_NS = 6;
constexpr int I8P_HALF = I8_TK / 2;                               // plane rows each CTA loads
constexpr int I8P_STAGE_BYTES = I8_A_BYTES + I8P_HALF * I8_SLAB;  // 32 KB per CTA
struct I8PBars {
    uint64_t full[I8P_NS], empty[I8P_NS], tfull[2], tempty[2];
    uint32_t tmem_base;
};

// Plane rows of tile kt that slab l (128 ids) can touch: MMA N = nn rows starting row_off rows into the tile.
//  * the last tile holds only n - 256 kt rows of data (rounded up to the UMMA N granularity of 16);
//  * slab 2kt+1 lies right of the diagonal for the tile's first 128 rows (the split kernel leaves that 128 x 128
//    block zero), so only rows 128.. of the tile are multiplied and accumulated into TMEM columns 128.. .
// The two CTAs supply nn/2 rows each (tcgen05 cta_group::2 takes the first half of N from the leader's shared
// memory, the second from the peer's), so CTA r loads the 128-row box starting at row_off + r nn/2 and the MMA reads
// its first nn/2 rows.  TMEM columns that a trimmed MMA does not write keep stale integers; the epilogue multiplies
// them by dosage bytes of ids >= n, which are zero.  n=2000: 65.3 instead of 72 slab-MMAs per plane (-9 %).
__device__ __forceinline__ void i8p_shape(const I8Params &p, int kt, int l, int &row_off, int &nn) {
    row_off = 0; nn = I8_TK;
    if (!p.trim) return;
    const int rows = p.n - kt * I8_TK;
    if (rows < I8_TK) nn = (rows + 15) & ~15;
    if (l == 2 * kt + 1) { row_off = I8_TK / 2; nn -= I8_TK / 2; }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(I8_THREADS, 1)
i8_scan_pair_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_ph,
                    const I8Params p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    I8PBars *bars = reinterpret_cast<I8PBars *>(smem + I8P_NS * I8P_STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < I8P_NS; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&bars->tfull[b], 1); mbar_init(&bars->tempty[b], 8); }
        mbar_fence_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                         smem_u32(&bars->tmem_base)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                       // both CTAs: barriers initialised, TMEM allocated
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    // The producer and issuer warps run their loops CONVERGED and hand only the single-thread instructions to one
    // elected lane (elect_one, tc5.cuh): addresses and descriptors then live in uniform registers.
    if (warp == 0) {
        int stage = 0; unsigned phase = 0;
        for (int item = cluster_id; item < p.total_items; item += n_clusters) {
            const int ii = i8_find_first(p.item_prefix, p.n_i, item);
            const int jt = 2 * (p.jt0[ii] + (item - p.item_prefix[ii])) + rank;
            for (int s = 0; s < p.n_planes; ++s) {
                const int prow = (ii * p.n_planes + s) * p.np;
                for (int kt = 0; kt < p.n_kt; ++kt) {
                    const int nl = i8_slabs_of(kt, p.n_slab);
                    for (int l = 0; l < nl; ++l) {
                        int row_off, nn;
                        i8p_shape(p, kt, l, row_off, nn);
                        mbar_wait(&bars->empty[stage], phase ^ 1u);
                        if (elect_one()) {
                            if (rank == 0) mbar_expect_tx(&bars->full[stage], 2 * I8P_STAGE_BYTES);
                            unsigned char *sa = smem + stage * I8P_STAGE_BYTES;
                            tma_load_2d_pair(sa, &map_g, l * I8_SLAB, jt * I8_TJ, &bars->full[stage]);
                            tma_load_2d_pair(sa + I8_A_BYTES, &map_ph, l * I8_SLAB,
                                             prow + kt * I8_TK + row_off + rank * (nn >> 1), &bars->full[stage]);
                        }
                        if (++stage == I8P_NS) { stage = 0; phase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {
            int stage = 0; unsigned phase = 0;
            int abuf = 0; unsigned aphase = 0;
            for (int item = cluster_id; item < p.total_items; item += n_clusters) {
                for (int s = 0; s < p.n_planes; ++s) {
                    for (int kt = 0; kt < p.n_kt; ++kt) {
                        const int nl = i8_slabs_of(kt, p.n_slab);
                        mbar_wait(&bars->tempty[abuf], aphase ^ 1u);       // 8 epilogue warps of the pair drained it
                        tc_fence_after();
                        const uint32_t tmem_d = tmem_base + (uint32_t)(abuf * I8_TK);
                        for (int l = 0; l < nl; ++l) {
                            mbar_wait(&bars->full[stage], phase);
                            tc_fence_after();
                            const uint32_t sa = smem_base + (uint32_t)(stage * I8P_STAGE_BYTES);
                            const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + I8_A_BYTES);
                            int row_off, nn;
                            i8p_shape(p, kt, l, row_off, nn);
                            const uint32_t idesc = i8p_idesc(nn);
                            if (elect_one()) {
#pragma unroll
                                for (int k = 0; k < I8_SLAB / 32; ++k)
                                    umma_i8_pair(tmem_d + (uint32_t)row_off, da + (uint64_t)(2 * k),
                                                 db + (uint64_t)(2 * k), idesc, (uint32_t)((l | k) != 0));
                                tc_commit_pair(&bars->empty[stage]);
                                if (l == nl - 1) tc_commit_pair(&bars->tfull[abuf]);
                            }
                            if (++stage == I8P_NS) { stage = 0; phase ^= 1u; }
                        }
                        if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
                    }
                }
            }
        }
    } else {
        const int q = warp & 3;
        int abuf = 0; unsigned aphase = 0;
        for (int item = cluster_id; item < p.total_items; item += n_clusters) {
            const int ii = i8_find_first(p.item_prefix, p.n_i, item);
            const int jt = 2 * (p.jt0[ii] + (item - p.item_prefix[ii])) + rank;
            const int j = jt * I8_TJ + q * 32 + lane;
            const signed char *arow = p.g8 + (size_t)(j < p.m ? j : p.m - 1) * p.ld;
            double qacc = 0.0, wgt = 1.0;
            for (int s = 0; s < p.n_planes; ++s) {
                wgt *= 0.0078125;
                long long plane_sum = 0;                                    // exact
                for (int kt = 0; kt < p.n_kt; ++kt) {
                    int4 abytes[I8_TK / 16];
                    i8_load_dosage(arow + kt * I8_TK, abytes);
                    mbar_wait_warp(&bars->tfull[abuf], aphase);
                    tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(abuf * I8_TK);
                    plane_sum += i8_drain_tile(taddr, abytes);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_leader(&bars->tempty[abuf]);
                    if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
                }
                qacc += (double)plane_sum * wgt;                         // the only FP64 work per plane
            }
            i8_finish_pair(p, ii, j, lane, qacc);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                       // nobody frees TMEM / exits while the peer may still touch this CTA
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// ---- dual-plane CTA-pair variant (variant 4) ---------------------------------------------------------------------
// ncu on the kernel above: 97.5 GB of TMA loads in 7.1 ms = 13.7 TB/s, the L2 -> SM ceiling of the chip (the
// guide's ~6300 B/clk LTS cap); trimming MMAs (variant 3) left the time unchanged.  A 256 x 256 tile per CTA pair
// needs 64 B/clk/SM at the full MMA rate, 148 SMs 9.5 KB/clk -- the kernel is bound by operand bytes, not by the
// tensor pipe.  This variant multiplies every genotype tile with the tiles of TWO digit planes (s, s+1) while it is in
// shared memory: 48 KB per CTA and stage feed 8 instead of 4 MMAs (47 B/clk/SM).  The two accumulators take all
// 512 TMEM columns, so there is no second buffer: accumulator a of the next row tile starts when group a of the
// epilogue warps (two groups of four warps, one per accumulator, draining in parallel) has released it, which
// exposes about one drain minus one slab of MMAs per row tile.  Both groups leave their exact integer plane sums in
// shared memory; group 0 adds the planes in the same order as the other variants (same bits) and finishes the pair.
constexpr int I8Q_NS = 4;
constexpr int I8Q_THREADS = 320;                                  // producer, issuer, 2 x 4 epilogue warps
constexpr int I8Q_STAGE_BYTES = I8_A_BYTES + 2 * I8P_HALF * I8_SLAB;   // 48 KB per CTA
struct I8QBars {
    uint64_t full[I8Q_NS], empty[I8Q_NS], tfull[2], tempty[2];
    uint32_t tmem_base;
    long long handoff[2][I8_MAX_PLANES][I8_TJ];                  // [item parity][plane][TMEM lane]: exact plane sums
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(I8Q_THREADS, 1)
i8_scan_pair2_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_ph,
                     const I8Params p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    I8QBars *bars = reinterpret_cast<I8QBars *>(smem + I8Q_NS * I8Q_STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    const int n_sp = p.n_planes >> 1;                             // plane pairs (n_planes is even)

    if (threadIdx.x == 0) {
        for (int s = 0; s < I8Q_NS; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&bars->tfull[b], 1); mbar_init(&bars->tempty[b], 8); }
        mbar_fence_init();
    }
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

    if (warp == 0) {
        // ================================ TMA producer (both CTAs; converged warp, elected lane issues) ====
        int stage = 0; unsigned phase = 0;
        for (int item = cluster_id; item < p.total_items; item += n_clusters) {
            const int ii = i8_find_first(p.item_prefix, p.n_i, item);
            const int jt = 2 * (p.jt0[ii] + (item - p.item_prefix[ii])) + rank;
            for (int sp = 0; sp < n_sp; ++sp) {
                const int prow = (ii * p.n_planes + 2 * sp) * p.np;         // plane 2sp; plane 2sp+1 is np rows on
                for (int kt = 0; kt < p.n_kt; ++kt) {
                    const int nl = i8_slabs_of(kt, p.n_slab);
                    for (int l = 0; l < nl; ++l) {
                        int row_off, nn;
                        i8p_shape(p, kt, l, row_off, nn);
                        const int brow = prow + kt * I8_TK + row_off + rank * (nn >> 1);
                        mbar_wait(&bars->empty[stage], phase ^ 1u);
                        if (elect_one()) {
                            if (rank == 0) mbar_expect_tx(&bars->full[stage], 2 * I8Q_STAGE_BYTES);
                            unsigned char *sa = smem + stage * I8Q_STAGE_BYTES;
                            tma_load_2d_pair(sa, &map_g, l * I8_SLAB, jt * I8_TJ, &bars->full[stage]);
                            tma_load_2d_pair(sa + I8_A_BYTES, &map_ph, l * I8_SLAB, brow, &bars->full[stage]);
                            tma_load_2d_pair(sa + I8_A_BYTES + I8P_HALF * I8_SLAB, &map_ph, l * I8_SLAB, brow + p.np,
                                             &bars->full[stage]);
                        }
                        if (++stage == I8Q_NS) { stage = 0; phase ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader CTA; converged warp, elected lane issues) =====
        if (rank == 0) {
            int stage = 0; unsigned phase = 0;
            unsigned aphase = 0;                                  // both accumulators are used once per row tile
            for (int item = cluster_id; item < p.total_items; item += n_clusters) {
                for (int sp = 0; sp < n_sp; ++sp) {
                    for (int kt = 0; kt < p.n_kt; ++kt) {
                        const int nl = i8_slabs_of(kt, p.n_slab);
                        for (int l = 0; l < nl; ++l) {
                            mbar_wait(&bars->full[stage], phase);
                            tc_fence_after();
                            const uint32_t sa = smem_base + (uint32_t)(stage * I8Q_STAGE_BYTES);
                            const uint64_t da = umma_desc_sw128(sa);
                            int row_off, nn;
                            i8p_shape(p, kt, l, row_off, nn);
                            const uint32_t idesc = i8p_idesc(nn);
#pragma unroll
                            for (int a = 0; a < 2; ++a) {
                                if (l == 0) {                     // epilogue group a drained accumulator a
                                    mbar_wait(&bars->tempty[a], aphase ^ 1u);
                                    tc_fence_after();
                                }
                                const uint64_t db = umma_desc_sw128(sa + I8_A_BYTES + a * I8P_HALF * I8_SLAB);
                                const uint32_t tmem_d = tmem_base + (uint32_t)(a * I8_TK + row_off);
                                if (elect_one()) {
#pragma unroll
                                    for (int k = 0; k < I8_SLAB / 32; ++k)
                                        umma_i8_pair(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                                                     (uint32_t)((l | k) != 0));
                                    if (l == nl - 1) tc_commit_pair(&bars->tfull[a]);
                                    if (a == 1) tc_commit_pair(&bars->empty[stage]);
                                }
                            }
                            if (++stage == I8Q_NS) { stage = 0; phase ^= 1u; }
                        }
                        aphase ^= 1u;
                    }
                }
            }
        }
    } else {
        // ================================ epilogue: group a = (warp - 2) / 4 drains accumulator a ========
        const int q = warp & 3;                                   // TMEM lane quarter this warp may read
        const int grp = (warp - 2) >> 2;
        unsigned aphase = 0;
        int parity = 0;
        for (int item = cluster_id; item < p.total_items; item += n_clusters, parity ^= 1) {
            const int ii = i8_find_first(p.item_prefix, p.n_i, item);
            const int jt = 2 * (p.jt0[ii] + (item - p.item_prefix[ii])) + rank;
            const int j = jt * I8_TJ + q * 32 + lane;
            const signed char *arow = p.g8 + (size_t)(j < p.m ? j : p.m - 1) * p.ld;
            for (int sp = 0; sp < n_sp; ++sp) {
                long long plane_sum = 0;                          // exact; plane 2sp + grp
                for (int kt = 0; kt < p.n_kt; ++kt) {
                    int4 abytes[I8_TK / 16];
                    i8_load_dosage(arow + kt * I8_TK, abytes);
                    mbar_wait_warp(&bars->tfull[grp], aphase);
                    tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * I8_TK);
                    plane_sum += i8_drain_tile(taddr, abytes);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_leader(&bars->tempty[grp]);
                    aphase ^= 1u;
                }
                bars->handoff[parity][2 * sp + grp][q * 32 + lane] = plane_sum;
            }
            named_bar_sync(1, 256);                               // the eight epilogue warps of this CTA
            if (grp == 0) {
                double qacc = 0.0, wgt = 1.0;
                for (int s = 0; s < p.n_planes; ++s) {            // planes in the order 0, 1, 2, ... like the other variants
                    wgt *= 0.0078125;                             // 128^-(s+1)
                    qacc += (double)bars->handoff[parity][s][q * 32 + lane] * wgt;
                }
                i8_finish_pair(p, ii, j, lane, qacc);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// Issue-rate microbenchmark of the int8 MMA (the roofline denominator of the engine): one thread per CTA issues
// back-to-back 128x256x32 MMAs on a resident (pseudo-random) smem tile pair, alternating two TMEM accumulators.
__global__ void __launch_bounds__(128, 1) i8_peak_kernel(int iters) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    for (int k = threadIdx.x; k < I8_STAGE_BYTES / 4; k += blockDim.x)
        reinterpret_cast<unsigned *>(smem)[k] = (unsigned)(k * 2654435761u) ^ (blockIdx.x * 40503u);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&tbase_slot)),
                     "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tbase_slot;
    if (threadIdx.x == 0) {
        const uint32_t sa = smem_u32(smem);
        const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + I8_A_BYTES);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_i8(tbase + (uint32_t)((it & 1) * I8_TK), da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), I8_IDESC,
                        (uint32_t)(it > 1));
        }
        tc_commit(&bar);
        mbar_wait(&bar, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tbase), "r"(512) : "memory");
}

// ------------------------------------------------------------------------------------------ operand builders
// integer code of the second SNP: additive dosage {0,miss->0,1,2}; dominance indicator {0,miss->0,1,0}
__global__ void i8_dosage_kernel(const unsigned char *__restrict__ bed, int n, long long m, long long row_bytes,
                                 int coding, signed char *__restrict__ g8, int ld) {
    const long long groups = ld / 16;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m * groups) return;
    const long long snp = t / groups;
    const int k0 = (int)(t % groups) * 16;
    const unsigned tab = coding == GMAT_B200_CODING_DOM ? 0x00010000u : 0x02010000u;   // byte[code]
    unsigned w = 0;
    if (k0 / 4 + 3 < row_bytes) w = *reinterpret_cast<const unsigned *>(bed + snp * row_bytes + k0 / 4);
    unsigned out[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        unsigned o = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int k = k0 + g * 4 + c;
            const unsigned code = (w >> (2 * (g * 4 + c))) & 3u;
            const unsigned val = k < n ? ((tab >> (8 * code)) & 0xffu) : 0u;
            o |= val << (8 * c);
        }
        out[g] = o;
    }
    *reinterpret_cast<uint4 *>(g8 + snp * ld + k0) = make_uint4(out[0], out[1], out[2], out[3]);
}

// B_i = diag(u_i) A diag(u_i) / scale_i -> S signed base-128 digits.  Lower block-triangle of 128 x 128 blocks,
// diagonal blocks halved (the kernel doubles the sum); everything else in the plane buffer stays zero.
template <int S>
__global__ void __launch_bounds__(256)
i8_split_kernel(const double *__restrict__ A, int n, const double *__restrict__ U, int n_i, int i_per_block,
                const double *__restrict__ scale, int np, int ld, signed char *__restrict__ planes) {
    int pk = 0, rem = blockIdx.x;                       // blockIdx.x -> (pk, pl), pl <= pk
    while (rem > pk) { rem -= pk + 1; ++pk; }
    const int pl = rem;
    const int cg = threadIdx.x & 7, r0 = threadIdx.x >> 3;
    const int l0 = pl * 128 + cg * 16;
    const double half = pk == pl ? 0.5 : 1.0;
    const double big = 72057594037927936.0;             // 128^8 = 2^56
    double top = 1.0;
    for (int s = S; s < 8; ++s) top *= 128.0;
    const double mul = half * (big / top);              // 128^S (x 1/2 on the diagonal blocks)
    const int i_begin = blockIdx.y * i_per_block;
    const int i_end = min(n_i, i_begin + i_per_block);
    for (int pass = 0; pass < 4; ++pass) {
        const int k = pk * 128 + pass * 32 + r0;
        double a[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) a[c] = (k < n && l0 + c < n) ? A[(size_t)k * n + l0 + c] : 0.0;
        for (int ii = i_begin; ii < i_end; ++ii) {
            const double *u = U + (size_t)ii * n;
            const double uk = k < n ? u[k] * (mul / scale[ii]) : 0.0;
            unsigned dg[S][4];
#pragma unroll
            for (int s = 0; s < S; ++s) dg[s][0] = dg[s][1] = dg[s][2] = dg[s][3] = 0u;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const double ul = l0 + c < n ? u[l0 + c] : 0.0;
                long long t = __double2ll_rn(uk * a[c] * ul);
#pragma unroll
                for (int s = S - 1; s >= 1; --s) {
                    const long long d = ((t + 64) & 127) - 64;          // balanced digit in [-64, 63]
                    t = (t - d) >> 7;
                    dg[s][c >> 2] |= ((unsigned)d & 0xffu) << (8 * (c & 3));
                }
                dg[0][c >> 2] |= ((unsigned)t & 0xffu) << (8 * (c & 3));  // |t| <= 127 by the choice of scale
            }
#pragma unroll
            for (int s = 0; s < S; ++s) {
                signed char *dst = planes + ((size_t)(ii * S + s) * np + k) * ld + l0;
                *reinterpret_cast<uint4 *>(dst) = make_uint4(dg[s][0], dg[s][1], dg[s][2], dg[s][3]);
            }
        }
    }
}

}  // namespace gb

using namespace gb;

extern "C" long long gmat_b200_i8_ld(long long n) { return round_up(n, I8_TK); }

extern "C" int gmat_b200_i8_dosage(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                   int coding, signed char *g8_dev, long long ld, void *stream) {
    GB_CHECK_ARG(bed_dev && g8_dev && n > 0 && m > 0, "null buffer or empty shape");
    GB_CHECK_ARG(ld == round_up(n, I8_TK) && row_bytes % 4 == 0, "ld must be gmat_b200_i8_ld(n)");
    GB_CHECK_ARG(coding == GMAT_B200_CODING_ADD || coding == GMAT_B200_CODING_DOM, "coding");
    const long long work = m * (ld / 16);
    i8_dosage_kernel<<<(unsigned)((work + 255) / 256), 256, 0, (cudaStream_t)stream>>>(bed_dev, (int)n, m, row_bytes,
                                                                                      coding, g8_dev, (int)ld);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_i8_split(const double *pvp_dev, long long n, const double *u_dev, long long n_i,
                                  const double *scale_dev, int n_planes, signed char *planes_dev, void *stream) {
    GB_CHECK_ARG(pvp_dev && u_dev && scale_dev && planes_dev && n > 0 && n_i > 0, "null buffer or empty shape");
    GB_CHECK_ARG(n_planes >= 2 && n_planes <= I8_MAX_PLANES, "n_planes must be in [2, 8]");
    const int ld = (int)round_up(n, I8_TK), np = ld;
    const int nb = (int)((n + 127) / 128);
    const int blocks = nb * (nb + 1) / 2;
    // enough CTAs for the chip; each keeps its 16 A values per thread across its share of the first SNPs
    int ysplit = (int)((2LL * sm_count() + blocks - 1) / blocks);
    if (ysplit > n_i) ysplit = (int)n_i;
    if (ysplit < 1) ysplit = 1;
    const int per = (int)((n_i + ysplit - 1) / ysplit);
    dim3 grid(blocks, (unsigned)((n_i + per - 1) / per));
    cudaStream_t st = (cudaStream_t)stream;
#define GB_SPLIT(SV) i8_split_kernel<SV><<<grid, 256, 0, st>>>(pvp_dev, (int)n, u_dev, (int)n_i, per, scale_dev, np, ld, planes_dev)
    switch (n_planes) {
        case 2: GB_SPLIT(2); break; case 3: GB_SPLIT(3); break; case 4: GB_SPLIT(4); break; case 5: GB_SPLIT(5); break;
        case 6: GB_SPLIT(6); break; case 7: GB_SPLIT(7); break; default: GB_SPLIT(8); break;
    }
#undef GB_SPLIT
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_i8_scan(const signed char *g8_dev, long long n, long long m, const signed char *planes_dev,
                                 int n_planes, const long long *ilist_dev, long long n_i,
                                 const int *item_prefix_dev, const int *jt0_dev, long long total_items,
                                 const double *scale_dev, const double *s_dev, const double *upy_dev,
                                 const double *t1_dev, const double *e_dev, const double *offs_dev,
                                 int pair_space, double chi_cut, gmat_b200_hit *hits_dev, long long cap,
                                 unsigned long long *count_dev, double *q_debug_dev, int wide, void *stream) {
    GB_CHECK_ARG(g8_dev && planes_dev && ilist_dev && item_prefix_dev && jt0_dev && scale_dev && s_dev && upy_dev &&
                 t1_dev && e_dev && offs_dev && count_dev, "null buffer");
    GB_CHECK_ARG(n > 0 && m > 0 && n_i > 0 && total_items >= 0 && total_items < (1LL << 31), "shape");
    GB_CHECK_ARG(n_planes >= 2 && n_planes <= I8_MAX_PLANES, "n_planes must be in [2, 8]");
    GB_CHECK_ARG(pair_space == GMAT_B200_PAIRS_UPPER || pair_space == GMAT_B200_PAIRS_ALL, "pair_space");
    GB_CHECK_ARG(cap == 0 || hits_dev, "hits buffer");
    if (total_items == 0) return GMAT_B200_OK;
    cudaStream_t st = (cudaStream_t)stream;
    I8Params p;
    p.n = (int)n; p.m = (int)m; p.ld = (int)round_up(n, I8_TK); p.np = p.ld;
    p.n_planes = n_planes; p.n_kt = p.ld / I8_TK; p.n_slab = (int)((n + I8_SLAB - 1) / I8_SLAB);
    p.n_i = (int)n_i; p.pair_space = pair_space;
    GB_CHECK_ARG((long long)n_i * n_planes * p.np < (1LL << 31), "plane buffer has too many rows for one tensor map");
    p.total_items = (int)total_items; p.item_prefix = item_prefix_dev; p.jt0 = jt0_dev;
    p.ilist = ilist_dev; p.g8 = g8_dev; p.scale = scale_dev; p.s_i = s_dev; p.upy = upy_dev; p.t1 = t1_dev; p.e = e_dev;
    p.offs = offs_dev; p.chi_cut = chi_cut; p.hits = hits_dev; p.cap = cap; p.count = count_dev; p.q_debug = q_debug_dev;

    CUtensorMap map_g, map_p;
    int rc = make_map(&map_g, g8_dev, p.ld, m, p.ld, I8_TJ);
    if (rc == GMAT_B200_OK) rc = make_map(&map_p, planes_dev, p.ld, (long long)n_i * n_planes * p.np, p.ld, I8_TK);
    if (rc != GMAT_B200_OK) return rc;
    const int grid = (int)(total_items < sm_count() ? total_items : sm_count());
    GB_CHECK_ARG(wide == 0 || wide == 2 || wide == 3 || wide == 4,
                 "variant must be 0 (one CTA per 128-SNP tile), 2 (CTA pair), 3 (CTA pair, trimmed MMAs) or 4 (CTA "
                 "pair, trimmed MMAs, two digit planes per genotype tile)");
    GB_CHECK_ARG(wide != 4 || n_planes % 2 == 0, "variant 4 multiplies the digit planes two at a time: n_planes must be even");
    p.trim = wide >= 3;
    if (wide == 4) {
        CUtensorMap map_ph;
        rc = make_map(&map_ph, planes_dev, p.ld, (long long)n_i * n_planes * p.np, p.ld, I8P_HALF);
        if (rc != GMAT_B200_OK) return rc;
        const int smem_bytes = I8Q_NS * I8Q_STAGE_BYTES + 1024 + (int)sizeof(I8QBars);
        GB_CUDA(cudaFuncSetAttribute(i8_scan_pair2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        const int pairs_of_sms = sm_count() / 2;
        const int grid2 = 2 * (int)(total_items < pairs_of_sms ? total_items : pairs_of_sms);
        i8_scan_pair2_kernel<<<grid2, I8Q_THREADS, smem_bytes, st>>>(map_g, map_ph, p);
    } else if (wide >= 2) {
        CUtensorMap map_ph;                                       // plane map with 128-row boxes (one half per CTA)
        rc = make_map(&map_ph, planes_dev, p.ld, (long long)n_i * n_planes * p.np, p.ld, I8P_HALF);
        if (rc != GMAT_B200_OK) return rc;
        const int smem_bytes = I8P_NS * I8P_STAGE_BYTES + 1024 + (int)sizeof(I8PBars);
        GB_CUDA(cudaFuncSetAttribute(i8_scan_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        const int pairs_of_sms = sm_count() / 2;
        const int grid2 = 2 * (int)(total_items < pairs_of_sms ? total_items : pairs_of_sms);
        i8_scan_pair_kernel<<<grid2, I8_THREADS, smem_bytes, st>>>(map_g, map_ph, p);
    } else {
        const int smem_bytes = I8_NS * I8_STAGE_BYTES + 1024 + (int)sizeof(I8Bars);
        GB_CUDA(cudaFuncSetAttribute(i8_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        i8_scan_kernel<<<grid, I8_THREADS, smem_bytes, st>>>(map_g, map_p, p);
    }
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_i8_peak(int iters, double *tops) {
    GB_CHECK_ARG(iters > 0 && tops, "iters/tops");
    const int smem_bytes = I8_STAGE_BYTES + 1024;
    GB_CUDA(cudaFuncSetAttribute(i8_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    cudaEvent_t e0, e1;
    GB_CUDA(cudaEventCreate(&e0));
    GB_CUDA(cudaEventCreate(&e1));
    const int grid = sm_count();
    i8_peak_kernel<<<grid, 128, smem_bytes>>>(iters / 8 + 1);          // warm-up
    GB_CUDA(cudaEventRecord(e0));
    i8_peak_kernel<<<grid, 128, smem_bytes>>>(iters);
    GB_CUDA(cudaEventRecord(e1));
    GB_CUDA(cudaEventSynchronize(e1));
    GB_LAUNCH_CHECK();
    float ms = 0.f;
    GB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tops = 2.0 * grid * (double)iters * 4.0 * I8_TJ * I8_TK * 32.0 / (ms * 1e-3) / 1e12;
    return GMAT_B200_OK;
}
