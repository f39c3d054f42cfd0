// INT8 tensor-core engine of the exact epistasis test (sm_100a: tcgen05.mma kind::i8, TMEM, TMA).
//
// The exact test of the reference (gmat/remma/remma_epiAA.py:72-77) needs, per SNP pair, var = w' A w with
// A = Z'PZ (n x n, FP64) and w = u_i o v_j.  With v_j = a_j - o_j (a_j the integer dosage / heterozygote
// indicator, o_j = 2p_j or 2p_j q_j) and B_i = diag(u_i) A diag(u_i):
//
//     var_ij = a_j' B_i a_j  -  2 o_j a_j.r_i  +  o_j^2 s_i,      r_i = u_i o (A u_i),  s_i = sum(r_i)
//     eff_ij = a_j.(u_i o Z'Py) - o_j (u_i . Z'Py)
//
// Only a_j' B_i a_j costs n^2 per pair, and a_j is an exact small-integer vector.  B_i / scale_i is split error-free
// into S signed base-256 digit planes (int8: every digit in [-128, 127]); "plane x genotype" products are exact
// int8 x int8 -> int32 tensor-core GEMMs, recombined in FP64.  S planes carry 8S - 1 bits of max|B_i| (S = 6: 47
// bits, S = 7: 55 bits -- finer than the FP64 rounding of B_i itself).  The two O(n)-per-pair terms a_j.r_i and
// a_j.(u_i o Z'Py) ride in the SAME tensor-core stream: r_i and u_i o Z'Py are split into digits as well (a high and
// a low part, 2S digits, so they lose nothing against FP64) and stored as four extra rows of every plane, below the
// n rows of B_i; their integer dot products with a_j come out of the accumulator as four extra TMEM columns.
// Nothing on this path is a library call: decode of the first SNPs, A u_i, the reductions, the digit split and the
// scan are the kernels of this file (remma_epiAA.py:72-74 is what they replace).
//
// The FP64 DMMA kernel (exact.cu) issues 2 n^2 FP64 flop per pair at 37 TFLOP/s; this path issues S n^2 / 2
// int8 MACs per pair (block-lower-triangle of the symmetric B_i, diagonal blocks halved, result doubled) on a
// pipe that is two orders of magnitude faster.
//
// A-posteriori guard: the split drops at most half a unit of the last digit per element of B_i, i.e. a random-walk
// error of ~0.41 * 256^-S * scale_i * sum_k a_jk^2 on a_j' B_i a_j.  Every pair checks 4 * 256^-S * scale_i * sum a^2
// (ten standard deviations) against recheck_tol * |var_ij|; a pair that fails (a variance that is tiny against
// max|u_i|^2 max|A|: two very rare variants) is not decided here but appended to a re-check list, which the caller
// evaluates with the FP64 pair kernel (exact.cu).  So the number of planes is a speed knob, not an accuracy risk.
#include "tc5.cuh"

namespace gb {

constexpr int I8_THREADS = 192;
constexpr int I8_MAX_PLANES = 7;          // 8 base-256 digits would not fit the int64 the split rounds into
constexpr int I8_SIDE = 8;                // doubles per first SNP in the side table (see i8_side_kernel)
constexpr int I8_XROWS = 4;               // extra plane rows: r hi, r lo, (u o py) hi, (u o py) lo

__host__ __device__ inline long long i8_xrow(long long n) { return round_up(n, 4); }            // first extra row
__host__ __device__ inline long long i8_rows(long long n) { return round_up(i8_xrow(n) + I8_XROWS, I8_TK); }

struct I8Params {
    int n, m, ld, np, n_planes, n_kt, n_slab;
    int n_i, total_items, pair_space;
    int xkt, xcol;               // row tile and TMEM column of the four extra rows
    const long long *ilist;
    const int *item_prefix;      // [n_i+1] first work item of every first SNP
    const int *jt0;              // [n_i]   first j tile of every first SNP
    const signed char *g8;       // [m][ld] integer coded second SNPs (the MMA operand, read by TMA)
    const unsigned char *g2;     // [m][ld/4] the same values packed 2 bits each (the epilogue's copy: 16 registers per 256 ids)
    const double *side;          // [n_i][I8_SIDE]
    const double *offs;          // [m]   o_j
    const double *asq;           // [m]   sum_k a_jk^2
    double chi_cut, wlo, recheck_tol;
    gmat_b200_hit *hits;
    long long cap;
    unsigned long long *count;
    int *recheck;                // [recheck_cap][2] pairs the split cannot decide (see header comment)
    long long recheck_cap;
    unsigned long long *recheck_count;
    double *q_debug;             // optional [n_i][m]: a_j' B_i a_j
    double *e_debug;             // optional [n_i][m]: eff_ij (the rows form of the pair-list engine reads both)
    int trim;                    // issue only the plane rows that can be non-zero (i8p_shape)
};

// Epilogue of one accumulator tile for one TMEM lane (= one second SNP): sum_k a_jk * Y[j][k] over the 256 plane
// rows of the tile, all in INTEGER registers.  No FP64 instruction may sit in this loop: DADD / I2F.F64 issue into
// the pipe the tensor MMAs occupy (ncu: 37 % of all stall samples sat on the one DADD of the first version and the
// tensor pipe idled 37 % of the time).  |Y| <= 2*128*n, so a 32-column partial fits int32 for n < 1.3e5.
__device__ __forceinline__ int i8_dot_chunk(const int (&v)[32], unsigned w0, unsigned w1) {
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;                           // four independent chains
#pragma unroll
    for (int q = 0; q < 16; q += 4) {
        c0 += (int)((w0 >> (2 * q)) & 3u) * v[q];
        c1 += (int)((w0 >> (2 * q + 2)) & 3u) * v[q + 1];
        c2 += (int)((w0 >> (2 * q + 4)) & 3u) * v[q + 2];
        c3 += (int)((w0 >> (2 * q + 6)) & 3u) * v[q + 3];
    }
#pragma unroll
    for (int q = 0; q < 16; q += 4) {
        c0 += (int)((w1 >> (2 * q)) & 3u) * v[16 + q];
        c1 += (int)((w1 >> (2 * q + 2)) & 3u) * v[17 + q];
        c2 += (int)((w1 >> (2 * q + 4)) & 3u) * v[18 + q];
        c3 += (int)((w1 >> (2 * q + 6)) & 3u) * v[19 + q];
    }
    return (c0 + c1) + (c2 + c3);                                 // |.| <= 32 * 2 * 2*128*n: int32 for n < 1.3e5
}
// The TMEM load of chunk ch+1 is in flight while chunk ch is multiplied (two register buffers).  a[] = the 256
// integer codes of this lane's SNP for the row tile, 2 bits each (16 per word).
__device__ __forceinline__ long long i8_drain_tile(uint32_t taddr, const uint4 (&a)[I8_TK / 64]) {
    const unsigned aw[16] = {a[0].x, a[0].y, a[0].z, a[0].w, a[1].x, a[1].y, a[1].z, a[1].w,
                             a[2].x, a[2].y, a[2].z, a[2].w, a[3].x, a[3].y, a[3].z, a[3].w};
    long long total = 0;
    int v0[32], v1[32];
    tmem_ld32(taddr, v0);
#pragma unroll
    for (int ch = 0; ch < I8_TK / 32; ch += 2) {
        tmem_ld_wait();
        tmem_ld32(taddr + (ch + 1) * 32, v1);
        total += i8_dot_chunk(v0, aw[2 * ch], aw[2 * ch + 1]);
        tmem_ld_wait();
        if (ch + 2 < I8_TK / 32) tmem_ld32(taddr + (ch + 2) * 32, v0);
        total += i8_dot_chunk(v1, aw[2 * ch + 2], aw[2 * ch + 3]);
    }
    return total;
}
// the 256 codes (64 bytes) of this lane's SNP for one row tile: issued BEFORE the wait on the accumulator, so the L2
// latency (the first version prefetched one chunk ahead and stalled on it: 23 % of all samples) is hidden.  Packed
// 2 bits per code they take 16 registers; as bytes (the MMA operand's layout) they took 64 and pushed the kernel to
// 247 registers -- two such warps fill an SM sub-partition's register file, and no operand kernel of the next batch
// could co-reside with the scan.
__device__ __forceinline__ void i8_load_dosage(const unsigned char *row, uint4 (&a)[I8_TK / 64]) {
    const uint4 *ap = reinterpret_cast<const uint4 *>(row);
#pragma unroll
    for (int q = 0; q < I8_TK / 64; ++q) a[q] = __ldg(ap + q);
}

// The same for operand rows whose integer values do not fit two bits (the pair-list engine below: sums of two
// genotype vectors, values 0..6): codes packed FOUR bits each, 32 registers per 256 ids.
__device__ __forceinline__ int i8_dot_chunk4(const int (&v)[32], unsigned w0, unsigned w1, unsigned w2, unsigned w3) {
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    const unsigned w[4] = {w0, w1, w2, w3};
#pragma unroll
    for (int g = 0; g < 4; ++g) {
#pragma unroll
        for (int q = 0; q < 8; q += 4) {
            c0 += (int)((w[g] >> (4 * q)) & 15u) * v[8 * g + q];
            c1 += (int)((w[g] >> (4 * q + 4)) & 15u) * v[8 * g + q + 1];
            c2 += (int)((w[g] >> (4 * q + 8)) & 15u) * v[8 * g + q + 2];
            c3 += (int)((w[g] >> (4 * q + 12)) & 15u) * v[8 * g + q + 3];
        }
    }
    return (c0 + c1) + (c2 + c3);                // rows hold values <= 6: |.| <= 32 * 6 * (6*128*n), int32 for n < 1.45e4
}
__device__ __forceinline__ long long i8_drain_tile4(uint32_t taddr, const uint4 (&a)[I8_TK / 32]) {
    long long total = 0;
    int v0[32], v1[32];
    tmem_ld32(taddr, v0);
#pragma unroll
    for (int ch = 0; ch < I8_TK / 32; ch += 2) {
        tmem_ld_wait();
        tmem_ld32(taddr + (ch + 1) * 32, v1);
        total += i8_dot_chunk4(v0, a[ch].x, a[ch].y, a[ch].z, a[ch].w);
        tmem_ld_wait();
        if (ch + 2 < I8_TK / 32) tmem_ld32(taddr + (ch + 2) * 32, v0);
        total += i8_dot_chunk4(v1, a[ch + 1].x, a[ch + 1].y, a[ch + 1].z, a[ch + 1].w);
    }
    return total;
}
__device__ __forceinline__ void i8_load_dosage4(const unsigned char *row, uint4 (&a)[I8_TK / 32]) {
    const uint4 *ap = reinterpret_cast<const uint4 *>(row);
#pragma unroll
    for (int q = 0; q < I8_TK / 32; ++q) a[q] = __ldg(ap + q);
}
template <int BITS> struct I8Codes;
template <> struct I8Codes<2> {
    uint4 a[I8_TK / 64];
    __device__ __forceinline__ void load(const unsigned char *row) { i8_load_dosage(row, a); }
    __device__ __forceinline__ long long drain(uint32_t taddr) const { return i8_drain_tile(taddr, a); }
};
template <> struct I8Codes<4> {
    uint4 a[I8_TK / 32];
    __device__ __forceinline__ void load(const unsigned char *row) { i8_load_dosage4(row, a); }
    __device__ __forceinline__ long long drain(uint32_t taddr) const { return i8_drain_tile4(taddr, a); }
};

// Finish one pair (first SNP ii of the launch, second SNP j).  qacc = sum_s 256^-(s+1) sum_k a_jk Y_s[j][k];
// t1acc / eacc = the digit sums of a_j.r_i and a_j.(u_i o py) in units of their scales.  The FP64 side terms, chi,
// the a-posteriori guard, threshold, warp-ballot compaction, 32-byte hit record (remma_epiAA.py:75-82).  Called by all
// 32 lanes of an epilogue warp.
__device__ __forceinline__ void i8_finish_pair(const I8Params &p, int ii, int j, int lane, double qacc, double t1acc,
                                               double eacc) {
    const int i = (int)p.ilist[ii];
    const bool valid = j < p.m && (p.pair_space == GMAT_B200_PAIRS_ALL || j > i);
    const int jc = j < p.m ? j : p.m - 1;
    const double *sd = p.side + (size_t)ii * I8_SIDE;
    const double quad = 2.0 * sd[0] * qacc;                         // a_j' B_i a_j
    const double o = p.offs[jc];
    const double asq = p.asq[jc];                                   // sign bit: the SNP is coded by 2 - a
    double var = quad - 2.0 * o * (sd[3] * t1acc) + o * o * sd[1];
    double eff = sd[4] * eacc - o * sd[2];
    if (signbit(asq)) eff = -eff;                                   // w' = -w: var is the same, eff changes sign
    bool plain = true;
    if (o != o) { var = 0.0; eff = 0.0; plain = false; }           // monomorphic second SNP: w == 0 exactly
    else if (o > 1.0e300) { var = o - o; eff = var; plain = false; }   // missing genotype in the second SNP: NaN, as numpy
    const double chi = eff * eff / var;
    if (p.q_debug != nullptr && j < p.m) p.q_debug[(size_t)ii * p.m + jc] = quad;
    if (p.e_debug != nullptr && j < p.m) p.e_debug[(size_t)ii * p.m + jc] = eff;
    // what the split may have dropped (ten standard deviations) against the variance itself; NaN compares false
    const bool weak = valid && plain && sd[5] * fabs(asq) > p.recheck_tol * fabs(var);
    gmat_b200_hit h;
    h.snp_0 = i; h.snp_1 = j; h.eff = eff; h.var = var; h.chi = chi;
    const bool hit = valid && !weak && (p.chi_cut < 0.0 || chi > p.chi_cut);
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
    const unsigned wballot = __ballot_sync(0xffffffffu, weak);
    if (wballot) {                                                  // rare: two very rare variants
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(p.recheck_count, (unsigned long long)__popc(wballot));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (weak) {
            const unsigned long long pos = base + __popc(wballot & ((1u << lane) - 1u));
            if ((long long)pos < p.recheck_cap) { p.recheck[2 * pos] = i; p.recheck[2 * pos + 1] = j; }
        }
    }
}

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

// ---- the scan kernel: CTA pairs, tcgen05.mma.cta_group::2 -----------------------------------------------------
// Persistent CTA pairs (cluster of 2 = one TPC), 192 threads per CTA: warp 0 = TMA producer, warp 1 of the leader =
// tcgen05.mma issuer (one elected lane), warps 2..5 = epilogue (TMEM -> registers -> a_jk-weighted row sums, kept per
// thread: the TMEM lane IS the pair, so no cross-lane reduction exists anywhere).  A work item is (first SNP, 256
// second SNPs); the two CTAs own the two 128-SNP tiles of the item and HALF of every plane tile each (128 of its 256
// rows); one MMA of the leader CTA computes M = 256 SNPs x N = 256 plane rows x K = 32 ids across both SMs, each
// tensor core reading its own genotype tile and the two plane halves.  Accumulators are double-buffered in TMEM
// (2 x 256 columns per CTA).
constexpr int I8P_NS = 6;
constexpr int I8P_HALF = I8_TK / 2;                               // plane rows each CTA loads
constexpr int I8P_STAGE_BYTES = I8_A_BYTES + I8P_HALF * I8_SLAB;  // 32 KB per CTA
struct I8PBars {
    uint64_t full[I8P_NS], empty[I8P_NS], tfull[2], tempty[2];
    uint32_t tmem_base;
};

// Plane rows of tile kt that slab l (128 ids) can touch: MMA N = nn rows starting row_off rows into the tile.
//  * the last tile holds only the rows up to the four extra rows (rounded up to the UMMA N granularity of 16);
//  * slab 2kt+1 lies right of the diagonal for the tile's first 128 rows (the split kernel leaves that 128 x 128
//    block zero), so only rows 128.. of the tile are multiplied and accumulated into TMEM columns 128.. .  (The
//    extra rows need every slab: when slab 2kt+1 exists in their tile, n > 256 kt + 128, so they sit in rows 128.. .)
// The two CTAs supply nn/2 rows each (tcgen05 cta_group::2 takes the first half of N from the leader's shared
// memory, the second from the peer's), so CTA r loads the 128-row box starting at row_off + r nn/2 and the MMA reads
// its first nn/2 rows.  TMEM columns that a trimmed MMA does not write keep stale integers; the epilogue multiplies
// them by dosage bytes of ids >= n, which are zero.  n=2000: 65.3 instead of 72 slab-MMAs per plane (-9 %).
__device__ __forceinline__ void i8p_shape(const I8Params &p, int kt, int l, int &row_off, int &nn) {
    row_off = 0; nn = I8_TK;
    if (!p.trim) return;
    const int rows = p.xkt * I8_TK + p.xcol + I8_XROWS - kt * I8_TK;
    if (rows < I8_TK) nn = (rows + 15) & ~15;
    if (l == 2 * kt + 1) { row_off = I8_TK / 2; nn -= I8_TK / 2; }
}

// BITS = width of the epilogue's packed copy of the operand rows (p.g2): 2 for genotype codes, 4 for the integer rows of
// the pair-list engine.
template <int BITS>
__device__ __forceinline__ void i8_scan_pair_body(const CUtensorMap &map_g, const CUtensorMap &map_ph, const I8Params &p) {
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
            const unsigned char *arow = p.g2 + (size_t)(j < p.m ? j : p.m - 1) * (size_t)(p.ld * BITS / 8);
            double qacc = 0.0, wgt = 1.0;
            double rh = 0.0, rl = 0.0, eh = 0.0, el = 0.0;                  // digit sums of the four extra rows
            for (int s = 0; s < p.n_planes; ++s) {
                wgt *= 0.00390625;                                          // 256^-(s+1)
                long long plane_sum = 0;                                    // exact
                int x[4] = {0, 0, 0, 0};
                for (int kt = 0; kt < p.n_kt; ++kt) {
                    I8Codes<BITS> codes;
                    codes.load(arow + kt * (I8_TK * BITS / 8));
                    mbar_wait_warp(&bars->tfull[abuf], aphase);
                    tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(abuf * I8_TK);
                    plane_sum += codes.drain(taddr);
                    if (kt == p.xkt) {                                      // a_j . digits of r_i and of u_i o py
                        tmem_ld4(taddr + (uint32_t)p.xcol, x);
                        tmem_ld_wait();
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_leader(&bars->tempty[abuf]);
                    if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
                }
                qacc += (double)plane_sum * wgt;                         // the only FP64 work per plane
                rh += (double)x[0] * wgt; rl += (double)x[1] * wgt;
                eh += (double)x[2] * wgt; el += (double)x[3] * wgt;
            }
            i8_finish_pair(p, ii, j, lane, qacc, rh + p.wlo * rl, eh + p.wlo * el);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                       // nobody frees TMEM / exits while the peer may still touch this CTA
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// (192 registers: 2 x 192 x 32 leave 4096 of a sub-partition's 16384 for a warp of an operand kernel of the next batch)
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(192)
i8_scan_pair_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_ph,
                    const I8Params p) {
    i8_scan_pair_body<2>(map_g, map_ph, p);
}
// the same kernel over integer operand rows with 4-bit epilogue codes (pair-list engine; no operand kernel runs beside it)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(I8_THREADS, 1)
i8_scan_rows_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_ph,
                    const I8Params p) {
    i8_scan_pair_body<4>(map_g, map_ph, p);
}

// ------------------------------------------------------------------------------------------ operand builders
// integer code of the second SNP: additive dosage {0,miss->0,1,2}; dominance indicator {0,miss->0,1,0}
// With genotype counts (the exact engine, additive coding) a SNP whose counted allele is the MAJOR one (more 11 than 00
// homozygotes) is coded by the other allele, a' = 2 - a: w_ij only changes sign (v' = a' - (2 - o) = -v), but
// sum_k a'_jk^2 -- what the error of the digit-plane split scales with -- is up to 35x smaller (p = 0.95: 3.7 n -> 0.1 n).
// i8_offs_kernel applies the same rule to o_j and sum a^2 and marks the SNP (sign of asq) so that eff gets its sign back.
__device__ __forceinline__ bool i8_flipped(const int *__restrict__ counts, long long j, int coding_additive) {
    return coding_additive && counts[4 * j + 3] > counts[4 * j];
}
constexpr unsigned I8_TAB_ADD = 0x02010000u, I8_TAB_DOM = 0x00010000u, I8_TAB_ADD_FLIPPED = 0x00010002u;

__global__ void i8_dosage_kernel(const unsigned char *__restrict__ bed, int n, long long m, long long row_bytes,
                                 unsigned tab, const int *__restrict__ counts, signed char *__restrict__ g8,
                                 unsigned char *__restrict__ g2, int ld) {
    const long long groups = ld / 16;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m * groups) return;
    const long long snp = t / groups;
    if (counts != nullptr && i8_flipped(counts, snp, tab == I8_TAB_ADD)) tab = I8_TAB_ADD_FLIPPED;
    const int k0 = (int)(t % groups) * 16;
    unsigned w = 0;
    if (k0 / 4 + 3 < row_bytes) w = *reinterpret_cast<const unsigned *>(bed + snp * row_bytes + k0 / 4);
    unsigned out[4], packed = 0;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        unsigned o = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int k = k0 + g * 4 + c;
            const unsigned code = (w >> (2 * (g * 4 + c))) & 3u;
            const unsigned val = k < n ? ((tab >> (8 * code)) & 0xffu) : 0u;
            o |= val << (8 * c);
            packed |= val << (2 * (g * 4 + c));
        }
        out[g] = o;
    }
    *reinterpret_cast<uint4 *>(g8 + snp * ld + k0) = make_uint4(out[0], out[1], out[2], out[3]);
    if (g2 != nullptr) *reinterpret_cast<unsigned *>(g2 + snp * (ld >> 2) + (k0 >> 2)) = packed;
}

// Per second SNP: o_j = -lut[j][code 0] with the two flags of the engine, and sum_k a_jk^2 (from the exact genotype
// counts).  A SNP whose integer code is the same for everybody has w == 0 exactly (eff = var = 0, chi = NaN in the
// reference); the expanded form would return rounding noise instead, so it is flagged o_j = NaN.  A missing call
// makes the coded column NaN in the reference's numpy path (remma_epiAA.py:59-61 on read_plink's NaN), so every
// statistic of such a SNP is NaN and it is never reported: flagged o_j = +inf.
__global__ void i8_offs_kernel(const int *__restrict__ counts, const double *__restrict__ lutV, long long m, int coding,
                               int minor_allele, double *__restrict__ offs, double *__restrict__ asq) {
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m) return;
    const int c0 = counts[4 * j], c1 = counts[4 * j + 1], c2 = counts[4 * j + 2], c3 = counts[4 * j + 3];
    const int v3 = coding == GMAT_B200_CODING_DOM ? 0 : 2;         // integer value of code 11
    int hi = -1, lo = 9;
    if (c0 > 0) { hi = max(hi, 0); lo = min(lo, 0); }
    if (c2 > 0) { hi = max(hi, 1); lo = min(lo, 1); }
    if (c3 > 0) { hi = max(hi, v3); lo = min(lo, v3); }
    double o = -lutV[4 * j];
    double a2 = (double)c2 + (double)(v3 * v3) * (double)c3;
    const bool flip = minor_allele && i8_flipped(counts, j, coding != GMAT_B200_CODING_DOM);
    if (flip) { o = 2.0 - o; a2 = (double)c2 + 4.0 * (double)c0; }   // a' = 2 - a (see i8_dosage_kernel)
    if (hi <= lo) o = __longlong_as_double(0x7ff8000000000000LL);
    if (c1 > 0) o = __longlong_as_double(0x7ff0000000000000LL);
    offs[j] = o;
    asq[j] = flip ? -a2 : a2;                                        // sign bit = flipped code: eff changes sign
}

// max |x| over a device array, as the bit pattern of a non-negative double (monotone in the value; a NaN wins)
__global__ void absmax_kernel(const double *__restrict__ x, long long len, unsigned long long *out) {
    double mx = 0.0;
    bool bad = false;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < len; k += (long long)gridDim.x * blockDim.x) {
        const double v = fabs(x[k]);
        bad |= v != v;
        mx = fmax(mx, v);
    }
    unsigned long long bits = bad ? 0x7ff8000000000000ULL : (unsigned long long)__double_as_longlong(mx);
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, d);
        bits = other > bits ? other : bits;
    }
    if ((threadIdx.x & 31) == 0 && bits) atomicMax(out, bits);
}

// coded first SNPs: u[ii][k] = lutU[i][code of individual k]
__global__ void i8_first_decode_kernel(const unsigned char *__restrict__ bed, int n, long long row_bytes,
                                       const double *__restrict__ lutU, const long long *__restrict__ ilist,
                                       double *__restrict__ u) {
    const int ii = blockIdx.x;
    const long long i = ilist[ii];
    const unsigned char *row = bed + i * row_bytes;
    const double l0 = lutU[4 * i], l1 = lutU[4 * i + 1], l2 = lutU[4 * i + 2], l3 = lutU[4 * i + 3];
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        const int c = code_at(row, k);
        u[(size_t)ii * n + k] = c == 0 ? l0 : (c == 1 ? l1 : (c == 2 ? l2 : l3));
    }
}

// r[ii][k] = u[ii][k] * sum_l A[k][l] u[ii][l]   (A symmetric: row k of A).  Tile = 32 rows k x 64 first SNPs, the l
// axis in steps of 32 through shared memory; every thread owns 2 x 4 outputs and sums over l in ascending order, so
// the result does not depend on the launch shape.  n^2 flop per first SNP against S n^2 / 2 int8 MACs per PAIR.
constexpr int AU_BK = 32, AU_BI = 64, AU_BL = 32;
__global__ void __launch_bounds__(256) i8_au_kernel(const double *__restrict__ A, int n, const double *__restrict__ u,
                                                    int n_i, double *__restrict__ r) {
    __shared__ double As[AU_BL][AU_BK + 1];
    __shared__ double Us[AU_BL][AU_BI + 1];
    const int k0 = blockIdx.x * AU_BK, i0 = blockIdx.y * AU_BI;
    const int tk = (threadIdx.x >> 4) * 2, ti = (threadIdx.x & 15) * 4;
    double acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
    for (int l0 = 0; l0 < n; l0 += AU_BL) {
        {   // A tile: 32 rows x 32 columns, 4 per thread
            const int row = threadIdx.x >> 3, c0 = (threadIdx.x & 7) * 4;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int k = k0 + row, l = l0 + c0 + c;
                As[c0 + c][row] = (k < n && l < n) ? A[(size_t)k * n + l] : 0.0;
            }
        }
        {   // u tile: 64 first SNPs x 32 columns, 8 per thread
            const int row = threadIdx.x >> 2, c0 = (threadIdx.x & 3) * 8;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int ii = i0 + row, l = l0 + c0 + c;
                Us[c0 + c][row] = (ii < n_i && l < n) ? u[(size_t)ii * n + l] : 0.0;
            }
        }
        __syncthreads();
#pragma unroll 8
        for (int l = 0; l < AU_BL; ++l) {
            const double a0 = As[l][tk], a1 = As[l][tk + 1];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const double uv = Us[l][ti + c];
                acc[0][c] = fma(a0, uv, acc[0][c]);
                acc[1][c] = fma(a1, uv, acc[1][c]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int k = k0 + tk + a, ii = i0 + ti + c;
            if (k < n && ii < n_i) r[(size_t)ii * n + k] = u[(size_t)ii * n + k] * acc[a][c];
        }
}

// Balanced base-256 digits of t, all S at once: with e_s = d_s + 128 in [0, 255], sum_s e_s 256^s = t + 128 sum_s 256^s,
// so the BYTES of t + 0x80..80 are the e_s, and e_s - 128 as an int8 bit pattern is e_s ^ 0x80: one 64-bit add and one
// 64-bit xor give every digit (byte s = digit s, least significant first; |t| <= 126.5 * 256^(S-1) by the choice of
// scale, so nothing carries out of byte S-1).  The shift-and-mask loop this replaces cost ~8 instructions per digit.
template <int S>
__device__ __forceinline__ unsigned long long i8_digits(long long t) {
    constexpr unsigned long long BIAS = 0x8080808080808080ULL >> (8 * (8 - S));
    return ((unsigned long long)t + BIAS) ^ BIAS;
}
// digit bytes of 4 neighbouring columns -> one word per plane (plane q = digit S-1-q: most significant first)
template <int S>
__device__ __forceinline__ void i8_pack4(const unsigned long long (&e)[4], int g, unsigned (&dg)[S][4]) {
#pragma unroll
    for (int q = 0; q < S; ++q) {
        const int b = S - 1 - q;
        const unsigned x0 = b < 4 ? (unsigned)e[0] : (unsigned)(e[0] >> 32), x1 = b < 4 ? (unsigned)e[1] : (unsigned)(e[1] >> 32);
        const unsigned x2 = b < 4 ? (unsigned)e[2] : (unsigned)(e[2] >> 32), x3 = b < 4 ? (unsigned)e[3] : (unsigned)(e[3] >> 32);
        const unsigned sel = (unsigned)(b & 3) | ((unsigned)(4 + (b & 3)) << 4);
        dg[q][g] = __byte_perm(__byte_perm(x0, x1, sel), __byte_perm(x2, x3, sel), 0x5410);
    }
}

// Per first SNP (one block): the scalars of the side table and the four extra plane rows.
//   side[0] = scale_i   >= max|B_i| / 0.49      (max|u_i|^2 max|A| / 0.49)
//   side[1] = s_i = sum_k r_ik,   side[2] = u_i . py
//   side[3] = scale_r (max|r_i| / 0.49),  side[4] = scale_e (max|u_i o py| / 0.49)
//   side[5] = 4 * 256^-S * scale_i   (the a-posteriori guard's error unit; 0 for a monomorphic first SNP)
// Sums run in a fixed order (per-thread strided partials, then a tree), so they are run-to-run identical.  A first SNP
// with a missing call has NaN in u_i (lut[i][code 01]); s_i and u_i.py are then NaN and so is every statistic.
// Extra rows (xrow .. xrow+3 of every plane): y = x / scale as t + f, t = rn(y 256^S) (the "high" S digits),
// f = y 256^S - t exactly, |f| <= 1/2, and rn(f 256^S / 2) (the "low" S digits): y = 256^-S (t + 2 256^-S t_low).
constexpr int I8_PREP_THREADS = 128;    // operand kernels: small CTAs (<= 16 K registers) so that they co-reside with the
                                        // persistent scan kernel (248 registers x 192 threads per SM) of the previous batch
template <int S>
__global__ void __launch_bounds__(I8_PREP_THREADS)
i8_side_kernel(const double *__restrict__ u, const double *__restrict__ r, const double *__restrict__ py, int n,
               const unsigned long long *__restrict__ amax_bits, int np, int ld, int xrow,
               double *__restrict__ side, signed char *__restrict__ planes) {
    constexpr int NT = I8_PREP_THREADS;
    __shared__ double red[5][NT];
    __shared__ double sc[2];
    const int ii = blockIdx.x, t = threadIdx.x;
    const double *ui = u + (size_t)ii * n, *ri = r + (size_t)ii * n;
    double s = 0.0, upy = 0.0, rmax = 0.0, emax = 0.0, umax = 0.0;
    for (int k = t; k < n; k += NT) {
        const double uu = ui[k], rr = ri[k], e = uu * py[k];
        s += rr; upy += e;
        rmax = fmax(rmax, fabs(rr)); emax = fmax(emax, fabs(e)); umax = fmax(umax, fabs(uu));
    }
    red[0][t] = s; red[1][t] = upy; red[2][t] = rmax; red[3][t] = emax; red[4][t] = umax;
    __syncthreads();
    for (int d = NT / 2; d >= 1; d >>= 1) {
        if (t < d) {
            red[0][t] += red[0][t + d]; red[1][t] += red[1][t + d];
            red[2][t] = fmax(red[2][t], red[2][t + d]); red[3][t] = fmax(red[3][t], red[3][t + d]);
            red[4][t] = fmax(red[4][t], red[4][t + d]);
        }
        __syncthreads();
    }
    double pw = 1.0;
#pragma unroll
    for (int q = 0; q < S; ++q) pw *= 256.0;                       // 256^S
    if (t == 0) {
        const double amax = __longlong_as_double((long long)*amax_bits);
        const double um = red[4][0];
        double *sd = side + (size_t)ii * I8_SIDE;
        const double scale = fmax(um * um * (amax / 0.49), 1e-200);
        sd[0] = scale; sd[1] = red[0][0]; sd[2] = red[1][0];
        const double scale_r = red[2][0] > 0.0 ? red[2][0] / 0.49 : 1.0;
        const double scale_e = red[3][0] > 0.0 ? red[3][0] / 0.49 : 1.0;
        sd[3] = scale_r; sd[4] = scale_e;
        sd[5] = um > 0.0 ? 4.0 * scale / pw : 0.0;
        sd[6] = 0.0; sd[7] = 0.0;
        sc[0] = pw / scale_r; sc[1] = pw / scale_e;
    }
    __syncthreads();
    const double mr = sc[0], me = sc[1];
    for (int g = t; g < ld / 16; g += NT) {
        const int k0 = g * 16;
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            unsigned hi[S][4], lo[S][4];
#pragma unroll
            for (int g4 = 0; g4 < 4; ++g4) {
                unsigned long long eh[4], el[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int k = k0 + 4 * g4 + c;
                    double x = 0.0;
                    if (k < n) x = which == 0 ? ri[k] * mr : (ui[k] * py[k]) * me;
                    const long long th = __double2ll_rn(x);
                    const double f = x - (double)th;                // exact, |f| <= 1/2
                    eh[c] = i8_digits<S>(th);
                    el[c] = i8_digits<S>(__double2ll_rn(f * (0.5 * pw)));
                }
                i8_pack4<S>(eh, g4, hi);
                i8_pack4<S>(el, g4, lo);
            }
#pragma unroll
            for (int q = 0; q < S; ++q) {
                signed char *base = planes + ((size_t)(ii * S + q) * np + xrow + 2 * which) * ld + k0;
                *reinterpret_cast<uint4 *>(base) = make_uint4(hi[q][0], hi[q][1], hi[q][2], hi[q][3]);
                *reinterpret_cast<uint4 *>(base + ld) = make_uint4(lo[q][0], lo[q][1], lo[q][2], lo[q][3]);
            }
        }
    }
}

// B_i = diag(u_i) A diag(u_i) / scale_i -> S signed base-256 digits.  Lower block-triangle of 128 x 128 blocks,
// diagonal blocks halved (the kernel doubles the sum); rows >= n are not touched (they hold the extra rows of
// i8_side_kernel and zeros), everything else in the plane buffer stays zero.
template <int S>
__global__ void __launch_bounds__(I8_PREP_THREADS)
i8_split_kernel(const double *__restrict__ A, int n, const double *__restrict__ U, int n_i, int i_per_block,
                const double *__restrict__ side, int np, int ld, signed char *__restrict__ planes) {
    int pk = 0, rem = blockIdx.x;                       // blockIdx.x -> (pk, pl), pl <= pk
    while (rem > pk) { rem -= pk + 1; ++pk; }
    const int pl = rem;
    const int cg = threadIdx.x & 7, r0 = threadIdx.x >> 3;
    const int l0 = pl * 128 + cg * 16;
    double mul = pk == pl ? 0.5 : 1.0;                  // 256^S (x 1/2 on the diagonal blocks)
#pragma unroll
    for (int q = 0; q < S; ++q) mul *= 256.0;
    const int i_begin = blockIdx.y * i_per_block;
    const int i_end = min(n_i, i_begin + i_per_block);
    for (int pass = 0; pass < 128 / (I8_PREP_THREADS / 8); ++pass) {
        const int k = pk * 128 + pass * (I8_PREP_THREADS / 8) + r0;
        if (k >= n) continue;
        double a[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) a[c] = l0 + c < n ? A[(size_t)k * n + l0 + c] : 0.0;
        for (int ii = i_begin; ii < i_end; ++ii) {
            const double *u = U + (size_t)ii * n;
            const double uk = u[k] * (mul / side[(size_t)ii * I8_SIDE]);
            unsigned dg[S][4];
#pragma unroll
            for (int g4 = 0; g4 < 4; ++g4) {
                unsigned long long e[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int col = 4 * g4 + c;
                    const double ul = l0 + col < n ? u[l0 + col] : 0.0;
                    e[c] = i8_digits<S>(__double2ll_rn(uk * a[col] * ul));
                }
                i8_pack4<S>(e, g4, dg);
            }
#pragma unroll
            for (int s = 0; s < S; ++s) {
                signed char *dst = planes + ((size_t)(ii * S + s) * np + k) * ld + l0;
                *reinterpret_cast<uint4 *>(dst) = make_uint4(dg[s][0], dg[s][1], dg[s][2], dg[s][3]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------ pair lists
// Exact test of an explicit pair list (remma_epiAA_pair.py:66-87) on the int8 tensor cores.  A list's pairs rarely
// share a first SNP, so the planes cannot be those of diag(u_i) A diag(u_i); they are the planes of A itself (a "first
// SNP" with u = 1, built once per model by the operand kernels above), and the operand rows are INTEGER vectors per
// pair.  w_p = (a - o_i) o (b - o_j) is not one (a, b: the integer codes of the two SNPs), but
//     w = q - o_j a - o_i b + o_i o_j 1,      q = a o b,
// so w'Aw needs the bilinear forms of q, a, b, 1.  Those with 1 are dot products with t1 = A 1; the others follow from
// QUADRATIC forms by polarisation, x'Ay = (Q(x + y) - Q(x) - Q(y)) / 2:  four rows per pair -- q, q + a, q + b, a + b
// (values 0..6: int8 operand, 4-bit epilogue codes) -- plus one row per distinct first / second SNP of the chunk (a, b).
// The scan kernel computes Q(x) = x'Ax (lower triangle, doubled) and x.py for every row; i8_pairs_combine_kernel puts
// var, eff, chi together and applies the a-posteriori guard (the split error of every Q, weighted), weak pairs go to
// the FP64 pair kernel like the scan's.  Cost: 4 rows x S n^2 / 2 int8 MACs per pair against n^2 issued FP64 FMAs.
__global__ void i8_pair_rows_kernel(const unsigned char *__restrict__ bed, int n, long long row_bytes, unsigned tabU,
                                    unsigned tabV, const int *__restrict__ pairs, int n_pairs,
                                    const int *__restrict__ snpU, int nU, const int *__restrict__ snpV, int nV,
                                    signed char *__restrict__ x8, unsigned char *__restrict__ x4, int ld) {
    const long long groups = ld / 16;
    const long long rows = 4LL * n_pairs + nU + nV;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rows * groups) return;
    const long long r = t / groups;
    const int k0 = (int)(t % groups) * 16;
    int i = -1, j = -1, mode;
    if (r < 4LL * n_pairs) { mode = (int)(r & 3); i = pairs[2 * (r >> 2)]; j = pairs[2 * (r >> 2) + 1]; }
    else if (r < 4LL * n_pairs + nU) { mode = 4; i = snpU[r - 4LL * n_pairs]; }
    else { mode = 5; j = snpV[r - 4LL * n_pairs - nU]; }
    unsigned wi = 0, wj = 0;
    if (k0 / 4 + 3 < row_bytes) {
        if (i >= 0) wi = *reinterpret_cast<const unsigned *>(bed + (size_t)i * row_bytes + k0 / 4);
        if (j >= 0) wj = *reinterpret_cast<const unsigned *>(bed + (size_t)j * row_bytes + k0 / 4);
    }
    unsigned out8[4] = {0, 0, 0, 0}, out4[2] = {0, 0};
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        const unsigned a = (tabU >> (8 * ((wi >> (2 * c)) & 3u))) & 0xffu;
        const unsigned b = (tabV >> (8 * ((wj >> (2 * c)) & 3u))) & 0xffu;
        unsigned v = mode == 0 ? a * b : mode == 1 ? a * b + a : mode == 2 ? a * b + b : mode == 3 ? a + b
                   : mode == 4 ? a : b;
        if (k0 + c >= n) v = 0;
        out8[c >> 2] |= v << (8 * (c & 3));
        out4[c >> 3] |= v << (4 * (c & 7));
    }
    *reinterpret_cast<uint4 *>(x8 + r * ld + k0) = make_uint4(out8[0], out8[1], out8[2], out8[3]);
    *reinterpret_cast<uint2 *>(x4 + r * (ld >> 1) + (k0 >> 1)) = make_uint2(out4[0], out4[1]);
}

// per operand row: sum x^2 (exact) and x . t1 (t1 = A 1; fixed order), one warp per row
__global__ void __launch_bounds__(256) i8_rows_dots_kernel(const signed char *__restrict__ x8, int ld, int n, long long rows,
                                                           const double *__restrict__ t1, double *__restrict__ asq,
                                                           double *__restrict__ xt1) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    const unsigned *row = reinterpret_cast<const unsigned *>(x8 + r * ld);
    double acc = 0.0;
    int sq = 0;
    for (int w = lane; w * 4 < n; w += 32) {
        const unsigned x = __ldg(row + w);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int v = (int)((x >> (8 * c)) & 0xffu), k = 4 * w + c;
            sq += v * v;
            if (k < n) acc += (double)v * t1[k];
        }
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, d);
        sq += __shfl_xor_sync(0xffffffffu, sq, d);
    }
    if (lane == 0) { asq[r] = (double)sq; xt1[r] = acc; }
}

// var, eff, chi of every pair from the rows' Q(x) = x'Ax, E(x) = x.py, T(x) = x.t1 (see above); weak[p] = 1 when the
// planes cannot resolve the variance (the caller re-evaluates the pair in FP64)
__global__ void i8_pairs_combine_kernel(const int *__restrict__ pairs, int n_pairs, const int *__restrict__ invU, int nU,
                                        const int *__restrict__ invV, const double *__restrict__ lutU,
                                        const double *__restrict__ lutV, const double *__restrict__ Q,
                                        const double *__restrict__ E, const double *__restrict__ T,
                                        const double *__restrict__ asq, const double *__restrict__ side,
                                        double recheck_tol, gmat_b200_hit *__restrict__ out,
                                        unsigned char *__restrict__ weak) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const int i = pairs[2 * p], j = pairs[2 * p + 1];
    const long long r0 = 4LL * p, ra = 4LL * n_pairs + invU[p], rb = 4LL * n_pairs + nU + invV[p];
    const double oi = -lutU[4 * (size_t)i], oj = -lutV[4 * (size_t)j];
    const double c2 = -oj, c3 = -oi, c4 = oi * oj;
    const double Q0 = Q[r0], Q1 = Q[r0 + 1], Q2 = Q[r0 + 2], Q3 = Q[r0 + 3], Qa = Q[ra], Qb = Q[rb];
    const double Bqa = 0.5 * ((Q1 - Q0) - Qa), Bqb = 0.5 * ((Q2 - Q0) - Qb), Bab = 0.5 * ((Q3 - Qa) - Qb);
    const double s11 = side[1], sp = side[2];                     // 1'A1 and 1.py (side table of the u = 1 "first SNP")
    const double var = Q0 + c2 * c2 * Qa + c3 * c3 * Qb + 2.0 * (c2 * Bqa + c3 * Bqb + c2 * c3 * Bab) +
                       2.0 * c4 * (T[r0] + c2 * T[ra] + c3 * T[rb]) + c4 * c4 * s11;
    const double eff = E[r0] + c2 * E[ra] + c3 * E[rb] + c4 * sp;
    gmat_b200_hit h;
    h.snp_0 = i; h.snp_1 = j; h.eff = eff; h.var = var; h.chi = eff * eff / var;
    out[p] = h;
    // what the split may have dropped: side[5] * sum x^2 per row, weighted by the row's coefficient in var
    const double bound = side[5] * (fabs(1.0 - c2 - c3) * asq[r0] + fabs(c2) * asq[r0 + 1] + fabs(c3) * asq[r0 + 2] +
                                    fabs(c2 * c3) * asq[r0 + 3] + fabs(c2 * c2 - c2 - c2 * c3) * asq[ra] +
                                    fabs(c3 * c3 - c3 - c2 * c3) * asq[rb]);
    weak[p] = !(bound <= recheck_tol * fabs(var)) ? 1 : 0;        // NaN-safe: anything odd is re-evaluated
}

int launch_i8_dosage(const unsigned char *bed_dev, int n, long long m, long long row_bytes, unsigned tab, const int *counts_dev,
                     signed char *g8_dev, unsigned char *g2_dev, int ld, cudaStream_t st) {
    const long long work = m * (ld / 16);
    i8_dosage_kernel<<<(unsigned)((work + 255) / 256), 256, 0, st>>>(bed_dev, n, m, row_bytes, tab, counts_dev, g8_dev, g2_dev, ld);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

}  // namespace gb

using namespace gb;

extern "C" long long gmat_b200_i8_ld(long long n) { return i8_rows(n); }

extern "C" int gmat_b200_i8_dosage(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                   int coding, const int *counts_dev, signed char *g8_dev, unsigned char *g2_dev,
                                   long long ld, void *stream) {
    GB_CHECK_ARG(bed_dev && g8_dev && n > 0 && m > 0, "null buffer or empty shape");
    GB_CHECK_ARG(ld >= n && ld % 16 == 0 && row_bytes % 4 == 0, "ld must be a multiple of 16 not below n");
    GB_CHECK_ARG(coding == GMAT_B200_CODING_ADD || coding == GMAT_B200_CODING_DOM, "coding");
    // byte[code]: additive dosage {0, miss -> 0, 1, 2}; dominance indicator {0, miss -> 0, 1, 0}; with counts the
    // additive code of a SNP with more 11 than 00 homozygotes is {2, 0, 1, 0} (the other allele is counted)
    return launch_i8_dosage(bed_dev, (int)n, m, row_bytes, coding == GMAT_B200_CODING_DOM ? I8_TAB_DOM : I8_TAB_ADD,
                            counts_dev, g8_dev, g2_dev, (int)ld, (cudaStream_t)stream);
}

extern "C" int gmat_b200_i8_offs(const int *counts_dev, const double *lutV_dev, long long m, int coding,
                                 int minor_allele, double *offs_dev, double *asq_dev, void *stream) {
    GB_CHECK_ARG(counts_dev && lutV_dev && offs_dev && asq_dev && m > 0, "null buffer or empty shape");
    GB_CHECK_ARG(coding == GMAT_B200_CODING_ADD || coding == GMAT_B200_CODING_DOM, "coding");
    i8_offs_kernel<<<(unsigned)((m + 255) / 256), 256, 0, (cudaStream_t)stream>>>(counts_dev, lutV_dev, m, coding,
                                                                                 minor_allele, offs_dev, asq_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_absmax(const double *x_dev, long long len, double *out_dev, void *stream) {
    GB_CHECK_ARG(x_dev && out_dev && len > 0, "null buffer or empty array");
    cudaStream_t st = (cudaStream_t)stream;
    GB_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(double), st));
    const long long want = (len + 1023) / 1024;
    const int grid = (int)(want < 4LL * sm_count() ? want : 4LL * sm_count());
    absmax_kernel<<<grid, 256, 0, st>>>(x_dev, len, reinterpret_cast<unsigned long long *>(out_dev));
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

#define GB_PLANES_SWITCH(n_planes, MACRO)                                                                     \
    switch (n_planes) {                                                                                       \
        case 2: MACRO(2); break; case 3: MACRO(3); break; case 4: MACRO(4); break; case 5: MACRO(5); break;   \
        case 6: MACRO(6); break; default: MACRO(7); break;                                                    \
    }

extern "C" int gmat_b200_i8_first(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                  const double *lutU_dev, const double *py_dev, const double *pvp_dev,
                                  const double *amax_dev, const long long *ilist_dev, long long n_i, int n_planes,
                                  double *u_dev, double *r_dev, double *side_dev, signed char *planes_dev,
                                  void *stream) {
    GB_CHECK_ARG(py_dev && pvp_dev && amax_dev && u_dev && r_dev && side_dev && planes_dev, "null buffer");
    GB_CHECK_ARG(ilist_dev == nullptr || (bed_dev && lutU_dev), "null buffer");
    GB_CHECK_ARG(n > 0 && m > 0 && n_i > 0, "empty shape");
    GB_CHECK_ARG(n_planes >= 2 && n_planes <= I8_MAX_PLANES, "n_planes must be in [2, 7]");
    cudaStream_t st = (cudaStream_t)stream;
    const int ld = (int)i8_rows(n), np = ld, xrow = (int)i8_xrow(n);
    if (ilist_dev != nullptr) {                  // NULL: u_dev already holds the n_i vectors (the pair-list engine: u = 1)
        i8_first_decode_kernel<<<(unsigned)n_i, 256, 0, st>>>(bed_dev, (int)n, row_bytes, lutU_dev, ilist_dev, u_dev);
        GB_LAUNCH_CHECK();
    }
    dim3 grid((unsigned)((n + AU_BK - 1) / AU_BK), (unsigned)((n_i + AU_BI - 1) / AU_BI));
    i8_au_kernel<<<grid, 256, 0, st>>>(pvp_dev, (int)n, u_dev, (int)n_i, r_dev);
    GB_LAUNCH_CHECK();
    const unsigned long long *amax_bits = reinterpret_cast<const unsigned long long *>(amax_dev);
#define GB_SIDE(SV) i8_side_kernel<SV><<<(unsigned)n_i, I8_PREP_THREADS, 0, st>>>(u_dev, r_dev, py_dev, (int)n, amax_bits, np, ld, xrow, side_dev, planes_dev)
    GB_PLANES_SWITCH(n_planes, GB_SIDE)
#undef GB_SIDE
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_i8_split(const double *pvp_dev, long long n, const double *u_dev, long long n_i,
                                  const double *side_dev, int n_planes, signed char *planes_dev, void *stream) {
    GB_CHECK_ARG(pvp_dev && u_dev && side_dev && planes_dev && n > 0 && n_i > 0, "null buffer or empty shape");
    GB_CHECK_ARG(n_planes >= 2 && n_planes <= I8_MAX_PLANES, "n_planes must be in [2, 7]");
    const int ld = (int)i8_rows(n), np = ld;
    const int nb = (int)((n + 127) / 128);
    const int blocks = nb * (nb + 1) / 2;
    // enough CTAs for the chip; each keeps its 16 A values per thread across its share of the first SNPs
    int ysplit = (int)((2LL * sm_count() + blocks - 1) / blocks);
    if (ysplit > n_i) ysplit = (int)n_i;
    if (ysplit < 1) ysplit = 1;
    const int per = (int)((n_i + ysplit - 1) / ysplit);
    dim3 grid(blocks, (unsigned)((n_i + per - 1) / per));
    cudaStream_t st = (cudaStream_t)stream;
#define GB_SPLIT(SV) i8_split_kernel<SV><<<grid, I8_PREP_THREADS, 0, st>>>(pvp_dev, (int)n, u_dev, (int)n_i, per, side_dev, np, ld, planes_dev)
    GB_PLANES_SWITCH(n_planes, GB_SPLIT)
#undef GB_SPLIT
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_i8_scan(const signed char *g8_dev, const unsigned char *g2_dev, long long n, long long m,
                                 const signed char *planes_dev,
                                 int n_planes, const long long *ilist_dev, long long n_i,
                                 const int *item_prefix_dev, const int *jt0_dev, long long total_items,
                                 const double *side_dev, const double *offs_dev, const double *asq_dev,
                                 int pair_space, double chi_cut, gmat_b200_hit *hits_dev, long long cap,
                                 unsigned long long *count_dev, int *recheck_dev, long long recheck_cap,
                                 unsigned long long *recheck_count_dev, double recheck_tol, double *q_debug_dev,
                                 int variant, void *stream) {
    GB_CHECK_ARG(g8_dev && g2_dev && planes_dev && ilist_dev && item_prefix_dev && jt0_dev && side_dev && offs_dev && asq_dev &&
                 count_dev && recheck_count_dev, "null buffer");
    GB_CHECK_ARG(n > 0 && m > 0 && n_i > 0 && total_items >= 0 && total_items < (1LL << 31), "shape");
    GB_CHECK_ARG(n_planes >= 2 && n_planes <= I8_MAX_PLANES, "n_planes must be in [2, 7]");
    GB_CHECK_ARG(pair_space == GMAT_B200_PAIRS_UPPER || pair_space == GMAT_B200_PAIRS_ALL, "pair_space");
    GB_CHECK_ARG(cap == 0 || hits_dev, "hits buffer");
    GB_CHECK_ARG(recheck_cap == 0 || recheck_dev, "re-check buffer");
    GB_CHECK_ARG(variant == 2 || variant == 3, "variant must be 2 (full 256-row MMAs) or 3 (trimmed MMAs)");
    if (total_items == 0) return GMAT_B200_OK;
    cudaStream_t st = (cudaStream_t)stream;
    I8Params p;
    p.n = (int)n; p.m = (int)m; p.ld = (int)i8_rows(n); p.np = p.ld;
    p.n_planes = n_planes; p.n_kt = p.ld / I8_TK; p.n_slab = (int)((n + I8_SLAB - 1) / I8_SLAB);
    p.n_i = (int)n_i; p.pair_space = pair_space;
    p.xkt = (int)(i8_xrow(n) / I8_TK); p.xcol = (int)(i8_xrow(n) % I8_TK);
    GB_CHECK_ARG((long long)n_i * n_planes * p.np < (1LL << 31), "plane buffer has too many rows for one tensor map");
    p.total_items = (int)total_items; p.item_prefix = item_prefix_dev; p.jt0 = jt0_dev;
    p.ilist = ilist_dev; p.g8 = g8_dev; p.g2 = g2_dev; p.side = side_dev; p.offs = offs_dev; p.asq = asq_dev;
    p.chi_cut = chi_cut; p.hits = hits_dev; p.cap = cap; p.count = count_dev;
    p.recheck = recheck_dev; p.recheck_cap = recheck_cap; p.recheck_count = recheck_count_dev;
    p.recheck_tol = recheck_tol;
    p.wlo = 2.0;
    for (int q = 0; q < n_planes; ++q) p.wlo *= 0.00390625;      // 2 * 256^-S: weight of the low digit rows
    p.q_debug = q_debug_dev;
    p.e_debug = nullptr;
    p.trim = variant >= 3;

    CUtensorMap map_g, map_ph;                                    // plane map with 128-row boxes (one half per CTA)
    int rc = make_map(&map_g, g8_dev, p.ld, m, p.ld, I8_TJ);
    if (rc == GMAT_B200_OK) rc = make_map(&map_ph, planes_dev, p.ld, (long long)n_i * n_planes * p.np, p.ld, I8P_HALF);
    if (rc != GMAT_B200_OK) return rc;
    const int smem_bytes = I8P_NS * I8P_STAGE_BYTES + 1024 + (int)sizeof(I8PBars);
    GB_CUDA(cudaFuncSetAttribute(i8_scan_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    // the whole 228 KB as shared memory: with the next smaller carveout (196 KB) nothing is left beside this kernel's
    // 194 KB and the operand kernels of the next batch could not co-reside with it (they then ran AFTER the scan:
    // 2 ms of every 11)
    GB_CUDA(cudaFuncSetAttribute(i8_scan_pair_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 cudaSharedmemCarveoutMaxShared));
    const int pairs_of_sms = sm_count() / 2;
    const int grid2 = 2 * (int)(total_items < pairs_of_sms ? total_items : pairs_of_sms);
    i8_scan_pair_kernel<<<grid2, I8_THREADS, smem_bytes, st>>>(map_g, map_ph, p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

// ---- pair lists on the int8 tensor cores (see i8_pair_rows_kernel) ------------------------------------------------
extern "C" int gmat_b200_i8_pair_rows(const unsigned char *bed_dev, long long n, long long row_bytes, int codingU,
                                      int codingV, const int *pairs_dev, long long n_pairs, const int *snpU_dev,
                                      long long nU, const int *snpV_dev, long long nV, const double *t1_dev,
                                      signed char *x8_dev, unsigned char *x4_dev, double *asq_dev, double *xt1_dev,
                                      void *stream) {
    GB_CHECK_ARG(bed_dev && pairs_dev && snpU_dev && snpV_dev && t1_dev && x8_dev && x4_dev && asq_dev && xt1_dev,
                 "null buffer");
    GB_CHECK_ARG(n > 0 && n_pairs > 0 && nU > 0 && nV > 0 && row_bytes % 4 == 0, "shape");
    GB_CHECK_ARG((codingU == GMAT_B200_CODING_ADD || codingU == GMAT_B200_CODING_DOM) &&
                 (codingV == GMAT_B200_CODING_ADD || codingV == GMAT_B200_CODING_DOM), "coding");
    const long long rows = 4 * n_pairs + nU + nV;
    GB_CHECK_ARG(rows < (1LL << 31), "too many operand rows for one launch: split the list");
    const int ld = (int)i8_rows(n);
    cudaStream_t st = (cudaStream_t)stream;
    const long long work = rows * (ld / 16);
    i8_pair_rows_kernel<<<(unsigned)((work + 255) / 256), 256, 0, st>>>(
        bed_dev, (int)n, row_bytes, codingU == GMAT_B200_CODING_DOM ? I8_TAB_DOM : I8_TAB_ADD,
        codingV == GMAT_B200_CODING_DOM ? I8_TAB_DOM : I8_TAB_ADD, pairs_dev, (int)n_pairs, snpU_dev, (int)nU, snpV_dev,
        (int)nV, x8_dev, x4_dev, ld);
    GB_LAUNCH_CHECK();
    i8_rows_dots_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(x8_dev, ld, (int)n, rows, t1_dev, asq_dev, xt1_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_i8_rows_scan(const signed char *x8_dev, const unsigned char *x4_dev, long long n,
                                      long long rows, const signed char *planes_dev, int n_planes,
                                      const double *side_dev, const double *zeros_dev, const double *asq_dev,
                                      const long long *ilist_dev, const int *item_prefix_dev, const int *jt0_dev,
                                      double *q_dev, double *e_dev, unsigned long long *count_dev,
                                      unsigned long long *recheck_count_dev, void *stream) {
    GB_CHECK_ARG(x8_dev && x4_dev && planes_dev && side_dev && zeros_dev && asq_dev && ilist_dev && item_prefix_dev &&
                 jt0_dev && q_dev && e_dev && count_dev && recheck_count_dev, "null buffer");
    GB_CHECK_ARG(n > 0 && rows > 0 && rows < (1LL << 31), "shape");
    GB_CHECK_ARG(n_planes >= 2 && n_planes <= I8_MAX_PLANES, "n_planes must be in [2, 7]");
    GB_CHECK_ARG(n < 14000, "4-bit epilogue sums fit int32 only for n < 14000");
    cudaStream_t st = (cudaStream_t)stream;
    I8Params p;
    p.n = (int)n; p.m = (int)rows; p.ld = (int)i8_rows(n); p.np = p.ld;
    p.n_planes = n_planes; p.n_kt = p.ld / I8_TK; p.n_slab = (int)((n + I8_SLAB - 1) / I8_SLAB);
    p.n_i = 1; p.pair_space = GMAT_B200_PAIRS_ALL;
    p.xkt = (int)(i8_xrow(n) / I8_TK); p.xcol = (int)(i8_xrow(n) % I8_TK);
    p.total_items = (int)((rows + 2 * I8_TJ - 1) / (2 * I8_TJ)); p.item_prefix = item_prefix_dev; p.jt0 = jt0_dev;
    p.ilist = ilist_dev; p.g8 = x8_dev; p.g2 = x4_dev; p.side = side_dev; p.offs = zeros_dev; p.asq = asq_dev;
    p.chi_cut = (double)INFINITY;   // +inf: no hit records, only q / e
    p.hits = nullptr; p.cap = 0; p.count = count_dev;
    p.recheck = nullptr; p.recheck_cap = 0; p.recheck_count = recheck_count_dev;
    p.recheck_tol = p.chi_cut;                                     // the guard is applied per PAIR by the combine kernel
    p.wlo = 2.0;
    for (int q = 0; q < n_planes; ++q) p.wlo *= 0.00390625;
    p.q_debug = q_dev; p.e_debug = e_dev;
    p.trim = 1;
    CUtensorMap map_g, map_ph;
    int rc = make_map(&map_g, x8_dev, p.ld, rows, p.ld, I8_TJ);
    if (rc == GMAT_B200_OK) rc = make_map(&map_ph, planes_dev, p.ld, (long long)n_planes * p.np, p.ld, I8P_HALF);
    if (rc != GMAT_B200_OK) return rc;
    const int smem_bytes = I8P_NS * I8P_STAGE_BYTES + 1024 + (int)sizeof(I8PBars);
    GB_CUDA(cudaFuncSetAttribute(i8_scan_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    const int pairs_of_sms = sm_count() / 2;
    const int grid2 = 2 * (p.total_items < pairs_of_sms ? p.total_items : pairs_of_sms);
    i8_scan_rows_kernel<<<grid2, I8_THREADS, smem_bytes, st>>>(map_g, map_ph, p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}

extern "C" int gmat_b200_i8_pairs_combine(const int *pairs_dev, long long n_pairs, const int *invU_dev, long long nU,
                                          const int *invV_dev, const double *lutU_dev, const double *lutV_dev,
                                          const double *q_dev, const double *e_dev, const double *xt1_dev,
                                          const double *asq_dev, const double *side_dev, double recheck_tol,
                                          gmat_b200_hit *out_dev, unsigned char *weak_dev, void *stream) {
    GB_CHECK_ARG(pairs_dev && invU_dev && invV_dev && lutU_dev && lutV_dev && q_dev && e_dev && xt1_dev && asq_dev &&
                 side_dev && out_dev && weak_dev, "null buffer");
    GB_CHECK_ARG(n_pairs > 0 && nU > 0, "shape");
    i8_pairs_combine_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        pairs_dev, (int)n_pairs, invU_dev, (int)nU, invV_dev, lutU_dev, lutV_dev, q_dev, e_dev, xt1_dev, asq_dev,
        side_dev, recheck_tol, out_dev, weak_dev);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
