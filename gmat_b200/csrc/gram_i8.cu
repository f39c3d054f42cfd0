// GRM on the INT8 tensor cores (sm_100a: tcgen05.mma cta_group::2 kind::i8, TMEM, TMA).
//
// agmat / dgmat_as (gmat/gmatrix/gmatrix.py:53-66, 116-130): K = X X' / scale with X[k][s] = a_ks - o_s, a the
// integer dosage (additive) or the heterozygote indicator (dominance), o_s = 2 p_s or 2 p_s q_s.  Expanding,
//
//     (X X')_kl = sum_s a_ks a_ls  -  c_k  -  c_l  +  d ,      c_k = sum_s o_s a_ks ,   d = sum_s o_s^2
//
// and the only n^2 m term, G G' with G = [a_ks], is a product of small integers: ONE exact int8 x int8 -> int32
// GEMM (|sum| <= 4 m), no digit planes needed.  c (n x m FP64 multiply-adds) and d are cheap.  The FP64 DMMA
// kernel (grm.cu) needs n(n+1)m FP64 flop at 34 TFLOP/s; this kernel does n(n+1)m/2 int8 MACs at > 1 PMAC/s.
//
// Layout: G is re-coded once per call from the 2-bit store into Gt[id][snp] int8 (K = SNP contiguous, what the
// K-major UMMA operands need), zero padded to 256 ids.  Tiles of 256 x 256 ids (lower block triangle) x all SNPs.
// Schedule (gram_segment): Gt is far larger than L2 at scale (5 GB at n = 10,000 x m = 500,000), so WHICH slabs the
// CTA pairs touch at the same time decides the DRAM traffic.  Full waves of tiles run in lockstep -- wave w = tiles
// [w C, (w+1) C) for C CTA pairs, every pair sweeps the SNP axis of ITS tile from slab 0, all at the same pace, so
// the ~40 id bands a wave needs are read from DRAM once and shared through L2 (the first version cut the (tile,
// slab) space into equal contiguous ranges from the start: every pair sat at a different SNP position, each
// re-read its bands from DRAM, 210 GB of DRAM reads = 32 of the 38 ms at that shape).  Only the tiles of the last,
// partial wave are cut stream-K style into equal contiguous (tile, 128-SNP slab) ranges; their parts are added with
// integer atomics -- exact, hence reproducible whatever the order.
#include "tc5.cuh"

namespace gb {

constexpr int GI_NS = 6;
constexpr int GI_THREADS = 192;                                     // producer, issuer, 4 epilogue warps
constexpr int GI_HALF = I8_TK / 2;
constexpr int GI_STAGE_BYTES = I8_A_BYTES + GI_HALF * I8_SLAB;      // 32 KB per CTA: its 128 rows + half of the columns

struct GramBars {
    uint64_t full[GI_NS], empty[GI_NS], tfull[2], tempty[2];
    uint32_t tmem_base;
};

struct GramParams {
    int n, np, ms, n_slab, nt;
    int tiles;                     // nt (nt + 1) / 2 lower-triangle tiles
    int *S;                        // [np][np] int32, zero-initialised
};

// idx-th piece of work of CTA pair c out of C: tile t, slabs [l0, l0 + cnt).  Pieces 0 .. W-1 (W = tiles / C) are the
// whole tiles t = idx C + c of the lockstep waves; then the (tile, slab) units of the remaining tiles % C tiles, cut
// into C equal contiguous ranges, one piece per tile a range touches.  Returns false past the last piece.
__device__ __forceinline__ bool gram_segment(const GramParams &p, int c, int C, int idx, int &t, int &l0, int &cnt) {
    const int W = p.tiles / C;
    if (idx < W) { t = idx * C + c; l0 = 0; cnt = p.n_slab; return true; }
    const long long units = (long long)(p.tiles - W * C) * p.n_slab;
    const long long u0 = units * c / C, u1 = units * (c + 1) / C;
    long long u = u0;
    for (int k = W; ; ++k) {                                   // at most a few pieces: walk them
        if (u >= u1) return false;
        const int rem_t = (int)(u / p.n_slab), l = (int)(u % p.n_slab);
        const int len = (int)min((long long)(p.n_slab - l), u1 - u);
        if (k == idx) { t = W * C + rem_t; l0 = l; cnt = len; return true; }
        u += len;
    }
}

__device__ __forceinline__ void gram_tile_of(int t, int &rt, int &ct) {
    int r = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
    while ((r + 1) * (r + 2) / 2 <= t) ++r;
    while (r * (r + 1) / 2 > t) --r;
    rt = r; ct = t - r * (r + 1) / 2;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GI_THREADS, 1)
i8_gram_pair_kernel(const __grid_constant__ CUtensorMap map_gt, const GramParams p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    GramBars *bars = reinterpret_cast<GramBars *>(smem + GI_NS * GI_STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < GI_NS; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
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

    // Producer and issuer warps run their loops converged; one elected lane issues the single-thread instructions
    // (elect_one, tc5.cuh), so addresses and descriptors stay in uniform registers.
    if (warp == 0) {
        // ================================ TMA producer (both CTAs) ===============================
        int stage = 0; unsigned phase = 0;
        int t, l0, cnt;
        for (int idx = 0; gram_segment(p, cluster_id, n_clusters, idx, t, l0, cnt); ++idx) {
            int rt, ct;
            gram_tile_of(t, rt, ct);
            for (int l = l0; l < l0 + cnt; ++l) {
                mbar_wait(&bars->empty[stage], phase ^ 1u);
                if (elect_one()) {
                    if (rank == 0) mbar_expect_tx(&bars->full[stage], 2 * GI_STAGE_BYTES);
                    unsigned char *sa = smem + stage * GI_STAGE_BYTES;
                    tma_load_2d_pair(sa, &map_gt, l * I8_SLAB, rt * I8_TK + rank * I8_TJ, &bars->full[stage]);
                    tma_load_2d_pair(sa + I8_A_BYTES, &map_gt, l * I8_SLAB, ct * I8_TK + rank * GI_HALF,
                                     &bars->full[stage]);
                }
                if (++stage == GI_NS) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader CTA) ================================
        if (rank == 0) {
            int stage = 0; unsigned phase = 0;
            int abuf = 0; unsigned aphase = 0;
            int t, l0, cnt;
            for (int idx = 0; gram_segment(p, cluster_id, n_clusters, idx, t, l0, cnt); ++idx) {
                mbar_wait(&bars->tempty[abuf], aphase ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(abuf * I8_TK);
                for (int l = 0; l < cnt; ++l) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + (uint32_t)(stage * GI_STAGE_BYTES);
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + I8_A_BYTES);
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < I8_SLAB / 32; ++k)
                            umma_i8_pair(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), I8P_IDESC,
                                         (uint32_t)((l | k) != 0));
                        tc_commit_pair(&bars->empty[stage]);
                        if (l == cnt - 1) tc_commit_pair(&bars->tfull[abuf]);
                    }
                    if (++stage == GI_NS) { stage = 0; phase ^= 1u; }
                }
                if (++abuf == 2) { abuf = 0; aphase ^= 1u; }
            }
        }
    } else {
        // ================================ epilogue (both CTAs): TMEM -> S ========================
        const int q = warp & 3;
        int abuf = 0; unsigned aphase = 0;
        int t, l0, cnt;
        for (int idx = 0; gram_segment(p, cluster_id, n_clusters, idx, t, l0, cnt); ++idx) {
            int rt, ct;
            gram_tile_of(t, rt, ct);
            const bool whole = cnt == p.n_slab;                       // the tile was not cut: plain stores
            const int k = rt * I8_TK + rank * I8_TJ + q * 32 + lane;  // this lane's row of S
            int *dst = p.S + (size_t)k * p.np + ct * I8_TK;
            mbar_wait_warp(&bars->tfull[abuf], aphase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(abuf * I8_TK);
#pragma unroll 1
            for (int ch = 0; ch < I8_TK / 32; ++ch) {
                int v[32];
                tmem_ld32(taddr + ch * 32, v);
                tmem_ld_wait();
                if (k < p.n) {
                    if (whole) {
#pragma unroll
                        for (int g = 0; g < 8; ++g)
                            *reinterpret_cast<int4 *>(dst + ch * 32 + 4 * g) =
                                make_int4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
                    } else {
#pragma unroll
                        for (int c = 0; c < 32; ++c) atomicAdd(dst + ch * 32 + c, v[c]);
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

// 2-bit store [snp][id] -> Gt[id][snp - s0] int8 (zero outside [0,n) x [s0,s1)); block = 128 SNPs x 128 ids.
// Pass 1: the block's 128 x 32 bytes of codes into shared memory (16-byte loads).  Pass 2: thread (sg, ig) turns the
// code words of 4 SNPs x 16 ids into 16 words "4 SNPs of one id" (a 4 x 16 byte transpose in registers) and stores
// them conflict-free (row stride 33 words).  Pass 3: full 128-byte lines of Gt are written, 8 lanes per id row.
// The first version (64 x 256 tile, byte stores into shared memory, 64-byte row pieces) ran at 1.3 TB/s.
__global__ void __launch_bounds__(256)
i8_dosage_t_kernel(const unsigned char *__restrict__ bed, int n, long long row_bytes, long long s0, long long s1,
                   unsigned table, signed char *__restrict__ gt, long long mld) {
    __shared__ unsigned raw[128][9];                               // codes of 128 ids per SNP (8 words) + 1 pad
    __shared__ unsigned tile[128][33];                             // [id][4-SNP group] + 1 pad
    const long long sblk = (long long)blockIdx.x * 128;
    const int k0 = blockIdx.y * 128;
    {
        const int sl = threadIdx.x >> 1, half = threadIdx.x & 1;
        const long long snp = s0 + sblk + sl;
        const long long off = (long long)(k0 / 4) + half * 16;
        uint4 w = make_uint4(0u, 0u, 0u, 0u);
        if (snp < s1 && off + 16 <= row_bytes) w = *reinterpret_cast<const uint4 *>(bed + snp * row_bytes + off);
        raw[sl][half * 4 + 0] = w.x; raw[sl][half * 4 + 1] = w.y; raw[sl][half * 4 + 2] = w.z; raw[sl][half * 4 + 3] = w.w;
    }
    __syncthreads();
    {
        const int sg = threadIdx.x & 31, ig = threadIdx.x >> 5;   // SNPs 4 sg .. 4 sg + 3, ids 16 ig .. 16 ig + 15
        unsigned w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) w[q] = raw[4 * sg + q][ig];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            unsigned o = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q) o |= ((table >> (8 * ((w[q] >> (2 * e)) & 3u))) & 0xffu) << (8 * q);
            tile[16 * ig + e][sg] = (k0 + 16 * ig + e < n) ? o : 0u;
        }
    }
    __syncthreads();
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
        const int row = pass * 32 + (threadIdx.x >> 3), seg = threadIdx.x & 7;
        const uint4 v = make_uint4(tile[row][4 * seg], tile[row][4 * seg + 1], tile[row][4 * seg + 2],
                                   tile[row][4 * seg + 3]);
        *reinterpret_cast<uint4 *>(gt + (size_t)(k0 + row) * mld + sblk + 16 * seg) = v;
    }
}

// c[k] = sum_s o_s a_ks (k < n) and d = sum_s o_s^2 (slot n), o_s = -lut[s][code 0], in a fixed summation order.
// Gt (n x m bytes: 5 GB at the C2 shape) is streamed ONCE: block (chunk, row group) stages the chunk's 4096 offsets
// in shared memory -- transposed, ot[b][i] = o[4 i + b], so that lanes reading consecutive 4-byte words of a row hit
// consecutive doubles -- and its 8 warps walk 128 rows, a warp per row, 128 coalesced bytes per load.  Partials go
// to part[chunk][k]; i8_rowsum_kernel adds the chunks in order.  (The first version ran one block per row over all
// SNPs and re-read the whole offset vector through L2 for every row: 4 ms at n = 10,000 x m = 500,000.)
constexpr int GI_DOT_CHUNK = 4096;
constexpr int GI_DOT_ROWS = 128;
__global__ void __launch_bounds__(256)
i8_rowdot_kernel(const signed char *__restrict__ gt, long long mld, const double *__restrict__ lut, long long s0,
                 long long ms, int n, double *__restrict__ part) {
    __shared__ double ot[4][GI_DOT_CHUNK / 4];
    __shared__ double red[8];
    const long long c0 = (long long)blockIdx.x * GI_DOT_CHUNK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int t = threadIdx.x; t < GI_DOT_CHUNK; t += 256)
        ot[t & 3][t >> 2] = c0 + t < ms ? -lut[(s0 + c0 + t) * 4] : 0.0;
    __syncthreads();
    double *out = part + (size_t)blockIdx.x * (n + 1);
    if (blockIdx.y == 0) {                                        // d: this chunk's share of sum o^2
        double acc = 0.0;
        for (int t = threadIdx.x; t < GI_DOT_CHUNK; t += 256) { const double o = ot[t & 3][t >> 2]; acc += o * o; }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        if (lane == 0) red[warp] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < 8; ++w) t += red[w];
            out[n] = t;
        }
    }
    const int words = (int)(min((long long)GI_DOT_CHUNK, mld - c0) >> 2);   // mld is a multiple of 128
    for (int r = warp; r < GI_DOT_ROWS; r += 8) {
        const int k = blockIdx.y * GI_DOT_ROWS + r;
        if (k >= n) break;
        const unsigned *row = reinterpret_cast<const unsigned *>(gt + (size_t)k * mld + c0);
        double acc = 0.0;
#pragma unroll 4
        for (int i = lane; i < words; i += 32) {
            const unsigned w = __ldg(row + i);
            acc += (double)(int)(w & 0xffu) * ot[0][i] + (double)(int)((w >> 8) & 0xffu) * ot[1][i] +
                   (double)(int)((w >> 16) & 0xffu) * ot[2][i] + (double)(int)(w >> 24) * ot[3][i];
        }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        if (lane == 0) out[k] = acc;
    }
}
__global__ void i8_rowsum_kernel(const double *__restrict__ part, int chunks, int n, double *__restrict__ c_out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n) return;
    double t = 0.0;
    for (int ch = 0; ch < chunks; ++ch) t += part[(size_t)ch * (n + 1) + k];
    c_out[k] = t;
}

// K = ((S - c 1' - 1 c') + d) / scale with the diagonal inflation, from the lower triangle of S, written symmetrically.
// One block per 32 x 32 tile of the lower block triangle: S is read coalesced ONCE (the first version read S[l][k] for
// the upper half: a column walk, 8 x the sectors), both K[k][l] and its mirror K[l][k] leave as full rows through a
// shared-memory transpose; the same expression, (S_hl - c_h) - c_l + d with h >= l, gives both entries of a pair.
__global__ void __launch_bounds__(256)
i8_gram_finish_kernel(const int *__restrict__ S, int np, const double *__restrict__ c, int n,
                      double scale, double small_val, int accumulate, double *__restrict__ K) {
    __shared__ double t[32][33];
    int bk = 0, rem = blockIdx.x;                                   // blockIdx.x -> (bk, bl), bl <= bk
    {
        bk = (int)((sqrtf(8.0f * (float)rem + 1.0f) - 1.0f) * 0.5f);
        while ((long long)(bk + 1) * (bk + 2) / 2 <= rem) ++bk;
        while ((long long)bk * (bk + 1) / 2 > rem) --bk;
        rem -= bk * (bk + 1) / 2;
    }
    const int bl = rem, k0 = bk * 32, l0 = bl * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const double d = c[n];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int k = k0 + ty + 8 * i, l = l0 + tx;
        double v = 0.0;
        if (k < n && l <= k) {
            v = ((double)S[(size_t)k * np + l] - c[k]) - c[l] + d;
            if (!accumulate) {
                v /= scale;
                if (k == l) v += v * small_val;
            }
            const size_t idx = (size_t)k * n + l;
            if (accumulate) K[idx] += v; else K[idx] = v;
        }
        t[ty + 8 * i][tx] = v;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {                                   // the mirror: row l0 + .., column k0 + ..
        const int l = l0 + ty + 8 * i, k = k0 + tx;
        if (k < n && l < k) {
            const size_t idx = (size_t)l * n + k;
            const double v = t[tx][ty + 8 * i];
            if (accumulate) K[idx] += v; else K[idx] = v;
        }
    }
}

// =====================================================================================================================
// The same GRM with FP4 operands: tcgen05.mma kind::mxf4 (block-scaled E2M1, unit scales), twice the int8 MMA rate.
//
// G holds only {0, 1, 2}: every entry is EXACTLY representable in E2M1 (FP4: 0, 0.5, 1, 1.5, 2, 3, 4, 6), every product
// is an integer <= 4 and the FP32 accumulator is exact as long as a sum stays below 2^24 (4 m < 2^24: SNP slices up to
// 4.19 M; beyond that the caller uses the int8 kernel).  So G G' is ONE exact FP4 x FP4 -> FP32 GEMM at the rate of
// the narrowest type the tensor cores have, on half the operand bytes of the int8 kernel (Gt is 2.5 GB instead of 5 GB
// at n = 10,000 x m = 500,000).  The block scale factors the instruction demands (one UE8M0 byte per 32 elements and
// row, read from tensor memory) are all 2^0: a region of TMEM is filled with 0x7F bytes once per CTA (tcgen05.st) and
// every MMA points at it.  The K order of the packed nibbles is irrelevant (both operands are rows of the same Gt).
//
// c_k = sum_s o_s a_ks rides in the same GEMM (the trick of ozaki.cu's side rows): o_s / 4 is cut into 18 balanced
// base-8 digits in [-4, 4] -- all E2M1 values -- stored as 18 extra "individuals" (rows n .. n+17 of Gt); row n+q of
// the product holds sum_s a_ks d_q,s exactly, and c_k = 4 sum_q 8^-(q+1) S[n+q][k] to 54 bits.  The separate pass over
// Gt that formed c (i8_rowdot_kernel: 5 GB) is gone.
constexpr int GF_DIGITS = 18;
constexpr int GF_SF_COL = I8_TK;                                   // TMEM columns [256, 320): the unit scale factors
constexpr int GF_SF_COLS = 64;
struct GramF4Bars {
    uint64_t full[GI_NS], empty[GI_NS], tfull, tempty, sfready;
    uint32_t tmem_base;
};
// instruction descriptor of tcgen05.mma.kind::mxf4.block_scale (cute/arch/mma_sm100_desc.hpp, InstrDescriptorBlockScaled):
// a/b format E2M1 (1) at [7,10) / [10,13), both K-major, N>>3 at [17,23), scale format UE8M0 (1) at bit 23,
// M>>4 at [24,29), scale-factor ids 0, K = 64 (bit 31 = 0)
constexpr uint32_t GF_IDESC = (1u << 7) | (1u << 10) | ((uint32_t)(I8_TK >> 3) << 17) | (1u << 23) |
                              ((uint32_t)((2 * I8_TJ) >> 4) << 24);

__device__ __forceinline__ void umma_f4_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t tsfa, uint32_t tsfb, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tsfa), "r"(tsfb)
        : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, uint32_t value) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, "
        "%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};\n" ::"r"(taddr),
        "r"(value)
        : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GI_THREADS, 1)
f4_gram_pair_kernel(const __grid_constant__ CUtensorMap map_gt, const GramParams p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                           ~static_cast<uintptr_t>(1023));
    GramF4Bars *bars = reinterpret_cast<GramF4Bars *>(smem + GI_NS * GI_STAGE_BYTES);
    const uint32_t smem_base = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < GI_NS; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
        mbar_init(&bars->tfull, 1); mbar_init(&bars->tempty, 8); mbar_init(&bars->sfready, 8);
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
        // ================================ TMA producer (both CTAs) ===============================
        int stage = 0; unsigned phase = 0;
        int t, l0, cnt;
        for (int idx = 0; gram_segment(p, cluster_id, n_clusters, idx, t, l0, cnt); ++idx) {
            int rt, ct;
            gram_tile_of(t, rt, ct);
            for (int l = l0; l < l0 + cnt; ++l) {
                mbar_wait(&bars->empty[stage], phase ^ 1u);
                if (elect_one()) {
                    if (rank == 0) mbar_expect_tx(&bars->full[stage], 2 * GI_STAGE_BYTES);
                    unsigned char *sa = smem + stage * GI_STAGE_BYTES;
                    tma_load_2d_pair(sa, &map_gt, l * I8_SLAB, rt * I8_TK + rank * I8_TJ, &bars->full[stage]);
                    tma_load_2d_pair(sa + I8_A_BYTES, &map_gt, l * I8_SLAB, ct * I8_TK + rank * GI_HALF,
                                     &bars->full[stage]);
                }
                if (++stage == GI_NS) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader CTA) ================================
        if (rank == 0) {
            int stage = 0; unsigned phase = 0;
            unsigned aphase = 0;
            int t, l0, cnt;
            mbar_wait(&bars->sfready, 0u);                           // the unit scale factors are in TMEM (both CTAs)
            tc_fence_after();
            const uint32_t tsf = tmem_base + (uint32_t)GF_SF_COL;
            for (int idx = 0; gram_segment(p, cluster_id, n_clusters, idx, t, l0, cnt); ++idx) {
                mbar_wait(&bars->tempty, aphase ^ 1u);
                tc_fence_after();
                for (int l = 0; l < cnt; ++l) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + (uint32_t)(stage * GI_STAGE_BYTES);
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + I8_A_BYTES);
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < I8_SLAB / 32; ++k)       // K = 64 nibbles = 32 bytes per MMA
                            umma_f4_pair(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), GF_IDESC, tsf,
                                         tsf + 32u, (uint32_t)((l | k) != 0));
                        tc_commit_pair(&bars->empty[stage]);
                        if (l == cnt - 1) tc_commit_pair(&bars->tfull);
                    }
                    if (++stage == GI_NS) { stage = 0; phase ^= 1u; }
                }
                aphase ^= 1u;
            }
        }
    } else {
        // ================================ epilogue (both CTAs): TMEM -> S ========================
        const int q = warp & 3;
        {   // unit scale factors: UE8M0 2^0 = 0x7F in every byte of TMEM columns [256, 320) of this warp's 32 lanes
            const uint32_t tq = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)GF_SF_COL;
            tmem_st32(tq, 0x7F7F7F7Fu);
            tmem_st32(tq + 32u, 0x7F7F7F7Fu);
            asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&bars->sfready);
        }
        unsigned aphase = 0;
        int t, l0, cnt;
        for (int idx = 0; gram_segment(p, cluster_id, n_clusters, idx, t, l0, cnt); ++idx) {
            int rt, ct;
            gram_tile_of(t, rt, ct);
            const bool whole = cnt == p.n_slab;                       // the tile was not cut: plain stores
            const int k = rt * I8_TK + rank * I8_TJ + q * 32 + lane;  // this lane's row of S
            int *dst = p.S + (size_t)k * p.np + ct * I8_TK;
            mbar_wait_warp(&bars->tfull, aphase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
            for (int ch = 0; ch < I8_TK / 32; ++ch) {
                int v[32];
                tmem_ld32(taddr + ch * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 32; ++c) v[c] = __float2int_rn(__int_as_float(v[c]));   // exact integers in FP32
                if (k < p.np) {
                    if (whole) {
#pragma unroll
                        for (int g = 0; g < 8; ++g)
                            *reinterpret_cast<int4 *>(dst + ch * 32 + 4 * g) =
                                make_int4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
                    } else {
#pragma unroll
                        for (int c = 0; c < 32; ++c) atomicAdd(dst + ch * 32 + c, v[c]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&bars->tempty);
            aphase ^= 1u;
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// 2-bit store [snp][id] -> Gt4[id][snp - s0] as packed E2M1 nibbles (8 SNPs per 32-bit word; zero outside [0,n) x
// [s0,s1)).  Block = 256 SNPs x 512 ids: pass 1 reads 128 contiguous bytes of every SNP row (whole lines: the int8
// kernel's 32-byte pieces made DRAM fetch 4.9 GB for a 1.25 GB store); then, for each of the four 128-id quarters,
// thread (sg = lane, ig = warp) turns the code words of 8 SNPs x 16 ids into 16 words "8 SNPs of one id" and stores
// them conflict-free (row stride 33 words), and full 128-byte lines of Gt4 are written, 8 lanes per id row.
// The 8 x 16 block is transposed WORD-PARALLEL: each code word is widened to nibble fields (three shift-or-mask
// steps), the E2M1 nibble of every code is formed by bit logic on all eight fields at once (code 10 -> 0x2 = 1.0,
// code 11 -> 0x4 = 2.0 additive / 0 dominance, codes 00 and 01 -> 0), and the 8 x 8 nibble matrix is transposed by a
// three-stage butterfly -- ~3 instructions per genotype instead of the ~7 of a per-nibble extract / look-up / insert.
__device__ __forceinline__ unsigned f4_expand8(unsigned x16, bool dominance) {
    unsigned x = x16 & 0xFFFFu;                                     // eight 2-bit codes
    x = (x | (x << 8)) & 0x00FF00FFu;
    x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u;                              // each code in the low 2 bits of its nibble
    const unsigned h = (x >> 1) & 0x11111111u, l = x & 0x11111111u;
    const unsigned one = (h & ~l) << 1;                             // code 10 -> nibble 0x2
    return dominance ? one : (one | ((h & l) << 2));                // code 11 -> nibble 0x4
}
__device__ __forceinline__ void f4_transpose8(unsigned (&a)[8]) {   // 8 x 8 nibble transpose, rows = words
#pragma unroll
    for (int i = 0; i < 4; ++i) { const unsigned t = ((a[i] >> 16) ^ a[i + 4]) & 0x0000FFFFu; a[i + 4] ^= t; a[i] ^= t << 16; }
#pragma unroll
    for (int i = 0; i < 8; i += 4)
#pragma unroll
        for (int j = 0; j < 2; ++j) { const unsigned t = ((a[i + j] >> 8) ^ a[i + j + 2]) & 0x00FF00FFu; a[i + j + 2] ^= t; a[i + j] ^= t << 8; }
#pragma unroll
    for (int i = 0; i < 8; i += 2) { const unsigned t = ((a[i] >> 4) ^ a[i + 1]) & 0x0F0F0F0Fu; a[i + 1] ^= t; a[i] ^= t << 4; }
}
__global__ void __launch_bounds__(256)
f4_dosage_t_kernel(const unsigned char *__restrict__ bed, int n, long long row_bytes, long long s0, long long s1,
                   int dominance, unsigned char *__restrict__ gt, long long mld) {
    extern __shared__ unsigned f4_smem[];
    // raw: the code words of SNP row r live at [r & 7][r >> 3][0..31] with plane stride 1060 and row stride 33 words:
    // both the row-wise writes of pass 1 and the "8 rows of one word column" reads of pass 2 are bank-conflict free
    constexpr int RAWQ = 1060;
    unsigned *raw = f4_smem;                                                               // 8 x 1060 words
    unsigned (*tile)[33] = reinterpret_cast<unsigned (*)[33]>(f4_smem + 8 * RAWQ);         // [128 ids][32 words + pad]
    const long long sblk = (long long)blockIdx.x * 256;
    const int kblk = blockIdx.y * 512;
    {
        const long long snp = s0 + sblk + threadIdx.x;
        const long long off = kblk / 4;
        unsigned *dst = raw + (threadIdx.x & 7) * RAWQ + (threadIdx.x >> 3) * 33;
#pragma unroll
        for (int v = 0; v < 8; ++v) {
            uint4 w = make_uint4(0u, 0u, 0u, 0u);
            if (snp < s1 && off + 16 * v + 16 <= row_bytes)
                w = *reinterpret_cast<const uint4 *>(bed + snp * row_bytes + off + 16 * v);
            dst[4 * v + 0] = w.x; dst[4 * v + 1] = w.y; dst[4 * v + 2] = w.z; dst[4 * v + 3] = w.w;
        }
    }
    __syncthreads();
    const int sg = threadIdx.x & 31, ig = threadIdx.x >> 5;       // SNPs 8 sg .. 8 sg + 7, ids 16 ig .. 16 ig + 15 of a quarter
    for (int quarter = 0; quarter < 4; ++quarter) {
        const int k0 = kblk + quarter * 128;
        unsigned lo[8], hi[8];                                     // rows = SNPs, nibble columns = ids 0..7 / 8..15
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const unsigned w = raw[q * RAWQ + sg * 33 + quarter * 8 + ig];
            lo[q] = f4_expand8(w, dominance != 0);
            hi[q] = f4_expand8(w >> 16, dominance != 0);
        }
        f4_transpose8(lo);                                         // rows = ids, nibble columns = SNPs
        f4_transpose8(hi);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            tile[16 * ig + e][sg] = (k0 + 16 * ig + e < n) ? lo[e] : 0u;
            tile[16 * ig + 8 + e][sg] = (k0 + 16 * ig + 8 + e < n) ? hi[e] : 0u;
        }
        __syncthreads();
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) {
            const int row = pass * 32 + (threadIdx.x >> 3), seg = threadIdx.x & 7;
            const uint4 v = make_uint4(tile[row][4 * seg], tile[row][4 * seg + 1], tile[row][4 * seg + 2],
                                       tile[row][4 * seg + 3]);
            *reinterpret_cast<uint4 *>(gt + (size_t)(k0 + row) * mld + (sblk >> 1) + 16 * seg) = v;
        }
        __syncthreads();
    }
}

// rows n .. n+17 of Gt4: the balanced base-8 digits of o_s / 4 (o_s = -lut[s][code 0] in [0, 2]) as E2M1 nibbles, most
// significant digit first; one thread per word (8 SNPs).  E2M1 of an integer v in [-4, 4]: sign 0x8, |v| -> 0,2,4,5,6.
__global__ void f4_offs_rows_kernel(const double *__restrict__ lut, long long s0, long long ms, int n, long long mld,
                                    unsigned char *__restrict__ gt) {
    const long long wi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (wi * 8 >= ms) return;
    unsigned out[GF_DIGITS];
#pragma unroll
    for (int q = 0; q < GF_DIGITS; ++q) out[q] = 0u;
    for (int e = 0; e < 8; ++e) {
        const long long s = wi * 8 + e;
        if (s >= ms) break;
        const double y = -lut[(s0 + s) * 4] * 0.25;                 // |y| <= 0.5
        long long t = __double2ll_rn(y * 18014398509481984.0);      // 8^18 = 2^54
#pragma unroll
        for (int q = GF_DIGITS - 1; q >= 0; --q) {
            long long d = q > 0 ? ((t + 4) & 7) - 4 : t;            // balanced digit in [-4, 3]; the top one is |t| <= 4
            t = (t - d) >> 3;
            const unsigned mag = (unsigned)(d < 0 ? -d : d);
            const unsigned nib = (d < 0 ? 8u : 0u) | (mag < 3u ? 2u * mag : mag + 2u);
            out[q] |= nib << (4 * e);
        }
    }
#pragma unroll
    for (int q = 0; q < GF_DIGITS; ++q)
        *reinterpret_cast<unsigned *>(gt + (size_t)(n + q) * mld + wi * 4) = out[q];
}
// c[k] = 4 sum_q 8^-(q+1) S[n+q][k] (k < n, smallest digit first) and c[n] = d = sum_s o_s^2 (one block, fixed order)
__global__ void __launch_bounds__(1024)
f4_c_kernel(const int *__restrict__ S, int np, int n, const double *__restrict__ lut, long long s0, long long ms,
            double *__restrict__ c) {
    if (blockIdx.x == gridDim.x - 1) {
        __shared__ double red[1024];
        double acc = 0.0;
        for (long long s = threadIdx.x; s < ms; s += 1024) { const double o = lut[(s0 + s) * 4]; acc += o * o; }
        red[threadIdx.x] = acc;
        __syncthreads();
        for (int d = 512; d >= 1; d >>= 1) {
            if ((int)threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
            __syncthreads();
        }
        if (threadIdx.x == 0) c[n] = red[0];
        return;
    }
    const int k = blockIdx.x * 1024 + threadIdx.x;
    if (k >= n) return;
    double acc = 0.0, w = 4.0 / 18014398509481984.0;                // 4 * 8^-18
    for (int q = GF_DIGITS - 1; q >= 0; --q) {
        acc += (double)S[(size_t)(n + q) * np + k] * w;
        w *= 8.0;
    }
    c[k] = acc;
}

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_grm_i8(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                const double *lut_dev, int coding, long long snp_begin, long long snp_end,
                                double scale, double small_val, int accumulate, double *K_dev, void *stream) {
    GB_CHECK_ARG(bed_dev && lut_dev && K_dev, "null buffer");
    GB_CHECK_ARG(n > 0 && m > 0 && row_bytes % 16 == 0 && row_bytes * 4 >= n, "shape");
    GB_CHECK_ARG(snp_begin >= 0 && snp_begin <= snp_end && snp_end <= m, "snp range");
    GB_CHECK_ARG(coding == GMAT_B200_CODING_ADD || coding == GMAT_B200_CODING_DOM, "coding");
    GB_CHECK_ARG(accumulate || scale != 0.0, "scale");
    GB_CHECK_ARG(snp_end - snp_begin < (1LL << 28) && n < (1 << 20), "shape too large for int32 sums / tensor maps");
    cudaStream_t st = (cudaStream_t)stream;
    const long long ms = snp_end - snp_begin;
    const int np = (int)round_up(n, I8_TK);
    const long long mld = round_up(ms > 0 ? ms : 1, 128);
    keep_scratch_pool();
    signed char *gt = nullptr; int *S = nullptr; double *c = nullptr, *part = nullptr;
    const int chunks = (int)((mld + GI_DOT_CHUNK - 1) / GI_DOT_CHUNK);
    GB_CUDA(cudaMallocAsync((void **)&gt, (size_t)np * mld, st));
    GB_CUDA(cudaMallocAsync((void **)&S, (size_t)np * np * sizeof(int), st));
    GB_CUDA(cudaMallocAsync((void **)&c, (size_t)(n + 1) * sizeof(double), st));
    GB_CUDA(cudaMallocAsync((void **)&part, (size_t)chunks * (n + 1) * sizeof(double), st));
    GB_CUDA(cudaMemsetAsync(S, 0, (size_t)np * np * sizeof(int), st));
    int rc = GMAT_B200_OK;
    if (ms > 0) {
        const unsigned table = coding == GMAT_B200_CODING_DOM ? 0x00010000u : 0x02010000u;
        dim3 tg((unsigned)(mld / 128), (unsigned)(np / 128));
        i8_dosage_t_kernel<<<tg, 256, 0, st>>>(bed_dev, (int)n, row_bytes, snp_begin, snp_end, table, gt, mld);
        count_launch();
        GramParams p;
        p.n = (int)n; p.np = np; p.ms = (int)ms; p.n_slab = (int)((ms + I8_SLAB - 1) / I8_SLAB); p.nt = np / I8_TK;
        p.tiles = p.nt * (p.nt + 1) / 2; p.S = S;
        CUtensorMap map_gt;
        rc = make_map(&map_gt, gt, ms, np, mld, I8_TJ);
        if (rc == GMAT_B200_OK) {
            const int smem_bytes = GI_NS * GI_STAGE_BYTES + 1024 + (int)sizeof(GramBars);
            cudaError_t e = cudaFuncSetAttribute(i8_gram_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 smem_bytes);
            if (e != cudaSuccess) { set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); rc = GMAT_B200_ERR_CUDA; }
            else {
                long long pairs = sm_count() / 2;
                if (pairs > (long long)p.tiles * p.n_slab) pairs = (long long)p.tiles * p.n_slab;
                i8_gram_pair_kernel<<<(unsigned)(2 * pairs), GI_THREADS, smem_bytes, st>>>(map_gt, p);
                count_launch();
            }
        }
    }
    if (rc == GMAT_B200_OK) {
        dim3 dg((unsigned)chunks, (unsigned)((n + GI_DOT_ROWS - 1) / GI_DOT_ROWS));
        i8_rowdot_kernel<<<dg, 256, 0, st>>>(gt, mld, lut_dev, snp_begin, ms, (int)n, part);
        count_launch();
        i8_rowsum_kernel<<<(unsigned)((n + 256) / 256), 256, 0, st>>>(part, chunks, (int)n, c);
        count_launch();
        {
            const long long nb32 = (n + 31) / 32;
            i8_gram_finish_kernel<<<(unsigned)(nb32 * (nb32 + 1) / 2), 256, 0, st>>>(S, np, c, (int)n, scale, small_val,
                                                                                    accumulate, K_dev);
        }
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("kernel launch failed: %s", cudaGetErrorString(e)); rc = GMAT_B200_ERR_CUDA; }
    }
    cudaFreeAsync(gt, st);
    cudaFreeAsync(S, st);
    cudaFreeAsync(c, st);
    cudaFreeAsync(part, st);
    return rc;
}

extern "C" int gmat_b200_grm_f4(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                                const double *lut_dev, int coding, long long snp_begin, long long snp_end,
                                double scale, double small_val, int accumulate, double *K_dev, void *stream) {
    GB_CHECK_ARG(bed_dev && lut_dev && K_dev, "null buffer");
    GB_CHECK_ARG(n > 0 && m > 0 && row_bytes % 16 == 0 && row_bytes * 4 >= n, "shape");
    GB_CHECK_ARG(snp_begin >= 0 && snp_begin <= snp_end && snp_end <= m, "snp range");
    GB_CHECK_ARG(coding == GMAT_B200_CODING_ADD || coding == GMAT_B200_CODING_DOM, "coding");
    GB_CHECK_ARG(accumulate || scale != 0.0, "scale");
    // the FP32 accumulator must hold sum a a' <= 4 ms (and |sum a d| <= 8 ms for the digit rows) exactly
    GB_CHECK_ARG(snp_end - snp_begin < (1LL << 21) && n < (1 << 20),
                 "SNP slice too long for exact FP32 sums (use gmat_b200_grm_i8)");
    cudaStream_t st = (cudaStream_t)stream;
    const long long ms = snp_end - snp_begin;
    const int np = (int)round_up(n + GF_DIGITS, I8_TK);
    const long long mld = round_up(ms > 0 ? (ms + 1) / 2 : 1, 128);       // bytes per row of Gt4: 256 SNPs per 128 bytes
    keep_scratch_pool();
    unsigned char *gt = nullptr; int *S = nullptr; double *c = nullptr;
    GB_CUDA(cudaMallocAsync((void **)&gt, (size_t)np * mld, st));
    GB_CUDA(cudaMallocAsync((void **)&S, (size_t)np * np * sizeof(int), st));
    GB_CUDA(cudaMallocAsync((void **)&c, (size_t)(n + 1) * sizeof(double), st));
    GB_CUDA(cudaMemsetAsync(S, 0, (size_t)np * np * sizeof(int), st));
    int rc = GMAT_B200_OK;
    if (ms > 0) {
        const int smem_t = (8 * 1060 + 128 * 33) * 4;
        cudaError_t e = cudaFuncSetAttribute(f4_dosage_t_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_t);
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); rc = GMAT_B200_ERR_CUDA; }
        if (rc == GMAT_B200_OK) {
            dim3 tg((unsigned)((2 * mld) / 256), (unsigned)((np + 511) / 512));
            // rows of the last 512-id block beyond np would fall outside Gt4: the kernel is launched over ceil(np / 512)
            // blocks and Gt4 is allocated with np rows, so guard by allocating the tail (np is a multiple of 256 only)
            if (np % 512 != 0) {
                cudaFreeAsync(gt, st);
                GB_CUDA(cudaMallocAsync((void **)&gt, (size_t)round_up(np, 512) * mld, st));
            }
            f4_dosage_t_kernel<<<tg, 256, smem_t, st>>>(bed_dev, (int)n, row_bytes, snp_begin, snp_end,
                                                         coding == GMAT_B200_CODING_DOM ? 1 : 0, gt, mld);
            count_launch();
            f4_offs_rows_kernel<<<(unsigned)((ms / 8 + 256) / 256), 256, 0, st>>>(lut_dev, snp_begin, ms, (int)n, mld, gt);
            count_launch();
            GramParams p;
            p.n = (int)n; p.np = np; p.ms = (int)ms; p.n_slab = (int)((ms + 2 * I8_SLAB - 1) / (2 * I8_SLAB));
            p.nt = np / I8_TK; p.tiles = p.nt * (p.nt + 1) / 2; p.S = S;
            CUtensorMap map_gt;
            rc = make_map(&map_gt, gt, (ms + 1) / 2, np, mld, I8_TJ);
            if (rc == GMAT_B200_OK) {
                const int smem_bytes = GI_NS * GI_STAGE_BYTES + 1024 + (int)sizeof(GramF4Bars);
                e = cudaFuncSetAttribute(f4_gram_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
                if (e != cudaSuccess) { set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); rc = GMAT_B200_ERR_CUDA; }
                else {
                    long long pairs = sm_count() / 2;
                    if (pairs > (long long)p.tiles * p.n_slab) pairs = (long long)p.tiles * p.n_slab;
                    f4_gram_pair_kernel<<<(unsigned)(2 * pairs), GI_THREADS, smem_bytes, st>>>(map_gt, p);
                    count_launch();
                }
            }
        }
    }
    if (rc == GMAT_B200_OK) {
        if (ms > 0) {
            f4_c_kernel<<<(unsigned)((n + 1023) / 1024 + 1), 1024, 0, st>>>(S, np, (int)n, lut_dev, snp_begin, ms, c);
            count_launch();
        } else {
            GB_CUDA(cudaMemsetAsync(c, 0, (size_t)(n + 1) * sizeof(double), st));
        }
        {
            const long long nb32 = (n + 31) / 32;
            i8_gram_finish_kernel<<<(unsigned)(nb32 * (nb32 + 1) / 2), 256, 0, st>>>(S, np, c, (int)n, scale, small_val,
                                                                                    accumulate, K_dev);
        }
        count_launch();
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("kernel launch failed: %s", cudaGetErrorString(e)); rc = GMAT_B200_ERR_CUDA; }
    }
    cudaFreeAsync(gt, st);
    cudaFreeAsync(S, st);
    cudaFreeAsync(c, st);
    return rc;
}
