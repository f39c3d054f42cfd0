// Genomic relationship matrix  K = X X' / scale  with X[ind][snp] = lut[snp][code] generated on the fly
// from the 2-bit genotype store; FP64 tensor cores (DMMA.8x8x4); SYRK-style: only lower-triangle
// 128x128 tiles are computed, then mirrored; fused epilogue: / scale, diag += diag*small_val.
//
// Reference behaviour restated (never copied):
//   agmat:    centring, kin = dot/scale, diag inflation ... gmat/gmatrix/gmatrix.py:53-66
//   dgmat_as: dominance coding and scale .................. gmat/gmatrix/gmatrix.py:116-130
//
// B200 design (same recipe as the exact-test kernel, exact.cu):
//  * FP64 tensor path = warp-level DMMA.8x8x4; 8 consumer warps (2 x 4, warp tile 64 x 32 individuals = 8 x 4
//    accumulator tiles) + 4 producer warps whose registers go to the consumers (setmaxnreg).
//  * K dimension = SNPs, in stages of 64.  A stage is 6 KB of shared memory: the 2-bit codes of the tile's 128
//    row individuals and 128 column individuals for 64 SNPs (64 x 32 B each, straight from the SNP-major .bed
//    rows by cp.async.bulk -- no decode pass, no doubles in shared memory) and the 64 SNPs' 4-entry value
//    tables (2 KB, one bulk copy).  full[s] completes on the TMA byte count, empty[s] on one arrival per
//    consumer warp; 4-deep ring, no CTA-wide barrier anywhere in the K loop.
//  * Operand fragments are built in registers: lane (g, t) owns SNP 4*k4 + t of the step, reads its 16 B + 8 B
//    of codes (broadcast LDS.128 / LDS.64) and fetches each of its 8 + 4 fragment elements with one LDS.64
//    from that SNP's value table (the 4 SNPs of a step share one 128 B line: conflict-free).
//  * stream-K work split: the (tile, stage) iteration space is cut into gridDim.x equal contiguous ranges, so
//    every SM works to the end whatever the tile count (n = 10,000: 3160 tiles = 21.35 waves of 148).  A tile cut
//    by a range boundary is finished by two CTAs, each adding its partial sum with red.global.add.f64 into the
//    zero-initialised output; the launch keeps every tile in at most two parts, so the sum order cannot vary and
//    the result is run-to-run reproducible.
#include "common.cuh"

namespace gb {

constexpr int GR_T = 128;           // output tile edge (individuals)
constexpr int GR_GS = 64;           // SNPs per ring stage
constexpr int GR_NS = 4;            // ring depth
constexpr int GR_CW = 8;            // consumer warps (2 x 4)
constexpr int GR_PW = 4;            // producer warps (one per SM sub-partition): 128 lanes, one row piece each per stage

struct GrmParams {
    const unsigned char *bed;
    long long n, m, row_bytes;
    const double *lut;
    long long snp_begin, snp_end;
    double scale, small_val;
    int accumulate;
    double *K;
    long long nblk;
    int split2;                     // 0: cost-balanced stream-K ranges, 1: two CTAs per tile, 2: one tile per CTA
};

struct GrmSmem {
    unsigned char cA[GR_NS][GR_GS][32];            // codes of the tile's row individuals, one 32 B piece per SNP
    unsigned char cB[GR_NS][GR_GS][32];            // codes of the tile's column individuals
    double lut[GR_NS][GR_GS][4];                   // value table of each SNP of the stage
    uint64_t full[GR_NS], empty[GR_NS];
};

__device__ __forceinline__ void grm_tile_of(long long tile, long long &mi, long long &ni) {
    mi = (long long)((sqrt(8.0 * (double)tile + 1.0) - 1.0) * 0.5);      // tile -> (mi, ni), ni <= mi
    while (mi * (mi + 1) / 2 > tile) --mi;
    while ((mi + 1) * (mi + 2) / 2 <= tile) ++mi;
    ni = tile - mi * (mi + 1) / 2;
}

// One ring stage (64 SNPs) of DMMAs for the first MTV of the warp's 8 row m-tiles.  MTV < 8 only in the last block
// row, whose rows past n are padding (n = 10,000: 16 of its 128 rows exist).
template <int MTV, int NTL>
__device__ __forceinline__ void grm_stage_mma(double (&acc)[8][NTL][2], const unsigned char (*cA)[32],
                                              const unsigned char (*cB)[32], const double (*lut)[4], int wm, int wn,
                                              int t, int sh0, int sh1) {
#pragma unroll 4
    for (int k4 = 0; k4 < GR_GS / 4; ++k4) {
        const int r = k4 * 4 + t;                    // this lane's SNP of the step
        const uint4 wa = *reinterpret_cast<const uint4 *>(&cA[r][wm * 16]);
        const uint2 wb = *reinterpret_cast<const uint2 *>(&cB[r][wn * 8]);
        const char *lb = reinterpret_cast<const char *>(&lut[r][0]);
        const unsigned wav[4] = {wa.x, wa.y, wa.z, wa.w};
        const unsigned wbv[2] = {wb.x, wb.y};
        double af[MTV], bf[NTL];
        // individual 8a + g of the warp's rows: code bits 16*(a&1) + 2g of word a >> 1
#pragma unroll
        for (int a = 0; a < MTV; ++a)
            af[a] = *reinterpret_cast<const double *>(lb + (((wav[a >> 1] >> ((a & 1) ? sh1 : sh0)) & 3u) << 3));
#pragma unroll
        for (int b = 0; b < NTL; ++b)
            bf[b] = *reinterpret_cast<const double *>(lb + (((wbv[b >> 1] >> ((b & 1) ? sh1 : sh0)) & 3u) << 3));
#pragma unroll
        for (int a = 0; a < MTV; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
}

// m-tiles (in pairs) of warp row wm of block row mi that hold an individual below n
__device__ __forceinline__ int grm_mtv(long long n, long long mi, int wm) {
    const long long left = n - (mi * GR_T + wm * 64);
    return left >= 64 ? 8 : (left <= 0 ? 0 : (int)((left + 15) / 16) * 2);
}

__global__ void __launch_bounds__((GR_CW + GR_PW) * 32, 1) grm_kernel(const GrmParams p) {
    constexpr int MT = 8, NTL = 4;                   // warp tile 64 x 32
    extern __shared__ __align__(128) unsigned char smem_raw[];
    GrmSmem &S = *reinterpret_cast<GrmSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long nstage = (p.snp_end - p.snp_begin + GR_GS - 1) / GR_GS;
    const long long ntiles = p.nblk * (p.nblk + 1) / 2;
    // stream-K: this CTA's contiguous share of the ntiles x nstage units (split2: two CTAs per tile, half each)
    long long u_begin, u_end;
    if (p.split2 == 2) {                             // one whole tile per CTA
        u_begin = (long long)blockIdx.x * nstage;
        u_end = u_begin + nstage;
    } else if (p.split2 == 1) {
        const long long tile = blockIdx.x >> 1, half = (nstage + 1) / 2;
        u_begin = tile * nstage + ((blockIdx.x & 1) ? half : 0);
        u_end = tile * nstage + ((blockIdx.x & 1) ? nstage : half);
    } else {
        // equal COST per CTA: a stage of a full tile costs 16 (m-tile pairs x 2 warp rows), a stage of the last
        // block row only what its non-padding m-tiles cost; tiles are numbered row-major, the last row last
        const long long nfull = ntiles - p.nblk, ufull = nfull * nstage;
        // (floor of 4: a stage costs its barrier round trip and 129 bulk copies however few DMMAs it has)
        long long w_last = grm_mtv(p.n, p.nblk - 1, 0) + grm_mtv(p.n, p.nblk - 1, 1);            // 2 .. 16
        if (w_last < 4) w_last = 4;
        const long long total = 16 * ufull + w_last * p.nblk * nstage;
        auto unit_at = [&](long long c) {            // first unit whose cost prefix reaches c
            if (c <= 16 * ufull) return (c + 15) / 16;
            return ufull + (c - 16 * ufull + w_last - 1) / w_last;
        };
        u_begin = unit_at(total / gridDim.x * blockIdx.x + total % gridDim.x * blockIdx.x / gridDim.x);
        u_end = (blockIdx.x + 1 == gridDim.x)
                    ? ntiles * nstage
                    : unit_at(total / gridDim.x * (blockIdx.x + 1) + total % gridDim.x * (blockIdx.x + 1) / gridDim.x);
    }

    if (tid == 0) {
        for (int s = 0; s < GR_NS; ++s) {
            mbar_init(&S.full[s], 1);                // the expect_tx arrival of producer lane 0
            mbar_init(&S.empty[s], GR_CW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp >= GR_CW) {
        // =============================== producer ========================================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
        // all four producer warps work: 128 lanes, one 32 B row piece each per stage (64 SNPs x 2 operands); a single
        // warp issuing all 129 copies took ~2300 clocks per stage, which is more than a stage of a padded
        // last-row tile needs for its DMMAs
        const int pl = (warp - GR_CW) * 32 + lane;   // 0..127
        const int prow = pl & (GR_GS - 1), pop = pl >> 6;           // SNP of the stage, operand (0: rows, 1: columns)
        int stage = 0;
        unsigned use_odd = 0;
        bool first_pass = true;
        for (long long u = u_begin; u < u_end;) {
            const long long tile = u / nstage, st_begin = u - tile * nstage;
            const long long st_end = (u_end - u) < (nstage - st_begin) ? st_begin + (u_end - u) : nstage;
            u += st_end - st_begin;
            long long mi, ni;
            grm_tile_of(tile, mi, ni);
            // bytes of a row piece: 32, or what is left of the (16 B padded) row for the last block
            const long long offA = mi * (GR_T / 4), offB = ni * (GR_T / 4);
            const unsigned bytesA = (unsigned)((p.row_bytes - offA) < 32 ? (p.row_bytes - offA) : 32);
            const unsigned bytesB = (unsigned)((p.row_bytes - offB) < 32 ? (p.row_bytes - offB) : 32);
            for (long long st = st_begin; st < st_end; ++st) {
                if (!first_pass) mbar_wait(&S.empty[stage], use_odd ^ 1u);
                fence_proxy_async();                 // consumers' LDS of this stage precede the async refill
                const long long s0 = p.snp_begin + st * GR_GS;
                const long long left = p.snp_end - s0;
                const int nvalid = left < GR_GS ? (int)left : GR_GS;
                if (pl == 0) {
                    // table rows of SNPs past the end are zero (plain stores, released by the arrive below).  The
                    // other lanes' copies may complete before this expect_tx: the phase cannot, it still waits for
                    // this arrival, and the transaction count is allowed to run negative in between.
                    for (int r = nvalid; r < GR_GS; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) S.lut[stage][r][c] = 0.0;
                    mbar_expect_tx(&S.full[stage], (unsigned)GR_GS * (bytesA + bytesB) + (unsigned)nvalid * 32u);
                    tma_bulk_g2s(S.lut[stage], p.lut + s0 * 4, (unsigned)nvalid * 32u, &S.full[stage]);
                }
                {
                    // SNPs past the end re-read the last valid row; their value table is zero
                    const long long snp = s0 + (prow < nvalid ? prow : nvalid - 1);
                    const unsigned char *row = p.bed + snp * p.row_bytes;
                    if (pop == 0) tma_bulk_g2s(S.cA[stage][prow], row + offA, bytesA, &S.full[stage]);
                    else tma_bulk_g2s(S.cB[stage][prow], row + offB, bytesB, &S.full[stage]);
                }
                if (++stage == GR_NS) { stage = 0; use_odd ^= 1u; first_pass = false; }
            }
        }
        return;
    }

    // =================================== consumer warps =================================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n");
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp >> 2, wn = warp & 3;
    const int sh0 = 2 * g, sh1 = 16 + 2 * g;
    int stage = 0;
    unsigned fphase = 0;
    for (long long u = u_begin; u < u_end;) {
        const long long tile = u / nstage, st_begin = u - tile * nstage;
        const long long st_end = (u_end - u) < (nstage - st_begin) ? st_begin + (u_end - u) : nstage;
        u += st_end - st_begin;
        const bool partial = (st_end - st_begin) != nstage;
        long long mi, ni;
        grm_tile_of(tile, mi, ni);
        const int mtv = grm_mtv(p.n, mi, wm);
        double acc[MT][NTL][2];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

        for (long long st = st_begin; st < st_end; ++st) {
            mbar_wait(&S.full[stage], fphase);
            if (mtv == 8) grm_stage_mma<8, NTL>(acc, S.cA[stage], S.cB[stage], S.lut[stage], wm, wn, t, sh0, sh1);
            else if (mtv == 6) grm_stage_mma<6, NTL>(acc, S.cA[stage], S.cB[stage], S.lut[stage], wm, wn, t, sh0, sh1);
            else if (mtv == 4) grm_stage_mma<4, NTL>(acc, S.cA[stage], S.cB[stage], S.lut[stage], wm, wn, t, sh0, sh1);
            else if (mtv == 2) grm_stage_mma<2, NTL>(acc, S.cA[stage], S.cB[stage], S.lut[stage], wm, wn, t, sh0, sh1);
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.empty[stage]);
            if (++stage == GR_NS) { stage = 0; fphase ^= 1u; }
        }

        // ---- epilogue ---------------------------------------------------------------------------
#pragma unroll
        for (int a = 0; a < MT; ++a) {
            const long long r = mi * GR_T + wm * 64 + a * 8 + g;
#pragma unroll
            for (int b = 0; b < NTL; ++b) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const long long col = ni * GR_T + wn * 32 + b * 8 + 2 * t + c;
                    if (r < p.n && col < p.n && (mi != ni || col <= r)) {
                        double v = acc[a][b][c];
                        if (!p.accumulate) {
                            v = v / p.scale;
                            if (col == r) v = v + v * p.small_val;       // gmatrix.py:64-66
                        }
                        if (partial || p.accumulate) {                    // output is zero-initialised / accumulated into
                            atomicAdd(&p.K[r * p.n + col], v);
                            if (col != r) atomicAdd(&p.K[col * p.n + r], v);
                        } else {
                            p.K[r * p.n + col] = v;
                            if (col != r) p.K[col * p.n + r] = v;
                        }
                    }
                }
            }
        }
    }
}

}  // namespace gb

using namespace gb;

extern "C" int gmat_b200_grm(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                             const double *lut_dev, long long snp_begin, long long snp_end, double scale,
                             double small_val, int accumulate, double *K_dev, void *stream) {
    GB_CHECK_ARG(bed_dev && lut_dev && K_dev, "NULL pointer");
    GB_CHECK_ARG(n > 0 && m > 0 && n < (1LL << 30), "sizes");
    GB_CHECK_ARG(row_bytes >= (n + 3) / 4 && row_bytes % 16 == 0, "row_bytes");
    GB_CHECK_ARG((reinterpret_cast<uintptr_t>(bed_dev) & 15) == 0, "bed_dev alignment");
    GB_CHECK_ARG(0 <= snp_begin && snp_begin <= snp_end && snp_end <= m, "snp range");
    GB_CHECK_ARG(accumulate || scale != 0.0, "scale");
    GrmParams p;
    p.bed = bed_dev; p.n = n; p.m = m; p.row_bytes = row_bytes; p.lut = lut_dev;
    p.snp_begin = snp_begin; p.snp_end = snp_end; p.scale = scale; p.small_val = small_val;
    p.accumulate = accumulate; p.K = K_dev;
    p.nblk = (n + GR_T - 1) / GR_T;
    auto kern = grm_kernel;
    const size_t smem = sizeof(GrmSmem);
    GB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long ntiles = p.nblk * (p.nblk + 1) / 2;
    const long long nstage = (snp_end - snp_begin + GR_GS - 1) / GR_GS;
    if (nstage == 0) {                               // empty SNP range: K = 0 (or unchanged when accumulating)
        if (!accumulate) GB_CUDA(cudaMemsetAsync(K_dev, 0, sizeof(double) * (size_t)n * (size_t)n, (cudaStream_t)stream));
        return GMAT_B200_OK;
    }
    // stream-K grid.  ntiles >= SMs: one CTA per SM, equal contiguous unit ranges (a range is at least one tile
    // long, so a tile has at most two parts and the two partial sums commute: reproducible).  2*ntiles <= SMs: two
    // CTAs per tile, half the stages each.  In between: one CTA per tile (a third part would make the atomic sum
    // order-dependent).
    long long grid = sm_count();
    p.split2 = 0;
    if (ntiles < grid) {
        if (2 * ntiles <= grid && nstage >= 2) { grid = 2 * ntiles; p.split2 = 1; }
        else { grid = ntiles; p.split2 = 2; }
    }
    if (!accumulate) GB_CUDA(cudaMemsetAsync(K_dev, 0, sizeof(double) * (size_t)n * (size_t)n, (cudaStream_t)stream));
    kern<<<(unsigned)grid, (GR_CW + GR_PW) * 32, smem, (cudaStream_t)stream>>>(p);
    GB_LAUNCH_CHECK();
    return GMAT_B200_OK;
}
