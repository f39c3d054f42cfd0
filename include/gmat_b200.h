/*
 * gmat_b200.h -- C-ABI of the B200-native REMMAX pair-scan / GRM engine.
 *
 * Plain C, plain pointers and sizes, no torch / CUDA types in any signature.  Two groups:
 *
 *  (1) DROP-IN symbols: the seven functions GMAT's own cffi extension modules export
 *      (`_cremma_epi_eff_cpu`, gmat/remma/_build.py:8-31; `_cread_plink_bed`,
 *      gmat/process_plink/_build.py:8).  Same names, argument order, types and file side effects,
 *      HOST pointers, so the reference Python (`lib.remma_epiAA_eff_cpu(...)`,
 *      gmat/remma/remma_epiAA/remma_epiAA_eff.py:80) can bind this library unchanged.
 *      Difference on purpose: on failure they return a negative status (see gmat_b200_last_error)
 *      instead of calling exit(1) (gmat/remma/_remma_epi_eff_cpu.c:18-21,115-118).
 *
 *  (2) gmat_b200_* symbols: work the reference does in numpy (exact test, GRM, decode, coding),
 *      operating on DEVICE pointers owned by the caller (torch tensors: tensor.data_ptr()) on a
 *      caller-supplied CUDA stream (cudaStream_t passed as void*; NULL = default stream).
 *      No callee-owned memory crosses the boundary; hit buffers are caller-allocated with an
 *      atomic counter.  All functions return GMAT_B200_OK (1) or a negative error.
 *
 * Genotype store layout (device): `bed` = the PLINK1 SNP-major payload (file bytes after the 3-byte
 * magic), one row per SNP, `row_bytes` >= ceil(n/4) bytes apart (pad rows to a multiple of 16 bytes
 * for aligned vector loads), 4 individuals per byte, low bits first, 2-bit code c:
 *   00 -> 0, 01 -> missing, 10 -> 1, 11 -> 2   (count of the .bim A2 allele;
 *   gmat/process_plink/_read_plink_bed.c:36-37: value = (c*c+c)/6).
 * A coded column (additive x = g-2p, dominance d = h-2p(1-p)) is never materialised: it is the
 * 4-entry table lut[snp][c] (double[m][4]) applied to the 2-bit codes inside the kernels.
 */
#ifndef GMAT_B200_H
#define GMAT_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define GMAT_B200_OK              1
#define GMAT_B200_ERR_ARG        -1   /* bad argument */
#define GMAT_B200_ERR_IO         -2   /* cannot open/read/write a file, bad .bed magic */
#define GMAT_B200_ERR_CUDA       -3   /* CUDA runtime error (no device, launch failure, OOM) */
#define GMAT_B200_ERR_OVERFLOW   -4   /* hit buffer too small: count > cap (count is still exact) */

/* coding selectors for gmat_b200_build_lut */
#define GMAT_B200_CODING_RAW      0   /* 0, 1/3(missing), 1, 2          _read_plink_bed.c:36-37 */
#define GMAT_B200_CODING_ADD      1   /* g - 2p                          remma_epiAA.py:59-61    */
#define GMAT_B200_CODING_DOM      2   /* h - 2p(1-p), h = g with 2 -> 0  remma_epiDD.py:59-65    */
/* how p is formed */
#define GMAT_B200_FREQ_NUMPY      0   /* p = (sum g) / (2n), sum exact       remma_epiAA.py:59   */
#define GMAT_B200_FREQ_CLOOP      1   /* p += g_k/(2n) term by term in k     _remma_epi_eff_cpu.c:103-107 */

/* flags for gmat_b200_epi_eff */
#define GMAT_B200_EFF_GE          1   /* emit when |eff| >= cut (first AD direction, _remma_epi_eff_cpu.c:238) */
#define GMAT_B200_EFF_SWAP        2   /* report (j,i) instead of (i,j)   (second AD direction, :245-247)       */

/* pair-space selectors for the exact scan */
#define GMAT_B200_PAIRS_UPPER     0   /* j in (i, m)   remma_epiAA.py:72, remma_epiDD.py:76 */
#define GMAT_B200_PAIRS_ALL       1   /* j in [0, m)   remma_epiAD.py:77 (incl. j == i)      */

typedef struct gmat_b200_hit {
    int    snp_0;      /* first column of the reference output files  */
    int    snp_1;
    double eff;        /* w' Z'Py                                     */
    double var;        /* w' Z'PZ w   (0 for the approximate screen)  */
    double chi;        /* eff^2/var   (0 for the approximate screen)  */
} gmat_b200_hit;       /* 32 bytes */

/* ------------------------------------------------------------------ runtime / errors ------- */
int         gmat_b200_version(void);
const char *gmat_b200_last_error(void);            /* thread-local, valid until the next call   */
/* which engine the last drop-in screen symbol (group 1) ran on this thread, and why when the int8 path was not
 * taken: e.g. "int8 tensor-core prefilter + FP64 re-test (3 base-32 planes: missing calls)" or "FP64 DMMA kernel
 * (margin not small against the cut)".  The fast paths never narrow silently.                                   */
const char *gmat_b200_last_engine_note(void);
int         gmat_b200_device_count(void);          /* >= 0, or negative error                   */
int         gmat_b200_set_device(int device);
int         gmat_b200_device_info(int *sm_count, int *cc_major, int *cc_minor, long long *hbm_bytes);
long long   gmat_b200_kernel_launches(void);       /* kernels launched by this library so far   */

/* ------------------------------------------------------------------ (1) drop-in symbols ---- */
/* gmat/process_plink/_read_plink_bed.c:5 -- marker_mat is HOST double[num_snp][num_id]        */
int read_plink_bed(char *bed_file, long long num_id, long long num_snp, double *marker_mat);

/* gmat/remma/_remma_epi_eff_cpu.c:91,260,443 -- all pointers HOST; writes text file out_file   */
int remma_epiAA_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                        long long len_snp_lst_0, double *pymat, double eff_cut, char *out_file);
int remma_epiAD_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                        long long len_snp_lst_0, double *pymat, double eff_cut, char *out_file);
int remma_epiDD_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                        long long len_snp_lst_0, double *pymat, double eff_cut, char *out_file);
/* gmat/remma/_remma_epi_eff_cpu.c:171,354,526 -- eff_cut is a 111-entry table indexed by
 * freq[snp_lst_0[i]]*10 + freq[j] (:154)                                                       */
int remma_epiAA_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                            long long len_snp_lst_0, double *pymat, long long *freq, double *eff_cut,
                            char *out_file);
int remma_epiAD_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                            long long len_snp_lst_0, double *pymat, long long *freqA, long long *freqD,
                            double *eff_cut, char *out_file);
int remma_epiDD_maf_eff_cpu(char *bed_file, long long num_id, long long num_snp, long long *snp_lst_0,
                            long long len_snp_lst_0, double *pymat, long long *freq, double *eff_cut,
                            char *out_file);

/* ------------------------------------------------------------------ (2) device-pointer API -- */
/* Read "<bed_file>.bed" (prefix without extension, like the reference), validate the magic
 * (6c 1b 01), and return the payload size; with host_buf != NULL also copy the payload into it
 * re-strided to row_bytes (zero padded).  Host-side I/O helper, no GPU work.                    */
int gmat_b200_bed_read_host(const char *bed_file, long long num_id, long long num_snp,
                            long long row_bytes, unsigned char *host_buf);

/* per-SNP genotype counts: counts[snp][c] = number of individuals k < n with code c            */
int gmat_b200_bed_stats(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                        int *counts_dev /* [m][4] */, void *stream);

/* per-SNP value tables for a coding; freq_dev (optional, may be NULL) receives p[m]            */
int gmat_b200_build_lut(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                        const int *counts_dev, int coding, int freq_mode,
                        double *lut_dev /* [m][4] */, double *freq_dev /* [m] or NULL */, void *stream);

/* materialise doubles: out[snp][k] = lut[snp][code]  (lut_dev == NULL: raw (c*c+c)/6, i.e. the
 * exact output of read_plink_bed).  HBM-bound: 0.25 B read + 8 B written per genotype.          */
int gmat_b200_decode(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                     const double *lut_dev, double *out_dev /* [m][n] */, void *stream);

/* inverse of decode for imputed matrices: geno[snp][k] in {0,1,2} (NaN/other -> missing)        */
int gmat_b200_pack(const double *geno_dev /* [m][n] */, long long n, long long m, long long row_bytes,
                   unsigned char *bed_dev_out, void *stream);

/* Approximate screen (gmat/remma/_remma_epi_eff_cpu.c:61-86): for i in ilist, j in (i, m):
 * eff = sum_k lutU[i][c_ik] * lutV[j][c_jk] * py[k]; emit a hit when |eff| > cut (>= with
 * GMAT_B200_EFF_GE), cut = eff_cut or, when cut_table_dev != NULL,
 * cut_table[binI[i]*10 + binJ[j]] (111 doubles).  hits beyond cap are counted, not stored.     */
int gmat_b200_epi_eff(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                      const double *lutU_dev, const double *lutV_dev, const double *py_dev,
                      const long long *ilist_dev, long long n_i, int flags, double eff_cut,
                      const double *cut_table_dev, const int *binI_dev, const int *binJ_dev,
                      gmat_b200_hit *hits_dev, long long cap, unsigned long long *count_dev,
                      void *stream);

/* The same screen with the INT8 tensor cores deciding which pairs are worth the FP64 arithmetic
 * (gmat_b200/csrc/screen_i8.cu).  Stage 1: eff = Q - o_j t_i - o_i (t_j - o_j P) with Q_ij = sum_k a_ik a_jk py_k
 * evaluated for every pair from S exact int8 GEMMs (py cut into S signed digits, tcgen05.mma cta_group::2 kind::i8)
 * in FP32; pairs with |eff~| >= cut - margin survive, margin a rigorous bound of everything stage 1 drops.
 * Stage 2: eff of the survivors in FP64 from the LUT values (the arithmetic of gmat_b200_epi_eff) and the reference's
 * threshold -- so the reported set and the reported values are FP64 results.
 * `tripled` selects the integer coding: 0 = the store has no missing call (a in {0,1,2} / {0,1,0}, S = 2 base-128
 * digits); 1 = missing calls present: the C screen does not impute, code 01 is the value 1/3
 * (_remma_epi_eff_cpu.c:42), so both operands carry 3a in {0,1,3,6} / {0,1,3,0} and py is cut into S = 3 base-32
 * digits (the products must stay int8).
 *
 * Prepared form (no host synchronisation inside; operands are caller-owned and can be kept across calls, the two
 * directions of the AD screen and the stages of the _approx pipelines):
 *   gmat_b200_screen_operand  g8[m][ld] integer codes of one coding (ld = gmat_b200_screen_ld(n)), t[m] = sum_k a_k py_k
 *   gmat_b200_screen_py       dig[3][ld] digits of Z'Py, stats[4] = {max|py|, sum py, ..}
 *   gmat_b200_epi_eff_i8p     the screen; cand_dev [cand_cap][2] ints + cand_count_dev: the survivor list of stage 1;
 *                             status_dev[0] after the call: 0 = done on the int8 path (hits/count as in
 *                             gmat_b200_epi_eff), 1 = the margin is not small against the cut (e.g. eff_cut = 0),
 *                             2 = more survivors than cand_cap -- in both cases nothing was emitted and the caller
 *                             runs gmat_b200_epi_eff (or retries with a larger candidate buffer).  The caller reads
 *                             status together with the hit count: one synchronisation per call, made by the caller. */
long long gmat_b200_screen_ld(long long n);
int gmat_b200_screen_operand(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                             int coding, int tripled, const double *py_dev, signed char *g8_dev /* [m][ld] */,
                             double *t_dev /* [m] */, void *stream);
int gmat_b200_screen_py(const double *py_dev, long long n, int tripled, signed char *dig_dev /* [3][ld] */,
                        double *stats_dev /* [4] */, void *stream);
int gmat_b200_epi_eff_i8p(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                          const double *lutU_dev, const double *lutV_dev, const signed char *g8U_dev,
                          const signed char *g8V_dev, const double *tU_dev, const double *tV_dev,
                          const signed char *dig_dev, const double *stats_dev, int tripled,
                          const double *py_dev, const long long *ilist_dev, long long n_i, int flags,
                          double eff_cut, const double *cut_table_dev, const int *binI_dev,
                          const int *binJ_dev, gmat_b200_hit *hits_dev, long long cap,
                          unsigned long long *count_dev, int *cand_dev, long long cand_cap,
                          unsigned long long *cand_count_dev, int *status_dev, void *stream);
/* One-shot form: builds the operands in stream-ordered scratch, runs the prepared screen, reads the status back (it
 * synchronises the stream once) and runs gmat_b200_epi_eff by itself when the int8 path declined; *engine_used
 * (optional) = 1 when the int8 path produced the result, 0 after a fallback.                                    */
int gmat_b200_epi_eff_i8(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                         const double *lutU_dev, const double *lutV_dev, int codingU, int codingV, int tripled,
                         const double *py_dev, const long long *ilist_dev, long long n_i, int flags, double eff_cut,
                         const double *cut_table_dev, const int *binI_dev, const int *binJ_dev,
                         gmat_b200_hit *hits_dev, long long cap, unsigned long long *count_dev,
                         int *engine_used, void *stream);

/* Z'PZ re-tiled for the exact kernel: lower block-triangle, 128-row blocks, DMMA fragment order.
 * gmat_b200_pvp_packed_elems(n) doubles are needed.                                            */
long long gmat_b200_pvp_packed_elems(long long n);
/* tile shape of the exact kernel (rows of A per block, pairs per tile, resident CTAs per SM); informational */
int gmat_b200_exact_config(int *bm, int *bn, int *ctas_per_sm);
int gmat_b200_pack_pvp(const double *pvp_dev /* [n][n] row-major, symmetric */, long long n,
                       double *packed_dev, void *stream);

/* Exact test over a scan (remma_epiAA.py:71-82): for i in ilist, j in (i,m) or [0,m):
 * w = lutU[i][c_i.] o lutV[j][c_j.]; eff = w'py; var = w' Z'PZ w; chi = eff^2/var;
 * emit when chi > chi_cut (strict) -- or every pair when chi_cut < 0.                           */
int gmat_b200_epi_exact_scan(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                             const double *lutU_dev, const double *lutV_dev, const double *py_dev,
                             const double *pvp_packed_dev, const long long *ilist_dev, long long n_i,
                             int pair_space, double chi_cut,
                             gmat_b200_hit *hits_dev, long long cap, unsigned long long *count_dev,
                             void *stream);

/* Allocation-free variant of gmat_b200_epi_exact_scan.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same results.
 * The scan's only scratch (its tile table: 2*ceil(n_i/bi)+1 int64) lives in a file-scope
 * static __device__ array sized for GMAT_B200_NOALLOC_MAX_FIRST first SNPs (1 MiB + 8 bytes, held
 * by every GPU that loads the library until exit); the call makes no cudaMalloc / cudaFree /
 * cudaMallocAsync.  n_i > GMAT_B200_NOALLOC_MAX_FIRST returns GMAT_B200_ERR_ARG with a message in
 * gmat_b200_last_error().  All calls on a device share the array (it keeps the last call's
 * contents; every entry a call reads is rewritten by that call first, nothing needs zeroing):
 * calls whose kernels OVERLAP on one GPU (two streams, two threads) race -- serialise them.
 * As for gmat_b200_epi_exact_scan, the caller zeroes *count_dev.                                */
#define GMAT_B200_NOALLOC_MAX_FIRST 65536
long long gmat_b200_noalloc_max_first(void);
int gmat_b200_epi_exact_scan_noalloc(const unsigned char *bed_dev, long long n, long long m,
                                     long long row_bytes, const double *lutU_dev, const double *lutV_dev,
                                     const double *py_dev, const double *pvp_packed_dev,
                                     const long long *ilist_dev, long long n_i, int pair_space,
                                     double chi_cut, gmat_b200_hit *hits_dev, long long cap,
                                     unsigned long long *count_dev, void *stream);

/* Exact test over an explicit pair list (remma_epiAA_pair.py:66-87): out[p] gets eff/var/chi of
 * pair p (all pairs are written, dense, in list order).                                        */
int gmat_b200_epi_exact_pairs(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                              const double *lutU_dev, const double *lutV_dev, const double *py_dev,
                              const double *pvp_packed_dev, const int *pairs_dev /* [n_pairs][2] */,
                              long long n_pairs, gmat_b200_hit *out_dev, void *stream);

/* GRM (gmat/gmatrix/gmatrix.py:63-66,127-130): K = X X' / scale with X[k][snp] = lut[snp][c],
 * then diag(K) += diag(K)*small_val.  K_dev is [n][n] row-major, fully (symmetrically) written.
 * snp range [snp_begin, snp_end) lets ranks own K-slices (partial sums, scale applied by caller
 * when accumulate != 0: K_dev += partial, no scale, no diagonal inflation).                    */
int gmat_b200_grm(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                  const double *lut_dev, long long snp_begin, long long snp_end,
                  double scale, double small_val, int accumulate, double *K_dev, void *stream);

/* ---- INT8 tensor-core engine of the exact test (tcgen05.mma kind::i8; see gmat_b200/csrc/ozaki.cu) --------
 * Same statistics as gmat_b200_epi_exact_scan (remma_epiAA.py:72-82).  With v_j = a_j - o_j (a_j the integer
 * dosage or heterozygote indicator of the second SNP) and B_i = diag(u_i) Z'PZ diag(u_i):
 *   var_ij = a_j' B_i a_j - 2 o_j (a_j . r_i) + o_j^2 s_i ,  r_i = u_i o (Z'PZ u_i), s_i = sum r_i
 *   eff_ij = a_j . (u_i o Z'Py) - o_j (u_i . Z'Py)
 * a_j' B_i a_j is evaluated on the INT8 tensor cores from an error-free split of B_i/scale_i into n_planes signed
 * base-256 digit planes (exact int8 x int8 -> int32 products, recombined in FP64; S planes = 8S-1 bits of
 * max|B_i|, n_planes in [2, 7]).  a_j.r_i and a_j.(u_i o Z'Py) come out of the same MMAs: r_i and u_i o Z'Py are
 * split into 2S digits each (high and low part: no loss against FP64) and stored as four extra rows of every plane.
 * Every step below is a kernel of this library (no library GEMM, no host arithmetic); all calls are asynchronous on
 * `stream`.  A SNP with a missing call has NaN statistics in the reference's numpy path: such second SNPs are flagged
 * offs = +inf (eff = var = chi = NaN), monomorphic ones offs = NaN (eff = var = 0, chi = NaN), by gmat_b200_i8_offs. */
long long gmat_b200_i8_ld(long long n);      /* bytes per row AND rows per plane of the int8 operands: n rounded up
                                              * to 4, plus the 4 extra rows, rounded up to 256                   */
/* integer code of every SNP: g8[snp][k] = dosage {0,1,2} (ADD) or heterozygote indicator (DOM), 0 for a missing
 * call and for k >= n; ld a multiple of 16, >= n.  g2_dev (optional, [m][ld/4]): the same values packed 2 bits each,
 * low bits first (what the scan kernel's epilogue reads).
 * counts_dev (optional, [m][4] genotype counts; what the exact engine passes): an additive SNP with more 11 than 00
 * homozygotes is coded by the OTHER allele, a' = 2 - a.  v'_j = a'_j - (2 - o_j) = -v_j, so var_ij is unchanged and
 * eff_ij changes sign (gmat_b200_i8_offs marks the SNP, the scan restores the sign); sum_k a'^2 -- which the error of
 * the digit-plane split scales with -- is up to 35x smaller.  NULL: the plain dosage for every SNP.              */
int gmat_b200_i8_dosage(const unsigned char *bed_dev, long long n, long long m, long long row_bytes, int coding,
                        const int *counts_dev /* [m][4] or NULL */, signed char *g8_dev /* [m][ld] */,
                        unsigned char *g2_dev /* [m][ld/4] or NULL */, long long ld, void *stream);
/* per second SNP: offs[j] = o_j = -lutV[j][0] with the two flags above, |asq[j]| = sum_k a_jk^2 (from the counts);
 * for an additive SNP coded by the other allele (see gmat_b200_i8_dosage with counts) offs[j] = 2 - o_j, the sums are
 * those of a' = 2 - a, and asq[j] carries a NEGATIVE sign as the mark the scan kernel reads                       */
int gmat_b200_i8_offs(const int *counts_dev /* [m][4] */, const double *lutV_dev /* [m][4] */, long long m,
                      int coding, int minor_allele /* 1 when gmat_b200_i8_dosage was given the counts, else 0 */,
                      double *offs_dev /* [m] */, double *asq_dev /* [m] */, void *stream);
/* *out_dev = max |x[k]| (NaN if any x is NaN)                                                                    */
int gmat_b200_absmax(const double *x_dev, long long len, double *out_dev, void *stream);
/* Per batch of n_i first SNPs (ilist_dev): u[ii][k] = lutU[i][code] (coded first SNPs), r = u o (Z'PZ u)
 * (scratch, [n_i][n]), the side table side[ii][8] = {scale_i, s_i, u_i.Z'Py, scale_r, scale_e, guard unit, 0, 0}
 * and the four extra rows (n rounded up to 4, +0..3) of every plane of every first SNP.  amax_dev = max|Z'PZ|
 * (gmat_b200_absmax).  planes_dev: [n_i][n_planes][ld][ld] int8, ZERO-initialised by the caller once.
 * ilist_dev == NULL: u_dev already holds the n_i vectors (bed_dev / lutU_dev unused) -- the pair-list engine below
 * passes one vector of ones and gets the planes of Z'PZ itself, r = Z'PZ 1, side = {max|Z'PZ|/0.49, 1'Z'PZ 1, 1.Z'Py, ..}. */
int gmat_b200_i8_first(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                       const double *lutU_dev, const double *py_dev, const double *pvp_dev /* [n][n] */,
                       const double *amax_dev, const long long *ilist_dev, long long n_i, int n_planes,
                       double *u_dev /* [n_i][n] */, double *r_dev /* [n_i][n] */, double *side_dev /* [n_i][8] */,
                       signed char *planes_dev, void *stream);
/* rows 0..n-1 of every plane: the digits of B_i / side[ii][0], lower block-triangle of 128 x 128 blocks, diagonal
 * blocks halved (blocks above the block diagonal are never written and stay zero)                              */
int gmat_b200_i8_split(const double *pvp_dev /* [n][n] */, long long n, const double *u_dev, long long n_i,
                       const double *side_dev, int n_planes, signed char *planes_dev, void *stream);
/* Work items of one launch are (first SNP ii, tile of 256 second SNPs, computed by a CTA PAIR -- cluster of 2,
 * tcgen05 cta_group::2): item_prefix_dev[n_i+1] = first item of every first SNP, jt0_dev[n_i] = its first tile
 * (UPPER: (i+1)/256, ALL: 0), total_items = item_prefix[n_i]; all device arrays, so that successive launches queue
 * without a host synchronisation.  Hits are appended at *count_dev as in gmat_b200_epi_exact_scan (count keeps
 * counting beyond cap).  A-posteriori guard: a pair whose variance is so small against scale_i that the digits the
 * split dropped could matter (side[ii][5] * asq[j] > recheck_tol * |var|) is NOT decided here; it is appended to
 * recheck_dev ([recheck_cap][2] ints, counter *recheck_count_dev, counting beyond the cap) for the caller to
 * evaluate with gmat_b200_epi_exact_pairs.  q_debug_dev (optional, [n_i][m]) receives a_j' B_i a_j.
 * variant 3 issues only the plane rows that can be non-zero (last row tile trimmed, the zero 128 x 128 block right
 * of the diagonal skipped) -- what the engine uses; variant 2 issues full 256-row MMAs: same results bit for bit. */
int gmat_b200_i8_scan(const signed char *g8_dev, const unsigned char *g2_dev, long long n, long long m,
                      const signed char *planes_dev, int n_planes, const long long *ilist_dev, long long n_i,
                      const int *item_prefix_dev, const int *jt0_dev, long long total_items,
                      const double *side_dev, const double *offs_dev, const double *asq_dev,
                      int pair_space, double chi_cut, gmat_b200_hit *hits_dev, long long cap,
                      unsigned long long *count_dev, int *recheck_dev, long long recheck_cap,
                      unsigned long long *recheck_count_dev, double recheck_tol, double *q_debug_dev,
                      int variant, void *stream);

/* ---- exact test of an explicit PAIR LIST on the int8 tensor cores (remma_epiAA_pair.py:66-87) -------------------
 * A list's pairs rarely share a first SNP, so the digit planes are those of Z'PZ itself (gmat_b200_i8_first with
 * ilist_dev == NULL and u = 1, then gmat_b200_i8_split with n_i = 1: once per model) and the operand rows are integer
 * vectors per pair.  With a, b the integer codes of the two SNPs, q = a o b and o_i, o_j the offsets (-lut[.][0]):
 *   w = (a - o_i) o (b - o_j) = q - o_j a - o_i b + o_i o_j 1,
 * so w'Z'PZ w needs the bilinear forms of q, a, b and 1.  Those with 1 are dot products with t1 = Z'PZ 1; the others
 * follow from quadratic forms by polarisation, x'Ay = (Q(x+y) - Q(x) - Q(y))/2.  Operand rows of a chunk of n_pairs
 * pairs: 4p .. 4p+3 = q, q+a, q+b, a+b of pair p; then one row per DISTINCT first SNP (a, nU rows, snpU_dev) and per
 * distinct second SNP (b, nV rows, snpV_dev).  gmat_b200_i8_pair_rows builds them as int8 (x8: the MMA operand,
 * [rows][ld]) and packed 4 bits each (x4: [rows][ld/2], the epilogue's copy; values 0..6) together with
 * asq[row] = sum x^2 and xt1[row] = x . t1.  gmat_b200_i8_rows_scan runs the scan kernel over the rows:
 * q[row] = x'Z'PZ x, e[row] = x . Z'Py (zeros_dev: [rows] zeros; ilist_dev = {0}, item_prefix_dev = {0, ceil(rows/256)},
 * jt0_dev = {0}; count_dev / recheck_count_dev stay 0).  gmat_b200_i8_pairs_combine writes eff, var, chi of every pair
 * (invU/invV[p] = index of the pair's SNPs in snpU/snpV) and weak[p] = 1 where the split error of the Q's, weighted by
 * their coefficients, exceeds recheck_tol * |var|: those pairs are for gmat_b200_epi_exact_pairs.  n < 14000.       */
int gmat_b200_i8_pair_rows(const unsigned char *bed_dev, long long n, long long row_bytes, int codingU, int codingV,
                           const int *pairs_dev /* [n_pairs][2] */, long long n_pairs, const int *snpU_dev,
                           long long nU, const int *snpV_dev, long long nV, const double *t1_dev /* [n] */,
                           signed char *x8_dev, unsigned char *x4_dev, double *asq_dev, double *xt1_dev,
                           void *stream);
int gmat_b200_i8_rows_scan(const signed char *x8_dev, const unsigned char *x4_dev, long long n, long long rows,
                           const signed char *planes_dev, int n_planes, const double *side_dev,
                           const double *zeros_dev, const double *asq_dev, const long long *ilist_dev,
                           const int *item_prefix_dev, const int *jt0_dev, double *q_dev, double *e_dev,
                           unsigned long long *count_dev, unsigned long long *recheck_count_dev, void *stream);
int gmat_b200_i8_pairs_combine(const int *pairs_dev, long long n_pairs, const int *invU_dev, long long nU,
                               const int *invV_dev, const double *lutU_dev, const double *lutV_dev,
                               const double *q_dev, const double *e_dev, const double *xt1_dev,
                               const double *asq_dev, const double *side_dev, double recheck_tol,
                               gmat_b200_hit *out_dev, unsigned char *weak_dev, void *stream);

/* The same GRM on the INT8 tensor cores (gmat_b200/csrc/gram_i8.cu): X X' = G G' - c 1' - 1 c' + d with G the
 * integer dosage / heterozygote indicator matrix, so the n^2 m term is ONE exact int8 x int8 -> int32 GEMM
 * (tcgen05.mma cta_group::2 kind::i8); c_k = sum_s o_s a_ks and d = sum_s o_s^2 in FP64.  Same arguments and
 * semantics as gmat_b200_grm plus the coding (GMAT_B200_CODING_ADD / _DOM) the LUT was built for; the SNP range
 * must not contain a missing genotype (callers use gmat_b200_grm then).  Scratch (n x m bytes + 4 n^2 bytes)
 * comes from the stream-ordered pool.                                                                          */
int gmat_b200_grm_i8(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                     const double *lut_dev, int coding, long long snp_begin, long long snp_end,
                     double scale, double small_val, int accumulate, double *K_dev, void *stream);

/* The same GRM with FP4 operands (gmat_b200/csrc/gram_i8.cu, f4_gram_pair_kernel): the entries of G are {0, 1, 2},
 * all exactly representable in E2M1, so G G' is one exact tcgen05.mma kind::mxf4 GEMM (unit block scales, FP32
 * accumulation exact below 2^24) at twice the int8 MMA rate on half the operand bytes; c_k = sum_s o_s a_ks comes out
 * of the same GEMM from 18 extra rows holding the base-8 digits of o_s.  Same arguments and semantics as
 * gmat_b200_grm_i8; the SNP range must not contain a missing genotype and must be shorter than 2^21 SNPs (callers use
 * gmat_b200_grm_i8 / gmat_b200_grm beyond that).                                                                 */
int gmat_b200_grm_f4(const unsigned char *bed_dev, long long n, long long m, long long row_bytes,
                     const double *lut_dev, int coding, long long snp_begin, long long snp_end,
                     double scale, double small_val, int accumulate, double *K_dev, void *stream);

/* Host text writer, byte-identical to np.savetxt(path, mat) as used by output_mat (gmat/gmatrix/gmatrix.py:10-14:
 * '%.18e', ' ' delimiter, '\n'): rows formatted in parallel by `threads` host threads (<= 0: all).  No GPU work. */
int gmat_b200_savetxt_host(const char *path, const double *mat /* HOST [rows][cols] */, long long rows,
                           long long cols, int threads);
/* The two list formats of output_mat (gmat/gmatrix/gmatrix.py:15-29: pandas to_csv of the lower triangle in
 * np.tril_indices order, sep ' ', no header): "<row> <col> <val>" with 1-based indices when ids == NULL
 * ('row_col_val'), "<id_r> <id_c> <val>" otherwise ('id_id_val'; ids = the n labels as pandas prints them,
 * concatenated, label k = ids[id_off[k] .. id_off[k+1])).  <val> is repr(float) -- shortest round-trip digits in
 * CPython's layout, NaN as the empty string -- so the files are byte-identical to the reference's; rows are
 * formatted in parallel by `threads` host threads (<= 0: all).  No GPU work.                                   */
int gmat_b200_write_tril_host(const char *path, const double *mat /* HOST [n][n] */, long long n,
                              const char *ids, const long long *id_off, int threads);
/* repr(float) of one value into out32 (NUL-terminated, <= 32 bytes): the formatter of the list writer, exported so
 * that it can be pinned against CPython on its own.                                                            */
int gmat_b200_repr_host(double v, char *out32);

#ifdef __cplusplus
}
#endif
#endif /* GMAT_B200_H */
