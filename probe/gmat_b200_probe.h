/* gmat_b200_probe.h -- measurement scaffolding (probe/libgmat_b200_probe.so), NOT part of the product C-ABI.
 * Issue-rate microbenchmarks used as roofline denominators by bench.py and scripts/. */
#ifndef GMAT_B200_PROBE_H
#define GMAT_B200_PROBE_H
#ifdef __cplusplus
extern "C" {
#endif
const char *gmat_b200_last_error(void);
int         gmat_b200_set_device(int device);
/* int8 tensor-core issue-rate microbenchmark (tcgen05.mma kind::i8 128x256x32 on resident tiles, all SMs): TOPS  */
int gmat_b200_i8_peak(int iters, double *tops);
/* the same for the FP4 MMA of the GRM (kind::mxf4.block_scale 128x256x64, unit scale factors): TOPS                */
int gmat_b200_f4_peak(int iters, double *tops);
/* FP64 issue-rate microbenchmarks (used by bench.py for the roofline denominator).
 * kind 0: DMMA (mma.sync m8n8k4 f64) ; kind 1: DFMA.  Writes TFLOP/s.  */
int gmat_b200_fp64_peak(int kind, int iters, double *tflops);
/* DMMA issue-model probe: `ctas_per_sm` CTAs of `warps` warps per SM, each warp an (ia x ib) register tile of
 * independent m8n8k4 accumulators with distinct A/B fragments, no memory traffic.  Writes TFLOP/s.           */
int gmat_b200_dmma_probe(int warps, int ia, int ib, int ctas_per_sm, int iters, double *tflops);
/* same, with every k-step's fragments re-loaded from shared memory (shape 0: 4x4 tile, B as LDS.64; 1: 4x4, B as
 * LDS.128; 2: 8x4, LDS.64; 3: 8x4, LDS.128; 4: 4x2, LDS.128): the LDS + DMMA issue mix without any barrier.   */
int gmat_b200_dmma_lds_probe(int warps, int shape, int ctas_per_sm, int iters, double *tflops);
/* 8x4 tile (32 DMMAs per step) + nmul DMULs per step; mode 0 grouped/independent, 1 spread/independent,
 * 2 grouped and feeding the step's B fragments.  Writes the DMMA-only TFLOP/s.                               */
int gmat_b200_dmma_dmul_probe(int warps, int nmul, int mode, int iters, double *tflops);
#ifdef __cplusplus
}
#endif
#endif
