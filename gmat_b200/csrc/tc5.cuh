// tcgen05 / TMEM / TMA building blocks shared by the int8 tensor-core kernels (ozaki.cu, gram_i8.cu): PTX wrappers,
// the K-major 128-byte-swizzle shared-memory matrix descriptor, instruction descriptors, tensor-map creation.
#pragma once
#include "common.cuh"
#include <cuda.h>

namespace gb {

constexpr int I8_TJ = 128;        // second SNPs per tile  (UMMA M, TMEM lanes)
constexpr int I8_TK = 256;        // plane rows per tile   (UMMA N, TMEM columns)
constexpr int I8_SLAB = 128;      // ids per pipeline stage = one 128-byte swizzle atom of int8
constexpr int I8_NS = 4;          // smem ring stages
constexpr int I8_A_BYTES = I8_TJ * I8_SLAB;            // 16 KB genotype tile
constexpr int I8_B_BYTES = I8_TK * I8_SLAB;            // 32 KB plane tile
constexpr int I8_STAGE_BYTES = I8_A_BYTES + I8_B_BYTES;

// ------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}
// one lane of a CONVERGED warp (elect.sync): the single-thread tcgen05 instructions are issued under this predicate
// while the loop around them stays warp-uniform, so that ptxas keeps descriptors / addresses in uniform registers
// instead of re-deriving them per MMA through an ELECT + R2UR.BROADCAST + BRA.U.ANY waterfall (12 dependent
// instructions per MMA when the whole loop runs under `if (lane == 0)`: the issuing thread, not the tensor pipe,
// was the limit of the int8 kernels -- profiles/r01_kernels_v8.md)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred P;\n"
        "elect.sync _|P, 0xffffffff;\n"
        "selp.b32 %0, 1, 0, P;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], int8 x int8 -> int32, issued by ONE thread for the CTA
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major operand tile, rows of 128 bytes, 128-byte swizzle (what TMA SWIZZLE_128B writes): 8-row groups
// are 1024 bytes apart (SBO); LBO is unused for swizzled K-major tiles; descriptor version 1 = sm_100.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(1024u >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// instruction descriptor: D = S32 (2) at [4,6), A = INT8 (1) at [7,10), B = INT8 (1) at [10,13), both K-major,
// N>>3 at [17,23), M>>4 at [24,29)
constexpr uint32_t I8_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(I8_TK >> 3) << 17) |
                              ((uint32_t)(I8_TJ >> 4) << 24);

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, int (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
// four consecutive 32-bit columns of this warp's 32 TMEM lanes
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, int (&v)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];\n"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

constexpr uint32_t I8P_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(I8_TK >> 3) << 17) |
                               ((uint32_t)((2 * I8_TJ) >> 4) << 24);
// the same with N = nn plane rows (a multiple of 16 in [16, 256] for M = 256)
__device__ __forceinline__ uint32_t i8p_idesc(int nn) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nn >> 3) << 17) | ((uint32_t)((2 * I8_TJ) >> 4) << 24);
}
constexpr uint32_t I8P_PEER_MASK = 0xFEFFFFFFu;                   // shared::cluster address of the same offset in CTA 0

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// TMA load whose completion bytes are signalled on the LEADER CTA's barrier (executed by both CTAs)
__device__ __forceinline__ void tma_load_2d_pair(void *smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar) & I8P_PEER_MASK)
        : "memory");
}
__device__ __forceinline__ void umma_i8_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the barrier at this offset in BOTH CTAs once the MMAs issued so far have retired
__device__ __forceinline__ void tc_commit_pair(uint64_t *bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(
            smem_u32(bar)), "h"((unsigned short)3)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];\n" ::"r"(smem_u32(bar) & I8P_PEER_MASK) : "memory");
}

// ------------------------------------------------------------------------------------------ host side
// 2-bit store -> int8 codes g8[snp][k] = byte code_of(c) of `tab` (byte c = the value of 2-bit code c), 0 for k >= n;
// g2 (optional): the same values 2 bits each.  Defined in ozaki.cu.
int launch_i8_dosage(const unsigned char *bed_dev, int n, long long m, long long row_bytes, unsigned tab,
                     const int *counts_dev /* nullable: per-SNP minor-allele coding of the additive table */,
                     signed char *g8_dev, unsigned char *g2_dev, int ld, cudaStream_t st);

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

// int8 matrix [rows][ld bytes], box = 128 bytes x box_rows, 128-byte swizzle, out-of-range rows read as zero
inline int make_map(CUtensorMap *map, const void *base, long long inner, long long rows, long long ld, int box_rows) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (enc == nullptr) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return GMAT_B200_ERR_CUDA; }
    cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld};
    cuuint32_t box[2] = {(cuuint32_t)I8_SLAB, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return GMAT_B200_ERR_CUDA; }
    return GMAT_B200_OK;
}

}  // namespace gb
