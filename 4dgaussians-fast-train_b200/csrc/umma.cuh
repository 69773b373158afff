// tcgen05 / TMEM / mbarrier primitives (inline PTX, sm_100a) used by the fused deformation kernels.
//
// Canonical tile.  Every on-chip operand tile (activations, gradients, pre-formatted weights) is stored as
//     tile[col/4][row][col%4]      float, `rows` rows, i.e. 16-byte units of 4 consecutive columns of one row,
//     tile_off(row, col, rows) = ((col >> 2) * rows + row) * 4 + (col & 3)
// One copy of such a tile can feed the tensor core in two ways (no-swizzle "interleave" descriptors):
//   * K-major operand   (MMA rows = tile rows, reduction over tile columns):  8 rows x 16 B core matrices are
//     contiguous (128 B), row groups advance by SBO = 128 B, column chunks by LBO = rows * 16 B;
//   * MN-major operand  (MMA rows = tile columns, reduction over tile rows):  4 columns x 8 rows core matrices,
//     column groups advance by SBO = rows * 16 B, row groups of 8 by LBO = 128 B.
// A thread that owns one row writes 16-byte vectors at consecutive addresses across the warp (conflict free).
//
// 3xTF32: x = hi + lo with hi = x rounded to TF32 and lo = (x - hi) rounded to TF32;
// a*b ~= hi_a*hi_b + lo_a*hi_b + hi_a*lo_b accumulated in FP32 in TMEM (error ~2^-23 per product).
// NOTE: the plain no-swizzle MN-major view described above does NOT work for 32-bit operands (the MMA
// silently produces zeros, measured); MN-major TF32 operands use the swizzled tile32 layout below.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace fdgs_umma {

__host__ __device__ __forceinline__ int tile_off(int row, int col, int rows) { return (((col >> 2) * rows + row) << 2) + (col & 3); }

// MN-major TF32 operands must use the 128-byte swizzle with 32-byte base (layout type 1): a "tile32" stores
//     tile32[col/32][row][32 columns], 128-byte rows, the 32-byte chunk index of a row XORed with (row & 3)
// (Swizzle<2,5,2> on the byte address: bits [7,9) into bits [5,7)); rows = reduction index; base 512-byte aligned.
__host__ __device__ __forceinline__ int tile32_off(int row, int col, int rows) {
    return ((col >> 5) * rows + row) * 32 + (((((col >> 3) & 3) ^ (row & 3)) << 3) | (col & 7));
}

__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
    lo = x - hi;
}
// Round-to-nearest split: hi = rna_tf32(x), lo = rna_tf32(x - hi).  Both are exactly representable in TF32, so
// the tensor core's truncation of its inputs (measured, tools/umma_probe.cu) is a no-op and the dropped
// remainder is <= 2^-24 |x|: a 3-pass product is FP32-accurate.
__device__ __forceinline__ float rna_tf32(float x) {
    // cvt.rna.tf32.f32 is emulated on sm_100a (IADD 0x1000, |x| < inf test, select, mask = 4 instructions); the
    // inf/nan guard is dropped here: add half an ulp of the 10-bit mantissa to the magnitude, then truncate.
    return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}
__device__ __forceinline__ void split_tf32_rn(float x, float& hi, float& lo) {
    hi = rna_tf32(x);
    lo = rna_tf32(x - hi);
}

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    } while (!ok);
}
// 1-D bulk async copy global -> shared (TMA engine), completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}

// L2 prefetch of a contiguous global range (multiple of 16 bytes): decouples DRAM latency from the shared-memory ring depth
__device__ __forceinline__ void bulk_prefetch_l2(const void* src_gmem, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}

// ---- fences --------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (tensor core / TMA reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMEM ----------------------------------------------------------------------------------------------
// whole warp; writes the base address (lane 0, column 0 of the allocation) to *dst_smem
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// 32 lanes x 32-bit: thread i of the warp gets / provides N consecutive columns of lane (base_lane + i)
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&v)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                   "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
                   "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- descriptors -----------------------------------------------------------------------------------------
// shared-memory matrix descriptor: start >> 4 [0,14), LBO >> 4 [16,30), SBO >> 4 [32,46), version = 1 [46,48),
// layout type [61,64) = 0 (no swizzle)
__device__ __forceinline__ uint64_t make_sdesc(uint32_t start_bytes, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((start_bytes >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// K-major view of a canonical tile with `rows` rows, starting at tile column k (multiple of 4)
__device__ __forceinline__ uint64_t desc_kmajor(const float* tile, int rows, int k) {
    return make_sdesc(smem_addr(tile) + (uint32_t)(k >> 2) * (uint32_t)rows * 16u, (uint32_t)rows * 16u, 128u);
}
// same, but the MMA rows start at tile row r0 (multiple of 8)
__device__ __forceinline__ uint64_t desc_kmajor_rows(const float* tile, int rows, int r0, int k) {
    return make_sdesc(smem_addr(tile) + ((uint32_t)(k >> 2) * (uint32_t)rows + (uint32_t)r0) * 16u, (uint32_t)rows * 16u, 128u);
}
// MN-major view: MMA rows = tile columns starting at column c0 (multiple of 4), reduction over tile rows from p (multiple of 8)
__device__ __forceinline__ uint64_t desc_mnmajor(const float* tile, int rows, int p, int c0 = 0) {
    return make_sdesc(smem_addr(tile) + ((uint32_t)(c0 >> 2) * (uint32_t)rows + (uint32_t)p) * 16u, 128u, (uint32_t)rows * 16u);
}
// MN-major view of a tile32 (layout type 1 = SWIZZLE_128B_BASE32B): MMA rows = tile columns from c0 (multiple of 32),
// reduction over tile rows from p (multiple of 8).  LBO = stride between 32-column blocks, SBO = 4 rows = 512 B.
__device__ __forceinline__ uint64_t desc_mn32(const float* tile, int rows, int p, int c0) {
    return make_sdesc(smem_addr(tile) + ((uint32_t)(c0 >> 5) * (uint32_t)rows + (uint32_t)p) * 128u, (uint32_t)rows * 128u, 512u) |
           (1ull << 61);
}
// instruction descriptor, kind::tf32, FP32 accumulate: c_format=F32 [4,6), a/b format TF32=2 [7,10)/[10,13),
// a_major [15], b_major [16] (0 = K-major, 1 = MN-major), N >> 3 [17,23), M >> 4 [24,29)
__host__ __device__ __forceinline__ uint32_t make_idesc_tf32(int M, int N, int a_mn, int b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}

// ---- MMA (one thread issues) ---------------------------------------------------------------------------------
__device__ __forceinline__ void mma_tf32_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, int accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, int accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// One lane of a converged warp (elect.sync).  The MMA issuer runs its loop with the WHOLE warp and issues under this
// predicate: the compiler then keeps descriptors / TMEM addresses in uniform registers and emits back-to-back UTCHMMA.
// A loop entered by `if (lane == 0)` instead costs ~100 cycles per MMA in per-thread address arithmetic and
// vector->uniform register "waterfalls" (measured, tools/umma_issue_probe.cu: N=128 SS 101 -> 69 cycles, N=16 TS 68 -> 36).
__device__ __forceinline__ uint32_t elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p;
}
// all previously issued MMAs of this thread arrive on `bar` when complete (implies fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(bar)) : "memory");
}

}  // namespace fdgs_umma
