// Issue-rate probe for tcgen05.mma kind::tf32 (sm_100a): how many cycles does ONE MMA cost when the loop around it is
// written the way the fused deformation kernels write it?  Not part of the product.
//   style 0: `if (tid == 0)` single-thread loop, descriptors rebuilt per MMA (what deform_tc.cu did first)
//   style 1: whole warp runs the loop, MMAs under elect.sync, descriptors advanced by adding constants
//   flags  : +commit per 6 MMAs (c), +tcgen05.fence::after_thread_sync and a try_wait on a completed barrier (w)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_issue_probe umma_issue_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../4dgaussians-fast-train_b200/csrc/umma.cuh"

using namespace fdgs_umma;

__device__ __forceinline__ uint32_t elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p;
}

// A: 128 x 128 canonical tile (hi, lo) in smem or TMEM; B: ring of 4 slots, each [hi: N x 16][lo: N x 16]
template <int STYLE, bool TS, bool COMMIT, bool WAIT>
__global__ void __launch_bounds__(128, 1) probe(int N, int iters, long long* out)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[8];
    __shared__ uint32_t tmem_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    float* A_hi = reinterpret_cast<float*>(smem);
    float* A_lo = A_hi + 128 * 128;
    float* ring = A_lo + 128 * 128;                 // 4 x 4096 floats
    for (int i = tid; i < 2 * 128 * 128 + 4 * 4096; i += 128) A_hi[i] = 0.f;
    if (warp == 0) tmem_alloc(&tmem_s, 512);
    if (tid == 0) { for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1); fence_mbar_init(); }
    fence_async_smem();
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tmem_raw = tmem_s;
    if (tid == 0) mbar_arrive(&bars[6]);            // bars[6]: phase 0 complete -> try_wait(parity 0) returns at once
    __syncthreads();
    long long t0 = clock64();
    if (warp == 0) {
        const uint32_t idesc = make_idesc_tf32(128, N, 0, 0);
        if (STYLE == 0) {
            if (tid == 0) {
                const uint32_t tmem = tmem_raw;
                uint32_t g = 0;
                for (int it = 0; it < iters; ++it)
                    for (int c = 0; c < 8; ++c, ++g) {
                        const uint32_t slot = g & 3;
                        if (WAIT) { mbar_wait(&bars[6], 0); tc_fence_after(); }
                        const float* bh = ring + slot * 4096;
                        const float* bl = bh + 16 * N;
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            const int k = c * 16 + ks * 8;
                            const int first = (c == 0 && ks == 0) ? 0 : 1;
                            const uint64_t dbh = desc_kmajor(bh, N, ks * 8), dbl = desc_kmajor(bl, N, ks * 8);
                            if (!TS) {
                                mma_tf32_ss(tmem, desc_kmajor(A_lo, 128, k), dbh, idesc, first);
                                mma_tf32_ss(tmem, desc_kmajor(A_hi, 128, k), dbl, idesc, 1);
                                mma_tf32_ss(tmem, desc_kmajor(A_hi, 128, k), dbh, idesc, 1);
                            } else {
                                mma_tf32_ts(tmem, tmem + 256 + k, dbh, idesc, first);
                                mma_tf32_ts(tmem, tmem + 128 + k, dbl, idesc, 1);
                                mma_tf32_ts(tmem, tmem + 128 + k, dbh, idesc, 1);
                            }
                        }
                        if (COMMIT) mma_commit(&bars[slot]);
                    }
                mma_commit(&bars[7]);
            }
        } else {
            const uint32_t tmem = __shfl_sync(0xffffffffu, tmem_raw, 0);
            const uint32_t leader = elect_one();
            // descriptor low words; the high word (LBO/SBO/version) is constant per operand
            const uint64_t a_hi0 = desc_kmajor(A_hi, 128, 0), a_lo0 = desc_kmajor(A_lo, 128, 0);
            const uint32_t a_step = (128u * 32u) >> 4;            // 8 columns of a 128-row tile, in 16-byte units
            const uint32_t b_step = ((uint32_t)N * 32u) >> 4;
            uint32_t g = 0;
            for (int it = 0; it < iters; ++it) {
#pragma unroll 1
                for (int c = 0; c < 8; ++c, ++g) {
                    const uint32_t slot = g & 3;
                    if (WAIT) { mbar_wait(&bars[6], 0); tc_fence_after(); }
                    const float* bh = ring + slot * 4096;
                    const uint64_t dbh0 = desc_kmajor(bh, N, 0), dbl0 = desc_kmajor(bh + 16 * N, N, 0);
                    if (leader) {
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            const uint32_t ka = (uint32_t)(c * 2 + ks);
                            const int first = (c == 0 && ks == 0) ? 0 : 1;
                            const uint64_t dbh = dbh0 + ks * b_step, dbl = dbl0 + ks * b_step;
                            if (!TS) {
                                mma_tf32_ss(tmem, a_lo0 + ka * a_step, dbh, idesc, first);
                                mma_tf32_ss(tmem, a_hi0 + ka * a_step, dbl, idesc, 1);
                                mma_tf32_ss(tmem, a_hi0 + ka * a_step, dbh, idesc, 1);
                            } else {
                                mma_tf32_ts(tmem, tmem + 256 + ka * 8, dbh, idesc, first);
                                mma_tf32_ts(tmem, tmem + 128 + ka * 8, dbl, idesc, 1);
                                mma_tf32_ts(tmem, tmem + 128 + ka * 8, dbh, idesc, 1);
                            }
                        }
                        if (COMMIT) mma_commit(&bars[slot]);
                    }
                    __syncwarp();
                }
            }
            if (leader) mma_commit(&bars[7]);
            __syncwarp();
        }
        mbar_wait(&bars[7], 0);
        tc_fence_after();
    }
    __syncthreads();
    if (tid == 0) out[blockIdx.x] = clock64() - t0;
    if (warp == 0) tmem_free(tmem_raw, 512);
}

template <int STYLE, bool TS, bool COMMIT, bool WAIT>
static void run(const char* name, int N, int grid) {
    const int iters = 100;
    long long* d; cudaMalloc(&d, sizeof(long long) * grid);
    const size_t smem = (2 * 128 * 128 + 4 * 4096) * 4 + 1024;
    auto k = probe<STYLE, TS, COMMIT, WAIT>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<grid, 128, smem>>>(N, iters, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    printf("%-44s N=%3d grid=%3d: %s  %.1f cycles/MMA (floor %d)\n", name, N, grid, cudaGetErrorString(e),
           (double)h[0] / (iters * 48.0), 128 * N / 256);
    cudaFree(d);
}

int main() {
    for (int grid : {1, 148}) {
        run<0, false, false, false>("SS thread0", 128, grid);
        run<0, false, true, false>("SS thread0 +commit", 128, grid);
        run<0, false, true, true>("SS thread0 +commit +wait/fence", 128, grid);
        run<1, false, true, true>("SS elect   +commit +wait/fence", 128, grid);
        run<1, false, false, false>("SS elect", 128, grid);
        run<0, true, true, true>("TS thread0 +commit +wait/fence", 128, grid);
        run<1, true, true, true>("TS elect   +commit +wait/fence", 128, grid);
        run<0, true, true, true>("TS thread0 +commit +wait/fence", 16, grid);
        run<1, true, true, true>("TS elect   +commit +wait/fence", 16, grid);
        run<1, true, false, false>("TS elect", 16, grid);
        run<0, true, true, true>("TS thread0 +commit +wait/fence", 48, grid);
        run<1, true, true, true>("TS elect   +commit +wait/fence", 48, grid);
        run<1, true, true, true>("TS elect   +commit +wait/fence", 32, grid);
        run<1, true, true, true>("TS elect   +commit +wait/fence", 64, grid);
    }
    return 0;
}
