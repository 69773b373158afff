// TMEM epilogue-traffic probe (sm_100a): cycles for 8 warps to read a 128 x 128 FP32 accumulator (tcgen05.ld) and write
// two 128 x 128 operand tiles back (tcgen05.st), alone and while N=128 SS MMAs run.  Not part of the product.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_probe tmem_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../4dgaussians-fast-train_b200/csrc/umma.cuh"
using namespace fdgs_umma;

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// MODE 0: ld16+wait x4 | 1: + st16 x8 + wait::st | 2: ld32+wait x2 + st32 x4 | 3: all four ld16 issued, one wait, then st
template <int MODE, bool MMA>
__global__ void __launch_bounds__(320, 1) probe(int iters, long long* out, float* sink)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[4];
    __shared__ uint32_t tmem_s;
    __shared__ int stop;
    const int tid = threadIdx.x, warp = tid >> 5;
    float* A_hi = reinterpret_cast<float*>(smem);
    float* Bt = A_hi + 128 * 128;
    for (int i = tid; i < 2 * 128 * 128; i += 320) A_hi[i] = 0.f;
    if (warp == 9) tmem_alloc(&tmem_s, 512);
    if (tid == 0) { for (int i = 0; i < 4; ++i) mbar_init(&bars[i], 1); fence_mbar_init(); stop = 0; }
    fence_async_smem();
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tmem = tmem_s;
    if (warp == 9) {
        if (MMA) {
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
            const uint32_t leader = elect_one();
            const uint32_t idesc = make_idesc_tf32(128, 128, 0, 0);
            const uint64_t a0 = desc_kmajor(A_hi, 128, 0), b0 = desc_kmajor(Bt, 128, 0);
            uint32_t ph = 0;
            while (!*(volatile int*)&stop) {
                if (leader) {
#pragma unroll
                    for (int k = 0; k < 16; ++k) mma_tf32_ss(tm + 128, a0 + k * 256u, b0 + k * 256u, idesc, 1);
                    mma_commit(&bars[0]);
                }
                __syncwarp();
                mbar_wait(&bars[0], ph); ph ^= 1;
            }
        }
    } else if (warp < 8) {
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int col0 = (tid >> 7) * 64;
        float acc = 0.f;
        epi_bar();
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            if (MODE == 0 || MODE == 1) {
#pragma unroll 1
                for (int c0 = 0; c0 < 64; c0 += 16) {
                    uint32_t v[16];
                    tmem_ld16(lane_base + col0 + c0, v);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 16; ++j) acc += __uint_as_float(v[j]);
                    if (MODE == 1) { tmem_st16(lane_base + 256 + col0 + c0, v); tmem_st16(lane_base + 384 + col0 + c0, v); }
                }
                if (MODE == 1) tmem_wait_st();
            } else if (MODE == 2) {
#pragma unroll 1
                for (int c0 = 0; c0 < 64; c0 += 32) {
                    uint32_t v[32];
                    tmem_ld32(lane_base + col0 + c0, v);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 32; ++j) acc += __uint_as_float(v[j]);
                    tmem_st32(lane_base + 256 + col0 + c0, v); tmem_st32(lane_base + 384 + col0 + c0, v);
                }
                tmem_wait_st();
            } else {
                uint32_t v[4][16];
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld16(lane_base + col0 + c * 16, v[c]);
                tmem_wait_ld();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) acc += __uint_as_float(v[c][j]);
                    tmem_st16(lane_base + 256 + col0 + c * 16, v[c]); tmem_st16(lane_base + 384 + col0 + c * 16, v[c]);
                }
                tmem_wait_st();
            }
            tc_fence_before();
            epi_bar();
            tc_fence_after();
        }
        long long t1 = clock64();
        if (tid == 0) { out[blockIdx.x] = t1 - t0; stop = 1; }
        sink[blockIdx.x * 320 + tid] = acc;
    }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    if (warp == 9) tmem_free(tmem, 512);
}

template <int MODE, bool MMA>
static void run(const char* name) {
    const int iters = 200, grid = 148;
    long long* d; float* sink; cudaMalloc(&d, 8 * grid); cudaMalloc(&sink, 4 * grid * 320);
    const size_t smem = 2 * 128 * 128 * 4 + 1024;
    auto k = probe<MODE, MMA>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<grid, 320, smem>>>(iters, d, sink);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, 8 * grid, cudaMemcpyDeviceToHost);
    printf("%-52s %s  %.0f cycles per 128x128 pass\n", name, cudaGetErrorString(e), (double)h[0] / iters);
    cudaFree(d); cudaFree(sink);
}

int main() {
    run<0, false>("ld16+wait x4");
    run<1, false>("ld16+wait x4, st16 x8, wait::st");
    run<2, false>("ld32+wait x2, st32 x4, wait::st");
    run<3, false>("ld16 x4, one wait, st16 x8, wait::st");
    run<0, true>("ld16+wait x4                          | MMA busy");
    run<1, true>("ld16+wait x4, st16 x8, wait::st       | MMA busy");
    run<2, true>("ld32+wait x2, st32 x4, wait::st       | MMA busy");
    run<3, true>("ld16 x4, one wait, st16 x8, wait::st  | MMA busy");
    return 0;
}
