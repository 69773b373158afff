// Scatter-add probe (sm_100a): 128-byte rows added to random 128-byte-aligned places of an L2-resident buffer, either with
// red.global.add.v4.f32 from registers (what the sampler backward does today) or with one TMA bulk reduction per row
// (cp.reduce.async.bulk.global.shared::cta.add.f32) from shared memory.  Not part of the product.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_probe red_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

__device__ __forceinline__ uint32_t hash(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// MODE 0: each group of 8 lanes adds one 128-byte row (8 x red.v4); MODE 1: each thread stages a row in smem, 1 bulk reduce
template <int MODE, int DEPTH>
__global__ void __launch_bounds__(256) probe(float* buf, uint32_t nrows_buf, int rows_per_thread) {
    extern __shared__ __align__(128) float sm[];
    const int tid = threadIdx.x;
    const uint32_t gid = blockIdx.x * 256 + tid;
    if (MODE == 0) {
        const int sub = tid & 7;
        for (int i = 0; i < rows_per_thread * 8; ++i) {            // 8 lanes share a row: same number of rows per CTA
            const uint32_t r = hash((gid >> 3) * 7919u + i) % nrows_buf;
            float* p = buf + (size_t)r * 32 + sub * 4;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(1.f), "f"(2.f), "f"(3.f), "f"(4.f) : "memory");
        }
    } else if (MODE == 2) {
        const int sub = tid & 3;                                   // 4 lanes share a 64-byte half row at its own random place
        for (int i = 0; i < rows_per_thread * 8; ++i) {
            const uint32_t r = hash((gid >> 2) * 7919u + i) % (nrows_buf * 2);
            float* p = buf + (size_t)r * 16 + sub * 4;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(1.f), "f"(2.f), "f"(3.f), "f"(4.f) : "memory");
        }
    } else if (MODE == 3) {
        for (int i = 0; i < rows_per_thread * 8; ++i) {            // every lane its own random 16 bytes
            const uint32_t r = hash(gid * 7919u + i) % (nrows_buf * 8);
            float* p = buf + (size_t)r * 4;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(1.f), "f"(2.f), "f"(3.f), "f"(4.f) : "memory");
        }
    } else {
        float* row = sm + (size_t)tid * 32 * DEPTH;
        for (int i = 0; i < rows_per_thread; ++i) {
            float* rw = row + (i % DEPTH) * 32;
            if (i >= DEPTH) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(DEPTH - 1) : "memory");
#pragma unroll
            for (int k = 0; k < 8; ++k) reinterpret_cast<float4*>(rw)[k] = make_float4(1.f, 2.f, 3.f, 4.f);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            const uint32_t r = hash(gid * 7919u + i) % nrows_buf;
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 128;"
                         ::"l"(buf + (size_t)r * 32), "r"(smem_u32(rw)) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

template <int MODE, int DEPTH>
static void run(const char* name, float* buf, uint32_t nrows_buf) {
    const int grid = 148 * 2, rpt = 72;                            // 148*2*256*72 = 5.46 M rows of 128 B (one DyNeRF view)
    const size_t smem = MODE == 1 ? (size_t)256 * 128 * DEPTH : 0;
    auto k = probe<MODE, DEPTH>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<grid, 256, smem>>>(buf, nrows_buf, rpt);
    cudaEventRecord(e0);
    k<<<grid, 256, smem>>>(buf, nrows_buf, rpt);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double rows = (double)grid * 256 * rpt;
    printf("%-46s %s  %.1f us for %.2f M rows (%.0f GB/s payload)\n", name, cudaGetErrorString(e), ms * 1e3, rows / 1e6, rows * 128 / ms / 1e6);
}

int main() {
    for (uint32_t mb : {10u, 1u}) {
        const uint32_t nrows = mb * 1024 * 1024 / 128;
        float* buf; cudaMalloc(&buf, (size_t)nrows * 128); cudaMemset(buf, 0, (size_t)nrows * 128);
        printf("target buffer %u MB\n", mb);
        run<0, 1>("red.global.add.v4.f32 (8 lanes per row)", buf, nrows);
        run<2, 1>("red.global.add.v4.f32 (4 lanes per 64 B)", buf, nrows);
        run<3, 1>("red.global.add.v4.f32 (each lane random 16 B)", buf, nrows);
        run<1, 1>("bulk reduce, 1 row in flight per thread", buf, nrows);
        cudaFree(buf);
    }
    return 0;
}
