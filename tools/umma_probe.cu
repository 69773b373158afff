// Standalone probe of the tcgen05 primitives the fused deformation kernels are built on (sm_100a).
// Not part of the product: it validates, on a real B200, the operand layouts / descriptors used by
// csrc/umma.cuh before the fused kernels rely on them.
//   mode 0: SS, A and B K-major, no swizzle, single-pass TF32
//   mode 1: SS, 3xTF32 split (hi*hi + lo*hi + hi*lo)
//   mode 2: TS, A written to TMEM by tcgen05.st, 3xTF32
//   mode 3: SS, A and B MN-major (reduction over the row index of the stored tiles), 3xTF32
//   mode 4: throughput loop (SS, N=128)  mode 5: throughput loop (TS, N=32)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_probe umma_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <vector>
#include "../4dgaussians-fast-train_b200/csrc/umma.cuh"

using namespace fdgs_umma;

constexpr int M = 128;

// tile in smem: [k/4][rows][4] floats (see umma.cuh "canonical tile")
__global__ void __launch_bounds__(128, 1)
probe(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int N, int K, int mode, int iters,
      long long* cycles)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int KA = (mode == 3) ? 128 : K;           // mode 3: tiles are [points=K][features]; A tile has 128 feature columns
    float* As_hi = reinterpret_cast<float*>(smem);
    float* As_lo = As_hi + 128 * 128;
    float* Bs_hi = As_lo + 128 * 128;
    float* Bs_lo = Bs_hi + 128 * 64;                // B tiles hold at most 8192 floats

    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tmem = tmem_base_s;

    // ---- stage operands ------------------------------------------------------------------------
    const bool sw = (mode == 10 || mode == 11);        // MN-major SWIZZLE_128B_BASE32B staging
    const bool a_pt = (mode == 3 || mode == 6 || mode == 8 || mode == 9 || sw);      // A given as [K points][128 feats]
    const bool b_pt = (mode == 3 || mode == 7 || mode == 8 || mode == 11);     // B given as [K points][N feats]
    if (!a_pt) {
        for (int i = tid; i < M * K; i += 128) { const int r = i / K, k = i % K; float hi, lo; split_tf32(A[i], hi, lo);
            As_hi[tile_off(r, k, M)] = hi; As_lo[tile_off(r, k, M)] = lo; }
    } else {
        for (int i = tid; i < K * 128; i += 128) { const int p = i / 128, f = i % 128; float hi, lo; split_tf32(A[i], hi, lo);
            As_hi[tile_off(p, f, K)] = hi; As_lo[tile_off(p, f, K)] = lo; }
    }
    if (!b_pt) {
        for (int i = tid; i < N * K; i += 128) { const int r = i / K, k = i % K; float hi, lo; split_tf32(B[i], hi, lo);
            Bs_hi[tile_off(r, k, N)] = hi; Bs_lo[tile_off(r, k, N)] = lo; }
    } else {
        for (int i = tid; i < K * N; i += 128) { const int p = i / N, f = i % N; float hi, lo; split_tf32(B[i], hi, lo);
            Bs_hi[tile_off(p, f, K)] = hi; Bs_lo[tile_off(p, f, K)] = lo; }
    }
    (void)KA;
    if (mode == 2 || mode == 5) {
        // A into TMEM: lane = row, column = k.  hi at columns [256, 256+K), lo at [384, 384+K)
        const uint32_t row_addr = tmem + ((uint32_t)(warp * 32) << 16);
        for (int k0 = 0; k0 < K; k0 += 8) {
            uint32_t h[8], l[8];
            for (int j = 0; j < 8; ++j) { float hi, lo; split_tf32(A[(size_t)tid * K + k0 + j], hi, lo); h[j] = __float_as_uint(hi); l[j] = __float_as_uint(lo); }
            tmem_st8(row_addr + 256 + k0, h);
            tmem_st8(row_addr + 384 + k0, l);
        }
        tmem_wait_st();
    }
    if (sw) {
        __syncthreads();
        // tile32[f/32][p][(f%32) ^ ((f>>2)&3)]
        for (int i = tid; i < K * 128; i += 128) { const int p = i / 128, f = i % 128; float hi, lo; split_tf32(A[i], hi, lo);
            As_hi[tile32_off(p, f, K)] = hi; As_lo[tile32_off(p, f, K)] = lo; }
        if (mode == 11)
            for (int i = tid; i < K * N; i += 128) { const int p = i / N, f = i % N; float hi, lo; split_tf32(B[i], hi, lo);
                Bs_hi[tile32_off(p, f, K)] = hi; Bs_lo[tile32_off(p, f, K)] = lo; }
    }
    fence_async_smem();
    tc_fence_before(); __syncthreads(); tc_fence_after();

    long long t0 = clock64();
    if (tid == 0) {
        const uint32_t idesc = make_idesc_tf32(M, N, a_pt ? 1 : 0, b_pt ? 1 : 0);
        for (int it = 0; it < iters; ++it) {
            if (mode == 0) {
                for (int k = 0; k < K; k += 8)
                    mma_tf32_ss(tmem, desc_kmajor(As_hi, M, k), desc_kmajor(Bs_hi, N, k), idesc, k > 0);
            } else if (mode == 1 || mode == 4) {
                for (int k = 0; k < K; k += 8) {
                    mma_tf32_ss(tmem, desc_kmajor(As_lo, M, k), desc_kmajor(Bs_hi, N, k), idesc, k > 0);
                    mma_tf32_ss(tmem, desc_kmajor(As_hi, M, k), desc_kmajor(Bs_lo, N, k), idesc, 1);
                    mma_tf32_ss(tmem, desc_kmajor(As_hi, M, k), desc_kmajor(Bs_hi, N, k), idesc, 1);
                }
            } else if (mode == 2 || mode == 5) {
                for (int k = 0; k < K; k += 8) {
                    mma_tf32_ts(tmem, tmem + 384 + k, desc_kmajor(Bs_hi, N, k), idesc, k > 0);
                    mma_tf32_ts(tmem, tmem + 256 + k, desc_kmajor(Bs_lo, N, k), idesc, 1);
                    mma_tf32_ts(tmem, tmem + 256 + k, desc_kmajor(Bs_hi, N, k), idesc, 1);
                }
            } else if (sw) {
                for (int p = 0; p < K; p += 8) {
                    const uint64_t dah = desc_mn32(As_hi, K, p, 0), dal = desc_mn32(As_lo, K, p, 0);
                    const uint64_t dbh = mode == 11 ? desc_mn32(Bs_hi, K, p, 0) : desc_kmajor(Bs_hi, N, p);
                    const uint64_t dbl = mode == 11 ? desc_mn32(Bs_lo, K, p, 0) : desc_kmajor(Bs_lo, N, p);
                    mma_tf32_ss(tmem, dal, dbh, idesc, p > 0);
                    mma_tf32_ss(tmem, dah, dbl, idesc, 1);
                    mma_tf32_ss(tmem, dah, dbh, idesc, 1);
                }
            } else if (mode >= 6) {
                // single-pass layout checks: 6 = A MN-major only, 7 = B MN-major only, 8 = both with LBO/SBO swapped, 9 = A MN swapped
                for (int p = 0; p < K; p += 8) {
                    uint64_t da = a_pt ? desc_mnmajor(As_hi, K, p) : desc_kmajor(As_hi, M, p);
                    uint64_t db = b_pt ? desc_mnmajor(Bs_hi, K, p) : desc_kmajor(Bs_hi, N, p);
                    if (mode == 8 || mode == 9) da = make_sdesc(smem_addr(As_hi) + (uint32_t)p * 16u, (uint32_t)K * 16u, 128u);
                    if (mode == 8) db = make_sdesc(smem_addr(Bs_hi) + (uint32_t)p * 16u, (uint32_t)K * 16u, 128u);
                    mma_tf32_ss(tmem, da, db, idesc, p > 0);
                }
            } else if (mode == 3) {
                // D[m][n] = sum_p X[p][m] Y[p][n]: A = X tile read MN-major (M = features), K = points
                for (int p = 0; p < K; p += 8) {
                    mma_tf32_ss(tmem, desc_mnmajor(As_lo, K, p), desc_mnmajor(Bs_hi, K, p), idesc, p > 0);
                    mma_tf32_ss(tmem, desc_mnmajor(As_hi, K, p), desc_mnmajor(Bs_lo, K, p), idesc, 1);
                    mma_tf32_ss(tmem, desc_mnmajor(As_hi, K, p), desc_mnmajor(Bs_hi, K, p), idesc, 1);
                }
            }
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    long long t1 = clock64();
    if (tid == 0 && cycles) *cycles = t1 - t0;

    // ---- read the accumulator back: warp w owns lanes [32w, 32w+32) ------------------------------------
    const uint32_t row_addr = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c0 = 0; c0 < N; c0 += 8) {
        uint32_t v[8];
        tmem_ld8(row_addr + c0, v);
        tmem_wait_ld();
        for (int j = 0; j < 8; ++j) D[(size_t)tid * N + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_free(tmem, 512);
    (void)lane;
}

static float tf32_trunc(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }
static float tf32_rna(float x) { uint32_t u; memcpy(&u, &x, 4); u += 0x1000u; u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }

int main() {
    const size_t smem_bytes = (2 * 128 * 128 + 2 * 128 * 64) * 4;
    cudaError_t ea = cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (ea != cudaSuccess) { printf("set attr: %s\n", cudaGetErrorString(ea)); return 1; }
    srand(1);
    auto rnd = [] { return (float)rand() / RAND_MAX * 2.f - 1.f; };
    struct Case { int mode, N, K; };
    Case cases[] = {{0, 128, 32}, {0, 48, 128}, {0, 8, 16}, {1, 64, 128}, {1, 32, 32}, {1, 128, 64}, {2, 64, 128}, {2, 32, 64}, 
                    {10, 64, 8}, {10, 64, 128}, {10, 32, 64}, {11, 64, 8}, {11, 64, 128}, {11, 32, 128}, {11, 128, 64}};
    for (auto c : cases) {
        const int N = c.N, K = c.K;
        const bool a_pt = (c.mode == 3 || c.mode == 6 || c.mode == 8 || c.mode == 9 || c.mode >= 10), b_pt = (c.mode == 3 || c.mode == 7 || c.mode == 8 || c.mode == 11);
        const int arows = a_pt ? K : M, acols = a_pt ? 128 : K;
        const int brows = b_pt ? K : N, bcols = b_pt ? N : K;
        std::vector<float> A((size_t)arows * acols), B((size_t)brows * bcols), D((size_t)M * N);
        for (auto& x : A) x = rnd();
        for (auto& x : B) x = rnd();
        float *dA, *dB, *dD;
        cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4);
        cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
        cudaMemset(dD, 0, D.size() * 4);
        probe<<<1, 128, smem_bytes>>>(dA, dB, dD, N, K, c.mode, 1, nullptr);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d N %d K %d: CUDA error %s\n", c.mode, N, K, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
        double e_full = 0, e_trunc = 0, e_rna = 0, mx = 0;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) {
                double s = 0, st = 0, sr = 0;
                for (int k = 0; k < K; ++k) {
                    const float a = a_pt ? A[(size_t)k * 128 + m] : A[(size_t)m * K + k];
                    const float b = b_pt ? B[(size_t)k * N + n] : B[(size_t)n * K + k];
                    s += (double)a * b; st += (double)tf32_trunc(a) * tf32_trunc(b); sr += (double)tf32_rna(a) * tf32_rna(b);
                }
                const double d = D[(size_t)m * N + n];
                e_full = fmax(e_full, fabs(d - s)); e_trunc = fmax(e_trunc, fabs(d - st)); e_rna = fmax(e_rna, fabs(d - sr));
                mx = fmax(mx, fabs(s));
            }
        if (c.mode >= 3) printf("   D[0][0..3] = %g %g %g %g   D[5][0..1] = %g %g\n", D[0], D[1], D[2], D[3], D[5 * N], D[5 * N + 1]);
        printf("mode %d N %3d K %3d: max|D| %.3f  err vs exact %.3e  vs tf32-trunc inputs %.3e  vs tf32-rna inputs %.3e\n",
               c.mode, N, K, mx, e_full, e_trunc, e_rna);
        cudaFree(dA); cudaFree(dB); cudaFree(dD);
    }
    // throughput
    for (int mode : {4, 5}) {
        const int N = mode == 4 ? 128 : 32, K = mode == 4 ? 64 : 128, iters = 200;
        float *dA, *dB, *dD; long long* dc;
        cudaMalloc(&dA, 128 * 128 * 4); cudaMalloc(&dB, 128 * 128 * 4); cudaMalloc(&dD, 128 * 128 * 4); cudaMalloc(&dc, 8);
        cudaMemset(dA, 0, 128 * 128 * 4); cudaMemset(dB, 0, 128 * 128 * 4);
        probe<<<1, 128, smem_bytes>>>(dA, dB, dD, N, K, mode, iters, dc);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        long long cyc = 0; cudaMemcpy(&cyc, dc, 8, cudaMemcpyDeviceToHost);
        const double macs = (double)iters * 3.0 * 128 * N * K;
        printf("mode %d (%s N=%d): %s, %lld cycles, %.1f TF32 MAC/cycle/SM (floor 2048)\n", mode, mode == 4 ? "SS" : "TS", N,
               cudaGetErrorString(e), cyc, macs / (double)cyc);
        cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dc);
    }
    return 0;
}
