// Backward rasterizer for sm_100a.
//   k_blend_bwd      : reverse-order blend backward; per-pixel partials are reduced across the
//                      warp with a 12-shuffle transposed butterfly and accumulated per
//                      (tile, Gaussian) in shared memory; one global reduction per instance.
//   k_preprocess_bwd : fused cov2D / mean2D->mean3D / SH / cov3D->(scale, rot) chain.
// Replaces CudaRasterizer::Rasterizer::backward
// (RAST/cuda_rasterizer/rasterizer_impl.cu:343-437, backward.cu:20-557).
#include "raster_math.cuh"

#include "prof.cuh"

namespace {

constexpr unsigned FULL = 0xffffffffu;
// 128 records per staged batch and 6 resident CTAs per SM (72 registers): measured 621 us per DyNeRF view against 711
// (256 / no bound, 73 registers but 34 KB of shared memory per CTA), 714 (128 / no bound: 90 registers), 755 (128 / 8 CTAs:
// 64 registers) on the same B200 (profiles/r02_ab_blend_bwd.txt)
#ifndef BLEND_BWD_BATCH
#define BLEND_BWD_BATCH 128
#endif
#ifndef BLEND_BWD_MINB
#define BLEND_BWD_MINB 6
#endif
constexpr int BATCH = BLEND_BWD_BATCH;     // records per staged batch
constexpr int NACC = 9;          // colour(3), mean2D(2), conic(3), opacity(1)
#ifndef BLEND_BWD_NPX
#define BLEND_BWD_NPX 2
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// Same conservative rectangle test as the forward blend (see raster_fwd.cu:rect_culled).
__device__ __forceinline__ bool rect_culled(float mx, float my, float A, float B, float C, float cull,
                                            float px0, float py0, float px1, float py1) {
    const float dx_lo = mx - px1, dx_hi = mx - px0;
    const float dy_lo = my - py1, dy_hi = my - py0;
    const bool in_x = (dx_lo <= 0.f) && (dx_hi >= 0.f);
    const bool in_y = (dy_lo <= 0.f) && (dy_hi >= 0.f);
    if (in_x && in_y) return false;
    float qmin = 3.0e38f;
    const float invC = 1.0f / C;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        const float dx = e ? dx_hi : dx_lo;
        const float dy = fminf(dy_hi, fmaxf(dy_lo, -B * dx * invC));
        qmin = fminf(qmin, A * dx * dx + 2.f * B * dx * dy + C * dy * dy);
    }
    const float invA = 1.0f / A;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        const float dy = e ? dy_hi : dy_lo;
        const float dx = fminf(dx_hi, fmaxf(dx_lo, -B * dy * invA));
        qmin = fminf(qmin, A * dx * dx + 2.f * B * dx * dy + C * dy * dy);
    }
    const float ax = fmaxf(fabsf(dx_lo), fabsf(dx_hi)), ay = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
    const float mag = fabsf(A) * ax * ax + fabsf(C) * ay * ay + 2.f * fabsf(B) * ax * ay;
    const float margin = 1e-3f + 1e-3f * fabsf(cull) + 2e-5f * mag;
    return (-0.5f * qmin) < (cull - margin);
}

// Reduce 10 per-lane slots (9 accumulators + a zero pad) across the warp with 12 shuffles (5+3+2+1+1): a transposed
// butterfly whose halves are 5 | 3 (of 5, padded to 6) | 2 (of 3, padded to 4) | 1.  On return lane L holds in v[0] the warp
// total of slot ((L>>4)&1)*5 + ((L>>3)&1)*3 + ((L>>2)&1)*2 + ((L>>1)&1); lanes whose path ran through a pad hold 0.
__device__ __forceinline__ void warp_reduce10(float (&v)[10], int lane) {
    {
        const bool up = lane & 16;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const float send = up ? v[k] : v[k + 5];
            const float keep = up ? v[k + 5] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 16);
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float hv = (k + 3 < 5) ? v[k + 3] : 0.f;
            const float send = up ? v[k] : hv;
            const float keep = up ? hv : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 8);
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float hv = (k + 2 < 3) ? v[k + 2] : 0.f;
            const float send = up ? v[k] : hv;
            const float keep = up ? hv : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 4);
        }
    }
    {
        const bool up = lane & 2;
        const float send = up ? v[0] : v[1];
        const float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(FULL, send, 2);
    }
    v[0] += __shfl_xor_sync(FULL, v[0], 1);
}

// The same transposed butterfly inside each HALF-warp (16 lanes, 11 shuffles: 5 + 3 + 2 + 1): on return lane L holds in v[0] the
// half-warp total of slot b3*5 + b2*3 + b1*2 + b0 (b_k = bit k of L) when that path did not run through a pad, i.e. for
// (b2, b1) != (1, 1) and (b2, b1, b0) != (0, 1, 1): ten lanes of sixteen (the tenth slot is the zero pad).
__device__ __forceinline__ void half_reduce10(float (&v)[10], int lane) {
    {
        const bool up = lane & 8;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const float send = up ? v[k] : v[k + 5];
            const float keep = up ? v[k + 5] : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 8);
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float hv = (k + 3 < 5) ? v[k + 3] : 0.f;
            const float send = up ? v[k] : hv;
            const float keep = up ? hv : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 4);
        }
    }
    {
        const bool up = lane & 2;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float hv = (k + 2 < 3) ? v[k + 2] : 0.f;
            const float send = up ? v[k] : hv;
            const float keep = up ? hv : v[k];
            v[k] = keep + __shfl_xor_sync(FULL, send, 2);
        }
    }
    {
        const bool up = lane & 1;
        const float send = up ? v[0] : v[1];
        const float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(FULL, send, 1);
    }
}

// (BLEND_BWD_STREAMS = 2: one record stream per half-warp; measured SLOWER, 736 vs 620 us per DyNeRF view - the default stays
//  one stream per warp.  profiles/r02_ab_blend_streams.txt)
#ifndef BLEND_BWD_SPARSE
#define BLEND_BWD_SPARSE 0          // 0 = always reduce with the butterfly (default).  N > 0: skip the reduction when no lane had a
                                    // contributing pixel and let up to N hitting lanes add their partials themselves - measured SLOWER at
                                    // every N (620 us -> 656 / 676 / 701 / 815 at N = 1 / 2 / 3 / 6): the extra ballot and the divergent
                                    // branch cost more than the butterflies they save
#endif
#ifndef BLEND_BWD_STREAMS
#define BLEND_BWD_STREAMS 1
#endif

// NPX pixels per thread (rows py and py + 4): a warp covers an 8 x (4 NPX) sub-rectangle of the tile, so one cull test,
// one 12-shuffle reduction and one set of shared-memory atomics serve 32 NPX pixels of a hit.
template <int NPX>
__global__ void __launch_bounds__(256 / NPX, BLEND_BWD_MINB)
k_blend_bwd(const uint2* __restrict__ ranges, const FdgsRecord* __restrict__ records, int W, int H,
            const float* __restrict__ bg_color, const float* __restrict__ final_Ts,
            const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
            float* __restrict__ dL_dmean2D /*[P,3]*/, float* __restrict__ dL_dconic /*[P,4]*/,
            float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors /*[P,3]*/)
{
    constexpr int NT = 256 / NPX;
    __shared__ __align__(128) FdgsRecord srec[2][BATCH];
    __shared__ float acc[BATCH][NACC];
    __shared__ __align__(8) uint64_t bars[2];
    __shared__ uint32_t s_maxc;

    const int tiles_x = (W + FDGS_TILE - 1) / FDGS_TILE;
    const int tile = blockIdx.y * tiles_x + blockIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int wx = wid & 1, wy = wid >> 1;
    const int bx0 = blockIdx.x * FDGS_TILE + wx * 8, by0 = blockIdx.y * FDGS_TILE + wy * 4 * NPX;
    const int px = bx0 + (lane & 7);
    const float pxf = (float)px;
    const float cx0 = (float)bx0, cy0 = (float)by0;
    const float cx1 = (float)min(bx0 + 7, W - 1), cy1 = (float)min(by0 + 4 * NPX - 1, H - 1);
#if BLEND_BWD_STREAMS == 2
    // two record streams per warp: half-warp h owns the 8 x (2 NPX) block of rows [by0 + 2 NPX h, by0 + 2 NPX (h + 1)), a thread
    // its column and the rows rp, rp + 2, ... of that block; each half walks only the records that survive ITS block's test
    const int half = lane >> 4, rp = (lane >> 3) & 1;
    const float cyA1 = (float)min(by0 + 2 * NPX - 1, H - 1), cyB0 = (float)(by0 + 2 * NPX);
    const bool blockB = by0 + 2 * NPX < H;
#endif
    const size_t HW = (size_t)H * W;
    const uint2 range = ranges[tile];

    float pyf[NPX], T_final[NPX], T[NPX], dLp0[NPX], dLp1[NPX], dLp2[NPX], bg_dot[NPX];
    float ar0[NPX], ar1[NPX], ar2[NPX], last_alpha[NPX], lc0[NPX], lc1[NPX], lc2[NPX];
    uint32_t last_contributor[NPX];
    uint32_t lmax = 0;
#pragma unroll
    for (int j = 0; j < NPX; ++j) {
#if BLEND_BWD_STREAMS == 2
        const int py = by0 + 2 * NPX * half + rp + 2 * j;
#else
        const int py = by0 + (lane >> 3) + 4 * j;
#endif
        const bool inside = px < W && py < H;
        const size_t pix_id = (size_t)py * W + px;
        pyf[j] = (float)py;
        T_final[j] = inside ? final_Ts[pix_id] : 0.f;
        last_contributor[j] = inside ? n_contrib[pix_id] : 0u;
        lmax = max(lmax, last_contributor[j]);
        dLp0[j] = dLp1[j] = dLp2[j] = 0.f;
        if (inside) { dLp0[j] = dL_dpixels[pix_id]; dLp1[j] = dL_dpixels[HW + pix_id]; dLp2[j] = dL_dpixels[2 * HW + pix_id]; }
        float bd = 0.f;
        bd += bg_color[0] * dLp0[j]; bd += bg_color[1] * dLp1[j]; bd += bg_color[2] * dLp2[j];
        bg_dot[j] = bd;
        T[j] = T_final[j];
        ar0[j] = ar1[j] = ar2[j] = 0.f;
        last_alpha[j] = lc0[j] = lc1[j] = lc2[j] = 0.f;
    }
    const float ddelx_dx = 0.5 * W, ddely_dy = 0.5 * H;

    if (threadIdx.x == 0) {
        s_maxc = 0;
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < BATCH * NACC; i += NT) (&acc[0][0])[i] = 0.f;
    __syncthreads();
    const uint32_t wmax = __reduce_max_sync(FULL, lmax);
#if BLEND_BWD_STREAMS == 2
    const uint32_t hmaxA = __reduce_max_sync(FULL, half ? 0u : lmax), hmaxB = __reduce_max_sync(FULL, half ? lmax : 0u);
#endif
    if (lane == 0 && wmax) atomicMax(&s_maxc, wmax);
    __syncthreads();
    const int total = (int)min(s_maxc, range.y - range.x);     // records beyond the last contributor are dead
    if (total == 0) return;
    const int nb = (total + BATCH - 1) / BATCH;
    // batch b (b = 0 is the BACK of the list) covers positions [lo_b, hi_b)
    auto lo_of = [&](int b) { return max(0, total - (b + 1) * BATCH); };
    auto hi_of = [&](int b) { return total - b * BATCH; };
    auto issue = [&](int b) {
        const int lo = lo_of(b), hi = hi_of(b);
        const uint32_t bytes = (uint32_t)(hi - lo) * (uint32_t)sizeof(FdgsRecord);
        mbar_expect_tx(&bars[b & 1], bytes);
        bulk_g2s(&srec[b & 1][0], records + range.x + lo, bytes, &bars[b & 1]);
    };
    if (threadIdx.x == 0) issue(0);

    for (int b = 0; b < nb; ++b) {
        if (threadIdx.x == 0 && b + 1 < nb) issue(b + 1);
        mbar_wait(&bars[b & 1], (uint32_t)((b >> 1) & 1));
        const FdgsRecord* rec = srec[b & 1];
        const int lo = lo_of(b), cnt = hi_of(b) - lo;

        if (wmax > (uint32_t)lo) {                        // warp-uniform: something here can still matter
            for (int c1 = cnt; c1 > 0; c1 -= 32) {        // chunks of 32, from the back
                const int j = c1 - 1 - lane;              // lane 0 tests the last record of the chunk
                // gradient partials of record jj for this thread's pixels -> v[0..8]
                auto eval = [&](int jj, uint32_t pos, float (&v)[10]) {
#pragma unroll
                    for (int pj = 0; pj < NPX; ++pj) {
                        if (pos < last_contributor[pj]) {
                            const float4 q0 = reinterpret_cast<const float4*>(&rec[jj])[0];
                            const float4 q1 = reinterpret_cast<const float4*>(&rec[jj])[1];
                            float dx = q0.x - pxf;                 // opaque per pixel: same FMA contraction of `power`
                            asm volatile("" : "+f"(dx));           // (and hence the same alpha tests) as the forward
                            const float dy = q0.y - pyf[pj];
                            const float power = -0.5f * (q0.z * dx * dx + q1.x * dy * dy) - q0.w * dx * dy;
                            if (!(power > 0.0f)) {
                                const float G = expf(power);
                                const float alpha = fminf(0.99f, q1.y * G);
                                if (!(alpha < 1.0f / 255.0f)) {
                                    const float4 q2 = reinterpret_cast<const float4*>(&rec[jj])[2];
                                    const float inv1ma = 1.f / (1.f - alpha);           // one division for both uses
                                    T[pj] = T[pj] * inv1ma;
                                    const float dchannel_dcolor = alpha * T[pj];
                                    float dL_dalpha = 0.0f;
                                    ar0[pj] = last_alpha[pj] * lc0[pj] + (1.f - last_alpha[pj]) * ar0[pj]; lc0[pj] = q2.x;
                                    dL_dalpha += (q2.x - ar0[pj]) * dLp0[pj];
                                    ar1[pj] = last_alpha[pj] * lc1[pj] + (1.f - last_alpha[pj]) * ar1[pj]; lc1[pj] = q2.y;
                                    dL_dalpha += (q2.y - ar1[pj]) * dLp1[pj];
                                    ar2[pj] = last_alpha[pj] * lc2[pj] + (1.f - last_alpha[pj]) * ar2[pj]; lc2[pj] = q2.z;
                                    dL_dalpha += (q2.z - ar2[pj]) * dLp2[pj];
                                    v[0] += dchannel_dcolor * dLp0[pj];
                                    v[1] += dchannel_dcolor * dLp1[pj];
                                    v[2] += dchannel_dcolor * dLp2[pj];
                                    dL_dalpha *= T[pj];
                                    last_alpha[pj] = alpha;
                                    dL_dalpha += (-T_final[pj] * inv1ma) * bg_dot[pj];
                                    const float dL_dG = q1.y * dL_dalpha;
                                    const float gdx = G * dx, gdy = G * dy;
                                    const float dG_ddelx = -gdx * q0.z - gdy * q0.w;
                                    const float dG_ddely = -gdy * q1.x - gdx * q0.w;
                                    v[3] += dL_dG * dG_ddelx * ddelx_dx;
                                    v[4] += dL_dG * dG_ddely * ddely_dy;
                                    v[5] += -0.5f * gdx * dx * dL_dG;
                                    v[6] += -0.5f * gdx * dy * dL_dG;
                                    v[7] += -0.5f * gdy * dy * dL_dG;
                                    v[8] += G * dL_dalpha;
                                }
                            }
                        }
                    }
                };
#if BLEND_BWD_STREAMS == 2
                bool keepA = false, keepB = false;
                if (j >= 0) {
                    const float4 q0 = reinterpret_cast<const float4*>(&rec[j])[0];
                    const float4 q1 = reinterpret_cast<const float4*>(&rec[j])[1];
                    if (!(q1.y <= 0.f)) {
                        const float cull = rec[j].cull;
                        keepA = ((uint32_t)(lo + j) < hmaxA) && !rect_culled(q0.x, q0.y, q0.z, q0.w, q1.x, cull, cx0, cy0, cx1, cyA1);
                        keepB = blockB && ((uint32_t)(lo + j) < hmaxB) &&
                                !rect_culled(q0.x, q0.y, q0.z, q0.w, q1.x, cull, cx0, cyB0, cx1, cy1);
                    }
                }
                unsigned mA = __ballot_sync(FULL, keepA), mB = __ballot_sync(FULL, keepB);
                unsigned mine = half ? mB : mA;
                const int hslot = ((lane >> 3) & 1) * 5 + ((lane >> 2) & 1) * 3 + ((lane >> 1) & 1) * 2 + (lane & 1);
                const bool hvalid = !(((lane >> 2) & 1) && ((lane >> 1) & 1)) && !(!((lane >> 2) & 1) && ((lane >> 1) & 1) && (lane & 1)) &&
                                    hslot < NACC;
                while (mA | mB) {
                    const int l = mine ? __ffs(mine) - 1 : 0;     // ascending lane == descending position
                    const int jj = c1 - 1 - l;
                    float v[10];
#pragma unroll
                    for (int k = 0; k < 10; ++k) v[k] = 0.f;
                    if (mine) eval(jj, (uint32_t)(lo + jj), v);
                    half_reduce10(v, lane);
                    if (mine && hvalid && v[0] != 0.f) atomicAdd(&acc[jj][hslot], v[0]);
                    mine &= mine - 1;                             // 0 stays 0
                    mA &= mA - 1; mB &= mB - 1;
                }
#else
                bool keep = false;
                if (j >= 0) {
                    const float4 q0 = reinterpret_cast<const float4*>(&rec[j])[0];
                    const float4 q1 = reinterpret_cast<const float4*>(&rec[j])[1];
                    keep = ((uint32_t)(lo + j) < wmax) && !(q1.y <= 0.f) &&
                           !rect_culled(q0.x, q0.y, q0.z, q0.w, q1.x, rec[j].cull, cx0, cy0, cx1, cy1);
                }
                unsigned m = __ballot_sync(FULL, keep);
                while (m) {
                    const int l = __ffs(m) - 1;           // ascending lane == descending position
                    m &= m - 1;
                    const int jj = c1 - 1 - l;
                    float v[10];
#pragma unroll
                    for (int k = 0; k < 10; ++k) v[k] = 0.f;
                    eval(jj, (uint32_t)(lo + jj), v);
#if BLEND_BWD_SPARSE
                    // A surviving record usually touches a handful of the warp's 64 pixels (the cull is a conservative
                    // rectangle test): v[8] (the opacity partial, G * dL/dalpha with G > 0) is non-zero exactly on the lanes that
                    // had a contributing pixel.  No such lane: nothing to add.  Up to BLEND_BWD_SPARSE lanes: they add their nine
                    // partials to the shared accumulator themselves (9 predicated atomics) instead of the 12-shuffle butterfly.
                    const unsigned hitm = __ballot_sync(FULL, v[8] != 0.f || v[0] != 0.f || v[3] != 0.f);
                    if (hitm == 0u) continue;
                    if (__popc(hitm) <= BLEND_BWD_SPARSE) {
                        if ((hitm >> lane) & 1u) {
#pragma unroll
                            for (int k = 0; k < NACC; ++k) if (v[k] != 0.f) atomicAdd(&acc[jj][k], v[k]);
                        }
                        continue;
                    }
#endif
                    warp_reduce10(v, lane);
                    const int slot = ((lane >> 4) & 1) * 5 + ((lane >> 3) & 1) * 3 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
                    if (!(lane & 1) && slot < NACC && v[0] != 0.f) atomicAdd(&acc[jj][slot], v[0]);
                }
#endif
            }
        }
        __syncthreads();                                   // batch fully accumulated
        for (int r0 = 0; r0 < BATCH; r0 += NT) {
            const int r = r0 + (int)threadIdx.x;
            if (r < cnt) {
                const uint32_t gid = rec[r].id;
                float a[NACC];
#pragma unroll
                for (int k = 0; k < NACC; ++k) { a[k] = acc[r][k]; acc[r][k] = 0.f; }
                if (a[0] != 0.f) atomicAdd(&dL_dcolors[3 * (size_t)gid + 0], a[0]);
                if (a[1] != 0.f) atomicAdd(&dL_dcolors[3 * (size_t)gid + 1], a[1]);
                if (a[2] != 0.f) atomicAdd(&dL_dcolors[3 * (size_t)gid + 2], a[2]);
                if (a[3] != 0.f) atomicAdd(&dL_dmean2D[3 * (size_t)gid + 0], a[3]);
                if (a[4] != 0.f) atomicAdd(&dL_dmean2D[3 * (size_t)gid + 1], a[4]);
                if (a[5] != 0.f) atomicAdd(&dL_dconic[4 * (size_t)gid + 0], a[5]);
                if (a[6] != 0.f) atomicAdd(&dL_dconic[4 * (size_t)gid + 1], a[6]);
                if (a[7] != 0.f) atomicAdd(&dL_dconic[4 * (size_t)gid + 3], a[7]);
                if (a[8] != 0.f) atomicAdd(&dL_dopacity[gid], a[8]);
            }
        }
        __syncthreads();                                   // srec[b&1] and acc are free again
    }
}

// d/dv of v/|v| applied to an upstream vector (reference: dnormvdv, auxiliary.h:99-110).
__device__ __forceinline__ float3 normalize_vjp(float3 v, float3 dv) {
    const float s2 = v.x * v.x + v.y * v.y + v.z * v.z;
    const float inv = 1.0f / sqrtf(s2 * s2 * s2);
    float3 r;
    r.x = ((s2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * inv;
    r.y = (-v.x * v.y * dv.x + (s2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * inv;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (s2 - v.z * v.z) * dv.z) * inv;
    return r;
}

// Fused K9 + K10 (backward.cu:144-274 and :346-396): one Gaussian.  sh_row != nullptr: shared-memory row holding the SH
// coefficients on entry and their gradient on return (see k_preprocess_bwd).
__device__ __forceinline__ void
preprocess_bwd_one(int idx, int D, int M, const float* __restrict__ means3D, const int* __restrict__ radii,
                   const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                   const float* __restrict__ scales, const float* __restrict__ rotations, float scale_modifier,
                   const float* __restrict__ cov3Ds, const float* V, const float* Pm, const float* cam,
                   float fx, float fy, float tanfovx, float tanfovy,
                   const float* __restrict__ dL_dmean2D, const float* __restrict__ dL_dconic,
                   const float* __restrict__ dL_dcolor,
                   float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh,
                   float* __restrict__ dL_dscales, float* __restrict__ dL_drots, float* sh_row,
                   int activate, const float4* __restrict__ conic_opacity, float* __restrict__ dL_dopacity)
{
    const size_t i3 = 3 * (size_t)idx;

    if (!(radii[idx] > 0)) {      // invisible: the reference leaves its zero-initialised outputs untouched
        dL_dmeans3D[i3] = 0.f; dL_dmeans3D[i3 + 1] = 0.f; dL_dmeans3D[i3 + 2] = 0.f;
#pragma unroll
        for (int k = 0; k < 6; ++k) dL_dcov3D[6 * (size_t)idx + k] = 0.f;
        if (sh_row) {
            for (int k = 0; k < M * 3; ++k) sh_row[k] = 0.f;       // staged: the CTA writes the rows out together
        } else if (dL_dsh) {
            float4* d4 = reinterpret_cast<float4*>(dL_dsh + (size_t)idx * M * 3);
            if ((M * 3) % 4 == 0 && ((uintptr_t)d4 % 16) == 0) {
                for (int k = 0; k < (M * 3) / 4; ++k) d4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                for (int k = 0; k < M * 3; ++k) dL_dsh[(size_t)idx * M * 3 + k] = 0.f;
            }
        }
        dL_dscales[i3] = 0.f; dL_dscales[i3 + 1] = 0.f; dL_dscales[i3 + 2] = 0.f;
        reinterpret_cast<float4*>(dL_drots)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        return;                   // (dL_dopacity of an invisible Gaussian is 0 already: nothing blended it)
    }
    if (activate) {               // dL/d(opacity logit) in place: sigmoid VJP with the activated opacity kept by the forward
        dL_dopacity[idx] = fdgs_act_opacity_vjp(dL_dopacity[idx], conic_opacity[idx].w);
    }

    const float3 mean = make_float3(means3D[i3], means3D[i3 + 1], means3D[i3 + 2]);
    const float* c6 = cov3Ds + 6 * (size_t)idx;
    float cc[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) cc[k] = c6[k];

    // ---- conic -> cov2D -> (cov3D, mean) -----------------------------------------------
    M3 T; float3 t; float xmul, ymul;
    const float3 cov = cov2d_project(mean, fx, fy, tanfovx, tanfovy, cc, V, &T, &t, &xmul, &ymul);
    const float a = cov.x, b = cov.y, c = cov.z;
    const float gA = dL_dconic[4 * (size_t)idx], gB = dL_dconic[4 * (size_t)idx + 1], gC = dL_dconic[4 * (size_t)idx + 3];
    const float denom = a * c - b * b;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float da = 0.f, db = 0.f, dc = 0.f;
    const float u0 = T.m[0][0], u1 = T.m[0][1], u2 = T.m[0][2];
    const float v0 = T.m[1][0], v1 = T.m[1][1], v2 = T.m[1][2];
    float dcov[6];
    if (denom2inv != 0) {
        da = denom2inv * (-c * c * gA + 2 * b * c * gB + (denom - a * c) * gC);
        dc = denom2inv * (-a * a * gC + 2 * a * b * gB + (denom - a * c) * gA);
        db = denom2inv * 2 * (b * c * gA - (denom + 2 * b * b) * gB + a * b * gC);
        dcov[0] = u0 * u0 * da + u0 * v0 * db + v0 * v0 * dc;
        dcov[3] = u1 * u1 * da + u1 * v1 * db + v1 * v1 * dc;
        dcov[5] = u2 * u2 * da + u2 * v2 * db + v2 * v2 * dc;
        dcov[1] = 2 * u0 * u1 * da + (u0 * v1 + u1 * v0) * db + 2 * v0 * v1 * dc;
        dcov[2] = 2 * u0 * u2 * da + (u0 * v2 + u2 * v0) * db + 2 * v0 * v2 * dc;
        dcov[4] = 2 * u2 * u1 * da + (u1 * v2 + u2 * v1) * db + 2 * v1 * v2 * dc;
    } else {
#pragma unroll
        for (int k = 0; k < 6; ++k) dcov[k] = 0.f;
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) dL_dcov3D[6 * (size_t)idx + k] = dcov[k];

    // V u and V v with V the symmetric 3-D covariance
    const float Vu0 = u0 * cc[0] + u1 * cc[1] + u2 * cc[2];
    const float Vu1 = u0 * cc[1] + u1 * cc[3] + u2 * cc[4];
    const float Vu2 = u0 * cc[2] + u1 * cc[4] + u2 * cc[5];
    const float Vv0 = v0 * cc[0] + v1 * cc[1] + v2 * cc[2];
    const float Vv1 = v0 * cc[1] + v1 * cc[3] + v2 * cc[4];
    const float Vv2 = v0 * cc[2] + v1 * cc[4] + v2 * cc[5];
    const float du0 = 2 * Vu0 * da + Vv0 * db, du1 = 2 * Vu1 * da + Vv1 * db, du2 = 2 * Vu2 * da + Vv2 * db;
    const float dv0 = 2 * Vv0 * dc + Vu0 * db, dv1 = 2 * Vv1 * dc + Vu1 * db, dv2 = 2 * Vv2 * dc + Vu2 * db;
    // u_r = V[4r] J00 + V[4r+2] J02 ; v_r = V[4r+1] J11 + V[4r+2] J12
    const float dJ00 = V[0] * du0 + V[4] * du1 + V[8] * du2;
    const float dJ02 = V[2] * du0 + V[6] * du1 + V[10] * du2;
    const float dJ11 = V[1] * dv0 + V[5] * dv1 + V[9] * dv2;
    const float dJ12 = V[2] * dv0 + V[6] * dv1 + V[10] * dv2;
    const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    const float dtx = xmul * -fx * tz2 * dJ02;
    const float dty = ymul * -fy * tz2 * dJ12;
    const float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t.x) * tz3 * dJ02 + (2 * fy * t.y) * tz3 * dJ12;
    float3 dmean = make_float3(V[0] * dtx + V[1] * dty + V[2] * dtz,
                               V[4] * dtx + V[5] * dty + V[6] * dtz,
                               V[8] * dtx + V[9] * dty + V[10] * dtz);

    // ---- mean2D -> mean3D through the perspective divide -------------------------------
    {
        const float4 mh = xform44(mean, Pm);
        const float mw = 1.0f / (mh.w + 0.0000001f);
        const float mul1 = mh.x * mw * mw, mul2 = mh.y * mw * mw;
        const float g2x = dL_dmean2D[i3], g2y = dL_dmean2D[i3 + 1];
        dmean.x += (Pm[0] * mw - Pm[3] * mul1) * g2x + (Pm[1] * mw - Pm[3] * mul2) * g2y;
        dmean.y += (Pm[4] * mw - Pm[7] * mul1) * g2x + (Pm[5] * mw - Pm[7] * mul2) * g2y;
        dmean.z += (Pm[8] * mw - Pm[11] * mul1) * g2x + (Pm[9] * mw - Pm[11] * mul2) * g2y;
    }

    // ---- colour -> SH (+ view-direction dependence on the mean) ------------------------
    if (shs) {
        const float3 dir_o = make_float3(mean.x - cam[0], mean.y - cam[1], mean.z - cam[2]);
        const float len = sqrtf(dir_o.x * dir_o.x + dir_o.y * dir_o.y + dir_o.z * dir_o.z);
        const float x = dir_o.x / len, y = dir_o.y / len, z = dir_o.z / len;
        const uint8_t cl = clamped[idx];
        float g[3];
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) g[ch] = ((cl >> ch) & 1) ? 0.f : dL_dcolor[i3 + ch];
        // staged: the row holds this Gaussian's SH coefficients and is overwritten in place by their gradient
        const float* sh = sh_row ? sh_row : shs + (size_t)idx * M * 3;
        float* dsh = sh_row ? sh_row : dL_dsh + (size_t)idx * M * 3;
        // basis values and their x/y/z derivatives for the active degree
        float bas[16], bx[16], by[16], bz[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) { bas[k] = 0.f; bx[k] = 0.f; by[k] = 0.f; bz[k] = 0.f; }
        bas[0] = FDGS_SH_C0;
        int nact = 1;
        if (D > 0) {
            nact = 4;
            bas[1] = -FDGS_SH_C1 * y; by[1] = -FDGS_SH_C1;
            bas[2] = FDGS_SH_C1 * z;  bz[2] = FDGS_SH_C1;
            bas[3] = -FDGS_SH_C1 * x; bx[3] = -FDGS_SH_C1;
            if (D > 1) {
                nact = 9;
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                bas[4] = FDGS_SH_C2_0 * xy;                    bx[4] = FDGS_SH_C2_0 * y; by[4] = FDGS_SH_C2_0 * x;
                bas[5] = FDGS_SH_C2_1 * yz;                    by[5] = FDGS_SH_C2_1 * z; bz[5] = FDGS_SH_C2_1 * y;
                bas[6] = FDGS_SH_C2_2 * (2.f * zz - xx - yy);  bx[6] = FDGS_SH_C2_2 * 2.f * -x; by[6] = FDGS_SH_C2_2 * 2.f * -y; bz[6] = FDGS_SH_C2_2 * 4.f * z;
                bas[7] = FDGS_SH_C2_3 * xz;                    bx[7] = FDGS_SH_C2_3 * z; bz[7] = FDGS_SH_C2_3 * x;
                bas[8] = FDGS_SH_C2_4 * (xx - yy);             bx[8] = FDGS_SH_C2_4 * 2.f * x; by[8] = FDGS_SH_C2_4 * 2.f * -y;
                if (D > 2) {
                    nact = 16;
                    bas[9] = FDGS_SH_C3_0 * y * (3.f * xx - yy);
                    bx[9] = FDGS_SH_C3_0 * 6.f * xy; by[9] = FDGS_SH_C3_0 * 3.f * (xx - yy);
                    bas[10] = FDGS_SH_C3_1 * xy * z;
                    bx[10] = FDGS_SH_C3_1 * yz; by[10] = FDGS_SH_C3_1 * xz; bz[10] = FDGS_SH_C3_1 * xy;
                    bas[11] = FDGS_SH_C3_2 * y * (4.f * zz - xx - yy);
                    bx[11] = FDGS_SH_C3_2 * -2.f * xy; by[11] = FDGS_SH_C3_2 * (-3.f * yy + 4.f * zz - xx); bz[11] = FDGS_SH_C3_2 * 8.f * yz;
                    bas[12] = FDGS_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
                    bx[12] = FDGS_SH_C3_3 * -6.f * xz; by[12] = FDGS_SH_C3_3 * -6.f * yz; bz[12] = FDGS_SH_C3_3 * 3.f * (2.f * zz - xx - yy);
                    bas[13] = FDGS_SH_C3_4 * x * (4.f * zz - xx - yy);
                    bx[13] = FDGS_SH_C3_4 * (-3.f * xx + 4.f * zz - yy); by[13] = FDGS_SH_C3_4 * -2.f * xy; bz[13] = FDGS_SH_C3_4 * 8.f * xz;
                    bas[14] = FDGS_SH_C3_5 * z * (xx - yy);
                    bx[14] = FDGS_SH_C3_5 * 2.f * xz; by[14] = FDGS_SH_C3_5 * -2.f * yz; bz[14] = FDGS_SH_C3_5 * (xx - yy);
                    bas[15] = FDGS_SH_C3_6 * x * (xx - 3.f * yy);
                    bx[15] = FDGS_SH_C3_6 * 3.f * (xx - yy); by[15] = FDGS_SH_C3_6 * -6.f * xy;
                }
            }
        }
        float ddx = 0.f, ddy = 0.f, ddz = 0.f;      // dL/d(dir)
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k < nact) {
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    const float s = sh[3 * k + ch];
                    dsh[3 * k + ch] = bas[k] * g[ch];
                    const float sg = s * g[ch];
                    ddx += bx[k] * sg; ddy += by[k] * sg; ddz += bz[k] * sg;
                }
            }
        }
        for (int k = nact; k < M; ++k) { dsh[3 * k] = 0.f; dsh[3 * k + 1] = 0.f; dsh[3 * k + 2] = 0.f; }
        const float3 dm = normalize_vjp(dir_o, make_float3(ddx, ddy, ddz));
        dmean.x += dm.x; dmean.y += dm.y; dmean.z += dm.z;
    }
    dL_dmeans3D[i3] = dmean.x; dL_dmeans3D[i3 + 1] = dmean.y; dL_dmeans3D[i3 + 2] = dmean.z;

    // ---- cov3D -> scale, rotation ------------------------------------------------------
    if (scales) {
        float3 sc = make_float3(scales[i3], scales[i3 + 1], scales[i3 + 2]);
        const float4 q_in = reinterpret_cast<const float4*>(rotations)[idx];
        float4 q = q_in;
        if (activate) {           // the inputs are the raw deformation outputs: redo exp / normalize (forward did the same)
            sc = make_float3(fdgs_act_scale(sc.x), fdgs_act_scale(sc.y), fdgs_act_scale(sc.z));
            q = fdgs_act_quat(q_in);
        }
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        const M3 R = quat_to_m3(r, x, y, z);
        const float s[3] = { scale_modifier * sc.x, scale_modifier * sc.y, scale_modifier * sc.z };
        // M[c][r] = s_r R[c][r];  dSigma symmetric with halved off-diagonals
        M3 Mm, dS;
#pragma unroll
        for (int cI = 0; cI < 3; ++cI)
#pragma unroll
            for (int rI = 0; rI < 3; ++rI) Mm.m[cI][rI] = s[rI] * R.m[cI][rI];
        dS.m[0][0] = dcov[0]; dS.m[0][1] = 0.5f * dcov[1]; dS.m[0][2] = 0.5f * dcov[2];
        dS.m[1][0] = 0.5f * dcov[1]; dS.m[1][1] = dcov[3]; dS.m[1][2] = 0.5f * dcov[4];
        dS.m[2][0] = 0.5f * dcov[2]; dS.m[2][1] = 0.5f * dcov[4]; dS.m[2][2] = dcov[5];
        M3 dM = m3_mul(Mm, dS);          // dL/dM = 2 M dSigma
#pragma unroll
        for (int cI = 0; cI < 3; ++cI)
#pragma unroll
            for (int rI = 0; rI < 3; ++rI) dM.m[cI][rI] *= 2.0f;
        // n[i][j] = dL_dM^T [i][j] = dM[j][i]
        float n[3][3];
        float ds[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            ds[i] = R.m[0][i] * dM.m[0][i] + R.m[1][i] * dM.m[1][i] + R.m[2][i] * dM.m[2][i];
#pragma unroll
            for (int j = 0; j < 3; ++j) n[i][j] = dM.m[j][i] * s[i];
        }
        if (activate) {           // exp VJP: dL/d(log scale) = dL/ds * s
            ds[0] = fdgs_act_scale_vjp(ds[0], sc.x); ds[1] = fdgs_act_scale_vjp(ds[1], sc.y); ds[2] = fdgs_act_scale_vjp(ds[2], sc.z);
        }
        dL_dscales[i3] = ds[0]; dL_dscales[i3 + 1] = ds[1]; dL_dscales[i3 + 2] = ds[2];
        float4 dq;
        dq.x = 2 * z * (n[0][1] - n[1][0]) + 2 * y * (n[2][0] - n[0][2]) + 2 * x * (n[1][2] - n[2][1]);
        dq.y = 2 * y * (n[1][0] + n[0][1]) + 2 * z * (n[2][0] + n[0][2]) + 2 * r * (n[1][2] - n[2][1]) - 4 * x * (n[2][2] + n[1][1]);
        dq.z = 2 * x * (n[1][0] + n[0][1]) + 2 * r * (n[2][0] - n[0][2]) + 2 * z * (n[1][2] + n[2][1]) - 4 * y * (n[2][2] + n[0][0]);
        dq.w = 2 * r * (n[0][1] - n[1][0]) + 2 * x * (n[2][0] + n[0][2]) + 2 * y * (n[1][2] + n[2][1]) - 4 * z * (n[1][1] + n[0][0]);
        if (activate) dq = fdgs_act_quat_vjp(q_in, q, dq);     // normalize VJP: gradient w.r.t. the raw quaternion
        reinterpret_cast<float4*>(dL_drots)[idx] = dq;
    } else {
        dL_dscales[i3] = 0.f; dL_dscales[i3 + 1] = 0.f; dL_dscales[i3 + 2] = 0.f;
        reinterpret_cast<float4*>(dL_drots)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// Thread per Gaussian.  A thread's 3M SH floats (read) and their gradient (written) are contiguous, so thread-per-Gaussian
// accesses touch 32 different lines per instruction: with sh_staged the CTA's [256][3M] block goes through shared memory
// (row stride 3M + 1) with coalesced float4 loads and stores, the gradient overwriting the coefficients in place.
__global__ void __launch_bounds__(256)
k_preprocess_bwd(int P, int D, int M, const float* __restrict__ means3D, const int* __restrict__ radii,
                 const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
                 const float* __restrict__ scales, const float* __restrict__ rotations, float scale_modifier,
                 const float* __restrict__ cov3Ds, const float* __restrict__ viewmatrix,
                 const float* __restrict__ projmatrix, float fx, float fy, float tanfovx, float tanfovy,
                 const float* __restrict__ campos,
                 const float* __restrict__ dL_dmean2D, const float* __restrict__ dL_dconic,
                 const float* __restrict__ dL_dcolor,
                 float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh,
                 float* __restrict__ dL_dscales, float* __restrict__ dL_drots, int sh_staged,
                 int activate, const float4* __restrict__ conic_opacity, float* __restrict__ dL_dopacity)
{
    __shared__ float V[16], Pm[16], cam[3];
    extern __shared__ float s_sh[];
    if (threadIdx.x < 16) { V[threadIdx.x] = viewmatrix[threadIdx.x]; Pm[threadIdx.x] = projmatrix[threadIdx.x]; }
    if (threadIdx.x < 3) cam[threadIdx.x] = campos[threadIdx.x];
    const int rowf = 3 * M;
    const long long base = (long long)blockIdx.x * blockDim.x * rowf;
    const long long avail = (long long)P * rowf - base;
    const int n = (int)(avail < (long long)blockDim.x * rowf ? avail : (long long)blockDim.x * rowf);
    constexpr int RF = 48;                       // sh_staged implies M == 16: a compile-time row length keeps the index
                                                 // arithmetic below free of integer divisions (8 per float4 before)
    if (sh_staged) {
        const float4* src = reinterpret_cast<const float4*>(shs + base);
        for (int i = threadIdx.x; i * 4 < n; i += blockDim.x) {
            const float4 v = src[i];
            const int r = i / (RF / 4), c = (i - r * (RF / 4)) * 4;      // a float4 never straddles two rows (RF % 4 == 0)
            float* d = s_sh + r * (RF + 1) + c;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
    }
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < P)
        preprocess_bwd_one(idx, D, M, means3D, radii, shs, clamped, scales, rotations, scale_modifier, cov3Ds, V, Pm, cam, fx, fy,
                           tanfovx, tanfovy, dL_dmean2D, dL_dconic, dL_dcolor, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drots,
                           sh_staged ? s_sh + threadIdx.x * (rowf + 1) : nullptr, activate, conic_opacity, dL_dopacity);
    if (sh_staged) {
        __syncthreads();
        float4* dst = reinterpret_cast<float4*>(dL_dsh + base);
        for (int i = threadIdx.x; i * 4 < n; i += blockDim.x) {
            const int r = i / (RF / 4), c = (i - r * (RF / 4)) * 4;
            const float* d = s_sh + r * (RF + 1) + c;
            dst[i] = make_float4(d[0], d[1], d[2], d[3]);
        }
    }
}

}  // namespace

extern "C" int fdgs_raster_backward(const FdgsRasterBackward* a, void* stream_) {
    if (!a || a->P < 0 || a->W <= 0 || a->H <= 0) return FDGS_E_BADARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int P = a->P;
    if (P == 0) return FDGS_OK;
    if (!a->means3D || !a->radii || !a->geom_ws || !a->img_ws || !a->dL_dpix || !a->viewmatrix || !a->projmatrix ||
        !a->campos || !a->background) return FDGS_E_BADARG;
    if (!a->dL_dmeans2D || !a->dL_dconic || !a->dL_dopacity || !a->dL_dcolors || !a->dL_dmeans3D || !a->dL_dcov3D ||
        !a->dL_dscales || !a->dL_drots) return FDGS_E_BADARG;
    if (a->shs && !a->dL_dsh) return FDGS_E_BADARG;
    if (a->R > 0 && !a->bin_ws) return FDGS_E_BADARG;
    const int gx = (a->W + FDGS_TILE - 1) / FDGS_TILE, gy = (a->H + FDGS_TILE - 1) / FDGS_TILE;
    FdgsGeom g = FdgsGeom::carve(const_cast<void*>(a->geom_ws), P);
    FdgsImg im = FdgsImg::carve(const_cast<void*>(a->img_ws), a->W, a->H);
    FdgsBin bn = FdgsBin::carve(const_cast<void*>(a->bin_ws), a->R);
    const float focal_y = a->H / (2.0f * a->tanfovy);
    const float focal_x = a->W / (2.0f * a->tanfovx);

    FDGS_CUDA(cudaMemsetAsync(a->dL_dmeans2D, 0, sizeof(float) * 3 * (size_t)P, stream));
    FDGS_CUDA(cudaMemsetAsync(a->dL_dconic, 0, sizeof(float) * 4 * (size_t)P, stream));
    FDGS_CUDA(cudaMemsetAsync(a->dL_dopacity, 0, sizeof(float) * (size_t)P, stream));
    FDGS_CUDA(cudaMemsetAsync(a->dL_dcolors, 0, sizeof(float) * 3 * (size_t)P, stream));
    if (a->R > 0) {
        dim3 grid(gx, gy);
        { FDGS_PROF("k_blend_bwd", stream);
        k_blend_bwd<BLEND_BWD_NPX><<<grid, 256 / BLEND_BWD_NPX, 0, stream>>>(im.ranges, bn.records, a->W, a->H, a->background, im.final_T, im.n_contrib,
                                              a->dL_dpix, a->dL_dmeans2D, a->dL_dconic, a->dL_dopacity, a->dL_dcolors);
        }
        FDGS_LAUNCH_CHECK(a->debug, stream);
    }
    const float* cov3D_ptr = a->cov3D_precomp ? a->cov3D_precomp : g.cov3D;
    const int sh_staged = (a->shs && a->dL_dsh && a->M == 16 && (((uintptr_t)a->shs | (uintptr_t)a->dL_dsh) & 15) == 0) ? 1 : 0;
    const size_t sh_smem = sh_staged ? (size_t)256 * (3 * a->M + 1) * sizeof(float) : 0;
    if (sh_staged) FDGS_CUDA(cudaFuncSetAttribute(k_preprocess_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh_smem));
    { FDGS_PROF("k_preprocess_bwd", stream);
    k_preprocess_bwd<<<(P + 255) / 256, 256, sh_smem, stream>>>(
        P, a->D, a->M, a->means3D, a->radii, a->shs, g.clamped, a->scales, a->rotations, a->scale_modifier, cov3D_ptr,
        a->viewmatrix, a->projmatrix, focal_x, focal_y, a->tanfovx, a->tanfovy, a->campos, a->dL_dmeans2D, a->dL_dconic,
        a->dL_dcolors, a->dL_dmeans3D, a->dL_dcov3D, a->dL_dsh, a->dL_dscales, a->dL_drots, sh_staged,
        (a->activate && a->scales && !a->cov3D_precomp) ? 1 : 0, g.conic_opacity, a->dL_dopacity);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);
    return FDGS_OK;
}
