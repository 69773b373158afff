// Forward rasterizer for sm_100a: preprocess + tile histogram, tile scan, tile-bucketed
// scatter, per-tile segmented sort + record packing, and the front-to-back blend with
// bulk-async (TMA) record staging and warp-level culling / early termination.
//
// Replaces CudaRasterizer::Rasterizer::forward
// (RAST/cuda_rasterizer/rasterizer_impl.cu:198-339, forward.cu:155-379).
#include "raster_math.cuh"
#include <cstdint>
#include <cstdio>
#include "prof.cuh"

namespace {

constexpr unsigned FULL = 0xffffffffu;

// ---------------------------------------------------------------------------
// Warp-cooperative iteration over the tile rectangles of the 32 Gaussians a warp
// owns.  Small rects are walked by their own lane; rects with more than 32 tiles
// are walked by the whole warp (one tile per lane) so that a few screen-filling
// Gaussians do not serialise a warp (reference: per-thread serial loop,
// rasterizer_impl.cu:95-109).  All 32 lanes must call this.
template <class Op>
__device__ __forceinline__ void for_each_tile(int x0, int y0, int x1, int y1, uint32_t payload_lo, uint32_t payload_hi,
                                              int gx, Op op) {
    const int lane = threadIdx.x & 31;
    const int w = x1 - x0, h = y1 - y0;
    const int n = w * h;
    const bool big = n > 32;
    if (!big) {
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) op(y * gx + x, payload_lo, payload_hi);
    }
    unsigned m = __ballot_sync(FULL, big);
    while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const int sx0 = __shfl_sync(FULL, x0, src), sy0 = __shfl_sync(FULL, y0, src);
        const int sw = __shfl_sync(FULL, w, src), sn = __shfl_sync(FULL, n, src);
        const uint32_t plo = __shfl_sync(FULL, payload_lo, src), phi = __shfl_sync(FULL, payload_hi, src);
        for (int i = lane; i < sn; i += 32) {
            const int ty = sy0 + i / sw, tx = sx0 + i % sw;
            op(ty * gx + tx, plo, phi);
        }
    }
}

// ---------------------------------------------------------------------------
// K1: per-Gaussian preprocess + per-tile instance histogram.
__global__ void __launch_bounds__(256)
k_preprocess(int P, int D, int M,
             const float* __restrict__ means3D, const float* __restrict__ scales, float scale_modifier,
             const float* __restrict__ rotations, const float* __restrict__ opacities,
             const float* __restrict__ shs, const float* __restrict__ cov3D_precomp,
             const float* __restrict__ colors_precomp,
             const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
             const float* __restrict__ campos,
             int W, int H, float tanfovx, float tanfovy, float focal_x, float focal_y,
             int gx, int gy, int prefiltered,
             int* __restrict__ radii, FdgsGeom g, uint32_t* __restrict__ tile_count, int sh_staged, int activate)
{
    __shared__ float sV[16], sPm[16], sCam[3];
    extern __shared__ float s_sh[];              // [256][SH_STRIDE] when the SH block is staged (sh_staged)
    if (threadIdx.x < 16) { sV[threadIdx.x] = viewmatrix[threadIdx.x]; sPm[threadIdx.x] = projmatrix[threadIdx.x]; }
    if (threadIdx.x < 3) sCam[threadIdx.x] = campos[threadIdx.x];
    // A thread's 3M SH floats are contiguous, so thread-per-Gaussian loads touch 32 different lines per instruction.
    // Stage the CTA's [256][3M] block with coalesced float4 loads instead (row stride 3M + 1: conflict-free reads).
    if (sh_staged) {
        const int rowf = 3 * M;
        const long long base = (long long)blockIdx.x * blockDim.x * rowf;
        const long long avail = (long long)P * rowf - base;
        const int n = (int)(avail < (long long)blockDim.x * rowf ? avail : (long long)blockDim.x * rowf);
        const float4* src = reinterpret_cast<const float4*>(shs + base);          // base is a multiple of 4 floats
        for (int i = threadIdx.x; i * 4 < n; i += blockDim.x) {
            const float4 v = src[i];                                              // n is a multiple of 4 (rowf = 48)
            constexpr int RF = 48;               // sh_staged implies M == 16: compile-time row length, no integer divisions
            const int r = i / (RF / 4), c = (i - r * (RF / 4)) * 4;              // a float4 never straddles two rows
            float* d = s_sh + r * (RF + 1) + c;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
    }
    __syncthreads();

    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    int rx0 = 0, ry0 = 0, rx1 = 0, ry1 = 0;     // empty rect unless the Gaussian survives
    int my_radius_i = 0;
    uint32_t touched = 0;

    if (idx < P) {
        const float3 p = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
        const float3 p_view = xform43(p, sV);
        bool alive = p_view.z > 0.2f;                       // near cull (auxiliary.h:154)
        if (!alive && prefiltered) {
            printf("Point is filtered although prefiltered is set. This shouldn't happen!");
            __trap();
        }
        if (alive) {
            const float4 p_hom = xform44(p, sPm);
            const float p_w = 1.0f / (p_hom.w + 0.0000001f);
            const float3 p_proj = make_float3(p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w);

            float c6[6];
            if (cov3D_precomp != nullptr) {
#pragma unroll
                for (int i = 0; i < 6; ++i) c6[i] = cov3D_precomp[6 * (size_t)idx + i];
            } else {
                float3 s = make_float3(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]);
                float4 q = reinterpret_cast<const float4*>(rotations)[idx];
                if (activate) {         // raw deformation outputs: exp / normalize here (gaussian_renderer/__init__.py:97-98)
                    s = make_float3(fdgs_act_scale(s.x), fdgs_act_scale(s.y), fdgs_act_scale(s.z));
                    q = fdgs_act_quat(q);
                }
                cov3d_from_scale_rot(s, scale_modifier, q, c6);
#pragma unroll
                for (int i = 0; i < 6; ++i) g.cov3D[6 * (size_t)idx + i] = c6[i];
            }
            const float3 cov = cov2d_project(p, focal_x, focal_y, tanfovx, tanfovy, c6, sV);
            const float det = cov.x * cov.z - cov.y * cov.y;
            if (det != 0.0f) {
                const float det_inv = 1.f / det;
                const float3 conic = make_float3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);
                const float mid = 0.5f * (cov.x + cov.z);
                const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
                const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
                const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
                const float2 pix = make_float2(fdgs_ndc2pix(p_proj.x, W), fdgs_ndc2pix(p_proj.y, H));
                int x0, y0, x1, y1;
                fdgs_tile_rect(pix.x, pix.y, (int)my_radius, gx, gy, x0, y0, x1, y1);
                // note: the reference hands the float radius to getRect(int) => truncation == (int)
                if ((x1 - x0) * (y1 - y0) != 0) {
                    if (colors_precomp == nullptr) {
                        float3 dir = make_float3(p.x - sCam[0], p.y - sCam[1], p.z - sCam[2]);
                        const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
                        dir.x = dir.x / len; dir.y = dir.y / len; dir.z = dir.z / len;
                        const float* sh = sh_staged ? s_sh + threadIdx.x * (3 * M + 1) : shs + (size_t)idx * M * 3;
                        float rgb[3];
                        uint8_t clampbits = 0;
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            const float v = sh_eval_channel(D, dir.x, dir.y, dir.z, [&](int k) { return sh[3 * k + c]; });
                            if (v < 0) clampbits |= (uint8_t)(1u << c);
                            rgb[c] = fmaxf(v, 0.0f);
                        }
                        g.rgb[3 * (size_t)idx + 0] = rgb[0];
                        g.rgb[3 * (size_t)idx + 1] = rgb[1];
                        g.rgb[3 * (size_t)idx + 2] = rgb[2];
                        g.clamped[idx] = clampbits;
                    }
                    g.depths[idx] = p_view.z;
                    g.means2D[idx] = pix;
                    g.conic_opacity[idx] = make_float4(conic.x, conic.y, conic.z,
                                                       activate ? fdgs_act_opacity(opacities[idx]) : opacities[idx]);
                    my_radius_i = (int)my_radius;
                    rx0 = x0; ry0 = y0; rx1 = x1; ry1 = y1;
                    touched = (uint32_t)((y1 - y0) * (x1 - x0));
                }
            }
        }
        radii[idx] = my_radius_i;
        g.tiles_touched[idx] = touched;
    }
    // histogram: one count per (Gaussian, tile) instance
    for_each_tile(rx0, ry0, rx1, ry1, 0u, 0u, gx,
                  [&](int tile, uint32_t, uint32_t) { atomicAdd(&tile_count[tile], 1u); });
}

// ---------------------------------------------------------------------------
// K2: exclusive scan over the per-tile counts (T <= ~10^4: one CTA), tile ranges, R, max.
__global__ void __launch_bounds__(1024)
k_tile_scan(int T, const uint32_t* __restrict__ tile_count, uint32_t* __restrict__ tile_start,
            uint2* __restrict__ ranges, uint32_t* __restrict__ scalars, uint32_t* __restrict__ large_tiles, uint32_t cap)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry_s, max_s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) { carry_s = 0; max_s = 0; }
    __syncthreads();
    uint32_t local_max = 0;
    for (int base = 0; base < T; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t c = (i < T) ? tile_count[i] : 0u;
        local_max = max(local_max, c);
        uint32_t v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(FULL, v, o);
            if (lane >= o) v += n;
        }
        if (lane == 31) warp_sums[wid] = v;
        __syncthreads();
        if (wid == 0) {
            uint32_t s = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(FULL, s, o);
                if (lane >= o) s += n;
            }
            warp_sums[lane] = s;
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        const uint32_t incl = carry + v + (wid > 0 ? warp_sums[wid - 1] : 0u);
        if (i < T) {
            const uint32_t start = incl - c;
            tile_start[i] = start;
            ranges[i] = c ? make_uint2(start, incl) : make_uint2(0u, 0u);
            if (c > FDGS_SORT_SMEM_CAP) large_tiles[atomicAdd(&scalars[2], 1u)] = (uint32_t)i;    // scalars were zeroed by the memset
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = incl;
        __syncthreads();
    }
    atomicMax(&max_s, local_max);
    __syncthreads();
    if (threadIdx.x == 0) {
        scalars[0] = carry_s;
        scalars[1] = max_s;
        scalars[3] = carry_s > cap ? 1u : 0u;
    }
    // One-shot mode (fdgs_raster_forward): the caller sized the instance workspace from an ESTIMATE of R.  When the real
    // count does not fit, every tile range is emptied and the large-tile list cleared, so that the scatter / sort / blend
    // kernels queued behind this one touch nothing; the host finds scalars[3] = 1 when it reads the counts back.
    if (carry_s > cap) {
        for (int i = threadIdx.x; i < T; i += 1024) ranges[i] = make_uint2(0u, 0u);
        if (threadIdx.x == 0) scalars[2] = 0u;
    }
}

// ---------------------------------------------------------------------------
// K3: scatter (depth_bits, id) into the tile buckets (order inside a bucket is arbitrary;
// the segmented sort below restores the reference's stable order).
__global__ void __launch_bounds__(256)
k_scatter(int P, const int* __restrict__ radii, FdgsGeom g, int gx, int gy,
          const uint32_t* __restrict__ tile_start, uint32_t* __restrict__ tile_cursor,
          uint64_t* __restrict__ bucket, const uint32_t* __restrict__ scalars)
{
    if (scalars[3]) return;                        // instance workspace too small (one-shot mode): see k_tile_scan
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    uint32_t dbits = 0;
    if (idx < P) {
        const int r = radii[idx];
        if (r > 0) {
            const float2 p = g.means2D[idx];
            fdgs_tile_rect(p.x, p.y, r, gx, gy, x0, y0, x1, y1);
            dbits = __float_as_uint(g.depths[idx]);
        }
    }
    const int w = x1 - x0, n = w * (y1 - y0);
    const uint64_t key = ((uint64_t)dbits << 32) | (uint32_t)idx;
    const bool big = n > 32;
    // Rects of up to 32 tiles (all but a few screen-filling Gaussians): the thread's cursor atomics are issued in batches of
    // 8 BEFORE any of their results is used.  One atomic per loop iteration left the kernel waiting for a ~1 us round trip per
    // instance (7 % issue utilisation, ncu: 157 long-scoreboard stalls per issued instruction).
    if (!big) {
        for (int base = 0; base < n; base += 8) {
            uint32_t pos[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int i = base + k;
                if (i < n) {
                    const int tile = (y0 + i / w) * gx + (x0 + i % w);
                    pos[k] = tile_start[tile] + atomicAdd(&tile_cursor[tile], 1u);
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) if (base + k < n) bucket[pos[k]] = key;
        }
    }
    // larger rects: the whole warp walks them, one tile per lane
    unsigned m = __ballot_sync(FULL, big);
    while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const int sx0 = __shfl_sync(FULL, x0, src), sy0 = __shfl_sync(FULL, y0, src);
        const int sw = __shfl_sync(FULL, w, src), sn = __shfl_sync(FULL, n, src);
        const uint32_t klo = __shfl_sync(FULL, (uint32_t)key, src), khi = __shfl_sync(FULL, (uint32_t)(key >> 32), src);
        for (int i = lane; i < sn; i += 32) {
            const int tile = (sy0 + i / sw) * gx + (sx0 + i % sw);
            bucket[tile_start[tile] + atomicAdd(&tile_cursor[tile], 1u)] = ((uint64_t)khi << 32) | klo;
        }
    }
}

// ---------------------------------------------------------------------------
// K4: per-tile segmented sort by (depth_bits, id) — equal to the reference's stable radix sort of
// (tile | depth) keys emitted in ascending id order (cub::DeviceRadixSort::SortPairs,
// rasterizer_impl.cu:304-309) — fused with the packing of the sorted instance records.
//   * buckets of up to FDGS_SORT_SMEM_CAP instances (every tile of an ordinary view): bitonic network on the
//     64-bit keys in shared memory, +inf virtual padding, power-of-two index arithmetic, and warp-level barriers for
//     every stage whose partners are <= 32 apart (a warp's 32 pair slots then cover one aligned run of 64 keys);
//   * larger buckets (clustered scenes): k_tile_sort_large below — a stable LSD radix sort in global memory.
__device__ __forceinline__ void cmpx(uint64_t* a, int i, int j) {
    const uint64_t x = a[i], y = a[j];
    if (x > y) { a[i] = y; a[j] = x; }
}

// blockDim.x threads, n <= 2 * MAXPAIRS keys.  `wide` stages (partner distance >= 64) need block barriers.
__device__ __forceinline__ void bitonic_sort_smem(uint64_t* a, int n) {
    int lg = 1;
    while ((1 << lg) < n) ++lg;                                   // npad = 1 << lg >= 2
    const int half = 1 << (lg - 1);
    bool prev_wide = true;                                         // the load into shared memory was a block-wide step
    for (int lk = 1; lk <= lg; ++lk) {                             // k = 1 << lk
        const int lhk = lk - 1, hk = 1 << lhk, k = 1 << lk;
        {   // flip step: (blk * k + off, blk * k + k - 1 - off), off < k / 2
            const bool wide = hk > 32;
            if (wide || prev_wide) __syncthreads(); else __syncwarp();
            for (int i = threadIdx.x; i < half; i += blockDim.x) {
                const int off = i & (hk - 1);
                const int lo = ((i >> lhk) << lk) + off, hi = lo + (k - 1 - 2 * off);
                if (hi < n) cmpx(a, lo, hi);
            }
            prev_wide = wide;
        }
        for (int lj = lhk - 1; lj >= 0; --lj) {                    // j = 1 << lj
            const int j = 1 << lj;
            const bool wide = j > 32;
            if (wide || prev_wide) __syncthreads(); else __syncwarp();
            for (int i = threadIdx.x; i < half; i += blockDim.x) {
                const int lo = ((i >> lj) << (lj + 1)) + (i & (j - 1)), hi = lo + j;
                if (hi < n) cmpx(a, lo, hi);
            }
            prev_wide = wide;
        }
    }
    __syncthreads();
}

// sorted keys a[0..n) of `tile` -> packed blend records (+ the reference-layout parity lists)
__device__ __forceinline__ void pack_records(const uint64_t* a, int n, uint32_t tile, uint32_t first, const FdgsGeom& g,
                                             const float* __restrict__ colors, FdgsRecord* __restrict__ records,
                                             uint64_t* __restrict__ keys_out, uint32_t* __restrict__ point_list_out) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const uint64_t k = a[i];
        const uint32_t id = (uint32_t)k;
        const float2 xy = g.means2D[id];
        const float4 co = g.conic_opacity[id];
        FdgsRecord rec;
        rec.x = xy.x; rec.y = xy.y; rec.ca = co.x; rec.cb = co.y;
        rec.cc = co.z; rec.opacity = co.w; rec.depth = __uint_as_float((uint32_t)(k >> 32)); rec.id = id;
        rec.r = colors[3 * (size_t)id]; rec.g = colors[3 * (size_t)id + 1]; rec.b = colors[3 * (size_t)id + 2];
        // cull threshold: a pixel can only pass alpha >= 1/255 if power >= ln(1/(255*opacity)).
        rec.cull = (co.w <= 0.f) ? 0.f : -logf(255.0f * co.w);   // NaN opacity => NaN => never culled
        float4* dst = reinterpret_cast<float4*>(records + first + i);
        const float4* src = reinterpret_cast<const float4*>(&rec);
        dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
        if (keys_out) keys_out[first + i] = ((uint64_t)tile << 32) | (k >> 32);
        if (point_list_out) point_list_out[first + i] = id;
    }
}

__global__ void __launch_bounds__(256)
k_tile_sort(const uint2* __restrict__ ranges, const uint64_t* __restrict__ bucket, FdgsGeom g,
            const float* __restrict__ colors, FdgsRecord* __restrict__ records,
            uint64_t* __restrict__ keys_out, uint32_t* __restrict__ point_list_out)
{
    __shared__ __align__(16) uint64_t skeys[FDGS_SORT_SMEM_CAP];
    const uint32_t tile = blockIdx.x;
    const uint2 rg = ranges[tile];
    const int n = (int)(rg.y - rg.x);
    if (n == 0 || n > FDGS_SORT_SMEM_CAP) return;                  // large buckets: k_tile_sort_large
    const uint64_t* seg = bucket + rg.x;
    for (int i = threadIdx.x; i < n; i += blockDim.x) skeys[i] = seg[i];
    if (n > 1) bitonic_sort_smem(skeys, n); else __syncthreads();
    pack_records(skeys, n, tile, rg.x, g, colors, records, keys_out, point_list_out);
}

// Buckets above the shared-memory cap: one CTA of 1024 threads per large tile (persistent over the device-built list),
// stable LSD radix sort with 8-bit digits on the 64-bit (depth, id) key, ping-ponging between the bucket segment and the
// tile's own (not yet written) slice of the record array, which is 6x as large.  A pass whose digit is the same for
// every key (the id bytes above log2 P, the exponent byte of the depth) is skipped after its histogram.
constexpr int LARGE_THREADS = 1024;
__global__ void __launch_bounds__(LARGE_THREADS)
k_tile_sort_large(const uint2* __restrict__ ranges, const uint32_t* __restrict__ large_tiles, const uint32_t* __restrict__ scalars,
                  uint64_t* __restrict__ bucket, FdgsGeom g, const float* __restrict__ colors, FdgsRecord* __restrict__ records,
                  uint64_t* __restrict__ keys_out, uint32_t* __restrict__ point_list_out)
{
    __shared__ uint32_t wc[32][256];          // per-warp digit counts of one 1024-key chunk -> exclusive prefix over warps
    __shared__ uint32_t hist[256], base[256];
    __shared__ uint32_t wsum[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t nlarge = scalars[2];
    for (uint32_t li = blockIdx.x; li < nlarge; li += gridDim.x) {
        const uint32_t tile = large_tiles[li];
        const uint2 rg = ranges[tile];
        const int n = (int)(rg.y - rg.x);
        uint64_t* src = bucket + rg.x;
        uint64_t* dst = reinterpret_cast<uint64_t*>(records + rg.x);          // n * 48 bytes, 16-byte aligned
        for (int pass = 0; pass < 8; ++pass) {
            const int shift = pass * 8;
            if (tid < 256) hist[tid] = 0;
            __syncthreads();
            for (int i = tid; i < n; i += LARGE_THREADS) atomicAdd(&hist[(uint32_t)(src[i] >> shift) & 255u], 1u);
            __syncthreads();
            const int trivial = __syncthreads_or(tid < 256 && hist[tid] == (uint32_t)n);
            if (trivial) continue;                                             // uniform: every thread sees the same value
            if (tid < 256) {                                                   // exclusive scan of the 256 bins
                const uint32_t h = hist[tid];
                uint32_t v = h;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, v, o); if (lane >= o) v += t; }
                if (lane == 31) wsum[warp] = v;
                __syncwarp();
                asm volatile("bar.sync 2, 256;" ::: "memory");
                uint32_t carry = 0;
                for (int w = 0; w < warp; ++w) carry += wsum[w];
                base[tid] = carry + v - h;
            }
            __syncthreads();
            for (int c0 = 0; c0 < n; c0 += LARGE_THREADS) {
#pragma unroll
                for (int m = 0; m < 8; ++m) (&wc[0][0])[tid + LARGE_THREADS * m] = 0;
                const int i = c0 + tid;
                const bool valid = i < n;
                const uint64_t key = valid ? src[i] : 0ull;
                const uint32_t d = (uint32_t)(key >> shift) & 255u;
                const unsigned peers = __match_any_sync(FULL, valid ? d : (0x1000u + (uint32_t)lane));
                const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
                __syncthreads();
                if (valid && rank == 0) wc[warp][d] = __popc(peers);
                __syncthreads();
                uint32_t tot = 0;
                if (tid < 256) {
                    for (int w = 0; w < 32; ++w) { const uint32_t t = wc[w][tid]; wc[w][tid] = tot; tot += t; }
                }
                __syncthreads();
                if (valid) dst[base[d] + wc[warp][d] + rank] = key;
                __syncthreads();
                if (tid < 256) base[tid] += tot;
            }
            __syncthreads();
            uint64_t* t = src; src = dst; dst = t;
        }
        if (src != bucket + rg.x) {            // the sorted keys sit in the record slice that packing overwrites: move them
            for (int i = tid; i < n; i += LARGE_THREADS) bucket[rg.x + i] = src[i];
            __syncthreads();
        }
        pack_records(bucket + rg.x, n, tile, rg.x, g, colors, records, keys_out, point_list_out);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------
// mbarrier / bulk-async-copy (TMA 1-D) helpers.
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

constexpr int BATCH = 256;     // records per staged batch (12 KB)

// Conservative test "no pixel of the rectangle [px0,px1]x[py0,py1] can reach alpha >= 1/255".
// power(d) = -0.5 (A dx^2 + C dy^2) - B dx dy <= 0 is maximal at the point of the rectangle
// closest to the mean in the metric of the conic; we bound it from above by evaluating the
// quadratic at the per-axis clamped nearest point along each edge (exact minimiser of a convex
// quadratic over a box lies on the boundary unless the mean is inside).  A safety margin
// absorbs FP32 rounding of both this bound and the per-pixel evaluation, so a culled record
// is one the reference's per-pixel tests (forward.cu:340-348) would reject for every pixel.
__device__ __forceinline__ bool rect_culled(float mx, float my, float A, float B, float C, float cull,
                                            float px0, float py0, float px1, float py1) {
    // offsets d = mean - pixel; pixel in [px0,px1] => dx in [mx-px1, mx-px0]
    const float dx_lo = mx - px1, dx_hi = mx - px0;
    const float dy_lo = my - py1, dy_hi = my - py0;
    const bool in_x = (dx_lo <= 0.f) && (dx_hi >= 0.f);
    const bool in_y = (dy_lo <= 0.f) && (dy_hi >= 0.f);
    if (in_x && in_y) return false;
    // q(dx,dy) = A dx^2 + 2 B dx dy + C dy^2 ; minimise over the box boundary.
    float qmin = 3.0e38f;
    // vertical edges: dx fixed, dy* = clamp(-B dx / C)
    {
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
    }
    // magnitude of the terms entering the per-pixel evaluation, for the rounding margin
    const float ax = fmaxf(fabsf(dx_lo), fabsf(dx_hi)), ay = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
    const float mag = fabsf(A) * ax * ax + fabsf(C) * ay * ay + 2.f * fabsf(B) * ax * ay;
    const float margin = 1e-3f + 1e-3f * fabsf(cull) + 2e-5f * mag;
    // culled iff best possible power = -0.5 qmin is safely below the alpha threshold ln(1/(255 o))
    return (-0.5f * qmin) < (cull - margin);   // false on NaN => keep
}

// K7: front-to-back alpha blend of one 16x16 tile per CTA.  8 warps, each owning an 8x4 pixel block split into two 4x4
// blocks, one per HALF-WARP.  Sorted records arrive in shared memory through double-buffered bulk async copies.  For each
// chunk of 32 records every lane tests ONE record against both 4x4 blocks (conservative rectangle test); each half-warp then
// walks only ITS OWN survivors, front to back: in one step the two halves evaluate two different records (two broadcast
// shared-memory reads), so a warp needs max(|S_A|, |S_B|) steps instead of |S_A u S_B| — small splats touch few 4x4 blocks.
// Per pixel the surviving records and their order are unchanged, so the result is bit-identical to a per-pixel walk.
// A warp whose 32 pixels are saturated stops blending.
// (BLEND_FWD_STREAMS = 2 below is that variant; MEASURED SLOWER than one stream per warp - 356 vs 302 us per DyNeRF view,
//  324 M vs ~280 M warp instructions: the steps saved do not pay for the second ballot / mask bookkeeping and the second
//  shared-memory wavefront of every record read - so the default is the single stream: the warp culls a chunk of 32 records
//  against its 8x4 block with one record per lane and walks the survivors together.  profiles/r02_ab_blend_streams.txt)
#ifndef BLEND_FWD_STREAMS
#define BLEND_FWD_STREAMS 1
#endif
__global__ void __launch_bounds__(256)
k_blend_fwd(const uint2* __restrict__ ranges, const FdgsRecord* __restrict__ records, int W, int H,
            const float* __restrict__ bg_color, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
            float* __restrict__ out_color, float* __restrict__ out_depth)
{
    __shared__ __align__(128) FdgsRecord srec[2][BATCH];
    __shared__ __align__(8) uint64_t bars[2];

    const int tiles_x = (W + FDGS_TILE - 1) / FDGS_TILE;
    const int tile = blockIdx.y * tiles_x + blockIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // warp (wx, wy) in a 2 x 4 arrangement of 8x4 pixel blocks
    const int wx = wid & 1, wy = wid >> 1;
    const int bx0 = blockIdx.x * FDGS_TILE + wx * 8, by0 = blockIdx.y * FDGS_TILE + wy * 4;
#if BLEND_FWD_STREAMS == 2
    const int half = lane >> 4;                                  // half-warp h owns the 4x4 block at x offset 4h
    const int px = bx0 + half * 4 + (lane & 3), py = by0 + ((lane >> 2) & 3);
#else
    const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
#endif
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    // pixel block bounds for the cull test (clipped to the image)
    const float cx0 = (float)bx0, cy0 = (float)by0;
    const float cx1 = (float)min(bx0 + 7, W - 1), cy1 = (float)min(by0 + 3, H - 1);
    const float cxm0 = (float)min(bx0 + 3, W - 1), cxm1 = (float)(bx0 + 4);       // right edge of block A, left edge of block B
    const bool blockB = bx0 + 4 < W;
    const bool warp_has_pixels = (bx0 < W) && (by0 < H);

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int nb = (total + BATCH - 1) / BATCH;

    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto issue = [&](int b) {
        const int cnt = min(BATCH, total - b * BATCH);
        const uint32_t bytes = (uint32_t)cnt * (uint32_t)sizeof(FdgsRecord);
        mbar_expect_tx(&bars[b & 1], bytes);
        bulk_g2s(&srec[b & 1][0], records + range.x + (size_t)b * BATCH, bytes, &bars[b & 1]);
    };
    if (threadIdx.x == 0 && nb > 0) issue(0);

    float T = 1.0f;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, Dp = 0.f;
    uint32_t last_contributor = 0;
    bool done = !inside;

    // one pixel x record evaluation (forward.cu:335-366)
    auto eval = [&](const FdgsRecord* rec, int jj, int b) {
        const float4 q0 = reinterpret_cast<const float4*>(&rec[jj])[0];
        const float4 q1 = reinterpret_cast<const float4*>(&rec[jj])[1];
        const float dx = q0.x - pxf, dy = q0.y - pyf;
        const float power = -0.5f * (q0.z * dx * dx + q1.x * dy * dy) - q0.w * dx * dy;
        if (!(power > 0.0f)) {
            const float alpha = fminf(0.99f, q1.y * expf(power));
            if (!(alpha < 1.0f / 255.0f)) {
                const float test_T = T * (1 - alpha);
                if (test_T < 0.0001f) {
                    done = true;
                } else {
                    const float4 q2 = reinterpret_cast<const float4*>(&rec[jj])[2];
                    C0 += q2.x * alpha * T;
                    C1 += q2.y * alpha * T;
                    C2 += q2.z * alpha * T;
                    Dp += q1.z * alpha * T;
                    T = test_T;
                    last_contributor = (uint32_t)(b * BATCH + jj + 1);
                }
            }
        }
    };

    for (int b = 0; b < nb; ++b) {
        if (threadIdx.x == 0 && b + 1 < nb) issue(b + 1);
        mbar_wait(&bars[b & 1], (uint32_t)((b >> 1) & 1));
        const FdgsRecord* rec = srec[b & 1];
        const int cnt = min(BATCH, total - b * BATCH);

        if (warp_has_pixels && !__all_sync(FULL, done)) {
            for (int c0 = 0; c0 < cnt; c0 += 32) {
                const int j = c0 + lane;
#if BLEND_FWD_STREAMS == 2
                bool keepA = false, keepB = false;
                if (j < cnt) {
                    const float4 q0 = reinterpret_cast<const float4*>(&rec[j])[0];
                    const float4 q1 = reinterpret_cast<const float4*>(&rec[j])[1];
                    const float cull = rec[j].cull;
                    if (!(q1.y <= 0.f)) {
                        keepA = !rect_culled(q0.x, q0.y, q0.z, q0.w, q1.x, cull, cx0, cy0, cxm0, cy1);
                        keepB = blockB && !rect_culled(q0.x, q0.y, q0.z, q0.w, q1.x, cull, cxm1, cy0, cx1, cy1);
                    }
                }
                unsigned mA = __ballot_sync(FULL, keepA), mB = __ballot_sync(FULL, keepB);
                unsigned mine = half ? mB : mA;
                while (mA | mB) {
                    if (mine && !done) eval(rec, c0 + (__ffs(mine) - 1), b);
                    mine &= mine - 1;                 // 0 stays 0
                    mA &= mA - 1; mB &= mB - 1;
                    if (__all_sync(FULL, done)) { mA = 0; mB = 0; c0 = cnt; }
                }
#else
                bool keep = false;
                if (j < cnt) {
                    const float4 q0 = reinterpret_cast<const float4*>(&rec[j])[0];
                    const float4 q1 = reinterpret_cast<const float4*>(&rec[j])[1];
                    const float cull = rec[j].cull;
                    keep = !(q1.y <= 0.f) && !rect_culled(q0.x, q0.y, q0.z, q0.w, q1.x, cull, cx0, cy0, cx1, cy1);
                }
                unsigned m = __ballot_sync(FULL, keep);
                while (m) {
                    const int jj = c0 + (__ffs(m) - 1);
                    m &= m - 1;
                    if (!done) eval(rec, jj, b);
                    if (__all_sync(FULL, done)) { m = 0; c0 = cnt; }
                }
#endif
            }
        }
        const int ndone = __syncthreads_count(done);
        if (ndone == 256) {
            if (b + 1 < nb) mbar_wait(&bars[(b + 1) & 1], (uint32_t)(((b + 1) >> 1) & 1));   // drain in-flight copy
            break;
        }
    }

    if (inside) {
        const size_t pix_id = (size_t)py * W + px;
        const size_t HW = (size_t)H * W;
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        out_color[pix_id] = C0 + T * bg_color[0];
        out_color[HW + pix_id] = C1 + T * bg_color[1];
        out_color[2 * HW + pix_id] = C2 + T * bg_color[2];
        out_depth[pix_id] = Dp;
    }
}

// Diagnostic (bench / roofline, never on the training path): the blend's algorithmic work of one rendered view.
// counts[0] += sum over pixels of n_contrib  = pixel x instance pairs a per-pixel walk has to TEST (forward.cu:321-366)
// counts[1] += pairs that CONTRIBUTE (alpha >= 1/255 before the pixel saturates) = the pairs the backward differentiates
__global__ void __launch_bounds__(256)
k_count_pairs(const uint2* __restrict__ ranges, const FdgsRecord* __restrict__ records, int W, int H,
              const uint32_t* __restrict__ n_contrib, unsigned long long* __restrict__ counts)
{
    const int tiles_x = (W + FDGS_TILE - 1) / FDGS_TILE;
    const int tile = blockIdx.y * tiles_x + blockIdx.x;
    const int px = blockIdx.x * FDGS_TILE + (threadIdx.x & 15), py = blockIdx.y * FDGS_TILE + (threadIdx.x >> 4);
    unsigned long long walked = 0, hits = 0;
    if (px < W && py < H) {
        const uint2 rg = ranges[tile];
        const uint32_t n = n_contrib[(size_t)py * W + px];
        walked = n;
        for (uint32_t i = 0; i < n && rg.x + i < rg.y; ++i) {
            const FdgsRecord r = records[rg.x + i];
            const float dx = r.x - (float)px, dy = r.y - (float)py;
            const float power = -0.5f * (r.ca * dx * dx + r.cc * dy * dy) - r.cb * dx * dy;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, r.opacity * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            ++hits;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        walked += __shfl_xor_sync(FULL, walked, o);
        hits += __shfl_xor_sync(FULL, hits, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&counts[0], walked); atomicAdd(&counts[1], hits); }
}

// checkFrustum (rasterizer_impl.cu:54-66)
__global__ void k_mark_visible(int P, const float* __restrict__ means3D, const float* __restrict__ V, uint8_t* present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float3 p = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
    const float3 pv = xform43(p, V);
    present[idx] = pv.z > 0.2f ? 1 : 0;
}

int check_forward_args(const FdgsRasterForward* a) {
    if (!a) return FDGS_E_BADARG;
    if (a->P < 0 || a->W <= 0 || a->H <= 0) return FDGS_E_BADARG;
    if (a->P == 0) return FDGS_OK;
    if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->campos || !a->background) return FDGS_E_BADARG;
    if (!a->out_color || !a->out_depth || !a->radii || !a->geom_ws || !a->img_ws || !a->counts_host) return FDGS_E_BADARG;
    if ((a->shs == nullptr) == (a->colors_precomp == nullptr)) return FDGS_E_BADARG;
    const bool has_sr = a->scales != nullptr && a->rotations != nullptr;
    if (has_sr == (a->cov3D_precomp != nullptr)) return FDGS_E_BADARG;
    if (a->shs && (a->M <= 0 || a->D < 0 || a->D > 3 || (a->D + 1) * (a->D + 1) > a->M)) return FDGS_E_BADARG;
    if (a->geom_bytes < FdgsGeom::bytes(a->P) || a->img_bytes < FdgsImg::bytes(a->W, a->H)) return FDGS_E_WORKSPACE;
    if (a->activate && a->cov3D_precomp) return FDGS_E_BADARG;      // raw inputs only make sense with scales / rotations
    return FDGS_OK;
}

}  // namespace

extern "C" {

size_t fdgs_raster_geom_bytes(int P) { return FdgsGeom::bytes(P); }
size_t fdgs_raster_img_bytes(int W, int H) { return FdgsImg::bytes(W, H); }
size_t fdgs_raster_bin_bytes(int64_t R) { return FdgsBin::bytes(R); }

// K1 + K2 (preprocess with the tile histogram, tile scan).  cap: capacity (instances) of the workspace the render stage will
// use, or 0xffffffff when the caller sizes it from the exact count afterwards.
static int launch_plan(const FdgsRasterForward* a, uint32_t cap, cudaStream_t stream) {
    const int gx = (a->W + FDGS_TILE - 1) / FDGS_TILE, gy = (a->H + FDGS_TILE - 1) / FDGS_TILE;
    const int T = gx * gy;
    FdgsGeom g = FdgsGeom::carve(a->geom_ws, a->P);
    FdgsImg im = FdgsImg::carve(a->img_ws, a->W, a->H);
    const float focal_y = a->H / (2.0f * a->tanfovy);
    const float focal_x = a->W / (2.0f * a->tanfovx);
    // tile_count, tile_cursor, tile_start, scalars are contiguous: one memset
    FDGS_CUDA(cudaMemsetAsync(im.tile_count, 0, (size_t)((char*)im.scalars - (char*)im.tile_count) + 32, stream));
    const int sh_staged = (a->shs && !a->colors_precomp && a->M == 16 && ((uintptr_t)a->shs & 15) == 0) ? 1 : 0;
    const size_t sh_smem = sh_staged ? (size_t)256 * (3 * a->M + 1) * sizeof(float) : 0;
    if (sh_staged) FDGS_CUDA(cudaFuncSetAttribute(k_preprocess, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh_smem));
    { FDGS_PROF("k_preprocess", stream);
    k_preprocess<<<(a->P + 255) / 256, 256, sh_smem, stream>>>(
        a->P, a->D, a->M, a->means3D, a->scales, a->scale_modifier, a->rotations, a->opacities, a->shs,
        a->cov3D_precomp, a->colors_precomp, a->viewmatrix, a->projmatrix, a->campos, a->W, a->H,
        a->tanfovx, a->tanfovy, focal_x, focal_y, gx, gy, a->prefiltered, a->radii, g, im.tile_count, sh_staged,
        (a->activate && !a->cov3D_precomp) ? 1 : 0);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);
    { FDGS_PROF("k_tile_scan", stream);
    k_tile_scan<<<1, 1024, 0, stream>>>(T, im.tile_count, im.tile_start, im.ranges, im.scalars, im.large_tiles, cap);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);
    // counts_host: [0] = R, [1] = largest tile bucket, [2] = tiles above the shared-memory sort cap, [3] = overflow flag
    FDGS_CUDA(cudaMemcpyAsync(a->counts_host, im.scalars, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
    return FDGS_OK;
}

// K3 - K7 (bucket scatter, segmented sort + record packing, blend).  cap: instances bin_ws was carved for.
// large: 1 = some bucket is above the shared-memory cap, 0 = none, -1 = unknown on the host (launch the large-bucket kernel
// anyway: it reads the device-side list and exits when that is empty).
static int launch_render(const FdgsRasterForward* a, int64_t cap, int large, cudaStream_t stream) {
    const int gx = (a->W + FDGS_TILE - 1) / FDGS_TILE, gy = (a->H + FDGS_TILE - 1) / FDGS_TILE;
    const int T = gx * gy;
    FdgsGeom g = FdgsGeom::carve(a->geom_ws, a->P);
    FdgsImg im = FdgsImg::carve(a->img_ws, a->W, a->H);
    FdgsBin bn = FdgsBin::carve(a->bin_ws, cap);
    if (cap > 0) {
        { FDGS_PROF("k_scatter", stream);
        k_scatter<<<(a->P + 255) / 256, 256, 0, stream>>>(a->P, a->radii, g, gx, gy, im.tile_start, im.tile_cursor, bn.bucket, im.scalars);
        }
        FDGS_LAUNCH_CHECK(a->debug, stream);
        const float* colors = a->colors_precomp ? a->colors_precomp : g.rgb;
        { FDGS_PROF("k_tile_sort", stream);
        k_tile_sort<<<T, 256, 0, stream>>>(im.ranges, bn.bucket, g, colors, bn.records, a->keys_out, a->point_list_out);
        }
        FDGS_LAUNCH_CHECK(a->debug, stream);
        if (large != 0) {
            FDGS_PROF("k_tile_sort_large", stream);
            k_tile_sort_large<<<148, LARGE_THREADS, 0, stream>>>(im.ranges, im.large_tiles, im.scalars, bn.bucket, g, colors,
                                                                 bn.records, a->keys_out, a->point_list_out);
        }
        FDGS_LAUNCH_CHECK(a->debug, stream);
    }
    dim3 grid(gx, gy);
    { FDGS_PROF("k_blend_fwd", stream);
    k_blend_fwd<<<grid, 256, 0, stream>>>(im.ranges, bn.records, a->W, a->H, a->background, im.final_T, im.n_contrib,
                                          a->out_color, a->out_depth);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);
    return FDGS_OK;
}

static int zero_outputs(const FdgsRasterForward* a, cudaStream_t stream) {     // P == 0: rasterize_points.cu:68-82
    const size_t HW = (size_t)a->W * a->H;
    FDGS_CUDA(cudaMemsetAsync(a->out_color, 0, 3 * HW * sizeof(float), stream));
    FDGS_CUDA(cudaMemsetAsync(a->out_depth, 0, HW * sizeof(float), stream));
    return FDGS_OK;
}

int fdgs_raster_forward_plan(const FdgsRasterForward* a, void* stream_) {
    int rc = check_forward_args(a);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    for (int i = 0; i < 4; ++i) a->counts_host[i] = 0;
    if (a->P == 0) return FDGS_OK;
    rc = launch_plan(a, 0xffffffffu, stream);
    if (rc) return rc;
    FDGS_CUDA(cudaStreamSynchronize(stream));
    return FDGS_OK;
}

int fdgs_raster_forward_render(const FdgsRasterForward* a, int64_t R, void* stream_) {
    int rc = check_forward_args(a);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (a->P == 0) return zero_outputs(a, stream);
    if (R < 0) return FDGS_E_BADARG;
    if (R > 0 && (!a->bin_ws || a->bin_bytes < FdgsBin::bytes(R))) return FDGS_E_WORKSPACE;
    return launch_render(a, R, a->counts_host[1] > FDGS_SORT_SMEM_CAP ? 1 : 0, stream);
}

int fdgs_raster_forward(const FdgsRasterForward* a, int64_t capacity, void* stream_) {
    int rc = check_forward_args(a);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (a->P == 0) { for (int i = 0; i < 4; ++i) a->counts_host[i] = 0; return zero_outputs(a, stream); }
    if (capacity <= 0 || capacity > 0xfffffff0LL) return FDGS_E_BADARG;
    if (!a->bin_ws || a->bin_bytes < FdgsBin::bytes(capacity)) return FDGS_E_WORKSPACE;
    rc = launch_plan(a, (uint32_t)capacity, stream);       // queues the asynchronous read-back of the counts; no host sync
    if (rc) return rc;
    return launch_render(a, capacity, -1, stream);
}

int fdgs_raster_copy_ranges(const void* img_ws, int W, int H, uint32_t* ranges_out, void* stream_) {
    if (!img_ws || !ranges_out || W <= 0 || H <= 0) return FDGS_E_BADARG;
    FdgsImg im = FdgsImg::carve(const_cast<void*>(img_ws), W, H);
    const size_t T = (size_t)((W + FDGS_TILE - 1) / FDGS_TILE) * ((H + FDGS_TILE - 1) / FDGS_TILE);
    FDGS_CUDA(cudaMemcpyAsync(ranges_out, im.ranges, T * sizeof(uint2), cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
    return FDGS_OK;
}

int fdgs_raster_count_pairs(const void* img_ws, const void* bin_ws, int64_t R, int W, int H, unsigned long long* counts,
                            void* stream_) {
    if (!img_ws || !counts || W <= 0 || H <= 0 || R < 0) return FDGS_E_BADARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    FDGS_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(unsigned long long), stream));
    if (R == 0) return FDGS_OK;
    if (!bin_ws) return FDGS_E_BADARG;
    FdgsImg im = FdgsImg::carve(const_cast<void*>(img_ws), W, H);
    FdgsBin bn = FdgsBin::carve(const_cast<void*>(bin_ws), R);
    dim3 grid((W + FDGS_TILE - 1) / FDGS_TILE, (H + FDGS_TILE - 1) / FDGS_TILE);
    k_count_pairs<<<grid, 256, 0, stream>>>(im.ranges, bn.records, W, H, im.n_contrib, counts);
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

int fdgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                      uint8_t* present, void* stream_) {
    (void)projmatrix;
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return FDGS_E_BADARG;
    if (P == 0) return FDGS_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    { FDGS_PROF("k_mark_visible", stream);
    k_mark_visible<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream_>>>(P, means3D, viewmatrix, present);
    }
    FDGS_LAUNCH_CHECK(0, (cudaStream_t)stream_);
    return FDGS_OK;
}

}  // extern "C"
