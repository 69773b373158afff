// Shared device/host helpers for the fdgs sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/fdgs.h"

#define FDGS_TILE 16                 // tile edge in pixels (reference: BLOCK_X/BLOCK_Y, RAST/cuda_rasterizer/config.h)
#define FDGS_TILE_PIX (FDGS_TILE * FDGS_TILE)
#define FDGS_ALIGN 256
#define FDGS_SORT_SMEM_CAP 2048      // tile buckets up to this many instances are sorted in shared memory

static inline size_t fdgs_align_up(size_t x) { return (x + (FDGS_ALIGN - 1)) & ~(size_t)(FDGS_ALIGN - 1); }

// One packed, sorted instance record (48 B = 3 x float4).  Written once by the
// per-tile sort, then streamed by both blend kernels with bulk async copies.
struct __align__(16) FdgsRecord {
    float x, y, ca, cb;      // pixel-space mean, conic A, B
    float cc, opacity, depth; uint32_t id;   // conic C, opacity, view depth, Gaussian id
    float r, g, b, cull;     // colour, cull = -2*ln(255*opacity) style threshold (see blend kernels)
};
static_assert(sizeof(FdgsRecord) == 48, "record must be 48 bytes");

// ---- typed views over the caller-owned workspaces ---------------------------
struct FdgsGeom {          // per-Gaussian, lives from forward to backward
    float* depths;         // [P]
    float2* means2D;       // [P]
    float* cov3D;          // [6P]
    float4* conic_opacity; // [P]
    float* rgb;            // [3P]
    uint8_t* clamped;      // [P] bit c set = channel c clamped at 0
    uint32_t* tiles_touched; // [P]
    static size_t bytes(int P) {
        size_t p = (size_t)(P > 0 ? P : 1);
        return fdgs_align_up(p * 4) + fdgs_align_up(p * 8) + fdgs_align_up(p * 24) + fdgs_align_up(p * 16) +
               fdgs_align_up(p * 12) + fdgs_align_up(p) + fdgs_align_up(p * 4);
    }
    static FdgsGeom carve(void* base, int P) {
        size_t p = (size_t)(P > 0 ? P : 1);
        char* c = (char*)base;
        FdgsGeom g;
        g.depths = (float*)c;            c += fdgs_align_up(p * 4);
        g.means2D = (float2*)c;          c += fdgs_align_up(p * 8);
        g.cov3D = (float*)c;             c += fdgs_align_up(p * 24);
        g.conic_opacity = (float4*)c;    c += fdgs_align_up(p * 16);
        g.rgb = (float*)c;               c += fdgs_align_up(p * 12);
        g.clamped = (uint8_t*)c;         c += fdgs_align_up(p);
        g.tiles_touched = (uint32_t*)c;  c += fdgs_align_up(p * 4);
        return g;
    }
};

struct FdgsImg {           // per-pixel / per-tile
    float* final_T;        // [HW]
    uint32_t* n_contrib;   // [HW]
    uint2* ranges;         // [T]  (start,end) into the sorted list; (0,0) for empty tiles
    uint32_t* tile_count;  // [T]
    uint32_t* tile_cursor; // [T]
    uint32_t* tile_start;  // [T]
    uint32_t* scalars;     // [8]: 0 = R, 1 = max tile count, 2 = number of tiles with more than FDGS_SORT_SMEM_CAP instances
    uint32_t* large_tiles; // [T] ids of those tiles (arbitrary order)
    static size_t bytes(int W, int H) {
        size_t hw = (size_t)W * H;
        size_t T = (size_t)((W + FDGS_TILE - 1) / FDGS_TILE) * ((H + FDGS_TILE - 1) / FDGS_TILE);
        if (hw == 0) hw = 1;
        if (T == 0) T = 1;
        return 2 * fdgs_align_up(hw * 4) + fdgs_align_up(T * 8) + 4 * fdgs_align_up(T * 4) + fdgs_align_up(32);
    }
    static FdgsImg carve(void* base, int W, int H) {
        size_t hw = (size_t)W * H;
        size_t T = (size_t)((W + FDGS_TILE - 1) / FDGS_TILE) * ((H + FDGS_TILE - 1) / FDGS_TILE);
        if (hw == 0) hw = 1;
        if (T == 0) T = 1;
        char* c = (char*)base;
        FdgsImg s;
        s.final_T = (float*)c;          c += fdgs_align_up(hw * 4);
        s.n_contrib = (uint32_t*)c;     c += fdgs_align_up(hw * 4);
        s.ranges = (uint2*)c;           c += fdgs_align_up(T * 8);
        s.tile_count = (uint32_t*)c;    c += fdgs_align_up(T * 4);
        s.tile_cursor = (uint32_t*)c;   c += fdgs_align_up(T * 4);
        s.tile_start = (uint32_t*)c;    c += fdgs_align_up(T * 4);
        s.scalars = (uint32_t*)c;       c += fdgs_align_up(32);
        s.large_tiles = (uint32_t*)c;   c += fdgs_align_up(T * 4);
        return s;
    }
};

struct FdgsBin {           // per-instance
    uint64_t* bucket;      // [R] (depth_bits << 32 | id), tile-bucketed, unsorted inside a tile
    FdgsRecord* records;   // [R] sorted, packed
    static size_t bytes(int64_t R) {
        size_t r = (size_t)(R > 0 ? R : 1);
        return fdgs_align_up(r * 8) + fdgs_align_up(r * sizeof(FdgsRecord));
    }
    static FdgsBin carve(void* base, int64_t R) {
        size_t r = (size_t)(R > 0 ? R : 1);
        char* c = (char*)base;
        FdgsBin b;
        b.bucket = (uint64_t*)c;        c += fdgs_align_up(r * 8);
        b.records = (FdgsRecord*)c;     c += fdgs_align_up(r * sizeof(FdgsRecord));
        return b;
    }
};

// ---- error plumbing ---------------------------------------------------------
#define FDGS_LAUNCH_CHECK(debug, stream)                                   \
    do {                                                                   \
        cudaError_t e__ = cudaGetLastError();                              \
        if (e__ == cudaSuccess && (debug)) e__ = cudaStreamSynchronize(stream); \
        if (e__ != cudaSuccess) return (int)e__;                           \
    } while (0)

#define FDGS_CUDA(call)                                                    \
    do {                                                                   \
        cudaError_t e__ = (call);                                          \
        if (e__ != cudaSuccess) return (int)e__;                           \
    } while (0)

// ---- small device helpers ---------------------------------------------------
#ifdef __CUDACC__
// Pixel-centre mapping.  The reference evaluates this in double because of its
// 1.0 / 0.5 literals (RAST/cuda_rasterizer/auxiliary.h:41-44); the rounding to
// float happens once at the end, which decides tile rects => keep it in double.
__device__ __forceinline__ float fdgs_ndc2pix(float v, int S) {
    return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5);
}

// Tile rectangle of a splat (reference: getRect, auxiliary.h:46-56).
__device__ __forceinline__ void fdgs_tile_rect(float px, float py, int radius, int gx, int gy,
                                               int& x0, int& y0, int& x1, int& y1) {
    const float r = (float)radius;
    x0 = min(gx, max(0, (int)((px - r) / (float)FDGS_TILE)));
    y0 = min(gy, max(0, (int)((py - r) / (float)FDGS_TILE)));
    // left-to-right float sums: ((p + r) + 16) - 1, not p + r + 15
    x1 = min(gx, max(0, (int)((((px + r) + (float)FDGS_TILE) - 1.0f) / (float)FDGS_TILE)));
    y1 = min(gy, max(0, (int)((((py + r) + (float)FDGS_TILE) - 1.0f) / (float)FDGS_TILE)));
}
#endif
