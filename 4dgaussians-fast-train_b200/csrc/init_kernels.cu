// Init-time native pieces of the reference, as CUDA hash-grid kernels behind the C ABI:
//   fdgs_knn_mean_dist2   simple_knn's distCUDA2 (submodules/simple-knn/simple_knn.cu:65-215, scene/gaussian_model.py:145:
//                         initial scales = sqrt of the mean squared distance to the 3 nearest neighbours)
//   fdgs_voxel_downsample Instant4D grid pruning (utils/grid_pruning.py:16-41 = open3d's PointCloud::VoxelDownSample:
//                         one averaged point / colour per occupied voxel)
// Both run once per training run: they are written for clarity and exactness, not tuned.
#include "fdgs_common.cuh"
#include "prof.cuh"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <mutex>

namespace {

constexpr unsigned FULL = 0xffffffffu;

// ---- bounding box -----------------------------------------------------------------------------------------------------
// floats mapped to integers whose order is the floats' order, so that atomicMin / atomicMax apply
__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void k_bbox_init(int* bb) { if (threadIdx.x < 3) { bb[threadIdx.x] = 0x7fffffff; bb[3 + threadIdx.x] = (int)0x80000000; } }

__global__ void __launch_bounds__(256)
k_bbox(int P, const float* __restrict__ pts, int* __restrict__ bb)
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x)
#pragma unroll
        for (int a = 0; a < 3; ++a) { const float v = pts[3 * (size_t)i + a]; mn[a] = fminf(mn[a], v); mx[a] = fmaxf(mx[a], v); }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(FULL, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(FULL, mx[a], o));
        }
        if ((threadIdx.x & 31) == 0) { atomicMin(&bb[a], f2ord(mn[a])); atomicMax(&bb[3 + a], f2ord(mx[a])); }
    }
}

// ---- uniform grid ------------------------------------------------------------------------------------------------------
struct Grid { int n[3]; float mn[3], inv[3], cell[3]; };

__device__ __forceinline__ Grid make_grid(const int* bb, int n) {
    Grid g;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = ord2f(bb[a]), hi = ord2f(bb[3 + a]);
        const float ext = fmaxf(hi - lo, 1e-20f);
        g.n[a] = n; g.mn[a] = lo; g.cell[a] = ext / (float)n; g.inv[a] = (float)n / ext;
    }
    return g;
}
__device__ __forceinline__ int cell_coord(const Grid& g, int a, float v) {
    const int c = (int)((v - g.mn[a]) * g.inv[a]);
    return min(g.n[a] - 1, max(0, c));
}

__global__ void __launch_bounds__(256)
k_knn_count(int P, const float* __restrict__ pts, const int* __restrict__ bb, int n, uint32_t* __restrict__ count, uint32_t* __restrict__ cell_of)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const Grid g = make_grid(bb, n);
    const int cx = cell_coord(g, 0, pts[3 * (size_t)i]), cy = cell_coord(g, 1, pts[3 * (size_t)i + 1]), cz = cell_coord(g, 2, pts[3 * (size_t)i + 2]);
    const uint32_t c = ((uint32_t)cz * n + cy) * n + cx;
    cell_of[i] = c;
    atomicAdd(&count[c], 1u);
}

// exclusive scan of `count` [m] into `start` [m + 1] by one CTA
__global__ void __launch_bounds__(1024)
k_scan_u32(int m, const uint32_t* __restrict__ count, uint32_t* __restrict__ start)
{
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry_s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < m; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t c = i < m ? count[i] : 0u;
        uint32_t v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, v, o); if (lane >= o) v += t; }
        if (lane == 31) wsum[wid] = v;
        __syncthreads();
        if (wid == 0) {
            uint32_t s = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, s, o); if (lane >= o) s += t; }
            wsum[lane] = s;
        }
        __syncthreads();
        const uint32_t incl = carry_s + v + (wid > 0 ? wsum[wid - 1] : 0u);
        if (i < m) start[i] = incl - c;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) start[m] = carry_s;
}

__global__ void __launch_bounds__(256)
k_knn_scatter(int P, const uint32_t* __restrict__ cell_of, const uint32_t* __restrict__ start, uint32_t* __restrict__ cursor,
              uint32_t* __restrict__ sorted)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t c = cell_of[i];
    sorted[start[c] + atomicAdd(&cursor[c], 1u)] = (uint32_t)i;
}

// the reference's k-best update (simple_knn.cu:134-148), K = 3, same squared-distance expression
__device__ __forceinline__ void update3(const float3 ref, const float3 p, float (&best)[3]) {
    const float dx = p.x - ref.x, dy = p.y - ref.y, dz = p.z - ref.z;
    float dist = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
#pragma unroll
    for (int j = 0; j < 3; ++j) if (best[j] > dist) { const float t = best[j]; best[j] = dist; dist = t; }
}

// One thread per point: cells are visited in growing Chebyshev shells around the point's cell until the third best
// distance cannot be beaten from outside the cube searched so far (exact 3-NN, like the reference's box pruning).
__global__ void __launch_bounds__(128)
k_knn_search(int P, const float* __restrict__ pts, const int* __restrict__ bb, int n, const uint32_t* __restrict__ start,
             const uint32_t* __restrict__ sorted, float* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const Grid g = make_grid(bb, n);
    const float3 me = make_float3(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2]);
    const int c0[3] = {cell_coord(g, 0, me.x), cell_coord(g, 1, me.y), cell_coord(g, 2, me.z)};
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    const float cell_min = fminf(g.cell[0], fminf(g.cell[1], g.cell[2]));
    for (int r = 0; r < n; ++r) {
        const int z0 = max(0, c0[2] - r), z1 = min(n - 1, c0[2] + r), y0 = max(0, c0[1] - r), y1 = min(n - 1, c0[1] + r);
        const int x0 = max(0, c0[0] - r), x1 = min(n - 1, c0[0] + r);
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                const bool face = (abs(z - c0[2]) == r) || (abs(y - c0[1]) == r);
                for (int x = x0; x <= x1; ++x) {
                    if (!face && abs(x - c0[0]) != r) continue;          // interior of the cube: searched by the earlier shells
                    const uint32_t c = ((uint32_t)z * n + y) * n + x;
                    for (uint32_t k = start[c]; k < start[c + 1]; ++k) {
                        const uint32_t j = sorted[k];
                        if ((int)j == i) continue;
                        update3(me, make_float3(pts[3 * (size_t)j], pts[3 * (size_t)j + 1], pts[3 * (size_t)j + 2]), best);
                    }
                }
            }
        // every point outside the (2r+1)^3 cube of cells is at least r cells away along some axis (with a safety factor for
        // the rounding of the cell coordinates)
        const float reach = (float)r * cell_min * 0.999f;
        if (best[2] <= reach * reach) break;
        if (x0 == 0 && y0 == 0 && z0 == 0 && x1 == n - 1 && y1 == n - 1 && z1 == n - 1) break;     // whole grid searched
    }
    out[i] = (best[0] + best[1] + best[2]) / 3.0f;
}

// ---- voxel down-sampling ------------------------------------------------------------------------------------------------
// open3d PointCloud::VoxelDownSample: voxel_min_bound = min_bound - voxel_size / 2; index = floor((p - voxel_min_bound) /
// voxel_size) per axis (in double); output = mean position and mean colour of the points of every occupied voxel.
__device__ __forceinline__ uint64_t voxel_key(const double* vmin, double inv, const float* p) {
    uint64_t k = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const long long v = (long long)floor(((double)p[a] - vmin[a]) * inv);
        k |= ((uint64_t)(v & 0x1fffff)) << (21 * a);
    }
    return k;
}
__device__ __forceinline__ uint32_t hash64(uint64_t k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
    return (uint32_t)k;
}
constexpr uint64_t EMPTY_KEY = ~0ull;

struct VoxelTable { unsigned long long* keys; double* sums; uint32_t* counts; uint32_t cap; };   // sums [cap][6]

__global__ void __launch_bounds__(256)
k_voxel_accumulate(int P, const float* __restrict__ pts, const float* __restrict__ cols, const int* __restrict__ bb, double voxel,
                   VoxelTable t)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const double vmin[3] = {(double)ord2f(bb[0]) - voxel * 0.5, (double)ord2f(bb[1]) - voxel * 0.5, (double)ord2f(bb[2]) - voxel * 0.5};
    const uint64_t key = voxel_key(vmin, 1.0 / voxel, pts + 3 * (size_t)i);
    uint32_t slot = hash64(key) % t.cap;
    for (;;) {
        const unsigned long long prev = atomicCAS(&t.keys[slot], (unsigned long long)EMPTY_KEY, (unsigned long long)key);
        if (prev == EMPTY_KEY || prev == key) break;
        slot = slot + 1 == t.cap ? 0 : slot + 1;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        atomicAdd(&t.sums[6 * (size_t)slot + a], (double)pts[3 * (size_t)i + a]);
        if (cols) atomicAdd(&t.sums[6 * (size_t)slot + 3 + a], (double)cols[3 * (size_t)i + a]);
    }
    atomicAdd(&t.counts[slot], 1u);
}

__global__ void __launch_bounds__(256)
k_voxel_emit(VoxelTable t, uint32_t* __restrict__ n_out, float* __restrict__ out_pts, float* __restrict__ out_cols, int max_out)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= t.cap || t.keys[s] == EMPTY_KEY) return;
    const uint32_t o = atomicAdd(n_out, 1u);
    if ((int)o >= max_out) return;
    const double inv = 1.0 / (double)t.counts[s];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        out_pts[3 * (size_t)o + a] = (float)(t.sums[6 * (size_t)s + a] * inv);
        if (out_cols) out_cols[3 * (size_t)o + a] = (float)(t.sums[6 * (size_t)s + 3 + a] * inv);
    }
}

static int knn_cells_per_axis(int P) {
    int n = (int)cbrt((double)P / 2.0);
    return n < 1 ? 1 : (n > 128 ? 128 : n);
}
static uint32_t voxel_table_cap(int P) { return (uint32_t)(2 * (size_t)(P > 0 ? P : 1) + 64); }

}  // namespace

extern "C" {

size_t fdgs_knn_ws_bytes(int P) {
    const size_t n = (size_t)knn_cells_per_axis(P);
    const size_t cells = n * n * n;
    return fdgs_align_up(32) + 2 * fdgs_align_up((cells + 1) * 4) + fdgs_align_up(cells * 4) + 2 * fdgs_align_up((size_t)(P > 0 ? P : 1) * 4);
}

int fdgs_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* ws, size_t ws_bytes, void* stream_) {
    if (P < 0 || (P > 0 && (!points || !mean_dist2 || !ws))) return FDGS_E_BADARG;
    if (P == 0) return FDGS_OK;
    if (ws_bytes < fdgs_knn_ws_bytes(P)) return FDGS_E_WORKSPACE;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int n = knn_cells_per_axis(P);
    const size_t cells = (size_t)n * n * n;
    char* c = (char*)ws;
    int* bb = (int*)c;                 c += fdgs_align_up(32);
    uint32_t* count = (uint32_t*)c;    c += fdgs_align_up((cells + 1) * 4);
    uint32_t* start = (uint32_t*)c;    c += fdgs_align_up((cells + 1) * 4);
    uint32_t* cursor = (uint32_t*)c;   c += fdgs_align_up(cells * 4);
    uint32_t* cell_of = (uint32_t*)c;  c += fdgs_align_up((size_t)P * 4);
    uint32_t* sorted = (uint32_t*)c;
    FDGS_CUDA(cudaMemsetAsync(count, 0, (cells + 1) * 4, stream));
    FDGS_CUDA(cudaMemsetAsync(cursor, 0, cells * 4, stream));
    FDGS_PROF("k_knn", stream);
    k_bbox_init<<<1, 32, 0, stream>>>(bb);
    k_bbox<<<std::min((P + 255) / 256, 1184), 256, 0, stream>>>(P, points, bb);
    k_knn_count<<<(P + 255) / 256, 256, 0, stream>>>(P, points, bb, n, count, cell_of);
    k_scan_u32<<<1, 1024, 0, stream>>>((int)cells, count, start);
    k_knn_scatter<<<(P + 255) / 256, 256, 0, stream>>>(P, cell_of, start, cursor, sorted);
    k_knn_search<<<(P + 127) / 128, 128, 0, stream>>>(P, points, bb, n, start, sorted, mean_dist2);
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

// ---- the same 3-NN pass with NO caller workspace and NO allocation (include/fdgs.h: fdgs_knn_mean_dist2_static) -----------
// Scratch = one file-scope device array (below), FDGS_KNN_STATIC_MAX_POINTS points at most.  In host code `g_knn_static` is only
// a stand-in for the symbol: its device address is taken once per device with cudaGetSymbolAddress, with that device current.
}  // extern "C"  (the array and its set-up are file-local)

namespace {

constexpr int KNN_STATIC_MAX_N = 51;                                         // knn_cells_per_axis(262144) = 50
constexpr size_t KNN_STATIC_MAX_CELLS = (size_t)KNN_STATIC_MAX_N * KNN_STATIC_MAX_N * KNN_STATIC_MAX_N;
constexpr size_t knn_static_align(size_t b) { return (b + 255) & ~(size_t)255; }
constexpr size_t KNN_STATIC_BYTES = knn_static_align(32) + 2 * knn_static_align((KNN_STATIC_MAX_CELLS + 1) * 4) +
                                    knn_static_align(KNN_STATIC_MAX_CELLS * 4) +
                                    2 * knn_static_align((size_t)FDGS_KNN_STATIC_MAX_POINTS * 4);   // = 3 689 728 bytes

static __device__ __align__(256) unsigned char g_knn_static[KNN_STATIC_BYTES];

constexpr int KNN_STATIC_MAX_DEVICES = 64;
void* g_knn_static_addr[KNN_STATIC_MAX_DEVICES] = {};     // device address of g_knn_static on each device, filled on first use
std::mutex g_knn_static_mu;

// device address of the scratch array on the CURRENT device (0 = ok, else a cudaError_t / FDGS_E_*)
int knn_static_ptr(void** out) {
    int dev = -1;
    FDGS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= KNN_STATIC_MAX_DEVICES) return FDGS_E_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_knn_static_mu);
    if (!g_knn_static_addr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_knn_static);
        if (e != cudaSuccess) return (int)e;
        if (!p) return FDGS_E_UNSUPPORTED;
        g_knn_static_addr[dev] = p;
    }
    *out = g_knn_static_addr[dev];
    return FDGS_OK;
}

}  // namespace

extern "C" {

int fdgs_knn_static_max_points(void) { return FDGS_KNN_STATIC_MAX_POINTS; }

int fdgs_knn_mean_dist2_static(int P, const float* points, float* mean_dist2, void* stream_) {
    if (P < 0 || (P > 0 && (!points || !mean_dist2))) return FDGS_E_BADARG;
    if (P > FDGS_KNN_STATIC_MAX_POINTS) return FDGS_E_TOO_LARGE;
    if (P == 0) return FDGS_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int n = knn_cells_per_axis(P);
    const size_t cells = (size_t)n * n * n;
    if (n > KNN_STATIC_MAX_N || fdgs_knn_ws_bytes(P) > KNN_STATIC_BYTES) return FDGS_E_TOO_LARGE;
    void* base = nullptr;
    const int rc = knn_static_ptr(&base);
    if (rc) return rc;
    char* c = (char*)base;
    int* bb = (int*)c;                 c += fdgs_align_up(32);
    uint32_t* count = (uint32_t*)c;    c += fdgs_align_up((cells + 1) * 4);
    uint32_t* start = (uint32_t*)c;    c += fdgs_align_up((cells + 1) * 4);
    uint32_t* cursor = (uint32_t*)c;   c += fdgs_align_up(cells * 4);
    uint32_t* cell_of = (uint32_t*)c;  c += fdgs_align_up((size_t)P * 4);
    uint32_t* sorted = (uint32_t*)c;
    // the array keeps the previous call's contents: the two counters start at zero, everything else is overwritten before use
    FDGS_CUDA(cudaMemsetAsync(count, 0, (cells + 1) * 4, stream));
    FDGS_CUDA(cudaMemsetAsync(cursor, 0, cells * 4, stream));
    k_bbox_init<<<1, 32, 0, stream>>>(bb);
    k_bbox<<<std::min((P + 255) / 256, 1184), 256, 0, stream>>>(P, points, bb);
    k_knn_count<<<(P + 255) / 256, 256, 0, stream>>>(P, points, bb, n, count, cell_of);
    k_scan_u32<<<1, 1024, 0, stream>>>((int)cells, count, start);
    k_knn_scatter<<<(P + 255) / 256, 256, 0, stream>>>(P, cell_of, start, cursor, sorted);
    k_knn_search<<<(P + 127) / 128, 128, 0, stream>>>(P, points, bb, n, start, sorted, mean_dist2);
    FDGS_CUDA(cudaGetLastError());
    return FDGS_OK;
}

size_t fdgs_voxel_downsample_ws_bytes(int P) {
    const size_t cap = voxel_table_cap(P);
    return fdgs_align_up(32) + fdgs_align_up(cap * 8) + fdgs_align_up(cap * 48) + fdgs_align_up(cap * 4);
}

int fdgs_voxel_downsample(int P, const float* points, const float* colors, double voxel_size, float* out_points, float* out_colors,
                          int max_out, int32_t* n_out, void* ws, size_t ws_bytes, void* stream_) {
    if (P < 0 || !(voxel_size > 0.0) || !n_out || (P > 0 && (!points || !out_points || !ws))) return FDGS_E_BADARG;
    if ((colors == nullptr) != (out_colors == nullptr)) return FDGS_E_BADARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    FDGS_CUDA(cudaMemsetAsync(n_out, 0, 4, stream));
    if (P == 0) return FDGS_OK;
    if (ws_bytes < fdgs_voxel_downsample_ws_bytes(P)) return FDGS_E_WORKSPACE;
    VoxelTable t;
    t.cap = voxel_table_cap(P);
    char* c = (char*)ws;
    int* bb = (int*)c;                          c += fdgs_align_up(32);
    t.keys = (unsigned long long*)c;            c += fdgs_align_up((size_t)t.cap * 8);
    t.sums = (double*)c;                        c += fdgs_align_up((size_t)t.cap * 48);
    t.counts = (uint32_t*)c;
    FDGS_CUDA(cudaMemsetAsync(t.keys, 0xff, (size_t)t.cap * 8, stream));
    FDGS_CUDA(cudaMemsetAsync(t.sums, 0, (size_t)t.cap * 48, stream));
    FDGS_CUDA(cudaMemsetAsync(t.counts, 0, (size_t)t.cap * 4, stream));
    FDGS_PROF("k_voxel_downsample", stream);
    k_bbox_init<<<1, 32, 0, stream>>>(bb);
    k_bbox<<<std::min((P + 255) / 256, 1184), 256, 0, stream>>>(P, points, bb);
    k_voxel_accumulate<<<(P + 255) / 256, 256, 0, stream>>>(P, points, colors, bb, voxel_size, t);
    k_voxel_emit<<<(t.cap + 255) / 256, 256, 0, stream>>>(t, (uint32_t*)n_out, out_points, out_colors, max_out);
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

}  // extern "C"
