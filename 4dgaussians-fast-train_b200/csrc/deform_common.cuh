// Shared host/device pieces of the deformation kernels: plane layout, parameter pointer blocks,
// ATen-compatible coordinate mapping and the per-point HexPlane sampler.
#pragma once
#include "fdgs_common.cuh"

namespace fdgs_deform {

constexpr int TP = 128;      // points per tile
constexpr int TPP = 132;     // padded pitch of the k-major activation tiles (16-byte aligned rows)
constexpr int NT = 256;      // threads per CTA
constexpr int KC = 32;       // reduction chunk staged per step
constexpr int BP = 132;      // pitch of the staged weight chunk

__host__ __device__ constexpr int pa_of(int k) { return k < 3 ? 0 : (k < 5 ? 1 : 2); }
__host__ __device__ constexpr int pb_of(int k) { return k == 0 ? 1 : (k == 1 || k == 3) ? 2 : 3; }
static const int h_pa[6] = {0, 0, 0, 1, 1, 2};
static const int h_pb[6] = {1, 2, 3, 2, 3, 3};
static const int h_tslot[6] = {-1, -1, 0, -1, 1, 2};   // time-plane slot of plane k

struct Layout {
    int L, C, F, Wn, depth, head_mask, apply_rotation;
    int pw[FDGS_MAX_LEVELS][6], ph[FDGS_MAX_LEVELS][6];
    long long poff[FDGS_MAX_LEVELS][6];     // float offset of packed plane [H][W][C]
    long long toff[FDGS_MAX_LEVELS][3];     // float offset of per-view time row table [W][C] (slots: xt, yt, zt)
    long long plane_floats, trow_floats;
    float aabb[6];
};

static int make_layout(const FdgsDeformNet& n, Layout& lo) {
    if (n.levels < 1 || n.levels > FDGS_MAX_LEVELS) return FDGS_E_BADARG;
    if (!(n.channels == 4 || (n.channels % 8 == 0 && n.channels >= 8 && n.channels <= 64))) return FDGS_E_UNSUPPORTED;
    if (n.width != 32 && n.width != 64 && n.width != 128) return FDGS_E_UNSUPPORTED;
    lo.L = n.levels; lo.C = n.channels; lo.F = n.levels * n.channels; lo.Wn = n.width;
    lo.depth = n.depth < 1 ? 1 : n.depth;
    if (lo.depth > 8) return FDGS_E_UNSUPPORTED;
    if (lo.F > 128) return FDGS_E_UNSUPPORTED;
    lo.head_mask = n.head_mask; lo.apply_rotation = n.apply_rotation;
    long long off = 0, toff = 0;
    for (int l = 0; l < lo.L; ++l)
        for (int k = 0; k < 6; ++k) {
            lo.pw[l][k] = n.res[l][h_pa[k]];
            lo.ph[l][k] = n.res[l][h_pb[k]];
            if (lo.pw[l][k] < 2 || lo.ph[l][k] < 2) return FDGS_E_BADARG;
            lo.poff[l][k] = off;
            off += (long long)lo.pw[l][k] * lo.ph[l][k] * lo.C;
            if (h_tslot[k] >= 0) { lo.toff[l][h_tslot[k]] = toff; toff += (long long)lo.pw[l][k] * lo.C; }
        }
    lo.plane_floats = off; lo.trow_floats = toff;
    for (int i = 0; i < 6; ++i) lo.aabb[i] = n.aabb[i];
    return FDGS_OK;
}

struct NetPtrs {
    const float* planes[FDGS_MAX_LEVELS][6];
    const float* trunk_w[8]; const float* trunk_b[8];
    const float* w1[FDGS_NUM_HEADS]; const float* b1[FDGS_NUM_HEADS];
    const float* w2[FDGS_NUM_HEADS]; const float* b2[FDGS_NUM_HEADS];
};
struct GradPtrs {
    float* planes[FDGS_MAX_LEVELS][6];
    float* trunk_w[8]; float* trunk_b[8];
    float* w1[FDGS_NUM_HEADS]; float* b1[FDGS_NUM_HEADS];
    float* w2[FDGS_NUM_HEADS]; float* b2[FDGS_NUM_HEADS];
};

// Activations the training forward keeps for the backward (floats; rows padded to whole 128-point tiles, blocked layout
// [tile][row/32][col/4][row%32][4] so that 32 consecutive rows are one contiguous run of 32*cols floats).
// off_M: ReLU gate bit masks, uint32 [6 slots: trunk, head 0..4][Pp][4 words: column half 0 (2 words), half 1 (2 words)].
struct SaveLayout { long long Pp, off_F, off_H, off_RZ[FDGS_NUM_HEADS], off_M, total; };
static inline SaveLayout make_save_layout(const Layout& lo, int P) {
    SaveLayout s;
    s.Pp = (((long long)P + TP - 1) / TP) * TP;
    auto al = [](long long x) { return (x + 63) & ~63LL; };
    long long o = 0;
    s.off_F = o; o += al(s.Pp * lo.F);
    s.off_H = o; o += al(s.Pp * lo.Wn);
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) { s.off_RZ[h] = o; if ((lo.head_mask >> h) & 1) o += al(s.Pp * lo.Wn); }
    s.off_M = o; o += al(6 * s.Pp * 4);
    s.total = o;
    return s;
}

__host__ __device__ inline int head_out(int h) { return h == 0 ? 3 : h == 1 ? 3 : h == 2 ? 4 : h == 3 ? 1 : 48; }

// ATen grid_sampler coordinate mapping, align_corners=True, padding_mode='border'
// (ATen/native/GridSampler.h: grid_sampler_unnormalize + clip_coordinates).
__device__ __forceinline__ float unnorm_clip(float q, int size, float* gmul = nullptr) {
    float ix = ((q + 1.f) / 2) * (float)(size - 1);
    if (gmul) *gmul = (ix <= 0.f || ix >= (float)(size - 1)) ? 0.f : (float)(size - 1) * 0.5f;
    return fminf((float)(size - 1), fmaxf(ix, 0.f));
}

// ---- sampling ------------------------------------------------------------------------------------------
struct PlaneTap {            // one plane's bilinear tap for one point
    int i00, i01, i10, i11;  // element offsets (already * C) or -1 when out of range
    float w00, w01, w10, w11;
};

__device__ __forceinline__ void point_coords(const Layout& lo, const float* __restrict__ xyz, int p, float time, float q[4]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float mx = lo.aabb[a], mn = lo.aabb[3 + a];
        q[a] = (xyz[3 * (size_t)p + a] - mx) * (2.0f / (mn - mx)) - 1.0f;
    }
    q[3] = time;
}

// Forward sample of NCH consecutive channels [c0, c0+NCH) of level l for one point into prod[].
template <int NCH>
__device__ __forceinline__ void sample_level(const Layout& lo, int l, const float q[4], const float* __restrict__ packed,
                                             const float* __restrict__ trow, int c0, float (&prod)[NCH]) {
    const int C = lo.C;
#pragma unroll
    for (int c = 0; c < NCH; ++c) prod[c] = 1.f;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const int a = pa_of(k);
        const int W = lo.pw[l][k];
        const float ix = unnorm_clip(q[a], W);
        const float x0f = floorf(ix);
        const int x0 = (int)x0f;
        const float wx0 = (x0f + 1.f) - ix, wx1 = ix - x0f;
        const bool xin = x0 + 1 < W;
        float s[NCH];
        if (k == 2 || k == 4 || k == 5) {
            const float* T = trow + lo.toff[l][k == 2 ? 0 : (k == 4 ? 1 : 2)] + c0;
#pragma unroll
            for (int c4 = 0; c4 < NCH; c4 += 4) {
                const float4 v0 = *reinterpret_cast<const float4*>(T + (size_t)x0 * C + c4);
                float4 v1 = make_float4(0.f, 0.f, 0.f, 0.f);
                if (xin) v1 = *reinterpret_cast<const float4*>(T + (size_t)(x0 + 1) * C + c4);
                s[c4 + 0] = v0.x * wx0 + v1.x * wx1; s[c4 + 1] = v0.y * wx0 + v1.y * wx1;
                s[c4 + 2] = v0.z * wx0 + v1.z * wx1; s[c4 + 3] = v0.w * wx0 + v1.w * wx1;
            }
        } else {
            const int b = pb_of(k);
            const int H = lo.ph[l][k];
            const float iy = unnorm_clip(q[b], H);
            const float y0f = floorf(iy);
            const int y0 = (int)y0f;
            const float wy0 = (y0f + 1.f) - iy, wy1 = iy - y0f;
            const bool yin = y0 + 1 < H;
            const float nw = wx0 * wy0, ne = wx1 * wy0, sw = wx0 * wy1, se = wx1 * wy1;
            const float* Pp = packed + lo.poff[l][k] + c0;
            const size_t o00 = ((size_t)y0 * W + x0) * C;
#pragma unroll
            for (int c4 = 0; c4 < NCH; c4 += 4) {
                const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 v00 = *reinterpret_cast<const float4*>(Pp + o00 + c4);
                const float4 v01 = xin ? *reinterpret_cast<const float4*>(Pp + o00 + C + c4) : z4;
                const float4 v10 = yin ? *reinterpret_cast<const float4*>(Pp + o00 + (size_t)W * C + c4) : z4;
                const float4 v11 = (xin && yin) ? *reinterpret_cast<const float4*>(Pp + o00 + (size_t)W * C + C + c4) : z4;
                s[c4 + 0] = v00.x * nw + v01.x * ne + v10.x * sw + v11.x * se;
                s[c4 + 1] = v00.y * nw + v01.y * ne + v10.y * sw + v11.y * se;
                s[c4 + 2] = v00.z * nw + v01.z * ne + v10.z * sw + v11.z * se;
                s[c4 + 3] = v00.w * nw + v01.w * ne + v10.w * sw + v11.w * se;
            }
        }
#pragma unroll
        for (int c = 0; c < NCH; ++c) prod[c] *= s[c];
    }
}

}  // namespace fdgs_deform
