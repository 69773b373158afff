// Per-Gaussian projection maths shared by the forward and backward preprocess kernels.
//
// The expressions are written as plain FP32 arithmetic in the same evaluation
// order as the reference's glm-based code so that nvcc's default FMA contraction
// produces the same rounding sequence (radii, depth bits and tile rects have to be
// bit-exact, see DESIGN.md "numerics").  Matrices are column-major 3x3, m[c][r].
#pragma once
#include "fdgs_common.cuh"

struct M3 { float m[3][3]; };   // m[column][row]

// ---- the three activations between the deformation field and the rasterizer (gaussian_renderer/__init__.py:97-99:
// torch.exp, F.normalize(dim=1, eps=1e-12), torch.sigmoid) and their VJPs.  ONE definition, used by the stand-alone
// glue kernels (glue.cu) and by the preprocess kernels when they take the raw deformation outputs (`activate`), so both
// routes produce the same bits.
constexpr float FDGS_NORM_EPS = 1e-12f;
__device__ __forceinline__ float fdgs_act_scale(float raw) { return expf(raw); }
__device__ __forceinline__ float fdgs_act_opacity(float raw) { return 1.0f / (1.0f + expf(-raw)); }
__device__ __forceinline__ float fdgs_quat_norm(const float4 q) {
    return sqrtf(__fmaf_rn(q.w, q.w, __fmaf_rn(q.z, q.z, __fmaf_rn(q.y, q.y, __fmul_rn(q.x, q.x)))));
}
__device__ __forceinline__ float4 fdgs_act_quat(const float4 q) {
    const float d = fmaxf(fdgs_quat_norm(q), FDGS_NORM_EPS);
    return make_float4(q.x / d, q.y / d, q.z / d, q.w / d);
}
__device__ __forceinline__ float fdgs_act_scale_vjp(float g, float s) { return g * s; }
__device__ __forceinline__ float fdgs_act_opacity_vjp(float g, float o) { return g * o * (1.0f - o); }
// raw: the un-normalised quaternion, u = fdgs_act_quat(raw), g = dL/du
__device__ __forceinline__ float4 fdgs_act_quat_vjp(const float4 raw, const float4 u, const float4 g) {
    const float n = fdgs_quat_norm(raw);
    if (n >= FDGS_NORM_EPS) {             // d(q / n): (g - u (u . g)) / n
        const float dot = u.x * g.x + u.y * g.y + u.z * g.z + u.w * g.w;
        const float inv = 1.0f / n;
        return make_float4((g.x - u.x * dot) * inv, (g.y - u.y * dot) * inv, (g.z - u.z * dot) * inv, (g.w - u.w * dot) * inv);
    }
    // clamp_min(eps) is active: the denominator is the constant eps
    return make_float4(g.x / FDGS_NORM_EPS, g.y / FDGS_NORM_EPS, g.z / FDGS_NORM_EPS, g.w / FDGS_NORM_EPS);
}

// R = A * B with the left-to-right summation order of a textbook column-major product.
__device__ __forceinline__ M3 m3_mul(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int q = 0; q < 3; ++q)
            r.m[c][q] = a.m[0][q] * b.m[c][0] + a.m[1][q] * b.m[c][1] + a.m[2][q] * b.m[c][2];
    return r;
}
__device__ __forceinline__ M3 m3_T(const M3& a) {
    M3 r;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int q = 0; q < 3; ++q) r.m[c][q] = a.m[q][c];
    return r;
}

// x' = M p (+ translation), M stored transposed (row-vector convention, scene/cameras.py:61-65).
__device__ __forceinline__ float3 xform43(const float3 p, const float* __restrict__ m) {
    return make_float3(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                       m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                       m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]);
}
__device__ __forceinline__ float4 xform44(const float3 p, const float* __restrict__ m) {
    return make_float4(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                       m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                       m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
                       m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]);
}

// Rotation matrix columns from an UN-normalised quaternion (r,x,y,z)
// (reference keeps it un-normalised: RAST/cuda_rasterizer/forward.cu:127).
__device__ __forceinline__ M3 quat_to_m3(float r, float x, float y, float z) {
    M3 R;
    R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y - r * z);       R.m[0][2] = 2.f * (x * z + r * y);
    R.m[1][0] = 2.f * (x * y + r * z);       R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z - r * x);
    R.m[2][0] = 2.f * (x * z - r * y);       R.m[2][1] = 2.f * (y * z + r * x);       R.m[2][2] = 1.f - 2.f * (x * x + y * y);
    return R;
}

// Sigma = (S R)^T (S R), upper triangle (forward.cu:118-152).
__device__ __forceinline__ void cov3d_from_scale_rot(float3 s, float mod, float4 q, float* __restrict__ c6) {
    M3 S;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r) S.m[c][r] = (c == r) ? 1.0f : 0.0f;
    S.m[0][0] = mod * s.x; S.m[1][1] = mod * s.y; S.m[2][2] = mod * s.z;
    M3 R = quat_to_m3(q.x, q.y, q.z, q.w);
    M3 M = m3_mul(S, R);
    M3 Sg = m3_mul(m3_T(M), M);
    c6[0] = Sg.m[0][0]; c6[1] = Sg.m[0][1]; c6[2] = Sg.m[0][2];
    c6[3] = Sg.m[1][1]; c6[4] = Sg.m[1][2]; c6[5] = Sg.m[2][2];
}

// EWA projection of the 3-D covariance (forward.cu:74-113).  Returns (a, b, c) of the
// 2x2 screen covariance with the 0.3 low-pass already added; optionally the T matrix.
__device__ __forceinline__ float3 cov2d_project(const float3 mean, float fx, float fy, float tanx, float tany,
                                                const float* __restrict__ c6, const float* __restrict__ V,
                                                M3* Tout = nullptr, float3* tclamped = nullptr,
                                                float* xmul = nullptr, float* ymul = nullptr) {
    float3 t = xform43(mean, V);
    const float limx = 1.3f * tanx;
    const float limy = 1.3f * tany;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    if (xmul) *xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    if (ymul) *ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;

    M3 J;
    J.m[0][0] = fx / t.z; J.m[0][1] = 0.0f;     J.m[0][2] = -(fx * t.x) / (t.z * t.z);
    J.m[1][0] = 0.0f;     J.m[1][1] = fy / t.z; J.m[1][2] = -(fy * t.y) / (t.z * t.z);
    J.m[2][0] = 0.0f;     J.m[2][1] = 0.0f;     J.m[2][2] = 0.0f;
    M3 Wm;
    Wm.m[0][0] = V[0]; Wm.m[0][1] = V[4]; Wm.m[0][2] = V[8];
    Wm.m[1][0] = V[1]; Wm.m[1][1] = V[5]; Wm.m[1][2] = V[9];
    Wm.m[2][0] = V[2]; Wm.m[2][1] = V[6]; Wm.m[2][2] = V[10];
    M3 T = m3_mul(Wm, J);
    M3 Vrk;
    Vrk.m[0][0] = c6[0]; Vrk.m[0][1] = c6[1]; Vrk.m[0][2] = c6[2];
    Vrk.m[1][0] = c6[1]; Vrk.m[1][1] = c6[3]; Vrk.m[1][2] = c6[4];
    Vrk.m[2][0] = c6[2]; Vrk.m[2][1] = c6[4]; Vrk.m[2][2] = c6[5];
    M3 cov = m3_mul(m3_mul(m3_T(T), m3_T(Vrk)), T);
    cov.m[0][0] += 0.3f;
    cov.m[1][1] += 0.3f;
    if (Tout) *Tout = T;
    if (tclamped) *tclamped = t;
    return make_float3(cov.m[0][0], cov.m[0][1], cov.m[1][1]);
}

// Real spherical-harmonics basis constants (degree <= 3), standard published values.
#define FDGS_SH_C0 0.28209479177387814f
#define FDGS_SH_C1 0.4886025119029199f
#define FDGS_SH_C2_0 1.0925484305920792f
#define FDGS_SH_C2_1 -1.0925484305920792f
#define FDGS_SH_C2_2 0.31539156525252005f
#define FDGS_SH_C2_3 -1.0925484305920792f
#define FDGS_SH_C2_4 0.5462742152960396f
#define FDGS_SH_C3_0 -0.5900435899266435f
#define FDGS_SH_C3_1 2.890611442640554f
#define FDGS_SH_C3_2 -0.4570457994644658f
#define FDGS_SH_C3_3 0.3731763325901154f
#define FDGS_SH_C3_4 -0.4570457994644658f
#define FDGS_SH_C3_5 1.445305721320277f
#define FDGS_SH_C3_6 -0.5900435899266435f

// View-dependent colour of one channel from its SH coefficients (forward.cu:20-71);
// `s(k)` fetches coefficient k of this channel.  Returns the pre-clamp value.
template <class Fetch>
__device__ __forceinline__ float sh_eval_channel(int deg, float x, float y, float z, Fetch s) {
    float v = FDGS_SH_C0 * s(0);
    if (deg > 0) {
        v = v - FDGS_SH_C1 * y * s(1) + FDGS_SH_C1 * z * s(2) - FDGS_SH_C1 * x * s(3);
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            v = v + FDGS_SH_C2_0 * xy * s(4) + FDGS_SH_C2_1 * yz * s(5) + FDGS_SH_C2_2 * (2.0f * zz - xx - yy) * s(6) +
                FDGS_SH_C2_3 * xz * s(7) + FDGS_SH_C2_4 * (xx - yy) * s(8);
            if (deg > 2) {
                v = v + FDGS_SH_C3_0 * y * (3.0f * xx - yy) * s(9) + FDGS_SH_C3_1 * xy * z * s(10) +
                    FDGS_SH_C3_2 * y * (4.0f * zz - xx - yy) * s(11) +
                    FDGS_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * s(12) +
                    FDGS_SH_C3_4 * x * (4.0f * zz - xx - yy) * s(13) + FDGS_SH_C3_5 * z * (xx - yy) * s(14) +
                    FDGS_SH_C3_6 * x * (xx - 3.0f * yy) * s(15);
            }
        }
    }
    return v + 0.5f;
}
