// Per-iteration training glue around the hot path:
//   k_plane_reg      HexPlane regularisers, value + gradient in one pass over the plane tables
//                    (scene/regulation.py:22-28 compute_plane_smoothness, scene/gaussian_model.py:538-577
//                     _plane_regulation / _time_regulation / _l1_regulation / compute_regulation)
//   k_adam           fused dense Adam over any number of parameter tensors with per-tensor lr / step
//                    (torch.optim.Adam(l, lr=0.0, eps=1e-15), scene/gaussian_model.py:172-184; same FP32 operation
//                     order as torch's foreach implementation, so the results are bit-identical), optionally fused with
//   densification    max_radii2D / xyz_gradient_accum / denom updates of the visible Gaussians
//                    (train.py:346-349, scene/gaussian_model.py:521-523 add_densification_stats)
// All three are HBM-bound sweeps; they replace ~40 (regularisers), ~10 per parameter (Adam) and ~8 (stats) ATen launches.
#include "fdgs_common.cuh"
#include "prof.cuh"
#include <algorithm>
#include <cmath>
#include <cstdint>

namespace {

constexpr unsigned FULL = 0xffffffffu;

// ---------------------------------------------------------------------------------------------------------------------
// Plane regularisers.  One job = one plane [C][H][W] (W contiguous):
//   loss += w_smooth * mean_{c, h < H-2, w} (t[h+2] - 2 t[h+1] + t[h])^2  +  w_l1 * mean |1 - t|
//   grad += g * d loss / d t          (g = *grad_scale, a device scalar, or 1)
// A thread owns one element (c, h, w): it evaluates the three second differences its value takes part in
// (d2[h-2], d2[h-1], d2[h]; the five rows it needs are L1 / L2 hits: the neighbouring threads' loads).
struct RegJob { const float* t; float* g; int C, H, W; float w_smooth, w_l1; long long first; };
struct RegArgs { RegJob j[FDGS_MAX_LEVELS * 6]; int njobs; long long total; };

__device__ __forceinline__ float second_diff(const float a, const float b, const float c) {
    // (c - b) - (b - a): the reference's first_difference[1:] - first_difference[:-1]
    return __fsub_rn(__fsub_rn(c, b), __fsub_rn(b, a));
}

__global__ void __launch_bounds__(1024)
k_plane_reg(RegArgs a, const float* __restrict__ grad_scale, float* __restrict__ loss)
{
    // grid-stride: at most one CTA per SM, so the value is the sum of <= 148 block partials; the per-thread and per-block
    // partial sums are kept in double (the value is a sum of ~10^7 terms of very different magnitude)
    double part = 0.0;
    const float gs = grad_scale ? *grad_scale : 1.0f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.total; i += (long long)gridDim.x * blockDim.x) {
        int jb = 0;
        while (jb + 1 < a.njobs && i >= a.j[jb + 1].first) ++jb;
        const RegJob& J = a.j[jb];
        const long long e = i - J.first;
        const int h = (int)((e / J.W) % J.H);
        const float* col = J.t + (e - (long long)h * J.W);      // element (c, 0, w)
        const size_t S = (size_t)J.W;
        const float v = col[h * S];
        float grad = 0.f;
        if (J.w_smooth != 0.f && J.H >= 3) {
            const float n_inv = 1.0f / ((float)J.C * (float)(J.H - 2) * (float)J.W);
            const float k = 2.0f * J.w_smooth * n_inv;
            // d2[m] = t[m+2] - 2 t[m+1] + t[m], m in [0, H-3];  d loss / d t[h] = k * (d2[h-2] - 2 d2[h-1] + d2[h])
            float d_m2 = 0.f, d_m1 = 0.f, d_0 = 0.f;
            if (h >= 2) d_m2 = second_diff(col[(h - 2) * S], col[(h - 1) * S], v);
            if (h >= 1 && h + 1 < J.H) d_m1 = second_diff(col[(h - 1) * S], v, col[(h + 1) * S]);
            if (h + 2 < J.H) {
                d_0 = second_diff(v, col[(h + 1) * S], col[(h + 2) * S]);
                part += (double)(J.w_smooth * n_inv) * (double)(d_0 * d_0);
            }
            grad += k * ((d_m2 - d_m1) - (d_m1 - d_0));
        }
        if (J.w_l1 != 0.f) {
            const float n_inv = 1.0f / ((float)J.C * (float)J.H * (float)J.W);
            const float d = 1.0f - v;
            part += (double)(J.w_l1 * n_inv) * (double)fabsf(d);
            grad += J.w_l1 * n_inv * (d > 0.f ? -1.f : (d < 0.f ? 1.f : 0.f));     // d|1 - t| / dt = -sign(1 - t)
        }
        if (J.g) J.g[e] += gs * grad;
    }
    // block reduction of the loss value: shuffle, shared memory, one atomic per CTA
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL, part, o);
    __shared__ double wsum[32];
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x == 0 && loss) {
        double s = 0.0;
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) s += wsum[k];
        if (s != 0.0) atomicAdd(loss, (float)s);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Fused Adam.  One job = one parameter tensor.  Per element, in FP32 and in exactly torch's foreach order
// (torch/optim/adam.py _multi_tensor_adam, capturable=False, weight_decay=0, amsgrad=False, maximize=False):
//   m  = m + (1-b1) * (g - m)                 _foreach_lerp_         (weight < 0.5 branch of at::native::lerp)
//   v  = v * b2 ;  v = v + (1-b2) * (g * g)   _foreach_mul_, _foreach_addcmul_
//   d  = sqrt(v) / sqrt(1 - b2^t) + eps       _foreach_sqrt, _foreach_div_, _foreach_add_
//   p  = p + (-(lr / (1 - b1^t))) * (m / d)   _foreach_addcdiv_
// with the scalars computed in double on the host and rounded to float once, as the foreach kernels receive them.
struct AdamJob { float* p; const float* g; float* m; float* v; long long n; long long first4; float one_minus_b1, b2, one_minus_b2,
                 bc2_sqrt, eps, neg_step; int zero_grad; };
constexpr int ADAM_MAX_JOBS = 64;
struct AdamArgs { AdamJob j[ADAM_MAX_JOBS]; int njobs; long long total4; };

__device__ __forceinline__ void adam_one(const AdamJob& J, float& p, const float g, float& m, float& v) {
    m = __fmaf_rn(J.one_minus_b1, __fsub_rn(g, m), m);
    v = __fmul_rn(v, J.b2);
    v = __fmaf_rn(J.one_minus_b2, __fmul_rn(g, g), v);
    const float d = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), J.bc2_sqrt), J.eps);
    p = __fmaf_rn(J.neg_step, __fdiv_rn(m, d), p);
}

__global__ void __launch_bounds__(256)
k_adam(AdamArgs a)
{
    // work items are groups of 4 consecutive floats of one tensor (float4 when aligned and complete)
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.total4; i += (long long)gridDim.x * blockDim.x) {
        int jb = 0;
        while (jb + 1 < a.njobs && i >= a.j[jb + 1].first4) ++jb;
        const AdamJob& J = a.j[jb];
        const long long e = (i - J.first4) * 4;
        const bool vec = (e + 4 <= J.n) && ((((uintptr_t)J.p | (uintptr_t)J.g | (uintptr_t)J.m | (uintptr_t)J.v) & 15) == 0);
        if (vec) {
            float4 p = *reinterpret_cast<float4*>(J.p + e);
            const float4 g = *reinterpret_cast<const float4*>(J.g + e);
            float4 m = *reinterpret_cast<float4*>(J.m + e);
            float4 v = *reinterpret_cast<float4*>(J.v + e);
            adam_one(J, p.x, g.x, m.x, v.x); adam_one(J, p.y, g.y, m.y, v.y);
            adam_one(J, p.z, g.z, m.z, v.z); adam_one(J, p.w, g.w, m.w, v.w);
            *reinterpret_cast<float4*>(J.p + e) = p;
            *reinterpret_cast<float4*>(J.m + e) = m;
            *reinterpret_cast<float4*>(J.v + e) = v;
            if (J.zero_grad) *reinterpret_cast<float4*>(const_cast<float*>(J.g) + e) = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            for (long long k = e; k < e + 4 && k < J.n; ++k) {
                float p = J.p[k], m = J.m[k], v = J.v[k];
                adam_one(J, p, J.g[k], m, v);
                J.p[k] = p; J.m[k] = m; J.v[k] = v;
                if (J.zero_grad) const_cast<float*>(J.g)[k] = 0.f;
            }
        }
    }
}

// Densification statistics of one iteration (train.py:346-349, gaussian_model.py:521-523), visible Gaussians only:
//   max_radii2D[i] = max(max_radii2D[i], radii[i]) ; xyz_gradient_accum[i] += || viewspace_grad[i, :2] || ; denom[i] += 1
__global__ void __launch_bounds__(256)
k_densify_stats(int P, const int32_t* __restrict__ radii, const uint8_t* __restrict__ visible,
                const float* __restrict__ vs_grad, float* __restrict__ max_radii2D, float* __restrict__ grad_accum,
                float* __restrict__ denom)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    const bool vis = visible ? visible[i] != 0 : r > 0;
    if (!vis) return;
    if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    if (grad_accum && vs_grad) {
        const float x = vs_grad[3 * (size_t)i], y = vs_grad[3 * (size_t)i + 1];
        grad_accum[i] += sqrtf(__fmaf_rn(y, y, __fmul_rn(x, x)));          // torch.norm(dim=-1) of the (x, y) pair
    }
    if (denom) denom[i] += 1.0f;
}

}  // namespace

extern "C" {

int fdgs_plane_regulation(const FdgsPlaneReg* a, void* stream_) {
    if (!a || a->levels < 0 || a->levels > FDGS_MAX_LEVELS || a->channels <= 0) return FDGS_E_BADARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    RegArgs ra;
    ra.njobs = 0;
    long long first = 0;
    for (int l = 0; l < a->levels; ++l)
        for (int k = 0; k < 6; ++k) {
            const bool time_plane = (k == 2 || k == 4 || k == 5);
            const float ws = time_plane ? a->time_smoothness_weight : a->plane_tv_weight;
            const float wl = time_plane ? a->l1_time_planes_weight : 0.f;
            if (ws == 0.f && wl == 0.f) continue;
            if (!a->planes[l][k]) return FDGS_E_BADARG;
            static const int pa[6] = {0, 0, 0, 1, 1, 2}, pb[6] = {1, 2, 3, 2, 3, 3};
            RegJob& J = ra.j[ra.njobs++];
            J.t = a->planes[l][k]; J.g = a->d_planes[l][k];
            J.C = a->channels; J.W = a->res[l][pa[k]]; J.H = a->res[l][pb[k]];
            if (J.W <= 0 || J.H <= 0) return FDGS_E_BADARG;
            J.w_smooth = ws; J.w_l1 = wl; J.first = first;
            first += (long long)J.C * J.H * J.W;
        }
    ra.total = first;
    if (ra.njobs == 0 || first == 0) return FDGS_OK;
    { FDGS_PROF("k_plane_reg", stream);
    k_plane_reg<<<(unsigned)std::min<long long>((first + 1023) / 1024, 148), 1024, 0, stream>>>(ra, a->grad_scale, a->loss);
    }
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

int fdgs_plane_smoothness(const float* plane, float* d_plane, int C, int H, int W, float weight, const float* grad_scale,
                          float* loss, void* stream_) {
    if (!plane || C <= 0 || H <= 0 || W <= 0) return FDGS_E_BADARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    RegArgs ra;
    ra.njobs = 1;
    ra.j[0] = RegJob{plane, d_plane, C, H, W, weight, 0.f, 0};
    ra.total = (long long)C * H * W;
    { FDGS_PROF("k_plane_reg", stream);
    k_plane_reg<<<(unsigned)std::min<long long>((ra.total + 1023) / 1024, 148), 1024, 0, stream>>>(ra, grad_scale, loss);
    }
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

int fdgs_adam_step(const FdgsAdamStep* a, void* stream_) {
    if (!a || a->ntensors < 0 || (a->ntensors > 0 && !a->t)) return FDGS_E_BADARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    int done = 0;
    while (done < a->ntensors) {
        AdamArgs aa;
        aa.njobs = 0;
        long long first4 = 0;
        for (; done < a->ntensors && aa.njobs < ADAM_MAX_JOBS; ++done) {
            const FdgsAdamTensor& t = a->t[done];
            if (t.n < 0 || (t.n > 0 && (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq)) || t.step < 1) return FDGS_E_BADARG;
            if (t.n == 0) continue;
            AdamJob& J = aa.j[aa.njobs++];
            J.p = t.param; J.g = t.grad; J.m = t.exp_avg; J.v = t.exp_avg_sq; J.n = t.n; J.first4 = first4;
            const double b1 = a->beta1, b2 = a->beta2;
            const double bc1 = 1.0 - std::pow(b1, (double)t.step), bc2 = 1.0 - std::pow(b2, (double)t.step);
            J.one_minus_b1 = (float)(1.0 - b1); J.b2 = (float)b2; J.one_minus_b2 = (float)(1.0 - b2);
            J.bc2_sqrt = (float)std::sqrt(bc2); J.eps = (float)a->eps; J.neg_step = (float)((t.lr / bc1) * -1.0);
            J.zero_grad = a->zero_grad;
            first4 += (t.n + 3) / 4;
        }
        aa.total4 = first4;
        if (first4 == 0) continue;
        long long blocks = (first4 + 255) / 256;
        if (blocks > 148LL * 16) blocks = 148LL * 16;
        { FDGS_PROF("k_adam", stream);
        k_adam<<<(unsigned)blocks, 256, 0, stream>>>(aa);
        }
        FDGS_LAUNCH_CHECK(0, stream);
    }
    return FDGS_OK;
}

int fdgs_densify_stats(int P, const int32_t* radii, const uint8_t* visible, const float* viewspace_grad, float* max_radii2D,
                       float* xyz_gradient_accum, float* denom, void* stream_) {
    if (P < 0 || (P > 0 && !radii)) return FDGS_E_BADARG;
    if (P == 0) return FDGS_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    { FDGS_PROF("k_densify_stats", stream);
    k_densify_stats<<<(P + 255) / 256, 256, 0, stream>>>(P, radii, visible, viewspace_grad, max_radii2D, xyz_gradient_accum, denom);
    }
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

}  // extern "C"
