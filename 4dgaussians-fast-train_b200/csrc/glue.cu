// Caller-side glue between the deformation field and the rasterizer (include/fdgs.h: fdgs_activate_*): the scale /
// rotation / opacity activations of gaussian_renderer/__init__.py:97-99 (torch.exp, F.normalize(dim=1, eps=1e-12),
// torch.sigmoid; scene/gaussian_model.py:47-58) and their backward, one launch each way instead of ~15 ATen kernels.
#include "raster_math.cuh"
#include "prof.cuh"
#include <cstdint>

namespace {

__global__ void __launch_bounds__(256)
k_activate_fwd(int P, const float* __restrict__ s_raw, const float* __restrict__ r_raw, const float* __restrict__ o_raw,
               float* __restrict__ s, float* __restrict__ r, float* __restrict__ o)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if (s) {
#pragma unroll
        for (int k = 0; k < 3; ++k) s[3 * (size_t)i + k] = fdgs_act_scale(s_raw[3 * (size_t)i + k]);
    }
    if (r) reinterpret_cast<float4*>(r)[i] = fdgs_act_quat(reinterpret_cast<const float4*>(r_raw)[i]);
    if (o) o[i] = fdgs_act_opacity(o_raw[i]);
}

__global__ void __launch_bounds__(256)
k_activate_bwd(int P, const float* __restrict__ r_raw, const float* __restrict__ s, const float* __restrict__ r,
               const float* __restrict__ o, const float* __restrict__ gs, const float* __restrict__ gr,
               const float* __restrict__ go, float* __restrict__ ds, float* __restrict__ dr, float* __restrict__ dopa)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if (ds) {
#pragma unroll
        for (int k = 0; k < 3; ++k) ds[3 * (size_t)i + k] = fdgs_act_scale_vjp(gs[3 * (size_t)i + k], s[3 * (size_t)i + k]);
    }
    if (dr)
        reinterpret_cast<float4*>(dr)[i] = fdgs_act_quat_vjp(reinterpret_cast<const float4*>(r_raw)[i], reinterpret_cast<const float4*>(r)[i],
                                                             reinterpret_cast<const float4*>(gr)[i]);
    if (dopa) dopa[i] = fdgs_act_opacity_vjp(go[i], o[i]);
}

// L1 loss: grid-stride float4 sweep, per-thread partial -> warp shuffle -> one atomic per warp.
__global__ void __launch_bounds__(256)
k_l1_fwd(const float* __restrict__ a, const float* __restrict__ b, long long n, float scale, float* __restrict__ loss)
{
    const long long n4 = n >> 2;
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 x = reinterpret_cast<const float4*>(a)[i], y = reinterpret_cast<const float4*>(b)[i];
        acc += (fabsf(x.x - y.x) + fabsf(x.y - y.y)) + (fabsf(x.z - y.z) + fabsf(x.w - y.w));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) acc += fabsf(a[(n4 << 2) + threadIdx.x] - b[(n4 << 2) + threadIdx.x]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += part[w];
        atomicAdd(loss, t * scale);
    }
}

__device__ __forceinline__ float sgn(float v) { return (float)((v > 0.f) - (v < 0.f)); }

__global__ void __launch_bounds__(256)
k_l1_bwd(const float* __restrict__ a, const float* __restrict__ b, long long n, float scale, const float* __restrict__ g,
         float* __restrict__ d_a)
{
    const float k = *g * scale;
    const long long n4 = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 x = reinterpret_cast<const float4*>(a)[i], y = reinterpret_cast<const float4*>(b)[i];
        reinterpret_cast<float4*>(d_a)[i] = make_float4(k * sgn(x.x - y.x), k * sgn(x.y - y.y), k * sgn(x.z - y.z), k * sgn(x.w - y.w));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const long long i = (n4 << 2) + threadIdx.x;
        d_a[i] = k * sgn(a[i] - b[i]);
    }
}

}  // namespace

extern "C" {

int fdgs_activate_forward(const FdgsActivateForward* a, void* stream_) {
    if (!a || a->P < 0) return FDGS_E_BADARG;
    if ((a->scales && !a->scales_raw) || (a->rotations && !a->rotations_raw) || (a->opacity && !a->opacity_raw)) return FDGS_E_BADARG;
    if (a->P == 0) return FDGS_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    { FDGS_PROF("k_activate_fwd", stream);
    k_activate_fwd<<<(a->P + 255) / 256, 256, 0, stream>>>(a->P, a->scales_raw, a->rotations_raw, a->opacity_raw, a->scales, a->rotations, a->opacity);
    }
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

int fdgs_activate_backward(const FdgsActivateBackward* a, void* stream_) {
    if (!a || a->P < 0) return FDGS_E_BADARG;
    if ((a->d_scales_raw && (!a->g_scales || !a->scales)) || (a->d_rotations_raw && (!a->g_rotations || !a->rotations || !a->rotations_raw)) ||
        (a->d_opacity_raw && (!a->g_opacity || !a->opacity))) return FDGS_E_BADARG;
    if (a->P == 0) return FDGS_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    { FDGS_PROF("k_activate_bwd", stream);
    k_activate_bwd<<<(a->P + 255) / 256, 256, 0, stream>>>(a->P, a->rotations_raw, a->scales, a->rotations, a->opacity, a->g_scales, a->g_rotations,
                                                           a->g_opacity, a->d_scales_raw, a->d_rotations_raw, a->d_opacity_raw);
    }
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

static int l1_grid(long long n) {
    const long long want = ((n >> 2) + 255) / 256;
    return (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
}

int fdgs_l1_loss_forward(const float* a, const float* b, long long n, float scale, float* loss, void* stream_) {
    if (!a || !b || !loss || n < 0) return FDGS_E_BADARG;
    if ((((uintptr_t)a | (uintptr_t)b) & 15) != 0) return FDGS_E_BADARG;       // float4 sweeps
    if (n == 0) return FDGS_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    { FDGS_PROF("k_l1_fwd", stream);
    k_l1_fwd<<<l1_grid(n), 256, 0, stream>>>(a, b, n, scale, loss);
    }
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

int fdgs_l1_loss_backward(const float* a, const float* b, long long n, float scale, const float* g, float* d_a, void* stream_) {
    if (!a || !b || !g || !d_a || n < 0) return FDGS_E_BADARG;
    if ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)d_a) & 15) != 0) return FDGS_E_BADARG;
    if (n == 0) return FDGS_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    { FDGS_PROF("k_l1_bwd", stream);
    k_l1_bwd<<<l1_grid(n), 256, 0, stream>>>(a, b, n, scale, g, d_a);
    }
    FDGS_LAUNCH_CHECK(0, stream);
    return FDGS_OK;
}

}  // extern "C"
