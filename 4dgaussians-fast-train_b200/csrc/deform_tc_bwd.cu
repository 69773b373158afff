// Tensor-core (tcgen05 + TMEM) backward of the fused HexPlane sampler + deformation MLP for sm_100a.
//
// Two kernels per view, both 3xTF32 (FP32-accurate), exchanging FP32 activations through a caller-owned
// scratch buffer in HBM (1.9 GB at P = 300k, Wn = 128, five heads: cheap against 6.5 TB/s):
//
//   k_deform_bwd_data   persistent CTA per SM, 128-point tiles, TMEM lane = point (same skeleton as the forward):
//        feat = sample(planes)            -> S_F            h  = feat W0^T (A smem)     relu(h) -> S_H, smem (hi/lo)
//        per head j: z = relu(h) W1_j^T (A smem) ; relu(z) -> S_RZ_j ; g_z = (g_o W2_j) [z>0] (CUDA cores) -> S_GZ_j
//                    and hi/lo into TMEM ;  g_rh += g_z W1_j   (A from TMEM, B = W1_j^T chunks)
//        g_h = g_rh [h>0] -> S_GH, TMEM ; g_feat = g_h W0 (A from TMEM) ; sampler backward (vector atomics) ; d_xyz
//   k_deform_bwd_weights   split-K "A^T B over points" GEMMs on MN-major (swizzled tile32) operands:
//        dW1_j = S_GZ_j^T S_H   db1_j = colsum S_GZ_j      dW2_j^T = S_RZ_j^T g_o_j   db2_j = colsum g_o_j
//        dW0   = S_GH^T  S_F    db0   = colsum S_GH
//     one CTA per (job, slice of points); FP32 rows are split hi/lo in registers on their way to shared memory;
//     partial results are reduced into the caller's gradient buffers with red.global.add.
//
// Replaces autograd over scene/deformation.py:45-148 + scene/hexplane.py:73-183 (ATen grid_sampler_2d_backward,
// addmm backward: ~400 launches per view in the reference).
#include "deform_common.cuh"
#include "umma.cuh"
#include "prof.cuh"
#include <cstdio>
#include <cstdlib>

namespace fdgs_deform_tc_bwd {
using namespace fdgs_deform;
using namespace fdgs_umma;

// Saved-activation mode: the g_relu_h += g_z W1 GEMM of a head starts on the first column chunk of g_z while the epilogue
// warps still produce the second one, and the next head's g_z goes into a chunk as soon as the tensor core has read it
// (1: default; 0: one head at a time - epilogue and tensor core alternate - for A/B).
#ifndef FDGS_BWD_PIPE
#define FDGS_BWD_PIPE 1
#endif
constexpr int WACC_CTAS = 192;     // weight kernel: CTAs that can have a private accumulation buffer (>= SMs of the device)
constexpr int KCH = 16;            // K columns per streamed weight chunk
constexpr int NTHREADS = 320;      // 8 sampler/epilogue warps + producer + MMA
constexpr int NEPI = 256;

struct Plan {
    int Wn, F, nheads;
    int heads[FDGS_NUM_HEADS];     // active heads that received an upstream gradient
    int nchunks_main;              // chunks of 32*Wn floats per tile: W0, then per head W1_j, W2_j^T (K padded to 16), W1_j^T
    int k2chunks[FDGS_NUM_HEADS];  // W2_j^T chunks of plan slot j: 1 (out <= 4) or 3 (out = 48)
    int cc[FDGS_NUM_HEADS];        // plan slot j (saved mode, out <= 4): g_relu_z = g_o W2 on the CUDA cores, chunk-pipelined (below)
    int seq_n;                     // main chunks streamed per tile, in MMA order (all of them, or - when the forward kept
    short seq[112];                // the activations - only the W2_j^T and W1_j^T chunks)
    int nchunks_gf;                // chunks of 32*F floats (W0^T)
    long long wprep_floats;
    // scratch (floats), Pp = ntiles * 128 rows each
    long long Pp;
    long long off_wprep, off_F, off_H, off_GH, off_GF, off_RZ[FDGS_NUM_HEADS], off_GZ[FDGS_NUM_HEADS], off_wacc, total_floats;
};

static inline bool tc_supported(const Layout& lo) {
    return lo.depth == 1 && (lo.Wn == 64 || lo.Wn == 128) && lo.F % KCH == 0 && lo.F >= KCH && lo.F <= 64 && lo.C >= 8;
}

static inline Plan make_plan(const Layout& lo, int P, const bool has_grad[FDGS_NUM_HEADS], bool saved = false) {
    Plan p;
    p.Wn = lo.Wn; p.F = lo.F; p.nheads = 0;
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) if (((lo.head_mask >> h) & 1) && has_grad[h]) p.heads[p.nheads++] = h;
    for (int h = p.nheads; h < FDGS_NUM_HEADS; ++h) p.heads[h] = 0;
    p.nchunks_main = lo.F / KCH + p.nheads * 2 * (lo.Wn / KCH);
    for (int j = 0; j < FDGS_NUM_HEADS; ++j) {
        p.k2chunks[j] = j < p.nheads ? (head_out(p.heads[j]) + KCH - 1) / KCH : 0;
        p.nchunks_main += p.k2chunks[j];
    }
    for (int j = 0; j < FDGS_NUM_HEADS; ++j)
        p.cc[j] = (FDGS_BWD_PIPE && saved && j < p.nheads && head_out(p.heads[j]) <= 4) ? 1 : 0;
    p.nchunks_gf = lo.Wn / KCH;
    p.wprep_floats = (long long)p.nchunks_main * 32 * lo.Wn + (long long)p.nchunks_gf * 32 * lo.F;
    {
        int c = 0, n = 0;
        for (int k = 0; k < lo.F / KCH; ++k, ++c) if (!saved) p.seq[n++] = (short)c;                 // W0
        for (int j = 0; j < p.nheads; ++j) {
            for (int k = 0; k < lo.Wn / KCH; ++k, ++c) if (!saved) p.seq[n++] = (short)c;            // W1_j
            for (int k = 0; k < p.k2chunks[j]; ++k, ++c) if (!p.cc[j]) p.seq[n++] = (short)c;        // W2_j^T
            if (saved && FDGS_BWD_PIPE) {
                // W1_j^T in the order the pipelined g_relu_h GEMM consumes it: the epilogue threads deliver g_z in two
                // column chunks per half (thread = row, column half hf), so K chunk 0 = first half of both halves
                const int nk = lo.Wn / KCH, hk = nk / 2, qk = hk / 2 > 0 ? hk / 2 : 1;
                for (int ch = 0; ch < 2; ++ch)
                    for (int hf = 0; hf < 2; ++hf)
                        for (int k = 0; k < qk && ch * qk + k < hk; ++k) p.seq[n++] = (short)(c + hf * hk + ch * qk + k);
                c += nk;
            } else {
                for (int k = 0; k < lo.Wn / KCH; ++k, ++c) p.seq[n++] = (short)c;                    // W1_j^T
            }
        }
        p.seq_n = n;
    }
    const long long ntiles = ((long long)P + TP - 1) / TP;
    p.Pp = ntiles * TP;
    auto al = [](long long x) { return (x + 63) & ~63LL; };
    long long o = 0;
    p.off_wprep = o; o += al(p.wprep_floats);
    p.off_F = o;  o += al(p.Pp * lo.F);
    p.off_H = o;  o += al(p.Pp * lo.Wn);
    p.off_GH = o; o += al(p.Pp * lo.Wn);
    p.off_GF = o; o += al(p.Pp * lo.F);          // g_feat rows [Pp][F] for the sampler-backward kernel (SAVED path)
    for (int j = 0; j < FDGS_NUM_HEADS; ++j) {
        p.off_RZ[j] = o; if (j < p.nheads) o += al(p.Pp * lo.Wn);
        p.off_GZ[j] = o; if (j < p.nheads) o += al(p.Pp * lo.Wn);
    }
    p.off_wacc = o; o += (long long)WACC_CTAS * 128 * 128;      // per-CTA FP32 sums of the weight kernel (WArgs::priv)
    p.total_floats = o;
    return p;
}

// ---- weight formatting ----------------------------------------------------------------------------------------
// A section is one [N rows][K] operand stored as K/16 chunks; a chunk = hi tile then lo tile, each the canonical
// K-major tile [4 column groups][N][4] (umma.cuh).  element(n, k) = T ? W[k*ld + n] : W[n*ld + k].
struct PrepSec { const float* W; int ld, N, K, T; long long off; int Kvalid; };   // columns k >= Kvalid are zero padding
struct PrepArgs { PrepSec s[2 + 3 * FDGS_NUM_HEADS]; int nsec; };

__global__ void k_deform_wprep_bwd(PrepArgs a, float* __restrict__ out) {
    for (int si = 0; si < a.nsec; ++si) {
        const PrepSec s = a.s[si];
        const int total = s.N * s.K;
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
            int n, k; float w;
            if (s.T) { k = i / s.N; n = i % s.N; w = k < s.Kvalid ? s.W[(size_t)k * s.ld + n] : 0.f; }     // coalesced read either way
            else     { n = i / s.K; k = i % s.K; w = k < s.Kvalid ? s.W[(size_t)n * s.ld + k] : 0.f; }
            float hi, lo;
            split_tf32_rn(w, hi, lo);
            float* c = out + s.off + (long long)(k / KCH) * 32 * s.N;
            const int o = tile_off(n, k % KCH, s.N);
            c[o] = hi;
            c[16 * s.N + o] = lo;
        }
    }
}

struct Net {
    const float* wprep;
    const float* b0;
    const float* b1[FDGS_NUM_HEADS];
    const float* w2[FDGS_NUM_HEADS];       // original [out][Wn]
};

struct IO {
    int P;
    const float* xyz;
    const float* g[FDGS_NUM_HEADS];        // upstream grads [P][out]
    float* d_xyz;
    float *S_F, *S_H, *S_GH;
    float* G_F;                            // dL/dfeat rows [Pp][F] (SAVED path: the sampler backward is its own kernel)
    float* S_RZ[FDGS_NUM_HEADS];           // indexed by plan slot j
    float* S_GZ[FDGS_NUM_HEADS];
    // activations kept by the forward (SAVED kernels read these instead of recomputing), same blocked layout
    const float *V_H;
    const float* V_RZ[FDGS_NUM_HEADS];     // indexed by plan slot j
    const uint32_t* V_M;                   // ReLU gate masks [6: trunk, head id 0..4][Pp][4]
    long long Pp;
};

struct Smem {
    int slots;
    int off_hs_hi, off_hs_lo, off_w2, off_b0, off_b1, off_red, off_bars, off_ring;
    size_t bytes;
};

static inline Smem make_smem(const Layout& lo, bool saved = false) {
    Smem s;
    auto al = [](int x) { return (x + 255) & ~255; };
    const int Wn = lo.Wn;
    int o = 0;
    // relu(h) hi / lo as K-major canonical tiles (feat hi / lo and, later, g_feat alias their heads); when the forward
    // kept the activations only the g_feat staging tile remains
    s.off_hs_hi = o; o += al(128 * (saved ? lo.F : Wn));
    s.off_hs_lo = o; o += saved ? 0 : al(128 * Wn);
    s.off_w2 = o; o += (saved && FDGS_BWD_PIPE) ? al(4 * Wn * 4) : 0;   // W2 of the <= 4-output heads as [head][column][4]
    s.off_b0 = o; o += al(Wn);
    s.off_b1 = o; o += al(FDGS_NUM_HEADS * Wn);
    s.off_red = o; o += al(2 * 128 * 3);
    s.off_bars = o; o += 256;
    s.off_ring = o;
    const int slot_floats = 32 * Wn;
    const int budget = (227 * 1024 - 1024) / 4 - o;
    s.slots = budget / slot_floats;
    if (s.slots > 8) s.slots = 8;
    s.bytes = (size_t)(o + s.slots * slot_floats) * 4 + 1024;
    return s;
}

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// Optional phase timing of CTA 0 / thread 0 (build with -DFDGS_PHASE_TIMING; prints cycle totals at kernel end).
#ifdef FDGS_PHASE_TIMING
#define PT_DECL long long pt_[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; long long pt_t = clock64();
#define PT(i) { const long long n_ = clock64(); pt_[i] += n_ - pt_t; pt_t = n_; }
#define PT_PRINT if (blockIdx.x == 0 && tid == 0) printf("bwd_data phases (cycles, CTA0): sample / chunk0 %lld | wait_trunk / wait_free1 %lld | trunk_epi %lld | wait_z / wait_free0 %lld | head_epi / chunk1 %lld | wait_grh %lld | gh_epi %lld | wait_gf %lld | samp_bwd %lld | tail %lld\n", pt_[0], pt_[1], pt_[2], pt_[3], pt_[4], pt_[5], pt_[6], pt_[7], pt_[8], pt_[9]);
#else
#define PT_DECL
#define PT(i)
#define PT_PRINT
#endif

// TMEM column map (512 columns): accumulator (trunk h, head z, g_feat), g_z / g_h hi, lo, g_relu_h accumulator
constexpr uint32_t TM_ACC = 0, TM_GHI = 128, TM_GLO = 256, TM_GRH = 384;
constexpr int B_WFULL = 0, B_WEMPTY = 8, B_FEATFULL = 16, B_ACCFULL = 17, B_HFULL = 18, B_GZFULL = 19, B_GRHFULL = 20,
              B_GHFULL = 21, B_GFFULL = 22, B_GOFULL = 23, B_GZFULL1 = 24, B_FREE1 = 25, B_COUNT = 26;
// saved mode with FDGS_BWD_PIPE: g_z column chunk c full = B_GZFULL / B_GZFULL1, chunk c read by the tensor core = B_GRHFULL / B_FREE1

__device__ __forceinline__ int w2_row0(int h) { return h == 0 ? 0 : h == 1 ? 3 : h == 2 ? 6 : h == 3 ? 10 : 11; }

// ---- sampler backward for 4 channels [c0, c0+4) of level l of one point -------------------------------------
// df = dL/dfeat of those channels.  Recomputes the six bilinear samples, scatters dL/dplane with vector atomics
// into the channel-last gradient planes / time rows, and accumulates dL/dq (normalised coordinates) into dq[3].
__device__ __forceinline__ void sample_bwd4(const Layout& lo, int l, const float q[4], const float* __restrict__ packed,
                                            const float* __restrict__ trow, float* __restrict__ gpacked,
                                            float* __restrict__ gtrow, int c0, const float (&df)[4], float (&dq)[3]) {
    const int C = lo.C;
    float s[6][4], gx_c[6][4], gy_c[6][4], gmx[6], gmy[6];
    int x0a[6], y0a[6];
    float wx0a[6], wx1a[6], wy0a[6], wy1a[6];
    bool xina[6], yina[6];
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const int a = pa_of(k);
        const int W = lo.pw[l][k];
        const float ix = unnorm_clip(q[a], W, &gmx[k]);
        const float x0f = floorf(ix);
        const int x0 = (int)x0f;
        const float wx0 = (x0f + 1.f) - ix, wx1 = ix - x0f;
        const bool xin = x0 + 1 < W;
        x0a[k] = x0; wx0a[k] = wx0; wx1a[k] = wx1; xina[k] = xin;
        if (k == 2 || k == 4 || k == 5) {
            const float* T = trow + lo.toff[l][k == 2 ? 0 : (k == 4 ? 1 : 2)] + c0;
            const float4 v0 = *reinterpret_cast<const float4*>(T + (size_t)x0 * C);
            const float4 v1 = xin ? *reinterpret_cast<const float4*>(T + (size_t)(x0 + 1) * C) : z4;
            const float a0[4] = {v0.x, v0.y, v0.z, v0.w}, a1[4] = {v1.x, v1.y, v1.z, v1.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) { s[k][c] = a0[c] * wx0 + a1[c] * wx1; gx_c[k][c] = a1[c] - a0[c]; gy_c[k][c] = 0.f; }
            gmy[k] = 0.f; y0a[k] = 0; wy0a[k] = 0.f; wy1a[k] = 0.f; yina[k] = false;
        } else {
            const int b = pb_of(k);
            const int H = lo.ph[l][k];
            const float iy = unnorm_clip(q[b], H, &gmy[k]);
            const float y0f = floorf(iy);
            const int y0 = (int)y0f;
            const float wy0 = (y0f + 1.f) - iy, wy1 = iy - y0f;
            const bool yin = y0 + 1 < H;
            y0a[k] = y0; wy0a[k] = wy0; wy1a[k] = wy1; yina[k] = yin;
            const float* Pp = packed + lo.poff[l][k] + c0;
            const size_t o00 = ((size_t)y0 * W + x0) * C;
            const float4 v00 = *reinterpret_cast<const float4*>(Pp + o00);
            const float4 v01 = xin ? *reinterpret_cast<const float4*>(Pp + o00 + C) : z4;
            const float4 v10 = yin ? *reinterpret_cast<const float4*>(Pp + o00 + (size_t)W * C) : z4;
            const float4 v11 = (xin && yin) ? *reinterpret_cast<const float4*>(Pp + o00 + (size_t)W * C + C) : z4;
            const float a00[4] = {v00.x, v00.y, v00.z, v00.w}, a01[4] = {v01.x, v01.y, v01.z, v01.w};
            const float a10[4] = {v10.x, v10.y, v10.z, v10.w}, a11[4] = {v11.x, v11.y, v11.z, v11.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                s[k][c] = a00[c] * (wx0 * wy0) + a01[c] * (wx1 * wy0) + a10[c] * (wx0 * wy1) + a11[c] * (wx1 * wy1);
                gx_c[k][c] = (a01[c] - a00[c]) * wy0 + (a11[c] - a10[c]) * wy1;
                gy_c[k][c] = (a10[c] - a00[c]) * wx0 + (a11[c] - a01[c]) * wx1;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float ds[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float pr = df[c];
#pragma unroll
            for (int j = 0; j < 6; ++j) if (j != k) pr *= s[j][c];
            ds[c] = pr;
        }
        float gx = 0.f, gy = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) { gx += ds[c] * gx_c[k][c]; gy += ds[c] * gy_c[k][c]; }
        dq[pa_of(k)] += gx * gmx[k];
        if (!(k == 2 || k == 4 || k == 5)) dq[pb_of(k)] += gy * gmy[k];
        const int W = lo.pw[l][k];
        if (k == 2 || k == 4 || k == 5) {
            float* T = gtrow + lo.toff[l][k == 2 ? 0 : (k == 4 ? 1 : 2)] + c0;
            atomicAdd(reinterpret_cast<float4*>(T + (size_t)x0a[k] * C),
                      make_float4(ds[0] * wx0a[k], ds[1] * wx0a[k], ds[2] * wx0a[k], ds[3] * wx0a[k]));
            if (xina[k])
                atomicAdd(reinterpret_cast<float4*>(T + (size_t)(x0a[k] + 1) * C),
                          make_float4(ds[0] * wx1a[k], ds[1] * wx1a[k], ds[2] * wx1a[k], ds[3] * wx1a[k]));
        } else {
            float* Gp = gpacked + lo.poff[l][k] + c0;
            const size_t o00 = ((size_t)y0a[k] * W + x0a[k]) * C;
            const float nw = wx0a[k] * wy0a[k], ne = wx1a[k] * wy0a[k], sw = wx0a[k] * wy1a[k], se = wx1a[k] * wy1a[k];
            atomicAdd(reinterpret_cast<float4*>(Gp + o00), make_float4(ds[0] * nw, ds[1] * nw, ds[2] * nw, ds[3] * nw));
            if (xina[k])
                atomicAdd(reinterpret_cast<float4*>(Gp + o00 + C), make_float4(ds[0] * ne, ds[1] * ne, ds[2] * ne, ds[3] * ne));
            if (yina[k])
                atomicAdd(reinterpret_cast<float4*>(Gp + o00 + (size_t)W * C), make_float4(ds[0] * sw, ds[1] * sw, ds[2] * sw, ds[3] * sw));
            if (xina[k] && yina[k])
                atomicAdd(reinterpret_cast<float4*>(Gp + o00 + (size_t)W * C + C), make_float4(ds[0] * se, ds[1] * se, ds[2] * se, ds[3] * se));
        }
    }
}

// ---- upstream gradient g_o hi/lo -> TMEM (A operand of g_relu_z = g_o W2) ----------------------------------------------
// The thread's g_o columns are [hf*24, hf*24+24) of 48 for dshs, and all 8 (zero padded) on the hf == 0 warps otherwise.
// go_load is issued early (global latency hides behind the barrier waits), go_store writes the split values to TMEM.
__device__ __forceinline__ void go_load(float (&x)[24], int hf, int NO, bool valid, const float* __restrict__ gsrc) {
#pragma unroll
    for (int j = 0; j < 24; ++j) x[j] = 0.f;
    if (!valid) return;
    if (NO == 48) {
#pragma unroll
        for (int c = 0; c < 24; c += 4) {
            const float4 v = *reinterpret_cast<const float4*>(gsrc + hf * 24 + c);
            x[c] = v.x; x[c + 1] = v.y; x[c + 2] = v.z; x[c + 3] = v.w;
        }
    } else if (hf == 0) {
#pragma unroll
        for (int j = 0; j < 4; ++j) if (j < NO) x[j] = gsrc[j];
    }
}
__device__ __forceinline__ void go_store(uint32_t lane_base, int hf, int NO, const float (&x)[24]) {
    if (NO == 48) {
#pragma unroll
        for (int c = 0; c < 24; c += 8) {
            uint32_t hi[8], lw[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { float a, b; split_tf32_rn(x[c + j], a, b); hi[j] = __float_as_uint(a); lw[j] = __float_as_uint(b); }
            tmem_st8(lane_base + TM_GHI + hf * 24 + c, hi);
            tmem_st8(lane_base + TM_GLO + hf * 24 + c, lw);
        }
    } else if (hf == 0) {
        uint32_t hi[8], lw[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { float a, b; split_tf32_rn(x[j], a, b); hi[j] = __float_as_uint(a); lw[j] = __float_as_uint(b); }
        tmem_st8(lane_base + TM_GHI, hi);
        tmem_st8(lane_base + TM_GLO, lw);
    }
    tmem_wait_st();
}

// ---- pass 1 (recompute mode): z = relu(acc + b1) -> scratch, ReLU gate bits ----------------------------------------------
template <int NH>
__device__ __forceinline__ void head_pass1(uint32_t lane_base, int col0, const float* __restrict__ b1s, float* __restrict__ rz_row,
                                           uint32_t (&zgate)[NH / 32 + 1]) {
#pragma unroll
    for (int c0 = 0; c0 < NH; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(lane_base + TM_ACC + col0 + c0, v);
        tmem_wait_ld();
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            float z[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                z[u] = fmaxf(__uint_as_float(v[j + u]) + b1s[col0 + c0 + j + u], 0.f);
                m |= (z[u] > 0.f ? 1u : 0u) << (j + u);
            }
            __stcs(reinterpret_cast<float4*>(rz_row + ((col0 + c0 + j) >> 2) * 128), make_float4(z[0], z[1], z[2], z[3]));
        }
        zgate[c0 / 32] = m;
    }
}

// ---- ReLU gate bits of this thread's NH columns from an activation the forward kept (blocked row pointer) ---------------
template <int NH>
__device__ __forceinline__ void gate_from_saved(const float* __restrict__ row, int col0, uint32_t (&gate)[NH / 32 + 1]) {
#pragma unroll
    for (int c0 = 0; c0 < NH; c0 += 32) {
        float4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __ldcs(reinterpret_cast<const float4*>(row + ((col0 + c0 + 4 * j) >> 2) * 128));
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            m |= ((v[j].x > 0.f ? 1u : 0u) | (v[j].y > 0.f ? 2u : 0u) | (v[j].z > 0.f ? 4u : 0u) | (v[j].w > 0.f ? 8u : 0u)) << (4 * j);
        gate[c0 / 32] = m;
    }
}

// ---- pass 2: g_z = (g_o W2)[z > 0] (accumulator) -> scratch, hi/lo -> TMEM (A operand of g_relu_h += g_z W1) -----------
template <int NH>
__device__ __forceinline__ void head_pass2(uint32_t lane_base, int col0, float* __restrict__ gz_row, const uint32_t (&zgate)[NH / 32 + 1],
                                           int cb = 0, int ce = NH) {          // columns [cb, ce) of this thread's NH
#pragma unroll
    for (int c0 = 0; c0 < NH; c0 += 16) {
        if (c0 < cb || c0 >= ce) continue;
        uint32_t v[16], hi[16], lw[16];
        tmem_ld16(lane_base + TM_ACC + col0 + c0, v);
        tmem_wait_ld();
        const uint32_t m = zgate[c0 / 32] >> (c0 & 31);
        float g[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            g[j] = ((m >> j) & 1u) ? __uint_as_float(v[j]) : 0.f;
            float a, b;
            split_tf32_rn(g[j], a, b);
            hi[j] = __float_as_uint(a); lw[j] = __float_as_uint(b);
        }
#pragma unroll
        for (int j = 0; j < 16; j += 4)
            __stcs(reinterpret_cast<float4*>(gz_row + ((col0 + c0 + j) >> 2) * 128), make_float4(g[j], g[j + 1], g[j + 2], g[j + 3]));
        tmem_st16(lane_base + TM_GHI + col0 + c0, hi);
        tmem_st16(lane_base + TM_GLO + col0 + c0, lw);
    }
    tmem_wait_st();
}

// Register-lean sampler backward for 4 channels of one level of one point (the standalone kernel; sample_bwd4 above needs
// ~170 registers, one 256-thread CTA per SM; measured 280 us; this one 222 us at two CTAs per SM, 244 at three, 309 at four).  Two sweeps over the six planes: the first keeps only the suffix products
// of the samples, the second re-reads the taps (L1 / L2 hits), forms dL/dsample = df * prefix * suffix, scatters and
// accumulates dL/dq.  Same arithmetic as sample_bwd4 up to the association of the 5-factor products.
struct PlaneTaps {
    float s[4], gx[4], gy[4];
    float wx0, wx1, wy0, wy1, gmx, gmy;
    int x0, y0;
    bool xin, yin;
};
template <bool GRAD>
__device__ __forceinline__ void plane_taps(const Layout& lo, int l, int k, const float q[4], const float* __restrict__ packed,
                                           const float* __restrict__ trow, int c0, PlaneTaps& t) {
    const int C = lo.C;
    const int W = lo.pw[l][k];
    const float ix = unnorm_clip(q[pa_of(k)], W, &t.gmx);
    const float x0f = floorf(ix);
    t.x0 = (int)x0f;
    t.wx0 = (x0f + 1.f) - ix; t.wx1 = ix - x0f;
    t.xin = t.x0 + 1 < W;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k == 2 || k == 4 || k == 5) {
        const float* T = trow + lo.toff[l][k == 2 ? 0 : (k == 4 ? 1 : 2)] + c0;
        const float4 v0 = *reinterpret_cast<const float4*>(T + (size_t)t.x0 * C);
        const float4 v1 = t.xin ? *reinterpret_cast<const float4*>(T + (size_t)(t.x0 + 1) * C) : z4;
        const float a0[4] = {v0.x, v0.y, v0.z, v0.w}, a1[4] = {v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            t.s[c] = a0[c] * t.wx0 + a1[c] * t.wx1;
            if (GRAD) { t.gx[c] = a1[c] - a0[c]; t.gy[c] = 0.f; }
        }
        t.gmy = 0.f; t.y0 = 0; t.wy0 = 0.f; t.wy1 = 0.f; t.yin = false;
    } else {
        const int H = lo.ph[l][k];
        const float iy = unnorm_clip(q[pb_of(k)], H, &t.gmy);
        const float y0f = floorf(iy);
        t.y0 = (int)y0f;
        t.wy0 = (y0f + 1.f) - iy; t.wy1 = iy - y0f;
        t.yin = t.y0 + 1 < H;
        const float* Pp = packed + lo.poff[l][k] + c0;
        const size_t o00 = ((size_t)t.y0 * W + t.x0) * C;
        const float4 v00 = *reinterpret_cast<const float4*>(Pp + o00);
        const float4 v01 = t.xin ? *reinterpret_cast<const float4*>(Pp + o00 + C) : z4;
        const float4 v10 = t.yin ? *reinterpret_cast<const float4*>(Pp + o00 + (size_t)W * C) : z4;
        const float4 v11 = (t.xin && t.yin) ? *reinterpret_cast<const float4*>(Pp + o00 + (size_t)W * C + C) : z4;
        const float a00[4] = {v00.x, v00.y, v00.z, v00.w}, a01[4] = {v01.x, v01.y, v01.z, v01.w};
        const float a10[4] = {v10.x, v10.y, v10.z, v10.w}, a11[4] = {v11.x, v11.y, v11.z, v11.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            t.s[c] = a00[c] * (t.wx0 * t.wy0) + a01[c] * (t.wx1 * t.wy0) + a10[c] * (t.wx0 * t.wy1) + a11[c] * (t.wx1 * t.wy1);
            if (GRAD) {
                t.gx[c] = (a01[c] - a00[c]) * t.wy0 + (a11[c] - a10[c]) * t.wy1;
                t.gy[c] = (a10[c] - a00[c]) * t.wx0 + (a11[c] - a01[c]) * t.wx1;
            }
        }
    }
}

__device__ __forceinline__ void sample_bwd4_lean(const Layout& lo, int l, const float q[4], const float* __restrict__ packed,
                                                 const float* __restrict__ trow, float* __restrict__ gpacked,
                                                 float* __restrict__ gtrow, int c0, const float (&df)[4], float (&dq)[3]) {
    const int C = lo.C;
    float suf[6][4];
    {
        float run[4] = {1.f, 1.f, 1.f, 1.f};
#pragma unroll
        for (int k = 5; k >= 0; --k) {
            PlaneTaps t;
            plane_taps<false>(lo, l, k, q, packed, trow, c0, t);
#pragma unroll
            for (int c = 0; c < 4; ++c) { suf[k][c] = run[c]; run[c] *= t.s[c]; }
        }
    }
    float pre[4] = {df[0], df[1], df[2], df[3]};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        PlaneTaps t;
        plane_taps<true>(lo, l, k, q, packed, trow, c0, t);
        float ds[4];
        float gx = 0.f, gy = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            ds[c] = pre[c] * suf[k][c];
            pre[c] *= t.s[c];
            gx += ds[c] * t.gx[c]; gy += ds[c] * t.gy[c];
        }
        dq[pa_of(k)] += gx * t.gmx;
        const int W = lo.pw[l][k];
        if (k == 2 || k == 4 || k == 5) {
            float* T = gtrow + lo.toff[l][k == 2 ? 0 : (k == 4 ? 1 : 2)] + c0;
            atomicAdd(reinterpret_cast<float4*>(T + (size_t)t.x0 * C), make_float4(ds[0] * t.wx0, ds[1] * t.wx0, ds[2] * t.wx0, ds[3] * t.wx0));
            if (t.xin)
                atomicAdd(reinterpret_cast<float4*>(T + (size_t)(t.x0 + 1) * C), make_float4(ds[0] * t.wx1, ds[1] * t.wx1, ds[2] * t.wx1, ds[3] * t.wx1));
        } else {
            dq[pb_of(k)] += gy * t.gmy;
            float* Gp = gpacked + lo.poff[l][k] + c0;
            const size_t o00 = ((size_t)t.y0 * W + t.x0) * C;
            const float nw = t.wx0 * t.wy0, ne = t.wx1 * t.wy0, sw = t.wx0 * t.wy1, se = t.wx1 * t.wy1;
            atomicAdd(reinterpret_cast<float4*>(Gp + o00), make_float4(ds[0] * nw, ds[1] * nw, ds[2] * nw, ds[3] * nw));
            if (t.xin)
                atomicAdd(reinterpret_cast<float4*>(Gp + o00 + C), make_float4(ds[0] * ne, ds[1] * ne, ds[2] * ne, ds[3] * ne));
            if (t.yin)
                atomicAdd(reinterpret_cast<float4*>(Gp + o00 + (size_t)W * C), make_float4(ds[0] * sw, ds[1] * sw, ds[2] * sw, ds[3] * sw));
            if (t.xin && t.yin)
                atomicAdd(reinterpret_cast<float4*>(Gp + o00 + (size_t)W * C + C), make_float4(ds[0] * se, ds[1] * se, ds[2] * se, ds[3] * se));
        }
    }
}

// Sampler backward as its own kernel (SAVED path): LP adjacent lanes per point take consecutive channel quads of its
// dL/dfeat row (one coalesced 128-byte read per point), recompute the six bilinear samples, scatter dL/dplane with float4
// reductions and reduce dL/dxyz over the point's lanes.
__global__ void __launch_bounds__(256, 2)
k_deform_sample_bwd(Layout lo, int P, const float* __restrict__ xyz, const float* __restrict__ GFrows, const float* __restrict__ packed,
                    const float* __restrict__ trow, float* __restrict__ gpacked, float* __restrict__ gtrow, float* __restrict__ d_xyz)
{
    const int F = lo.F, C = lo.C;
    const int items = F >> 2;
    const int LP = (items & 7) == 0 ? 8 : 4, IPT = items / LP, PPW = 32 / LP;
    const int lane = threadIdx.x & 31, sg = lane / LP, sq = lane % LP;
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long sp_ = gw * PPW + sg;
    const bool sval = sp_ < P;
    float q[4] = {0.f, 0.f, 0.f, 0.f};
    if (sval) point_coords(lo, xyz, (int)sp_, 0.f, q);
    float dq[3] = {0.f, 0.f, 0.f};
    for (int it = 0; it < IPT; ++it) {
        const int item = sq + it * LP;
        const int l = (item * 4) / C, c0 = (item * 4) % C;
        if (sval) {
            const float4 d4 = __ldcs(reinterpret_cast<const float4*>(GFrows + (size_t)sp_ * F + item * 4));
            const float df[4] = {d4.x, d4.y, d4.z, d4.w};
            sample_bwd4_lean(lo, l, q, packed, trow, gpacked, gtrow, c0, df, dq);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        dq[a] += __shfl_xor_sync(0xffffffffu, dq[a], 1);
        dq[a] += __shfl_xor_sync(0xffffffffu, dq[a], 2);
        if (LP == 8) dq[a] += __shfl_xor_sync(0xffffffffu, dq[a], 4);
    }
    if (sq == 0 && sval) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float mx = lo.aabb[a], mn = lo.aabb[3 + a];
            d_xyz[3 * (size_t)sp_ + a] = dq[a] * (2.0f / (mn - mx));
        }
    }
}

// =====================================================================================================================
template <int WN, bool SAVED>
__global__ void __launch_bounds__(NTHREADS, 1)
k_deform_bwd_data(Layout lo, Plan plan, Smem sp, Net net, IO io, const float* __restrict__ packed,
                  const float* __restrict__ trow, float* __restrict__ gpacked, float* __restrict__ gtrow, int ntiles)
{
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment by OFFSET from the shared array (not by integer casts of the pointer): the compiler keeps the
    // shared address space and emits LDS / STS with 32-bit addresses instead of generic LD / ST
    float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_addr(smem_raw) & 1023u)) & 1023u));
    float* HS_hi = sm + sp.off_hs_hi;
    float* HS_lo = sm + sp.off_hs_lo;
    float* F_hi = HS_hi;                       // the feat tile lives in the head of the relu(h) buffers
    float* F_lo = HS_lo;
    float* B0s = sm + sp.off_b0;
    float* B1s = sm + sp.off_b1;
    float* red = sm + sp.off_red;
    float* W2s = sm + sp.off_w2;               // saved mode: W2 of the <= 4-output heads, [head][column][4]
    float* ring = sm + sp.off_ring;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + sp.off_bars);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + B_COUNT + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = sp.slots;
    constexpr int NH = WN / 2;
    constexpr int SLOT = 32 * WN;
    const int F = lo.F;

    if (tid == 0) {
        for (int i = 0; i < B_COUNT; ++i) mbar_init(&bars[i], 1);
        fence_mbar_init();
    }
    if (warp == 9) tmem_alloc(tmem_slot, 512);
    if (tid < NEPI) {
        for (int i = tid; i < WN; i += NEPI) B0s[i] = net.b0[i];
    }
    __syncthreads();
    if (tid < NEPI) {
        for (int j = 0; j < plan.nheads; ++j) {
            const int h = plan.heads[j];
            const int NO = head_out(h);
            for (int i = tid; i < WN; i += NEPI) B1s[h * WN + i] = net.b1[h][i];
            if (plan.cc[j])                                  // W2 [NO][WN] -> [column][4] (zero padded)
                for (int e = tid; e < WN * 4; e += NEPI) W2s[h * WN * 4 + e] = (e & 3) < NO ? net.w2[h][(e & 3) * WN + (e >> 2)] : 0.f;
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 8) {
        // ===== producer: the same chunk sequence for every tile ===========================================
        if (lane == 0) {
            uint32_t g = 0;
            const int nch = plan.seq_n + plan.nchunks_gf;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
                for (int c = 0; c < nch; ++c, ++g) {
                    const uint32_t slot = g % S;
                    const bool gf = c >= plan.seq_n;
                    const uint32_t bytes = (gf ? 32 * F : SLOT) * 4;
                    const float* src = net.wprep + (gf ? (size_t)plan.nchunks_main * SLOT + (size_t)(c - plan.seq_n) * 32 * F
                                                       : (size_t)plan.seq[c] * SLOT);
                    mbar_wait(&bars[B_WEMPTY + slot], ((g / S) & 1) ^ 1);
                    mbar_expect_tx(&bars[B_WFULL + slot], bytes);
                    bulk_g2s(ring + (size_t)slot * SLOT, src, bytes, &bars[B_WFULL + slot]);
                }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ===== MMA issuer: the whole warp runs the loop, one elected lane issues (see umma.cuh:elect_one) ============
        {
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
            const uint32_t leader = elect_one();
            const uint32_t idesc = make_idesc_tf32(128, WN, 0, 0);
            const uint32_t idesc_gf = make_idesc_tf32(128, F, 0, 0);
            const uint64_t s_hi0 = desc_kmajor(HS_hi, 128, 0), s_lo0 = desc_kmajor(HS_lo, 128, 0);   // feat aliases relu(h)
            const uint32_t ring_u = smem_addr(ring) >> 4;
            uint32_t slot = 0, wph = 0, ph_feat = 0, ph_h = 0, ph_gz = 0, ph_gz1 = 0, ph_gh = 0, ph_go = 0;
            // D[acc] (+)= A . B^T over `nck` streamed chunks (two 8-wide k-steps each, `ksteps` of them real) of an [N][K]
            // weight.  A: the shared-memory tile (ss) or TMEM columns ta_hi / ta_lo.  fresh: the first MMA overwrites D.
            auto gemm = [&](uint32_t acc, int nck, int N, uint32_t id, bool ss, uint32_t ta_hi, uint32_t ta_lo, int ksteps, bool fresh,
                            int kbase = 0) {       // kbase: first k-step (A operand column / 8) the streamed chunks refer to
                const uint64_t bd0 = make_sdesc(0u, (uint32_t)N * 16u, 128u);
                const uint32_t bstep = (uint32_t)N * 2u;
#pragma unroll 1
                for (int c = 0; c < nck; ++c) {
                    mbar_wait(&bars[B_WFULL + slot], wph);
                    tc_fence_after();
                    if (leader) {
                        const uint64_t dbh0 = bd0 + (ring_u + slot * (uint32_t)(SLOT / 4));
                        const uint64_t dbl0 = dbh0 + (uint32_t)N * 4u;
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            if (c * 2 + ks < ksteps) {
                                const uint32_t kk = (uint32_t)(kbase + c * 2 + ks);
                                const int first = (fresh && c == 0 && ks == 0) ? 0 : 1;
                                const uint64_t dbh = dbh0 + ks * bstep, dbl = dbl0 + ks * bstep;
                                if (ss) {
                                    const uint64_t ah = s_hi0 + kk * 256u, al = s_lo0 + kk * 256u;
                                    mma_tf32_ss(acc, al, dbh, id, first);
                                    mma_tf32_ss(acc, ah, dbl, id, 1);
                                    mma_tf32_ss(acc, ah, dbh, id, 1);
                                } else {
                                    mma_tf32_ts(acc, ta_lo + kk * 8u, dbh, id, first);
                                    mma_tf32_ts(acc, ta_hi + kk * 8u, dbl, id, 1);
                                    mma_tf32_ts(acc, ta_hi + kk * 8u, dbh, id, 1);
                                }
                            }
                        }
                        mma_commit(&bars[B_WEMPTY + slot]);
                    }
                    __syncwarp();
                    if (++slot == (uint32_t)S) { slot = 0; wph ^= 1; }
                }
            };
            auto commit = [&](int b) { if (leader) mma_commit(&bars[b]); __syncwarp(); };
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                if constexpr (!SAVED) {
                    // trunk: ACC = feat . W0^T
                    mbar_wait(&bars[B_FEATFULL], ph_feat); ph_feat ^= 1;
                    tc_fence_after();
                    gemm(tm + TM_ACC, F / KCH, WN, idesc, true, 0, 0, 1 << 20, true);
                    commit(B_ACCFULL);
                    mbar_wait(&bars[B_HFULL], ph_h); ph_h ^= 1;
                    tc_fence_after();
                }
                for (int i = 0; i < plan.nheads; ++i) {
                    if constexpr (!SAVED) {
                        // z_i: ACC = relu(h) . W1^T      (A from shared memory)
                        gemm(tm + TM_ACC, WN / KCH, WN, idesc, true, 0, 0, 1 << 20, true);
                        commit(B_ACCFULL);
                    }
                    // g_relu_z: ACC = g_o . W2       (A = g_o from TMEM, B = W2^T chunks, K padded to 8 / 48)
                    if (!plan.cc[i]) {
                        mbar_wait(&bars[B_GOFULL], ph_go); ph_go ^= 1;
                        tc_fence_after();
                        gemm(tm + TM_ACC, plan.k2chunks[i], WN, idesc, false, tm + TM_GHI, tm + TM_GLO,
                             head_out(plan.heads[i]) > 8 ? 6 : 1, true);
                        commit(B_ACCFULL);
                    }
                    if constexpr (SAVED && FDGS_BWD_PIPE) {
                        // g_relu_h += g_z . W1 in two K chunks (column chunk ch of both column halves), each released to the
                        // epilogue warps for the next head's g_z as soon as it has been read
                        constexpr int HK = WN / KCH / 2, QK = HK / 2;
#pragma unroll
                        for (int ch = 0; ch < 2; ++ch) {
                            mbar_wait(&bars[ch ? B_GZFULL1 : B_GZFULL], ch ? ph_gz1 : ph_gz);
                            if (ch) ph_gz1 ^= 1; else ph_gz ^= 1;
                            tc_fence_after();
                            gemm(tm + TM_GRH, QK, WN, idesc, false, tm + TM_GHI, tm + TM_GLO, 1 << 20, i == 0 && ch == 0, (ch * QK) * 2);
                            gemm(tm + TM_GRH, QK, WN, idesc, false, tm + TM_GHI, tm + TM_GLO, 1 << 20, false, (HK + ch * QK) * 2);
                            commit(ch ? B_FREE1 : B_GRHFULL);
                        }
                        continue;
                    }
                    // g_relu_h += g_z . W1          (A from TMEM, B = W1^T chunks)
                    mbar_wait(&bars[B_GZFULL], ph_gz); ph_gz ^= 1;
                    tc_fence_after();
                    gemm(tm + TM_GRH, WN / KCH, WN, idesc, false, tm + TM_GHI, tm + TM_GLO, 1 << 20, i == 0);
                    if constexpr (SAVED) commit(B_GRHFULL);    // per head: g_z in TMEM may be overwritten by the next g_o
                }
                if constexpr (!SAVED) commit(B_GRHFULL);
                // g_feat: ACC[0, F) = g_h . W0       (A from TMEM, B = W0^T chunks, N = F)
                mbar_wait(&bars[B_GHFULL], ph_gh); ph_gh ^= 1;
                tc_fence_after();
                gemm(tm + TM_ACC, WN / KCH, F, idesc_gf, false, tm + TM_GHI, tm + TM_GLO, 1 << 20, true);
                commit(B_GFFULL);
            }
        }
    } else {
        // ===== sampler + epilogues: thread = point row `row`, column half `hf` ==========================
        const int row = tid & 127, hf = tid >> 7;
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int col0 = hf * NH;
        uint32_t ph_acc = 0, ph_grh = 0, ph_gf = 0, ph_free1 = 0;
        const int C = lo.C, CH = C >> 1;
        // sampler mapping (see below): LP lanes per point, IPT channel quads per lane, PPW points per warp
        const int items = F >> 2;
        const int LP = (items & 7) == 0 ? 8 : 4, IPT = items / LP, PPW = 32 / LP, SPASS = TP / (8 * PPW);
        const int sg = lane / LP, sq = lane % LP;
        PT_DECL
        float* GF = HS_hi;                   // g_feat staging tile [F][128]: every MMA that read relu(h) has completed by then
        int pend_tile = -1, pend_pass = 0;
        // One quarter (8*PPW points) of a tile's sampler backward: plane-gradient scatter + d_xyz.  In SAVED mode these
        // passes are issued between the head steps of the following tile, where the epilogue warps would otherwise
        // wait for the tensor core.
        auto samp_pass = [&](int t, int pass) {
            const int srow = pass * (8 * PPW) + warp * PPW + sg;
            const int sp_ = t * TP + srow;
            const bool sval = sp_ < io.P;
            float q[4] = {0.f, 0.f, 0.f, 0.f};
            if (sval) point_coords(lo, io.xyz, sp_, 0.f, q);
            float dq[3] = {0.f, 0.f, 0.f};
            for (int it = 0; it < IPT; ++it) {
                const int item = sq + it * LP;
                const int l = (item * 4) / C, c0 = (item * 4) % C;
                const float4 d4 = *reinterpret_cast<const float4*>(GF + tile_off(srow, item * 4, 128));
                const float df[4] = {d4.x, d4.y, d4.z, d4.w};
                if (sval) sample_bwd4(lo, l, q, packed, trow, gpacked, gtrow, c0, df, dq);
            }
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                dq[a] += __shfl_xor_sync(0xffffffffu, dq[a], 1);
                dq[a] += __shfl_xor_sync(0xffffffffu, dq[a], 2);
                if (LP == 8) dq[a] += __shfl_xor_sync(0xffffffffu, dq[a], 4);
            }
            if (sq == 0 && sval) {
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const float mx = lo.aabb[a], mn = lo.aabb[3 + a];
                    io.d_xyz[3 * (size_t)sp_ + a] = dq[a] * (2.0f / (mn - mx));
                }
            }
        };
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const int p = tile * TP + row;
            const bool valid = p < io.P;
            uint32_t gate[NH / 32 + 1];
            if constexpr (SAVED) {
                const uint32_t* mp = io.V_M + (size_t)p * 4 + hf * 2;
                gate[0] = __ldcs(mp);
                if (NH > 32) gate[1] = __ldcs(mp + 1);
            } else {
            // ---- HexPlane sampling -> feat tile (hi / lo) + scratch ----------------------------------------
            // Sampler mapping: LP adjacent lanes share one point and take consecutive channel quads, so a warp-wide
            // gather touches 32/LP points x (64-byte taps) instead of 32 unrelated lines (the L1 tag stage, one line
            // per cycle, bounded the old thread-per-point mapping).
            for (int pass = 0; pass < SPASS; ++pass) {
                const int srow = pass * (8 * PPW) + warp * PPW + sg;
                const int sp_ = tile * TP + srow;
                const bool sval = sp_ < io.P;
                float q[4] = {0.f, 0.f, 0.f, 0.f};
                if (sval) point_coords(lo, io.xyz, sp_, 0.f, q);
                for (int it = 0; it < IPT; ++it) {
                    const int item = sq + it * LP;
                    const int l = (item * 4) / C, c0 = (item * 4) % C;
                    float prod[4] = {0.f, 0.f, 0.f, 0.f};
                    if (sval) sample_level<4>(lo, l, q, packed, trow, c0, prod);
                    float4 hi, lw;
                    split_tf32_rn(prod[0], hi.x, lw.x); split_tf32_rn(prod[1], hi.y, lw.y);
                    split_tf32_rn(prod[2], hi.z, lw.z); split_tf32_rn(prod[3], hi.w, lw.w);
                    const int o = tile_off(srow, item * 4, 128);
                    *reinterpret_cast<float4*>(F_hi + o) = hi;
                    *reinterpret_cast<float4*>(F_lo + o) = lw;
                    __stcs(reinterpret_cast<float4*>(io.S_F + (size_t)tile * TP * F + (srow >> 5) * 32 * F + (item * 32 + (srow & 31)) * 4),
                           make_float4(prod[0], prod[1], prod[2], prod[3]));
                }
            }
            fence_async_smem();
            epi_bar();
            if (tid == 0) mbar_arrive(&bars[B_FEATFULL]);
            PT(0)

            // ---- trunk epilogue: relu(h + b0) -> scratch, shared memory hi / lo; gate bits kept in registers ------
            mbar_wait(&bars[B_ACCFULL], ph_acc); ph_acc ^= 1;
            tc_fence_after();
            PT(1)
#pragma unroll
            for (int c0 = 0; c0 < NH; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(lane_base + TM_ACC + col0 + c0, v);
                tmem_wait_ld();
                uint32_t m = 0;
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    float x[4], a[4], b[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        x[u] = fmaxf(__uint_as_float(v[j + u]) + B0s[col0 + c0 + j + u], 0.f);
                        m |= (x[u] > 0.f ? 1u : 0u) << (j + u);
                        split_tf32_rn(x[u], a[u], b[u]);
                    }
                    const int o = tile_off(row, col0 + c0 + j, 128);
                    *reinterpret_cast<float4*>(HS_hi + o) = make_float4(a[0], a[1], a[2], a[3]);
                    *reinterpret_cast<float4*>(HS_lo + o) = make_float4(b[0], b[1], b[2], b[3]);
                    __stcs(reinterpret_cast<float4*>(io.S_H + (size_t)tile * TP * WN + (row >> 5) * 32 * WN + (((col0 + c0 + j) >> 2) * 32 + (row & 31)) * 4),
                           make_float4(x[0], x[1], x[2], x[3]));
                }
                gate[c0 / 32] = m;
            }
            fence_async_smem();
            tc_fence_before();
            epi_bar();
            if (tid == 0) mbar_arrive(&bars[B_HFULL]);
            PT(2)

            }
            // ---- heads ---------------------------------------------------------------------------------------------
            for (int i = 0; i < plan.nheads; ++i) {
                const int h = plan.heads[i];
                const int NO = head_out(h);
                const float* gsrc = io.g[h] + (size_t)(valid ? p : 0) * NO;
                // blocked scratch: [tile][row / 32][col / 4][row % 32][4]  ->  + (col >> 2) * 128
                const size_t rowoff = (size_t)tile * TP * WN + (row >> 5) * 32 * WN + (row & 31) * 4;
                float* gz_row = io.S_GZ[i] + rowoff;
                uint32_t zgate[NH / 32 + 1];
                if constexpr (SAVED && FDGS_BWD_PIPE) {
                    const uint32_t* mp = io.V_M + ((size_t)(1 + h) * io.Pp + p) * 4 + hf * 2;
                    zgate[0] = __ldcs(mp);
                    if (NH > 32) zgate[1] = __ldcs(mp + 1);
                    if (plan.cc[i]) {
                        // ---- <= 4 outputs: g_z = (g_o W2)[z > 0] on the CUDA cores, delivered in two column chunks ---------
                        // Chunk ch is written as soon as the tensor core has read the previous head's chunk ch, and the
                        // g_relu_h GEMM starts on it while the second chunk is still being produced: epilogue and tensor
                        // core overlap instead of taking turns.
                        float go4[4] = {0.f, 0.f, 0.f, 0.f};
                        if (valid) {
#pragma unroll
                            for (int o = 0; o < 4; ++o) if (o < NO) go4[o] = gsrc[o];
                        }
                        constexpr int CW = NH / 2;
                        const float* w2h = W2s + (h * WN + col0) * 4;
#pragma unroll
                        for (int ch = 0; ch < 2; ++ch) {
                            if (ch) { PT(0) }                           // (phase timing: chunk 0 produced)
                            if (i > 0) {                                // the previous head's chunk ch has been read
                                if (ch) { mbar_wait(&bars[B_FREE1], ph_free1); ph_free1 ^= 1; }
                                else { mbar_wait(&bars[B_GRHFULL], ph_grh); ph_grh ^= 1; }
                                tc_fence_after();
                            }
                            if (ch == 0) { PT(3) } else { PT(1) }       // (waits for the tensor core: chunk 0 / chunk 1)
#pragma unroll
                            for (int c0 = ch * CW; c0 < (ch + 1) * CW; c0 += 16) {
                                uint32_t hi[16], lw[16];
                                float g[16];
                                const uint32_t m = zgate[c0 / 32] >> (c0 & 31);
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    const float4 w = *reinterpret_cast<const float4*>(w2h + (c0 + j) * 4);
                                    float acc = go4[0] * w.x;
                                    acc = fmaf(go4[1], w.y, acc); acc = fmaf(go4[2], w.z, acc); acc = fmaf(go4[3], w.w, acc);
                                    g[j] = ((m >> j) & 1u) ? acc : 0.f;
                                    float a, b;
                                    split_tf32_rn(g[j], a, b);
                                    hi[j] = __float_as_uint(a); lw[j] = __float_as_uint(b);
                                }
                                if (gz_row) {
#pragma unroll
                                    for (int j = 0; j < 16; j += 4)
                                        __stcs(reinterpret_cast<float4*>(gz_row + ((col0 + c0 + j) >> 2) * 128), make_float4(g[j], g[j + 1], g[j + 2], g[j + 3]));
                                }
                                tmem_st16(lane_base + TM_GHI + col0 + c0, hi);
                                tmem_st16(lane_base + TM_GLO + col0 + c0, lw);
                            }
                            tmem_wait_st();
                            tc_fence_before();
                            epi_bar();
                            if (tid == 0) mbar_arrive(&bars[ch ? B_GZFULL1 : B_GZFULL]);
                        }
                        PT(4)
                        continue;
                    }
                    // ---- 48 outputs: g_o W2 on the tensor core; needs the whole g_z buffer (g_o is staged in its head) --------
                    float gox[24];
                    go_load(gox, hf, NO, valid, gsrc);
                    if (i > 0) {
                        mbar_wait(&bars[B_GRHFULL], ph_grh); ph_grh ^= 1;
                        mbar_wait(&bars[B_FREE1], ph_free1); ph_free1 ^= 1;
                        tc_fence_after();
                    }
                    PT(3)
                    go_store(lane_base, hf, NO, gox);
                    tc_fence_before();
                    epi_bar();
                    if (tid == 0) mbar_arrive(&bars[B_GOFULL]);
                    mbar_wait(&bars[B_ACCFULL], ph_acc); ph_acc ^= 1;  // accumulator = g_o . W2
                    tc_fence_after();
                    head_pass2<NH>(lane_base, col0, gz_row, zgate, 0, NH / 2);          // chunk 0: the GEMM starts on it
                    tc_fence_before();
                    epi_bar();
                    if (tid == 0) mbar_arrive(&bars[B_GZFULL]);
                    head_pass2<NH>(lane_base, col0, gz_row, zgate, NH / 2, NH);
                    tc_fence_before();
                    epi_bar();
                    if (tid == 0) mbar_arrive(&bars[B_GZFULL1]);
                    PT(4)
                    continue;
                }
                float gox[24];
                go_load(gox, hf, NO, valid, gsrc);
                if constexpr (SAVED) {
                    const uint32_t* mp = io.V_M + ((size_t)(1 + h) * io.Pp + p) * 4 + hf * 2;
                    zgate[0] = __ldcs(mp);
                    if (NH > 32) zgate[1] = __ldcs(mp + 1);
                    if (i > 0) {                                    // g_relu_h += g_z W1 of the previous head has read its g_z
                        mbar_wait(&bars[B_GRHFULL], ph_grh); ph_grh ^= 1;
                        tc_fence_after();
                    }
                    PT(3)
                } else {
                    mbar_wait(&bars[B_ACCFULL], ph_acc); ph_acc ^= 1;
                    tc_fence_after();
                    PT(3)
                    head_pass1<NH>(lane_base, col0, B1s + h * WN, io.S_RZ[i] + rowoff, zgate);
                }
                go_store(lane_base, hf, NO, gox);
                tc_fence_before();
                epi_bar();
                if (tid == 0) mbar_arrive(&bars[B_GOFULL]);        // z drained from the accumulator, g_o staged in TMEM
                mbar_wait(&bars[B_ACCFULL], ph_acc); ph_acc ^= 1;  // accumulator = g_o . W2
                tc_fence_after();
                head_pass2<NH>(lane_base, col0, gz_row, zgate);
                tc_fence_before();
                epi_bar();
                if (tid == 0) mbar_arrive(&bars[B_GZFULL]);
                if (pend_tile >= 0 && pend_pass < SPASS) samp_pass(pend_tile, pend_pass++);     // overlaps g_relu_h += g_z W1
                PT(4)
            }

            // ---- g_h = g_relu_h [h > 0] -> scratch, TMEM hi / lo -----------------------------------------------------
            mbar_wait(&bars[B_GRHFULL], ph_grh); ph_grh ^= 1;
            if constexpr (SAVED && FDGS_BWD_PIPE) { mbar_wait(&bars[B_FREE1], ph_free1); ph_free1 ^= 1; }   // last head's second chunk
            tc_fence_after();
            PT(5)
#pragma unroll
            for (int c0 = 0; c0 < NH; c0 += 16) {
                uint32_t v[16], hi[16], lw[16];
                tmem_ld16(lane_base + TM_GRH + col0 + c0, v);
                tmem_wait_ld();
                const uint32_t m = gate[c0 / 32] >> (c0 & 31);
                float gh[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    gh[j] = ((m >> j) & 1u) ? __uint_as_float(v[j]) : 0.f;
                    float a, b;
                    split_tf32_rn(gh[j], a, b);
                    hi[j] = __float_as_uint(a); lw[j] = __float_as_uint(b);
                }
#pragma unroll
                for (int j = 0; j < 16; j += 4)
                    __stcs(reinterpret_cast<float4*>(io.S_GH + (size_t)tile * TP * WN + (row >> 5) * 32 * WN + (((col0 + c0 + j) >> 2) * 32 + (row & 31)) * 4),
                           make_float4(gh[j], gh[j + 1], gh[j + 2], gh[j + 3]));
                tmem_st16(lane_base + TM_GHI + col0 + c0, hi);
                tmem_st16(lane_base + TM_GLO + col0 + c0, lw);
            }
            tmem_wait_st();
            tc_fence_before();
            epi_bar();
            if (tid == 0) mbar_arrive(&bars[B_GHFULL]);
            PT(6)

            // ---- g_feat -> sampler backward -----------------------------------------------------------------------------
            mbar_wait(&bars[B_GFFULL], ph_gf); ph_gf ^= 1;
            tc_fence_after();
            PT(7)
            if constexpr (SAVED) {
                // g_feat rows go to memory (128 B per point): the sampler backward - float4 reductions into the plane
                // gradients - runs as its own full-occupancy kernel (k_deform_sample_bwd).  Inside this kernel its
                // reductions ran at half the rate a dedicated kernel reaches (8 warps per SM, the LSU shared with the
                // tensor core's operand reads; tools/red_probe.cu) and were 44 % of the kernel.
                for (int l = 0; l < lo.L; ++l)
                    for (int cc = 0; cc < CH; cc += 4) {
                        const int c0 = l * C + hf * CH + cc;
                        uint32_t v[4];
                        tmem_ld4(lane_base + TM_ACC + (uint32_t)c0, v);            // warp-uniform address: all lanes execute
                        tmem_wait_ld();
                        if (valid)
                            __stcs(reinterpret_cast<float4*>(io.G_F + (size_t)p * F + c0),
                                   make_float4(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]), __uint_as_float(v[3])));
                    }
                tc_fence_before();
                epi_bar();
            } else {
            // accumulator rows (TMEM lane = point) -> shared memory, so that the sampler mapping can read any point.
            while (pend_tile >= 0 && pend_pass < SPASS) samp_pass(pend_tile, pend_pass++);
            epi_bar();
            for (int l = 0; l < lo.L; ++l)
                for (int cc = 0; cc < CH; cc += 4) {
                    const int c0 = l * C + hf * CH + cc;
                    uint32_t v[4];
                    tmem_ld4(lane_base + TM_ACC + (uint32_t)c0, v);                // warp-uniform address: all lanes execute
                    tmem_wait_ld();
                    *reinterpret_cast<float4*>(GF + tile_off(row, c0, 128)) =
                        make_float4(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]), __uint_as_float(v[3]));
                }
            tc_fence_before();
            epi_bar();
            pend_tile = tile; pend_pass = 0;
            while (pend_pass < SPASS) samp_pass(pend_tile, pend_pass++);          // the staging buffer aliases relu(h): scatter now
            epi_bar();
            }
            PT(8)
            PT(9)
        }
        while (pend_tile >= 0 && pend_pass < SPASS) samp_pass(pend_tile, pend_pass++);
        PT_PRINT
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 9) tmem_free(tmem, 512);
}

// =====================================================================================================================
// Weight gradients: C[m][n] (+)= sum_p A[p][m] * B[p][n]   (M = 128 zero-padded, N = nb rounded up to 16)
struct WJob {
    const float* A; int lda, ma, rowsA;      // A [rowsA valid rows][lda]; columns >= ma read as 0
    const float* B; int ldb, nb, rowsB;
    float* C; int ldc, transposed;           // transposed: C[n * ldc + m]
    float* biasA;                            // += column sums of A [ma]   (or null)
    float* biasB;                            // += column sums of B [nb]   (or null)
    int slices;                              // CTAs sharing this job
    int blkA, blkB;                          // operand in the blocked scratch layout [row/32][col/4][row%32][4]: the 32 rows
                                             // of a stage are one contiguous run of 32*ld floats (one bulk copy)
    int cuda_core;                           // nb <= 4 (the output layers of the pos / scale / rotation / opacity heads): the
                                             // staging warps accumulate C on the CUDA cores straight from the raw stage —
                                             // 4 nb FMAs per loaded float4 instead of the TF32 split, two tile stores and an
                                             // MMA padded to N = 16 (the job then costs about a third of a tensor-core one)
};
constexpr int MAXJOBS = 2 * FDGS_NUM_HEADS + 1;
struct WArgs {
    WJob j[MAXJOBS]; int njobs; long long rows;    // rows = Pp (multiple of 128)
    // Block-wise accumulation of the tensor-core jobs: the TMEM accumulator adds with truncation, so a CTA that sums its
    // whole slice (586 stages = 18.7 k points at 300 k Gaussians) in it drifts (1.2e-4 of the reference on dW0, measured);
    // every `flush` stages the accumulator is drained into a per-CTA FP32 buffer `priv` ([CTA][n][m], plain read-add-write:
    // only this CTA touches it) and the next block starts fresh in the other of two TMEM accumulators.  0: one block.
    int flush; float* priv;
};

constexpr int WK = 32;             // points per stage
constexpr int TSTAGES = 2;         // formatted (hi/lo tile32) operand stages feeding the tensor core
constexpr int RSTAGES = 3;         // raw FP32 stages filled by bulk async copies
constexpr int NSW = 16;            // staging / epilogue warps
constexpr int W_THREADS = (NSW + 2) * 32;   // + MMA warp + bulk-copy producer warp
constexpr int WB_TFULL = 0, WB_TEMPTY = 2, WB_RFULL = 4, WB_REMPTY = 7, WB_DONE = 10 /* +1: one per TMEM accumulator */, WB_COUNT = 12;

__device__ __forceinline__ float4 load4(const float* __restrict__ base, int ld, long long r, int c, int ncols, long long nrows) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r >= nrows || c >= ncols) return v;
    const float* p = base + (size_t)r * ld + c;
    if ((ld & 3) == 0 && c + 3 < ncols) return *reinterpret_cast<const float4*>(p);
    v.x = p[0];
    if (c + 1 < ncols) v.y = p[1];
    if (c + 2 < ncols) v.z = p[2];
    if (c + 3 < ncols) v.w = p[3];
    return v;
}

// Drains one accumulation block of a tensor-core job (WArgs::flush): TMEM accumulator -> per-CTA FP32 sums (block 0 stores,
// the middle blocks add, the last block adds and sends the sum to the caller's gradient with one reduction per element).
// Every thread meets the same (m, n) elements in every drain, so the per-CTA buffer needs no synchronisation.
// NOT inlined on purpose: inlined into k_deform_bwd_weights its registers and spills changed the allocation of the stage loop,
// which is sensitive to it (535 -> 600-690 us per view with identical loop source, measured).
#ifndef FDGS_W_BLOCKMAP
#define FDGS_W_BLOCKMAP 0   // 1: re-derive the staging threads' row / column mapping per accumulation block (see the stage loop)
#endif
#ifndef FDGS_W_DRAIN
#define FDGS_W_DRAIN 1      // 0: inline, 16-column groups; 1: inline, 8-column groups; 2: the function below; 3: as 1, loop not unrolled
#endif
struct WDrain { uint32_t lane_base; int m, ma, nb, nb0, nb1, transposed, ldc; float* C; float* priv; bool first, last; };
__device__ __noinline__ void w_drain(uint64_t* done_bar, uint32_t parity, WDrain d)
{
    mbar_wait(done_bar, parity);
    tc_fence_after();
    // 16-column TMEM loads (two per drain at N = 128): a tcgen05.ld issued while the tensor core works on the other
    // accumulator waits for the MMAs in flight
    for (int c0 = d.nb0; c0 < d.nb1; c0 += 16) {                    // warp-uniform bounds; columns beyond N16 are never used
        // the running sums first, all 16 loads in flight together (read one by one between the stores below they serialise on
        // the L2 round trip: 23 k cycles per drain, measured)
        float pv[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int n = c0 + j;
            pv[j] = (!d.first && d.m < d.ma && n < d.nb1 && n < d.nb) ? __ldcg(d.priv + n * 128 + d.m) : 0.f;
        }
        uint32_t v[16];
        tmem_ld16(d.lane_base + c0, v);
        tmem_wait_ld();
        if (d.m < d.ma) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int n = c0 + j;
                if (n < d.nb1 && n < d.nb) {
                    const float x = __uint_as_float(v[j]) + pv[j];
                    if (d.last) atomicAdd(d.transposed ? &d.C[(size_t)n * d.ldc + d.m] : &d.C[(size_t)d.m * d.ldc + n], x);
                    else __stcg(d.priv + n * 128 + d.m, x);          // [n][m]: a warp's lanes are consecutive m
                }
            }
        }
    }
    tc_fence_before();                                              // the loads are complete before this warp's next TFULL arrive
}

// Pipeline per CTA:  producer warp: cp.async.bulk of the stage's 32 rows of A (and B) -> raw ring (mbarrier tx bytes)
//                    8 staging warps: raw (or, for row-major g_o, global) -> TF32 hi/lo split -> swizzled tile32 stage
//                    MMA thread: 4 x 3 tcgen05.mma (K = 8 points each) per stage into one TMEM accumulator
__global__ void __launch_bounds__(W_THREADS, 1)
k_deform_bwd_weights(WArgs args)
{
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment by OFFSET from the shared array (not by integer casts of the pointer): the compiler keeps the
    // shared address space and emits LDS / STS with 32-bit addresses instead of generic LD / ST
    float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_addr(smem_raw) & 1023u)) & 1023u));
    __shared__ __align__(8) uint64_t bars[WB_COUNT + 1];
    __shared__ uint32_t tmem_slot;
    __shared__ float csumA[128], csumB[128];
    float* cc_acc = sm;                        // C [128][4] of a CUDA-core job (nb <= 4): the head of the tile-stage area,
                                               // which such a job never uses (static shared memory is at its limit)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef FDGS_PHASE_TIMING
    const long long t_start_ = clock64();
#endif
    int jb = 0, first = 0;
    for (; jb < args.njobs; ++jb) { if ((int)blockIdx.x < first + args.j[jb].slices) break; first += args.j[jb].slices; }
    if (jb >= args.njobs) return;
    const WJob J = args.j[jb];
    const int slice = blockIdx.x - first;
    const long long nst = args.rows / WK;
    const long long s_begin = nst * slice / J.slices, s_end = nst * (slice + 1) / J.slices;
    const int N16 = (J.nb + 15) & ~15;
    const int nblkB = (J.nb + 31) >> 5;
    const int stage_floats = 2 * 4096 + 2 * nblkB * 1024;         // A_hi[4][32][32], A_lo, B_hi[nblkB][32][32], B_lo
    // raw stage: the 32 rows of A (blocked) and of B (blocked, or row-major [32][ldb] for the upstream gradients g_o)
    const int rawA_f = 32 * J.lda, rawB_f = 32 * J.ldb;
    const long long fullB = J.blkB ? nst : (long long)J.rowsB / WK;   // stages whose 32 B rows all exist in memory
    const int raw_floats = (rawA_f + rawB_f + 255) & ~255;
    float* raw0 = sm + (size_t)TSTAGES * stage_floats;

    if (tid == 0) {
        for (int i = 0; i < WB_COUNT; ++i) {
            const bool eight = (i >= WB_TFULL && i < WB_TFULL + TSTAGES) || (i >= WB_REMPTY && i < WB_REMPTY + RSTAGES);
            mbar_init(&bars[i], eight ? NSW : 1);
        }
        fence_mbar_init();
    }
    if (tid < 128) { csumA[tid] = 0.f; csumB[tid] = 0.f; }
    for (int e = tid; e < 128 * 4; e += W_THREADS) cc_acc[e] = 0.f;
    if (warp == NSW) tmem_alloc(&tmem_slot, 256);       // two accumulators of up to 128 columns (see WArgs::flush)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_slot;

    if (warp == NSW + 1) {
        // ===== producer ====================================================================================
        if (lane == 0) {
            uint32_t g = 0;
            constexpr int PFD = 8;                                  // stages prefetched into L2 ahead of the ring
            for (long long s = s_begin; s < s_end && s < s_begin + PFD; ++s) {
                bulk_prefetch_l2(J.A + (size_t)s * rawA_f, (uint32_t)rawA_f * 4);
                if (s < fullB) bulk_prefetch_l2(J.B + (size_t)s * rawB_f, (uint32_t)rawB_f * 4);
            }
            for (long long s = s_begin; s < s_end; ++s, ++g) {
                const uint32_t slot = g % RSTAGES;
                if (s + PFD < s_end) {
                    bulk_prefetch_l2(J.A + (size_t)(s + PFD) * rawA_f, (uint32_t)rawA_f * 4);
                    if (s + PFD < fullB) bulk_prefetch_l2(J.B + (size_t)(s + PFD) * rawB_f, (uint32_t)rawB_f * 4);
                }
                const bool withB = s < fullB;                       // a partial last stage of g_o is read with checked loads
                mbar_wait(&bars[WB_REMPTY + slot], ((g / RSTAGES) & 1) ^ 1);
                mbar_expect_tx(&bars[WB_RFULL + slot], (uint32_t)(rawA_f + (withB ? rawB_f : 0)) * 4);
                float* dst = raw0 + (size_t)slot * raw_floats;
                bulk_g2s(dst, J.A + (size_t)s * rawA_f, (uint32_t)rawA_f * 4, &bars[WB_RFULL + slot]);
                if (withB) bulk_g2s(dst + rawA_f, J.B + (size_t)s * rawB_f, (uint32_t)rawB_f * 4, &bars[WB_RFULL + slot]);
            }
        }
        __syncwarp();
    } else if (warp == NSW) {
        // ===== MMA issuer ==================================================================================
        // Single-thread loop on purpose: this kernel is shared-memory bound, and the elected-lane issuer of the data
        // kernels (12 MMAs back to back) bunches the operand fetches against the staging warps: 690 -> 761 us, measured.
        if (lane == 0 && !J.cuda_core) {
            const uint32_t idesc = make_idesc_tf32(128, N16, 1, 1);
            uint32_t g = 0;
            const uint32_t FL = args.flush > 0 ? (uint32_t)args.flush : 0xffffffffu;   // stages per accumulation block
            uint32_t blk = 0, gi = 0;
            for (long long s = s_begin; s < s_end; ++s, ++g) {
                const uint32_t slot = g % TSTAGES;
                mbar_wait(&bars[WB_TFULL + slot], (g / TSTAGES) & 1);
                tc_fence_after();
                const float* Ah = sm + (size_t)slot * stage_floats;
                const float* Al = Ah + 4096;
                const float* Bh = Al + 4096;
                const float* Bl = Bh + nblkB * 1024;
                const uint32_t tacc = tmem + (blk & 1u) * 128u;     // block b accumulates in accumulator b & 1 (drained by the
                                                                    // staging warps when they have staged block b + 1)
#pragma unroll
                for (int p = 0; p < WK; p += 8) {
                    const int acc = (gi == 0 && p == 0) ? 0 : 1;
                    mma_tf32_ss(tacc, desc_mn32(Al, WK, p, 0), desc_mn32(Bh, WK, p, 0), idesc, acc);
                    mma_tf32_ss(tacc, desc_mn32(Ah, WK, p, 0), desc_mn32(Bl, WK, p, 0), idesc, 1);
                    mma_tf32_ss(tacc, desc_mn32(Ah, WK, p, 0), desc_mn32(Bh, WK, p, 0), idesc, 1);
                }
                mma_commit(&bars[WB_TEMPTY + slot]);
                // block complete: one barrier per accumulator (a single barrier could be lapped - block b + 1 may complete
                // before the staging warps look at block b - and the wait would then never return)
                if (gi == FL - 1 || s == s_end - 1) mma_commit(&bars[WB_DONE + (blk & 1u)]);
                if (++gi == FL) { gi = 0; ++blk; }
            }
        }
        __syncwarp();
    } else {
        // ===== staging =====================================================================================
        // NSW warps; a warp instruction covers 4 column groups x 8 rows; a thread's column group is the same in
        // every pass and stage, so its column sums (bias gradients) stay in registers.
        constexpr int NI = 1024 / (NSW * 32);                        // float4s of A per thread per stage
        const int gprB = nblkB * 8;                                 // float4 groups per staged B row (8, 16 or 32)
        // Bank conflicts: a 128-bit shared-memory access is served one quarter-warp (8 lanes) at a time, so the 8 lanes of a
        // quarter must touch 8 different 16-byte bank groups in BOTH layouts.  Raw stage [col/4][32 rows][4]: the group is
        // row % 8.  Swizzled tile32: the group is ((col/8 % 4) ^ (row % 4)) * 2 + (col/4 % 2).  Lanes 0-3 of a quarter take
        // rows r..r+3 of column group cg, lanes 4-7 rows r+4..r+7 of column group cg ^ 1 (the neighbouring quarter does
        // the opposite): conflict-free on both sides (8 rows of one column group made every tile store 2-way conflicted,
        // 65 M of the kernel's 96 M shared-memory wavefronts, ncu).
        const int sub = lane >> 3, row8 = lane & 7;
        const int u0 = warp * 4 + sub;                              // 0 .. 4*NSW-1
        const int cgA = (u0 & 31) ^ (row8 >> 2);
        int cgB, rowA[NI], rowB[NI];
        bool actB[NI];
#pragma unroll
        for (int i = 0; i < NI; ++i) rowA[i] = ((i * (4 * NSW) + u0) >> 5) * 8 + row8;
        if (J.blkB) {
            cgB = (u0 % gprB) ^ (row8 >> 2);
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                const int u = i * (4 * NSW) + u0;
                actB[i] = u < gprB * 4;
                rowB[i] = ((u / gprB) * 8 + row8) & (WK - 1);
            }
        } else {
            cgB = tid % gprB;
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                const int idx = i * (NSW * 32) + tid;
                actB[i] = idx < gprB * WK;
                rowB[i] = (idx / gprB) & (WK - 1);
            }
        }
        const bool onA = cgA * 4 < J.ma, onB = cgB * 4 < J.nb;
        const bool wantA = J.biasA != nullptr, wantB = J.biasB != nullptr;
        int oA[NI], oB[NI];
#pragma unroll
        for (int i = 0; i < NI; ++i) { oA[i] = tile32_off(rowA[i], cgA * 4, WK); oB[i] = tile32_off(rowB[i], cgB * 4, WK); }
        float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t g = 0;
        if (J.cuda_core) {
            // ---- C[m][n] = sum_p A[p][m] B[p][n] for n < nb <= 4 on the CUDA cores -------------------------------------
            // A thread meets every (row, its 4 columns) of the slice exactly once; threads with column group 0 also
            // take the column sums of B (the bias gradient).
            float acc[4][4];
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int o = 0; o < 4; ++o) acc[c][o] = 0.f;
            for (long long s = s_begin; s < s_end; ++s, ++g) {
                const uint32_t rslot = g % RSTAGES;
                const float* rawA = raw0 + (size_t)rslot * raw_floats;
                const float* rawB = rawA + rawA_f;
                mbar_wait(&bars[WB_RFULL + rslot], (g / RSTAGES) & 1);
#pragma unroll
                for (int i = 0; i < NI; ++i) {
                    const float4 a4 = onA ? *reinterpret_cast<const float4*>(rawA + (cgA * 32 + rowA[i]) * 4) : z4;
                    float b[4] = {0.f, 0.f, 0.f, 0.f};
                    if (s < fullB) {
                        const float* r = rawB + rowA[i] * J.ldb;
#pragma unroll
                        for (int o = 0; o < 4; ++o) if (o < J.nb) b[o] = r[o];
                    } else {
                        const long long row = s * WK + rowA[i];
                        if (row < J.rowsB) {
#pragma unroll
                            for (int o = 0; o < 4; ++o) if (o < J.nb) b[o] = J.B[(size_t)row * J.ldb + o];
                        }
                    }
                    const float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
                    for (int c = 0; c < 4; ++c)
#pragma unroll
                        for (int o = 0; o < 4; ++o) acc[c][o] = fmaf(av[c], b[o], acc[c][o]);
                    if (wantB && cgA == 0) { sb[0] += b[0]; sb[1] += b[1]; sb[2] += b[2]; sb[3] += b[3]; }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[WB_REMPTY + rslot]);
            }
            // rows of a column group are spread over 4 adjacent lanes (lanes 0-3 / 4-7 of a quarter-warp hold different
            // column groups, see above) and over warps w, w + 8
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    float v = acc[c][o];
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    acc[c][o] = v;
                }
            if (wantB) {          // warp-uniform condition: every lane takes part in the full-mask shuffles
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    float v = sb[o];
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    if (cgA == 0 && (row8 & 3) == 0 && o < J.nb) atomicAdd(&csumB[o], v);
                }
            }
            if ((row8 & 3) == 0 && onA) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int o = 0; o < 4; ++o) if (o < J.nb) atomicAdd(&cc_acc[(cgA * 4 + c) * 4 + o], acc[c][o]);
            }
            asm volatile("bar.sync 1, %0;" ::"n"(NSW * 32) : "memory");
            if (s_end > s_begin) {
                for (int e = tid; e < 128 * 4; e += NSW * 32) {
                    const int m = e >> 2, n = e & 3;
                    if (m < J.ma && n < J.nb)
                        atomicAdd(J.transposed ? &J.C[(size_t)n * J.ldc + m] : &J.C[(size_t)m * J.ldc + n], cc_acc[e]);
                }
                if (wantB && tid < J.nb) atomicAdd(&J.biasB[tid], csumB[tid]);
            }
        } else {
        // ---- block-wise accumulation (WArgs::flush): drain a completed block's TMEM accumulator -------------------------------
        const unsigned long long FLs = args.flush > 0 ? (unsigned long long)args.flush : 0xffffffffull;
        const long long len = s_end - s_begin;
        const uint32_t nblk = len > 0 ? (uint32_t)(((unsigned long long)len + FLs - 1) / FLs) : 0u;
        uint32_t next_drain = 0;
        float* priv = args.priv ? args.priv + (size_t)blockIdx.x * (128 * 128) : nullptr;
        // Every thread meets the same (m, n) elements in every drain, so the per-CTA buffer needs no synchronisation: block 0
        // stores, the middle blocks add, the last block adds and sends the sum to the caller's gradient with one reduction.
        auto drain = [&](uint32_t b) {
            constexpr int NGc = NSW / 4;                             // warps per TMEM lane quarter: split the columns
            const int per = ((N16 + NGc - 1) / NGc + 7) & ~7;
            const int nb0 = (warp >> 2) * per, nb1 = nb0 + per < N16 ? nb0 + per : N16;
#if FDGS_W_DRAIN == 2
            WDrain d;
            d.lane_base = tmem + (b & 1u) * 128u + ((uint32_t)((warp & 3) * 32) << 16);
            d.m = (warp & 3) * 32 + lane; d.ma = J.ma; d.nb = J.nb; d.nb0 = nb0; d.nb1 = nb1;
            d.transposed = J.transposed; d.ldc = J.ldc; d.C = J.C; d.priv = priv;
            d.first = b == 0; d.last = b + 1 == nblk;
            w_drain(&bars[WB_DONE + (b & 1u)], (b >> 1) & 1u, d);    // the (b / 2)-th completion of accumulator b & 1
#else
            mbar_wait(&bars[WB_DONE + (b & 1u)], (b >> 1) & 1u);     // the (b / 2)-th completion of accumulator b & 1
            tc_fence_after();
            const bool first = b == 0, last = b + 1 == nblk;
            const int m = (warp & 3) * 32 + lane;
            const uint32_t lane_base = tmem + (b & 1u) * 128u + ((uint32_t)((warp & 3) * 32) << 16);
            constexpr int GW = (FDGS_W_DRAIN == 0) ? 16 : 8;         // columns per TMEM load
#if FDGS_W_DRAIN == 3
#pragma unroll 1
#endif
            for (int c0 = nb0; c0 < nb1; c0 += GW) {                // warp-uniform bounds; columns beyond N16 are never used
                // the running sums first, all GW loads in flight together (read one by one between the stores below they
                // serialise on the L2 round trip - the compiler cannot move a load above a store to the same array: 23 k
                // cycles per drain, measured)
                float pv[GW];
#pragma unroll
                for (int j = 0; j < GW; ++j) {
                    const int n = c0 + j;
                    pv[j] = (!first && m < J.ma && n < nb1 && n < J.nb) ? __ldcg(priv + n * 128 + m) : 0.f;
                }
                uint32_t v[GW];
                if (GW == 16) tmem_ld16(lane_base + c0, reinterpret_cast<uint32_t(&)[16]>(v[0]));
                else tmem_ld8(lane_base + c0, reinterpret_cast<uint32_t(&)[8]>(v[0]));
                tmem_wait_ld();
                if (m < J.ma) {
#pragma unroll
                    for (int j = 0; j < GW; ++j) {
                        const int n = c0 + j;
                        if (n < nb1 && n < J.nb) {
                            const float x = __uint_as_float(v[j]) + pv[j];
                            if (last) atomicAdd(J.transposed ? &J.C[(size_t)n * J.ldc + m] : &J.C[(size_t)m * J.ldc + n], x);
                            else __stcg(priv + n * 128 + m, x);      // [n][m]: a warp's lanes are consecutive m
                        }
                    }
                }
            }
            tc_fence_before();                                      // the loads are complete before this warp's next TFULL arrive
#endif
        };
        // Outer loop over accumulation blocks, so that the drain code stays out of the stage loop (inlined there it cost the
        // loop 24 % - 1390 -> 1717 cycles per stage).  Block b - 2 is drained before block b is staged: the tensor core
        // finished it long ago (no wait), and its accumulator is not written again before the first stage of block b, whose
        // tiles this warp has not delivered yet.
        for (long long blk_s = s_begin; blk_s < s_end; blk_s += (long long)FLs) {
        const long long blk_e = blk_s + (long long)FLs < s_end ? blk_s + (long long)FLs : s_end;
        if (blk_s > s_begin + (long long)FLs) drain(next_drain++);
#if FDGS_W_BLOCKMAP
        // The thread's stage mapping is derived again for every block (behind an opaque copy of its inputs, so that the compiler
        // does not hoist it): kept alive across the drain above, it takes the registers the stage loop needs for its invariants
        // (365 instead of 311 instructions per stage, SASS; 338 this way).
        int u0b = u0, row8b = row8, tidb = tid;
        asm volatile("" : "+r"(u0b), "+r"(row8b), "+r"(tidb));
        const int cgA_b = (u0b & 31) ^ (row8b >> 2);
        int cgB_b, rowA_b[NI], rowB_b[NI], oA_b[NI], oB_b[NI];
        bool actB_b[NI];
#pragma unroll
        for (int i = 0; i < NI; ++i) rowA_b[i] = ((i * (4 * NSW) + u0b) >> 5) * 8 + row8b;
        if (J.blkB) {
            cgB_b = (u0b % gprB) ^ (row8b >> 2);
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                const int u = i * (4 * NSW) + u0b;
                actB_b[i] = u < gprB * 4;
                rowB_b[i] = ((u / gprB) * 8 + row8b) & (WK - 1);
            }
        } else {
            cgB_b = tidb % gprB;
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                const int idx = i * (NSW * 32) + tidb;
                actB_b[i] = idx < gprB * WK;
                rowB_b[i] = (idx / gprB) & (WK - 1);
            }
        }
        const bool onA_b = cgA_b * 4 < J.ma, onB_b = cgB_b * 4 < J.nb;
#pragma unroll
        for (int i = 0; i < NI; ++i) { oA_b[i] = tile32_off(rowA_b[i], cgA_b * 4, WK); oB_b[i] = tile32_off(rowB_b[i], cgB_b * 4, WK); }
#else
        const int cgA_b = cgA, cgB_b = cgB;
        const bool onA_b = onA, onB_b = onB;
        const int (&rowA_b)[NI] = rowA; const int (&rowB_b)[NI] = rowB; const int (&oA_b)[NI] = oA; const int (&oB_b)[NI] = oB;
        const bool (&actB_b)[NI] = actB;
#endif
        for (long long s = blk_s; s < blk_e; ++s, ++g) {
            const uint32_t rslot = g % RSTAGES, tslot = g % TSTAGES;
            const float* rawA = raw0 + (size_t)rslot * raw_floats;
            const float* rawB = rawA + rawA_f;
            float4 va[NI], vb[NI];
            mbar_wait(&bars[WB_RFULL + rslot], (g / RSTAGES) & 1);
#pragma unroll
            for (int i = 0; i < NI; ++i) va[i] = onA_b ? *reinterpret_cast<const float4*>(rawA + (cgA_b * 32 + rowA_b[i]) * 4) : z4;
            if (J.blkB) {
#pragma unroll
                for (int i = 0; i < NI; ++i) vb[i] = (actB_b[i] && onB_b) ? *reinterpret_cast<const float4*>(rawB + (cgB_b * 32 + rowB_b[i]) * 4) : z4;
            } else if (s < fullB) {                                  // row-major [32][ldb] rows of g_o in the raw stage
#pragma unroll
                for (int i = 0; i < NI; ++i) {
                    vb[i] = z4;
                    if (actB_b[i] && onB_b) {
                        const float* r = rawB + rowB_b[i] * J.ldb + cgB_b * 4;
                        const int c = cgB_b * 4;
                        vb[i].x = r[0];
                        if (c + 1 < J.nb) vb[i].y = r[1];
                        if (c + 2 < J.nb) vb[i].z = r[2];
                        if (c + 3 < J.nb) vb[i].w = r[3];
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < NI; ++i) vb[i] = actB_b[i] ? load4(J.B, J.ldb, s * WK + rowB_b[i], cgB_b * 4, J.nb, J.rowsB) : z4;
            }
            float4 ahi[NI], alo[NI], bhi[NI], blo[NI];
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                split_tf32_rn(va[i].x, ahi[i].x, alo[i].x); split_tf32_rn(va[i].y, ahi[i].y, alo[i].y);
                split_tf32_rn(va[i].z, ahi[i].z, alo[i].z); split_tf32_rn(va[i].w, ahi[i].w, alo[i].w);
                if (wantA) { sa[0] += va[i].x; sa[1] += va[i].y; sa[2] += va[i].z; sa[3] += va[i].w; }
                split_tf32_rn(vb[i].x, bhi[i].x, blo[i].x); split_tf32_rn(vb[i].y, bhi[i].y, blo[i].y);
                split_tf32_rn(vb[i].z, bhi[i].z, blo[i].z); split_tf32_rn(vb[i].w, bhi[i].w, blo[i].w);
                if (wantB) { sb[0] += vb[i].x; sb[1] += vb[i].y; sb[2] += vb[i].z; sb[3] += vb[i].w; }
            }
            __syncwarp();                                            // every lane has consumed its raw values
            if (lane == 0) mbar_arrive(&bars[WB_REMPTY + rslot]);
            mbar_wait(&bars[WB_TEMPTY + tslot], ((g / TSTAGES) & 1) ^ 1);
            float* Ah = sm + (size_t)tslot * stage_floats;
            float* Bh = Ah + 8192;
            const int lofsB = nblkB * 1024;
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                *reinterpret_cast<float4*>(Ah + oA_b[i]) = ahi[i];
                *reinterpret_cast<float4*>(Ah + 4096 + oA_b[i]) = alo[i];
                if (actB_b[i]) {
                    *reinterpret_cast<float4*>(Bh + oB_b[i]) = bhi[i];
                    *reinterpret_cast<float4*>(Bh + lofsB + oB_b[i]) = blo[i];
                }
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars[WB_TFULL + tslot]);
        }
        }   // accumulation blocks
        // ---- bias gradients ------------------------------------------------------------------------------------
        if (wantA && onA) {
#pragma unroll
            for (int u = 0; u < 4; ++u) atomicAdd(&csumA[cgA * 4 + u], sa[u]);
        }
        if (wantB && onB) {
#pragma unroll
            for (int u = 0; u < 4; ++u) atomicAdd(&csumB[cgB * 4 + u], sb[u]);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(NSW * 32) : "memory");
        if (tid < 128) {
            if (wantA && tid < J.ma) atomicAdd(&J.biasA[tid], csumA[tid]);
            if (wantB && tid < J.nb) atomicAdd(&J.biasB[tid], csumB[tid]);
        }
        // ---- remaining accumulator blocks -> per-CTA sums -> global (+=) -------------------------------------------------
        while (next_drain < nblk) drain(next_drain++);
        }   // tensor-core path
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == NSW) tmem_free(tmem, 256);
#ifdef FDGS_PHASE_TIMING
    if (tid == 0) printf("bwd_weights cta %d job %d slice %d/%d stages %lld cycles %lld\n", (int)blockIdx.x, jb, slice, J.slices, s_end - s_begin, clock64() - t_start_);
#endif
}

// =====================================================================================================================
static int sm_count() {          // of the CURRENT device (one process may drive several)
    static int n[64] = {0};
    int dev = 0; cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!n[dev]) { cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev); if (n[dev] <= 0) n[dev] = 148; }
    const int lim = fdgs_sm_limit_value();        // fdgs_set_sm_limit: leave SMs to a concurrent kernel (the NCCL all-reduce)
    return (lim > 0 && lim < n[dev]) ? lim : n[dev];
}

bool supported(const Layout& lo) { return tc_supported(lo); }

size_t scratch_bytes(const Layout& lo, int P) {
    if (!tc_supported(lo) || P <= 0) return 0;
    bool all[FDGS_NUM_HEADS] = {true, true, true, true, true};
    const Plan p = make_plan(lo, P, all);
    return fdgs_align_up(sizeof(float) * (size_t)p.total_floats);
}

// gpacked / gtrow: zeroed channel-last plane-gradient scratch (as in the CUDA-core path).
int backward(const Layout& lo, const FdgsDeformBackward* a, const float* packed, const float* trow, float* gpacked, float* gtrow,
             cudaStream_t stream) {
    const float* gsrc[FDGS_NUM_HEADS] = {a->g_xyz, a->g_scales, a->g_rotations, a->g_opacity, a->g_shs};
    bool has[FDGS_NUM_HEADS];
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) has[h] = gsrc[h] != nullptr;
    const bool saved = a->saved_ws != nullptr;
    const Plan plan = make_plan(lo, a->P, has, saved);
    if (plan.nheads == 0) {
        FDGS_CUDA(cudaMemsetAsync(a->d_xyz_grid, 0, sizeof(float) * 3 * (size_t)a->P, stream));
        return FDGS_OK;
    }
    if (!a->scratch || a->scratch_bytes < sizeof(float) * (size_t)plan.total_floats) return FDGS_E_WORKSPACE;
    const Smem sp = make_smem(lo, saved);
    if (sp.slots < 3) return FDGS_E_SMEM;
    float* S = (float*)a->scratch;
    const SaveLayout sl = make_save_layout(lo, a->P);
    if (saved && a->saved_bytes < sizeof(float) * (size_t)sl.total) return FDGS_E_WORKSPACE;
    const float* V = (const float*)a->saved_ws;
    const int Wn = lo.Wn, F = lo.F;

    // 1. weights -> hi/lo K-major chunks in consumption order: W0 | per head (W1, W1^T) | W0^T
    PrepArgs pa;
    int ns = 0;
    long long off = 0;
    pa.s[ns++] = PrepSec{a->net.trunk_w[0], F, Wn, F, 0, off, F}; off += (long long)(F / KCH) * 32 * Wn;
    for (int j = 0; j < plan.nheads; ++j) {
        const float* w1 = a->net.head_w1[plan.heads[j]];
        pa.s[ns++] = PrepSec{w1, Wn, Wn, Wn, 0, off, Wn}; off += (long long)(Wn / KCH) * 32 * Wn;
        const int NO = head_out(plan.heads[j]);                                                       // (n = z, k = o) = W2[o][z]
        pa.s[ns++] = PrepSec{a->net.head_w2[plan.heads[j]], Wn, Wn, plan.k2chunks[j] * KCH, 1, off, NO};
        off += (long long)plan.k2chunks[j] * 32 * Wn;
        pa.s[ns++] = PrepSec{w1, Wn, Wn, Wn, 1, off, Wn}; off += (long long)(Wn / KCH) * 32 * Wn;  // (n = h, k = z) = W1[z][h]
    }
    pa.s[ns++] = PrepSec{a->net.trunk_w[0], F, F, Wn, 1, off, Wn}; off += (long long)(Wn / KCH) * 32 * F;  // (n = f, k = h) = W0[h][f]
    pa.nsec = ns;
    if (off != plan.wprep_floats) return FDGS_E_BADARG;
    { FDGS_PROF("k_deform_wprep_bwd", stream);
    k_deform_wprep_bwd<<<96, 256, 0, stream>>>(pa, S + plan.off_wprep);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);

    // 2. data gradients
    Net net;
    net.wprep = S + plan.off_wprep; net.b0 = a->net.trunk_b[0];
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) { net.b1[h] = a->net.head_b1[h]; net.w2[h] = a->net.head_w2[h]; }
    IO io;
    io.P = a->P; io.xyz = a->xyz; io.d_xyz = a->d_xyz_grid;
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) io.g[h] = gsrc[h];
    io.S_F = S + plan.off_F; io.S_H = S + plan.off_H; io.S_GH = S + plan.off_GH; io.G_F = S + plan.off_GF;
    for (int j = 0; j < FDGS_NUM_HEADS; ++j) { io.S_RZ[j] = S + plan.off_RZ[j]; io.S_GZ[j] = S + plan.off_GZ[j]; }
    io.V_H = nullptr; io.V_M = nullptr; io.Pp = plan.Pp;
    for (int j = 0; j < FDGS_NUM_HEADS; ++j) io.V_RZ[j] = nullptr;
    if (saved) {        // the forward kept feat, relu(h), relu(z_j): nothing is recomputed and the weight kernel reads them in place
        io.V_H = V + sl.off_H;
        io.V_M = reinterpret_cast<const uint32_t*>(V + sl.off_M);
        io.S_F = const_cast<float*>(V + sl.off_F); io.S_H = const_cast<float*>(V + sl.off_H);
        for (int j = 0; j < plan.nheads; ++j) { io.V_RZ[j] = V + sl.off_RZ[plan.heads[j]]; io.S_RZ[j] = const_cast<float*>(io.V_RZ[j]); }
    }
    const int ntiles = (a->P + TP - 1) / TP;
    const int grid = ntiles < sm_count() ? ntiles : sm_count();
#define FDGS_LAUNCH_BWD_DATA(WN_, SV_)                                                                                          \
    do {                                                                                                                         \
        FDGS_CUDA(cudaFuncSetAttribute(k_deform_bwd_data<WN_, SV_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp.bytes)); \
        FDGS_PROF(SV_ ? "k_deform_bwd_data_saved" : "k_deform_bwd_data", stream);                                                \
        k_deform_bwd_data<WN_, SV_><<<grid, NTHREADS, sp.bytes, stream>>>(lo, plan, sp, net, io, packed, trow, gpacked, gtrow, ntiles); \
    } while (0)
    if (Wn == 128) { if (saved) FDGS_LAUNCH_BWD_DATA(128, true); else FDGS_LAUNCH_BWD_DATA(128, false); }
    else           { if (saved) FDGS_LAUNCH_BWD_DATA(64, true);  else FDGS_LAUNCH_BWD_DATA(64, false); }
#undef FDGS_LAUNCH_BWD_DATA
    FDGS_LAUNCH_CHECK(a->debug, stream);
    if (saved) {
        const int lp = ((F >> 2) & 7) == 0 ? 8 : 4;
        const long long warps = ((long long)a->P + (32 / lp) - 1) / (32 / lp);
        FDGS_PROF("k_deform_sample_bwd", stream);
        k_deform_sample_bwd<<<(unsigned)((warps + 7) / 8), 256, 0, stream>>>(lo, a->P, a->xyz, io.G_F, packed, trow, gpacked, gtrow, a->d_xyz_grid);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);

    // 3. weight gradients: jobs x point slices, one CTA each
    WArgs wa;
    int nj = 0;
    float* dw1[FDGS_NUM_HEADS] = {a->d_head_w1[0], a->d_head_w1[1], a->d_head_w1[2], a->d_head_w1[3], a->d_head_w1[4]};
    double cost[MAXJOBS];
    // per-stage cost of a CUDA-core dW2 job relative to a tensor-core dW1 job (FDGS_DW2_COST overrides, for tuning)
    // measured (cycles per 32-point stage, FDGS_PHASE_TIMING): dW1 1718, dW2 on the CUDA cores 889 (HBM-bound: ~1000 when all four
    // such jobs stream at once), dW2 (shs) 1543, dW0 1455
    double cc_cost = getenv("FDGS_DW2_COST") ? atof(getenv("FDGS_DW2_COST")) : 0.65, shs_cost = 0.90, dw0_cost = 0.85;
    if (const char* e = getenv("FDGS_W_COSTS")) {            // "cc,shs,dw0" relative to a dW1 job (tuning)
        double a_ = 0, b_ = 0, c_ = 0;
        if (sscanf(e, "%lf,%lf,%lf", &a_, &b_, &c_) == 3 && a_ > 0 && b_ > 0 && c_ > 0) { cc_cost = a_; shs_cost = b_; dw0_cost = c_; }
    }
    for (int j = 0; j < plan.nheads; ++j) {
        const int h = plan.heads[j];
        const int NO = head_out(h);
        wa.j[nj] = WJob{io.S_GZ[j], Wn, Wn, (int)plan.Pp, io.S_H, Wn, Wn, (int)plan.Pp, dw1[h], Wn, 0, a->d_head_b1[h], nullptr, 1, 1, 1, 0};
        cost[nj++] = 1.0;                                 // relative per-stage cost, measured (cycles/stage 1750 : 2050 : 1530)
        const int cc = (NO <= 4 && !getenv("FDGS_DW2_TENSOR")) ? 1 : 0;
        wa.j[nj] = WJob{io.S_RZ[j], Wn, Wn, (int)plan.Pp, gsrc[h], NO, NO, a->P, a->d_head_w2[h], Wn, 1, nullptr, a->d_head_b2[h], 1, 1, 0, cc};
        cost[nj++] = cc ? cc_cost : shs_cost;
    }
    wa.j[nj] = WJob{io.S_GH, Wn, Wn, (int)plan.Pp, io.S_F, F, F, (int)plan.Pp, a->d_trunk_w[0], F, 0, a->d_trunk_b[0], nullptr, 1, 1, 1, 0};
    cost[nj++] = dw0_cost;
    wa.njobs = nj; wa.rows = plan.Pp;
    // distribute the SMs over the jobs in proportion to the bytes they stream
    double tot = 0; for (int i = 0; i < nj; ++i) tot += cost[i];
    const long long nst = plan.Pp / WK;
    int used = 0;
    for (int i = 0; i < nj; ++i) {
        int s = (int)(cost[i] / tot * sm_count());
        if (s < 1) s = 1;
        if ((long long)s > nst) s = (int)nst;
        wa.j[i].slices = s; used += s;
    }
    for (int i = 0; used < sm_count() && i < nj; i += 2) if (wa.j[i].slices < nst) { wa.j[i].slices++; used++; }
    // accumulation block of the tensor-core jobs (stages of 32 points); FDGS_W_FLUSH=0: one block per CTA (round-1 behaviour)
    wa.flush = getenv("FDGS_W_FLUSH") ? atoi(getenv("FDGS_W_FLUSH")) : 32;
    wa.priv = S + plan.off_wacc;
    if (used > WACC_CTAS || (wa.flush > 0 && wa.flush < 8)) wa.flush = 0;
    const int nblk_max = (Wn + 31) / 32;
    const size_t wsm = ((size_t)TSTAGES * (2 * 4096 + 2 * nblk_max * 1024) + (size_t)RSTAGES * (64 * Wn)) * 4 + 1024;
    FDGS_CUDA(cudaFuncSetAttribute(k_deform_bwd_weights, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm));
    { FDGS_PROF("k_deform_bwd_weights", stream);
    k_deform_bwd_weights<<<used, W_THREADS, wsm, stream>>>(wa);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);
    return FDGS_OK;
}

}  // namespace fdgs_deform_tc_bwd
