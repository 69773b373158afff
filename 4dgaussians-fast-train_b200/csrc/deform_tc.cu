// Tensor-core (tcgen05 + TMEM) implementation of the fused HexPlane sampler + deformation MLP for sm_100a.
//
// One persistent CTA per SM, 128 points per tile, warp-specialised:
//   warps 0-7  sampler + epilogues (thread = point row; TMEM lane = row; two column halves)
//   warp  8    producer: streams pre-formatted weight chunks global(L2) -> shared with 1-D bulk async copies
//   warp  9    MMA issuer: one thread issues every tcgen05.mma and commits to mbarriers
//
// GEMMs run as 3xTF32 (hi*hi + lo*hi + hi*lo, FP32 accumulate in TMEM) so that results match the reference's
// FP32 cuBLAS/ATen path to ~1e-6 (BASELINE north_star: images within 1e-5); precision==1 runs one pass.
//   trunk   h = feat W0^T          A = feat tile in shared memory (K-major), B = W0 chunks, D -> TMEM ACC0
//   relu(h) split hi/lo is written back to TMEM and is the A operand (from TMEM) of every head's first GEMM
//   head j  z = relu(h) W1_j^T     D -> TMEM ACC[(j+1)&1]  (double buffered: MMA of head j+1 overlaps epilogue j)
//           o = relu(z) W2_j^T     out <= 48 columns: done on the CUDA cores inside the epilogue
// Replaces scene/deformation.py:45-148 + scene/hexplane.py:73-183 (reference: ~200 ATen launches per view).
#include "deform_common.cuh"
#include "umma.cuh"
#include "prof.cuh"

namespace fdgs_deform_tc {
using namespace fdgs_deform;
using namespace fdgs_umma;

constexpr int KCH = 16;            // K columns per streamed weight chunk
constexpr int NTHREADS = 320;      // 8 epilogue warps + producer + MMA
constexpr int NEPI = 256;

// Pre-formatted weights ("wprep" workspace), floats:
//   a chunk holds KCH=16 reduction columns of an [N rows][K] weight as a hi tile then a lo tile, each the canonical
//   K-major tile [4 column groups][N][4]  (see umma.cuh) => 32*N floats, one contiguous bulk copy.
//   forward section : W0 (N = Wn, K = F) chunks, then for each active head W1_j (N = Wn, K = Wn) chunks.
struct TcPlan {
    int Wn, F, nheads;
    int heads[FDGS_NUM_HEADS];
    int fwd_chunks;                // per tile
    long long fwd_floats;
};

static inline TcPlan make_plan(const Layout& lo) {
    TcPlan p;
    p.Wn = lo.Wn; p.F = lo.F; p.nheads = 0;
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) if ((lo.head_mask >> h) & 1) p.heads[p.nheads++] = h;
    p.fwd_chunks = lo.F / KCH + p.nheads * (lo.Wn / KCH);
    p.fwd_floats = (long long)p.fwd_chunks * 32 * lo.Wn;
    return p;
}

static inline bool tc_supported(const Layout& lo) {
    return lo.depth == 1 && (lo.Wn == 64 || lo.Wn == 128) && lo.F % KCH == 0 && lo.F >= KCH && lo.F <= 64 && lo.C >= 8;
}

struct PrepArgs {
    const float* w0;                       // [Wn][F]
    const float* w1[FDGS_NUM_HEADS];       // [Wn][Wn]
    int heads[FDGS_NUM_HEADS];
    int nheads, Wn, F;
};

// one thread per weight element: split (round-to-nearest) and scatter into its chunk
__global__ void k_deform_wprep(PrepArgs a, float* __restrict__ out) {
    const int Wn = a.Wn, F = a.F;
    const int n0 = Wn * F;
    const int total = n0 + a.nheads * Wn * Wn;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const float* W; int K, idx; long long chunk0;
        if (i < n0) { W = a.w0; K = F; idx = i; chunk0 = 0; }
        else { const int j = (i - n0) / (Wn * Wn); W = a.w1[a.heads[j]]; K = Wn; idx = (i - n0) - j * Wn * Wn; chunk0 = F / KCH + (long long)j * (Wn / KCH); }
        const int n = idx / K, k = idx % K;
        float hi, lo;
        split_tf32_rn(W[idx], hi, lo);
        float* c = out + (chunk0 + k / KCH) * 32 * Wn;
        const int o = tile_off(n, k % KCH, Wn);
        c[o] = hi;
        c[16 * Wn + o] = lo;
    }
}

struct TcNet {
    const float* wprep;
    const float* b0;
    const float* b1[FDGS_NUM_HEADS];
    const float* w2[FDGS_NUM_HEADS];       // original [out][Wn]
    const float* b2[FDGS_NUM_HEADS];
};

struct TcFwdIO {
    int P, sh_coeffs;
    const float *xyz, *scales, *rot, *opacity, *shs;
    float *o_xyz, *o_scales, *o_rot, *o_opacity, *o_shs;
    // training: activations kept for the backward pass (blocked layout [tile][row/32][col/4][row%32][4]); null = none
    float *S_F, *S_H;
    float* S_RZ[FDGS_NUM_HEADS];     // by head id
};

// shared-memory plan (floats unless noted), all offsets multiples of 256 floats
struct SmemPlan {
    int slots;
    int off_ring, off_fhi, off_flo, off_w2, off_b0, off_b1, off_b2, off_osum, off_bars;
    size_t bytes;
};

static inline SmemPlan make_smem_plan(const Layout& lo) {
    SmemPlan s;
    auto al = [](int x) { return (x + 255) & ~255; };
    const int Wn = lo.Wn, F = lo.F;
    int o = 0;
    s.off_fhi = o; o += al(F * 128);
    s.off_flo = o; o += al(F * 128);
    s.off_w2 = o; o += al(59 * Wn);
    s.off_b0 = o; o += al(Wn);
    s.off_b1 = o; o += al(FDGS_NUM_HEADS * Wn);
    s.off_b2 = o; o += al(64);
    s.off_osum = o; o += al(48 * 128);
    s.off_bars = o; o += 256;
    s.off_ring = o;
    const int slot_floats = 32 * Wn;
    const int budget = (227 * 1024 - 1024) / 4 - o;
    s.slots = budget / slot_floats;
    if (s.slots > 12) s.slots = 12;
    s.bytes = (size_t)(o + s.slots * slot_floats) * 4 + 1024;     // + alignment slack
    return s;
}

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// TMEM column map (512 columns allocated): accumulators at 0 and 128, relu(h) hi at 256, lo at 384
constexpr uint32_t TM_ACC0 = 0, TM_ACC1 = 128, TM_HHI = 256, TM_HLO = 384;

// barrier slots inside the `bars` block
constexpr int B_WFULL = 0, B_WEMPTY = 16, B_ACCFULL = 32, B_ACCEMPTY = 34, B_FEATFULL = 36, B_HFULL = 37, B_COUNT = 38;

__device__ __forceinline__ int w2_row0(int h) { return h == 0 ? 0 : h == 1 ? 3 : h == 2 ? 6 : h == 3 ? 10 : 11; }

template <int WN>
__global__ void __launch_bounds__(NTHREADS, 1)
k_deform_fwd_tc(Layout lo, TcPlan plan, SmemPlan sp, TcNet net, TcFwdIO io, const float* __restrict__ packed,
                const float* __restrict__ trow, int ntiles, int passes)
{
    extern __shared__ uint8_t smem_raw[];
    float* sm = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float* F_hi = sm + sp.off_fhi;
    float* F_lo = sm + sp.off_flo;
    float* W2s = sm + sp.off_w2;
    float* B0s = sm + sp.off_b0;
    float* B1s = sm + sp.off_b1;
    float* B2s = sm + sp.off_b2;
    float* osum = sm + sp.off_osum;
    float* ring = sm + sp.off_ring;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + sp.off_bars);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + B_COUNT + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = sp.slots;
    constexpr int NH = WN / 2;                 // accumulator columns per epilogue thread
    constexpr int SLOT = 32 * WN;              // floats per ring slot

    // ---- one-time setup -------------------------------------------------------------------------------
    if (tid == 0) {
        for (int i = 0; i < B_COUNT; ++i) mbar_init(&bars[i], 1);
        fence_mbar_init();
    }
    if (warp == 9) tmem_alloc(tmem_slot, 512);
    if (tid < NEPI) {
        for (int i = tid; i < WN; i += NEPI) B0s[i] = net.b0[i];
        for (int j = 0; j < plan.nheads; ++j) {
            const int h = plan.heads[j];
            const int NO = head_out(h);
            for (int i = tid; i < WN; i += NEPI) B1s[h * WN + i] = net.b1[h][i];
            for (int i = tid; i < NO * WN; i += NEPI) W2s[w2_row0(h) * WN + i] = net.w2[h][i];
            for (int i = tid; i < NO; i += NEPI) B2s[w2_row0(h) + i] = net.b2[h][i];
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 8) {
        // ===== producer: weight chunks, the same sequence for every tile ==================================
        if (lane == 0) {
            uint32_t g = 0;
            const uint32_t bytes = SLOT * 4;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
                for (int c = 0; c < plan.fwd_chunks; ++c, ++g) {
                    const uint32_t slot = g % S;
                    mbar_wait(&bars[B_WEMPTY + slot], ((g / S) & 1) ^ 1);
                    mbar_expect_tx(&bars[B_WFULL + slot], bytes);
                    bulk_g2s(ring + (size_t)slot * SLOT, net.wprep + (size_t)c * SLOT, bytes, &bars[B_WFULL + slot]);
                }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ===== MMA issuer ===================================================================================
        if (lane == 0) {
            const uint32_t idesc = make_idesc_tf32(128, WN, 0, 0);
            uint32_t g = 0, ph_feat = 0, ph_h = 0, pe0 = 1, pe1 = 1;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                // trunk: ACC0 = feat . W0^T   (A from shared memory)
                mbar_wait(&bars[B_FEATFULL], ph_feat); ph_feat ^= 1;
                mbar_wait(&bars[B_ACCEMPTY + 0], pe0); pe0 ^= 1;
                tc_fence_after();
                for (int c = 0; c < lo.F / KCH; ++c, ++g) {
                    const uint32_t slot = g % S;
                    mbar_wait(&bars[B_WFULL + slot], (g / S) & 1);
                    tc_fence_after();
                    const float* bh = ring + (size_t)slot * SLOT;
                    const float* bl = bh + 16 * WN;
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        const int k = c * KCH + ks * 8;
                        const int first = (c == 0 && ks == 0) ? 0 : 1;
                        if (passes == 3) {
                            mma_tf32_ss(tmem + TM_ACC0, desc_kmajor(F_lo, 128, k), desc_kmajor(bh, WN, ks * 8), idesc, first);
                            mma_tf32_ss(tmem + TM_ACC0, desc_kmajor(F_hi, 128, k), desc_kmajor(bl, WN, ks * 8), idesc, 1);
                            mma_tf32_ss(tmem + TM_ACC0, desc_kmajor(F_hi, 128, k), desc_kmajor(bh, WN, ks * 8), idesc, 1);
                        } else {
                            mma_tf32_ss(tmem + TM_ACC0, desc_kmajor(F_hi, 128, k), desc_kmajor(bh, WN, ks * 8), idesc, first);
                        }
                    }
                    mma_commit(&bars[B_WEMPTY + slot]);
                }
                mma_commit(&bars[B_ACCFULL + 0]);
                // heads: ACC[(i+1)&1] = relu(h) . W1^T   (A from TMEM)
                mbar_wait(&bars[B_HFULL], ph_h); ph_h ^= 1;
                tc_fence_after();
                for (int i = 0; i < plan.nheads; ++i) {
                    const int b = (i + 1) & 1;
                    if (b) { mbar_wait(&bars[B_ACCEMPTY + 1], pe1); pe1 ^= 1; }
                    else   { mbar_wait(&bars[B_ACCEMPTY + 0], pe0); pe0 ^= 1; }
                    tc_fence_after();
                    const uint32_t acc = tmem + (b ? TM_ACC1 : TM_ACC0);
                    for (int c = 0; c < WN / KCH; ++c, ++g) {
                        const uint32_t slot = g % S;
                        mbar_wait(&bars[B_WFULL + slot], (g / S) & 1);
                        tc_fence_after();
                        const float* bh = ring + (size_t)slot * SLOT;
                        const float* bl = bh + 16 * WN;
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            const uint32_t k = (uint32_t)(c * KCH + ks * 8);
                            const int first = (c == 0 && ks == 0) ? 0 : 1;
                            if (passes == 3) {
                                mma_tf32_ts(acc, tmem + TM_HLO + k, desc_kmajor(bh, WN, ks * 8), idesc, first);
                                mma_tf32_ts(acc, tmem + TM_HHI + k, desc_kmajor(bl, WN, ks * 8), idesc, 1);
                                mma_tf32_ts(acc, tmem + TM_HHI + k, desc_kmajor(bh, WN, ks * 8), idesc, 1);
                            } else {
                                mma_tf32_ts(acc, tmem + TM_HHI + k, desc_kmajor(bh, WN, ks * 8), idesc, first);
                            }
                        }
                        mma_commit(&bars[B_WEMPTY + slot]);
                    }
                    mma_commit(&bars[B_ACCFULL + b]);
                }
            }
        }
        __syncwarp();
    } else {
        // ===== sampler + epilogues: thread = point row `row`, column half `hf` ==========================
        const int row = tid & 127, hf = tid >> 7;
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int col0 = hf * NH;
        uint32_t pf0 = 0, pf1 = 0;
        const int C = lo.C;
        const int items = lo.F >> 2;
        const int LP = (items & 7) == 0 ? 8 : 4, IPT = items / LP, PPW = 32 / LP, SPASS = TP / (8 * PPW);
        const int sg = lane / LP, sq = lane % LP;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const int p0 = tile * TP;
            const int p = p0 + row;
            // ---- HexPlane sampling -> feat tile (hi / lo) ---------------------------------------------------
            // LP adjacent lanes share one point and take consecutive channel quads: a warp-wide gather then touches
            // 32/LP points x 64-byte taps instead of 32 unrelated lines (the L1 tag stage serves one line per cycle).
            for (int pass = 0; pass < SPASS; ++pass) {
                const int srow = pass * (8 * PPW) + warp * PPW + sg;
                const int sp_ = p0 + srow;
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
                    if (io.S_F)
                        __stcs(reinterpret_cast<float4*>(io.S_F + (size_t)tile * TP * lo.F + (srow >> 5) * 32 * lo.F + (item * 32 + (srow & 31)) * 4),
                               make_float4(prod[0], prod[1], prod[2], prod[3]));
                }
            }
            fence_async_smem();
            epi_bar();
            if (tid == 0) mbar_arrive(&bars[B_FEATFULL]);

            // ---- trunk epilogue: relu(h + b0) -> TMEM hi / lo --------------------------------------------------
            mbar_wait(&bars[B_ACCFULL + 0], pf0); pf0 ^= 1;
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < NH; c0 += 32) {
                uint32_t v[32], hi[32], lw[32];
                tmem_ld32(lane_base + TM_ACC0 + col0 + c0, v);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float x = fmaxf(__uint_as_float(v[j]) + B0s[col0 + c0 + j], 0.f);
                    float a, b;
                    split_tf32_rn(x, a, b);
                    hi[j] = __float_as_uint(a); lw[j] = __float_as_uint(b);
                    v[j] = __float_as_uint(x);
                }
                if (io.S_H) {
                    float* dst = io.S_H + (size_t)tile * TP * WN + (row >> 5) * 32 * WN + (row & 31) * 4;
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        __stcs(reinterpret_cast<float4*>(dst + ((col0 + c0 + j) >> 2) * 128),
                               make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
                }
                tmem_st32(lane_base + TM_HHI + col0 + c0, hi);
                tmem_st32(lane_base + TM_HLO + col0 + c0, lw);
            }
            tmem_wait_st();
            tc_fence_before();
            epi_bar();
            if (tid == 0) { mbar_arrive(&bars[B_HFULL]); mbar_arrive(&bars[B_ACCEMPTY + 0]); }

            // ---- heads -----------------------------------------------------------------------------------------
            for (int i = 0; i < plan.nheads; ++i) {
                const int h = plan.heads[i];
                const int b = (i + 1) & 1;
                const int NO = head_out(h);
                if (b) { mbar_wait(&bars[B_ACCFULL + 1], pf1); pf1 ^= 1; }
                else   { mbar_wait(&bars[B_ACCFULL + 0], pf0); pf0 ^= 1; }
                tc_fence_after();
                float z[NH];
#pragma unroll
                for (int c0 = 0; c0 < NH; c0 += 32) {
                    uint32_t v[32];
                    tmem_ld32(lane_base + (b ? TM_ACC1 : TM_ACC0) + col0 + c0, v);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 32; ++j) z[c0 + j] = fmaxf(__uint_as_float(v[j]) + B1s[h * WN + col0 + c0 + j], 0.f);
                    if (io.S_RZ[h]) {
                        float* dst = io.S_RZ[h] + (size_t)tile * TP * WN + (row >> 5) * 32 * WN + (row & 31) * 4;
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            __stcs(reinterpret_cast<float4*>(dst + ((col0 + c0 + j) >> 2) * 128),
                                   make_float4(z[c0 + j], z[c0 + j + 1], z[c0 + j + 2], z[c0 + j + 3]));
                    }
                }
                tc_fence_before();
                epi_bar();
                if (tid == 0) mbar_arrive(&bars[B_ACCEMPTY + b]);     // accumulator drained: next head may overwrite it

                // o = relu(z) . W2^T on the CUDA cores (out <= 48): this thread's half of the reduction
                const float* w2 = W2s + w2_row0(h) * WN + col0;
                float o[48];
#pragma unroll
                for (int n0 = 0; n0 < 48; n0 += 8) {
                    if (n0 < NO) {
                        float a[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u) a[u] = 0.f;
#pragma unroll
                        for (int k = 0; k < NH; k += 4) {
#pragma unroll
                            for (int u = 0; u < 8; ++u) {
                                if (n0 + u < NO) {
                                    const float4 w = *reinterpret_cast<const float4*>(w2 + (n0 + u) * WN + k);
                                    a[u] = fmaf(z[k], w.x, a[u]); a[u] = fmaf(z[k + 1], w.y, a[u]);
                                    a[u] = fmaf(z[k + 2], w.z, a[u]); a[u] = fmaf(z[k + 3], w.w, a[u]);
                                }
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 8; ++u) o[n0 + u] = a[u];
                    }
                }
                if (hf == 1) {
#pragma unroll
                    for (int n = 0; n < 48; ++n) if (n < NO) osum[n * 128 + row] = o[n];
                }
                epi_bar();
                if (hf == 0 && p < io.P) {
                    float* outp = h == 0 ? io.o_xyz : h == 1 ? io.o_scales : h == 2 ? io.o_rot : h == 3 ? io.o_opacity : io.o_shs;
                    const float* inp = h == 0 ? io.xyz : h == 1 ? io.scales : h == 2 ? io.rot : h == 3 ? io.opacity : io.shs;
                    const bool nobase = (h == 2 && lo.apply_rotation);
#pragma unroll
                    for (int n = 0; n < 48; ++n) {
                        if (n < NO) {
                            float base;
                            if (nobase) base = 0.f;
                            else if (h == 4) base = (io.sh_coeffs == 16) ? inp[(size_t)p * 48 + n] : inp[(size_t)p * io.sh_coeffs * 3 + (n % 3)];
                            else base = inp[(size_t)p * NO + n];
                            outp[(size_t)p * NO + n] = base + ((o[n] + osum[n * 128 + row]) + B2s[w2_row0(h) + n]);
                        }
                    }
                }
            }
        }
    }
    // ---- teardown ----------------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 9) tmem_free(tmem, 512);
}

static int sm_count() {
    static int n = 0;
    if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
    return n;
}

size_t save_bytes(const Layout& lo, int P) {
    if (!tc_supported(lo) || P <= 0) return 0;
    return fdgs_align_up(sizeof(float) * (size_t)make_save_layout(lo, P).total);
}

size_t wprep_bytes(const Layout& lo) {
    if (!tc_supported(lo)) return 256;
    const TcPlan p = make_plan(lo);
    return fdgs_align_up(sizeof(float) * (size_t)p.fwd_floats);
}

// Formats the weights (every call: they change with every optimizer step; ~0.7 MB) and runs the forward kernel.
int forward(const Layout& lo, const FdgsDeformForward* a, const float* packed, const float* trow, cudaStream_t stream) {
    const TcPlan plan = make_plan(lo);
    const SmemPlan sp = make_smem_plan(lo);
    if (sp.slots < 3) return FDGS_E_SMEM;
    if (!a->wprep_ws || a->wprep_bytes < wprep_bytes(lo)) return FDGS_E_WORKSPACE;
    float* wprep = (float*)a->wprep_ws;
    PrepArgs pa;
    pa.w0 = a->net.trunk_w[0]; pa.nheads = plan.nheads; pa.Wn = lo.Wn; pa.F = lo.F;
    for (int j = 0; j < FDGS_NUM_HEADS; ++j) { pa.w1[j] = a->net.head_w1[j]; pa.heads[j] = j < plan.nheads ? plan.heads[j] : 0; }
    { FDGS_PROF("k_deform_wprep", stream);
    k_deform_wprep<<<64, 256, 0, stream>>>(pa, wprep);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);

    TcNet net;
    net.wprep = wprep; net.b0 = a->net.trunk_b[0];
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) { net.b1[h] = a->net.head_b1[h]; net.w2[h] = a->net.head_w2[h]; net.b2[h] = a->net.head_b2[h]; }
    TcFwdIO io;
    io.P = a->P; io.sh_coeffs = a->sh_coeffs;
    io.xyz = a->xyz; io.scales = a->scales; io.rot = a->rotations; io.opacity = a->opacity; io.shs = a->shs;
    io.o_xyz = a->out_xyz; io.o_scales = a->out_scales; io.o_rot = a->out_rotations; io.o_opacity = a->out_opacity; io.o_shs = a->out_shs;
    io.S_F = nullptr; io.S_H = nullptr;
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) io.S_RZ[h] = nullptr;
    if (a->save_ws) {
        if (a->save_bytes < save_bytes(lo, a->P)) return FDGS_E_WORKSPACE;
        SaveLayout sl = make_save_layout(lo, a->P);
        float* S = (float*)a->save_ws;
        io.S_F = S + sl.off_F; io.S_H = S + sl.off_H;
        for (int h = 0; h < FDGS_NUM_HEADS; ++h) if ((lo.head_mask >> h) & 1) io.S_RZ[h] = S + sl.off_RZ[h];
    }
    const int ntiles = (a->P + TP - 1) / TP;
    const int grid = ntiles < sm_count() ? ntiles : sm_count();
    const int passes = a->precision == 1 ? 1 : 3;
    if (lo.Wn == 128) {
        FDGS_CUDA(cudaFuncSetAttribute(k_deform_fwd_tc<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp.bytes));
        { FDGS_PROF("k_deform_fwd_tc", stream);
        k_deform_fwd_tc<128><<<grid, NTHREADS, sp.bytes, stream>>>(lo, plan, sp, net, io, packed, trow, ntiles, passes);
        }
    } else {
        FDGS_CUDA(cudaFuncSetAttribute(k_deform_fwd_tc<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp.bytes));
        { FDGS_PROF("k_deform_fwd_tc", stream);
        k_deform_fwd_tc<64><<<grid, NTHREADS, sp.bytes, stream>>>(lo, plan, sp, net, io, packed, trow, ntiles, passes);
        }
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);
    return FDGS_OK;
}

bool supported(const Layout& lo) { return tc_supported(lo); }

}  // namespace fdgs_deform_tc
