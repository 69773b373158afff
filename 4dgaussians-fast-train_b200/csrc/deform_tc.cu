// Tensor-core (tcgen05 + TMEM) implementation of the fused HexPlane sampler + deformation MLP for sm_100a.
//
// One persistent CTA per SM, 128 points per tile, warp-specialised:
//   epilogue warps (8 or 16)  thread = point row = TMEM lane, 2 or 4 column groups
//   sampler warps (8)         HexPlane gathers of tile t+1 WHILE the epilogue warps work through the heads of tile t
//                             (LP adjacent lanes per point, see below); they write the feat tile once relu(h) is dead
//   producer warp             streams pre-formatted weight chunks global(L2) -> shared with 1-D bulk async copies
//   MMA warp                  one elected thread issues every tcgen05.mma and commits to mbarriers
//
// Every GEMM runs as 3xTF32 (hi*hi + lo*hi + hi*lo, FP32 accumulate in TMEM) so that results match the reference's
// FP32 cuBLAS/ATen path to ~1e-6 (BASELINE north_star: images within 1e-5); precision==1 runs one pass.
//   trunk   h = feat W0^T          A = feat tile (shared memory, K-major)            D -> TMEM ACC
//   relu(h) hi/lo -> shared memory (K-major): the A operand of every head's first GEMM
//   head j  z = relu(h) W1_j^T     D -> TMEM ACC ; relu(z) hi/lo -> TMEM (A operand of the output layer)
//           o = relu(z) W2_j^T     N = 16 / 48 (out padded), D -> TMEM OUT ; the next head's z GEMM overlaps its epilogue
//           heads with <= 4 outputs (pos, scales, rotations, opacity): o on the CUDA CORES inside the relu(z) pass
//           (FDGS_FWD_OUT_CC, see below) - no hi/lo split, no TMEM stores, no N = 16 GEMM to wait for
//   outputs are staged in shared memory and written (base + delta) with fully coalesced row-contiguous stores.
// Training calls also keep feat, relu(h), relu(z_j) in a blocked layout for the backward kernels (deform_tc_bwd.cu).
// Replaces scene/deformation.py:45-148 + scene/hexplane.py:73-183 (reference: ~200 ATen launches per view).
#include "deform_common.cuh"
#include "umma.cuh"
#include "prof.cuh"
#include <cstdio>
#ifdef FDGS_PHASE_TIMING
#define PT_DECL long long pt_[8] = {0,0,0,0,0,0,0,0}; long long pt_t = clock64();
#define PT(i) { const long long n_ = clock64(); pt_[i] += n_ - pt_t; pt_t = n_; }
#define PT_PRINT if (blockIdx.x == 0 && tid == 0) printf("fwd phases (cycles, epilogue thread 0 of CTA0): w2_load %lld | wait_trunk %lld | trunk_epi %lld | wait_z %lld | relu_pass %lld | wait_out %lld | out_pass %lld\n", pt_[0], pt_[1], pt_[2], pt_[3], pt_[4], pt_[5], pt_[6]);
#else
#define PT_DECL
#define PT(i)
#define PT_PRINT
#endif

namespace fdgs_deform_tc {
using namespace fdgs_deform;
using namespace fdgs_umma;

constexpr int KCH = 16;            // K columns per streamed weight chunk
// Epilogue / sampler warps: 4 per group of accumulator columns (a warp reaches only the 32 TMEM lanes of its quarter).
// Wn = 128 runs FOUR groups of 32 columns = 16 warps (4 per scheduler): the epilogue passes are latency-bound (phase timing
// and ncu: 34 % issue-active with 8 warps), FDGS_FWD_EPI16=0 builds the 8-warp layout (two groups of 64 columns) for A/B.
#ifndef FDGS_FWD_EPI16
#define FDGS_FWD_EPI16 1
#endif
template <int WN> struct EpiCfg {
    static constexpr int NG = (FDGS_FWD_EPI16 && WN == 128) ? 4 : 2;   // column groups
    static constexpr int NH = WN / NG;                                   // accumulator columns per epilogue thread
    static constexpr int NEPI = 128 * NG;                                // epilogue threads
    static constexpr int NEW = 4 * NG;                                   // epilogue warps
    static constexpr int NSMP = 8;                                       // sampler warps: NEW .. NEW + NSMP - 1
    static constexpr int WPROD = NEW + NSMP, WMMA = NEW + NSMP + 1;      // producer warp, MMA issuer warp
    static constexpr int NTHREADS = NEPI + NSMP * 32 + 64;
};
// Output layers with <= 4 columns on the CUDA cores (1) or as N = 16 tensor-core GEMMs like the 48-wide one (0, for A/B).
#ifndef FDGS_FWD_OUT_CC
#define FDGS_FWD_OUT_CC 1
#endif
constexpr int MAXCH = 96;          // chunks per tile: F/16 + heads * (Wn/16 + Wn/16)

// Pre-formatted weights ("wprep" workspace): a chunk holds KCH=16 reduction columns of an [N rows][K] weight as a hi
// tile then a lo tile, each the canonical K-major tile [4 column groups][N][4] (umma.cuh) => 32*N floats, one bulk copy.
// Memory order: W0 (N = Wn), then per active head W1_j (N = Wn) and W2_j (N = 16 or 48, rows >= out are zero).
// Stream order (coff[]) = MMA order: W0, W1_0, then per head i: W1_{i+1} (its GEMM runs ahead, into the other
// accumulator), W2_i.
struct TcPlan {
    int Wn, F, nheads;
    int heads[FDGS_NUM_HEADS];
    int nop[FDGS_NUM_HEADS];       // padded output width of plan slot j
    int cc[FDGS_NUM_HEADS];        // plan slot j: output layer on the CUDA cores (its W2 is not formatted or streamed)
    int nchunks;
    int coff[MAXCH];               // float offset of chunk c inside wprep
    int cfl[MAXCH];                // floats of chunk c
    long long floats;
};

static inline int out_pad(int h) { return head_out(h) > 16 ? 48 : 16; }
static inline int out_cc(int h) { return (FDGS_FWD_OUT_CC && head_out(h) <= 4) ? 1 : 0; }

static inline TcPlan make_plan(const Layout& lo) {
    TcPlan p;
    p.Wn = lo.Wn; p.F = lo.F; p.nheads = 0;
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) if ((lo.head_mask >> h) & 1) p.heads[p.nheads++] = h;
    long long o = (long long)(lo.F / KCH) * 32 * lo.Wn;
    long long off1[FDGS_NUM_HEADS], off2[FDGS_NUM_HEADS];
    for (int j = 0; j < FDGS_NUM_HEADS; ++j) {
        p.nop[j] = j < p.nheads ? out_pad(p.heads[j]) : 16;
        p.cc[j] = j < p.nheads ? out_cc(p.heads[j]) : 0;
        off1[j] = off2[j] = 0;
        if (j < p.nheads) {
            off1[j] = o; o += (long long)(lo.Wn / KCH) * 32 * lo.Wn;
            off2[j] = o; if (!p.cc[j]) o += (long long)(lo.Wn / KCH) * 32 * p.nop[j];
        }
    }
    p.floats = o;
    int c = 0;
    auto add = [&](long long at, int n, int N) { for (int i = 0; i < n; ++i) { p.coff[c] = (int)(at + (long long)i * 32 * N); p.cfl[c] = 32 * N; ++c; } };
    add(0, lo.F / KCH, lo.Wn);
    if (p.nheads > 0) add(off1[0], lo.Wn / KCH, lo.Wn);
    for (int j = 0; j < p.nheads; ++j) {
        if (j + 1 < p.nheads) add(off1[j + 1], lo.Wn / KCH, lo.Wn);
        if (!p.cc[j]) add(off2[j], lo.Wn / KCH, p.nop[j]);
    }
    p.nchunks = c;
    return p;
}

static inline bool tc_supported(const Layout& lo) {
    return lo.depth == 1 && (lo.Wn == 64 || lo.Wn == 128) && lo.F % KCH == 0 && lo.F >= KCH && lo.F <= 64 && lo.C >= 8;
}

struct PrepSec { const float* W; int ld, N, Nvalid, K; long long off; };
struct PrepArgs { PrepSec s[1 + 2 * FDGS_NUM_HEADS]; int nsec; };

// one thread per (padded) weight element: split (round-to-nearest) and scatter into its chunk
__global__ void k_deform_wprep(PrepArgs a, float* __restrict__ out) {
    for (int si = 0; si < a.nsec; ++si) {
        const PrepSec s = a.s[si];
        const int total = s.N * s.K;
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
            const int n = i / s.K, k = i % s.K;
            const float w = n < s.Nvalid ? s.W[(size_t)n * s.ld + k] : 0.f;
            float hi, lo;
            split_tf32_rn(w, hi, lo);
            float* c = out + s.off + (long long)(k / KCH) * 32 * s.N;
            const int o = tile_off(n, k % KCH, s.N);
            c[o] = hi;
            c[16 * s.N + o] = lo;
        }
    }
}

struct TcNet {
    const float* wprep;
    const float* b0;
    const float* b1[FDGS_NUM_HEADS];
    const float* b2[FDGS_NUM_HEADS];
    const float* w2[FDGS_NUM_HEADS];     // raw [NO][Wn] output weights (CUDA-core output layers)
};

struct TcFwdIO {
    int P, sh_coeffs;
    const float *xyz, *scales, *rot, *opacity, *shs;
    float *o_xyz, *o_scales, *o_rot, *o_opacity, *o_shs;
    // training: activations kept for the backward pass (blocked layout [tile][row/32][col/4][row%32][4]); null = none
    float *S_F, *S_H;
    float* S_RZ[FDGS_NUM_HEADS];     // by head id
    uint32_t* S_M;                   // ReLU gate masks [6][Pp][4]
    long long Pp;
};

// shared-memory plan (floats unless noted), all offsets multiples of 256 floats
struct SmemPlan {
    int slots;
    int off_hs_hi, off_hs_lo, off_b0, off_b1, off_b2, off_out, off_bars, off_ring;
    size_t bytes;
};

static inline SmemPlan make_smem_plan(const Layout& lo) {
    SmemPlan s;
    auto al = [](int x) { return (x + 255) & ~255; };
    const int Wn = lo.Wn;
    int o = 0;
    s.off_hs_hi = o; o += al(128 * Wn);        // relu(h) hi (K-major canonical tile); the feat tile aliases its head
    s.off_hs_lo = o; o += al(128 * Wn);
    s.off_b0 = o; o += al(Wn);
    s.off_b1 = o; o += al(FDGS_NUM_HEADS * Wn);
    s.off_b2 = o; o += al(64);
    s.off_out = o; o += al(52 * 128);          // output staging [row][out], row stride 52 for the 48-wide head
    s.off_bars = o; o += 256;
    s.off_ring = o;
    const int slot_floats = 32 * Wn;
    const int budget = (227 * 1024 - 1024) / 4 - o;
    s.slots = budget / slot_floats;
    if (s.slots > 8) s.slots = 8;
    s.bytes = (size_t)(o + s.slots * slot_floats) * 4 + 1024;     // + alignment slack
    return s;
}

template <int N> __device__ __forceinline__ void epi_bar_n() { asm volatile("bar.sync 1, %0;" ::"n"(N) : "memory"); }

// TMEM column map (512 columns allocated): two accumulators, relu(z) hi, lo.  Head i uses accumulator i & 1: z_i, then
// (once relu(z_i) has been read out of it) its first 16 / 48 columns take the output layer o_i.  z_{i+1} is computed into
// the other accumulator WHILE the epilogue warps turn z_i into relu(z_i) hi / lo.
constexpr uint32_t TM_A0 = 0, TM_A1 = 128, TM_ZHI = 256, TM_ZLO = 384;
constexpr int B_WFULL = 0, B_WEMPTY = 8, B_FEATFULL = 16, B_ACCFULL = 17 /* +1 */, B_HFULL = 19, B_RZFULL = 20, B_OUTFULL = 21,
              B_OUTFREE = 22, B_HSFREE = 23, B_TRUNKFULL = 24, B_COUNT = 25;

__device__ __forceinline__ int b2_row0(int h) { return h == 0 ? 0 : h == 1 ? 3 : h == 2 ? 6 : h == 3 ? 10 : 11; }

template <int WN>
__global__ void __launch_bounds__(EpiCfg<WN>::NTHREADS, 1)
k_deform_fwd_tc(Layout lo, TcPlan plan, SmemPlan sp, TcNet net, TcFwdIO io, const float* __restrict__ packed,
                const float* __restrict__ trow, int ntiles, int passes)
{
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment by OFFSET from the shared array (not by integer casts of the pointer): the compiler keeps the
    // shared address space and emits LDS / STS with 32-bit addresses instead of generic LD / ST
    float* sm = reinterpret_cast<float*>(smem_raw + ((1024u - (smem_addr(smem_raw) & 1023u)) & 1023u));
    float* HS_hi = sm + sp.off_hs_hi;
    float* HS_lo = sm + sp.off_hs_lo;
    float* F_hi = HS_hi;
    float* F_lo = HS_lo;
    float* B0s = sm + sp.off_b0;
    float* B1s = sm + sp.off_b1;
    float* B2s = sm + sp.off_b2;
    float* OUTs = sm + sp.off_out;
    // CUDA-core output layers (plan.cc): their W2 as [head][column][4] and the per-column-group partial sums, double
    // buffered.  Both live in the staging buffer, which only a tensor-core output layer (the 48-wide head) fills with rows:
    // W2s is therefore re-read at the start of every tile.
    float* W2s = OUTs;                                        // [4][WN][4]
    float* PARTs = OUTs + 4 * WN * 4;                         // [2][NG][128][4]
    float* ring = sm + sp.off_ring;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm + sp.off_bars);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + B_COUNT + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = sp.slots;
    constexpr int NG = EpiCfg<WN>::NG, NH = EpiCfg<WN>::NH, NEPI = EpiCfg<WN>::NEPI, NEW = EpiCfg<WN>::NEW;
    constexpr int NSMP = EpiCfg<WN>::NSMP, WPROD = EpiCfg<WN>::WPROD, WMMA = EpiCfg<WN>::WMMA;
    auto epi_bar = [] { epi_bar_n<EpiCfg<WN>::NEPI>(); };
    constexpr int SLOT = 32 * WN;              // floats per ring slot
    const int F = lo.F;

    // ---- one-time setup -------------------------------------------------------------------------------
    if (tid == 0) {
        for (int i = 0; i < B_COUNT; ++i) mbar_init(&bars[i], 1);
        fence_mbar_init();
    }
    if (warp == WMMA) tmem_alloc(tmem_slot, 512);
    if (tid < NEPI) {
        for (int i = tid; i < WN; i += NEPI) B0s[i] = net.b0[i];
        for (int j = 0; j < plan.nheads; ++j) {
            const int h = plan.heads[j];
            const int NO = head_out(h);
            for (int i = tid; i < WN; i += NEPI) B1s[h * WN + i] = net.b1[h][i];
            for (int i = tid; i < NO; i += NEPI) B2s[b2_row0(h) + i] = net.b2[h][i];
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == WPROD) {
        // ===== producer: weight chunks, the same sequence for every tile ==================================
        if (lane == 0) {
            uint32_t g = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
                for (int c = 0; c < plan.nchunks; ++c, ++g) {
                    const uint32_t slot = g % S;
                    const uint32_t bytes = (uint32_t)plan.cfl[c] * 4;
                    mbar_wait(&bars[B_WEMPTY + slot], ((g / S) & 1) ^ 1);
                    mbar_expect_tx(&bars[B_WFULL + slot], bytes);
                    bulk_g2s(ring + (size_t)slot * SLOT, net.wprep + plan.coff[c], bytes, &bars[B_WFULL + slot]);
                }
        }
        __syncwarp();
    } else if (warp == WMMA) {
        // ===== MMA issuer: the whole warp runs the loop, one elected lane issues (see umma.cuh:elect_one) ============
        {
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
            const uint32_t leader = elect_one();
            const uint32_t idesc = make_idesc_tf32(128, WN, 0, 0);
            const uint64_t f_hi0 = desc_kmajor(F_hi, 128, 0), f_lo0 = desc_kmajor(F_lo, 128, 0);   // feat and relu(h) alias
            const uint32_t ring_u = smem_addr(ring) >> 4;
            uint32_t slot = 0, wph = 0, ph_feat = 0, ph_h = 0, ph_rz = 0, ph_free = 0;
            // D[acc] (+)= A . B^T over `nck` streamed chunks of an [N][K] weight; A: shared-memory tile (SS) or TMEM columns
            auto gemm = [&](uint32_t acc, int nck, int N, uint32_t id, bool ss, uint32_t ta_hi, uint32_t ta_lo) {
                const uint64_t bd0 = make_sdesc(0u, (uint32_t)N * 16u, 128u);      // B descriptor without the start address
                const uint32_t bstep = (uint32_t)N * 2u;                           // 8 K-columns of an N-row tile, 16-byte units
#pragma unroll 1
                for (int c = 0; c < nck; ++c) {
                    mbar_wait(&bars[B_WFULL + slot], wph);
                    tc_fence_after();
                    if (leader) {
                        const uint64_t dbh0 = bd0 + (ring_u + slot * (uint32_t)(SLOT / 4));
                        const uint64_t dbl0 = dbh0 + (uint32_t)N * 4u;             // lo half follows the 16 hi columns
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            const uint32_t kk = (uint32_t)(c * 2 + ks);
                            const int first = (c == 0 && ks == 0) ? 0 : 1;
                            const uint64_t dbh = dbh0 + ks * bstep, dbl = dbl0 + ks * bstep;
                            if (ss) {
                                const uint64_t ah = f_hi0 + kk * 256u, al = f_lo0 + kk * 256u;
                                if (passes == 3) {
                                    mma_tf32_ss(acc, al, dbh, id, first);
                                    mma_tf32_ss(acc, ah, dbl, id, 1);
                                    mma_tf32_ss(acc, ah, dbh, id, 1);
                                } else {
                                    mma_tf32_ss(acc, ah, dbh, id, first);
                                }
                            } else {
                                if (passes == 3) {
                                    mma_tf32_ts(acc, ta_lo + kk * 8u, dbh, id, first);
                                    mma_tf32_ts(acc, ta_hi + kk * 8u, dbl, id, 1);
                                    mma_tf32_ts(acc, ta_hi + kk * 8u, dbh, id, 1);
                                } else {
                                    mma_tf32_ts(acc, ta_hi + kk * 8u, dbh, id, first);
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
            // The trunk GEMM of tile t+1 can start while the epilogue warps still drain the LAST head of tile t (the sampler
            // warps deliver the feat tile early): it goes to the accumulator that head does not use, once the head before
            // it (the previous user of that accumulator) has been drained (B_OUTFREE, arrived by every head but the last).
            const uint32_t trunk_acc = (plan.nheads & 1) ? TM_A1 : TM_A0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                mbar_wait(&bars[B_FEATFULL], ph_feat); ph_feat ^= 1;
                if (tile != (int)blockIdx.x && plan.nheads >= 2) { mbar_wait(&bars[B_OUTFREE], ph_free); ph_free ^= 1; }
                tc_fence_after();
                gemm(tm + trunk_acc, F / KCH, WN, idesc, true, 0, 0);                         // h = feat W0^T
                commit(B_TRUNKFULL);        // a barrier of its own: this GEMM can complete before the epilogue warps have
                                            // waited for the last z of the previous tile (a shared barrier would lap them)
                mbar_wait(&bars[B_HFULL], ph_h); ph_h ^= 1;
                tc_fence_after();
                if (plan.nheads > 0) {
                    gemm(tm + TM_A0, WN / KCH, WN, idesc, true, 0, 0);                        // z_0 = relu(h) W1_0^T
                    commit(B_ACCFULL);
                    if (plan.nheads == 1) commit(B_HSFREE);            // last reader of relu(h): the sampler warps may overwrite it
                }
                for (int i = 0; i < plan.nheads; ++i) {
                    if (i + 1 < plan.nheads) {
                        if (i >= 1) {                                  // o_{i-1} has been read out of this accumulator
                            mbar_wait(&bars[B_OUTFREE], ph_free); ph_free ^= 1;
                            tc_fence_after();
                        }
                        gemm(tm + (((i + 1) & 1) ? TM_A1 : TM_A0), WN / KCH, WN, idesc, true, 0, 0);   // z_{i+1}: runs ahead
                        commit(B_ACCFULL + ((i + 1) & 1));
                        if (i + 2 == plan.nheads) commit(B_HSFREE);
                    }
                    if (plan.cc[i]) continue;                          // o_i is computed by the epilogue warps themselves
                    mbar_wait(&bars[B_RZFULL], ph_rz); ph_rz ^= 1;     // z_i drained, relu(z_i) hi/lo in TMEM
                    tc_fence_after();
                    gemm(tm + ((i & 1) ? TM_A1 : TM_A0), WN / KCH, plan.nop[i], make_idesc_tf32(128, plan.nop[i], 0, 0), false,
                         tm + TM_ZHI, tm + TM_ZLO);                                           // o_i = relu(z_i) W2_i^T
                    commit(B_OUTFULL);
                }
            }
        }
    } else if (warp >= NEW) {
        // ===== sampler warps =====================================================================================
        // LP adjacent lanes share one point and take consecutive channel quads, so a warp-wide gather touches 32/LP points
        // x 64-byte taps instead of 32 unrelated lines (the L1 tag stage serves one line per cycle).  The gathers of tile
        // t+1 run while the epilogue warps and the tensor core work on tile t; the sampled values wait in nf[] until the
        // last GEMM that reads relu(h) of tile t has completed (B_HSFREE: the feat tile aliases that buffer).
        const int sw = warp - NEW, stid = tid - NEPI;
        const int C = lo.C;
        const int items = F >> 2;
        const int LP = (items & 7) == 0 ? 8 : 4, IPT = items / LP, PPW = 32 / LP, SPASS = TP / (NSMP * PPW);
        const int sg = lane / LP, sq = lane % LP;
        float4 nf[8];                                                // SPASS * IPT <= 8 float4 per thread
        uint32_t ph_free = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            for (int pass = 0; pass < SPASS; ++pass) {
                const int srow = pass * (NSMP * PPW) + sw * PPW + sg;
                const int sp_ = tile * TP + srow;
                const bool sval = sp_ < io.P;
                float q[4] = {0.f, 0.f, 0.f, 0.f};
                if (sval) point_coords(lo, io.xyz, sp_, 0.f, q);
                for (int it = 0; it < IPT; ++it) {
                    const int item = sq + it * LP;
                    const int l = (item * 4) / C, c0 = (item * 4) % C;
                    float prod[4] = {0.f, 0.f, 0.f, 0.f};
                    if (sval) sample_level<4>(lo, l, q, packed, trow, c0, prod);
                    nf[pass * IPT + it] = make_float4(prod[0], prod[1], prod[2], prod[3]);
                }
            }
            if (tile != (int)blockIdx.x) {                           // relu(h) of the previous tile is dead
                mbar_wait(&bars[B_HSFREE], ph_free); ph_free ^= 1;
                tc_fence_after();
            }
            // ---- feat tile (hi / lo) -> shared memory (A operand of the trunk GEMM), kept copy ---------------------
            for (int pass = 0; pass < SPASS; ++pass) {
                const int srow = pass * (NSMP * PPW) + sw * PPW + sg;
                for (int it = 0; it < IPT; ++it) {
                    const int item = sq + it * LP;
                    const float4 pr = nf[pass * IPT + it];
                    float4 hi, lw;
                    split_tf32_rn(pr.x, hi.x, lw.x); split_tf32_rn(pr.y, hi.y, lw.y);
                    split_tf32_rn(pr.z, hi.z, lw.z); split_tf32_rn(pr.w, hi.w, lw.w);
                    const int o = tile_off(srow, item * 4, 128);
                    *reinterpret_cast<float4*>(F_hi + o) = hi;
                    *reinterpret_cast<float4*>(F_lo + o) = lw;
                    if (io.S_F)
                        __stcs(reinterpret_cast<float4*>(io.S_F + (size_t)tile * TP * F + (srow >> 5) * 32 * F + (item * 32 + (srow & 31)) * 4), pr);
                }
            }
            fence_async_smem();
            asm volatile("bar.sync 2, %0;" ::"n"(NSMP * 32) : "memory");
            if (stid == 0) mbar_arrive(&bars[B_FEATFULL]);
        }
    } else {
        // ===== epilogues ===========================================================================================
        const int row = tid & 127, hf = tid >> 7;                    // epilogue mapping: TMEM lane = point row, column group
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int col0 = hf * NH;
        uint32_t ph_a0 = 0, ph_a1 = 0, ph_out = 0, ph_tr = 0;
        uint32_t plan_ccmask = 0;                                    // by head id
        for (int j = 0; j < plan.nheads; ++j) if (plan.cc[j]) plan_ccmask |= 1u << plan.heads[j];
        PT_DECL
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const int p0 = tile * TP;
            const size_t rowoff = (size_t)tile * TP * WN + (row >> 5) * 32 * WN + (row & 31) * 4;   // blocked save layout
            // W2 of the CUDA-core output layers: 4 * WN * 4 floats = 4 per thread (visible to the head loop through the
            // barrier that ends the trunk epilogue)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int e = tid + k * NEPI;
                const int hh = e / (WN * 4), c = (e >> 2) % WN, o = e & 3;
                W2s[e] = (((plan_ccmask >> hh) & 1) && o < head_out(hh)) ? __ldg(net.w2[hh] + o * WN + c) : 0.f;
            }
            PT(0)
            // ---- trunk epilogue: relu(h + b0) -> shared memory hi / lo (A operand of the heads), kept copy ----------
            mbar_wait(&bars[B_TRUNKFULL], ph_tr); ph_tr ^= 1;
            tc_fence_after();
            PT(1)
#pragma unroll 1
            for (int c0 = 0; c0 < NH; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(lane_base + ((plan.nheads & 1) ? TM_A1 : TM_A0) + col0 + c0, v);
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
                    if (io.S_H)
                        __stcs(reinterpret_cast<float4*>(io.S_H + rowoff + ((col0 + c0 + j) >> 2) * 128), make_float4(x[0], x[1], x[2], x[3]));
                }
                if (io.S_M) io.S_M[((size_t)(p0 + row)) * 4 + hf * (4 / NG) + (c0 >> 5)] = m;    // word layout the backward reads
            }
            fence_async_smem();
            tc_fence_before();
            epi_bar();
            if (tid == 0) mbar_arrive(&bars[B_HFULL]);
            if (tid == 0 && plan.nheads == 0) mbar_arrive(&bars[B_HSFREE]);      // otherwise: the MMA warp, after the last z GEMM
            PT(2)

            // ---- heads -----------------------------------------------------------------------------------------
            for (int i = 0; i < plan.nheads; ++i) {
                const int h = plan.heads[i];
                const int NO = head_out(h);
                const int OS = NO == 48 ? 52 : NO;                      // staging row stride
                // base values (the head's input rows of this tile, contiguous): thread t owns float4s t, t + NEPI, ... of the
                // tile's rows * NO floats; fetched ahead of the wait whose latency hides them
                float* outp = h == 0 ? io.o_xyz : h == 1 ? io.o_scales : h == 2 ? io.o_rot : h == 3 ? io.o_opacity : io.o_shs;
                const float* inp = h == 0 ? io.xyz : h == 1 ? io.scales : h == 2 ? io.rot : h == 3 ? io.opacity : io.shs;
                const int rows = io.P - p0 < TP ? io.P - p0 : TP;
                const int total = rows * NO;
                const bool vec = (total & 3) == 0 && !(h == 2 && lo.apply_rotation) && !(h == 4 && io.sh_coeffs != 16);
                const float4* src4 = reinterpret_cast<const float4*>(inp + (size_t)p0 * NO);
                float4 base0 = make_float4(0.f, 0.f, 0.f, 0.f);       // <= 4 outputs: rows * NO <= 512 floats, one float4 per thread
                if (vec && plan.cc[i] && tid * 4 < total) base0 = __ldcs(src4 + tid);
                // z = relu(acc + b1): kept copy, hi / lo -> TMEM (A operand of the output layer)
                const uint32_t acc = lane_base + ((i & 1) ? TM_A1 : TM_A0);
                if (i & 1) { mbar_wait(&bars[B_ACCFULL + 1], ph_a1); ph_a1 ^= 1; }
                else { mbar_wait(&bars[B_ACCFULL], ph_a0); ph_a0 ^= 1; }
                tc_fence_after();
                PT(3)
                float* rz = io.S_RZ[h] ? io.S_RZ[h] + rowoff : nullptr;
                uint32_t zm0 = 0u, zm1 = 0u;
#ifndef FDGS_RELU_UNROLL
#define FDGS_RELU_UNROLL 1
#endif
                constexpr int RELU_UNROLL = FDGS_RELU_UNROLL;
                if (plan.cc[i]) {
                    // ---- <= 4 outputs: o = relu(z) W2^T on the CUDA cores, inside the pass that drains z ------------------
                    // Each thread dots its NH columns of relu(z) with W2 (one broadcast LDS.128 per column) while it writes
                    // the kept copy; the NG partial sums of a row meet in shared memory (fixed order: deterministic).
                    float po[4] = {0.f, 0.f, 0.f, 0.f};
                    const float* w2h = W2s + (h * WN + col0) * 4;
#pragma unroll (RELU_UNROLL)
                    for (int c0 = 0; c0 < NH; c0 += 16) {
                        uint32_t v[16];
                        tmem_ld16(acc + col0 + c0, v);
                        tmem_wait_ld();
                        uint32_t mbits = 0u;
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float z = fmaxf(__uint_as_float(v[j]) + B1s[h * WN + col0 + c0 + j], 0.f);
                            const float4 w = *reinterpret_cast<const float4*>(w2h + (c0 + j) * 4);
                            po[0] = fmaf(z, w.x, po[0]); po[1] = fmaf(z, w.y, po[1]);
                            po[2] = fmaf(z, w.z, po[2]); po[3] = fmaf(z, w.w, po[3]);
                            v[j] = __float_as_uint(z);
                            mbits |= (z > 0.f ? 1u : 0u) << j;
                        }
                        if (c0 < 32) zm0 |= mbits << (c0 & 31); else zm1 |= mbits << (c0 & 31);
                        if (rz) {
#pragma unroll
                            for (int j = 0; j < 16; j += 4)
                                __stcs(reinterpret_cast<float4*>(rz + ((col0 + c0 + j) >> 2) * 128),
                                       make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
                        }
                    }
                    if (io.S_M) {
                        uint32_t* mp = io.S_M + ((size_t)(1 + h) * io.Pp + (p0 + row)) * 4 + hf * (4 / NG);
                        mp[0] = zm0;
                        if (NH > 32) mp[1] = zm1;
                    }
                    const float* pb = PARTs + (i & 1) * (NG * 512);
                    *reinterpret_cast<float4*>(PARTs + (i & 1) * (NG * 512) + (hf * 128 + row) * 4) = make_float4(po[0], po[1], po[2], po[3]);
                    tc_fence_before();
                    epi_bar();
                    if (tid == 0 && i + 1 < plan.nheads) mbar_arrive(&bars[B_OUTFREE]);          // accumulator free for z_{i+2} / the next trunk
                    PT(4)
                    auto delta = [&](int idx) {
                        const int r = idx / NO, n = idx - r * NO;
                        float d = B2s[b2_row0(h) + n];
#pragma unroll
                        for (int g = 0; g < NG; ++g) d += pb[(g * 128 + r) * 4 + n];
                        return d;
                    };
                    float* dst = outp + (size_t)p0 * NO;
                    if (vec) {                                        // rows * NO <= 512 floats: one float4 per thread at most
                        if (tid * 4 < total)
                            reinterpret_cast<float4*>(dst)[tid] = make_float4(base0.x + delta(tid * 4), base0.y + delta(tid * 4 + 1),
                                                                              base0.z + delta(tid * 4 + 2), base0.w + delta(tid * 4 + 3));
                    } else if (h == 2 && lo.apply_rotation) {
                        for (int idx = tid; idx < total; idx += NEPI) dst[idx] = delta(idx);
                    } else {
                        const float* src = inp + (size_t)p0 * NO;
                        for (int idx = tid; idx < total; idx += NEPI) dst[idx] = src[idx] + delta(idx);
                    }
                    PT(6)
                    continue;                                         // the partial sums are double buffered: no barrier here
                }
#pragma unroll (RELU_UNROLL)
                for (int c0 = 0; c0 < NH; c0 += 16) {
                    uint32_t v[16], hi[16], lw[16];
                    tmem_ld16(acc + col0 + c0, v);
                    tmem_wait_ld();
                    uint32_t mbits = 0u;
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float z = fmaxf(__uint_as_float(v[j]) + B1s[h * WN + col0 + c0 + j], 0.f);
                        float a, b;
                        split_tf32_rn(z, a, b);
                        hi[j] = __float_as_uint(a); lw[j] = __float_as_uint(b);
                        v[j] = __float_as_uint(z);
                        mbits |= (z > 0.f ? 1u : 0u) << j;
                    }
                    if (c0 < 32) zm0 |= mbits << (c0 & 31); else zm1 |= mbits << (c0 & 31);
                    if (rz) {
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                            __stcs(reinterpret_cast<float4*>(rz + ((col0 + c0 + j) >> 2) * 128),
                                   make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
                    }
                    tmem_st16(lane_base + TM_ZHI + col0 + c0, hi);
                    tmem_st16(lane_base + TM_ZLO + col0 + c0, lw);
                }
                if (io.S_M) {
                    uint32_t* mp = io.S_M + ((size_t)(1 + h) * io.Pp + (p0 + row)) * 4 + hf * (4 / NG);
                    mp[0] = zm0;
                    if (NH > 32) mp[1] = zm1;
                }
                tmem_wait_st();
                tc_fence_before();
                epi_bar();
                if (tid == 0) mbar_arrive(&bars[B_RZFULL]);
                PT(4)
                float4 base[6];                                       // fetched now: the latency hides behind the o_i GEMM
                if (vec) {
#pragma unroll
                    for (int k = 0; k < 6; ++k) {
                        const int i4 = tid + k * NEPI;
                        base[k] = (i4 * 4 < total) ? __ldcs(src4 + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }

                // o = relu(z) W2^T (+ b2): accumulator -> staging [row][NO] -> coalesced base + delta
                mbar_wait(&bars[B_OUTFULL], ph_out); ph_out ^= 1;
                tc_fence_after();
                PT(5)
                {
                    const int npt = plan.nop[i] / NG;                 // output columns per thread: 8 / 24 (two groups) or 4 / 12
                    // (row stride OS: 48 floats would put a warp's 32 rows on two banks, 52 spreads them over eight)
                    if (NG == 2) {
                        for (int c0 = 0; c0 < npt; c0 += 8) {
                            uint32_t v[8];
                            tmem_ld8(acc + hf * npt + c0, v);
                            tmem_wait_ld();
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const int n = hf * npt + c0 + j;
                                if (n < NO) OUTs[row * OS + n] = __uint_as_float(v[j]) + B2s[b2_row0(h) + n];
                            }
                        }
                    } else {
                        for (int c0 = 0; c0 < npt; c0 += 4) {
                            uint32_t v[4];
                            tmem_ld4(acc + hf * npt + c0, v);
                            tmem_wait_ld();
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int n = hf * npt + c0 + j;
                                if (n < NO) OUTs[row * OS + n] = __uint_as_float(v[j]) + B2s[b2_row0(h) + n];
                            }
                        }
                    }
                }
                tc_fence_before();
                epi_bar();
                if (tid == 0 && i + 1 < plan.nheads) mbar_arrive(&bars[B_OUTFREE]);      // accumulator free for z_{i+2} / the next trunk
                {
                    const bool nobase = (h == 2 && lo.apply_rotation);
                    float* dst = outp + (size_t)p0 * NO;
                    if (vec) {
                        float4* dst4 = reinterpret_cast<float4*>(dst);
                        const float4* o4 = reinterpret_cast<const float4*>(OUTs);
#pragma unroll
                        for (int k = 0; k < 6; ++k) {
                            const int i4 = tid + k * NEPI;
                            if (i4 * 4 < total) {
                                const float4 d = NO == 48 ? o4[(i4 / 12) * 13 + (i4 % 12)] : o4[i4];
                                dst4[i4] = make_float4(base[k].x + d.x, base[k].y + d.y, base[k].z + d.z, base[k].w + d.w);
                            }
                        }
                    } else if (h == 4 && io.sh_coeffs != 16) {        // M = 1 broadcast: base index (p, n % 3)
                        for (int idx = tid; idx < total; idx += NEPI) {
                            const int r = idx / NO, n = idx % NO;
                            dst[idx] = inp[(size_t)(p0 + r) * io.sh_coeffs * 3 + (n % 3)] + OUTs[r * OS + n];
                        }
                    } else if (nobase) {
                        for (int idx = tid; idx < total; idx += NEPI) dst[idx] = OUTs[(idx / NO) * OS + idx % NO];
                    } else {
                        const float* src = inp + (size_t)p0 * NO;
                        for (int idx = tid; idx < total; idx += NEPI) dst[idx] = src[idx] + OUTs[(idx / NO) * OS + idx % NO];
                    }
                }
                epi_bar();                                            // staging buffer free for the next head
                PT(6)
            }
        }
        PT_PRINT
    }
    // ---- teardown ----------------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == WMMA) tmem_free(tmem, 512);
}

static int sm_count() {          // of the CURRENT device (one process may drive several)
    static int n[64] = {0};
    int dev = 0; cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!n[dev]) { cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev); if (n[dev] <= 0) n[dev] = 148; }
    const int lim = fdgs_sm_limit_value();        // fdgs_set_sm_limit: leave SMs to a concurrent kernel (the NCCL all-reduce)
    return (lim > 0 && lim < n[dev]) ? lim : n[dev];
}

size_t save_bytes(const Layout& lo, int P) {
    if (!tc_supported(lo) || P <= 0) return 0;
    return fdgs_align_up(sizeof(float) * (size_t)make_save_layout(lo, P).total);
}

size_t wprep_bytes(const Layout& lo) {
    if (!tc_supported(lo)) return 256;
    const TcPlan p = make_plan(lo);
    return fdgs_align_up(sizeof(float) * (size_t)p.floats);
}

// Formats the weights (every call: they change with every optimizer step; ~0.8 MB) and runs the forward kernel.
int forward(const Layout& lo, const FdgsDeformForward* a, const float* packed, const float* trow, cudaStream_t stream) {
    const TcPlan plan = make_plan(lo);
    const SmemPlan sp = make_smem_plan(lo);
    if (sp.slots < 3 || plan.nchunks > MAXCH) return FDGS_E_SMEM;
    if (!a->wprep_ws || a->wprep_bytes < wprep_bytes(lo)) return FDGS_E_WORKSPACE;
    float* wprep = (float*)a->wprep_ws;
    const int Wn = lo.Wn, F = lo.F;
    PrepArgs pa;
    int ns = 0;
    long long off = 0;
    pa.s[ns++] = PrepSec{a->net.trunk_w[0], F, Wn, Wn, F, off}; off += (long long)(F / KCH) * 32 * Wn;
    for (int j = 0; j < plan.nheads; ++j) {
        const int h = plan.heads[j];
        pa.s[ns++] = PrepSec{a->net.head_w1[h], Wn, Wn, Wn, Wn, off}; off += (long long)(Wn / KCH) * 32 * Wn;
        if (plan.cc[j]) continue;                            // read raw by the kernel (CUDA-core output layer)
        pa.s[ns++] = PrepSec{a->net.head_w2[h], Wn, plan.nop[j], head_out(h), Wn, off}; off += (long long)(Wn / KCH) * 32 * plan.nop[j];
    }
    pa.nsec = ns;
    if (off != plan.floats) return FDGS_E_BADARG;
    { FDGS_PROF("k_deform_wprep", stream);
    k_deform_wprep<<<64, 256, 0, stream>>>(pa, wprep);
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);

    TcNet net;
    net.wprep = wprep; net.b0 = a->net.trunk_b[0];
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) { net.b1[h] = a->net.head_b1[h]; net.b2[h] = a->net.head_b2[h]; net.w2[h] = a->net.head_w2[h]; }
    TcFwdIO io;
    io.P = a->P; io.sh_coeffs = a->sh_coeffs;
    io.xyz = a->xyz; io.scales = a->scales; io.rot = a->rotations; io.opacity = a->opacity; io.shs = a->shs;
    io.o_xyz = a->out_xyz; io.o_scales = a->out_scales; io.o_rot = a->out_rotations; io.o_opacity = a->out_opacity; io.o_shs = a->out_shs;
    io.S_F = nullptr; io.S_H = nullptr; io.S_M = nullptr; io.Pp = 0;
    for (int h = 0; h < FDGS_NUM_HEADS; ++h) io.S_RZ[h] = nullptr;
    if (a->save_ws) {
        if (a->save_bytes < save_bytes(lo, a->P)) return FDGS_E_WORKSPACE;
        SaveLayout sl = make_save_layout(lo, a->P);
        float* S = (float*)a->save_ws;
        io.S_F = S + sl.off_F; io.S_H = S + sl.off_H;
        io.S_M = reinterpret_cast<uint32_t*>(S + sl.off_M); io.Pp = sl.Pp;
        for (int h = 0; h < FDGS_NUM_HEADS; ++h) if ((lo.head_mask >> h) & 1) io.S_RZ[h] = S + sl.off_RZ[h];
    }
    const int ntiles = (a->P + TP - 1) / TP;
    const int grid = ntiles < sm_count() ? ntiles : sm_count();
    const int passes = a->precision == 1 ? 1 : 3;
    if (lo.Wn == 128) {
        FDGS_CUDA(cudaFuncSetAttribute(k_deform_fwd_tc<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp.bytes));
        { FDGS_PROF("k_deform_fwd_tc", stream);
        k_deform_fwd_tc<128><<<grid, EpiCfg<128>::NTHREADS, sp.bytes, stream>>>(lo, plan, sp, net, io, packed, trow, ntiles, passes);
        }
    } else {
        FDGS_CUDA(cudaFuncSetAttribute(k_deform_fwd_tc<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp.bytes));
        { FDGS_PROF("k_deform_fwd_tc", stream);
        k_deform_fwd_tc<64><<<grid, EpiCfg<64>::NTHREADS, sp.bytes, stream>>>(lo, plan, sp, net, io, packed, trow, ntiles, passes);
        }
    }
    FDGS_LAUNCH_CHECK(a->debug, stream);
    return FDGS_OK;
}

bool supported(const Layout& lo) { return tc_supported(lo); }

}  // namespace fdgs_deform_tc
