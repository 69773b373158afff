/*
 * fdgs.h — C ABI of the B200-native 4D-Gaussian training hot path.
 *
 * This is the drop-in boundary.  Every entry point replaces one function of the
 * reference's torch/pybind glue (cited per function, paths relative to the
 * reference checkout; RAST = submodules/depth-diff-gaussian-rasterization).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes only; no torch / C++ types.
 *   - every `const float*` / `float*` is a DEVICE pointer to contiguous FP32
 *     unless the field name ends in `_host`.
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).
 *   - the library never allocates or frees device memory and never keeps a
 *     pointer past the call: outputs and workspaces are caller-owned.
 *   - return value: 0 = ok, <0 = FDGS_E_* argument/capacity error,
 *     >0 = cudaError_t raised by a launch (see fdgs_error_string).
 *   - `debug != 0` synchronises the stream after every launch and reports the
 *     first failing stage (reference: CHECK_CUDA, RAST/cuda_rasterizer/auxiliary.h:166-173).
 */
#ifndef FDGS_H_INCLUDED
#define FDGS_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FDGS_ABI_VERSION 1

enum {
    FDGS_OK = 0,
    FDGS_E_BADARG = -1,    /* null/inconsistent arguments (reference: AT_ERROR, RAST/rasterize_points.cu:57-59) */
    FDGS_E_WORKSPACE = -2, /* a caller-owned workspace is smaller than fdgs_*_bytes() says */
    FDGS_E_UNSUPPORTED = -3,
    FDGS_E_SMEM = -4,      /* configuration does not fit the 227 KB shared-memory budget */
    FDGS_E_TOO_LARGE = -5  /* problem size exceeds the fixed maximum of an allocation-free entry point (fdgs_*_static) */
};

/* ---- library info ------------------------------------------------------- */
int fdgs_abi_version(void);
const char* fdgs_error_string(int code);
/* sm_XY the library was compiled for (100 for sm_100a). */
int fdgs_compiled_arch(void);

/* Optional per-kernel CUDA-event timing (bench/roofline evidence).  enable(1) makes every kernel
 * launch of the library record a start/stop event pair on its stream; collect() synchronises the
 * device and writes one "kernel_name launches total_ms" line per kernel into buf, then clears.
 * fdgs_launch_count() = number of library kernels launched so far in this process.            */
/* The persistent deformation kernels run one CTA per SM and leave no room for a kernel of another stream.  n > 0 caps
 * their grids at n CTAs (e.g. SM count - 8 while a gradient all-reduce is in flight, fdgs.dp overlap); 0 = all SMs.  */
void fdgs_set_sm_limit(int n);
void fdgs_profile_enable(int on);
int fdgs_profile_collect(char* buf, size_t n);
unsigned long long fdgs_launch_count(void);

/* ---- rasterizer workspaces ----------------------------------------------
 * Replaces the three resize-callback byte buffers of the reference
 * (geomBuffer / binningBuffer / imgBuffer, RAST/rasterize_points.cu:27-33,74-79
 * and RAST/cuda_rasterizer/rasterizer_impl.cu:155-194).  They are typed,
 * caller-owned and must be 256-byte aligned.  geom/img/bin must be kept alive
 * (and unmodified) until the matching backward call.                        */
size_t fdgs_raster_geom_bytes(int P);
size_t fdgs_raster_img_bytes(int W, int H);
size_t fdgs_raster_bin_bytes(int64_t R);

typedef struct FdgsRasterForward {
    /* shapes / scalars (reference: RasterizeGaussiansCUDA args, RAST/rasterize_points.cu:35-56) */
    int32_t P, D, M, W, H;
    float tanfovx, tanfovy, scale_modifier;
    int32_t prefiltered, debug;
    /* inputs */
    const float* background;     /* [3] */
    const float* means3D;        /* [P,3] */
    const float* shs;            /* [P,M,3] or NULL */
    const float* colors_precomp; /* [P,3] or NULL */
    const float* opacities;      /* [P] */
    const float* scales;         /* [P,3] or NULL */
    const float* rotations;      /* [P,4] or NULL */
    const float* cov3D_precomp;  /* [P,6] or NULL */
    const float* viewmatrix;     /* [16], transposed world->view (scene/cameras.py:61-65) */
    const float* projmatrix;     /* [16] */
    const float* campos;         /* [3] */
    /* outputs */
    float* out_color;            /* [3,H,W] */
    float* out_depth;            /* [H,W] */
    int32_t* radii;              /* [P] */
    /* workspaces */
    void* geom_ws;  size_t geom_bytes;
    void* img_ws;   size_t img_bytes;
    void* bin_ws;   size_t bin_bytes;   /* only read by fdgs_raster_forward_render */
    /* pinned host scratch, 4 x int32: [0] = num_rendered R, [1] = most instances in one tile, [2] = tiles whose bucket is
     * above the shared-memory sort cap, [3] = 1 when R exceeded the capacity given to fdgs_raster_forward */
    int32_t* counts_host;
    /* optional parity/debug outputs (NULL to skip): sorted 64-bit tile|depth keys [R]
     * and sorted Gaussian ids [R], laid out exactly like the reference's
     * point_list_keys / point_list (RAST/cuda_rasterizer/rasterizer_impl.cu:181-194) */
    uint64_t* keys_out;
    uint32_t* point_list_out;
    /* 1 = `scales`, `rotations`, `opacities` are the RAW deformation outputs (log-scales, un-normalised quaternions, opacity
     * logits): the preprocess kernel applies the activations of gaussian_renderer/__init__.py:97-99 itself (exp,
     * F.normalize(dim=1, eps=1e-12), sigmoid - the same device functions fdgs_activate_forward uses, so the same bits), and the
     * activated [P,3] / [P,4] / [P,1] tensors never exist in memory.  Needs scales + rotations (not cov3D_precomp).          */
    int32_t activate;
} FdgsRasterForward;

/* Stage 1 of Rasterizer::forward (RAST/cuda_rasterizer/rasterizer_impl.cu:198-282):
 * per-Gaussian preprocess (cull, EWA projection, conic, radius, tile rect, SH->RGB),
 * per-tile instance histogram and tile ranges.  Synchronises `stream` once to
 * return num_rendered (the reference does the same blocking read at :282).
 * On return counts_host[0] = R.                                              */
int fdgs_raster_forward_plan(const FdgsRasterForward* a, void* stream);

/* Stage 2 (rasterizer_impl.cu:284-338): tile-bucketed key scatter, per-tile
 * segmented sort (== stable sort of tile|depth keys), record packing, and the
 * front-to-back blend.  `R` must be the value returned by the plan call and
 * bin_ws must hold fdgs_raster_bin_bytes(R).                                  */
int fdgs_raster_forward_render(const FdgsRasterForward* a, int64_t R, void* stream);

/* One-shot forward WITHOUT the blocking read of num_rendered (the reference stalls the host at rasterizer_impl.cu:282
 * because its binning buffer is sized from R; here the caller sizes bin_ws for `capacity` instances - an estimate, e.g.
 * 1.3 x the count of the previous render of this view - and every kernel of both stages is queued at once).  The counts
 * are copied to counts_host ASYNCHRONOUSLY: they are valid once the caller has synchronised with the stream (an event
 * recorded after this call).  If R > capacity the device empties every tile range (the queued scatter / sort / blend
 * kernels then touch nothing, the image is background) and sets counts_host[3] = 1: the caller must discard the result
 * and call again with a larger workspace (or use the plan / render pair).  geom_ws / img_ws / bin_ws go to
 * fdgs_raster_backward as usual, with FdgsRasterBackward.R = capacity (bin_ws is carved for `capacity`, not for R).      */
int fdgs_raster_forward(const FdgsRasterForward* a, int64_t capacity, void* stream);

/* Reads back the tile ranges (uint2 per tile, (0,0) for empty tiles exactly as
 * rasterizer_impl.cu:311-319 leaves them) into a device buffer — parity/debug. */
int fdgs_raster_copy_ranges(const void* img_ws, int W, int H, uint32_t* ranges_out, void* stream);

/* Diagnostic for the roofline of the blend kernels (bench only; not on the training path): after a forward call,
 * counts[0] = sum over pixels of n_contrib (pixel x instance pairs a per-pixel walk tests, RAST/cuda_rasterizer/
 * forward.cu:321-366), counts[1] = pairs that contribute (alpha >= 1/255 before saturation): what the backward
 * differentiates (backward.cu:459-555).  counts: 2 DEVICE uint64, zeroed by the call.                           */
int fdgs_raster_count_pairs(const void* img_ws, const void* bin_ws, int64_t R, int W, int H,
                            unsigned long long* counts, void* stream);

typedef struct FdgsRasterBackward {
    /* reference: RasterizeGaussiansBackwardCUDA args, RAST/rasterize_points.cu:119-142 */
    int32_t P, D, M, W, H;
    int64_t R;                   /* the instance count bin_ws was carved for: num_rendered after the plan / render pair,
                                    `capacity` after fdgs_raster_forward */
    float tanfovx, tanfovy, scale_modifier;
    int32_t debug;
    const float* background;
    const float* means3D;
    const float* shs;            /* or NULL */
    const float* colors_precomp; /* or NULL */
    const float* scales;         /* or NULL */
    const float* rotations;      /* or NULL */
    const float* cov3D_precomp;  /* or NULL */
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    const int32_t* radii;
    const void* geom_ws;
    const void* img_ws;
    const void* bin_ws;
    const float* dL_dpix;        /* [3,H,W] */
    /* outputs — fully written by the call (no pre-zeroing needed) */
    float* dL_dmeans2D;          /* [P,3]  (.z always 0, RAST/cuda_rasterizer/backward.cu:545-546) */
    float* dL_dconic;            /* [P,4]  scratch (dA,dB,unused,dC) */
    float* dL_dopacity;          /* [P] */
    float* dL_dcolors;           /* [P,3]  (this IS grad_colors_precomp when colors_precomp != NULL) */
    float* dL_dmeans3D;          /* [P,3] */
    float* dL_dcov3D;            /* [P,6]  (this IS grad_cov3Ds_precomp when cov3D_precomp != NULL) */
    float* dL_dsh;               /* [P,M,3] or NULL when M == 0 */
    float* dL_dscales;           /* [P,3] */
    float* dL_drots;             /* [P,4] */
    /* same flag as the forward call's: `scales` / `rotations` are the raw values, and dL_dscales / dL_drots / dL_dopacity come
     * back as gradients w.r.t. the RAW log-scales / quaternions / logits (the activations' VJPs are applied in the kernel) */
    int32_t activate;
} FdgsRasterBackward;

/* Rasterizer::backward (RAST/cuda_rasterizer/rasterizer_impl.cu:343-437):
 * reverse-order blend backward with warp-reduced accumulation, then the fused
 * cov2D / mean / SH / cov3D chain (backward.cu:144-396).                     */
int fdgs_raster_backward(const FdgsRasterBackward* a, void* stream);

/* Rasterizer::markVisible (rasterizer_impl.cu:141-153): present[i] = z_view > 0.2 */
int fdgs_mark_visible(int P, const float* means3D, const float* viewmatrix,
                      const float* projmatrix, uint8_t* present, void* stream);

/* ---- HexPlane deformation field -----------------------------------------
 * Replaces deform_network.forward -> Deformation.forward_dynamic
 * (scene/deformation.py:185-212, 97-148) with HexPlaneField.forward
 * (scene/hexplane.py:160-183) fused in.  All P points share one time value. */
#define FDGS_MAX_LEVELS 4
#define FDGS_HEAD_POS 0
#define FDGS_HEAD_SCALE 1
#define FDGS_HEAD_ROT 2
#define FDGS_HEAD_OPACITY 3
#define FDGS_HEAD_SHS 4
#define FDGS_NUM_HEADS 5

typedef struct FdgsDeformNet {
    int32_t levels;                 /* L = len(multires)            (scene/hexplane.py:131-151) */
    int32_t channels;               /* C = output_coordinate_dim                                  */
    int32_t width;                  /* Wn = net_width                (scene/deformation.py:45-65) */
    int32_t depth;                  /* D = defor_depth: number of trunk Linear layers = max(D,1)  */
    int32_t head_mask;              /* bit h set = head h active (= !no_dx/!no_ds/!no_dr/!no_do/!no_dshs) */
    int32_t apply_rotation;         /* scene/deformation.py:127-128 */
    int32_t res[FDGS_MAX_LEVELS][4];/* per level [rx,ry,rz,rt]                                    */
    float aabb[6];                  /* aabb[0..2] = "max", aabb[3..5] = "min" (scene/hexplane.py:118-119) */
    /* plane k of level l: [C, H_k, W_k] row-major, pair order (0,1),(0,2),(0,3),(1,2),(1,3),(2,3);
     * W_k = res[first], H_k = res[second]  (scene/hexplane.py:58-63) */
    const float* planes[FDGS_MAX_LEVELS][6];
    /* trunk: feature_out.0 (weight [Wn, C*L], bias [Wn]) then (depth-1) x (ReLU, Linear Wn->Wn) */
    const float* trunk_w[8];
    const float* trunk_b[8];
    /* heads: *_deform.1 (weight [Wn,Wn], bias [Wn]) and *_deform.3 (weight [out,Wn], bias [out]) */
    const float* head_w1[FDGS_NUM_HEADS];
    const float* head_b1[FDGS_NUM_HEADS];
    const float* head_w2[FDGS_NUM_HEADS];
    const float* head_b2[FDGS_NUM_HEADS];
} FdgsDeformNet;

typedef struct FdgsDeformForward {
    FdgsDeformNet net;
    int32_t P;
    int32_t sh_coeffs;              /* second dim of shs (16, or 1 with broadcast)   */
    float time;                     /* the view's timestamp (times_sel[:,0] is constant per view)   */
    const float* time_dev;          /* optional DEVICE scalar overriding `time` (avoids a host sync) */
    int32_t precision;              /* 0 = FP32-equivalent (tcgen05 3xTF32 split), 1 = single-pass TF32 (faster, ~1e-3),
                                       2 = FP32 CUDA-core kernels (no tensor cores)                   */
    int32_t planes_packed_valid;    /* 1 = packed_ws already holds the packed planes of these parameters */
    const float* xyz;               /* [P,3] */
    const float* scales;            /* [P,3] */
    const float* rotations;         /* [P,4] */
    const float* opacity;           /* [P,1] */
    const float* shs;               /* [P,sh_coeffs,3] */
    /* outputs; a NULL pointer means "head inactive: caller returns its input unchanged".
     * out = input + delta, except out_rotations = delta alone when net.apply_rotation
     * (the caller then applies the quaternion product, scene/deformation.py:127-128). */
    float* out_xyz;
    float* out_scales;
    float* out_rotations;
    float* out_opacity;
    float* out_shs;                 /* [P,16,3] */
    void* packed_ws; size_t packed_bytes; /* fdgs_deform_packed_bytes(): channel-last planes; reusable across
                                             views until the plane parameters change                     */
    void* trow_ws;   size_t trow_bytes;   /* fdgs_deform_trow_bytes(): this view's pre-blended time rows;
                                             must stay alive until the matching backward              */
    void* wprep_ws;  size_t wprep_bytes;  /* fdgs_deform_wprep_bytes(): MLP weights split hi/lo and tiled for the
                                             tensor cores; rewritten by every call (scratch)          */
    void* save_ws; size_t save_bytes;     /* optional, training: fdgs_deform_forward_save_bytes(net, P) bytes that receive
                                             the HexPlane features and the hidden activations relu(h), relu(z_j) for
                                             fdgs_deform_backward (saved_ws); NULL = keep nothing (inference)        */
    int32_t debug;
} FdgsDeformForward;

size_t fdgs_deform_packed_bytes(const FdgsDeformNet* net);
size_t fdgs_deform_trow_bytes(const FdgsDeformNet* net);
size_t fdgs_deform_wprep_bytes(const FdgsDeformNet* net);
size_t fdgs_deform_forward_save_bytes(const FdgsDeformNet* net, int P);   /* 0 = this net/precision keeps nothing */
int fdgs_deform_forward(const FdgsDeformForward* a, void* stream);

typedef struct FdgsDeformBackward {
    FdgsDeformNet net;
    int32_t P;
    float time;
    const float* time_dev;          /* optional DEVICE scalar overriding `time` */
    int32_t precision;
    const float* xyz;
    /* upstream grads w.r.t. the five outputs (NULL = head inactive / no grad) */
    const float* g_xyz;             /* [P,3] */
    const float* g_scales;          /* [P,3] */
    const float* g_rotations;       /* [P,4] */
    const float* g_opacity;         /* [P,1] */
    const float* g_shs;             /* [P,16,3] */
    /* outputs: grad wrt xyz THROUGH THE GRID only (identity paths are added by the caller),
     * and parameter grads ACCUMULATED (+=) into caller-owned buffers laid out like the params */
    float* d_xyz_grid;              /* [P,3] */
    float* d_planes[FDGS_MAX_LEVELS][6];
    float* d_trunk_w[8];
    float* d_trunk_b[8];
    float* d_head_w1[FDGS_NUM_HEADS];
    float* d_head_b1[FDGS_NUM_HEADS];
    float* d_head_w2[FDGS_NUM_HEADS];
    float* d_head_b2[FDGS_NUM_HEADS];
    const void* packed_ws;          /* the forward call's packed planes */
    const void* trow_ws;            /* the forward call's time rows of this view */
    void* ws; size_t ws_bytes;      /* fdgs_deform_backward_ws_bytes(): packed plane-grad scratch */
    void* scratch; size_t scratch_bytes; /* fdgs_deform_backward_scratch_bytes(net, P): FP32 activations exchanged between
                                       the two tensor-core backward kernels + formatted weights; contents are dead
                                       after the call (reuse it across views).  Required unless precision == 2 or the
                                       function returns 0 for this net (then the CUDA-core kernel runs).            */
    const void* saved_ws; size_t saved_bytes; /* the forward call's save_ws (same net, same P, same xyz); NULL = the
                                       backward recomputes the activations itself                                  */
    int32_t debug;
} FdgsDeformBackward;

size_t fdgs_deform_backward_ws_bytes(const FdgsDeformNet* net);
size_t fdgs_deform_backward_scratch_bytes(const FdgsDeformNet* net, int P);
int fdgs_deform_backward(const FdgsDeformBackward* a, void* stream);

/* ---- activations between the two operators (caller-side glue) -----------------------------------
 * Replaces the three activation calls of gaussian_renderer/__init__.py:97-99 on the deformation outputs
 *   scales = exp(scales_raw), rotations = F.normalize(rotations_raw, dim=1, eps=1e-12), opacity = sigmoid(opacity_raw)
 * (scene/gaussian_model.py:47-58 binds them as scaling_activation / rotation_activation / opacity_activation) and their
 * ~10 autograd kernels with one kernel each way.  All pointers: device, contiguous FP32, P rows.                        */
typedef struct {
    int32_t P;
    const float* scales_raw;        /* [P,3] */
    const float* rotations_raw;     /* [P,4] */
    const float* opacity_raw;       /* [P,1] */
    float* scales;                  /* [P,3] out */
    float* rotations;               /* [P,4] out */
    float* opacity;                 /* [P,1] out */
} FdgsActivateForward;
int fdgs_activate_forward(const FdgsActivateForward* a, void* stream);

typedef struct {
    int32_t P;
    const float* rotations_raw;     /* [P,4] forward input */
    const float* scales;            /* [P,3] forward outputs */
    const float* rotations;         /* [P,4] */
    const float* opacity;           /* [P,1] */
    const float* g_scales;          /* upstream grads w.r.t. the activated values; NULL = none */
    const float* g_rotations;
    const float* g_opacity;
    float* d_scales_raw;            /* [P,3] out (written, not accumulated); NULL with a NULL upstream grad */
    float* d_rotations_raw;         /* [P,4] */
    float* d_opacity_raw;           /* [P,1] */
} FdgsActivateBackward;
int fdgs_activate_backward(const FdgsActivateBackward* a, void* stream);

/* ---- L1 photometric loss of the training step (caller-side glue) ---------------------------------------------------------
 * utils/loss_utils.py l1_loss(network_output, gt) = mean |a - b| (train.py's Ll1), with its autograd:
 *   forward : *loss += scale * sum_i |a_i - b_i|          (*loss is a DEVICE scalar the caller zeroed; scale = 1 / n, or
 *                                                           1 / (n * views) for one view's share of a batch loss)
 *   backward: d_a_i = *g * scale * sign(a_i - b_i)        (*g = DEVICE scalar upstream gradient; sign(0) = 0 as in ATen)   */
int fdgs_l1_loss_forward(const float* a, const float* b, long long n, float scale, float* loss, void* stream);
int fdgs_l1_loss_backward(const float* a, const float* b, long long n, float scale, const float* g, float* d_a, void* stream);

/* ---- HexPlane regularisers ----------------------------------------------------------------------------
 * scene/gaussian_model.py:538-577 compute_regulation = plane_tv_weight * _plane_regulation() + time_smoothness_weight *
 * _time_regulation() + l1_time_planes_weight * _l1_regulation(), with scene/regulation.py:22-28 compute_plane_smoothness
 * (mean squared second difference along a plane's H axis) on planes {0,1,3} (spatial) and {2,4,5} (space-time) of every
 * level, and mean |1 - plane| on the space-time planes.  One kernel computes the value AND adds its gradient:
 *   *loss      += the regulariser value                                  (DEVICE scalar, caller-zeroed; NULL = skip)
 *   d_planes   += (*grad_scale) * d value / d plane                      (NULL entry = no gradient for that plane;
 *                                                                         grad_scale NULL = 1, as train.py:251-254 adds
 *                                                                         tv_loss to the loss with weight 1)
 * Plane k of level l is [channels, H_k, W_k] with W_k = res[l][first], H_k = res[l][second] of the pair order
 * (0,1),(0,2),(0,3),(1,2),(1,3),(2,3) — the layout FdgsDeformNet.planes uses.                                            */
typedef struct FdgsPlaneReg {
    int32_t levels, channels;
    int32_t res[FDGS_MAX_LEVELS][4];
    const float* planes[FDGS_MAX_LEVELS][6];
    float* d_planes[FDGS_MAX_LEVELS][6];
    float plane_tv_weight, time_smoothness_weight, l1_time_planes_weight;
    const float* grad_scale;
    float* loss;
} FdgsPlaneReg;
int fdgs_plane_regulation(const FdgsPlaneReg* a, void* stream);
/* compute_plane_smoothness (scene/regulation.py:22-28) of ONE plane [C,H,W]: *loss += weight * mean(second difference^2),
 * d_plane += (*grad_scale) * its gradient (either may be NULL).                                                         */
int fdgs_plane_smoothness(const float* plane, float* d_plane, int C, int H, int W, float weight, const float* grad_scale,
                          float* loss, void* stream);

/* ---- fused Adam + densification statistics -----------------------------------------------------------
 * torch.optim.Adam(l, lr=0.0, eps=1e-15) as scene/gaussian_model.py:172-184 builds it (betas (0.9, 0.999), no weight
 * decay, no amsgrad), one launch for any number of tensors, each with its own lr and step count (param groups xyz,
 * deformation, grid, f_dc, f_rest, opacity, scaling, rotation).  The FP32 operation order is that of torch's foreach
 * implementation, so parameters and both moments are BIT-IDENTICAL to torch.optim.Adam.step() on the same inputs.
 * `step` is the step count AFTER this update (>= 1).  zero_grad != 0 also clears the gradients in the same pass.          */
typedef struct FdgsAdamTensor {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int64_t n;
    double lr;                      /* the group's lr as the Python double torch would divide by the bias correction */
    int64_t step;
} FdgsAdamTensor;
typedef struct FdgsAdamStep {
    int32_t ntensors;
    int32_t zero_grad;
    double beta1, beta2, eps;
    const FdgsAdamTensor* t;        /* HOST array of ntensors entries */
} FdgsAdamStep;
int fdgs_adam_step(const FdgsAdamStep* a, void* stream);

/* train.py:346-349 + scene/gaussian_model.py:521-523: for the visible Gaussians of an iteration (visible[i] != 0, or
 * radii[i] > 0 when visible is NULL)   max_radii2D[i] = max(max_radii2D[i], radii[i]),
 * xyz_gradient_accum[i] += || viewspace_grad[i, :2] ||,  denom[i] += 1.   Any of the three outputs may be NULL.           */
int fdgs_densify_stats(int P, const int32_t* radii, const uint8_t* visible, const float* viewspace_grad /* [P,3] */,
                       float* max_radii2D, float* xyz_gradient_accum, float* denom, void* stream);

/* ---- init-time native pieces ---------------------------------------------------------------------------
 * simple_knn's distCUDA2 (submodules/simple-knn/simple_knn.cu:65-215; scene/gaussian_model.py:145 initialises the scales
 * from it): mean_dist2[i] = mean of the squared distances from point i to its 3 nearest other points.  Exact 3-NN on a
 * uniform hash grid (the reference prunes Morton-sorted boxes; the three distances, their expression and the mean are the
 * same, so the result is bit-identical).  ws: fdgs_knn_ws_bytes(P) bytes of device scratch.                               */
size_t fdgs_knn_ws_bytes(int P);
int fdgs_knn_mean_dist2(int P, const float* points /* [P,3] */, float* mean_dist2 /* [P] */, void* ws, size_t ws_bytes, void* stream);

/* Allocation-free variant of fdgs_knn_mean_dist2.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, same output (bit-identical to fdgs_knn_mean_dist2), but the scratch is a file-scope
 * `static __device__` array of the library instead of a caller workspace: the call performs no cudaMalloc / cudaFree and
 * calls no allocating library routine.
 *   - limit: P <= FDGS_KNN_STATIC_MAX_POINTS (= fdgs_knn_static_max_points()); a larger P returns FDGS_E_TOO_LARGE and
 *     launches nothing.  The array is 3 689 728 bytes and is held by every GPU that loads the library until exit.
 *   - the array is SHARED by all calls on a device and keeps the last call's contents (the call clears what must start at
 *     zero): two calls that overlap on the GPU (different streams, or different host threads) RACE - serialise them.
 *   - the device to use is the CURRENT device; `points` / `mean_dist2` are device pointers on it.                        */
#define FDGS_KNN_STATIC_MAX_POINTS 262144
int fdgs_knn_static_max_points(void);
int fdgs_knn_mean_dist2_static(int P, const float* points /* [P,3] */, float* mean_dist2 /* [P] */, void* stream);

/* Instant4D grid pruning (utils/grid_pruning.py:16-41 -> open3d PointCloud::VoxelDownSample): voxel index =
 * floor((p - (min_bound - voxel_size / 2)) / voxel_size) per axis, one output point per occupied voxel = the mean position
 * (and mean colour, when colors != NULL) of its points, accumulated in double like open3d.  The order of the output rows is
 * unspecified (open3d iterates an unordered_map).  *n_out (DEVICE int32) receives the number of occupied voxels; at most
 * max_out rows are written.  ws: fdgs_voxel_downsample_ws_bytes(P) bytes of device scratch.                             */
size_t fdgs_voxel_downsample_ws_bytes(int P);
int fdgs_voxel_downsample(int P, const float* points, const float* colors, double voxel_size, float* out_points,
                          float* out_colors, int max_out, int32_t* n_out, void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FDGS_H_INCLUDED */
