/*
 * d3g.h — C ABI of libd3g.so, the B200-native (sm_100a) hot path of Deformable-3D-Gaussians:
 * time-conditioned deformation MLP -> differentiable tile-based Gaussian rasterizer, fwd + bwd.
 *
 * This is the drop-in boundary.  Every entry point below replaces one piece of what the reference's
 * Python reaches through torch / its CUDA extension; reference citations are relative to
 * /root/reference.  The reference side binds these with ctypes; the in-tree
 * host mirror lives in deformable-3d-gaussians_b200/{rasterizer,deform,renderer}.py.
 *
 * Conventions
 *  - plain C: pointers, sizes, POD structs.  No torch / C++ types.  All `float*`/`void*` data
 *    pointers are DEVICE pointers unless the name ends in `_host`.
 *  - all work is enqueued on the caller's `stream` (a cudaStream_t passed as void*); no hidden
 *    synchronisation, no global state; safe from one thread per GPU.
 *  - every function returns 0 on success or a negative d3g_status; d3g_last_error() returns a
 *    thread-local message.  No exceptions, exit(), abort() or device traps cross the boundary.
 *  - tensors are fp32, contiguous, row-major with the shapes of the reference call sites.
 */
#ifndef D3G_H_
#define D3G_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum d3g_status {
  D3G_OK = 0,
  D3G_ERR_INVALID = -1,  /* bad argument combination (mirrors the reference's Python exceptions) */
  D3G_ERR_CUDA = -2,     /* a CUDA runtime call or launch failed; message has cudaGetErrorString */
  D3G_ERR_CAPACITY = -3, /* workspace too small for the request */
  D3G_ERR_UNSUPPORTED = -4
} d3g_status;

const char* d3g_last_error(void);
/* ABI version, bumped on any signature change. */
int d3g_version(void);

/* Optional per-kernel timing for benchmarks: while enabled, CUDA events are recorded on the caller's
 * stream around every kernel of the path.  d3g_prof_collect synchronises the device, adds up the
 * elapsed time and the launch count per kernel id (arrays of at least d3g_prof_count() entries) and
 * resets the recorder.  Process-global, not thread-safe; leave it off on the hot path. */
int d3g_prof_enable(int32_t on);
int d3g_prof_count(void);
const char* d3g_prof_name(int32_t id);
int d3g_prof_collect(double* total_ms, int32_t* launches, int32_t n);

/* ------------------------------------------------------------------------------------------------
 * Rasterizer.  Replaces diff_gaussian_rasterization.{GaussianRasterizationSettings,GaussianRasterizer}
 * as constructed at gaussian_renderer/__init__.py:53-68 and invoked at :115-124 (returning image,
 * radii, depth).  Source of the reference extension is absent (.SUBMODULES.json:9-15).
 * ---------------------------------------------------------------------------------------------- */

/* Field-for-field the NamedTuple built at gaussian_renderer/__init__.py:53-66.  bg / viewmatrix /
 * projmatrix / campos stay DEVICE tensors exactly as the reference passes them (no D2H copy). */
typedef struct d3g_raster_settings {
  int32_t image_height;
  int32_t image_width;
  float tanfovx;
  float tanfovy;
  const float* bg;         /* [3]   device */
  float scale_modifier;
  const float* viewmatrix; /* [4,4] device, world_view_transform as stored (transposed) */
  const float* projmatrix; /* [4,4] device, full_proj_transform as stored */
  int32_t sh_degree;       /* active degree, 0..3 */
  const float* campos;     /* [3]   device */
  int32_t prefiltered;     /* accepted, ignored (the reference always passes False, :64) */
  int32_t debug;           /* accepted; when non-zero every launch is followed by a stream sync + check */
  int32_t densify_mode;    /* semantics of means2D_densify.grad: 1 = same as means2D.grad,
                              2 = component-wise sum of |per-pixel term| (default), 3 = sum of per-pixel
                              2-norms in .x.  See DESIGN.md "open semantic #1". */
  int32_t binning_mode;    /* 0 (default): per-tile binning — count, scan over tiles, scatter, one shared-memory sort
                              per tile; 1: the reference's scheme — duplicateWithKeys + one 64-bit radix sort +
                              identifyTileRanges.  Sorted keys, point list and tile ranges are bit-identical; the
                              unsorted key/value lists and the per-Gaussian offsets exist only in mode 1. */
} d3g_raster_settings;

/* Workspace sizes (bytes).  geom: per-Gaussian state; img: per-pixel + per-tile state;
 * bin: key/value double buffers + sort scratch for `pair_capacity` (tile,Gaussian) pairs. */
int d3g_raster_geom_bytes(int32_t N, size_t* bytes);
int d3g_raster_img_bytes(int32_t H, int32_t W, size_t* bytes);
int d3g_raster_bin_bytes(uint64_t pair_capacity, size_t* bytes);

/* Stage 1 of forward (preprocessCUDA + inclusive scan of tiles_touched).  Exactly one of shs /
 * colors_precomp and exactly one of (scales,rotations) / cov3D_precomp must be non-NULL, as the
 * reference's wrapper enforces.  shs is [N, sh_coeffs, 3].  radii is the [N] int32 output tensor.
 * The number of (tile,Gaussian) pairs R is left in the geom header on the device and, when
 * num_rendered_host (pinned) is non-NULL, copied there asynchronously on `stream`. */
int d3g_raster_preprocess(const d3g_raster_settings* s, int32_t N, const float* means3D,
                          const float* shs, int32_t sh_coeffs, const float* colors_precomp,
                          const float* opacities, const float* scales, const float* rotations,
                          const float* cov3D_precomp, void* geom, int32_t* radii,
                          uint32_t* num_rendered_host, void* stream);

/* Stage 2 of forward (duplicateWithKeys -> 64-bit tile|depth radix sort -> identifyTileRanges ->
 * renderCUDA).  pair_capacity >= R; R itself is read on the device.  If R > pair_capacity the
 * overflow flag readable via d3g_raster_read_header is set and the image is left undefined.
 * out_color [3,H,W], out_depth [1,H,W]. */
int d3g_raster_render(const d3g_raster_settings* s, int32_t N, uint64_t pair_capacity, void* geom,
                      void* bin, void* img, const int32_t* radii, float* out_color, float* out_depth,
                      void* stream);

/* Backward (renderCUDA bwd -> computeCov2D bwd -> preprocess bwd).  Consumes the workspaces saved by
 * forward.  dL_ddepth may be NULL (the reference never back-propagates through depth).  Any output
 * pointer that does not apply to the chosen input mode may be NULL.  dL_dmeans2D / _densify are
 * [N,3] (xy populated, z = 0) because scene/gaussian_model.py:398-401 reads [:, :2].  All outputs
 * are fully written (no pre-zeroing needed). */
int d3g_raster_backward(const d3g_raster_settings* s, int32_t N, uint64_t pair_capacity,
                        const float* means3D, const float* shs, int32_t sh_coeffs,
                        const float* colors_precomp, const float* scales, const float* rotations,
                        const float* cov3D_precomp, const int32_t* radii, void* geom, void* bin,
                        void* img, const float* dL_dcolor, const float* dL_ddepth,
                        float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dmeans2D_densify,
                        float* dL_dshs, float* dL_dcolors_precomp, float* dL_dopacities,
                        float* dL_dscales, float* dL_drotations, float* dL_dcov3D_precomp,
                        void* stream);

/* The same two calls with the spherical-harmonics coefficients left in the reference's two parameters instead of
 * the [N,16,3] tensor its `get_features` concatenates every frame (scene/gaussian_model.py:111-115:
 * cat(_features_dc [N,1,3], _features_rest [N,15,3]); gaussian_renderer/__init__.py:112 passes the result as `shs`):
 * the forward reads both tensors in place (no 19 MB copy per frame at 100k Gaussians), and the backward ADDS
 * dL/d_features_dc and dL/d_features_rest to grad_features_dc / grad_features_rest — the parameters' .grad tensors,
 * which must hold the running sums already (zeros before the first frame) — instead of writing a [N,16,3] tensor
 * that autograd then splits and adds.  sh_coeffs must be 16 (max_sh_degree 3); features_rest and grad_features_rest
 * must be 16-byte aligned.  Colours, clamp flags and every other gradient are bit-identical to the calls above. */
int d3g_raster_preprocess_split(const d3g_raster_settings* s, int32_t N, const float* means3D,
                                const float* features_dc, const float* features_rest, int32_t sh_coeffs,
                                const float* opacities, const float* scales, const float* rotations,
                                const float* cov3D_precomp, void* geom, int32_t* radii,
                                uint32_t* num_rendered_host, void* stream);
int d3g_raster_backward_split(const d3g_raster_settings* s, int32_t N, uint64_t pair_capacity,
                              const float* means3D, const float* features_dc, const float* features_rest,
                              int32_t sh_coeffs, const float* scales, const float* rotations,
                              const float* cov3D_precomp, const int32_t* radii, void* geom, void* bin,
                              void* img, const float* dL_dcolor, const float* dL_ddepth,
                              float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dmeans2D_densify,
                              float* grad_features_dc, float* grad_features_rest, float* dL_dopacities,
                              float* dL_dscales, float* dL_drotations, float* dL_dcov3D_precomp,
                              void* stream);

/* Activation + deformation of the Gaussian parameters, fused (the elementwise work of
 * gaussian_renderer/__init__.py:69-97 and of the getters of scene/gaussian_model.py:58-78):
 *   means3D = xyz + d_xyz, or from_homogenous(d_xyz[N,4,4] . [xyz,1]) when is_6dof (utils/rigid_utils.py:86-107)
 *   opacity_out = sigmoid(opacity); scales_out = exp(scaling) + d_scaling; rotations_out =
 *   normalize(rotation) + d_rotation (d_scaling / d_rotation may be NULL = no delta).
 * The backward call turns the cotangents of the four outputs (any may be NULL = zero) into the
 * cotangents of the raw parameters (any output pointer may be NULL = not wanted); the cotangents of the
 * additive deltas are the incoming cotangents themselves, g_dxyz_6dof [N,4,4] is written only when is_6dof,
 * and g_xyz is written only when is_6dof (otherwise it equals g_means3D). */
int d3g_gaussians_prepare(int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                          const float* scaling, const float* d_scaling, const float* rotation,
                          const float* d_rotation, const float* opacity, float* means3D,
                          float* opacity_out, float* scales_out, float* rotations_out, void* stream);
int d3g_gaussians_prepare_backward(int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                                   const float* scaling, const float* rotation, const float* opacity,
                                   const float* g_means3D, const float* g_opacity_out,
                                   const float* g_scales_out, const float* g_rotations_out,
                                   float* g_xyz, float* g_dxyz_6dof, float* g_opacity,
                                   float* g_scaling, float* g_rotation, void* stream);
/* Same, but the four parameter cotangents are ADDED to acc_xyz / acc_opacity / acc_scaling / acc_rotation (the
 * parameters' .grad tensors holding the running sums; any may be NULL = not wanted) instead of being written to
 * fresh tensors that autograd then adds to .grad one launch each.  acc_xyz receives g_means3D itself when not
 * is_6dof.  g_dxyz_6dof (the cotangent handed on to the deformation network) is still written, not added. */
int d3g_gaussians_prepare_backward_into(int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                                        const float* scaling, const float* rotation, const float* opacity,
                                        const float* g_means3D, const float* g_opacity_out,
                                        const float* g_scales_out, const float* g_rotations_out,
                                        float* acc_xyz, float* g_dxyz_6dof, float* acc_opacity,
                                        float* acc_scaling, float* acc_rotation, void* stream);

/* Fused image loss of the training step (utils/loss_utils.py:18-19 l1_loss, :42-73 ssim with its 11x11 sigma-1.5
 * window, zero padding and C1/C2; combined as in train_baseline.py:126-127).  img, gt: [C,H,W] fp32 device.
 * out3 (device float[3]) receives {mean|img-gt|, mean SSIM, w_l1*L1 + w_ssim*SSIM + bias}; the reference's loss
 * is w_l1 = 1-lambda, w_ssim = -lambda, bias = lambda.  dL_dimg (nullable, [C,H,W]) receives the gradient of
 * out3[2] w.r.t. img.  workspace: d3g_image_loss_workspace_bytes(C,H,W) bytes of device memory. */
int d3g_image_loss_workspace_bytes(int32_t C, int32_t H, int32_t W, size_t* bytes);
int d3g_image_loss(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float w_l1, float w_ssim,
                   float bias, void* workspace, float* out3, float* dL_dimg, void* stream);
/* The gradient alone, from the workspace a d3g_image_loss call on the same img/gt left behind: dL_dimg =
 * upstream * d(out3[2])/d(img); upstream is a device scalar (the cotangent autograd hands to backward) or NULL = 1. */
int d3g_image_loss_backward(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float w_l1,
                            float w_ssim, const void* workspace, const float* upstream, float* dL_dimg,
                            void* stream);
/* d3g_image_loss for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same inputs,
 * outputs and results as d3g_image_loss, but the workspace is a static device array inside the library (24.1 MiB, held
 * by every GPU that loads the library until the process exits), so the call performs no cudaMalloc / cudaFree and needs
 * no allocator from its caller.  Limits: C*H*W <= 2^21 elements (3 x 800 x 800 fits) and
 * C * ceil(H/16) * ceil(W/32) <= 2^14 pixel blocks; larger shapes return D3G_ERR_CAPACITY with a message naming the
 * limit (d3g_image_loss_static_limits reports both limits and the array size).  All calls on a device share the one
 * array: calls that overlap on the GPU (different streams, or a graph replay concurrent with an eager call) race on it
 * — serialise them.  The array keeps the previous call's contents, and nothing in it has to start at zero (every word
 * a kernel reads is written earlier in the same call).  There is no static counterpart of d3g_image_loss_backward:
 * pass dL_dimg here to get the gradient in the same call. */
int d3g_image_loss_static_limits(size_t* max_elements, size_t* max_blocks, size_t* scratch_bytes);
int d3g_image_loss_static(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float w_l1,
                          float w_ssim, float bias, float* out3, float* dL_dimg, void* stream);

/* Fused multi-tensor Adam step: what `optimizer.step()` of scene/gaussian_model.py:136 and
 * scene/deform_model.py:295,337 (torch.optim.Adam, betas (0.9, 0.999), eps 1e-15, optional weight_decay, no amsgrad)
 * does for every parameter, in one launch.  betas / eps are the Python doubles torch uses (1 - beta and the bias
 * corrections are formed in double).  `step` is the 1-based step count AFTER this update (torch increments
 * before use); all pointers are fp32 device arrays of `numel` elements, updated in place. */
typedef struct d3g_adam_tensor {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  int64_t numel;
  float lr;
  float weight_decay;
  int64_t step;
} d3g_adam_tensor;
int d3g_adam_step(int32_t n_tensors, const d3g_adam_tensor* tensors, double beta1, double beta2, double eps,
                  void* stream);

/* Densification bookkeeping of one rendered frame in one launch: for every Gaussian with radii > 0 (the
 * `visibility_filter` of gaussian_renderer/__init__.py:131)
 *   max_radii2D = max(max_radii2D, radii)                                     (train.py:238-239)
 *   xyz_gradient_accum += ||viewspace_grad[:, :2]||, denom += 1               (scene/gaussian_model.py:398-401)
 * viewspace_grad: [N,3] (the .grad of `viewspace_points_densify` after backward) or NULL to update max_radii2D only;
 * xyz_gradient_accum / denom: [N,1] fp32; max_radii2D: [N] fp32 or NULL. */
int d3g_densification_stats(int32_t N, const float* viewspace_grad, const int32_t* radii,
                            float* xyz_gradient_accum, float* denom, float* max_radii2D, void* stream);

/* `simple_knn._C.distCUDA2` (submodules/simple-knn/spatial.cu:15-26 -> SimpleKNN::knn, simple_knn.cu:185-221; caller
 * scene/gaussian_model.py:105): mean_dist2[i] = mean of the three smallest squared distances from points[i] to the
 * points at the OTHER indices (exact; FLT_MAX stands in for a missing neighbour when P < 4, as in the reference).
 * points: [P,3] fp32 device, mean_dist2: [P] fp32 device.  Asynchronous on `stream`, no host round-trip, no allocation. */
int d3g_knn_mean_dist2(int32_t P, const float* points, float* mean_dist2, void* stream);
/* The same result (bit for bit) for large clouds: exact search over a uniform grid built on the device (bounding box,
 * counting sort by cell, shell-by-shell query) instead of all P^2 pairs.  workspace: d3g_knn_workspace_bytes(P) bytes
 * of device memory, contents irrelevant on entry.  Asynchronous on `stream`, no host round-trip. */
int d3g_knn_workspace_bytes(int32_t P, size_t* bytes);
int d3g_knn_mean_dist2_grid(int32_t P, const float* points, float* mean_dist2, void* workspace,
                            size_t workspace_bytes, void* stream);

/* Upstream `markVisible` (never called by the reference; provided for API completeness).
 * present[i] = 1 iff view-space z > 0.2. */
int d3g_mark_visible(int32_t N, const float* means3D, const float* viewmatrix, uint8_t* present,
                     void* stream);

/* Debug / parity access to the opaque workspaces (device -> host copies, synchronous on `stream`).
 * which: 0 means2D[N,2] f32 | 1 depths[N] f32 | 2 conic_opacity[N,4] f32 | 3 rgb[N,3] f32 |
 *        4 cov3D[N,6] f32 | 5 tiles_touched[N] u32 | 6 offsets[N] u32 | 7 clamped[N,3] u8 |
 *        8 sorted keys[R] u64 | 9 point_list[R] u32 | 10 ranges[tiles,2] u32 | 11 final_T[H*W] f32 |
 *        12 n_contrib[H*W] u32 | 13 unsorted keys[R] u64 | 14 unsorted values[R] u32
 * header_host receives {R, overflow}. */
int d3g_raster_read_header(const void* geom, uint32_t header_host[2], void* stream);
/* Same copy of {R, overflow}, enqueued on `stream` without synchronising (header_pinned must be pinned host
 * memory): lets a caller that sized pair_capacity ahead of time verify it later, off the critical path. */
int d3g_raster_header_async(const void* geom, uint32_t* header_pinned, void* stream);
int d3g_raster_read_state(int32_t which, int32_t N, int32_t H, int32_t W, uint64_t pair_capacity,
                          const void* geom, const void* bin, const void* img, void* dst_host,
                          size_t dst_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Deformation MLP.  Replaces DeformNetwork / DeformNetworkBaseline.forward
 * (utils/time_utils.py:102-127, :176-201) incl. the is_6dof se(3) branch (:114-121,
 * utils/rigid_utils.py:60-83), as called by DeformModel.step / DeformModelBaseline.step
 * (scene/deform_model.py:260-273, :323-324).  D=8, W=256, multires=10 (the only configuration the
 * reference instantiates, arguments/__init__.py:52-73); other shapes return D3G_ERR_UNSUPPORTED.
 * ---------------------------------------------------------------------------------------------- */
typedef struct d3g_mlp_config {
  int32_t is_blender; /* timenet 13->256->30 and t_multires=6 if set, else t_multires=10 no timenet */
  int32_t is_6dof;    /* branch_w / branch_v + exp_se3 instead of gaussian_warp */
  int32_t heads;      /* bit0: warp (always), bit1: rotation head, bit2: scaling head.
                         DeformNetworkBaseline = 7, DeformNetwork = 1 */
  int32_t D, W, multires;
} d3g_mlp_config;

/* Parameter pointers in the reference's state_dict layout (utils/time_utils.py:74-100):
 * weight [out,in] row-major, bias [out].  Unused entries NULL. */
typedef struct d3g_mlp_params {
  const float* timenet0_w; const float* timenet0_b; /* [256,13],[256]  (is_blender) */
  const float* timenet2_w; const float* timenet2_b; /* [30,256],[30] */
  const float* linear_w[8]; const float* linear_b[8]; /* [256,in_l],[256] */
  const float* warp_w; const float* warp_b;         /* gaussian_warp [3,256] (not 6dof) */
  const float* branch_w_w; const float* branch_w_b; /* [3,256] (6dof) */
  const float* branch_v_w; const float* branch_v_b; /* [3,256] (6dof) */
  const float* rot_w; const float* rot_b;           /* gaussian_rotation [4,256] */
  const float* scale_w; const float* scale_b;       /* gaussian_scaling [3,256] */
} d3g_mlp_params;

/* Same layout, mutable: gradients are ACCUMULATED (+=) like torch .grad; caller zeroes. */
typedef struct d3g_mlp_grads {
  float* timenet0_w; float* timenet0_b;
  float* timenet2_w; float* timenet2_b;
  float* linear_w[8]; float* linear_b[8];
  float* warp_w; float* warp_b;
  float* branch_w_w; float* branch_w_b;
  float* branch_v_w; float* branch_v_b;
  float* rot_w; float* rot_b;
  float* scale_w; float* scale_b;
} d3g_mlp_grads;

/* Bytes of the packed-weight buffer (tile images in streaming order, fwd + transposed bwd copies)
 * and of the workspace.  With save_for_backward = 0 the workspace only holds the per-row time features
 * and a status word; with 1 it also holds the activation / mask / gradient tile images that
 * d3g_mlp_backward consumes (about 16.8 KB per row). */
int d3g_mlp_packed_bytes(const d3g_mlp_config* cfg, size_t* bytes);
int d3g_mlp_workspace_bytes(const d3g_mlp_config* cfg, int64_t rows, int32_t save_for_backward,
                            size_t* bytes);
/* The same for the time-batched calls below (n_times blocks of n_points rows). */
int d3g_mlp_workspace_bytes_times(const d3g_mlp_config* cfg, int64_t n_points, int32_t n_times,
                                  int32_t save_for_backward, size_t* bytes);

/* Re-pack the parameters into the swizzled tile images the kernels stream (call after every
 * optimizer step; ~4 MB, one small kernel). */
int d3g_mlp_pack_weights(const d3g_mlp_config* cfg, const d3g_mlp_params* p, void* packed,
                         void* stream);

/* Forward for `rows` points.  xyz [rows,3].  t: [rows] with element stride t_stride (0 = one time
 * value shared by all rows, the `fid.expand(N,1)` pattern of train_baseline.py:110).
 * d_xyz [rows,3] or [rows,4,4] (6dof); d_rot [rows,4] / d_scale [rows,3] may be NULL when the
 * head bit is clear. */
int d3g_mlp_forward(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p,
                    int64_t rows, const float* xyz, const float* t, int64_t t_stride,
                    void* workspace, int32_t save_for_backward, float* d_xyz, float* d_rot,
                    float* d_scale, void* stream);

/* The multi-timestamp form of DeformModel.step (scene/deform_model.py:260-273 loops over the columns of
 * time_emb[N,T] and stacks the results): n_times time values, each applied to the same n_points points, in ONE
 * persistent launch (the weights are shared across timestamps).  t_values: [n_times] device.  Outputs are
 * [n_times, n_points, .] contiguous (torch.stack order).  workspace: d3g_mlp_workspace_bytes_times. */
int d3g_mlp_forward_times(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p,
                          int64_t n_points, int32_t n_times, const float* xyz, const float* t_values,
                          void* workspace, int32_t save_for_backward, float* d_xyz, float* d_rot,
                          float* d_scale, void* stream);

/* Backward.  dL_d* are the cotangents of forward's outputs (NULL = zero).  Parameter gradients are
 * accumulated into `g`.  dL_dxyz [rows,3] and dL_dt may be NULL (the baseline call sites pass xyz.detach(),
 * train.py:208, train_baseline.py:115; the ODE function of utils/time_utils.py:392-438 needs both).  dL_dt is
 * [rows] (one value per row) when forward was given per-row times (t_stride != 0) and a single float (the sum
 * over the rows) when t_stride == 0.  Per-row time values are fully supported (scene/deform_model.py:260-273
 * accepts any [N,T]; train_baseline.py:110-115 passes a materialised [N,1]): whether the rows actually differ
 * is decided on the device, so N copies of one value still take the shared-t fast path. */
int d3g_mlp_backward(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p,
                     int64_t rows, const float* xyz, const float* t, int64_t t_stride,
                     void* workspace, const float* dL_dxyz_out, const float* dL_drot,
                     const float* dL_dscale, const d3g_mlp_grads* g, float* dL_dxyz, float* dL_dt,
                     void* stream);

/* Backward of d3g_mlp_forward_times.  Cotangents are [n_times, n_points, .]; dL_dxyz [n_points,3] (nullable)
 * receives the sum over the time blocks, dL_dt [n_times] (nullable) the sum over the points of each block. */
int d3g_mlp_backward_times(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p,
                           int64_t n_points, int32_t n_times, const float* xyz, const float* t_values,
                           void* workspace, const float* dL_dxyz_out, const float* dL_drot,
                           const float* dL_dscale, const d3g_mlp_grads* g, float* dL_dxyz, float* dL_dt,
                           void* stream);

/* The MLP kernels are persistent (one CTA per SM, all of its shared memory).  A collective that should overlap them —
 * the all-reduce of the Gaussian-parameter gradients, which are complete before the MLP backward starts — needs SMs
 * of its own: n_sms > 0 makes the forward / dgrad / wgrad grids leave that many SMs free.  Process-wide; 0 = default. */
int d3g_mlp_sm_reserve(int32_t n_sms);

/* Synchronises `stream` and returns the kernels' status word (0 = ok, non-zero = an internal
 * pipeline wait timed out, i.e. a bug).  Tests and debug mode call it; the hot path does not. */
int d3g_mlp_read_status(const void* workspace, int32_t* status_host, void* stream);

/* Parity / debug: activation pattern (relu'(z) bits) of trunk layer 0..7 saved by a forward with
 * save_for_backward = 1.  dst_host[row*8 + c] bit j = (h_layer[row][c*32+j] > 0).  Synchronous. */
int d3g_mlp_read_masks(const d3g_mlp_config* cfg, const void* workspace, int64_t rows, int32_t layer,
                       uint32_t* dst_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* D3G_H_ */
