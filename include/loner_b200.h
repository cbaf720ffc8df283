/*
 * loner_b200.h -- C-ABI of libloner_b200.so: hand-written sm_100a CUDA kernels for LONER's mapping
 * hot path.  Plain pointers and sizes only; no torch/ATen types cross this boundary.
 *
 * The reference (esulimma/LONER) is 100% Python and has no FFI layer of its own; the seam this
 * library replaces is (a) the Python API of the third-party `tinycudann` extension and (b) the
 * torch-op bodies of src/models/{ray_sampling,rendering_tcnn,losses}.py and
 * src/mapping/optimizer.py::compute_loss.  Each entry point cites the reference code it replaces
 * (paths relative to the reference root).  The Python binding (ctypes) is loner_b200/_lib.py; the
 * binding a reference maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - every device buffer is allocated by the caller (PyTorch) and passed as a raw device pointer;
 *     the library never frees or retains caller memory.  The plan handle is the only library-owned
 *     object and owns no device memory.
 *   - all calls are asynchronous on `stream` (a cudaStream_t passed as void*; NULL = default stream).
 *   - return 0 on success, negative lb_status on failure; lb_last_error() gives the message of the
 *     last failure on the calling thread.  There is no CPU fallback and no backend dispatch.
 *   - layouts are row-major and contiguous: rays [N,13] = [o(3), d(3), viewdir(3), 0, 0, near, far]
 *     (reference src/common/ray_utils.py:313-315), per-sample arrays [N,S].
 *   - a plan is immutable and may be shared between threads; calls are not re-entrant on the same
 *     output buffers.
 *   - an empty batch (n_rays == 0 / n_points == 0 / n == 0) is a successful no-op and may pass NULL data
 *     buffers (an empty torch tensor has a NULL data pointer); handles (plan, cfg) must still be valid.
 */
#ifndef LONER_B200_H
#define LONER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define LB_API __attribute__((visibility("default")))
#else
#define LB_API
#endif

#define LB_VERSION 100
#define LB_MAX_LEVELS 32
#define LB_RAY_STRIDE 13

typedef enum {
    LB_OK = 0,
    LB_ERR_BAD_ARGUMENT = -1,
    LB_ERR_UNSUPPORTED = -2,
    LB_ERR_CUDA = -3,
} lb_status;

/* Hash-grid + MLP configuration: the keys of the reference's tcnn config dicts
 * (cfg/nerf_config/default_nerf_hash.yaml:14-31). */
typedef struct {
    int32_t n_levels;            /* 16 */
    int32_t n_features_per_level;/* 2 (only 2 is supported) */
    int32_t log2_hashmap_size;   /* 18 / 19 / 22 */
    int32_t base_resolution;     /* 16 */
    float   per_level_scale;     /* 2.0 (tcnn default when the YAML omits it) */
    int32_t mlp_width;           /* n_neurons: 64 (only 64 is supported) */
    int32_t mlp_hidden_layers;   /* 1 for the sigma head */
} lb_grid_config;

typedef struct lb_plan lb_plan;

/* Loss configuration: cfg/model_config/default_model_config.yaml:39-58 as read by
 * src/mapping/optimizer.py:711-848. */
typedef enum { LB_LOSS_L1_JS = 0, LB_LOSS_L2_JS = 1, LB_LOSS_L1_LOS = 2, LB_LOSS_L2_LOS = 3 } lb_loss_kind;

typedef struct {
    int32_t kind;             /* lb_loss_kind */
    float   scale_factor;     /* world_cube.scale_factor: metres per cube unit */
    float   min_js_score;     /* JS_loss.min_js_score */
    float   max_js_score;     /* JS_loss.max_js_score */
    float   js_alpha;         /* JS_loss.alpha */
    float   min_depth_eps;    /* loss.min_depth_eps (eps_min) */
    float   depth_eps;        /* fixed / decayed eps for the *_LOS kinds (caller applies the decay) */
    float   los_lambda;       /* loss.los_lambda (caller applies decay_los_lambda) */
    float   depthloss_lambda; /* loss.depthloss_lambda */
    float   n_opaque;         /* GLOBAL count of opaque rays (all ranks); <=0: use the local count */
    float   n_total_samples;  /* GLOBAL N*S for the mean over all samples; <=0: use local N*S */
} lb_loss_config;

/* ------------------------------------------------------------------------------------------ */
LB_API int         lb_version(void);
/* bit 0: built with -DLB_TEST_SWITCHES (fp32 CUDA-core MLP cross-check and the LONER_B200_* environment switches
 * exist); the shipped library returns 0: one implementation per operation, no run-time backend selection. */
LB_API int         lb_build_flags(void);
LB_API const char* lb_last_error(void);

/* Immutable level table (scale, resolution, offset, size, dense|hash) following tiny-cuda-nn's
 * GridEncoding constructor; replaces tcnn.Encoding/NetworkWithInputEncoding construction
 * (src/models/nerf_tcnn.py:35-40). */
LB_API int  lb_plan_create(const lb_grid_config* cfg, lb_plan** out);
LB_API void lb_plan_destroy(lb_plan* plan);
/* Overrides the device L2 size the plan was created with.  The encode launchers walk the levels in groups whose tables
 * fit 55 % of it (lb_encode.cu); tests use a tiny value to exercise the level-group passes on small tables. */
LB_API int  lb_plan_set_l2_bytes(lb_plan* plan, int64_t l2_bytes);
/* number of fp32 scalars: MLP matrices (W1[64,32] then W2[16,64], row-major [out,in]) and tables */
LB_API int  lb_plan_param_counts(const lb_plan* plan, int64_t* n_mlp, int64_t* n_table);
LB_API int  lb_plan_level(const lb_plan* plan, int level, float* scale, uint32_t* resolution,
                   uint32_t* offset, uint32_t* size, int32_t* is_dense);

/* ---- R0: ray construction ---------------------------------------------------------------- */
/* Pose 6-vectors [n_poses, 6] = [t, axis-angle] -> rot [n_poses, 9] (row-major R) and origin [n_poses, 3] =
 * (t + shift) / scale: Pose.get_transformation_matrix / tensor_to_transform (src/common/pose.py:148-152,
 * src/common/pose_utils.py:354-368: pytorch3d axis_angle_to_matrix) + the origin lines of build_lidar_rays
 * (src/common/ray_utils.py:293-296).  _bwd: d_pose6 [n_poses, 6] = chain rule through the same formulas (overwritten). */
LB_API int lb_pose_to_rot_origin(const float* pose6, const float* shift, float scale, int n_poses, float* rot,
                          float* origin, void* stream);
LB_API int lb_pose_to_rot_origin_bwd(const float* pose6, float scale, int n_poses, const float* d_rot,
                              const float* d_origin, float* d_pose6, void* stream);

/* LidarRayDirections.build_lidar_rays (src/common/ray_utils.py:269-322) incl. get_far_val (ray_utils.py:31-60)
 * and the index selection of src/mapping/optimizer.py:389-400, as called by KeyFrame.build_lidar_rays
 * (src/mapping/keyframe.py:75-105).  directions [3,n]; distances [n] in metres (NULL: every depth is
 * const_depth, the sky-ray case of keyframe.py:96-98); indices [m] int64 (NULL: identity, m <= n);
 * rot [9] = row-major R of the lidar pose, origin [3] = (t + shift) / scale (device pointers, computed by the
 * host from the pose 6-vector); near = r0/scale, far_max = r1/scale, inv_scale = 1/scale.
 * Outputs rays [m,13] = [o, d, -d, 0, 0, near, min(far_max, cube exit)], depths [m] in cube units.
 * mask_span (cube units; 0 = off): the reference DROPS rays with far <= near + 1/scale (ray_utils.py:321-322, a
 * boolean-mask gather: host sync + data-dependent shape).  With mask_span = 1/scale such a ray stays in the batch and
 * its depth is set to NaN -- the "masked ray" convention of lb_count_opaque / lb_composite_loss / lb_occ_grad: no loss
 * term, no gradient, not counted in either normaliser, i.e. the arithmetic of the batch without the ray. */
LB_API int lb_build_lidar_rays(const float* directions, const float* distances, const int64_t* indices, int n,
                        int m, const float* rot, const float* origin, float near, float far_max,
                        float inv_scale, float const_depth, float mask_span, float* rays, float* depths,
                        void* stream);
/* its backward: d_rays [m,13] -> d_rot [9] += , d_origin [3] +=  (the pose chain rule R(axis-angle), o(t)
 * of src/common/pose_utils.py:354-368 stays with the host's autograd). */
LB_API int lb_build_lidar_rays_bwd(const float* directions, const int64_t* indices, int n, int m,
                            const float* rot, const float* origin, float far_max, const float* d_rays,
                            float* d_rot, float* d_origin, void* stream);

/* CameraRayDirections.build_rays (src/common/ray_utils.py:168-206) as called by KeyFrame.build_camera_rays
 * (src/mapping/keyframe.py:108-128), with the pixel selection of src/mapping/optimizer.py:626-651 fused in.
 * directions [n,3] (camera frame, one per pixel), pix_i / pix_j [n] pixel grid coordinates (NULL: zeros),
 * image [n,n_colors] (NULL: no intensities), indices [m] int64 (NULL: identity); rot [9] / origin [3] of the
 * (detached) camera pose as for the LiDAR rays.  rays [m,13] = [o, d, -d, pix_i, pix_j, near, cube exit];
 * intensities [m,n_colors] = image[indices].  No backward: the reference detaches the camera pose. */
LB_API int lb_build_camera_rays(const float* directions, const float* pix_i, const float* pix_j, const float* image,
                         int n_colors, const int64_t* indices, int n, int m, const float* rot,
                         const float* origin, float near, float* rays, float* intensities, void* stream);

/* ---- R2: sampling ------------------------------------------------------------------------ */
/* OccGridRaySampler.get_samples (src/models/ray_sampling.py:53-92) incl. OccupancyGridModel.interpolate
 * (src/models/model_tcnn.py:126-134) and sample_pdf (src/models/rendering_tcnn.py:19-68).
 * occ [D,D,D] fp32 logits (x -> last dim); t_steps [S/2] = linspace(0,1,S/2); u1,u2 [N,S/2] uniform.
 * z_out [N,S] sorted.  Optional debug outputs (may be NULL): occ_bits [N,S/2] (prob>0),
 * inds [N,S/2] int64 searchsorted indices -- both bit-exact w.r.t. oracle/loner_oracle.py. */
LB_API int lb_sample_occ(const float* rays, const float* occ, int occ_dim, const float* t_steps,
                  const float* u1, const float* u2, int n_rays, int n_samples, float perturb,
                  float* z_out, uint8_t* occ_bits, int64_t* inds, void* stream);
/* UniformRaySampler.get_samples (src/models/ray_sampling.py:22-43). t_steps [S], u [N,S]. */
LB_API int lb_sample_uniform(const float* rays, const float* t_steps, const float* u, int n_rays,
                      int n_samples, float perturb, float* z_out, void* stream);

/* ---- R4: sigma field (tcnn NetworkWithInputEncoding / Encoding) -------------------------- */
/* tcnn.Encoding(HashGrid).forward (src/models/nerf_tcnn.py:40,64): x01 [P,3] in [0,1] -> enc [P,L*2]. */
LB_API int lb_encode_fwd(const lb_plan* plan, const float* x01, const float* table, int64_t n_points,
                  float* enc, void* stream);
/* its backward: d_enc [P,L*2] -> grad_table += scatter; d_x01 [P,3] (NULL to skip) */
LB_API int lb_encode_bwd(const lb_plan* plan, const float* x01, const float* table, const float* d_enc,
                  int64_t n_points, float* grad_table, float* d_x01, void* stream);

/* Fused sigma head: tcnn.NetworkWithInputEncoding(3->1) forward (src/models/nerf_tcnn.py:35-38,71-72)
 * fused with xyz = o + d*z and pos01 = (xyz+1)/2 (src/models/rendering_tcnn.py:390, nerf_tcnn.py:63).
 * Point source: either (rays,z) with P = n_rays*n_samples, or x01 [P,3] when rays == NULL
 * (then n_rays = P, n_samples = 1).  params = [mlp | table] flat fp32 (tcnn layout).
 * sigma [P]; enc [P,32] receives the encodings and must be kept for lb_sigma_bwd.  act [P] uint64 receives
 * the ReLU bit mask of the 64 hidden units; when non-NULL the MLP runs on the tcgen05 tensor cores (exact
 * tf32 hi/lo split operands, fp32 accumulate in TMEM) and act must be handed to lb_sigma_bwd; NULL selects
 * the fp32 CUDA-core MLP.  With (rays, z) on a plan for which lb_sigma_fwd_enc_optional() returns 1, gather and
 * MLP run as ONE kernel (the encodings are handed over in shared memory); enc may then be NULL for inference
 * (nothing is saved).  In every other case enc is required: it is the hand-off between the two kernels. */
LB_API int lb_sigma_fwd_enc_optional(const lb_plan* plan);
LB_API int lb_sigma_fwd(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                 const float* params, int n_rays, int n_samples, float* sigma, float* enc,
                 uint64_t* act, void* stream);
/* lb_sigma_fwd for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc): sigma-head INFERENCE (nothing is kept for a backward) with the encode -> MLP hand-off
 * and the ReLU masks in a fixed scratch array inside the library instead of caller-owned enc / act buffers.
 * NO COUNTERPART IN THE REFERENCE.
 *   - limit: P = n_rays * n_samples (or n_rays points with x01) <= LB_NOALLOC_MAX_POINTS; more is rejected with
 *     LB_ERR_BAD_ARGUMENT and a message naming the limit.  The scratch (136 B per point = 8.5 MiB) is a file-scope
 *     __device__ array: every GPU that loads the library holds it until the process exits.
 *   - calls no cudaMalloc / cudaFree / allocating library routine.  The first call on a device looks up the array's
 *     device address (cudaGetSymbolAddress); make that call before starting a stream capture.
 *   - all calls on a device share the array, which keeps the previous call's contents (nothing needs clearing: a call
 *     writes every element it reads).  Calls that OVERLAP on the GPU race on it: keep them on one stream or order them
 *     with events.
 * Same arguments and results as lb_sigma_fwd without enc / act; same kernels, so sigma is bit-identical to it. */
#define LB_NOALLOC_MAX_POINTS 65536
LB_API int lb_sigma_fwd_noalloc(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                         const float* params, int n_rays, int n_samples, float* sigma, void* stream);
/* Backward of the above: d_sigma [P] -> grad_params += (MLP grads, table scatter-add; NULL to skip);
 * d_rays [N,13] += (origin/direction gradients through the grid input) when non-NULL
 * (or d_x01 [P,3] when the point source is x01).  enc_inout = enc as written by lb_sigma_fwd; it is
 * CONSUMED: its contents are undefined on return (the two-kernel path overwrites it with dL/d_enc, the
 * one-kernel path leaves it alone).  act = the mask written by lb_sigma_fwd (NULL: fp32 CUDA-core MLP, which
 * recomputes the hidden layer). */
LB_API int lb_sigma_bwd(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                 const float* params, float* enc_inout, const uint64_t* act, const float* d_sigma,
                 int n_rays, int n_samples, float* grad_params, float* d_rays, float* d_x01, void* stream);

/* Forward-mode input derivative for pose refinement (the lidar pose is a leaf of the reference's graph,
 * src/mapping/optimizer.py:701-712,849: autograd carries dL/dsigma through tcnn's input gradient to the ray origins and
 * directions built in src/common/ray_utils.py:300-322).  lb_sigma_fwd_jvp = lb_sigma_fwd on the one-kernel path plus
 * dsigma_dx [P,3] = d sigma / d x01 of every sample point, formed inside the same kernel from the corner values the
 * gather already holds (no second table gather in the backward).  The backward is then lb_sigma_bwd with d_rays = d_x01 =
 * NULL plus, for the (rays, z) source, lb_rays_grad_from_jvp: d_rays[:,0:3] += sum_s d_sigma g / 2,
 * d_rays[:,3:6] += sum_s d_sigma z g / 2 (with x01 the caller forms d_x01 = d_sigma * dsigma_dx itself).
 * Covered plans only (lb_sigma_fwd_jvp_supported() = 1: tables within the L2 budget); others get LB_ERR_UNSUPPORTED and
 * use the d_rays / d_x01 outputs of lb_sigma_bwd.  enc / act may be NULL when no parameter gradient will be asked. */
LB_API int lb_sigma_fwd_jvp_supported(const lb_plan* plan);
LB_API int lb_sigma_fwd_jvp(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                     const float* params, int n_rays, int n_samples, float* sigma, float* enc,
                     uint64_t* act, float* dsigma_dx, void* stream);
LB_API int lb_rays_grad_from_jvp(const float* z, const float* dsigma_dx, const float* d_sigma, int n_rays, int n_samples,
                          float* d_rays, void* stream);

/* ---- R5/R6: compositing ------------------------------------------------------------------ */
/* raw2outputs, sigma-only (src/models/rendering_tcnn.py:219-295).  noise [N,S] already scaled by
 * raw_noise_std (NULL: none).  Outputs: weights [N,S], depth/opacity/variance [N] (variance may be NULL).
 * strategy 0 = 'default' (expected depth incl. far term), 1 = 'adjusted' (median-transmittance depth,
 * noise ignored; rendering_tcnn.py:70-214 effective semantics). */
LB_API int lb_composite_fwd(const float* sigma, const float* z, const float* rays, const float* noise,
                     int n_rays, int n_samples, int strategy, float* weights, float* depth,
                     float* opacity, float* variance, void* stream);
/* autograd backward of raw2outputs('default') for upstream grads g_w [N,S] (NULL=0), g_depth,
 * g_opacity, g_variance [N] (NULL=0): d_sigma [N,S]; d_rays [N,13] += (|d| and far terms) if non-NULL. */
LB_API int lb_composite_bwd(const float* sigma, const float* z, const float* rays, const float* noise,
                     const float* g_w, const float* g_depth, const float* g_opacity,
                     const float* g_variance, int n_rays, int n_samples, float* d_sigma,
                     float* d_rays, void* stream);

/* ---- L1-L5: fused compositing + JS dynamic-margin depth loss + analytic gradient ---------- */
/* Optimizer.compute_loss lidar branch (src/mapping/optimizer.py:719-848) with get_weights_gt
 * (src/models/losses.py:29-51), fused with raw2outputs and its backward.
 * Outputs: weights [N,S], depth/opacity/variance [N], per-ray js [N] (clamped score), eps [N],
 * loss_parts [4] += {sum_opaque (s*D-g)^2, sum_all |w-w_gt| (or squared), sum_opaque |A-1|, n_opaque_local},
 * d_sigma [N,S] = dL/dsigma for grad_output 1, d_rays [N,13] (overwritten; |d| and far terms).
 * cfg->n_opaque / n_total_samples carry the GLOBAL normalisers; when norm_dev (device float[2] =
 * {opaque-ray count, ray count}, filled by lb_count_opaque and summed over chunks / ranks) is non-NULL it
 * overrides them without a host round trip: n_opaque = norm_dev[0], n_total_samples = norm_dev[1] * S
 * (norm_dev[1] <= 0 keeps cfg->n_total_samples). */
LB_API int lb_composite_loss(const float* sigma, const float* z, const float* rays, const float* noise,
                      const float* depth_gt, const lb_loss_config* cfg, const float* norm_dev,
                      int n_rays, int n_samples,
                      float* weights, float* depth, float* opacity, float* variance, float* js,
                      float* eps, float* loss_parts, float* d_sigma, float* d_rays, void* stream);
/* count of opaque rays (depth_gt > 0 and depth_gt <= far; optimizer.py:724-727): count[0] += opaque rays,
 * count[1] += rays that are not masked (depth_gt != NaN; all n_rays when nothing is masked) -- device float[2]: the
 * norm_dev of lb_composite_loss */
LB_API int lb_count_opaque(const float* rays, const float* depth_gt, int n_rays, float* count, void* stream);

/* ---- O1/O2: optimiser step and occupancy update ------------------------------------------ */
/* torch.optim.Adam defaults, no weight decay (src/mapping/optimizer.py:255-265): in-place on p, m, v.
 * step is 1-based.  grad_scale multiplies g first (e.g. 1/world for averaged DP); zero_grad != 0 clears g. */
LB_API int lb_adam_step(float* p, float* g, float* m, float* v, int64_t n, double lr, double beta1,
                 double beta2, double eps, int step, float grad_scale, int zero_grad, void* stream);
/* _step_occupancy_grid (src/mapping/optimizer.py:897-908) with get_logits_grad (losses.py:54-62):
 * occ_grad [D,D,D] += trilinear scatter of the log-odds update at xyz = o + d*z; the caller applies
 * occ -= lr * occ_grad (lb_sgd_step) so that DP ranks can all-reduce occ_grad first. */
LB_API int lb_occ_grad(const float* rays, const float* z, const float* depth_gt, float scale_factor,
                int n_rays, int n_samples, int occ_dim, float* occ_grad, void* stream);
LB_API int lb_sgd_step(float* p, float* g, int64_t n, float lr, int zero_grad, void* stream);

/* ---- R4/R5 colour branch (decoupled RGB head) ------------------------------ */
/* tcnn.Encoding(3, {otype: SphericalHarmonics, degree: 4}) (src/models/nerf_tcnn.py:43,86;
 * cfg/nerf_config/default_nerf_hash.yaml:4-6): d01 [P,3] in [0,1] -> 16 real SH coefficients of 2*d01-1. */
LB_API int lb_sh4_fwd(const float* d01, int64_t n_points, float* out, void* stream);
/* input gradient of the above: d_d01 [P,3] = J^T d_out [P,16] */
LB_API int lb_sh4_bwd(const float* d01, const float* d_out, int64_t n_points, float* d_d01, void* stream);
/* tcnn.Network(n_in, n_out, {FullyFusedMLP, ReLU, n_neurons 64, n_hidden_layers}) forward
 * (src/models/nerf_tcnn.py:50-52,89-91; default_nerf_hash.yaml:8-13), one fused tcgen05 kernel.
 * x [P,n_in] with n_in in {16,32,48,64}; params in tcnn's layout: W0 [64,n_in], W1.. [64,64], W_out [16,64]
 * (rows >= n_out unused), row-major, no bias.  out [P,n_out]; out_act 0 = none, 1 = sigmoid (the torch.sigmoid
 * of nerf_tcnn.py:93 fused).  acts receives the post-ReLU activations for lb_net_bwd (NULL for inference):
 * n_hidden * ceil(P/128)*128 * 64 floats in a tile-transposed layout private to the two kernels. */
LB_API int lb_net_fwd(const float* x, int64_t n_points, int n_in, int n_hidden, int n_out, int out_act,
               const float* params, float* out, float* acts, void* stream);
/* backward of lb_net_fwd for upstream d_out [P,n_out] (`out` = forward output, needed when out_act = 1):
 * grad_params += dL/dparams (NULL: skip), d_x [P,n_in] = dL/dx (NULL: skip). */
LB_API int lb_net_bwd(const float* x, int64_t n_points, int n_in, int n_hidden, int n_out, int out_act,
               const float* params, const float* acts, const float* out, const float* d_out,
               float* grad_params, float* d_x, void* stream);
/* colour lines of raw2outputs (src/models/rendering_tcnn.py:284-289): rgb [N,C] = sum_s weights[N,S] colors[N,S,C]
 * (+ 1 - sum_s weights when white_bkgd); C <= 4. */
LB_API int lb_composite_rgb_fwd(const float* weights, const float* colors, int n_rays, int n_samples, int n_colors,
                         int white_bkgd, float* rgb, void* stream);
/* its backward for upstream g_rgb [N,C]: d_weights [N,S] and d_colors [N,S,C] (either may be NULL) */
LB_API int lb_composite_rgb_bwd(const float* weights, const float* colors, const float* g_rgb, int n_rays,
                         int n_samples, int n_colors, int white_bkgd, float* d_weights, float* d_colors,
                         void* stream);

/* ---- f3: eval helpers --------------------------------------------------------------------------- */
/* The mesher's max-weight voxel splat (analysis/mesher.py:137-180): for every sample point o + d*z of every ray with
 * ray_mask != 0 (NULL: all rays) that lies inside [bounds[0], bounds[n-1]] on all three axes, volume[voxel] =
 * max(volume[voxel], weight) with voxel = bucketize(x)*nz + bucketize(y)*nx*nz + bucketize(z) (torch.bucketize,
 * right=False, fp64 boundaries).  volume [ny*nx*nz] fp32 is accumulated into (caller zero-fills once per mesh);
 * weights must be >= 0 (compositing weights are). */
LB_API int lb_voxel_splat_max(const float* rays, const float* z, const float* weights, const uint8_t* ray_mask,
                       int n_rays, int n_samples, const double* x_bounds, int nx, const double* y_bounds, int ny,
                       const double* z_bounds, int nz, float* volume, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LONER_B200_H */
