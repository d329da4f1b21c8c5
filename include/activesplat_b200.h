/*
 * activesplat_b200.h -- C ABI of libactivesplat_b200.so
 *
 * The drop-in boundary for ActiveSplat's rasterizer hot path.  Everything here is plain C:
 * device pointers, sizes, a CUDA stream passed as void*.  No torch / C++ types cross it.
 * The library keeps no state between calls and allocates no device memory of its own: outputs
 * are caller-owned and scratch comes from caller-supplied allocator callbacks, exactly like the
 * reference (dgr = /root/reference/submodules/diff-gaussian-rasterization):
 *
 *   as_forward       replaces CudaRasterizer::Rasterizer::forward      dgr/cuda_rasterizer/rasterizer.h:33-60
 *                                                                       (impl rasterizer_impl.cu:198-342)
 *   as_backward      replaces CudaRasterizer::Rasterizer::backward     rasterizer.h:62-86 (impl :346-441)
 *   as_mark_visible  replaces CudaRasterizer::Rasterizer::markVisible  rasterizer.h:24-29 (impl :141-153)
 *   AsAllocFn        replaces std::function<char*(size_t)>             rasterizer.h:34-36, rasterize_points.cu:27-33
 *
 * All functions return >= 0 on success and a negative AS_ERR_* code on failure;
 * as_error_string() turns a code into text.  Every entry point is re-entrant and may be called
 * concurrently on different streams / devices.
 */
#ifndef ACTIVESPLAT_B200_H
#define ACTIVESPLAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AS_OK 0
#define AS_ERR_INVALID_ARG (-1)   /* null pointer / negative size / unsupported channel count */
#define AS_ERR_NO_COLORS (-2)     /* colors_precomp == NULL: "For non-RGB, provide precomputed Gaussian colors!" (rasterizer_impl.cu:244-247); SH evaluation is not on the ActiveSplat path (sh_degree 0 + colors_precomp always) */
#define AS_ERR_ALLOC (-3)         /* an allocator callback returned NULL */
#define AS_ERR_CUDA (-4)          /* a CUDA runtime call or (debug mode) a kernel failed; see as_last_cuda_error() */
#define AS_ERR_CAPACITY (-5)      /* a caller-sized buffer is too small (fused / no-sync entries) */
#define AS_ERR_STATIC_LIMIT (-6)  /* input exceeds the fixed size of the library's built-in scratch */

/* Scratch allocator: returns a DEVICE pointer to at least `bytes` bytes (128-byte aligned or better),
 * valid until the matching as_backward has run.  Mirrors the resize lambdas of rasterize_points.cu:27-33. */
typedef char* (*AsAllocFn)(void* ctx, size_t bytes);

/* Arguments of Rasterizer::forward (rasterizer.h:33-60), in the same order, as a POD.
 * All pointers are device pointers to contiguous fp32 unless stated.  Optional inputs are NULL
 * when absent (the reference passes the null data pointer of an empty CPU tensor). */
typedef struct AsForwardArgs {
    int P;                       /* number of Gaussians */
    int D;                       /* SH degree (must be 0 on this path) */
    int M;                       /* SH coefficients per Gaussian (ignored) */
    int channels;                /* 3 = reference NUM_CHANNELS (config.h:15); 6 = fused [rgb | z,1,z^2] bundle */
    const float* background;     /* [channels] */
    int width, height;
    const float* means3D;        /* [P,3] */
    const float* shs;            /* unused, may be NULL */
    const float* colors_precomp; /* [P,channels] */
    const float* opacities;      /* [P] */
    const float* scales;         /* [P,3] or NULL when cov3D_precomp is given */
    float scale_modifier;
    const float* rotations;      /* [P,4] (r,x,y,z), used as given (forward.cu:127) */
    const float* cov3D_precomp;  /* [P,6] or NULL */
    const float* viewmatrix;     /* [16] = w2c^T flattened (column-major w2c) */
    const float* projmatrix;     /* [16] = (P_gl w2c)^T flattened */
    const float* cam_pos;        /* [3], unused without SH */
    float tan_fovx, tan_fovy;
    int prefiltered;
    float* out_color;            /* [channels,H,W] */
    float* out_depth;            /* [1,H,W] median depth (15.0 where T never crosses 0.5) */
    float* out_opacity;          /* [1,H,W] 1 - T */
    int* radii;                  /* [P] int32, may be NULL */
    int debug;                   /* non-zero: synchronise and check after every stage (auxiliary.h:166-173) */
} AsForwardArgs;

/* Arguments of Rasterizer::backward (rasterizer.h:62-86).  Gradient buffers are caller-owned;
 * the library zero-fills what it needs itself (the caller does NOT have to pre-zero them). */
typedef struct AsBackwardArgs {
    int P, D, M, R;              /* R = num_rendered returned by as_forward */
    int channels;
    const float* background;
    int width, height;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    float tan_fovx, tan_fovy;
    const int* radii;            /* may be NULL: the copy kept in the geometry buffer is used */
    char* geom_buffer;           /* the three scratch buffers handed out during as_forward */
    char* binning_buffer;
    char* image_buffer;
    const float* dL_dpix;        /* [channels,H,W] */
    float* dL_dmean2D;           /* [P,3]  (x,y in NDC units, z = 0; backward.cu:460-461,545-546) */
    float* dL_dconic;            /* [P,4]  (x,y,w used) -- may be NULL */
    float* dL_dopacity;          /* [P] */
    float* dL_dcolor;            /* [P,channels] */
    float* dL_dmean3D;           /* [P,3] */
    float* dL_dcov3D;            /* [P,6] -- may be NULL unless cov3D_precomp is given */
    float* dL_dsh;               /* unused, may be NULL */
    float* dL_dscale;            /* [P,3] */
    float* dL_drot;              /* [P,4] */
    int debug;
} AsBackwardArgs;

/* ---- the three reference entry points ------------------------------------------------------ */

/* Returns num_rendered (>= 0).  Performs ONE host synchronisation on `stream` to learn
 * num_rendered before sizing the binning buffer, like rasterizer_impl.cu:283. */
int as_forward(const AsForwardArgs* args, AsAllocFn geometryBuffer, AsAllocFn binningBuffer,
               AsAllocFn imageBuffer, void* alloc_ctx, void* cuda_stream);

int as_backward(const AsBackwardArgs* args, void* cuda_stream);

/* present[i] = (w2c * p_i).z > 0.2  (auxiliary.h:154); present is uint8 (bool) [P]. */
int as_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                    uint8_t* present, void* cuda_stream);

/* ---- scratch sizing / layout (rasterizer_impl.h:21-72 `required<T>` / `fromChunk`) ---------- */
size_t as_geometry_bytes(int P);
size_t as_image_bytes(int width, int height);
size_t as_binning_bytes(int num_rendered, int channels);

/* Byte offsets inside the scratch buffers, for parity tests that compare intermediates.
 * (The layout is private to the library; offsets are relative to the pointer the allocator returned.) */
typedef struct AsScratchLayout {
    /* geometry buffer */
    size_t depths, means2D, conic_opacity, cov3D, tiles_touched, point_offsets, radii;
    /* binning buffer */
    size_t keys_unsorted, values_unsorted, keys_sorted, point_list;
    /* image buffer */
    size_t final_T, n_contrib, ranges;
} AsScratchLayout;
int as_scratch_layout(int P, int width, int height, int num_rendered, int channels, AsScratchLayout* out);

/* ---- fused SplaTAM frame step (additive; the three entry points above are unchanged) --------
 *
 * One call = transform_to_frame + activations + [z,1,z^2] features + the 6-channel bundle render
 * (reference: src/mapper/splatam/utils/slam_helpers.py:124-139,196-213,252-304 feeding two
 * Renderer calls at src/mapper/splatam/splatam.py:208,212), straight from the RAW parameters.
 * No host synchronisation: num_rendered stays on the device (status[0]); the binning buffer is sized
 * by the caller for `binning_capacity` instances (as_binning_bytes(capacity, 6)); if a frame needs more,
 * status[1] is set to 1, nothing is written out of bounds, and the frame's outputs are invalid
 * (re-run with a larger buffer).  The three scratch buffers are caller-owned and reused across
 * frames; geom_buffer must be zero-filled once when it is created (as_geometry_bytes(P) bytes). */
/* How as_frame_forward orders the (tile, depth) instances.  Both produce the reference's sorted list exactly.
 *   AS_BINNING_GLOBAL_SORT   radix sort of all 64-bit keys (CudaRasterizer: rasterizer_impl.cu:305-310); any scene.
 *   AS_BINNING_TILE_BUCKETS  per-tile buckets filled with atomics, each sorted by (depth, Gaussian id) by one CTA;
 *                            fastest while no tile holds more than a few thousand instances (status[3] reports the
 *                            largest bucket); bigger buckets are still sorted correctly, only slowly. */
#define AS_BINNING_GLOBAL_SORT 0
#define AS_BINNING_TILE_BUCKETS 1
typedef struct AsFrameArgs {
    int P, width, height;
    int isotropic;                 /* log_scales is [P,1] (rotations then pass through un-transformed) */
    const float* means3D;          /* [P,3] world frame */
    const float* rgb_colors;       /* [P,3] */
    const float* unnorm_rotations; /* [P,4] */
    const float* logit_opacities;  /* [P] */
    const float* log_scales;       /* [P,3] or [P,1] */
    const float* cam_unnorm_rot;   /* camera quaternion (r,x,y,z), element k at cam_unnorm_rot[k*cam_rot_stride] */
    int cam_rot_stride;
    const float* cam_trans;        /* camera translation, element k at cam_trans[k*cam_trans_stride] */
    int cam_trans_stride;
    const float* sil_w2c;          /* [16] row-major w2c of the [z,1,z^2] features (curr_data['w2c']); NULL = identity */
    const float* viewmatrix;       /* raster settings, as in AsForwardArgs */
    const float* projmatrix;
    const float* background;       /* [6] */
    float tan_fovx, tan_fovy, scale_modifier;
    char* geom_buffer;
    char* image_buffer;
    char* binning_buffer;
    int binning_capacity;
    uint32_t* status;              /* [4] device words: [0] num_rendered, [1] overflow flag, [2] visible Gaussians,
                                      [3] largest number of instances in one tile (binning_mode 1 only) */
    float* out_color;              /* [6,H,W]: rgb | depth, silhouette, depth^2 */
    float* out_depth;              /* [1,H,W] median depth */
    float* out_opacity;            /* [1,H,W] */
    int* radii;                    /* [P] or NULL */
    int debug;
    float* max_2D_radius;          /* [P] or NULL: updated in place, max(radius, .) where seen (splatam.py:296-298) */
    uint8_t* seen;                 /* [P] bool or NULL: radius > 0 */
    int binning_mode;              /* AS_BINNING_GLOBAL_SORT (0) or AS_BINNING_TILE_BUCKETS (1); same results */
    void* chunk_scratch;           /* NULL, or as_chunk_scratch_bytes(binning_capacity, width, height) bytes (16-byte
                                    * aligned): blend every tile's list in chunks of 1024 records, one CTA per chunk, and
                                    * combine -- for low-resolution wide-angle views with 10^4-10^5 records per tile,
                                    * where one serial walk per pixel strip leaves the GPU idle.  Images agree with the
                                    * single walk to rounding (T at a chunk start is a product of chunk products);
                                    * n_contrib is not produced, so no as_frame_backward may follow. */
} AsFrameArgs;
size_t as_chunk_scratch_bytes(int binning_capacity, int width, int height);

typedef struct AsFrameBackwardArgs {
    const AsFrameArgs* frame;      /* the arguments of the matching as_frame_forward */
    const float* dL_dpix;          /* [6,H,W] */
    const float* grad_scale;       /* device scalar multiplied into every output gradient; NULL = 1 */
    int gaussians_grad;            /* transform_to_frame(gaussians_grad=...) */
    int camera_grad;               /* transform_to_frame(camera_grad=...) */
    float* grad_params;            /* 14 P floats, the five gradient tensors one after the other, each in its parameter's
                                    * layout: means [P,3] at 0 | rgb [P,3] at 3P | rotation [P,4] at 6P | logit opacity
                                    * [P,1] at 10P | log scales [P,1 or 3] at 11P (the "flat segment layout") */
    float* grad_means2D;           /* [P,3] or NULL (NDC-unit screen gradient used by densification) */
    float* grad_cam;               /* 7 values, element k at grad_cam[k*grad_cam_stride]: d/d cam_unnorm_rot (4) | d/d cam_trans (3) */
    float* pose_scratch;           /* as_pose_scratch_floats(P) device floats (per-block partial sums; no init needed) */
    int accumulate;                /* 0: grad_params = (cleared first, then the visible Gaussians' rows are written);
                                    * 1: grad_params += (keyframe-batch accumulation); 2: =, into a buffer the caller has
                                    * already cleared (e.g. on another stream while the forward ran) */
    int grad_cam_stride;           /* 0 or 1 = packed */
} AsFrameBackwardArgs;

size_t as_pose_scratch_floats(int P);
int as_tile_sort_capacity(void);   /* bucket size up to which AS_BINNING_TILE_BUCKETS sorts a tile in shared memory */
int as_frame_forward(const AsFrameArgs* args, void* cuda_stream);
int as_frame_backward(const AsFrameBackwardArgs* args, void* cuda_stream);

/* Tracking loss of get_loss (splatam.py:213-247, use_l1): masked L1 sums and their gradient w.r.t. the
 * six rendered channels, in one sweep.  losses[8] = {weighted total, weighted im, weighted depth, mask count,
 * num_rendered, overflow flag, -, -}; the last two are copied from frame_status (AsFrameArgs.status) when it
 * is not NULL, so one read-back of `losses` tells the host everything about the frame; if the frame overflowed
 * its binning buffer, the three loss values are NaN.
 * scratch: as_loss_scratch_floats(w,h) floats.  mask_color: 1 = colour loss uses the depth mask (splatam.py:243-245). */
size_t as_loss_scratch_floats(int width, int height);
int as_tracking_loss(int width, int height, const float* color6, const float* gt_rgb, const float* gt_depth,
                     float w_im, float w_depth, float sil_thres, int use_sil_for_loss, int mask_color,
                     float* dL_dcolor6, float* losses, float* scratch, const uint32_t* frame_status,
                     void* cuda_stream);

/* Mapping loss of get_loss (splatam.py:239-249): w_im * (0.8 * mean|im - gt| + 0.2 * (1 - mean SSIM)) +
 * w_depth * mean over the valid-depth mask of |gt_depth - depth|, SSIM as calc_ssim (slam_external.py:54-97:
 * 11x11 Gaussian window, sigma 1.5, zero padding, c1 1e-4, c2 9e-4), with its gradient w.r.t. the six
 * rendered channels.  losses[8] as in as_tracking_loss, plus losses[6] = mean SSIM.
 * scratch: as_mapping_loss_scratch_floats(w,h) floats. */
size_t as_mapping_loss_scratch_floats(int width, int height);
int as_mapping_loss(int width, int height, const float* color6, const float* gt_rgb, const float* gt_depth,
                    float w_im, float w_depth, float* dL_dcolor6, float* losses, float* scratch,
                    const uint32_t* frame_status, void* cuda_stream);

/* torch.optim.Adam step (splatam.py:118-124: betas (0.9,0.999), eps 1e-15, no weight decay) over the five
 * parameter tensors at once; grad/exp_avg/exp_avg_sq are in the flat segment layout of AsFrameBackwardArgs.grad_params
 * (14 P floats each): the update is one elementwise stream over them.
 * lrs[5]: means3D, rgb_colors, unnorm_rotations, logit_opacities, log_scales.  step counts from 1. */
int as_adam_step(int P, int isotropic, float* means3D, float* rgb_colors, float* unnorm_rotations,
                 float* logit_opacities, float* log_scales, const float* grad_params, float* exp_avg,
                 float* exp_avg_sq, const float* lrs, float beta1, float beta2, float eps, int step,
                 void* cuda_stream);

/* The same update with the step count kept on the device: step_state is 8 floats of device memory, zeroed by the
 * caller when the optimizer is created (the reference re-creates its optimizer every mapping round,
 * src/mapper/splatam/__init__.py:440).  Each call first advances step_state[0] by one and derives that step's
 * bias corrections from it (in double, as torch does on the host), then updates.  The launch is identical every step,
 * so a mapping iteration that ends in it can be captured once in a CUDA graph and replayed. */
int as_adam_step_counted(int P, int isotropic, float* means3D, float* rgb_colors, float* unnorm_rotations,
                         float* logit_opacities, float* log_scales, const float* grad_params, float* exp_avg,
                         float* exp_avg_sq, const float* lrs, float beta1, float beta2, float eps, float* step_state,
                         void* cuda_stream);

/* as_adam_step_counted restricted to the flat elements [e_begin, e_end) of the segment layout (e_begin a multiple of
 * 4): a mapping step over several GPUs all-reduces the gradient in chunks and updates each chunk as soon as it has
 * arrived, while the next one is still on the wire.  Every chunk of a step sees the same step count; pass
 * advance_step != 0 with the LAST chunk only. */
int as_adam_step_range(int P, int isotropic, float* means3D, float* rgb_colors, float* unnorm_rotations,
                       float* logit_opacities, float* log_scales, const float* grad_params, float* exp_avg,
                       float* exp_avg_sq, const float* lrs, float beta1, float beta2, float eps, float* step_state,
                       long long e_begin, long long e_end, int advance_step, void* cuda_stream);

/* Adam over the camera-pose tensors of a tracking loop (pose-only iterations: learning rates 1e-3 on cam_unnorm_rots
 * [1,4,T] and 4e-3 on cam_trans [1,3,T], torch's default eps 1e-8: initialize_optimizer(tracking=True),
 * splatam.py:118-122, config/splatam/online_habitat_sim.py:37-45).  n_rots / n_trans floats, dense; exp_avg / exp_avg_sq
 * hold n_rots + n_trans floats; step_state as in as_adam_step_counted (zeroed when the optimizer is created). */
int as_pose_adam_step(float* cam_unnorm_rots, const float* grad_rots, int n_rots, float lr_rots, float* cam_trans,
                      const float* grad_trans, int n_trans, float lr_trans, float* exp_avg, float* exp_avg_sq,
                      float beta1, float beta2, float eps, float* step_state, void* cuda_stream);

/* Adam fused with the gradient exchange of keyframe-batched mapping, over peer memory (one process per GPU,
 * NVLink / NVSwitch).  Replaces "all-reduce the flat gradient, then as_adam_step on every rank": rank `rank` of
 * `world` (<= 8) owns the flat elements [e_begin, e_end) of the segment layout ((11 + 1|3) P elements in use); it reads
 * those elements of EVERY rank's gradient buffer through peer pointers, adds them in rank order, applies the Adam update
 * with optimizer state kept only for them (exp_avg / exp_avg_sq hold e_end - e_begin floats), and writes the new values
 * into the parameter tensors of every rank.
 *   params  host array of 5*world device pointers: params[5*r + t] = tensor t (means3D, rgb_colors, unnorm_rotations,
 *           logit_opacities, log_scales) of rank r as mapped into THIS process (cudaIpcOpenMemHandle or equivalent,
 *           peer access enabled); grads: host array of `world` pointers to the ranks' flat gradient buffers.
 * The caller orders the ranks around the call: every rank's gradient must be complete before any rank starts, and no
 * rank may read its parameters or overwrite its gradient before all ranks have finished (two stream-ordered barriers,
 * e.g. tiny NCCL all-reduces).  No reference counterpart: ActiveSplat is single-GPU. */
/* Lets kernels launched on the current device dereference memory of `peer_device` (cudaDeviceEnablePeerAccess;
 * already-enabled is not an error).  Needed once per (device, peer) before as_adam_step_peer. */
int as_enable_peer_access(int peer_device);
/* cudaIpcOpenMemHandle (64-byte handle from cudaIpcGetMemHandle in the exporting process) on the CURRENT device with
 * lazy peer access, so that this device's kernels can use the mapping; returns the base of the exported allocation. */
int as_ipc_open(const void* handle64, void** out_ptr);
int as_ipc_close(void* ptr);
int as_adam_step_peer(int P, int isotropic, int world, int rank, int e_begin, int e_end, float* const* params,
                      const float* const* grads, float* exp_avg, float* exp_avg_sq, const float* lrs, float beta1,
                      float beta2, float eps, int step, void* cuda_stream);

/* ---- the step after the GetOpacity renders: clustering of low-visibility pixels ----------------
 *
 * DBSCAN over the pixels of `batch` images whose value exceeds value_threshold, replacing
 * sklearn.cluster.DBSCAN(eps, min_samples).fit_predict(np.column_stack(np.where(img > thr))) as called by
 * get_invisibility_clusters (src/mapper/__init__.py:92-117; eps 5, min_samples 10, threshold 0.3) and
 * get_convexhull_volume (src/mapper/__init__.py:8-21; eps 5, min_samples 25, threshold 0.8).  The labels are
 * sklearn's, bit for bit: clusters numbered in row-major order of their first core pixel, border pixels joined to
 * the lowest-numbered cluster that reaches them, -1 = noise; -2 marks pixels that are not samples.
 *   values       [batch,H,W] float (device)            labels   [batch,H,W] int32 out
 *   scratch      [batch,H,W] int32 work space          n_clusters [batch] int32 out (may exceed max_clusters)
 *   stats_int    [batch,max_clusters,3] int32 out: member count, sum of rows, sum of columns
 *   stats_value  [batch,max_clusters] double out: sum of values over the members
 * eps <= 8.  One CTA per image; asynchronous on the stream. */
int as_dbscan_grid(int batch, int height, int width, const float* values, float value_threshold, float eps,
                   int min_samples, int* labels, int* scratch, int max_clusters, int* n_clusters, int* stats_int,
                   double* stats_value, void* cuda_stream);

/* Silhouette-driven densification: add_new_gaussians + get_pointcloud + initialize_new_params
 * (src/mapper/splatam/splatam.py:332-379, 25-75, 304-329) on the device, writing the new Gaussians behind row P of
 * parameter tensors that were allocated with room for `capacity` rows (the reference re-allocates all five tensors with
 * torch.cat on every call).  color6 is the rendered bundle of the current map at the frame's pose (as_frame_forward's
 * out_color: channel 3 = depth, 4 = silhouette).  Pixels are taken where the silhouette is below sil_thres, or where
 * the map lies behind the measured depth by more than twice the MEDIAN depth error (exact lower median, as
 * torch.median) on pixels with silhouette above the threshold and measured depth below 5 m; always only where the
 * measured depth is positive.  New rows, in row-major pixel order: mean = c2w * back-projection, colour = gt_rgb,
 * rotation (1,0,0,0), logit opacity 0, log scale log(sqrt((z / ((fx + fy) / 2))^2)), timestep = time_idx.
 *   result (8 x uint32, device): [0] number of new Gaussians, [1] 1 if they did not fit (nothing was written: grow
 *   the tensors and call again), [2] bits of the median depth error, [3] size of the mask before the positive-depth
 *   test (the reference does nothing at all when this is 0, :355).
 * Asynchronous on the stream; the caller reads result[0] to learn the new P. */
typedef struct AsDensifyArgs {
    int width, height;
    const float* color6;      /* [6,H,W] */
    const float* gt_rgb;      /* [3,H,W] */
    const float* gt_depth;    /* [1,H,W] */
    float sil_thres;
    float fx, fy, cx, cy;     /* curr_data['intrinsics'] */
    const float* c2w;         /* [4,4] row-major, device: inverse of the frame's w2c (:357-362) */
    int isotropic;            /* log_scales is [P,1] */
    float time_idx;
    int P, capacity;          /* rows in use / rows allocated in the tensors below */
    float* means3D;           /* [capacity,3] */
    float* rgb_colors;        /* [capacity,3] */
    float* unnorm_rotations;  /* [capacity,4] */
    float* logit_opacities;   /* [capacity,1] */
    float* log_scales;        /* [capacity,1 or 3] */
    float* timestep;          /* [capacity] or NULL (variables['timestep'], :376-377) */
    uint32_t* scratch;        /* as_densify_scratch_words(width, height) x uint32 */
    uint32_t* result;         /* 8 x uint32 */
} AsDensifyArgs;
size_t as_densify_scratch_words(int width, int height);
int as_densify(const AsDensifyArgs* args, void* cuda_stream);

/* The rest of get_convexhull_volume (src/mapper/__init__.py:29-86) for every cluster as_dbscan_grid found in a batch
 * of GLOBAL GetOpacity panoramas (eps 5, min_samples 25, threshold 0.8): the cluster's mask dilated with OpenCV's
 * 15x15 elliptic element (cv2.dilate, :37-38), its outer contour as cv2.findContours(RETR_EXTERNAL,
 * CHAIN_APPROX_SIMPLE) returns it (:41-42), the contour pixels back-projected as (x * rad(360 / width),
 * y * rad(vfov / height), depth[y,x]) skipping depth == 15 (the rasterizer's "no surface" median depth, :49-66), and
 * the volume of their 3-D convex hull (scipy.spatial.ConvexHull(...).volume, :69-78).
 *   labels, n_clusters   as written by as_dbscan_grid          depth    [batch,H,W] float: median-depth panoramas
 *   volumes   [batch,max_clusters] double out                   n_points [batch,max_clusters] int32 out: points that
 *             entered the hull (0: the reference skips the cluster, :60-61; negative: more than 1024 contour vertices,
 *             the hull was built from the first 1024)
 * The per-node answers of the service are then  invisibility = sum_c [n_points_c != 0] stats_value_c * volume_c  and
 * volume = sum_c volume_c (:81-90).  One CTA per (cluster, panorama); asynchronous on the stream. */
int as_cluster_hull_volumes(int batch, int height, int width, const int* labels, const int* n_clusters,
                            const float* depth, double vfov_deg, int max_clusters, double* volumes, int* n_points,
                            void* cuda_stream);

/* ---- stage-level entry points (used by the fused paths and by parity tests) ----------------- */

/* Stable ascending LSD radix sort (hand-written onesweep) of n (u64 key, u32 value) pairs on key bits
 * [0, end_bit).  Replaces cub::DeviceRadixSort::SortPairs as called at rasterizer_impl.cu:305-310.
 * temp must hold as_sort_temp_bytes(n) bytes.  keys_in/values_in are preserved. */
size_t as_sort_temp_bytes(int n);
int as_sort_pairs(const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out,
                  uint32_t* values_out, int n, int end_bit, void* temp, size_t temp_bytes,
                  void* cuda_stream);

/* Inclusive prefix sum of n u32 (single-pass decoupled look-back).  Replaces cub::DeviceScan::InclusiveSum
 * (rasterizer_impl.cu:279).  temp must hold as_scan_temp_bytes(n) bytes; total (last element) is also
 * written to *total_out (device u32) when total_out != NULL. */
size_t as_scan_temp_bytes(int n);
int as_inclusive_scan_u32(const uint32_t* in, uint32_t* out, int n, uint32_t* total_out, void* temp,
                          size_t temp_bytes, void* cuda_stream);

/* Prefix sum for callers that cannot allocate.  (No counterpart in the reference: there every caller
 * owns a torch allocator and CUB temp storage is carved out of it, rasterizer_impl.cu:166,279.)
 *
 * Same result as as_inclusive_scan_u32, but it takes no scratch argument and neither this function
 * nor anything it calls allocates device or host memory.  It is meant for hosts that have no
 * allocator to lend us: e.g. a robot's mapping process that split all of its device memory into
 * fixed arenas at start-up, or a loop with a latency budget in which cudaMalloc / cudaFree (which may
 * synchronise the device) are off limits.  The look-back state lives in a fixed array inside the
 * library image instead (about 16 KiB per GPU that loads the library, held until the process exits).
 *
 *   - n may be at most as_static_scan_max_n() = 4,194,304 elements; larger inputs are refused with
 *     AS_ERR_STATIC_LIMIT and nothing is launched.
 *   - The first call on each device looks up the array's device address (a host-side query) with that
 *     device current; make that first call before starting any stream capture.
 *   - All calls on a device share the one array.  Each call clears the part it needs on `cuda_stream`
 *     before use, so back-to-back calls on one stream are fine; calls that can overlap on the GPU
 *     (different streams, different host threads) race on it and must be serialised by the caller. */
int as_static_scan_max_n(void);
int as_inclusive_scan_u32_static(const uint32_t* in, uint32_t* out, int n, uint32_t* total_out,
                                 void* cuda_stream);

/* getHigherMsb (rasterizer_impl.cu:35-50) */
uint32_t as_higher_msb(uint32_t n);

/* ---- instrumentation (no counterpart in the reference; used by bench.py) ---------------------- */
enum {
    AS_STAGE_PREPROCESS = 0, AS_STAGE_SCAN, AS_STAGE_DUPLICATE, AS_STAGE_SORT, AS_STAGE_RANGES_GATHER,
    AS_STAGE_BLEND_FWD, AS_STAGE_BLEND_BWD, AS_STAGE_PREPROCESS_BWD, AS_STAGE_PROLOGUE, AS_STAGE_LOSS,
    AS_STAGE_ADAM, AS_STAGE_OTHER, AS_NUM_STAGES
};
/* Per-stage CUDA-event timing on the launching stream, process-wide.  Off by default. */
void as_profile_enable(int on);
/* Synchronises the recorded events, writes summed milliseconds and call counts per stage
 * (arrays of n_stages >= AS_NUM_STAGES entries) and clears the record. */
int as_profile_read(double* ms_per_stage, int* calls_per_stage, int n_stages);
/* Number of kernels this library has launched in this process so far. */
long long as_launch_count(void);

const char* as_error_string(int code);
/* cudaError_t value recorded by the last failing call on this host thread (0 = none). */
int as_last_cuda_error(void);
const char* as_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ACTIVESPLAT_B200_H */
