// common.cuh -- shared declarations of libactivesplat_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/activesplat_b200.h"

namespace asb {

// tile-bucket binning: every tile's counter is split TB_SUB ways (by Gaussian id) so that the atomics of a crowded
// tile do not all serialise on one address; a tile's bucket is its TB_SUB sub-buckets back to back
#ifndef ASB_TB_SUB
#define ASB_TB_SUB 8
#endif
constexpr int TB_SUB = ASB_TB_SUB;


constexpr int TILE_X = 16;  // dgr/cuda_rasterizer/config.h:16-17 (BLOCK_X, BLOCK_Y)
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;
constexpr int NUM_SMS = 148;  // B200

// ---------------------------------------------------------------------------------------------
// Scratch layout.  Same idea as rasterizer_impl.h:21-72 (bump allocation out of one byte buffer
// per state), different contents: 256-byte alignment, no CUB temp storage, and the binning state
// carries per-instance records in sorted order so the blend kernels stream them with bulk copies.
// ---------------------------------------------------------------------------------------------
struct Bump {
    char* p;
    explicit __host__ Bump(char* base) : p(base) {}
    template <typename T>
    __host__ T* take(size_t count, size_t align = 256) {
        uintptr_t a = (reinterpret_cast<uintptr_t>(p) + align - 1) & ~(uintptr_t)(align - 1);
        T* r = reinterpret_cast<T*>(a);
        p = reinterpret_cast<char*>(r + count);
        return r;
    }
};

struct GeometryState {  // per Gaussian
    float* depths;            // [P]
    float2* means2D;          // [P]
    float4* conic_opacity;    // [P]
    uint32_t* tiles_touched;  // [P]
    uint32_t* point_offsets;  // [P] inclusive scan of tiles_touched
    int* radii;               // [P] internal copy (used when the caller passes radii == NULL)
    float* gacc;              // [P,12] backward accumulators (conic x,y,w | opacity | mean2D x,y | colour[<=6])
    float* zsil;              // [P] depth in the silhouette frame (frame path: feature z of [z,1,z^2])
    uint32_t* vis_list;       // [P] compact list of visible Gaussian ids (frame path)
    uint32_t* scan_temp;      // look-back state of the scan
    uint32_t* num_rendered;   // [1] device copy of R
    static GeometryState fromChunk(char*& chunk, size_t P);
};

struct ImageState {  // per pixel / per tile
    float* final_T;       // [N]
    uint32_t* n_contrib;  // [N]
    uint2* ranges;        // [tiles]
    uint32_t* tile_count;   // [tiles] instances per tile (frame path, tile-bucket binning)
    uint32_t* tile_cursor;  // [tiles * TB_SUB] next free slot of each tile's sub-buckets
    uint32_t* tile_order;   // [tiles] tile ids, crowded tiles first: launch order of the blend kernels (tile buckets)
    static ImageState fromChunk(char*& chunk, size_t N, size_t tiles);
};

struct BinningState {  // per instance (Gaussian x tile)
    uint64_t* keys_unsorted;    // [R]
    uint64_t* keys_sorted;      // [R]
    uint32_t* values_unsorted;  // [R]
    uint32_t* point_list;       // [R] sorted Gaussian ids
    float4* inst_conic;         // [R] (conic.x, conic.y, conic.z, opacity) in sorted order
    float4* inst_geom;          // [R] (mean2D.x, mean2D.y, depth, id bits)
    float4* inst_feat;          // [R * (channels==3 ? 1 : 2)] features padded to float4
    char* sort_temp;
    size_t sort_temp_bytes;
    static BinningState fromChunk(char*& chunk, size_t R, int channels);
};

size_t scan_temp_bytes(int n);
size_t sort_temp_bytes(int n);
int static_scan_max_n();
cudaError_t static_scan_scratch(void** out);  // device address of the built-in scan scratch (current device)

// ---------------------------------------------------------------------------------------------
// Stage launchers (definitions in the .cu files).  All asynchronous on `st`.
// ---------------------------------------------------------------------------------------------
struct Camera {
    float view[16];
    float proj[16];
    int W, H, gx, gy;
    float tanfovx, tanfovy, focal_x, focal_y;
    float scale_modifier;
};

void launch_preprocess_forward(int P, const float* means3D, const float* scales, const float* rotations,
                               const float* opacities, const float* cov3D_precomp, const float* viewmatrix,
                               const float* projmatrix, int W, int H, float tanfovx, float tanfovy,
                               float scale_modifier, int* radii, const GeometryState& g, cudaStream_t st);

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t st);

void launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, uint32_t* total_out, uint32_t* temp,
                           cudaStream_t st);

void launch_duplicate_with_keys(int P, const GeometryState& g, const int* radii, int W, int H,
                                uint64_t* keys, uint32_t* values, cudaStream_t st);

// n_dev != NULL: the element count is read on the device (min(n, *n_dev)); n is then the capacity the
// grids are sized for, so no host synchronisation is needed to learn it.
void launch_sort_pairs(const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out,
                       uint32_t* values_out, int n, int end_bit, char* temp, cudaStream_t st,
                       const uint32_t* n_dev = nullptr);

void launch_ranges_and_gather(int R, int tiles, int channels, const uint64_t* keys_sorted,
                              const uint32_t* point_list, const GeometryState& g, const float* features,
                              const BinningState& b, uint2* ranges, cudaStream_t st);

int blend_chunk_count_bound(int capacity, int tiles);
size_t blend_chunk_scratch_bytes(int capacity, int tiles);
void launch_blend_forward_chunked(int W, int H, const uint2* ranges, const BinningState& b, int capacity, const float* bg,
                                  float* out_color, float* out_depth, float* out_opacity, float* final_T,
                                  uint32_t* n_contrib, void* scratch, cudaStream_t st);
void launch_blend_forward(int channels, int W, int H, const uint2* ranges, const BinningState& b,
                          const float* bg, float* out_color, float* out_depth, float* out_opacity,
                          float* final_T, uint32_t* n_contrib, cudaStream_t st, const uint32_t* tile_order = nullptr,
                          bool frame_features = false);  // frame_features: channels 3..5 are (z, 1, z^2) (fused frame path)

// gacc: [P,12] accumulators (conic x,y,w | opacity | mean2D x,y | colour[<=6]); must be zero on entry.
// mode (6 channels only): 0 = all 12 terms, 1 = the frame path with Gaussian gradients (the constant-1 channel of
// [z,1,z^2] is skipped), 2 = pose-only frame path (conic, mean2D and the z / z^2 terms; the rest is left undefined).
void launch_blend_backward(int channels, int W, int H, const uint2* ranges, const BinningState& b,
                           const float* bg, const float* final_T, const uint32_t* n_contrib,
                           const float* dL_dpix, float* gacc, cudaStream_t st, int mode = 0,
                           const uint32_t* tile_order = nullptr);

struct PreprocessBackwardOut {
    float* dL_dmean2D;   // [P,3] or NULL
    float* dL_dconic;    // [P,4] or NULL
    float* dL_dopacity;  // [P]
    float* dL_dcolor;    // [P,channels]
    float* dL_dmean3D;   // [P,3]
    float* dL_dcov3D;    // [P,6] or NULL
    float* dL_dscale;    // [P,3] or NULL
    float* dL_drot;      // [P,4] or NULL
};
void launch_preprocess_backward(int P, int channels, const float* means3D, const int* radii, const float* scales,
                                const float* rotations, const float* cov3D_precomp, float scale_modifier,
                                const float* viewmatrix, const float* projmatrix, int W, int H, float tanfovx,
                                float tanfovy, const float* cov3D_state, const float* gacc,
                                const PreprocessBackwardOut& o, cudaStream_t st);

constexpr int GACC_STRIDE = 12;

// ---- fused frame path (fused.cu) ----
// tile_count != NULL (zeroed by the caller): also counts the instances every tile will receive
void launch_frame_preprocess(const AsFrameArgs& a, const GeometryState& g, uint32_t* status, uint32_t* tile_count,
                             cudaStream_t st);
// tile-bucket binning: scan of the tile counts -> ranges, scatter into buckets, per-tile sort + record gather
void launch_frame_tile_binning(int P, const GeometryState& g, int W, int H, uint32_t capacity, const float* rgb,
                               const BinningState& b, const ImageState& img, uint32_t* status, cudaStream_t st);
void launch_frame_duplicate(int P, const GeometryState& g, int W, int H, uint32_t capacity, uint64_t* keys,
                            uint32_t* values, uint32_t* status, cudaStream_t st);
void launch_frame_ranges_gather(const uint32_t* n_dev, uint32_t capacity, int tiles, const GeometryState& g,
                                const float* rgb, const BinningState& b, uint2* ranges, cudaStream_t st);
void launch_frame_preprocess_backward(const AsFrameArgs& a, const GeometryState& g, const float* gscale,
                                      int gaussians_grad, int camera_grad, int accumulate, float* grad_params,
                                      float* grad_means2D, float* pose_acc, float* grad_cam, int grad_cam_stride,
                                      cudaStream_t st);
int tile_sort_capacity();
int loss_partial_blocks(int N);
size_t pose_scratch_floats(int P);
void launch_tracking_loss(int N, const float* c6, const float* gt_rgb, const float* gt_depth, float w_im,
                          float w_depth, float sil_thres, int use_sil, int mask_color, float* dL, float* partials,
                          float* losses, const uint32_t* status, cudaStream_t st);
size_t mapping_loss_scratch_floats(int W, int H);
void launch_mapping_loss(int W, int H, const float* c6, const float* gt_rgb, const float* gt_depth, float w_im,
                         float w_depth, float* dL, float* losses, float* scratch, const uint32_t* status,
                         cudaStream_t st);
void launch_adam_peer(int P, int e_begin, int e_end, int iso, int world, int rank, float* const* params, const float* const* grads,
                      const float lr[5], float* m, float* v, float beta1, float beta2, float eps, int step,
                      cudaStream_t st);
// DBSCAN on a pixel grid, one CTA per image (scoring.cu)
void launch_dbscan_grid(int batch, int H, int W, const float* values, float thr, float eps, int min_samples, int* labels,
                        int* scratch, int max_clusters, int* n_clusters, int* stats_i, double* stats_v, cudaStream_t st);
size_t densify_scratch_words(int W, int H);
void launch_densify(int W, int H, const float* c6, const float* gt_rgb, const float* gt_depth, float sil_thres, float fx,
                    float fy, float cx, float cy, const float* c2w, int iso, float time_idx, int P, int capacity,
                    float* means3D, float* rgb, float* rot, float* logit, float* log_scales, float* timestep,
                    uint32_t* scratch, uint32_t* result, cudaStream_t st);
int launch_cluster_hull(int batch, int H, int W, const int* labels, const int* n_clusters, const float* depth,
                        double vfov_deg, int max_clusters, double* volumes, int* n_points, cudaStream_t st);
void launch_adam(int P, int iso, float* const params[5], const float lr[5], const float* grad, float* m, float* v,
                 float beta1, float beta2, float eps, int step, cudaStream_t st);
void launch_pose_adam(float* rots, const float* g_rots, int n_rots, float lr_rots, float* trans, const float* g_trans,
                      int n_trans, float lr_trans, float* m, float* v, float beta1, float beta2, float eps,
                      float* step_state, cudaStream_t st);
void launch_adam_counted(int P, int iso, float* const params[5], const float lr[5], const float* grad, float* m,
                         float* v, float beta1, float beta2, float eps, float* step_state, size_t e_begin, size_t e_end,
                         int advance, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// mbarrier + 1-D bulk async copy (TMA engine; SASS UBLKCP)
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {   // completes a phase of a count-1 barrier without data
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    }
}
// vector float reduction to global memory (SASS REDG.E.ADD.F32x4)
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}
__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_volatile_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// a0*b0 + a1*b1 + a2*b2 with the roundings nvcc gives the reference's GLM/auxiliary.h expressions:
// fma(a2,b2, fma(a0,b0, a1*b1)).  Written with intrinsics so no compiler version can re-contract it.
__device__ __forceinline__ float dot3c(float a0, float b0, float a1, float b1, float a2, float b2) {
    return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}
#endif

}  // namespace asb
