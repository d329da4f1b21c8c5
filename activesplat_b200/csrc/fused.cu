// fused.cu -- the SplaTAM frame step around the rasterizer, fused (sm_100a).
//
// The reference runs these as dozens of torch-eager kernels per iteration (SURVEY 2.2 "F1-F5"):
//   F1 transform_to_frame                    src/mapper/splatam/utils/slam_helpers.py:252-304
//   F2 activations (sigmoid/exp/normalize)   slam_helpers.py:124-139
//   F3 [z, 1, z^2] features                  slam_helpers.py:196-213
//   F4 loss (masked L1, SSIM)                src/mapper/splatam/splatam.py:213-249, utils/slam_external.py:54-97
//   F5 Adam                                  splatam.py:118-124 (torch.optim.Adam, eps 1e-15)
// Here: frame_preprocess_kernel = F1+F2+F3+preprocessCUDA in one pass over the raw parameters;
// frame_preprocess_backward_kernel = computeCov2D/preprocess backward + the backward of F1-F3 down to the
// raw parameters and the camera pose (block-reduced); tracking_loss_kernel / mapping loss kernels = F4
// forward and backward in one or two sweeps over the image; adam_kernel = F5 over all five tensors.
// None of the frame path synchronises with the host: the instance count stays on the device and the
// binning buffer has a caller-chosen capacity (overflow is flagged, never written past).
#include "raster_math.cuh"

namespace asb {

namespace {

constexpr int FP_THREADS = 256;
#ifndef ASB_FP_GPB
#define ASB_FP_GPB 4
#endif
constexpr int FP_GPB = ASB_FP_GPB;   // Gaussians per thread of the frame preprocess (its cheap visibility screen)

struct Pose {        // shared-memory copy of the frame pose
    float q[4];      // normalised camera quaternion (r,x,y,z)
    float R[9];      // rotation of q (row-major, textbook)
    float t[3];
    float qnorm;     // |cam_unnorm_rot|
    float sil[4];    // row 2 of the silhouette w2c
};

__device__ __forceinline__ void quat_to_rot(const float q[4], float R[9]) {
    // build_rotation (slam_external.py:25-42) normalises once more
    const float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const float r = q[0] / n, x = q[1] / n, y = q[2] / n, z = q[3] / n;
    R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - r * z); R[2] = 2.f * (x * z + r * y);
    R[3] = 2.f * (x * y + r * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - r * x);
    R[6] = 2.f * (x * z - r * y); R[7] = 2.f * (y * z + r * x); R[8] = 1.f - 2.f * (x * x + y * y);
}

__device__ __forceinline__ void load_pose(Pose& ps, const float* cam_rot, int rs, const float* cam_trans, int ts,
                                          const float* sil_w2c) {
    float q[4];
#pragma unroll
    for (int k = 0; k < 4; k++) q[k] = __ldg(cam_rot + (size_t)k * rs);
    const float n = fmaxf(sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), 1e-12f);  // F.normalize
    ps.qnorm = n;
#pragma unroll
    for (int k = 0; k < 4; k++) ps.q[k] = q[k] / n;
    quat_to_rot(ps.q, ps.R);
#pragma unroll
    for (int k = 0; k < 3; k++) ps.t[k] = __ldg(cam_trans + (size_t)k * ts);
    if (sil_w2c) {
#pragma unroll
        for (int k = 0; k < 4; k++) ps.sil[k] = __ldg(sil_w2c + 8 + k);
    } else {
        ps.sil[0] = 0.f; ps.sil[1] = 0.f; ps.sil[2] = 1.f; ps.sil[3] = 0.f;
    }
}

// quat_mult (slam_helpers.py:21-28)
__device__ __forceinline__ void quat_mul(const float a[4], const float b[4], float o[4]) {
    o[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
    o[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
    o[2] = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
    o[3] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
}

// Activations + frame transform of one Gaussian (F1+F2).  rot_t is the un-normalised transformed rotation.
struct Activated {
    float mc[3];     // mean in the frame
    float qn[4];     // normalised own rotation (anisotropic only)
    float qn_norm;   // |unnorm_rotation|
    float rot_t[4];  // quat_mult(q_cam, qn)  (anisotropic)  or the raw rotation (isotropic)
    float rot_t_norm;
    float rot[4];    // normalised rotation fed to the rasterizer
    float scale[3];
    float opacity;
};

template <bool ISO>
__device__ __forceinline__ void activate(const Pose& ps, float px, float py, float pz, float4 qraw, const float* ls,
                                         float lo, Activated& a) {
    a.mc[0] = fmaf(ps.R[2], pz, fmaf(ps.R[1], py, ps.R[0] * px)) + ps.t[0];
    a.mc[1] = fmaf(ps.R[5], pz, fmaf(ps.R[4], py, ps.R[3] * px)) + ps.t[1];
    a.mc[2] = fmaf(ps.R[8], pz, fmaf(ps.R[7], py, ps.R[6] * px)) + ps.t[2];
    const float q[4] = {qraw.x, qraw.y, qraw.z, qraw.w};
    if (ISO) {
#pragma unroll
        for (int k = 0; k < 4; k++) a.rot_t[k] = q[k];
        a.qn_norm = 1.f;
    } else {
        a.qn_norm = fmaxf(sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), 1e-12f);
#pragma unroll
        for (int k = 0; k < 4; k++) a.qn[k] = q[k] / a.qn_norm;
        quat_mul(ps.q, a.qn, a.rot_t);
    }
    a.rot_t_norm = fmaxf(sqrtf(a.rot_t[0] * a.rot_t[0] + a.rot_t[1] * a.rot_t[1] + a.rot_t[2] * a.rot_t[2] +
                               a.rot_t[3] * a.rot_t[3]), 1e-12f);
#pragma unroll
    for (int k = 0; k < 4; k++) a.rot[k] = a.rot_t[k] / a.rot_t_norm;
    a.scale[0] = expf(ls[0]);
    a.scale[1] = ISO ? a.scale[0] : expf(ls[1]);
    a.scale[2] = ISO ? a.scale[0] : expf(ls[2]);
    a.opacity = 1.f / (1.f + expf(-lo));
}

__device__ __forceinline__ void load_rows(const float* __restrict__ src, int row0, int nrows, int width, float* sm) {
    const float* p = src + (size_t)row0 * width;
    const int nfl = nrows * width, nt = blockDim.x;
    if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const int nv = nfl >> 2;
        const float4* p4 = reinterpret_cast<const float4*>(p);
        for (int i = threadIdx.x; i < nv; i += nt) reinterpret_cast<float4*>(sm)[i] = __ldg(p4 + i);
        for (int i = (nv << 2) + threadIdx.x; i < nfl; i += nt) sm[i] = __ldg(p + i);
    } else {
        for (int i = threadIdx.x; i < nfl; i += nt) sm[i] = __ldg(p + i);
    }
}

// Look-back state of the scan fused into frame_preprocess_kernel: one u64 per block,
// flag (2 bits, 62..63) | visible count (24 bits, 32..55) | tiles touched (32 bits, 0..31).

constexpr uint64_t FPS_AGG = 1ull << 62;
constexpr uint64_t FPS_PREFIX = 2ull << 62;
constexpr uint64_t FPS_VALUE_MASK = (1ull << 56) - 1ull;

// ORDERED: point_offsets (the inclusive scan of tiles touched, rasterizer_impl.cu:279) and the visible list come out
// in Gaussian order through a decoupled look-back -- what the global sort's duplicateWithKeys needs.  With tile
// buckets nothing depends on that order, so blocks just reserve their span of the visible list with one atomic
// and add their totals to status[0] / status[2] (zeroed by the launcher): no block ever waits for another.
// NT threads = Gaussians per block.
// GPB Gaussians per thread: a block screens a slab of NT * GPB Gaussians, so that the exact path behind the screen (a
// fifth to a quarter of them survive) finds a candidate for nearly every thread instead of for two warps out of eight.
template <bool ISO, bool ORDERED, int NT, int GPB>
__global__ void __launch_bounds__(NT)
frame_preprocess_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ unnorm_rot,
                        const float* __restrict__ logit_op, const float* __restrict__ log_scales,
                        const float* __restrict__ cam_rot, int rs, const float* __restrict__ cam_trans, int ts,
                        const float* __restrict__ sil_w2c, const float* __restrict__ viewmatrix,
                        const float* __restrict__ projmatrix, int W, int H, int gx, int gy, float tanfovx,
                        float tanfovy, float focal_x, float focal_y, float scale_modifier, int* __restrict__ radii,
                        int* __restrict__ radii_internal, float2* __restrict__ means2D_out,
                        float* __restrict__ depths, float4* __restrict__ conic_opacity,
                        uint32_t* __restrict__ point_offsets, float* __restrict__ zsil,
                        uint32_t* __restrict__ vis_list, uint32_t* ticket, uint64_t* scan_status, uint32_t* status,
                        float* __restrict__ max_2D_radius, uint8_t* __restrict__ seen,
                        uint32_t* __restrict__ tile_count) {
    constexpr int SLAB = NT * GPB;
    __shared__ __align__(16) float s_means[SLAB * 3];
    __shared__ __align__(16) float s_ls[SLAB * (ISO ? 1 : 3)];
    __shared__ float s_view[16], s_proj[16];
    __shared__ Pose ps;
    __shared__ uint32_t s_block;
    __shared__ uint32_t s_wt[NT / 32], s_wv[NT / 32];
    __shared__ uint32_t s_pref_t, s_pref_v;
    __shared__ int s_rad[SLAB];
    __shared__ uint32_t s_til[SLAB];
    __shared__ uint16_t s_cand[SLAB];
    __shared__ uint32_t s_ncand;
    __shared__ float s_bound_k;
    // blocks take their slab in the order they start, so the look-back below can never wait on a
    // block that has not been scheduled yet
    if (ORDERED) {
        if (threadIdx.x == 0) s_block = atomicAdd(ticket, 1u);
        __syncthreads();
    }
    const uint32_t blk = ORDERED ? s_block : blockIdx.x;
    const int row0 = blk * SLAB;
    const int nrows = min(SLAB, P - row0);
    load_rows(means3D, row0, nrows, 3, s_means);
    load_rows(log_scales, row0, nrows, ISO ? 1 : 3, s_ls);
    if (threadIdx.x < 16) s_view[threadIdx.x] = __ldg(viewmatrix + threadIdx.x);
    else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = __ldg(projmatrix + threadIdx.x - 16);
    else if (threadIdx.x == 32) load_pose(ps, cam_rot, rs, cam_trans, ts, sil_w2c);
    else if (threadIdx.x == 33) s_ncand = 0;
    else if (threadIdx.x == 64) {
        // constant of the radius bound of phase A: |J|_F^2 tz^2 <= fx^2 (1 + (1.3 tanfovx)^2) + fy^2 (1 + (1.3 tanfovy)^2)
        // (cov2d clamps tx/tz, ty/tz: forward.cu:82-87), times |W|_2^2 <= largest Gershgorin row sum of W^T W
        float w[9];
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) w[3 * i + j] = __ldg(viewmatrix + 4 * j + i);  // rotation part (transposed layout)
        float gmax = 0.f;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            float rowsum = 0.f;
#pragma unroll
            for (int j = 0; j < 3; j++) rowsum += fabsf(w[i] * w[j] + w[3 + i] * w[3 + j] + w[6 + i] * w[6 + j]);
            gmax = fmaxf(gmax, rowsum);
        }
        const float lx = 1.3f * tanfovx, ly = 1.3f * tanfovy;
        s_bound_k = (focal_x * focal_x * (1.f + lx * lx) + focal_y * focal_y * (1.f + ly * ly)) * gmax;
    }
    if (ORDERED && blk == 0 && threadIdx.x == 0) status[1] = 0;  // overflow flag of this frame
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // ---- phase A (thread = GPB Gaussians, one after the other): cheap, conservative visibility test.  Usually most
    //      of the map is behind the camera or off the image; the expensive per-Gaussian path below then runs densely
    //      over the candidates instead of dragging every warp through it for a few scattered lanes.
    //      A Gaussian is dropped here only if the reference would give it radius 0 / no tiles: near cull
    //      (auxiliary.h:154), or an empty tile rectangle even for an upper bound of its radius:
    //      lambda_max(cov2D + 0.3 I) <= trace(J W S W^T J^T) + 0.92 <= |J|_F^2 |W|_2^2 (mod * s_max)^2 + 0.92.
#pragma unroll
    for (int g = 0; g < GPB; g++) {
        const int li = threadIdx.x + g * NT;   // slot of the slab
        bool cand = false;
        if (row0 + li < P) {
            const float px = s_means[3 * li], py = s_means[3 * li + 1], pz = s_means[3 * li + 2];
            const float m0 = fmaf(ps.R[2], pz, fmaf(ps.R[1], py, ps.R[0] * px)) + ps.t[0];  // as in activate()
            const float m1 = fmaf(ps.R[5], pz, fmaf(ps.R[4], py, ps.R[3] * px)) + ps.t[1];
            const float m2 = fmaf(ps.R[8], pz, fmaf(ps.R[7], py, ps.R[6] * px)) + ps.t[2];
            const float pvz = xform_row(s_view, 2, m0, m1, m2);
            if (pvz > 0.2f) {
                const float* ls = s_ls + (ISO ? 1 : 3) * li;
                const float smax = expf(ISO ? ls[0] : fmaxf(ls[0], fmaxf(ls[1], ls[2]))) * fabsf(scale_modifier);
                const float tr = s_bound_k * smax * smax / (pvz * pvz);
                const float rb = ceilf(3.f * sqrtf(tr * 1.001f + 0.93f)) + 1.f;
                const float hx = xform_row(s_proj, 0, m0, m1, m2), hy = xform_row(s_proj, 1, m0, m1, m2);
                const float hw = xform_row(s_proj, 3, m0, m1, m2);
                const float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
                const float pix_x = ndc2pix(__fmul_rn(hx, p_w), W), pix_y = ndc2pix(__fmul_rn(hy, p_w), H);
                cand = true;
                if (rb < 1e8f && pix_x == pix_x && pix_y == pix_y) {  // anything not finite goes down the exact path
                    int x0, y0, x1, y1;
                    get_rect(pix_x, pix_y, __float2int_rz(rb), gx, gy, x0, y0, x1, y1);
                    cand = (x1 - x0) * (y1 - y0) != 0;
                }
            }
        }
        s_rad[li] = 0;
        s_til[li] = 0;
        // candidates go to the list in arrival order (one shared-memory atomic per warp and round): nothing below
        // depends on their order
        const uint32_t bal = __ballot_sync(0xffffffffu, cand);
        uint32_t base = 0;
        if (lane == 0 && bal) base = atomicAdd(&s_ncand, (uint32_t)__popc(bal));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (cand) s_cand[base + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)li;
    }
    __syncthreads();
    // ---- phase B (thread = candidate): the exact per-Gaussian path ----
    const int ncand = (int)s_ncand;
    for (int c = threadIdx.x; c < ncand; c += NT) {
        const int slot = s_cand[c];
        const int cidx = row0 + slot;
        Activated a;
        activate<ISO>(ps, s_means[3 * slot], s_means[3 * slot + 1], s_means[3 * slot + 2],
                      __ldg(reinterpret_cast<const float4*>(unnorm_rot) + cidx), s_ls + (ISO ? 1 : 3) * slot,
                      __ldg(logit_op + cidx), a);
        if (xform_row(s_view, 2, a.mc[0], a.mc[1], a.mc[2]) > 0.2f) {
            float cov3D[6];
            cov3d_from_scale_rot(a.scale[0], a.scale[1], a.scale[2], scale_modifier,
                                 make_float4(a.rot[0], a.rot[1], a.rot[2], a.rot[3]), cov3D);
            const PreprocessResult r = preprocess_one(a.mc[0], a.mc[1], a.mc[2], cov3D, a.opacity, s_view, s_proj, W, H,
                                                      gx, gy, tanfovx, tanfovy, focal_x, focal_y);
            if (r.radius > 0) {
                s_rad[slot] = r.radius;
                s_til[slot] = r.tiles;
                depths[cidx] = r.depth;
                means2D_out[cidx] = r.mean2D;
                conic_opacity[cidx] = r.conic_opacity;
                zsil[cidx] = ps.sil[0] * a.mc[0] + ps.sil[1] * a.mc[1] + ps.sil[2] * a.mc[2] + ps.sil[3];
                if (tile_count) {  // tile-bucket binning: how many instances each tile will receive
                    int x0, y0, x1, y1;
                    get_rect(r.mean2D.x, r.mean2D.y, r.radius, gx, gy, x0, y0, x1, y1);
                    const uint32_t sub = (uint32_t)cidx & (TB_SUB - 1);
                    for (int y = y0; y < y1; y++)
                        for (int x = x0; x < x1; x++) atomicAdd(tile_count + ((y * gx + x) * TB_SUB + sub), 1u);
                }
            }
        }
    }
    __syncthreads();
    // per-Gaussian outputs, coalesced: slot threadIdx.x + g * NT
#pragma unroll
    for (int g = 0; g < GPB; g++) {
        const int li = threadIdx.x + g * NT, idx = row0 + li;
        if (idx < P) {
            const int rad = s_rad[li];
            if (radii) radii[idx] = rad;
            radii_internal[idx] = rad;
            // bookkeeping of get_loss (splatam.py:296-298): max_2D_radius[seen] = max(radius, .), seen = radius > 0
            if (max_2D_radius && rad > 0) max_2D_radius[idx] = fmaxf(max_2D_radius[idx], (float)rad);
            if (seen) seen[idx] = rad > 0 ? 1 : 0;
        }
    }
    // ---- fused prefix sums (replaces the separate InclusiveSum launch, rasterizer_impl.cu:279):
    //      tiles touched -> point_offsets, visible count -> compact list of visible Gaussians.
    //      Thread t owns the GPB consecutive slots t * GPB ..: a serial prefix inside the thread, then a warp scan of the
    //      threads' totals ----
    uint32_t my_t[GPB], my_v = 0, sum_t = 0, cnt_v = 0;   // inclusive tile sums, visibility bits
#pragma unroll
    for (int g = 0; g < GPB; g++) {
        const int li = threadIdx.x * GPB + g;
        sum_t += s_til[li];
        my_t[g] = sum_t;
        if (s_rad[li] > 0) { my_v |= 1u << g; cnt_v++; }
    }
    uint32_t incl = sum_t, incl_v = cnt_v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o), tv = __shfl_up_sync(0xffffffffu, incl_v, o);
        if (lane >= o) { incl += t; incl_v += tv; }
    }
    if (lane == 31) { s_wt[warp] = incl; s_wv[warp] = incl_v; }
    __syncthreads();
    uint32_t woff_t = 0, woff_v = 0, tot_t = 0, tot_v = 0;
#pragma unroll
    for (int w = 0; w < NT / 32; w++) {
        if (w < warp) { woff_t += s_wt[w]; woff_v += s_wv[w]; }
        tot_t += s_wt[w];
        tot_v += s_wv[w];
    }
    if (!ORDERED) {
        if (threadIdx.x == 0) {
            s_pref_v = tot_v ? atomicAdd(status + 2, tot_v) : 0u;  // my span of the visible list
            if (tot_t) atomicAdd(status + 0, tot_t);               // num_rendered
        }
    } else if (warp == 0) {
        // decoupled look-back, one warp wide: lane l inspects slab blk-1-l, the window slides back by 32
        const uint64_t mine = ((uint64_t)tot_v << 32) | tot_t;
        if (lane == 0 && blk > 0) st_volatile_u64(scan_status + blk, FPS_AGG | mine);
        uint64_t prefix = 0;
        int base = (int)blk - 1;
        while (base >= 0) {
            const int p = base - lane;
            uint64_t v = FPS_PREFIX;  // slabs before the first one: an empty inclusive prefix
            if (p >= 0) {
                do { v = ld_volatile_u64(scan_status + p); } while ((v >> 62) == 0);
            }
            const uint32_t is_prefix = __ballot_sync(0xffffffffu, (v >> 62) == 2);
            const int first = is_prefix ? (__ffs(is_prefix) - 1) : 31;
            // fields cannot carry into each other: tiles total < 2^32, visible total < 2^24
            uint64_t c = (lane <= first) ? (v & FPS_VALUE_MASK) : 0ull;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
            prefix += c;
            if (is_prefix) break;
            base -= 32;
        }
        if (lane == 0) {
            st_volatile_u64(scan_status + blk, FPS_PREFIX | (prefix + mine));
            s_pref_t = (uint32_t)prefix;
            s_pref_v = (uint32_t)(prefix >> 32);
            if ((int)blk == (P + SLAB - 1) / SLAB - 1) {  // the last slab knows the totals
                status[0] = (uint32_t)prefix + tot_t;          // num_rendered
                status[2] = (uint32_t)(prefix >> 32) + tot_v;  // number of visible Gaussians
            }
        }
    }
    __syncthreads();
    const uint32_t before_t = s_pref_t + woff_t + incl - sum_t;      // tiles of everything before this thread's slots
    uint32_t pos_v = s_pref_v + woff_v + incl_v - cnt_v;
#pragma unroll
    for (int g = 0; g < GPB; g++) {
        const int idx = row0 + threadIdx.x * GPB + g;
        if (idx < P) {
            if (ORDERED) point_offsets[idx] = before_t + my_t[g];
            if ((my_v >> g) & 1u) vis_list[pos_v++] = (uint32_t)idx;
        }
    }
}

// duplicateWithKeys (rasterizer_impl.cu:70-111) with a capacity guard instead of a host-sized buffer
__global__ void __launch_bounds__(256)
frame_duplicate_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                       const uint32_t* __restrict__ offsets, const int* __restrict__ radii, int gx, int gy,
                       uint32_t capacity, uint64_t* __restrict__ keys, uint32_t* __restrict__ values,
                       uint32_t* status) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const int r = radii[idx];
    if (r <= 0) return;
    if (offsets[idx] > capacity) {  // this Gaussian's slots end past the buffer: flag, write nothing
        status[1] = 1;
        return;
    }
    uint32_t off = idx == 0 ? 0u : offsets[idx - 1];
    const float2 p = means2D[idx];
    int x0, y0, x1, y1;
    get_rect(p.x, p.y, r, gx, gy, x0, y0, x1, y1);
    const uint64_t dbits = (uint64_t)__float_as_uint(depths[idx]);
    for (int y = y0; y < y1; y++)
        for (int x = x0; x < x1; x++) {
            keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
            values[off] = (uint32_t)idx;
            off++;
        }
}

// identifyTileRanges + record gather for the frame path: features are (rgb | z, 1, z^2) built on the fly
__global__ void __launch_bounds__(256)
frame_ranges_gather_kernel(const uint32_t* __restrict__ n_dev, uint32_t capacity, const uint64_t* __restrict__ keys,
                           const uint32_t* __restrict__ point_list, const float2* __restrict__ means2D,
                           const float4* __restrict__ conic_opacity, const float* __restrict__ rgb,
                           const float* __restrict__ zsil, float4* __restrict__ inst_conic,
                           float4* __restrict__ inst_geom, float4* __restrict__ inst_feat, uint2* __restrict__ ranges,
                           uint32_t tiles) {
    if (n_dev[1]) return;  // overflow: the key buffer is incomplete; leave every range empty
    const uint32_t R = min(*n_dev, capacity);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint64_t key = keys[i];
    const uint32_t cur = (uint32_t)(key >> 32);
    if (cur >= tiles) return;
    if (i == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
        if (cur != prev) {
            ranges[prev].y = i;
            ranges[cur].x = i;
        }
    }
    if (i == R - 1) ranges[cur].y = R;
    const uint32_t id = point_list[i];
    const float2 m = __ldg(means2D + id);
    const float4 co = __ldg(conic_opacity + id);
    inst_conic[i] = co;
    inst_geom[i] = make_float4(m.x, m.y, __uint_as_float((uint32_t)key), __uint_as_float(id));
    const float z = __ldg(zsil + id);
    const float* c = rgb + 3 * (size_t)id;
    inst_feat[2 * (size_t)i] = make_float4(__ldg(c), __ldg(c + 1), __ldg(c + 2), z);
    inst_feat[2 * (size_t)i + 1] = make_float4(1.f, z * z, pack_cull_extent(co), 0.f);
}

// ---- tile-bucket binning (AsFrameArgs.binning_mode = 1) -------------------------------------------------
// The reference sorts all R (tile, depth) keys globally (rasterizer_impl.cu:305-310).  The tile part of the key is
// known before any sorting, so here every tile gets its own bucket: the preprocess counts instances per tile, one
// block scans the counts into tile ranges, the scatter drops (depth bits << 32 | Gaussian id) into the tile's
// bucket in arrival order, and one CTA per tile sorts its bucket in shared memory and gathers the records.
// Sorting by (depth, id) reproduces the stable radix sort exactly: the reference emits instances in ascending
// Gaussian id, so equal (tile, depth) keys stay in id order.
constexpr int TS_THREADS = 256;
#ifndef ASB_TS_GATHER
#define ASB_TS_GATHER 2
#endif
#ifndef ASB_TS_MERGE
#define ASB_TS_MERGE 4
#endif
constexpr int TS_MERGE = ASB_TS_MERGE;     // runs whose lower bounds a thread searches together
constexpr int TS_GATHER = ASB_TS_GATHER;   // records a thread gathers per round (their scattered reads in flight together)
constexpr int TS_CAP = 3584;  // instances per tile sorted in shared memory (28 KB: 8 CTAs per SM); larger buckets are sorted in place (slow)

__global__ void __launch_bounds__(1024)
frame_tile_scan_kernel(const uint32_t* __restrict__ tile_count, int tiles, uint32_t capacity, uint2* __restrict__ ranges,
                       uint32_t* __restrict__ cursor, uint32_t* __restrict__ tile_order, uint32_t* status) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_max;
    __shared__ uint32_t s_cls[32], s_base[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_max = 0;
    __syncthreads();
    const bool overflow = status[0] > capacity;  // num_rendered (written by the preprocess) exceeds the buffer
    // every thread owns `per` consecutive tiles (TB_SUB sub-counters each): local sums, one block scan, write-out
    const int per = (tiles + 1023) / 1024;
    const int t0 = threadIdx.x * per, t1 = min(tiles, t0 + per);
    uint32_t mine = 0, biggest = 0;
    for (int t = t0; t < t1; t++) {
        const uint4* p = reinterpret_cast<const uint4*>(tile_count + (size_t)t * TB_SUB);
        uint32_t c = 0;
#pragma unroll
        for (int q = 0; q < TB_SUB / 4; q++) {
            const uint4 v = p[q];
            c += v.x + v.y + v.z + v.w;
        }
        mine += c;
        biggest = max(biggest, c);
    }
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, biggest);
    if (lane == 0) atomicMax(&s_max, wmax);
    __syncthreads();
    if (warp == 0) {
        uint32_t w = s_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        s_warp[lane] = wi - w;  // exclusive offset of each warp
    }
    __syncthreads();
    uint32_t run = s_warp[warp] + incl - mine;
    for (int t = t0; t < t1; t++) {
        const uint4* p = reinterpret_cast<const uint4*>(tile_count + (size_t)t * TB_SUB);
        const uint32_t start = run;
#pragma unroll
        for (int q = 0; q < TB_SUB / 4; q++) {
            const uint4 v = p[q];
            uint4 cur;
            cur.x = run; run += v.x;
            cur.y = run; run += v.y;
            cur.z = run; run += v.z;
            cur.w = run; run += v.w;
            reinterpret_cast<uint4*>(cursor + (size_t)t * TB_SUB)[q] = cur;
        }
        // untouched tiles keep {0,0} like identifyTileRanges leaves them (rasterizer_impl.cu:116-138)
        ranges[t] = (run > start && !overflow) ? make_uint2(start, run) : make_uint2(0u, 0u);
    }
    if (threadIdx.x == 0) {
        status[3] = s_max;
        if (overflow) status[1] = 1;
    }
    // Launch order for the blend kernels: crowded tiles first (32 size classes, counting sort), so that the last
    // CTAs of their grids are the short ones and the tail of the kernel is short too.
    if (threadIdx.x < 32) s_cls[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t width = s_max / 32u + 1u;
    for (int t = t0; t < t1; t++) {
        const uint2 r = ranges[t];
        atomicAdd(&s_cls[min(31u, (r.y - r.x) / width)], 1u);
    }
    __syncthreads();
    if (warp == 0) {
        const uint32_t c = s_cls[lane];
        uint32_t incl2 = c;  // inclusive sum over classes lane..31 (descending order of size)
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_down_sync(0xffffffffu, incl2, o);
            if (lane + o < 32) incl2 += v;
        }
        s_base[lane] = incl2 - c;
        s_cls[lane] = 0;
    }
    __syncthreads();
    for (int t = t0; t < t1; t++) {
        const uint2 r = ranges[t];
        const uint32_t c = min(31u, (r.y - r.x) / width);
        tile_order[s_base[c] + atomicAdd(&s_cls[c], 1u)] = (uint32_t)t;
    }
}

__global__ void __launch_bounds__(256)
frame_scatter_kernel(const uint32_t* __restrict__ vis_list, const float2* __restrict__ means2D,
                     const float* __restrict__ depths, const int* __restrict__ radii, int gx, int gy,
                     uint32_t* __restrict__ cursor, uint64_t* __restrict__ entries, const uint32_t* __restrict__ status) {
    if (status[1]) return;  // overflow: nothing may be written
    const uint32_t nvis = status[2];
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= nvis) return;
    const uint32_t idx = vis_list[slot];
    const float2 p = means2D[idx];
    int x0, y0, x1, y1;
    get_rect(p.x, p.y, radii[idx], gx, gy, x0, y0, x1, y1);
    const uint64_t e = ((uint64_t)__float_as_uint(depths[idx]) << 32) | idx;
    const uint32_t sub = idx & (TB_SUB - 1);
    // the slot of a tile comes back from an atomic: keep eight of them in flight before the dependent stores
    const int w = x1 - x0, total = w * (y1 - y0);
    int x = x0, y = y0;
    for (int base = 0; base < total; base += 8) {
        uint32_t pos[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (base + u < total) {
                pos[u] = atomicAdd(cursor + ((y * gx + x) * TB_SUB + sub), 1u);
                if (++x == x1) { x = x0; y++; }
            }
        }
#pragma unroll
        for (int u = 0; u < 8; u++)
            if (base + u < total) entries[pos[u]] = e;
    }
}

// compare-exchange network, ascending only (first step of every merge mirrors the upper half), so the
// missing tail of a bucket whose size is not a power of two behaves like +infinity and never moves.
// Compare-exchange t of a stage touches only the aligned 64-element block 2*32*(t/32).. when the stage's span is
// at most 64, and each warp always takes the same t's: those stages need only a warp-level barrier.
__device__ __forceinline__ void stage_barrier(bool block_wide) {
    if (block_wide) __syncthreads();
    else __syncwarp();
}

template <class Ptr>
__device__ __forceinline__ void bitonic_sort_ascending(Ptr a, int n, int nthreads) {
    int npow2 = 1;
    while (npow2 < n) npow2 <<= 1;
    for (int k = 2, lk = 1; k <= npow2; k <<= 1, lk++) {  // k = 1 << lk
        for (int t = threadIdx.x; t < npow2 / 2; t += nthreads) {
            const int blk = t >> (lk - 1), off = t & ((k >> 1) - 1);
            const int i = (blk << lk) + off, j = (blk << lk) + k - 1 - off;
            if (j < n) {
                const uint64_t x = a[i], y = a[j];
                if (x > y) { a[i] = y; a[j] = x; }
            }
        }
        stage_barrier(k > 64);  // what follows has a smaller span (or, for k == 2, span 4)
        for (int js = k >> 2, ljs = lk - 2; js > 0; js >>= 1, ljs--) {  // js = 1 << ljs
            for (int t = threadIdx.x; t < npow2 / 2; t += nthreads) {
                const int i = ((t >> ljs) << (ljs + 1)) + (t & (js - 1)), j = i + js;
                if (j < n) {
                    const uint64_t x = a[i], y = a[j];
                    if (x > y) { a[i] = y; a[j] = x; }
                }
            }
            // the next stage is either js/2 of this merge or, after js == 1, the span-2k flip of the next merge
            const int next_span = js > 1 ? js : 2 * k;
            stage_barrier(2 * js > 64 || next_span > 64);
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(TS_THREADS)
frame_tile_sort_gather_kernel(const uint2* __restrict__ ranges, uint64_t* __restrict__ entries,
                              uint32_t* __restrict__ point_list, const float2* __restrict__ means2D,
                              const float4* __restrict__ conic_opacity, const float* __restrict__ rgb,
                              const float* __restrict__ zsil, float4* __restrict__ inst_conic,
                              float4* __restrict__ inst_geom, float4* __restrict__ inst_feat,
                              const uint32_t* __restrict__ status, const uint32_t* __restrict__ tile_order) {
    __shared__ uint64_t s_e[TS_CAP];
    if (status[1]) return;
    // crowded buckets first (their sort costs the most): the grid's last CTAs are the short ones
    const uint2 range = ranges[tile_order[blockIdx.x]];
    const int n = (int)(range.y - range.x);
    if (n <= 0) return;
    uint64_t* ge = entries + range.x;
    const bool in_smem = n <= TS_CAP;
    const bool merged = n <= TS_CAP / 2;  // result in the upper half of s_e
    if (merged) {
        // Typical bucket (a few hundred entries): runs of 64 are sorted by one warp each with warp-level barriers
        // only, then every entry finds its final position as its rank in its own run plus, by binary search, the
        // number of smaller keys in every other run (keys are unique).  Two block barriers instead of one per stage
        // of a block-wide network, and no padding to a power of two beyond the last run.
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        const int nruns = (n + 63) >> 6;
        for (int i = threadIdx.x; i < nruns * 64; i += TS_THREADS) s_e[i] = i < n ? ge[i] : ~0ull;
        __syncthreads();
        for (int r = warp; r < nruns; r += TS_THREADS / 32) {
            const int base = 64 * r;  // shared memory is addressed by index throughout (no generic pointers)
            for (int k = 2, lk = 1; k <= 64; k <<= 1, lk++) {
                {
                    const int blk = lane >> (lk - 1), off = lane & ((k >> 1) - 1);
                    const int i = base + (blk << lk) + off, j = base + (blk << lk) + k - 1 - off;
                    const uint64_t x = s_e[i], y = s_e[j];
                    if (x > y) { s_e[i] = y; s_e[j] = x; }
                }
                __syncwarp();
                for (int js = k >> 2; js > 0; js >>= 1) {
                    const int ljs = 31 - __clz(js);
                    const int i = base + ((lane >> ljs) << (ljs + 1)) + (lane & (js - 1)), j = i + js;
                    const uint64_t x = s_e[i], y = s_e[j];
                    if (x > y) { s_e[i] = y; s_e[j] = x; }
                    __syncwarp();
                }
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += TS_THREADS) {
            const uint64_t key = s_e[i];
            // final position = number of smaller keys = sum over the runs of a lower bound (in the key's own run that is
            // its index).  TS_MERGE runs at a time, branch-free: their chains of dependent shared loads overlap.
            int pos = 0;
            for (int r0 = 0; r0 < nruns; r0 += TS_MERGE) {
                int lo[TS_MERGE];
                const uint64_t* run[TS_MERGE];
#pragma unroll
                for (int u = 0; u < TS_MERGE; u++) {
                    run[u] = s_e + 64 * min(r0 + u, nruns - 1);
                    lo[u] = 0;
                }
#pragma unroll
                for (int step = 32; step > 0; step >>= 1)
#pragma unroll
                    for (int u = 0; u < TS_MERGE; u++)
                        if (run[u][lo[u] + step - 1] < key) lo[u] += step;
#pragma unroll
                for (int u = 0; u < TS_MERGE; u++) {
                    lo[u] += run[u][lo[u]] < key ? 1 : 0;   // 63 -> 64 when the whole run is below (padding keys never are)
                    pos += r0 + u < nruns ? lo[u] : 0;
                }
            }
            s_e[TS_CAP / 2 + pos] = key;
        }
        __syncthreads();
    } else if (in_smem) {
        for (int i = threadIdx.x; i < n; i += TS_THREADS) s_e[i] = ge[i];
        __syncthreads();
        bitonic_sort_ascending(s_e, n, TS_THREADS);
    } else {
        __syncthreads();
        bitonic_sort_ascending(ge, n, TS_THREADS);  // global memory is coherent within the block across the barriers
    }
    // TS_GATHER records per thread and round: the scattered reads of all are in flight before any is stored
    for (int i0 = threadIdx.x; i0 < n; i0 += TS_GATHER * TS_THREADS) {
        uint64_t e[TS_GATHER];
        float2 m[TS_GATHER];
        float4 co[TS_GATHER];
        float z[TS_GATHER], c[TS_GATHER][3];
        bool on[TS_GATHER];
#pragma unroll
        for (int u = 0; u < TS_GATHER; u++) {
            const int i = i0 + u * TS_THREADS;
            on[u] = i < n;
            if (on[u]) {
                e[u] = merged ? s_e[TS_CAP / 2 + i] : (in_smem ? s_e[i] : ge[i]);
                const uint32_t id = (uint32_t)e[u];
                m[u] = __ldg(means2D + id);
                co[u] = __ldg(conic_opacity + id);
                z[u] = __ldg(zsil + id);
                const float* cp = rgb + 3 * (size_t)id;
                c[u][0] = __ldg(cp); c[u][1] = __ldg(cp + 1); c[u][2] = __ldg(cp + 2);
            }
        }
#pragma unroll
        for (int u = 0; u < TS_GATHER; u++) {
            if (!on[u]) continue;
            const uint32_t id = (uint32_t)e[u];
            const size_t o = (size_t)range.x + i0 + u * TS_THREADS;
            point_list[o] = id;
            inst_conic[o] = co[u];
            inst_geom[o] = make_float4(m[u].x, m[u].y, __uint_as_float((uint32_t)(e[u] >> 32)), __uint_as_float(id));
            inst_feat[2 * o] = make_float4(c[u][0], c[u][1], c[u][2], z[u]);
            inst_feat[2 * o + 1] = make_float4(1.f, z[u] * z[u], pack_cull_extent(co[u]), 0.f);
        }
    }
}

// Block-wide sum of NV per-thread values -> out[NV] (plain stores: one row of partials per block; a
// few thousand blocks adding into the same 16 words with atomics serialise in L2).
template <int NV>
__device__ __forceinline__ void block_reduce_store(float (&v)[NV], float* s_red /*[NV * 8]*/, float* out) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; k++) {
        float x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) s_red[k * 8 + warp] = x;
    }
    __syncthreads();
    if (threadIdx.x < NV) {
        float x = 0.f;
#pragma unroll
        for (int w = 0; w < FP_THREADS / 32; w++) x += s_red[threadIdx.x * 8 + w];
        out[threadIdx.x] = x;
    }
}

// Flat gradient layout shared by the frame backward, the all-reduce and Adam: one buffer of 14 P floats holding the
// five gradient tensors one after the other, each in its parameter's own [P,w] layout:
//   means [P,3] at 0 | rgb [P,3] at 3P | rotation [P,4] at 6P | logit opacity [P,1] at 10P | log scales [P,1|3] at 11P.
// (A first version interleaved them as [P,14] rows: Adam then spent its time mapping flat slots back to tensors, and
// autograd had to copy every per-tensor gradient out of the rows.)  Segment k of Gaussian i, component c:
__device__ __forceinline__ float* grad_slot(float* base, size_t P, int seg_off, int width, size_t i, int c) {
    return base + (size_t)seg_off * P + (size_t)width * i + c;
}

// Backward of frame_preprocess_kernel.  grad_params: the flat segment layout above.
// pose_acc ([blocks,16] partial sums): dL/dt (3) | dL/dR (9, row-major) | dL/dq_cam through quat_mult (4).
#ifndef ASB_PPB_MINBLOCKS
#define ASB_PPB_MINBLOCKS 6
#endif
// CAM: the pose gradient is wanted (tracking, bundle adjustment); the mapping instantiation carries no pose accumulators
template <bool ISO, bool CAM>
__global__ void __launch_bounds__(FP_THREADS, CAM ? 4 : ASB_PPB_MINBLOCKS)
frame_preprocess_backward_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ unnorm_rot,
                                 const float* __restrict__ logit_op, const float* __restrict__ log_scales,
                                 const float* __restrict__ cam_rot, int rs, const float* __restrict__ cam_trans, int ts,
                                 const float* __restrict__ sil_w2c, const float* __restrict__ viewmatrix,
                                 const float* __restrict__ projmatrix, float h_x, float h_y, float tanfovx,
                                 float tanfovy, float scale_modifier, const int* __restrict__ radii,
                                 const float* __restrict__ zsil, const uint32_t* __restrict__ vis_list,
                                 const uint32_t* __restrict__ status, float* __restrict__ gacc,
                                 const float* __restrict__ gscale, int gaussians_grad, int camera_grad_,
                                 int accumulate, float* __restrict__ grad_params, float* __restrict__ grad_means2D,
                                 float* __restrict__ pose_acc) {
    __shared__ float view[16], proj[16];
    __shared__ Pose ps;
    __shared__ float s_red[16 * 8];
    const int nvis = (int)status[2];
    // blocks past the compact visible list have no per-Gaussian work: they only zero the rows of culled Gaussians
    // (A) and leave at once -- no matrix loads, no barriers, no pose partials (pose_finalize_kernel reads the first
    // ceil(V/256) rows only).  With ~20 % of the map visible that is most of the grid.
    const bool has_dense_work = blockIdx.x * FP_THREADS < nvis;
    if (has_dense_work) {
        if (threadIdx.x < 16) view[threadIdx.x] = __ldg(viewmatrix + threadIdx.x);
        else if (threadIdx.x < 32) proj[threadIdx.x - 16] = __ldg(projmatrix + threadIdx.x - 16);
        else if (threadIdx.x == 32) load_pose(ps, cam_rot, rs, cam_trans, ts, sil_w2c);
    }
    // (A) rows of culled Gaussians are zero
    {
        const int my = blockIdx.x * FP_THREADS + threadIdx.x;
        if (my < P && !(radii[my] > 0)) {
            // (the flat parameter gradient was cleared as a whole by the launcher: one streaming memset instead of
            // fourteen strided scalar stores per culled Gaussian)
            if (grad_means2D) {
                grad_means2D[3 * (size_t)my] = 0.f;
                grad_means2D[3 * (size_t)my + 1] = 0.f;
                grad_means2D[3 * (size_t)my + 2] = 0.f;
            }
        }
    }
    if (!has_dense_work) return;
    __syncthreads();
    const float gs = gscale ? __ldg(gscale) : 1.f;
    constexpr bool camera_grad = CAM;
    (void)camera_grad_;
    float pacc[CAM ? 16 : 1];
#pragma unroll
    for (int k = 0; k < (CAM ? 16 : 1); k++) pacc[k] = 0.f;
    // (B) the long per-Gaussian path runs densely over the compact list of visible Gaussians that the
    //     forward wrote (visible ones are usually a scattered minority): the first ceil(V/256) blocks are
    //     full, the rest are done after (A)
    for (int slot = blockIdx.x * FP_THREADS + threadIdx.x; slot < nvis; slot += FP_THREADS * gridDim.x) {
    const int idx = (int)vis_list[slot];
    float out[14];
#pragma unroll
    for (int k = 0; k < 14; k++) out[k] = 0.f;
    float gm2x = 0.f, gm2y = 0.f;
    {
        float4* ga = reinterpret_cast<float4*>(gacc) + 3 * (size_t)idx;
        const float4 g0 = ga[0], g1 = ga[1], g2 = ga[2];  // conic x,y,w | op ; mean2D x,y | c0 c1 ; c2 c3 c4 c5
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
        ga[0] = zero; ga[1] = zero; ga[2] = zero;  // leave the accumulators clean for the next frame
        const float px = __ldg(means3D + 3 * (size_t)idx), py = __ldg(means3D + 3 * (size_t)idx + 1),
                    pz = __ldg(means3D + 3 * (size_t)idx + 2);
        float ls[3];
        ls[0] = __ldg(log_scales + (ISO ? 1 : 3) * (size_t)idx);
        ls[1] = ISO ? ls[0] : __ldg(log_scales + 3 * (size_t)idx + 1);
        ls[2] = ISO ? ls[0] : __ldg(log_scales + 3 * (size_t)idx + 2);
        Activated a;
        activate<ISO>(ps, px, py, pz, __ldg(reinterpret_cast<const float4*>(unnorm_rot) + idx), ls,
                      __ldg(logit_op + idx), a);
        float c3[6];
        const float4 rot4 = make_float4(a.rot[0], a.rot[1], a.rot[2], a.rot[3]);
        cov3d_from_scale_rot(a.scale[0], a.scale[1], a.scale[2], scale_modifier, rot4, c3);
        float dmean[3], dcov[6], dscale[3], drot[4];
        preprocess_backward_one(a.mc[0], a.mc[1], a.mc[2], c3, true, rot4, a.scale, scale_modifier, view, proj, h_x, h_y,
                                tanfovx, tanfovy, g0.x, g0.y, g0.z, g1.x, g1.y, dmean, dcov, dscale, drot);
        gm2x = g1.x; gm2y = g1.y;
        // [z,1,z^2] features (T20): colour-channel gradients 3 and 5 reach the mean through z
        const float z = __ldg(zsil + idx);
        const float gz = g2.y + 2.f * z * g2.w;
        dmean[0] += ps.sil[0] * gz; dmean[1] += ps.sil[1] * gz; dmean[2] += ps.sil[2] * gz;
        // rotation chain: rot = normalize(rot_t)
        float dot = 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) dot += a.rot[k] * drot[k];
        float g_rt[4];
#pragma unroll
        for (int k = 0; k < 4; k++) g_rt[k] = (drot[k] - a.rot[k] * dot) / a.rot_t_norm;
        float g_qraw[4];
        if (ISO) {
#pragma unroll
            for (int k = 0; k < 4; k++) g_qraw[k] = g_rt[k];
        } else {
            const float* q1 = ps.q;
            const float* q2 = a.qn;
            float g_qn[4];
            g_qn[0] = q1[0] * g_rt[0] + q1[1] * g_rt[1] + q1[2] * g_rt[2] + q1[3] * g_rt[3];
            g_qn[1] = -q1[1] * g_rt[0] + q1[0] * g_rt[1] + q1[3] * g_rt[2] - q1[2] * g_rt[3];
            g_qn[2] = -q1[2] * g_rt[0] - q1[3] * g_rt[1] + q1[0] * g_rt[2] + q1[1] * g_rt[3];
            g_qn[3] = -q1[3] * g_rt[0] + q1[2] * g_rt[1] - q1[1] * g_rt[2] + q1[0] * g_rt[3];
            float d2 = 0.f;
#pragma unroll
            for (int k = 0; k < 4; k++) d2 += q2[k] * g_qn[k];
#pragma unroll
            for (int k = 0; k < 4; k++) g_qraw[k] = (g_qn[k] - q2[k] * d2) / a.qn_norm;
            if constexpr (CAM) {
                pacc[12] += q2[0] * g_rt[0] + q2[1] * g_rt[1] + q2[2] * g_rt[2] + q2[3] * g_rt[3];
                pacc[13] += -q2[1] * g_rt[0] + q2[0] * g_rt[1] - q2[3] * g_rt[2] + q2[2] * g_rt[3];
                pacc[14] += -q2[2] * g_rt[0] + q2[3] * g_rt[1] + q2[0] * g_rt[2] - q2[1] * g_rt[3];
                pacc[15] += -q2[3] * g_rt[0] - q2[2] * g_rt[1] + q2[1] * g_rt[2] + q2[0] * g_rt[3];
            }
        }
        if constexpr (CAM) {
            pacc[0] += dmean[0]; pacc[1] += dmean[1]; pacc[2] += dmean[2];
#pragma unroll
            for (int i = 0; i < 3; i++) {
                pacc[3 + 3 * i + 0] += dmean[i] * px;
                pacc[3 + 3 * i + 1] += dmean[i] * py;
                pacc[3 + 3 * i + 2] += dmean[i] * pz;
            }
        }
        if (gaussians_grad) {
            // means: world-frame gradient = R^T dL/dmeans_cam
            out[0] = ps.R[0] * dmean[0] + ps.R[3] * dmean[1] + ps.R[6] * dmean[2];
            out[1] = ps.R[1] * dmean[0] + ps.R[4] * dmean[1] + ps.R[7] * dmean[2];
            out[2] = ps.R[2] * dmean[0] + ps.R[5] * dmean[1] + ps.R[8] * dmean[2];
            out[3] = g1.z; out[4] = g1.w; out[5] = g2.x;  // rgb
#pragma unroll
            for (int k = 0; k < 4; k++) out[6 + k] = g_qraw[k];
            out[10] = g0.w * a.opacity * (1.f - a.opacity);  // sigmoid'
            if (ISO) {
                out[11] = (dscale[0] + dscale[1] + dscale[2]) * a.scale[0];
            } else {
                out[11] = dscale[0] * a.scale[0]; out[12] = dscale[1] * a.scale[1]; out[13] = dscale[2] * a.scale[2];
            }
        }
    }
    {
        if (gaussians_grad && grad_params) {
            const size_t Pz = (size_t)P, i = (size_t)idx;
            auto put = [&](float* dst, float val) {   // keyframe-batch accumulation adds to what is there
                *dst = accumulate ? (*dst + val * gs) : (val * gs);
            };
#pragma unroll
            for (int c = 0; c < 3; c++) {
                put(grad_slot(grad_params, Pz, 0, 3, i, c), out[c]);
                put(grad_slot(grad_params, Pz, 3, 3, i, c), out[3 + c]);
            }
#pragma unroll
            for (int c = 0; c < 4; c++) put(grad_slot(grad_params, Pz, 6, 4, i, c), out[6 + c]);
            put(grad_slot(grad_params, Pz, 10, 1, i, 0), out[10]);
#pragma unroll
            for (int c = 0; c < (ISO ? 1 : 3); c++) put(grad_slot(grad_params, Pz, 11, ISO ? 1 : 3, i, c), out[11 + c]);
        }
        if (grad_means2D) {
            grad_means2D[3 * (size_t)idx] = gm2x * gs;
            grad_means2D[3 * (size_t)idx + 1] = gm2y * gs;
            grad_means2D[3 * (size_t)idx + 2] = 0.f;
        }
    }
    }  // slot loop
    if constexpr (CAM) {
#pragma unroll
        for (int k = 0; k < 16; k++) pacc[k] *= gs;
        block_reduce_store<16>(pacc, s_red, pose_acc + 16 * (size_t)blockIdx.x);
    }
}

// per-block pose partials [nblocks,16] -> gradients of cam_unnorm_rot (4) and cam_trans (3).
__global__ void __launch_bounds__(256)
pose_finalize_kernel(const float* __restrict__ cam_rot, int rs, const float* __restrict__ partials, int nblocks,
                     const uint32_t* __restrict__ status, float* grad_cam, int gstride) {
    __shared__ float s_part[16][17];
    __shared__ float pose_acc[16];
    // only the first ceil(V/256) blocks of the backward had visible Gaussians; the others wrote zeros
    nblocks = min(nblocks, (int)((status[2] + FP_THREADS - 1) / FP_THREADS));
    {
        const int k = threadIdx.x & 15, g = threadIdx.x >> 4;  // 16 groups stride over the blocks
        float x = 0.f;
        for (int b = g; b < nblocks; b += 16) x += partials[16 * (size_t)b + k];
        s_part[g][k] = x;
    }
    __syncthreads();
    if (threadIdx.x < 16) {
        float x = 0.f;
#pragma unroll
        for (int g = 0; g < 16; g++) x += s_part[g][threadIdx.x];
        pose_acc[threadIdx.x] = x;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    float qr[4];
#pragma unroll
    for (int k = 0; k < 4; k++) qr[k] = cam_rot[(size_t)k * rs];
    const float n = fmaxf(sqrtf(qr[0] * qr[0] + qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]), 1e-12f);
    float qc[4];
#pragma unroll
    for (int k = 0; k < 4; k++) qc[k] = qr[k] / n;
    // build_rotation normalises qc once more: qh = qc / |qc|
    const float n2 = sqrtf(qc[0] * qc[0] + qc[1] * qc[1] + qc[2] * qc[2] + qc[3] * qc[3]);
    const float r = qc[0] / n2, x = qc[1] / n2, y = qc[2] / n2, z = qc[3] / n2;
    const float* G = pose_acc + 3;  // dL/dR row-major
    float gh[4];
    gh[0] = 2.f * (-z * G[1] + y * G[2] + z * G[3] - x * G[5] - y * G[6] + x * G[7]);
    gh[1] = 2.f * (y * G[1] + z * G[2] + y * G[3] - 2.f * x * G[4] - r * G[5] + z * G[6] + r * G[7] - 2.f * x * G[8]);
    gh[2] = 2.f * (-2.f * y * G[0] + x * G[1] + r * G[2] + x * G[3] + z * G[5] - r * G[6] + z * G[7] - 2.f * y * G[8]);
    gh[3] = 2.f * (-2.f * z * G[0] - r * G[1] + x * G[2] + r * G[3] - 2.f * z * G[4] + y * G[5] + x * G[6] + y * G[7]);
    const float qh[4] = {r, x, y, z};
    float d = 0.f;
#pragma unroll
    for (int k = 0; k < 4; k++) d += qh[k] * gh[k];
    float gq[4];
#pragma unroll
    for (int k = 0; k < 4; k++) gq[k] = (gh[k] - qh[k] * d) / n2 + pose_acc[12 + k];
    float d2 = 0.f;
#pragma unroll
    for (int k = 0; k < 4; k++) d2 += qc[k] * gq[k];
#pragma unroll
    for (int k = 0; k < 4; k++) grad_cam[(size_t)k * gstride] = (gq[k] - qc[k] * d2) / n;
#pragma unroll
    for (int k = 0; k < 3; k++) grad_cam[(size_t)(4 + k) * gstride] = pose_acc[k];
}

// ------------------------------------------------------------------------------------------------
// F4 tracking loss (splatam.py:213-247, tracking branch with use_l1):
//   mask = gt_depth > 0 & !nan(depth) & !nan(depth_sq - depth^2) [& sil > sil_thres]
//   loss = w_im * sum_mask |gt_rgb - im| + w_depth * sum_mask |gt_depth - depth|
// One sweep: per-block partial sums + dL/d(colour6).  A second tiny kernel adds the partials.
// ------------------------------------------------------------------------------------------------
constexpr int LOSS_THREADS = 256;

__device__ __forceinline__ float sgn(float x) { return (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f); }

__global__ void __launch_bounds__(LOSS_THREADS)
tracking_loss_kernel(int N, const float* __restrict__ c6, const float* __restrict__ gt_rgb,
                     const float* __restrict__ gt_depth, float w_im, float w_depth, float sil_thres, int use_sil,
                     int mask_color, float* __restrict__ dL, float* __restrict__ partials) {
    float s_im = 0.f, s_d = 0.f, s_cnt = 0.f;
    for (int i = blockIdx.x * LOSS_THREADS + threadIdx.x; i < N; i += gridDim.x * LOSS_THREADS) {
        const float depth = c6[3 * (size_t)N + i], sil = c6[4 * (size_t)N + i], dsq = c6[5 * (size_t)N + i];
        const float gd = gt_depth[i];
        const float inv = dsq - depth * depth;
        bool m = (gd > 0.f) && !isnan(depth) && !isnan(inv);
        if (use_sil) m = m && (sil > sil_thres);
        const bool mc = mask_color ? m : true;
        float g_d = 0.f;
        if (m) {
            const float e = gd - depth;
            s_d += fabsf(e);
            g_d = -w_depth * sgn(e);
            s_cnt += 1.f;
        }
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
            float g = 0.f;
            if (mc) {
                const float e = gt_rgb[ch * (size_t)N + i] - c6[ch * (size_t)N + i];
                s_im += fabsf(e);
                g = -w_im * sgn(e);
            }
            dL[ch * (size_t)N + i] = g;
        }
        dL[3 * (size_t)N + i] = g_d;
        dL[4 * (size_t)N + i] = 0.f;
        dL[5 * (size_t)N + i] = 0.f;
    }
    __shared__ float s_red[3 * 8];
    float v[3] = {s_im, s_d, s_cnt};
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) s_red[k * 8 + warp] = x;
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        float x = 0.f;
#pragma unroll
        for (int w = 0; w < LOSS_THREADS / 32; w++) x += s_red[threadIdx.x * 8 + w];
        partials[blockIdx.x * 4 + threadIdx.x] = x;
    }
}

// losses[0] = weighted total, [1] = weighted im, [2] = weighted depth, [3] = mask count
__global__ void tracking_loss_finalize_kernel(int nblocks, const float* __restrict__ partials, float w_im,
                                              float w_depth, float* __restrict__ losses,
                                              const uint32_t* __restrict__ status) {
    __shared__ float s[3][32];
    float v[3] = {0.f, 0.f, 0.f};
    for (int b = threadIdx.x; b < nblocks; b += 32)
#pragma unroll
        for (int k = 0; k < 3; k++) v[k] += partials[b * 4 + k];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        s[k][0] = x;
    }
    if (threadIdx.x == 0) {
        const float li = w_im * s[0][0], ld = w_depth * s[1][0];
        losses[0] = li + ld;
        losses[1] = li;
        losses[2] = ld;
        losses[3] = s[2][0];
        if (status) {  // frame status rides along so that one read-back tells the host everything
            losses[4] = (float)status[0];
            losses[5] = (float)status[1];
            if (status[1]) {  // binning buffer overflowed: the frame is invalid -- make that impossible to miss
                losses[0] = CUDART_NAN_F; losses[1] = CUDART_NAN_F; losses[2] = CUDART_NAN_F;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// F5 Adam (torch.optim.Adam semantics: no amsgrad, no weight decay), all five tensors in one pass.
// grad/exp_avg/exp_avg_sq are [P,14] flat (means 3 | rgb 3 | rot 4 | logit 1 | log_scales 3).
// ------------------------------------------------------------------------------------------------
struct AdamPtrs {   // plain members (no arrays): indexing a kernel-parameter array at run time would spill it to local memory
    float *p0, *p1, *p2, *p3, *p4;
    float s0, s1, s2, s3, s4;   // lr[t] / bias_correction1 (plain lr[t] when the kernel is given a step_state)
};

#ifndef ASB_AD_THREADS
#define ASB_AD_THREADS 128
#endif
constexpr int AD_THREADS = ASB_AD_THREADS;
#ifndef ASB_AD_VEC
#define ASB_AD_VEC 2
#endif
constexpr int AD_VEC_PER_THREAD = ASB_AD_VEC;   // 128-bit groups per thread: that many independent load quadruples in flight

// arithmetic per element as torch.optim.Adam (single-tensor path): lerp_, mul_/addcmul_, sqrt / bias_correction2_sqrt
// + eps, addcdiv_
__device__ __forceinline__ void adam_one(float g, float& m, float& v, float& p, float step_size, float beta1, float beta2,
                                         float eps, float bc2_sqrt) {
    m = m + (g - m) * (1.f - beta1);          // exp_avg.lerp_(grad, 1 - beta1)
    v = v * beta2 + (1.f - beta2) * g * g;    // mul_(beta2).addcmul_(g, g, 1 - beta2)
    // sqrt is correctly rounded; the two quotients use the 2-ulp fast division (parity bar on the updated
    // parameters is 1e-5 relative; the IEEE division sequences tripled the kernel's instruction count)
    const float denom = __fdividef(__fsqrt_rn(v), bc2_sqrt) + eps;
    p = p - step_size * __fdividef(m, denom);  // addcdiv_(exp_avg, denom, -step_size)
}

// which tensor a flat element belongs to, its learning-rate-derived step size and where its parameter lives
struct AdamSeg {
    float* p;
    float step;
};
__device__ __forceinline__ AdamSeg adam_segment(const AdamPtrs& ap, size_t P, size_t e, const float (&st)[5]) {
    AdamSeg r;
    if (e < 3 * P) { r.p = ap.p0 + e; r.step = st[0]; }
    else if (e < 6 * P) { r.p = ap.p1 + (e - 3 * P); r.step = st[1]; }
    else if (e < 10 * P) { r.p = ap.p2 + (e - 6 * P); r.step = st[2]; }
    else if (e < 11 * P) { r.p = ap.p3 + (e - 10 * P); r.step = st[3]; }
    else { r.p = ap.p4 + (e - 11 * P); r.step = st[4]; }
    return r;
}

// Gradient, exp_avg and exp_avg_sq share the flat segment layout (see grad_slot) and every segment has its
// parameter's own layout, so the update is one elementwise stream over `n` = (11 + ws) P floats: 128-bit loads of
// g, m, v and the parameter, 128-bit stores of m, v and the parameter, four groups per thread in flight.  A group that
// straddles two tensors or is not 16-byte aligned in its parameter tensor (P not a multiple of 4) takes the scalar path.
// step_state (optional, device memory): [0] = steps taken so far, [7] = block ticket.  When given, every block derives
// the bias corrections of step [0]+1 itself (in double, as torch does on the host) and the last block to finish
// advances [0]: the launch is identical every step, so it can sit in a CUDA graph.  ap.s* are then plain learning rates.
__global__ void __launch_bounds__(AD_THREADS)
adam_kernel(int P_, int iso, AdamPtrs ap, const float* __restrict__ grad, float* __restrict__ m, float* __restrict__ v,
            float beta1, float beta2, float eps, float bc2_sqrt, float* __restrict__ step_state, size_t e_begin,
            size_t e_end, int advance) {
    __shared__ float s_sched[2];
    const size_t P = (size_t)P_;
    const size_t n = min((size_t)(iso ? 12 : 14) * P, e_end);   // this launch updates the flat elements [e_begin, n)
    float st[5] = {ap.s0, ap.s1, ap.s2, ap.s3, ap.s4};
    if (step_state) {
        if (threadIdx.x == 0) {
            const double step = (double)*reinterpret_cast<volatile float*>(step_state) + 1.0;
            s_sched[0] = (float)(1.0 - pow((double)beta1, step));           // bias_correction1
            s_sched[1] = (float)sqrt(1.0 - pow((double)beta2, step));       // sqrt(bias_correction2)
        }
        __syncthreads();
        const float bc1 = s_sched[0];
#pragma unroll
        for (int k = 0; k < 5; k++) st[k] = st[k] / bc1;                    // lr / bias_correction1
        bc2_sqrt = s_sched[1];
    }
    const size_t ngroups = (n + 3) / 4;
    const size_t g0 = e_begin / 4 + ((size_t)blockIdx.x * AD_THREADS) * AD_VEC_PER_THREAD + threadIdx.x;
    float4 gg[AD_VEC_PER_THREAD], mm[AD_VEC_PER_THREAD], vv[AD_VEC_PER_THREAD], pp[AD_VEC_PER_THREAD];
    AdamSeg seg[AD_VEC_PER_THREAD];
    bool fast[AD_VEC_PER_THREAD];
#pragma unroll
    for (int u = 0; u < AD_VEC_PER_THREAD; u++) {
        const size_t q = g0 + (size_t)u * AD_THREADS, e = 4 * q;
        fast[u] = false;
        if (q < ngroups && e + 3 < n) {
            seg[u] = adam_segment(ap, P, e, st);
            const AdamSeg last = adam_segment(ap, P, e + 3, st);
            fast[u] = (last.p == seg[u].p + 3) && ((reinterpret_cast<uintptr_t>(seg[u].p) & 15) == 0);
        }
        if (fast[u]) {
            gg[u] = __ldg(reinterpret_cast<const float4*>(grad + e));
            mm[u] = *reinterpret_cast<const float4*>(m + e);
            vv[u] = *reinterpret_cast<const float4*>(v + e);
            pp[u] = *reinterpret_cast<const float4*>(seg[u].p);
        }
    }
    // A group whose gradient and both moments are all zero does not move: m and v stay 0 and the parameter changes by
    // step_size * 0 / (0 + eps) = 0 exactly -- the Gaussians no keyframe of this optimizer has seen yet (four out of
    // five in a mapping round over a large map).  Nothing is written back for them: the result is the dense update
    // bit for bit, at two thirds of its memory traffic.  (Not loading their parameters either was measured slower: the
    // parameter load then waits for the other three, and the kernel is short of loads in flight as it is.)
    bool idle[AD_VEC_PER_THREAD];
#pragma unroll
    for (int u = 0; u < AD_VEC_PER_THREAD; u++) {
        idle[u] = false;
        if (fast[u]) {
            const float4 a = gg[u], b = mm[u], c = vv[u];
            idle[u] = a.x == 0.f && a.y == 0.f && a.z == 0.f && a.w == 0.f && b.x == 0.f && b.y == 0.f && b.z == 0.f &&
                      b.w == 0.f && c.x == 0.f && c.y == 0.f && c.z == 0.f && c.w == 0.f;
        }
    }
#pragma unroll
    for (int u = 0; u < AD_VEC_PER_THREAD; u++) {
        const size_t q = g0 + (size_t)u * AD_THREADS, e = 4 * q;
        if (fast[u] && idle[u]) continue;
        if (fast[u]) {
            const float s1 = seg[u].step;
            adam_one(gg[u].x, mm[u].x, vv[u].x, pp[u].x, s1, beta1, beta2, eps, bc2_sqrt);
            adam_one(gg[u].y, mm[u].y, vv[u].y, pp[u].y, s1, beta1, beta2, eps, bc2_sqrt);
            adam_one(gg[u].z, mm[u].z, vv[u].z, pp[u].z, s1, beta1, beta2, eps, bc2_sqrt);
            adam_one(gg[u].w, mm[u].w, vv[u].w, pp[u].w, s1, beta1, beta2, eps, bc2_sqrt);
            *reinterpret_cast<float4*>(m + e) = mm[u];
            *reinterpret_cast<float4*>(v + e) = vv[u];
            *reinterpret_cast<float4*>(seg[u].p) = pp[u];
        } else if (q < ngroups) {
            for (size_t k = e; k < min(e + 4, n); k++) {
                const AdamSeg sg = adam_segment(ap, P, k, st);
                float a = m[k], b = v[k], c = *sg.p;
                adam_one(grad[k], a, b, c, sg.step, beta1, beta2, eps, bc2_sqrt);
                m[k] = a; v[k] = b; *sg.p = c;
            }
        }
    }
    if (step_state && advance && threadIdx.x == 0) {
        // every block read the step count before taking its ticket; the last ticket advances it for the next launch
        __threadfence();
        int* ticket = reinterpret_cast<int*>(step_state + 7);
        if (atomicAdd(ticket, 1) == (int)gridDim.x - 1) {
            step_state[0] = *reinterpret_cast<volatile float*>(step_state) + 1.f;
            *ticket = 0;
        }
    }
}

static int adam_blocks(int P, int iso, size_t e_begin = 0, size_t e_end = ~(size_t)0) {
    const size_t n = std::min((size_t)(iso ? 12 : 14) * P, e_end);
    const size_t groups = n > e_begin ? (n + 3) / 4 - e_begin / 4 : 0;
    return (int)((groups + (size_t)AD_THREADS * AD_VEC_PER_THREAD - 1) / ((size_t)AD_THREADS * AD_VEC_PER_THREAD));
}

// ---- Adam fused with the gradient exchange over peer memory (NVLink / NVSwitch) ------------------------------
// Mapping over a keyframe batch: every rank holds the flat [P,14] gradient of ITS keyframes.  Instead of
// all-reducing that buffer (NCCL) and running the same Adam step on every rank, rank r owns the Gaussians
// [g_begin, g_end): its threads load those rows of every rank's gradient buffer straight over NVLink (peer pointers),
// add them in rank order, apply Adam with the optimizer state kept only for the owned rows, and store the new
// parameter values into the parameter tensors of every rank.  One kernel: reduce-scatter + Adam + all-gather.
struct PeerAdamArgs {
    float* params[8][5];    // [rank][tensor] as mapped into this process
    const float* grads[8];  // [rank] flat gradient buffers (segment layout)
    float* own[5];          // this rank's parameter tensors (= params[rank])
    float step_size[5];     // lr[t] / bias_correction1
    int world;
};

// One thread per flat element (segment layout, see grad_slot): neighbouring lanes touch neighbouring words of the
// gradient slabs AND of the parameter tensors, so the remote loads and, more importantly, the remote stores leave the SM
// as full sectors (a 128-bit-per-thread variant scattered its parameter stores across lanes and was 1.7x slower over
// NVLink).  Sums run in rank order and every element has exactly one owner, so all ranks end up with identical
// parameters.  The rank owns the flat elements [e_begin, e_end); m, v hold only those.
__global__ void __launch_bounds__(256)
adam_peer_kernel(size_t P, size_t e_begin, size_t e_end, PeerAdamArgs pa, float* __restrict__ m, float* __restrict__ v,
                 float beta1, float beta2, float eps, float bc2_sqrt) {
    const size_t le = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // element of the owned slab
    if (e_begin + le >= e_end) return;
    const size_t e = e_begin + le;
    float g = 0.f;
#pragma unroll
    for (int r = 0; r < 8; r++)
        if (r < pa.world) g += pa.grads[r][e];
    int t;
    size_t po;
    if (e < 3 * P) { t = 0; po = e; }
    else if (e < 6 * P) { t = 1; po = e - 3 * P; }
    else if (e < 10 * P) { t = 2; po = e - 6 * P; }
    else if (e < 11 * P) { t = 3; po = e - 10 * P; }
    else { t = 4; po = e - 11 * P; }
    float* mine = t == 0 ? pa.own[0] : t == 1 ? pa.own[1] : t == 2 ? pa.own[2] : t == 3 ? pa.own[3] : pa.own[4];
    const float ss = t == 0 ? pa.step_size[0] : t == 1 ? pa.step_size[1] : t == 2 ? pa.step_size[2]
                   : t == 3 ? pa.step_size[3] : pa.step_size[4];
    float mm = m[le], vv = v[le], pv = mine[po];
    adam_one(g, mm, vv, pv, ss, beta1, beta2, eps, bc2_sqrt);
    m[le] = mm;
    v[le] = vv;
#pragma unroll
    for (int r = 0; r < 8; r++) {  // unrolled: indexing a kernel parameter at run time would copy it to local memory
        if (r < pa.world) {
            float* dst = t == 0 ? pa.params[r][0] : t == 1 ? pa.params[r][1] : t == 2 ? pa.params[r][2]
                       : t == 3 ? pa.params[r][3] : pa.params[r][4];
            dst[po] = pv;
        }
    }
}

}  // namespace

// ---- host launchers --------------------------------------------------------------------------------
void launch_frame_preprocess(const AsFrameArgs& a, const GeometryState& g, uint32_t* status, uint32_t* tile_count,
                             cudaStream_t st) {
    const int P = a.P, W = a.width, H = a.height;
    const bool ordered = tile_count == nullptr;
    if (ordered)  // ticket + one look-back word per 256-Gaussian slab (same footprint as the stand-alone scan's temp)
        cudaMemsetAsync(g.scan_temp, 0, 16 + sizeof(uint64_t) * (size_t)((P + FP_THREADS - 1) / FP_THREADS), st);
    else          // the blocks add their totals into status[0] / status[2]; [1] is this frame's overflow flag
        cudaMemsetAsync(status, 0, 4 * sizeof(uint32_t), st);
    const float focal_y = H / (2.0f * a.tan_fovy), focal_x = W / (2.0f * a.tan_fovx);
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    constexpr int NT_ORDERED = FP_THREADS;  // 1024-thread blocks (a 4x shorter look-back chain) measured no faster
    const int nt = ordered ? NT_ORDERED : FP_THREADS;
    const int blocks = (P + nt * FP_GPB - 1) / (nt * FP_GPB);
#define ASB_PP(ISO_, ORD_, NT_)                                                                                     \
    frame_preprocess_kernel<ISO_, ORD_, NT_, FP_GPB><<<blocks, NT_, 0, st>>>(                                       \
        P, a.means3D, a.unnorm_rotations, a.logit_opacities, a.log_scales, a.cam_unnorm_rot, a.cam_rot_stride,      \
        a.cam_trans, a.cam_trans_stride, a.sil_w2c, a.viewmatrix, a.projmatrix, W, H, gx, gy, a.tan_fovx,           \
        a.tan_fovy, focal_x, focal_y, a.scale_modifier, a.radii, g.radii, g.means2D, g.depths, g.conic_opacity,     \
        g.point_offsets, g.zsil, g.vis_list, g.scan_temp, reinterpret_cast<uint64_t*>(g.scan_temp + 4), status,     \
        a.max_2D_radius, a.seen, tile_count)
    if (a.isotropic) {
        if (ordered) ASB_PP(true, true, NT_ORDERED);
        else ASB_PP(true, false, FP_THREADS);
    } else {
        if (ordered) ASB_PP(false, true, NT_ORDERED);
        else ASB_PP(false, false, FP_THREADS);
    }
#undef ASB_PP
}

void launch_frame_tile_binning(int P, const GeometryState& g, int W, int H, uint32_t capacity, const float* rgb,
                               const BinningState& b, const ImageState& img, uint32_t* status, cudaStream_t st) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y, tiles = gx * gy;
    frame_tile_scan_kernel<<<1, 1024, 0, st>>>(img.tile_count, tiles, capacity, img.ranges, img.tile_cursor,
                                               img.tile_order, status);
    if (capacity == 0) return;
    // one thread per slot of the compact visible list; the count is only known on the device
    frame_scatter_kernel<<<(P + 255) / 256, 256, 0, st>>>(g.vis_list, g.means2D, g.depths, g.radii, gx, gy,
                                                                 img.tile_cursor, b.keys_unsorted, status);
    frame_tile_sort_gather_kernel<<<tiles, TS_THREADS, 0, st>>>(img.ranges, b.keys_unsorted, b.point_list, g.means2D,
                                                                g.conic_opacity, rgb, g.zsil, b.inst_conic, b.inst_geom,
                                                                b.inst_feat, status, img.tile_order);
}

void launch_frame_duplicate(int P, const GeometryState& g, int W, int H, uint32_t capacity, uint64_t* keys,
                            uint32_t* values, uint32_t* status, cudaStream_t st) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    frame_duplicate_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, g.means2D, g.depths, g.point_offsets, g.radii, gx, gy,
                                                            capacity, keys, values, status);
}

void launch_frame_ranges_gather(const uint32_t* n_dev, uint32_t capacity, int tiles, const GeometryState& g,
                                const float* rgb, const BinningState& b, uint2* ranges, cudaStream_t st) {
    cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)tiles, st);
    if (capacity == 0) return;
    frame_ranges_gather_kernel<<<(capacity + 255) / 256, 256, 0, st>>>(n_dev, capacity, b.keys_sorted, b.point_list,
                                                                      g.means2D, g.conic_opacity, rgb, g.zsil,
                                                                      b.inst_conic, b.inst_geom, b.inst_feat, ranges,
                                                                      (uint32_t)tiles);
}

void launch_frame_preprocess_backward(const AsFrameArgs& a, const GeometryState& g, const float* gscale,
                                      int gaussians_grad, int camera_grad, int accumulate, float* grad_params,
                                      float* grad_means2D, float* pose_acc, float* grad_cam, int grad_cam_stride,
                                      cudaStream_t st) {
    const int P = a.P, W = a.width, H = a.height;
    const float h_y = H / (2.0f * a.tan_fovy), h_x = W / (2.0f * a.tan_fovx);
    const int blocks = (P + FP_THREADS - 1) / FP_THREADS;
    // accumulate: 0 = overwrite (the buffer is cleared here: rows of culled Gaussians are zero, the kernel writes the
    // visible ones), 1 = add to what is there, 2 = overwrite a buffer the caller has already cleared
    if (gaussians_grad && grad_params && accumulate == 0)
        cudaMemsetAsync(grad_params, 0, sizeof(float) * 14 * (size_t)P, st);
    accumulate = accumulate == 1 ? 1 : 0;
#define ASB_PPB(ISO_, CAM_)                                                                                         \
    frame_preprocess_backward_kernel<ISO_, CAM_><<<blocks, FP_THREADS, 0, st>>>(                                     \
        P, a.means3D, a.unnorm_rotations, a.logit_opacities, a.log_scales, a.cam_unnorm_rot, a.cam_rot_stride,       \
        a.cam_trans, a.cam_trans_stride, a.sil_w2c, a.viewmatrix, a.projmatrix, h_x, h_y, a.tan_fovx, a.tan_fovy,    \
        a.scale_modifier, g.radii, g.zsil, g.vis_list, a.status, g.gacc, gscale, gaussians_grad, camera_grad,        \
        accumulate, grad_params, grad_means2D, pose_acc)
    if (a.isotropic) {
        if (camera_grad) ASB_PPB(true, true);
        else ASB_PPB(true, false);
    } else {
        if (camera_grad) ASB_PPB(false, true);
        else ASB_PPB(false, false);
    }
#undef ASB_PPB
    if (camera_grad)
        pose_finalize_kernel<<<1, 256, 0, st>>>(a.cam_unnorm_rot, a.cam_rot_stride, pose_acc, blocks, a.status, grad_cam,
                                                grad_cam_stride > 0 ? grad_cam_stride : 1);
}

void launch_adam_peer(int P, int e_begin, int e_end, int iso, int world, int rank, float* const* params /*[world*5]*/,
                      const float* const* grads /*[world]*/, const float lr[5], float* m, float* v, float beta1,
                      float beta2, float eps, int step, cudaStream_t st) {
    (void)iso;
    if (e_end <= e_begin) return;
    PeerAdamArgs pa;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    for (int r = 0; r < 8; r++) {
        for (int t = 0; t < 5; t++) pa.params[r][t] = r < world ? params[5 * r + t] : nullptr;
        pa.grads[r] = r < world ? grads[r] : nullptr;
    }
    for (int t = 0; t < 5; t++) {
        pa.own[t] = params[5 * rank + t];
        pa.step_size[t] = (float)(lr[t] / bc1);
    }
    pa.world = world;
    const size_t threads = (size_t)(e_end - e_begin);
    adam_peer_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>((size_t)P, (size_t)e_begin, (size_t)e_end, pa, m, v,
                                                                      beta1, beta2, eps, bc2_sqrt);
}

int tile_sort_capacity() { return TS_CAP; }

size_t pose_scratch_floats(int P) { return 16 * (size_t)((P + FP_THREADS - 1) / FP_THREADS) + 16; }

int loss_partial_blocks(int N) {
    int b = (N + LOSS_THREADS - 1) / LOSS_THREADS;
    return b > NUM_SMS * 4 ? NUM_SMS * 4 : b;
}

void launch_tracking_loss(int N, const float* c6, const float* gt_rgb, const float* gt_depth, float w_im,
                          float w_depth, float sil_thres, int use_sil, int mask_color, float* dL, float* partials,
                          float* losses, const uint32_t* status, cudaStream_t st) {
    const int blocks = loss_partial_blocks(N);
    tracking_loss_kernel<<<blocks, LOSS_THREADS, 0, st>>>(N, c6, gt_rgb, gt_depth, w_im, w_depth, sil_thres, use_sil,
                                                          mask_color, dL, partials);
    tracking_loss_finalize_kernel<<<1, 32, 0, st>>>(blocks, partials, w_im, w_depth, losses, status);
}

// Bias corrections in double, like torch.optim.Adam computes them on the host (1 - beta ** step in Python floats).
void launch_adam(int P, int iso, float* const params[5], const float lr[5], const float* grad, float* m, float* v,
                 float beta1, float beta2, float eps, int step, cudaStream_t st) {
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    AdamPtrs ap;
    ap.p0 = params[0]; ap.p1 = params[1]; ap.p2 = params[2]; ap.p3 = params[3]; ap.p4 = params[4];
    ap.s0 = (float)(lr[0] / bc1); ap.s1 = (float)(lr[1] / bc1); ap.s2 = (float)(lr[2] / bc1);
    ap.s3 = (float)(lr[3] / bc1); ap.s4 = (float)(lr[4] / bc1);
    adam_kernel<<<adam_blocks(P, iso), AD_THREADS, 0, st>>>(P, iso, ap, grad, m, v, beta1, beta2, eps, bc2_sqrt, nullptr, 0,
                                                            ~(size_t)0, 0);
}

// One thread: step += 1 (kept as a float-encoded integer in state[0]; exact up to 2^24 steps), then the step sizes and
// sqrt(bias_correction2) of that step in double.
__global__ void adam_advance_kernel(float* state, float lr0, float lr1, float lr2, float lr3, float lr4, float beta1,
                                    float beta2) {
    const double step = (double)state[0] + 1.0;
    const double bc1 = 1.0 - pow((double)beta1, step);
    state[0] = (float)step;
    state[1] = (float)(lr0 / bc1); state[2] = (float)(lr1 / bc1); state[3] = (float)(lr2 / bc1);
    state[4] = (float)(lr3 / bc1); state[5] = (float)(lr4 / bc1);
    state[6] = (float)sqrt(1.0 - pow((double)beta2, step));
}

// Adam over the two camera-pose tensors (tracking: lr 1e-3 on cam_unnorm_rots, 4e-3 on cam_trans, default eps 1e-8,
// config/splatam/online_habitat_sim.py:37-45).  A few hundred floats at most: one small grid, the step count on the device.
__global__ void pose_adam_kernel(float* __restrict__ rots, const float* __restrict__ g_rots, int n_rots,
                                 float* __restrict__ trans, const float* __restrict__ g_trans, int n_trans,
                                 float* __restrict__ m, float* __restrict__ v, float beta1, float beta2, float eps,
                                 const float* __restrict__ sched) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rots + n_trans) return;
    const bool is_rot = i < n_rots;
    float* p = is_rot ? rots + i : trans + (i - n_rots);
    const float g = is_rot ? g_rots[i] : g_trans[i - n_rots];
    const float step_size = sched[is_rot ? 1 : 2], bc2_sqrt = sched[6];
    const float nm = m[i] + (g - m[i]) * (1.f - beta1);
    const float nv = v[i] * beta2 + (1.f - beta2) * g * g;
    m[i] = nm;
    v[i] = nv;
    *p = *p - step_size * (nm / (sqrtf(nv) / bc2_sqrt + eps));
}

void launch_pose_adam(float* rots, const float* g_rots, int n_rots, float lr_rots, float* trans, const float* g_trans,
                      int n_trans, float lr_trans, float* m, float* v, float beta1, float beta2, float eps,
                      float* step_state, cudaStream_t st) {
    adam_advance_kernel<<<1, 1, 0, st>>>(step_state, lr_rots, lr_trans, 0.f, 0.f, 0.f, beta1, beta2);
    const int n = n_rots + n_trans;
    if (n > 0)
        pose_adam_kernel<<<(n + 127) / 128, 128, 0, st>>>(rots, g_rots, n_rots, trans, g_trans, n_trans, m, v, beta1,
                                                          beta2, eps, step_state);
}

void launch_adam_counted(int P, int iso, float* const params[5], const float lr[5], const float* grad, float* m,
                         float* v, float beta1, float beta2, float eps, float* step_state, size_t e_begin, size_t e_end,
                         int advance, cudaStream_t st) {
    if (P <= 0 || adam_blocks(P, iso, e_begin, e_end) == 0) {   // nothing to update: the step still counts
        if (!advance) return;
        adam_advance_kernel<<<1, 1, 0, st>>>(step_state, lr[0], lr[1], lr[2], lr[3], lr[4], beta1, beta2);
        return;
    }
    AdamPtrs ap;
    ap.p0 = params[0]; ap.p1 = params[1]; ap.p2 = params[2]; ap.p3 = params[3]; ap.p4 = params[4];
    ap.s0 = lr[0]; ap.s1 = lr[1]; ap.s2 = lr[2]; ap.s3 = lr[3]; ap.s4 = lr[4];
    adam_kernel<<<adam_blocks(P, iso, e_begin, e_end), AD_THREADS, 0, st>>>(P, iso, ap, grad, m, v, beta1, beta2, eps, 1.f,
                                                                            step_state, e_begin, e_end, advance);
}

}  // namespace asb
