// preprocess.cu -- per-Gaussian stages of the rasterizer (sm_100a).
//
//   preprocess_forward_kernel   replaces preprocessCUDA<3> fwd   dgr/cuda_rasterizer/forward.cu:155-256
//   mark_visible_kernel         replaces checkFrustum            rasterizer_impl.cu:54-66
//   duplicate_with_keys_kernel  replaces duplicateWithKeys       rasterizer_impl.cu:70-111
//   ranges_gather_kernel        replaces identifyTileRanges      rasterizer_impl.cu:116-138 (+ gathers the
//                               per-instance records the blend kernels stream with bulk copies)
//   preprocess_backward_kernel  replaces computeCov2DCUDA + preprocessCUDA<3> bwd  backward.cu:144-274, 346-396
//
// The forward arithmetic is spelled with rounding-explicit intrinsics in the operation order that
// nvcc gives the reference's expressions, so radii / tile rects / depth key bits come out identical.
#include "common.cuh"

namespace asb {

namespace {

constexpr int PP_THREADS = 256;

__device__ __forceinline__ float xform_row(const float* m, int r, float x, float y, float z) {
    // auxiliary.h:58-66 : m[r]*x + m[4+r]*y + m[8+r]*z + m[12+r]
    return __fadd_rn(dot3c(m[r], x, m[4 + r], y, m[8 + r], z), m[12 + r]);
}

// auxiliary.h:41-44 ndc2Pix: double literals => fp64 evaluation
__device__ __forceinline__ float ndc2pix(float v, int S) {
    return (float)__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5);
}

// auxiliary.h:46-56 getRect
__device__ __forceinline__ void get_rect(float px, float py, int max_radius, int gx, int gy, int& x0, int& y0,
                                         int& x1, int& y1) {
    const float r = (float)max_radius;
    x0 = min(gx, max(0, __float2int_rz(__fmul_rn(__fsub_rn(px, r), 1.0f / TILE_X))));
    y0 = min(gy, max(0, __float2int_rz(__fmul_rn(__fsub_rn(py, r), 1.0f / TILE_Y))));
    x1 = min(gx, max(0, __float2int_rz(__fmul_rn(__fsub_rn(__fadd_rn(__fadd_rn(px, r), (float)TILE_X), 1.0f), 1.0f / TILE_X))));
    y1 = min(gy, max(0, __float2int_rz(__fmul_rn(__fsub_rn(__fadd_rn(__fadd_rn(py, r), (float)TILE_Y), 1.0f), 1.0f / TILE_Y))));
}

// forward.cu:118-152 computeCov3D.  q is used as given (not normalised, forward.cu:127).
__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float4 q, float* cov) {
    const float sx = __fmul_rn(mod, s0), sy = __fmul_rn(mod, s1), sz = __fmul_rn(mod, s2);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    float R[3][3];  // textbook rotation rows (GLM's R is its transpose)
    R[0][0] = __fmaf_rn(-2.f, __fmaf_rn(y, y, __fmul_rn(z, z)), 1.f);
    R[0][1] = __fmul_rn(2.f, __fmaf_rn(x, y, -__fmul_rn(r, z)));
    R[0][2] = __fmul_rn(2.f, __fmaf_rn(x, z, __fmul_rn(r, y)));
    R[1][0] = __fmul_rn(2.f, __fmaf_rn(x, y, __fmul_rn(r, z)));
    R[1][1] = __fmaf_rn(-2.f, __fmaf_rn(x, x, __fmul_rn(z, z)), 1.f);
    R[1][2] = __fmul_rn(2.f, __fmaf_rn(y, z, -__fmul_rn(r, x)));
    R[2][0] = __fmul_rn(2.f, __fmaf_rn(x, z, -__fmul_rn(r, y)));
    R[2][1] = __fmul_rn(2.f, __fmaf_rn(y, z, __fmul_rn(r, x)));
    R[2][2] = __fmaf_rn(-2.f, __fmaf_rn(x, x, __fmul_rn(y, y)), 1.f);
    float M[3][3];  // M_glm[j][k] = s_k * R[j][k]
#pragma unroll
    for (int j = 0; j < 3; j++) {
        M[j][0] = __fmul_rn(sx, R[j][0]);
        M[j][1] = __fmul_rn(sy, R[j][1]);
        M[j][2] = __fmul_rn(sz, R[j][2]);
    }
    cov[0] = dot3c(M[0][0], M[0][0], M[0][1], M[0][1], M[0][2], M[0][2]);
    cov[1] = dot3c(M[1][0], M[0][0], M[1][1], M[0][1], M[1][2], M[0][2]);
    cov[2] = dot3c(M[2][0], M[0][0], M[2][1], M[0][1], M[2][2], M[0][2]);
    cov[3] = dot3c(M[1][0], M[1][0], M[1][1], M[1][1], M[1][2], M[1][2]);
    cov[4] = dot3c(M[2][0], M[1][0], M[2][1], M[1][1], M[2][2], M[1][2]);
    cov[5] = dot3c(M[2][0], M[2][0], M[2][1], M[2][1], M[2][2], M[2][2]);
}

// forward.cu:74-98 / backward.cu:160-193: clamped camera-space point t and the two non-zero columns of
// GLM's T = W * J, as Tm[0][k] = T_glm[0][k], Tm[1][k] = T_glm[1][k].
__device__ __forceinline__ void cov2d_T(float mx, float my, float mz, float fx, float fy, float tanfovx,
                                        float tanfovy, const float* view, float t[3], float Tm[2][3],
                                        float& txtz, float& tytz) {
    t[0] = xform_row(view, 0, mx, my, mz);
    t[1] = xform_row(view, 1, mx, my, mz);
    t[2] = xform_row(view, 2, mx, my, mz);
    const float limx = __fmul_rn(1.3f, tanfovx), limy = __fmul_rn(1.3f, tanfovy);
    txtz = __fdiv_rn(t[0], t[2]);
    tytz = __fdiv_rn(t[1], t[2]);
    t[0] = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), t[2]);
    t[1] = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), t[2]);
    const float tz2 = __fmul_rn(t[2], t[2]);
    const float J00 = __fdiv_rn(fx, t[2]), J02 = __fdiv_rn(-__fmul_rn(fx, t[0]), tz2);
    const float J11 = __fdiv_rn(fy, t[2]), J12 = __fdiv_rn(-__fmul_rn(fy, t[1]), tz2);
#pragma unroll
    for (int i = 0; i < 3; i++) {
        Tm[0][i] = __fmaf_rn(view[4 * i + 2], J02, __fmul_rn(view[4 * i + 0], J00));
        Tm[1][i] = __fmaf_rn(view[4 * i + 2], J12, __fmul_rn(view[4 * i + 1], J11));
    }
}

// cov = transpose(T) * transpose(Vrk) * T, entries [0][0], [0][1], [1][1] (forward.cu:100-112)
__device__ __forceinline__ void cov2d_from_T(const float Tm[2][3], const float* c, float& a, float& b, float& cc) {
    const float V[3][3] = {{c[0], c[1], c[2]}, {c[1], c[3], c[4]}, {c[2], c[4], c[5]}};
    float A[3][2];
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 2; i++) A[j][i] = dot3c(Tm[i][0], V[0][j], Tm[i][1], V[1][j], Tm[i][2], V[2][j]);
    a = dot3c(A[0][0], Tm[0][0], A[1][0], Tm[0][1], A[2][0], Tm[0][2]);
    b = dot3c(A[0][1], Tm[0][0], A[1][1], Tm[0][1], A[2][1], Tm[0][2]);
    cc = dot3c(A[0][1], Tm[1][0], A[1][1], Tm[1][1], A[2][1], Tm[1][2]);
}

// Cooperative, vectorised load of a [n,3] fp32 AoS slab (256 rows per block) into shared memory.
// base is the block's first row; falls back to scalar loads when the slab is not 16-byte aligned.
__device__ __forceinline__ void load_rows3(const float* __restrict__ src, int row0, int nrows, float* sm) {
    const float* p = src + (size_t)row0 * 3;
    const int nfl = nrows * 3;
    if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const int nv = nfl >> 2;
        const float4* p4 = reinterpret_cast<const float4*>(p);
        for (int i = threadIdx.x; i < nv; i += PP_THREADS) reinterpret_cast<float4*>(sm)[i] = __ldg(p4 + i);
        for (int i = (nv << 2) + threadIdx.x; i < nfl; i += PP_THREADS) sm[i] = __ldg(p + i);
    } else {
        for (int i = threadIdx.x; i < nfl; i += PP_THREADS) sm[i] = __ldg(p + i);
    }
}

__global__ void __launch_bounds__(PP_THREADS)
preprocess_forward_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ scales,
                          const float* __restrict__ rotations, const float* __restrict__ opacities,
                          const float* __restrict__ cov3D_precomp, const float* __restrict__ viewmatrix,
                          const float* __restrict__ projmatrix, int W, int H, int gx, int gy, float tanfovx,
                          float tanfovy, float focal_x, float focal_y, float scale_modifier,
                          int* __restrict__ radii, int* __restrict__ radii_internal, float2* __restrict__ means2D_out,
                          float* __restrict__ depths, float4* __restrict__ conic_opacity,
                          uint32_t* __restrict__ tiles_touched) {
    __shared__ __align__(16) float s_means[PP_THREADS * 3];
    __shared__ __align__(16) float s_scales[PP_THREADS * 3];
    __shared__ float s_view[16], s_proj[16];
    const int row0 = blockIdx.x * PP_THREADS;
    const int nrows = min(PP_THREADS, P - row0);
    load_rows3(means3D, row0, nrows, s_means);
    if (scales) load_rows3(scales, row0, nrows, s_scales);
    if (threadIdx.x < 16) s_view[threadIdx.x] = __ldg(viewmatrix + threadIdx.x);
    else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = __ldg(projmatrix + threadIdx.x - 16);
    __syncthreads();
    const int idx = row0 + threadIdx.x;
    if (idx >= P) return;

    int my_radii = 0;
    uint32_t my_tiles = 0;
    const float px = s_means[3 * threadIdx.x], py = s_means[3 * threadIdx.x + 1], pz = s_means[3 * threadIdx.x + 2];
    // in_frustum (auxiliary.h:139-164): only the near test is live
    const float pvz = xform_row(s_view, 2, px, py, pz);
    if (pvz > 0.2f) {
        const float hx = xform_row(s_proj, 0, px, py, pz);
        const float hy = xform_row(s_proj, 1, px, py, pz);
        const float hw = xform_row(s_proj, 3, px, py, pz);
        const float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
        const float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
        float cov3D[6];
        if (cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 6; k++) cov3D[k] = __ldg(cov3D_precomp + 6 * (size_t)idx + k);
        } else {
            const float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
            cov3d_from_scale_rot(s_scales[3 * threadIdx.x], s_scales[3 * threadIdx.x + 1],
                                 s_scales[3 * threadIdx.x + 2], scale_modifier, q, cov3D);
        }
        float t[3], Tm[2][3], txtz, tytz, a, b, c;
        cov2d_T(px, py, pz, focal_x, focal_y, tanfovx, tanfovy, s_view, t, Tm, txtz, tytz);
        cov2d_from_T(Tm, cov3D, a, b, c);
        a = __fadd_rn(a, 0.3f);
        c = __fadd_rn(c, 0.3f);
        const float det = __fmaf_rn(a, c, -__fmul_rn(b, b));
        if (det != 0.0f) {
            const float det_inv = __frcp_rn(det);
            const float mid = __fmul_rn(0.5f, __fadd_rn(a, c));
            const float disc = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
            const float lambda1 = __fadd_rn(mid, disc), lambda2 = __fsub_rn(mid, disc);
            const float my_radius = ceilf(__fmul_rn(3.f, __fsqrt_rn(fmaxf(lambda1, lambda2))));
            const float pix_x = ndc2pix(projx, W), pix_y = ndc2pix(projy, H);
            int x0, y0, x1, y1;
            const int ir = __float2int_rz(my_radius);
            get_rect(pix_x, pix_y, ir, gx, gy, x0, y0, x1, y1);
            const uint32_t area = (uint32_t)(x1 - x0) * (uint32_t)(y1 - y0);
            if (area != 0) {
                my_radii = ir;
                my_tiles = area;
                depths[idx] = pvz;
                means2D_out[idx] = make_float2(pix_x, pix_y);
                conic_opacity[idx] = make_float4(__fmul_rn(c, det_inv), __fmul_rn(-b, det_inv),
                                                 __fmul_rn(a, det_inv), __ldg(opacities + idx));
            }
        }
    }
    if (radii) radii[idx] = my_radii;
    radii_internal[idx] = my_radii;
    tiles_touched[idx] = my_tiles;
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ viewmatrix,
                                    uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float x = __ldg(means3D + 3 * (size_t)idx), y = __ldg(means3D + 3 * (size_t)idx + 1),
                z = __ldg(means3D + 3 * (size_t)idx + 2);
    present[idx] = xform_row(viewmatrix, 2, x, y, z) > 0.2f ? 1 : 0;
}

__global__ void __launch_bounds__(256)
duplicate_with_keys_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                           const uint32_t* __restrict__ offsets, const int* __restrict__ radii, int gx, int gy,
                           uint64_t* __restrict__ keys, uint32_t* __restrict__ values) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const int r = radii[idx];
    if (r <= 0) return;
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

// identifyTileRanges + gather of the per-instance records in sorted order.
template <int C>
__global__ void __launch_bounds__(256)
ranges_gather_kernel(int R, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ point_list,
                     const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                     const float* __restrict__ features, float4* __restrict__ inst_conic,
                     float4* __restrict__ inst_geom, float4* __restrict__ inst_feat, uint2* __restrict__ ranges) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint64_t key = keys[i];
    const uint32_t cur = (uint32_t)(key >> 32);
    if (i == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
        if (cur != prev) {
            ranges[prev].y = (uint32_t)i;
            ranges[cur].x = (uint32_t)i;
        }
    }
    if (i == R - 1) ranges[cur].y = (uint32_t)R;

    const uint32_t id = point_list[i];
    const float2 m = __ldg(means2D + id);
    inst_conic[i] = __ldg(conic_opacity + id);
    // depth is the low word of the key (rasterizer_impl.cu:101-103)
    inst_geom[i] = make_float4(m.x, m.y, __uint_as_float((uint32_t)key), __uint_as_float(id));
    const float* f = features + (size_t)id * C;
    if (C == 3) {
        inst_feat[i] = make_float4(__ldg(f), __ldg(f + 1), __ldg(f + 2), 0.f);
    } else {
        inst_feat[2 * (size_t)i] = make_float4(__ldg(f), __ldg(f + 1), __ldg(f + 2), __ldg(f + 3));
        inst_feat[2 * (size_t)i + 1] = make_float4(__ldg(f + 4), __ldg(f + 5), 0.f, 0.f);
    }
}

// Fused computeCov2DCUDA (backward.cu:144-274) + preprocessCUDA bwd (backward.cu:346-396, :278-341).
// Every Gaussian writes its outputs (zeros when it was culled), so callers need not pre-zero them.
template <int C>
__global__ void __launch_bounds__(256)
preprocess_backward_kernel(int P, const float* __restrict__ means3D, const int* __restrict__ radii,
                           const float* __restrict__ scales, const float* __restrict__ rotations,
                           const float* __restrict__ cov3D_precomp, float scale_modifier,
                           const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix, float h_x,
                           float h_y, float tanfovx, float tanfovy, const float* __restrict__ gacc,
                           PreprocessBackwardOut o) {
    __shared__ float view[16], proj[16];
    if (threadIdx.x < 16) view[threadIdx.x] = __ldg(viewmatrix + threadIdx.x);
    else if (threadIdx.x < 32) proj[threadIdx.x - 16] = __ldg(projmatrix + threadIdx.x - 16);
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const bool vis = radii[idx] > 0;
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gacc) + 3 * (size_t)idx);      // conic x,y,w | opacity
    const float4 g1 = __ldg(reinterpret_cast<const float4*>(gacc) + 3 * (size_t)idx + 1);  // mean2D x,y | c0 c1
    const float4 g2 = __ldg(reinterpret_cast<const float4*>(gacc) + 3 * (size_t)idx + 2);  // c2 c3 c4 c5
    if (o.dL_dopacity) o.dL_dopacity[idx] = g0.w;
    if (o.dL_dmean2D) {
        o.dL_dmean2D[3 * (size_t)idx] = g1.x;
        o.dL_dmean2D[3 * (size_t)idx + 1] = g1.y;
        o.dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
    }
    if (o.dL_dconic) reinterpret_cast<float4*>(o.dL_dconic)[idx] = make_float4(g0.x, g0.y, 0.f, g0.z);
    if (o.dL_dcolor) {
        float* dc = o.dL_dcolor + (size_t)idx * C;
        dc[0] = g1.z; dc[1] = g1.w; dc[2] = g2.x;
        if (C == 6) { dc[3] = g2.y; dc[4] = g2.z; dc[5] = g2.w; }
    }
    float dmean[3] = {0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
    if (vis) {
        const float mx = __ldg(means3D + 3 * (size_t)idx), my = __ldg(means3D + 3 * (size_t)idx + 1),
                    mz = __ldg(means3D + 3 * (size_t)idx + 2);
        float c3[6];
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        float s[3] = {0.f, 0.f, 0.f};
        if (cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 6; k++) c3[k] = __ldg(cov3D_precomp + 6 * (size_t)idx + k);
        } else {
            q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
#pragma unroll
            for (int k = 0; k < 3; k++) s[k] = __ldg(scales + 3 * (size_t)idx + k);
            cov3d_from_scale_rot(s[0], s[1], s[2], scale_modifier, q, c3);  // same bits the forward used
        }
        // ---- computeCov2DCUDA ----
        float t[3], T[2][3], txtz, tytz, a, b, c;
        cov2d_T(mx, my, mz, h_x, h_y, tanfovx, tanfovy, view, t, T, txtz, tytz);
        const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        cov2d_from_T(T, c3, a, b, c);
        a += 0.3f;
        c += 0.3f;
        const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
        const float dcx = g0.x, dcy = g0.y, dcz = g0.z;
        const float denom = a * c - b * b;
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);  // backward.cu:203
        if (denom2inv != 0.f) {
            dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dcov[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
            dcov[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
            dcov[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
            dcov[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
            dcov[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
            dcov[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
        }
        float dL_dT[2][3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float tv0 = T[0][0] * V[k][0] + T[0][1] * V[k][1] + T[0][2] * V[k][2];
            const float tv1 = T[1][0] * V[k][0] + T[1][1] * V[k][1] + T[1][2] * V[k][2];
            dL_dT[0][k] = 2 * tv0 * dL_da + tv1 * dL_db;
            dL_dT[1][k] = 2 * tv1 * dL_dc + tv0 * dL_db;
        }
        // W_glm[k][i] = view[4*i + k]
        const float dL_dJ00 = view[0] * dL_dT[0][0] + view[4] * dL_dT[0][1] + view[8] * dL_dT[0][2];
        const float dL_dJ02 = view[2] * dL_dT[0][0] + view[6] * dL_dT[0][1] + view[10] * dL_dT[0][2];
        const float dL_dJ11 = view[1] * dL_dT[1][0] + view[5] * dL_dT[1][1] + view[9] * dL_dT[1][2];
        const float dL_dJ12 = view[2] * dL_dT[1][0] + view[6] * dL_dT[1][1] + view[10] * dL_dT[1][2];
        const float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t[0]) * tz3 * dL_dJ02 +
                             (2 * h_y * t[1]) * tz3 * dL_dJ12;
        // transformVec4x3Transpose (auxiliary.h:91-99)
        dmean[0] = view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz;
        dmean[1] = view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz;
        dmean[2] = view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz;

        // ---- preprocessCUDA backward: projection chain of dL_dmean2D (backward.cu:373-387) ----
        const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
        const float m_w = 1.0f / (m_hw + 0.0000001f);
        const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
        const float gx2 = g1.x, gy2 = g1.y;
        dmean[0] += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
        dmean[1] += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
        dmean[2] += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;

        // ---- computeCov3D backward (backward.cu:278-341); no normalisation Jacobian (:340) ----
        if (!cov3D_precomp) {
            const float r = q.x, x = q.y, y = q.z, z = q.w;
            const float Rt[3][3] = {
                {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
            const float sm[3] = {scale_modifier * s[0], scale_modifier * s[1], scale_modifier * s[2]};
            const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                                    {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                                    {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
            float D[3][3];  // D[k][j] = dL/dRt[j][k] (the reference's dL_dMt[k][j] after scaling by s_k)
#pragma unroll
            for (int k = 0; k < 3; k++) {
                float acc = 0.f;
#pragma unroll
                for (int j = 0; j < 3; j++) {
                    const float G = 2.0f * sm[k] * (Rt[0][k] * dS[j][0] + Rt[1][k] * dS[j][1] + Rt[2][k] * dS[j][2]);
                    acc += Rt[j][k] * G;
                    D[k][j] = sm[k] * G;
                }
                dscale[k] = acc;
            }
            drot[0] = 2 * z * (D[0][1] - D[1][0]) + 2 * y * (D[2][0] - D[0][2]) + 2 * x * (D[1][2] - D[2][1]);
            drot[1] = 2 * y * (D[1][0] + D[0][1]) + 2 * z * (D[2][0] + D[0][2]) + 2 * r * (D[1][2] - D[2][1]) - 4 * x * (D[2][2] + D[1][1]);
            drot[2] = 2 * x * (D[1][0] + D[0][1]) + 2 * r * (D[2][0] - D[0][2]) + 2 * z * (D[1][2] + D[2][1]) - 4 * y * (D[2][2] + D[0][0]);
            drot[3] = 2 * r * (D[0][1] - D[1][0]) + 2 * x * (D[2][0] + D[0][2]) + 2 * y * (D[1][2] + D[2][1]) - 4 * z * (D[1][1] + D[0][0]);
        }
    }
    if (o.dL_dmean3D) {
#pragma unroll
        for (int k = 0; k < 3; k++) o.dL_dmean3D[3 * (size_t)idx + k] = dmean[k];
    }
    if (o.dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; k++) o.dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
    }
    if (o.dL_dscale) {
#pragma unroll
        for (int k = 0; k < 3; k++) o.dL_dscale[3 * (size_t)idx + k] = dscale[k];
    }
    if (o.dL_drot) reinterpret_cast<float4*>(o.dL_drot)[idx] = make_float4(drot[0], drot[1], drot[2], drot[3]);
}

}  // namespace

void launch_preprocess_forward(int P, const float* means3D, const float* scales, const float* rotations,
                               const float* opacities, const float* cov3D_precomp, const float* viewmatrix,
                               const float* projmatrix, int W, int H, float tanfovx, float tanfovy,
                               float scale_modifier, int* radii, const GeometryState& g, cudaStream_t st) {
    if (P <= 0) return;
    const float focal_y = H / (2.0f * tanfovy);  // rasterizer_impl.cu:226-227
    const float focal_x = W / (2.0f * tanfovx);
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    preprocess_forward_kernel<<<(P + PP_THREADS - 1) / PP_THREADS, PP_THREADS, 0, st>>>(
        P, means3D, scales, rotations, opacities, cov3D_precomp, viewmatrix, projmatrix, W, H, gx, gy, tanfovx,
        tanfovy, focal_x, focal_y, scale_modifier, radii, g.radii, g.means2D, g.depths, g.conic_opacity,
        g.tiles_touched);
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t st) {
    if (P <= 0) return;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, viewmatrix, present);
}

void launch_duplicate_with_keys(int P, const GeometryState& g, const int* radii, int W, int H, uint64_t* keys,
                                uint32_t* values, cudaStream_t st) {
    if (P <= 0) return;
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    duplicate_with_keys_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, g.means2D, g.depths, g.point_offsets, radii, gx,
                                                                gy, keys, values);
}

void launch_ranges_and_gather(int R, int tiles, int channels, const uint64_t* keys_sorted,
                              const uint32_t* point_list, const GeometryState& g, const float* features,
                              const BinningState& b, uint2* ranges, cudaStream_t st) {
    cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)tiles, st);  // rasterizer_impl.cu:312
    if (R <= 0) return;
    const int blocks = (R + 255) / 256;
    if (channels == 3)
        ranges_gather_kernel<3><<<blocks, 256, 0, st>>>(R, keys_sorted, point_list, g.means2D, g.conic_opacity,
                                                         features, b.inst_conic, b.inst_geom, b.inst_feat, ranges);
    else
        ranges_gather_kernel<6><<<blocks, 256, 0, st>>>(R, keys_sorted, point_list, g.means2D, g.conic_opacity,
                                                         features, b.inst_conic, b.inst_geom, b.inst_feat, ranges);
}

void launch_preprocess_backward(int P, int channels, const float* means3D, const int* radii, const float* scales,
                                const float* rotations, const float* cov3D_precomp, float scale_modifier,
                                const float* viewmatrix, const float* projmatrix, int W, int H, float tanfovx,
                                float tanfovy, const float* /*cov3D_state*/, const float* gacc,
                                const PreprocessBackwardOut& o, cudaStream_t st) {
    if (P <= 0) return;
    const float h_y = H / (2.0f * tanfovy), h_x = W / (2.0f * tanfovx);
    const int blocks = (P + 255) / 256;
    if (channels == 3)
        preprocess_backward_kernel<3><<<blocks, 256, 0, st>>>(P, means3D, radii, scales, rotations, cov3D_precomp,
                                                               scale_modifier, viewmatrix, projmatrix, h_x, h_y,
                                                               tanfovx, tanfovy, gacc, o);
    else
        preprocess_backward_kernel<6><<<blocks, 256, 0, st>>>(P, means3D, radii, scales, rotations, cov3D_precomp,
                                                               scale_modifier, viewmatrix, projmatrix, h_x, h_y,
                                                               tanfovx, tanfovy, gacc, o);
}

}  // namespace asb
