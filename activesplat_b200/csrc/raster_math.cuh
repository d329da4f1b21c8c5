// raster_math.cuh -- per-Gaussian device math shared by preprocess.cu and fused.cu (sm_100a).
// The forward functions use rounding-explicit intrinsics in the operation order nvcc gives the
// reference's expressions (dgr/cuda_rasterizer/forward.cu, auxiliary.h), see DESIGN.md section 2.
#pragma once
#include <cuda_fp16.h>
#include <math_constants.h>

#include "common.cuh"

namespace asb {

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


// Culling level L of a Gaussian for the blend kernels: a pixel can pass their skip rules (power <= 0 and
// opacity * exp(power) >= 1/255; forward.cu:343-353) only if  A dx^2 + 2B dx dy + C dy^2 <= 2 ln(255 opacity) =: L.
// Stored with a small margin.  Negative = can never contribute; +inf = unknown, never culled.
__device__ __forceinline__ float pack_cull_extent(float4 co) {
    if (!(co.w >= 1.0f / 255.0f))  // alpha = min(0.99, op * G) < 1/255 for every G <= 1 (NaN opacity: keep)
        return (co.w == co.w) ? -1.f : CUDART_INF_F;
    const float det = co.x * co.z - co.y * co.y;
    const float L = 2.f * logf(255.f * co.w);
    if (!(det > 0.f) || !(co.x > 0.f) || !isfinite(det) || !(L >= 0.f) || !isfinite(L)) return CUDART_INF_F;
    return L * 1.001f + 1e-4f;
}

// Result of the forward per-Gaussian computation (forward.cu:155-256, colors_precomp path).
struct PreprocessResult {
    int radius;        // 0 = culled
    uint32_t tiles;    // tiles touched
    float depth;       // p_view.z
    float2 mean2D;
    float4 conic_opacity;
};

__device__ __forceinline__ PreprocessResult preprocess_one(float px, float py, float pz, const float* cov3D,
                                                           float opacity, const float* s_view, const float* s_proj,
                                                           int W, int H, int gx, int gy, float tanfovx, float tanfovy,
                                                           float focal_x, float focal_y) {
    PreprocessResult r;
    r.radius = 0;
    r.tiles = 0;
    r.depth = 0.f;
    r.mean2D = make_float2(0.f, 0.f);
    r.conic_opacity = make_float4(0.f, 0.f, 0.f, 0.f);
    // in_frustum (auxiliary.h:139-164): only the near test is live; caller checked pvz > 0.2
    const float pvz = xform_row(s_view, 2, px, py, pz);
    if (!(pvz > 0.2f)) return r;
    const float hx = xform_row(s_proj, 0, px, py, pz);
    const float hy = xform_row(s_proj, 1, px, py, pz);
    const float hw = xform_row(s_proj, 3, px, py, pz);
    const float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
    const float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
    float t[3], Tm[2][3], txtz, tytz, a, b, c;
    cov2d_T(px, py, pz, focal_x, focal_y, tanfovx, tanfovy, s_view, t, Tm, txtz, tytz);
    cov2d_from_T(Tm, cov3D, a, b, c);
    a = __fadd_rn(a, 0.3f);
    c = __fadd_rn(c, 0.3f);
    const float det = __fmaf_rn(a, c, -__fmul_rn(b, b));
    if (det == 0.0f) return r;
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
    if (area == 0) return r;
    r.radius = ir;
    r.tiles = area;
    r.depth = pvz;
    r.mean2D = make_float2(pix_x, pix_y);
    r.conic_opacity = make_float4(__fmul_rn(c, det_inv), __fmul_rn(-b, det_inv), __fmul_rn(a, det_inv), opacity);
    return r;
}

// computeCov2DCUDA (backward.cu:144-274) + preprocessCUDA bwd (backward.cu:346-396, :278-341) for one
// visible Gaussian.  g0 = (dL/dconic x,y,w | -), gm2 = dL/dmean2D (x,y).  has_sr: scale/rotation given.
__device__ __forceinline__ void preprocess_backward_one(float mx, float my, float mz, const float* c3, bool has_sr,
                                                        float4 q, const float* s, float scale_modifier,
                                                        const float* view, const float* proj, float h_x, float h_y,
                                                        float tanfovx, float tanfovy, float dcx, float dcy, float dcz,
                                                        float gx2, float gy2, float dmean[3], float dcov[6],
                                                        float dscale[3], float drot[4]) {
    float t[3], T[2][3], txtz, tytz, a, b, c;
    cov2d_T(mx, my, mz, h_x, h_y, tanfovx, tanfovy, view, t, T, txtz, tytz);
    const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    cov2d_from_T(T, c3, a, b, c);
    a += 0.3f;
    c += 0.3f;
    const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
    const float denom = a * c - b * b;
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);  // backward.cu:203
#pragma unroll
    for (int k = 0; k < 6; k++) dcov[k] = 0.f;
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

    // preprocessCUDA backward: projection chain of dL_dmean2D (backward.cu:373-387)
    const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
    const float m_w = 1.0f / (m_hw + 0.0000001f);
    const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
    dmean[0] += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
    dmean[1] += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
    dmean[2] += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;

#pragma unroll
    for (int k = 0; k < 3; k++) dscale[k] = 0.f;
#pragma unroll
    for (int k = 0; k < 4; k++) drot[k] = 0.f;
    // computeCov3D backward (backward.cu:278-341); no normalisation Jacobian (:340)
    if (has_sr) {
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

}  // namespace asb
