/*
 * raster_oracle.c -- CPU restatement of the reference rasterizer hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (activesplat_b200/) may call, link or
 * import this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs use it, and only as the checker / timed CPU baseline.
 *
 * It restates, in plain C (fp32, no fast-math, -ffp-contract=off), the algorithm of
 *   /root/reference/submodules/diff-gaussian-rasterization ("dgr/") :
 *     cuda_rasterizer/forward.cu, backward.cu, rasterizer_impl.cu, auxiliary.h, config.h
 * Each function cites the reference lines it follows.  The reference has no CPU code and no
 * golden vectors of its own (parity is pinned instead against outputs of the reference CUDA
 * build captured on a B200: tests/golden/ref_*.npz, made by oracle/make_golden.py).
 *
 * Forward per-Gaussian arithmetic (preprocess) is written with explicit fmaf() in exactly the
 * places where nvcc contracts the reference's expressions (rule observed in SASS:
 * fadd(fmul(a,b),c) -> fma(a,b,c), the LEFT product of a sum of two products is the one fused),
 * so that integer outputs derived from floats (radii, tile rects, keys) can be compared
 * bit-for-bit with a CUDA build.  Backward math is tolerance-checked and uses plain expressions.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BLOCK_X 16 /* dgr/cuda_rasterizer/config.h:16 */
#define BLOCK_Y 16 /* dgr/cuda_rasterizer/config.h:17 */
#define BLOCK_SIZE (BLOCK_X * BLOCK_Y)

/* a0*b0 + a1*b1 + a2*b2 as nvcc contracts it: fma(a2,b2, fma(a0,b0, a1*b1)) */
static inline float dot3c(float a0, float b0, float a1, float b1, float a2, float b2) {
    float t = a1 * b1;
    t = fmaf(a0, b0, t);
    return fmaf(a2, b2, t);
}

/* auxiliary.h:58-66 transformPoint4x3, one row r in {0,1,2}: m[r]x + m[4+r]y + m[8+r]z + m[12+r] */
static inline float xform_row(const float* m, int r, float x, float y, float z) {
    return dot3c(m[r], x, m[4 + r], y, m[8 + r], z) + m[12 + r];
}

/* CUDA float->int conversion saturates; C's is undefined out of range. */
static inline int f2i_sat(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return INT32_MAX;
    if (v <= -2147483648.0f) return INT32_MIN;
    return (int)v;
}
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

/* auxiliary.h:46-56 getRect */
static void get_rect(float px, float py, int max_radius, int gx, int gy, int* rmin, int* rmax) {
    float r = (float)max_radius;
    rmin[0] = imin(gx, imax(0, f2i_sat((px - r) / (float)BLOCK_X)));
    rmin[1] = imin(gy, imax(0, f2i_sat((py - r) / (float)BLOCK_Y)));
    rmax[0] = imin(gx, imax(0, f2i_sat((((px + r) + (float)BLOCK_X) - 1.0f) / (float)BLOCK_X)));
    rmax[1] = imin(gy, imax(0, f2i_sat((((py + r) + (float)BLOCK_Y) - 1.0f) / (float)BLOCK_Y)));
}

/* auxiliary.h:41-44 ndc2Pix -- the literals are double, so this is evaluated in fp64 */
static inline float ndc2pix(float v, int S) {
    return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5);
}

/* rasterizer_impl.cu:35-50 getHigherMsb */
uint32_t orc_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* forward.cu:118-152 computeCov3D (quaternion used as given, NOT normalised: forward.cu:127) */
static void cov3d_from_scale_rot(const float* s3, float mod, const float* q4, float* cov) {
    float sx = mod * s3[0], sy = mod * s3[1], sz = mod * s3[2];
    float r = q4[0], x = q4[1], y = q4[2], z = q4[3];
    /* textbook rotation rows (GLM's R is the transpose) */
    float R[3][3];
    R[0][0] = fmaf(-2.f, fmaf(y, y, z * z), 1.f);
    R[0][1] = 2.f * fmaf(x, y, -(r * z));
    R[0][2] = 2.f * fmaf(x, z, r * y);
    R[1][0] = 2.f * fmaf(x, y, r * z);
    R[1][1] = fmaf(-2.f, fmaf(x, x, z * z), 1.f);
    R[1][2] = 2.f * fmaf(y, z, -(r * x));
    R[2][0] = 2.f * fmaf(x, z, -(r * y));
    R[2][1] = 2.f * fmaf(y, z, r * x);
    R[2][2] = fmaf(-2.f, fmaf(x, x, y * y), 1.f);
    /* M = S * R_glm : M_glm[j][k] = s_k * Rtextbook[j][k]  (one rounding each) */
    float M[3][3];
    for (int j = 0; j < 3; j++) { M[j][0] = sx * R[j][0]; M[j][1] = sy * R[j][1]; M[j][2] = sz * R[j][2]; }
    /* Sigma = transpose(M) * M : Sigma[i][j] = sum_k M[i][k] M[j][k] */
    cov[0] = dot3c(M[0][0], M[0][0], M[0][1], M[0][1], M[0][2], M[0][2]);
    cov[1] = dot3c(M[1][0], M[0][0], M[1][1], M[0][1], M[1][2], M[0][2]);
    cov[2] = dot3c(M[2][0], M[0][0], M[2][1], M[0][1], M[2][2], M[0][2]);
    cov[3] = dot3c(M[1][0], M[1][0], M[1][1], M[1][1], M[1][2], M[1][2]);
    cov[4] = dot3c(M[2][0], M[1][0], M[2][1], M[1][1], M[2][2], M[1][2]);
    cov[5] = dot3c(M[2][0], M[2][0], M[2][1], M[2][1], M[2][2], M[2][2]);
}

/* Shared by forward (forward.cu:74-113) and backward (backward.cu:160-193): the clamped camera
 * point t, the two non-zero rows of T = W*J (GLM), stored as Tm[0][k] = T_glm[0][k], Tm[1][k] = T_glm[1][k]. */
static void cov2d_T(const float* mean, float fx, float fy, float tanfovx, float tanfovy,
                    const float* view, float t[3], float Tm[2][3], float* txtz_out, float* tytz_out) {
    t[0] = xform_row(view, 0, mean[0], mean[1], mean[2]);
    t[1] = xform_row(view, 1, mean[0], mean[1], mean[2]);
    t[2] = xform_row(view, 2, mean[0], mean[1], mean[2]);
    const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
    const float txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    if (txtz_out) *txtz_out = txtz;
    if (tytz_out) *tytz_out = tytz;
    const float J00 = fx / t[2], J02 = -(fx * t[0]) / (t[2] * t[2]);
    const float J11 = fy / t[2], J12 = -(fy * t[1]) / (t[2] * t[2]);
    /* W_glm[k][i] = view[4*i + k]; T_glm[0][i] = W[0][i]*J00 + W[1][i]*0 + W[2][i]*J02 */
    for (int i = 0; i < 3; i++) {
        Tm[0][i] = fmaf(view[4 * i + 2], J02, view[4 * i + 0] * J00);
        Tm[1][i] = fmaf(view[4 * i + 2], J12, view[4 * i + 1] * J11);
    }
}

/* cov = transpose(T) * transpose(Vrk) * T, entries [0][0], [0][1], [1][1] (forward.cu:100-112) */
static void cov2d_from_T(float Tm[2][3], const float* c, float* a, float* b, float* cc) {
    const float V[3][3] = {{c[0], c[1], c[2]}, {c[1], c[3], c[4]}, {c[2], c[4], c[5]}};
    /* A = Tt * Vrkt : A_glm[j][i] = sum_k T[i][k] V[k][j]; rows i = 0,1 are the non-zero ones */
    float A[3][2];
    for (int j = 0; j < 3; j++)
        for (int i = 0; i < 2; i++)
            A[j][i] = dot3c(Tm[i][0], V[0][j], Tm[i][1], V[1][j], Tm[i][2], V[2][j]);
    /* cov_glm[j][i] = sum_k A[k][i] * T[j][k] */
    *a = dot3c(A[0][0], Tm[0][0], A[1][0], Tm[0][1], A[2][0], Tm[0][2]);
    *b = dot3c(A[0][1], Tm[0][0], A[1][1], Tm[0][1], A[2][1], Tm[0][2]);
    *cc = dot3c(A[0][1], Tm[1][0], A[1][1], Tm[1][1], A[2][1], Tm[1][2]);
}

/*
 * forward.cu:155-256 preprocessCUDA (colors_precomp path; SH evaluation is unused by ActiveSplat)
 * with in_frustum (auxiliary.h:139-164).  Outputs that the reference leaves unwritten for culled
 * Gaussians (means2D, depths, cov3D, conic_opacity) are left untouched here as well.
 */
void orc_preprocess(int P, const float* means3D, const float* scales, float scale_modifier,
                    const float* rotations, const float* opacities, const float* cov3D_precomp,
                    const float* view, const float* proj, int W, int H, float tanfovx, float tanfovy,
                    int* radii, float* means2D, float* depths, float* cov3Ds, float* conic_opacity,
                    uint32_t* tiles_touched) {
    const float focal_y = H / (2.0f * tanfovy); /* rasterizer_impl.cu:226-227 */
    const float focal_x = W / (2.0f * tanfovx);
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
        const float* p = means3D + 3 * idx;
        /* in_frustum: only the near test survives (auxiliary.h:154) */
        const float pvz = xform_row(view, 2, p[0], p[1], p[2]);
        if (pvz <= 0.2f) continue;
        const float hx = xform_row(proj, 0, p[0], p[1], p[2]);
        const float hy = xform_row(proj, 1, p[0], p[1], p[2]);
        const float hw = xform_row(proj, 3, p[0], p[1], p[2]);
        const float p_w = 1.0f / (hw + 0.0000001f);
        const float projx = hx * p_w, projy = hy * p_w;

        const float* cov3D;
        if (cov3D_precomp) cov3D = cov3D_precomp + 6 * idx;
        else {
            cov3d_from_scale_rot(scales + 3 * idx, scale_modifier, rotations + 4 * idx, cov3Ds + 6 * idx);
            cov3D = cov3Ds + 6 * idx;
        }
        float t[3], Tm[2][3], a, b, c;
        cov2d_T(p, focal_x, focal_y, tanfovx, tanfovy, view, t, Tm, NULL, NULL);
        cov2d_from_T(Tm, cov3D, &a, &b, &c);
        a += 0.3f; c += 0.3f; /* forward.cu:110-111 */

        const float det = fmaf(a, c, -(b * b)); /* forward.cu:219 */
        if (det == 0.0f) continue;
        const float det_inv = 1.f / det;
        const float conx = c * det_inv, cony = -b * det_inv, conz = a * det_inv;

        const float mid = 0.5f * (a + c); /* forward.cu:229-232 */
        const float disc = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        const float lambda1 = mid + disc, lambda2 = mid - disc;
        const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
        const float pix_x = ndc2pix(projx, W), pix_y = ndc2pix(projy, H);
        int rmin[2], rmax[2];
        get_rect(pix_x, pix_y, f2i_sat(my_radius), gx, gy, rmin, rmax);
        if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;

        depths[idx] = pvz;
        radii[idx] = f2i_sat(my_radius);
        means2D[2 * idx] = pix_x;
        means2D[2 * idx + 1] = pix_y;
        conic_opacity[4 * idx + 0] = conx;
        conic_opacity[4 * idx + 1] = cony;
        conic_opacity[4 * idx + 2] = conz;
        conic_opacity[4 * idx + 3] = opacities[idx];
        tiles_touched[idx] = (uint32_t)((rmax[1] - rmin[1]) * (rmax[0] - rmin[0]));
    }
}

/* rasterizer_impl.cu:54-66 checkFrustum / :141-153 markVisible */
void orc_mark_visible(int P, const float* means3D, const float* view, uint8_t* present) {
    for (int i = 0; i < P; i++) {
        const float* p = means3D + 3 * i;
        present[i] = xform_row(view, 2, p[0], p[1], p[2]) > 0.2f;
    }
}

/* rasterizer_impl.cu:279 InclusiveSum; returns num_rendered (:283) */
int64_t orc_inclusive_scan(int P, const uint32_t* tiles_touched, uint32_t* offsets) {
    uint32_t s = 0;
    for (int i = 0; i < P; i++) { s += tiles_touched[i]; offsets[i] = s; }
    return (int64_t)s;
}

/* rasterizer_impl.cu:70-111 duplicateWithKeys */
void orc_duplicate_with_keys(int P, const float* means2D, const float* depths, const uint32_t* offsets,
                             const int* radii, int W, int H, uint64_t* keys, uint32_t* values) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        if (radii[idx] <= 0) continue;
        uint32_t off = idx == 0 ? 0 : offsets[idx - 1];
        int rmin[2], rmax[2];
        get_rect(means2D[2 * idx], means2D[2 * idx + 1], radii[idx], gx, gy, rmin, rmax);
        uint32_t dbits;
        memcpy(&dbits, depths + idx, 4);
        for (int y = rmin[1]; y < rmax[1]; y++)
            for (int x = rmin[0]; x < rmax[0]; x++) {
                uint64_t key = (uint64_t)(y * gx + x);
                key <<= 32;
                key |= dbits;
                keys[off] = key;
                values[off] = (uint32_t)idx;
                off++;
            }
    }
}

/* rasterizer_impl.cu:305-310 cub::DeviceRadixSort::SortPairs(begin_bit 0, end_bit) -- a stable
 * ascending sort on the low end_bit bits; restated as a stable LSD byte radix sort. */
void orc_sort_pairs(int64_t R, const uint64_t* keys_in, const uint32_t* vals_in,
                    uint64_t* keys_out, uint32_t* vals_out, int end_bit) {
    if (R <= 0) return;
    uint64_t* kb = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)R);
    uint32_t* vb = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)R);
    memcpy(keys_out, keys_in, sizeof(uint64_t) * (size_t)R);
    memcpy(vals_out, vals_in, sizeof(uint32_t) * (size_t)R);
    uint64_t *ks = keys_out, *kd = kb;
    uint32_t *vs = vals_out, *vd = vb;
    for (int shift = 0; shift < end_bit; shift += 8) {
        int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint32_t mask = (1u << bits) - 1u;
        int64_t count[257];
        memset(count, 0, sizeof(count));
        for (int64_t i = 0; i < R; i++) count[((ks[i] >> shift) & mask) + 1]++;
        for (int b = 0; b < 256; b++) count[b + 1] += count[b];
        for (int64_t i = 0; i < R; i++) {
            int64_t d = count[(ks[i] >> shift) & mask]++;
            kd[d] = ks[i]; vd[d] = vs[i];
        }
        uint64_t* tk = ks; ks = kd; kd = tk;
        uint32_t* tv = vs; vs = vd; vd = tv;
    }
    if (ks != keys_out) {
        memcpy(keys_out, ks, sizeof(uint64_t) * (size_t)R);
        memcpy(vals_out, vs, sizeof(uint32_t) * (size_t)R);
    }
    free(kb); free(vb);
}

/* rasterizer_impl.cu:312 (memset) + :116-138 identifyTileRanges.  ranges is uint2[tiles]. */
void orc_tile_ranges(int64_t R, const uint64_t* keys, int tiles, uint32_t* ranges) {
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)tiles);
    for (int64_t i = 0; i < R; i++) {
        uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
}

/*
 * forward.cu:261-397 renderCUDA.  C = number of feature channels (3 in the reference build;
 * 6 = the two SplaTAM passes [rgb | z,1,z^2] evaluated in one sweep, which is arithmetically the
 * same as two 3-channel sweeps because alpha/T do not depend on the features).
 * The per-pixel loop order, the skip rules and the termination rule are T10-T14 of SURVEY 8a'.
 */
void orc_blend_forward(int W, int H, int C, const uint32_t* ranges, const uint32_t* point_list,
                       const float* means2D, const float* features, const float* conic_opacity,
                       const float* depths, const float* bg, float* out_color, float* out_depth,
                       float* out_opacity, float* final_T, uint32_t* n_contrib) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; tile++) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int ly = 0; ly < BLOCK_Y; ly++)
            for (int lx = 0; lx < BLOCK_X; lx++) {
                const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
                if (px >= W || py >= H) continue;
                const int pix_id = W * py + px;
                const float pxf = (float)px, pyf = (float)py;
                float T = 1.0f, D = 15.0f; /* forward.cu:304,309 */
                float Cacc[8] = {0};
                uint32_t contributor = 0, last = 0;
                for (uint32_t k = r0; k < r1; k++) {
                    contributor++; /* forward.cu:335 */
                    const uint32_t id = point_list[k];
                    const float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                    const float* co = conic_opacity + 4 * id;
                    /* forward.cu:342 with nvcc's contraction */
                    const float q = fmaf(co[0] * dx, dx, (co[2] * dy) * dy);
                    const float power = fmaf(q, -0.5f, -((co[1] * dx) * dy));
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, co[3] * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) break; /* done = true: nothing further is blended */
                    for (int ch = 0; ch < C; ch++)
                        Cacc[ch] = fmaf(features[id * C + ch] * alpha, T, Cacc[ch]);
                    if (T > 0.5f && test_T < 0.5f) D = depths[id]; /* median depth, forward.cu:370-374 */
                    T = test_T;
                    last = contributor;
                }
                final_T[pix_id] = T;
                n_contrib[pix_id] = last;
                for (int ch = 0; ch < C; ch++)
                    out_color[(size_t)ch * H * W + pix_id] = fmaf(T, bg[ch], Cacc[ch]);
                out_depth[pix_id] = D;
                out_opacity[pix_id] = 1 - T;
            }
    }
}

static inline void atomic_addf(float* p, float v) {
#pragma omp atomic
    *p += v;
}

/*
 * backward.cu:399-557 renderCUDA (backward).  Gradient buffers must be zeroed by the caller
 * (rasterize_points.cu:155-163).  dL_dmean2D is [P,3] (z untouched), dL_dconic is [P,4] (x,y,w used).
 */
void orc_blend_backward(int W, int H, int C, const uint32_t* ranges, const uint32_t* point_list,
                        const float* bg, const float* means2D, const float* conic_opacity,
                        const float* colors, const float* final_Ts, const uint32_t* n_contrib,
                        const float* dL_dpixels, float* dL_dmean2D, float* dL_dconic2D,
                        float* dL_dopacity, float* dL_dcolors) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H; /* backward.cu:460-461 */
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; tile++) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int ly = 0; ly < BLOCK_Y; ly++)
            for (int lx = 0; lx < BLOCK_X; lx++) {
                const int px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
                if (px >= W || py >= H) continue;
                const int pix_id = W * py + px;
                const float pxf = (float)px, pyf = (float)py;
                const float T_final = final_Ts[pix_id];
                float T = T_final;
                const uint32_t last_contributor = n_contrib[pix_id];
                float accum_rec[8] = {0}, last_color[8] = {0}, dL_dpixel[8];
                float bg_dot_dpixel = 0;
                for (int ch = 0; ch < C; ch++) {
                    dL_dpixel[ch] = dL_dpixels[(size_t)ch * H * W + pix_id];
                    bg_dot_dpixel += bg[ch] * dL_dpixel[ch];
                }
                float last_alpha = 0;
                /* walk the tile list back to front (backward.cu:472-503) */
                uint32_t contributor = r1 - r0;
                for (uint32_t k = r1; k-- > r0;) {
                    contributor--;
                    if (contributor >= last_contributor) continue;
                    const uint32_t id = point_list[k];
                    const float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                    const float* co = conic_opacity + 4 * id;
                    const float q = fmaf(co[0] * dx, dx, (co[2] * dy) * dy);
                    const float power = fmaf(q, -0.5f, -((co[1] * dx) * dy));
                    if (power > 0.0f) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, co[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
                    for (int ch = 0; ch < C; ch++) {
                        const float c = colors[id * C + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        dL_dalpha += (c - accum_rec[ch]) * dL_dpixel[ch];
                        atomic_addf(&dL_dcolors[id * C + ch], dchannel_dcolor * dL_dpixel[ch]);
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel; /* backward.cu:531-534 */
                    /* clamp at 0.99 is straight-through (backward.cu:538) */
                    const float dL_dG = co[3] * dL_dalpha;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                    const float dG_ddely = -gdy * co[2] - gdx * co[1];
                    atomic_addf(&dL_dmean2D[3 * id + 0], dL_dG * dG_ddelx * ddelx_dx);
                    atomic_addf(&dL_dmean2D[3 * id + 1], dL_dG * dG_ddely * ddely_dy);
                    atomic_addf(&dL_dconic2D[4 * id + 0], -0.5f * gdx * dx * dL_dG);
                    atomic_addf(&dL_dconic2D[4 * id + 1], -0.5f * gdx * dy * dL_dG);
                    atomic_addf(&dL_dconic2D[4 * id + 3], -0.5f * gdy * dy * dL_dG);
                    atomic_addf(&dL_dopacity[id], G * dL_dalpha);
                }
            }
    }
}

/* backward.cu:144-274 computeCov2DCUDA.  Writes dL_dcov3D[P,6] and OVERWRITES dL_dmean3D (:273). */
void orc_cov2d_backward(int P, const float* means3D, const int* radii, const float* cov3Ds,
                        int W, int H, float tanfovx, float tanfovy, const float* view,
                        const float* dL_dconics, float* dL_dmeans, float* dL_dcov) {
    const float h_y = H / (2.0f * tanfovy), h_x = W / (2.0f * tanfovx);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        if (!(radii[idx] > 0)) continue;
        const float* c3 = cov3Ds + 6 * idx;
        const float dcx = dL_dconics[4 * idx], dcy = dL_dconics[4 * idx + 1], dcz = dL_dconics[4 * idx + 3];
        float t[3], Tm[2][3], txtz, tytz;
        cov2d_T(means3D + 3 * idx, h_x, h_y, tanfovx, tanfovy, view, t, Tm, &txtz, &tytz);
        const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        float a, b, c;
        cov2d_from_T(Tm, c3, &a, &b, &c);
        a += 0.3f; c += 0.3f;
        const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
        /* GLM T[0][k] = Tm[0][k], T[1][k] = Tm[1][k] */
        const float (*T)[3] = Tm;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f); /* backward.cu:203 */
        float* dc = dL_dcov + 6 * idx;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dc[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
            dc[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
            dc[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
            dc[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
            dc[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
            dc[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
        } else {
            for (int i = 0; i < 6; i++) dc[i] = 0;
        }
        float dL_dT[2][3];
        for (int k = 0; k < 3; k++) {
            const float tv0 = T[0][0] * V[k][0] + T[0][1] * V[k][1] + T[0][2] * V[k][2];
            const float tv1 = T[1][0] * V[k][0] + T[1][1] * V[k][1] + T[1][2] * V[k][2];
            dL_dT[0][k] = 2 * tv0 * dL_da + tv1 * dL_db;
            dL_dT[1][k] = 2 * tv1 * dL_dc + tv0 * dL_db;
        }
        /* W_glm[k][i] = view[4*i+k] (backward.cu:181-184) */
#define WG(k, i) view[4 * (i) + (k)]
        const float dL_dJ00 = WG(0, 0) * dL_dT[0][0] + WG(0, 1) * dL_dT[0][1] + WG(0, 2) * dL_dT[0][2];
        const float dL_dJ02 = WG(2, 0) * dL_dT[0][0] + WG(2, 1) * dL_dT[0][1] + WG(2, 2) * dL_dT[0][2];
        const float dL_dJ11 = WG(1, 0) * dL_dT[1][0] + WG(1, 1) * dL_dT[1][1] + WG(1, 2) * dL_dT[1][2];
        const float dL_dJ12 = WG(2, 0) * dL_dT[1][0] + WG(2, 1) * dL_dT[1][1] + WG(2, 2) * dL_dT[1][2];
#undef WG
        const float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t[0]) * tz3 * dL_dJ02 + (2 * h_y * t[1]) * tz3 * dL_dJ12;
        /* transformVec4x3Transpose (auxiliary.h:91-99) */
        dL_dmeans[3 * idx + 0] = view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz;
        dL_dmeans[3 * idx + 1] = view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz;
        dL_dmeans[3 * idx + 2] = view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz;
    }
}

/* backward.cu:346-396 preprocessCUDA (backward) + :278-341 computeCov3D (backward); SH path unused. */
void orc_preprocess_backward(int P, const float* means3D, const int* radii, const float* scales,
                             const float* rotations, float scale_modifier, const float* proj,
                             const float* dL_dmean2D, float* dL_dmeans, const float* dL_dcov3D,
                             float* dL_dscale, float* dL_drot) {
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        if (!(radii[idx] > 0)) continue;
        const float mx = means3D[3 * idx], my = means3D[3 * idx + 1], mz = means3D[3 * idx + 2];
        const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
        const float m_w = 1.0f / (m_hw + 0.0000001f);
        const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
        const float gx = dL_dmean2D[3 * idx], gy = dL_dmean2D[3 * idx + 1];
        dL_dmeans[3 * idx + 0] += (proj[0] * m_w - proj[3] * mul1) * gx + (proj[1] * m_w - proj[3] * mul2) * gy;
        dL_dmeans[3 * idx + 1] += (proj[4] * m_w - proj[7] * mul1) * gx + (proj[5] * m_w - proj[7] * mul2) * gy;
        dL_dmeans[3 * idx + 2] += (proj[8] * m_w - proj[11] * mul1) * gx + (proj[9] * m_w - proj[11] * mul2) * gy;
        if (!scales) continue;
        /* computeCov3D backward (backward.cu:278-341); no quaternion-normalisation Jacobian (:340) */
        const float* q = rotations + 4 * idx;
        const float r = q[0], x = q[1], y = q[2], z = q[3];
        const float Rt[3][3] = { /* textbook rotation rows */
            {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
            {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
            {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
        const float s[3] = {scale_modifier * scales[3 * idx], scale_modifier * scales[3 * idx + 1], scale_modifier * scales[3 * idx + 2]};
        const float* d = dL_dcov3D + 6 * idx;
        const float dS[3][3] = {{d[0], 0.5f * d[1], 0.5f * d[2]}, {0.5f * d[1], d[3], 0.5f * d[4]}, {0.5f * d[2], 0.5f * d[4], d[5]}};
        /* In textbook terms: Sigma = Rt diag(s^2) Rt^T, M_glm column j = s (.) Rt[j][:].
         * dL_dMt (GLM) [i][j] = dL/d(M_glm[j][i]) ... worked out element-wise:
         * G[k][j] := dL/d(s_k * Rt[j][k]) = 2 * sum_i dS[j][i] * s_k * Rt[i][k]                     */
        float G[3][3];
        for (int k = 0; k < 3; k++)
            for (int j = 0; j < 3; j++)
                G[k][j] = 2.0f * (s[k] * Rt[0][k] * dS[j][0] + s[k] * Rt[1][k] * dS[j][1] + s[k] * Rt[2][k] * dS[j][2]);
        for (int k = 0; k < 3; k++)
            dL_dscale[3 * idx + k] = Rt[0][k] * G[k][0] + Rt[1][k] * G[k][1] + Rt[2][k] * G[k][2];
        /* dL/dRt[j][k] = s_k * G[k][j];  name it D[k][j] to match the reference's dL_dMt[k][j] after *= s */
        float D[3][3];
        for (int k = 0; k < 3; k++)
            for (int j = 0; j < 3; j++) D[k][j] = s[k] * G[k][j];
        float* o = dL_drot + 4 * idx;
        o[0] = 2 * z * (D[0][1] - D[1][0]) + 2 * y * (D[2][0] - D[0][2]) + 2 * x * (D[1][2] - D[2][1]);
        o[1] = 2 * y * (D[1][0] + D[0][1]) + 2 * z * (D[2][0] + D[0][2]) + 2 * r * (D[1][2] - D[2][1]) - 4 * x * (D[2][2] + D[1][1]);
        o[2] = 2 * x * (D[1][0] + D[0][1]) + 2 * r * (D[2][0] - D[0][2]) + 2 * z * (D[1][2] + D[2][1]) - 4 * y * (D[2][2] + D[0][0]);
        o[3] = 2 * r * (D[0][1] - D[1][0]) + 2 * x * (D[2][0] + D[0][2]) + 2 * y * (D[1][2] + D[2][1]) - 4 * z * (D[1][1] + D[0][0]);
    }
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
    extern void omp_set_num_threads(int);
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_num_threads(void) {
#ifdef _OPENMP
    extern int omp_get_max_threads(void);
    return omp_get_max_threads();
#else
    return 1;
#endif
}
