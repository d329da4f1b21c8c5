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
#include "raster_math.cuh"

namespace asb {

namespace {

constexpr int PP_THREADS = 256;

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

    const float px = s_means[3 * threadIdx.x], py = s_means[3 * threadIdx.x + 1], pz = s_means[3 * threadIdx.x + 2];
    int my_radii = 0;
    uint32_t my_tiles = 0;
    if (xform_row(s_view, 2, px, py, pz) > 0.2f) {  // near cull first: culled Gaussians skip the covariance work
        float cov3D[6];
        if (cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 6; k++) cov3D[k] = __ldg(cov3D_precomp + 6 * (size_t)idx + k);
        } else {
            const float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
            cov3d_from_scale_rot(s_scales[3 * threadIdx.x], s_scales[3 * threadIdx.x + 1],
                                 s_scales[3 * threadIdx.x + 2], scale_modifier, q, cov3D);
        }
        const PreprocessResult r = preprocess_one(px, py, pz, cov3D, __ldg(opacities + idx), s_view, s_proj, W, H, gx,
                                                  gy, tanfovx, tanfovy, focal_x, focal_y);
        if (r.radius > 0) {
            my_radii = r.radius;
            my_tiles = r.tiles;
            depths[idx] = r.depth;
            means2D_out[idx] = r.mean2D;
            conic_opacity[idx] = r.conic_opacity;
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
    const float4 co = __ldg(conic_opacity + id);
    inst_conic[i] = co;
    // depth is the low word of the key (rasterizer_impl.cu:101-103)
    inst_geom[i] = make_float4(m.x, m.y, __uint_as_float((uint32_t)key), __uint_as_float(id));
    const float ext = pack_cull_extent(co);  // spare slot of the feature record: culling box for the blend kernels
    const float* f = features + (size_t)id * C;
    if (C == 3) {
        inst_feat[i] = make_float4(__ldg(f), __ldg(f + 1), __ldg(f + 2), ext);
    } else {
        inst_feat[2 * (size_t)i] = make_float4(__ldg(f), __ldg(f + 1), __ldg(f + 2), __ldg(f + 3));
        inst_feat[2 * (size_t)i + 1] = make_float4(__ldg(f + 4), __ldg(f + 5), ext, 0.f);
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
        preprocess_backward_one(mx, my, mz, c3, !cov3D_precomp, q, s, scale_modifier, view, proj, h_x, h_y, tanfovx,
                                tanfovy, g0.x, g0.y, g0.z, g1.x, g1.y, dmean, dcov, dscale, drot);
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
