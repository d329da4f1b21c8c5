// blend.cu -- per-tile alpha blending, forward and backward (sm_100a).
//
//   blend_forward_kernel<C>   replaces renderCUDA<3> (forward)   dgr/cuda_rasterizer/forward.cu:261-397
//   blend_backward_kernel<C>  replaces renderCUDA<3> (backward)  dgr/cuda_rasterizer/backward.cu:399-557
//
// One CTA per 16x16 tile, one thread per pixel (same as the reference).  What differs:
//   * the tile's instances were gathered into contiguous per-instance records in sorted order
//     (ranges_gather_kernel), so a batch is THREE contiguous slabs that one elected thread pulls
//     into shared memory with cp.async.bulk (TMA engine, SASS UBLKCP) on an mbarrier, double
//     buffered: the copy of batch k+1/k+2 overlaps the blending of batch k;
//   * C = 6 blends [rgb | z,1,z^2] in one sweep (the reference makes two full passes);
//   * backward: each warp reduces a Gaussian's 6+C gradient terms with a transposing butterfly
//     (16 shuffles instead of 80), adds them to a per-batch accumulator in shared memory, and the
//     CTA flushes once per (tile, instance) with 128-bit vector reductions (REDG.ADD.F32x4)
//     -- versus 9 scalar global atomics per (pixel, Gaussian) in the reference (backward.cu:523-554);
//   * backward skips the tail of the list that no pixel of the warp/tile reached in the forward.
// Per-pixel arithmetic, skip rules and termination are the reference's (SURVEY 8a' T10-T18) and are
// spelled with rounding-explicit intrinsics so alpha decisions match the reference build bit for bit.
#include <cuda_fp16.h>

#include "common.cuh"

namespace asb {

namespace {

constexpr int BL_THREADS = TILE_PIX;  // 256
constexpr int BL_BATCH = 128;
constexpr int BL_STAGES = 4;  // bulk-copy pipeline depth (power of two): batches b..b+3 are in flight
constexpr int FW_UNROLL = 8;  // instances whose alpha is evaluated together (ILP); divides BL_BATCH

__device__ __forceinline__ float eval_power(const float4& co, float dx, float dy) {
    // forward.cu:342: -0.5f*(A*dx*dx + C*dy*dy) - B*dx*dy, contracted as nvcc contracts it
    const float q = __fmaf_rn(__fmul_rn(co.x, dx), dx, __fmul_rn(__fmul_rn(co.z, dy), dy));
    return __fmaf_rn(q, -0.5f, -__fmul_rn(__fmul_rn(co.y, dx), dy));
}

template <int C>
struct BlendSmem {
    static constexpr int FV = (C == 3) ? 1 : 2;
    float4 conic[BL_STAGES][BL_BATCH];
    float4 geom[BL_STAGES][BL_BATCH];
    float4 feat[BL_STAGES][BL_BATCH * FV];
    uint64_t bar[BL_STAGES];
};

template <int C>
__device__ __forceinline__ void issue_batch(BlendSmem<C>& sm, int stage, const float4* __restrict__ inst_conic,
                                            const float4* __restrict__ inst_geom,
                                            const float4* __restrict__ inst_feat, uint32_t start, int cnt) {
    constexpr int FV = BlendSmem<C>::FV;
    const uint32_t bytes = (uint32_t)cnt * 16u;
    mbar_expect_tx(&sm.bar[stage], bytes * (2 + FV));
    bulk_g2s(sm.conic[stage], inst_conic + start, bytes, &sm.bar[stage]);
    bulk_g2s(sm.geom[stage], inst_geom + start, bytes, &sm.bar[stage]);
    bulk_g2s(sm.feat[stage], inst_feat + (size_t)start * FV, bytes * FV, &sm.bar[stage]);
}

// ---- per-warp culling ---------------------------------------------------------------------------
// A pair (pixel, Gaussian) can only pass the skip rules (power <= 0, alpha >= 1/255; forward.cu:343-353)
// inside the ellipse  A dx^2 + 2B dx dy + C dy^2 <= 2 ln(255 * opacity).  ranges_gather stores the
// ellipse's axis-aligned half-extents (rounded UP, with margin) per instance; a warp whose 16x2-pixel
// strip lies outside that box skips the instance without evaluating anything.  Skipped pairs would have
// been skipped by the rules anyway, so images, n_contrib and gradients are unchanged.
template <int C>
__device__ __forceinline__ float2 inst_extent(const BlendSmem<C>& sm, int s, int j) {
    constexpr int FV = BlendSmem<C>::FV;
    const float4 f = sm.feat[s][j * FV + FV - 1];
    const uint32_t bits = __float_as_uint(C == 3 ? f.w : f.z);
    const __half2 h = *reinterpret_cast<const __half2*>(&bits);
    return __half22float2(h);
}

// Compacts the indices of the batch's records that can touch this warp's pixel strip into the warp's
// private list (ascending record order) and returns how many there are.
template <int C>
__device__ __forceinline__ int cull_batch(const BlendSmem<C>& sm, int s, int limit, int lane, float strip_x0,
                                          float strip_yc, uint8_t* list) {
    int n = 0;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < BL_BATCH / 32; k++) {
        const int j = lane + 32 * k;
        bool ok = false;
        if (j < limit) {
            const float4 ge = sm.geom[s][j];
            const float2 e = inst_extent<C>(sm, s, j);
            ok = (fabsf(ge.y - strip_yc) <= e.y + 0.5f) && (ge.x + e.x >= strip_x0) &&
                 (ge.x - e.x <= strip_x0 + (float)(TILE_X - 1));
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, ok);
        if (ok) list[n + __popc(bal & lt)] = (uint8_t)j;
        n += __popc(bal);
    }
    __syncwarp();
    return n;
}

// WARPS warps per CTA: a CTA blends 2*WARPS pixel rows of one tile, so 8/WARPS CTAs share a tile (each
// stages the tile's records for itself).  Small images have few tiles; splitting them keeps all SMs busy
// and spreads a heavy tile over several SMs.  Per-pixel results do not depend on WARPS.
template <int C, int WARPS>
__global__ void __launch_bounds__(32 * WARPS)
blend_forward_kernel(const uint2* __restrict__ ranges, const float4* __restrict__ inst_conic,
                     const float4* __restrict__ inst_geom, const float4* __restrict__ inst_feat, int W, int H,
                     int gx, const float* __restrict__ bg, float* __restrict__ out_color,
                     float* __restrict__ out_depth, float* __restrict__ out_opacity, float* __restrict__ final_T,
                     uint32_t* __restrict__ n_contrib) {
    constexpr int FV = BlendSmem<C>::FV;
    constexpr int NT = 32 * WARPS, PARTS = (BL_THREADS / 32) / WARPS;
    __shared__ __align__(128) BlendSmem<C> sm;
    __shared__ uint8_t s_list[WARPS][BL_BATCH];
    const int tile = blockIdx.x / PARTS;
    const int tx = tile % gx, ty = tile / gx;
    const int lane = threadIdx.x & 31;
    const int strip = (blockIdx.x % PARTS) * WARPS + (threadIdx.x >> 5);  // which 16x2 strip of the tile
    const int px = tx * TILE_X + (lane & (TILE_X - 1));
    const int py = ty * TILE_Y + 2 * strip + (lane >> 4);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    // this warp's pixels: columns [tile x0, +15], rows 2*strip and 2*strip+1
    const float strip_x0 = (float)(tx * TILE_X);
    const float strip_yc = (float)(ty * TILE_Y + 2 * strip) + 0.5f;
    const uint2 range = ranges[tile];
    const int len = (int)(range.y - range.x);
    const int nb = (len + BL_BATCH - 1) / BL_BATCH;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int q = 0; q < BL_STAGES; q++) mbar_init(&sm.bar[q], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int q = 0; q < BL_STAGES && q < nb; q++)
            issue_batch<C>(sm, q, inst_conic, inst_geom, inst_feat, range.x + q * BL_BATCH,
                           min(BL_BATCH, len - q * BL_BATCH));
    }

    bool done = !inside;
    float T = 1.0f;
    float D = 15.0f;  // median depth default (forward.cu:309)
    uint32_t last_contributor = 0;
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ch++) acc[ch] = 0.f;

    int consumed = 0;
    for (int b = 0; b < nb; b++) {
        const int s = b & (BL_STAGES - 1);
        mbar_wait(&sm.bar[s], (uint32_t)((b / BL_STAGES) & 1));
        const int cnt = min(BL_BATCH, len - b * BL_BATCH);
        // The per-pixel loop of forward.cu:332-381, restructured so that the warp stays converged:
        // (0) the warp drops the records whose ellipse cannot reach its strip,
        // (1) alpha of up to FW_UNROLL surviving records is evaluated branch-free for all lanes
        //     (independent chains: the exp latency of one record hides behind the others),
        // (2) the order-dependent part (T, colour, median depth, termination) is applied in list order,
        //     only for the records that passed the skip rules.  Same arithmetic per pixel.
        bool warp_done = __all_sync(0xffffffffu, done);
        if (!warp_done) {
            uint8_t* list = s_list[threadIdx.x >> 5];
            const int n = cull_batch<C>(sm, s, cnt, lane, strip_x0, strip_yc, list);
#pragma unroll 1
            for (int t = 0; t < n && !warp_done; t += FW_UNROLL) {
                int jj[FW_UNROLL];
#pragma unroll
                for (int u = 0; u < FW_UNROLL; u++) jj[u] = (t + u < n) ? (int)list[t + u] : -1;
                float al[FW_UNROLL];
                uint32_t hit = 0;
#pragma unroll
                for (int u = 0; u < FW_UNROLL; u++) {
                    const int j = max(jj[u], 0);
                    const float4 co = sm.conic[s][j];
                    const float4 ge = sm.geom[s][j];
                    const float dx = __fsub_rn(ge.x, pxf), dy = __fsub_rn(ge.y, pyf);
                    const float power = eval_power(co, dx, dy);
                    const float alpha = fminf(0.99f, __fmul_rn(co.w, expf(power)));
                    al[u] = alpha;
                    // skip rules of forward.cu:343-353 (written so that NaN behaves as in the reference)
                    if (jj[u] >= 0 && !(power > 0.0f) && !(alpha < 1.0f / 255.0f)) hit |= 1u << u;
                }
                if (done) hit = 0;
                if (__any_sync(0xffffffffu, hit != 0)) {
                    // order-dependent part, branch-free: the only serial chain is T -> test_T -> T
                    // (one multiply, one compare, one select per record); everything else is predicated
                    // selects, so a warp with many hits is not slowed down by 8 divergent branches.
#pragma unroll
                    for (int u = 0; u < FW_UNROLL; u++) {
                        const int j = max(jj[u], 0);
                        const float alpha = al[u];
                        const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
                        const bool h = ((hit >> u) & 1u) && !done;
                        const bool stop = h && (test_T < 0.0001f);  // forward.cu:355-359: stop BEFORE blending it
                        const bool blend = h && !stop;
                        done = done || stop;
                        const float4 f0 = sm.feat[s][j * FV];
                        const float4 ge = sm.geom[s][j];
                        acc[0] = blend ? __fmaf_rn(__fmul_rn(f0.x, alpha), T, acc[0]) : acc[0];
                        acc[1] = blend ? __fmaf_rn(__fmul_rn(f0.y, alpha), T, acc[1]) : acc[1];
                        acc[2] = blend ? __fmaf_rn(__fmul_rn(f0.z, alpha), T, acc[2]) : acc[2];
                        if (C == 6) {
                            const float4 f1 = sm.feat[s][j * FV + FV - 1];
                            acc[3 % C] = blend ? __fmaf_rn(__fmul_rn(f0.w, alpha), T, acc[3 % C]) : acc[3 % C];
                            acc[4 % C] = blend ? __fmaf_rn(__fmul_rn(f1.x, alpha), T, acc[4 % C]) : acc[4 % C];
                            acc[5 % C] = blend ? __fmaf_rn(__fmul_rn(f1.y, alpha), T, acc[5 % C]) : acc[5 % C];
                        }
                        D = (blend && T > 0.5f && test_T < 0.5f) ? ge.z : D;  // forward.cu:370-374
                        last_contributor = blend ? (uint32_t)(b * BL_BATCH + j + 1) : last_contributor;  // :335,380
                        T = blend ? test_T : T;
                    }
                    warp_done = __all_sync(0xffffffffu, done);
                }
            }
            __syncwarp();  // the list is rewritten by the next batch
        }
        consumed = b + 1;
        // all 256 pixels finished => stop (forward.cu:315-317); the barrier also frees stage s for reuse
        if (__syncthreads_count(done) == NT) break;
        if (threadIdx.x == 0 && b + BL_STAGES < nb) {
            const int nstart = (b + BL_STAGES) * BL_BATCH;
            issue_batch<C>(sm, s, inst_conic, inst_geom, inst_feat, range.x + nstart, min(BL_BATCH, len - nstart));
        }
    }
    // copies that were issued but never consumed must land before this CTA's shared memory is released:
    // after an early exit at batch consumed-1, batches consumed .. consumed+BL_STAGES-2 are still in flight
    if (consumed >= 1)
        for (int q = consumed; q < nb && q < consumed - 1 + BL_STAGES; q++)
            mbar_wait(&sm.bar[q & (BL_STAGES - 1)], (uint32_t)((q / BL_STAGES) & 1));

    if (inside) {
        const size_t pix_id = (size_t)W * py + px;
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < C; ch++) out_color[(size_t)ch * H * W + pix_id] = __fmaf_rn(T, __ldg(bg + ch), acc[ch]);
        out_depth[pix_id] = D;
        out_opacity[pix_id] = __fsub_rn(1.f, T);
    }
}

// Sum 16 per-lane values over the warp.  On return lane l holds in v[0] the warp total of value (l >> 1).
__device__ __forceinline__ void warp_transpose_reduce16(float (&v)[16], int lane) {
    {
        const bool up = lane & 16;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float send = up ? v[k] : v[k + 8];
            const float keep = up ? v[k + 8] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const float send = up ? v[k] : v[k + 4];
            const float keep = up ? v[k + 4] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const float send = up ? v[k] : v[k + 2];
            const float keep = up ? v[k + 2] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
    }
    {
        const bool up = lane & 2;
        const float send = up ? v[0] : v[1];
        const float keep = up ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

constexpr int BW_UNROLL = 4;  // records whose alpha is evaluated together in the backward

template <int C, int WARPS>
__global__ void __launch_bounds__(32 * WARPS)
blend_backward_kernel(const uint2* __restrict__ ranges, const float4* __restrict__ inst_conic,
                      const float4* __restrict__ inst_geom, const float4* __restrict__ inst_feat, int W, int H,
                      int gx, const float* __restrict__ bg, const float* __restrict__ final_Ts,
                      const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                      float* __restrict__ gacc) {
    constexpr int FV = BlendSmem<C>::FV;
    constexpr int NV = 6 + C;  // live gradient terms per (pixel, Gaussian)
    constexpr int NT = 32 * WARPS, PARTS = (BL_THREADS / 32) / WARPS;
    __shared__ __align__(128) BlendSmem<C> sm;
    __shared__ __align__(16) float s_acc[BL_BATCH * 16];
    __shared__ uint8_t s_list[WARPS][BL_BATCH];
    __shared__ int s_max_contrib;

    const int tile = blockIdx.x / PARTS;
    const int tx = tile % gx, ty = tile / gx;
    const int lane = threadIdx.x & 31;
    const int strip = (blockIdx.x % PARTS) * WARPS + (threadIdx.x >> 5);
    const int px = tx * TILE_X + (lane & (TILE_X - 1));
    const int py = ty * TILE_Y + 2 * strip + (lane >> 4);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix_id = (size_t)W * py + px;
    const float strip_x0 = (float)(tx * TILE_X);
    const float strip_yc = (float)(ty * TILE_Y + 2 * strip) + 0.5f;
    const uint2 range = ranges[tile];

    const float T_final = inside ? final_Ts[pix_id] : 0.f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;

    if (threadIdx.x == 0) {
        s_max_contrib = 0;
#pragma unroll
        for (int q = 0; q < BL_STAGES; q++) mbar_init(&sm.bar[q], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const int warp_max = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0 && warp_max > 0) atomicMax(&s_max_contrib, warp_max);
    __syncthreads();
    // Entries at list index >= len are skipped by every pixel of the tile (backward.cu:489-490).
    const int len = s_max_contrib;
    const int nb = (len + BL_BATCH - 1) / BL_BATCH;
    // batch k covers list indices [hi_k - cnt_k, hi_k), hi_k = len - k*BATCH, walked back to front
    if (threadIdx.x == 0) {
        for (int q = 0; q < BL_STAGES && q < nb; q++) {
            const int hi = len - q * BL_BATCH, cnt = min(BL_BATCH, hi);
            issue_batch<C>(sm, q, inst_conic, inst_geom, inst_feat, range.x + (uint32_t)(hi - cnt), cnt);
        }
    }

    float accum_rec[C], last_color[C], dL_dpixel[C];
    float bg_dot_dpixel = 0.f;
#pragma unroll
    for (int ch = 0; ch < C; ch++) {
        accum_rec[ch] = 0.f;
        last_color[ch] = 0.f;
        dL_dpixel[ch] = inside ? __ldg(dL_dpixels + (size_t)ch * H * W + pix_id) : 0.f;
        bg_dot_dpixel += __ldg(bg + ch) * dL_dpixel[ch];
    }
    float last_alpha = 0.f;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;  // backward.cu:460-461

    for (int b = 0; b < nb; b++) {
        const int s = b & (BL_STAGES - 1);
        const int hi = len - b * BL_BATCH;
        const int cnt = min(BL_BATCH, hi);
        const int lo = hi - cnt;  // list index of record 0 of this batch
        // clear the batch accumulator (16 floats per instance)
        {
            float4* a4 = reinterpret_cast<float4*>(s_acc);
            for (int i = threadIdx.x; i < cnt * 4; i += NT) a4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        mbar_wait(&sm.bar[s], (uint32_t)((b / BL_STAGES) & 1));
        __syncthreads();

        // records at list index >= warp_max are skipped by the whole warp; so are records whose ellipse
        // cannot reach this warp's strip (see cull_batch)
        uint8_t* list = s_list[threadIdx.x >> 5];
        const int n = cull_batch<C>(sm, s, min(cnt, warp_max - lo), lane, strip_x0, strip_yc, list);
#pragma unroll 1
        for (int t = n; t > 0; t -= BW_UNROLL) {
            {
                int jj[BW_UNROLL];
#pragma unroll
                for (int u = 0; u < BW_UNROLL; u++) jj[u] = (t - 1 - u >= 0) ? (int)list[t - 1 - u] : -1;  // back to front
                // (1) branch-free alpha for the chunk
                float al[BW_UNROLL], Gs[BW_UNROLL];
                uint32_t act = 0;
#pragma unroll
                for (int u = 0; u < BW_UNROLL; u++) {
                    const int j = max(jj[u], 0);
                    const float4 co = sm.conic[s][j];
                    const float4 ge = sm.geom[s][j];
                    const float dx = __fsub_rn(ge.x, pxf), dy = __fsub_rn(ge.y, pyf);
                    const float power = eval_power(co, dx, dy);
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, __fmul_rn(co.w, G));
                    al[u] = alpha;
                    Gs[u] = G;
                    // the reference's `contributor` after its decrement is lo + j (backward.cu:486-490)
                    if (jj[u] >= 0 && (lo + j) < last_contributor && !(power > 0.0f) && !(alpha < 1.0f / 255.0f))
                        act |= 1u << u;
                }
                // (2) ordered part + warp reduction, one record at a time
#pragma unroll
                for (int u = 0; u < BW_UNROLL; u++) {
                    const bool active = act & (1u << u);
                    const uint32_t active_lanes = __ballot_sync(0xffffffffu, active);
                    if (active_lanes == 0) continue;
                    const int j = jj[u];
                    float v[16];
#pragma unroll
                    for (int q = 0; q < 16; q++) v[q] = 0.f;
                    if (active) {
                        const float4 co = sm.conic[s][j];
                        const float4 ge = sm.geom[s][j];
                        const float dx = __fsub_rn(ge.x, pxf), dy = __fsub_rn(ge.y, pyf);
                        const float alpha = al[u], G = Gs[u];
                        const float inv_1ma = 1.f / (1.f - alpha);  // shared by T / (1 - alpha) and the background term
                        T = T * inv_1ma;
                        const float dchannel_dcolor = alpha * T;
                        const float4 f0 = sm.feat[s][j * FV];
                        const float4 f1 = sm.feat[s][j * FV + FV - 1];
                        float col[C];
                        col[0] = f0.x; col[1] = f0.y; col[2] = f0.z;
                        if (C == 6) { col[3 % C] = f0.w; col[4 % C] = f1.x; col[5 % C] = f1.y; }
                        float dL_dalpha = 0.f;
#pragma unroll
                        for (int ch = 0; ch < C; ch++) {
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = col[ch];
                            dL_dalpha += (col[ch] - accum_rec[ch]) * dL_dpixel[ch];
                            v[6 + ch] = dchannel_dcolor * dL_dpixel[ch];
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;  // backward.cu:531-534
                        const float dL_dG = co.w * dL_dalpha;                     // straight-through clamp (:538)
                        const float gdx = G * dx, gdy = G * dy;
                        const float dG_ddelx = -gdx * co.x - gdy * co.y;
                        const float dG_ddely = -gdy * co.z - gdx * co.y;
                        v[0] = -0.5f * gdx * dx * dL_dG;
                        v[1] = -0.5f * gdx * dy * dL_dG;
                        v[2] = -0.5f * gdy * dy * dL_dG;
                        v[3] = G * dL_dalpha;
                        v[4] = dL_dG * dG_ddelx * ddelx_dx;
                        v[5] = dL_dG * dG_ddely * ddely_dy;
                    }
                    warp_transpose_reduce16(v, lane);
                    if (!(lane & 1) && (lane >> 1) < NV) atomicAdd(&s_acc[j * 16 + (lane >> 1)], v[0]);
                }
            }
        }
        __syncthreads();
        // flush: two threads per instance; 3 (C=6) or 2.25 (C=3) vector reductions per touched instance
        for (int i = threadIdx.x; i < 2 * cnt; i += NT) {
            const int inst = i >> 1, half = i & 1;
            {
                const uint32_t id = __float_as_uint(sm.geom[s][inst].w);
                float* dst = gacc + (size_t)id * GACC_STRIDE;
                const float4* a4 = reinterpret_cast<const float4*>(s_acc + inst * 16);
                if (half == 0) {
                    const float4 x = a4[0], y = a4[1];
                    if (x.x != 0.f || x.y != 0.f || x.z != 0.f || x.w != 0.f) red_add_v4(dst, x.x, x.y, x.z, x.w);
                    if (y.x != 0.f || y.y != 0.f || y.z != 0.f || y.w != 0.f) red_add_v4(dst + 4, y.x, y.y, y.z, y.w);
                } else {
                    const float4 z = a4[2];
                    if (z.x != 0.f || z.y != 0.f || z.z != 0.f || z.w != 0.f) red_add_v4(dst + 8, z.x, z.y, z.z, z.w);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && b + BL_STAGES < nb) {
            const int nhi = len - (b + BL_STAGES) * BL_BATCH, ncnt = min(BL_BATCH, nhi);
            issue_batch<C>(sm, s, inst_conic, inst_geom, inst_feat, range.x + (uint32_t)(nhi - ncnt), ncnt);
        }
    }
}

}  // namespace

// Warps per CTA: whole tiles (8 warps) when there are plenty of tiles, finer splits for small images.
static int pick_warps(int tiles) {
    if (tiles >= 3 * NUM_SMS) return 8;
    if (tiles * 2 >= 3 * NUM_SMS) return 4;
    if (tiles * 4 >= 3 * NUM_SMS) return 2;
    return 1;
}

template <int C>
static void launch_fwd_c(int warps, int tiles, cudaStream_t st, const uint2* ranges, const BinningState& b, int W, int H,
                         int gx, const float* bg, float* out_color, float* out_depth, float* out_opacity,
                         float* final_T, uint32_t* n_contrib) {
#define ASB_FWD(WP)                                                                                              \
    blend_forward_kernel<C, WP><<<tiles*(8 / WP), 32 * WP, 0, st>>>(ranges, b.inst_conic, b.inst_geom, b.inst_feat, W, \
                                                                    H, gx, bg, out_color, out_depth, out_opacity,    \
                                                                    final_T, n_contrib)
    if (warps == 8) ASB_FWD(8);
    else if (warps == 4) ASB_FWD(4);
    else if (warps == 2) ASB_FWD(2);
    else ASB_FWD(1);
#undef ASB_FWD
}

template <int C>
static void launch_bwd_c(int warps, int tiles, cudaStream_t st, const uint2* ranges, const BinningState& b, int W, int H,
                         int gx, const float* bg, const float* final_T, const uint32_t* n_contrib,
                         const float* dL_dpix, float* gacc) {
#define ASB_BWD(WP)                                                                                               \
    blend_backward_kernel<C, WP><<<tiles*(8 / WP), 32 * WP, 0, st>>>(ranges, b.inst_conic, b.inst_geom, b.inst_feat, W, \
                                                                     H, gx, bg, final_T, n_contrib, dL_dpix, gacc)
    if (warps == 8) ASB_BWD(8);
    else if (warps == 4) ASB_BWD(4);
    else if (warps == 2) ASB_BWD(2);
    else ASB_BWD(1);
#undef ASB_BWD
}

void launch_blend_forward(int channels, int W, int H, const uint2* ranges, const BinningState& b, const float* bg,
                          float* out_color, float* out_depth, float* out_opacity, float* final_T,
                          uint32_t* n_contrib, cudaStream_t st) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;
    if (tiles <= 0) return;
    const int warps = pick_warps(tiles);
    if (channels == 3)
        launch_fwd_c<3>(warps, tiles, st, ranges, b, W, H, gx, bg, out_color, out_depth, out_opacity, final_T, n_contrib);
    else
        launch_fwd_c<6>(warps, tiles, st, ranges, b, W, H, gx, bg, out_color, out_depth, out_opacity, final_T, n_contrib);
}

void launch_blend_backward(int channels, int W, int H, const uint2* ranges, const BinningState& b, const float* bg,
                           const float* final_T, const uint32_t* n_contrib, const float* dL_dpix, float* gacc,
                           cudaStream_t st) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;
    if (tiles <= 0) return;
    const int warps = pick_warps(tiles);
    if (channels == 3)
        launch_bwd_c<3>(warps, tiles, st, ranges, b, W, H, gx, bg, final_T, n_contrib, dL_dpix, gacc);
    else
        launch_bwd_c<6>(warps, tiles, st, ranges, b, W, H, gx, bg, final_T, n_contrib, dL_dpix, gacc);
}

}  // namespace asb
