// blend.cu -- per-tile alpha blending, forward and backward (sm_100a).
//
//   blend_forward_kernel<C>   replaces renderCUDA<3> (forward)   dgr/cuda_rasterizer/forward.cu:261-397
//   blend_backward_kernel<C>  replaces renderCUDA<3> (backward)  dgr/cuda_rasterizer/backward.cu:399-557
//
// One CTA per 16x16 tile, one thread per pixel (same as the reference).  What differs:
//   * the tile's instances were gathered into contiguous per-instance records in sorted order
//     (ranges_gather_kernel / frame_tile_sort_gather_kernel), so a batch is THREE contiguous slabs that one
//     thread pulls into shared memory with cp.async.bulk (TMA engine, SASS UBLKCP) on an mbarrier,
//     multi-buffered: the copy of the next batches overlaps the blending of this one.  The warps of a CTA are not
//     held together by a block barrier per batch: the last warp to leave a stage refills it;
//   * C = 6 blends [rgb | z,1,z^2] in one sweep (the reference makes two full passes);
//   * a warp owns a 16x2-pixel strip.  Per batch it keeps only the records whose ellipse of possible contribution
//     reaches the strip, with the exact set of the strip's pixels inside that ellipse (cull_batch); a bit-matrix
//     transpose hands every lane the records that can reach ITS pixel, and each lane walks its own list
//     (the reference drags all 256 threads of a tile through every record of the tile);
//   * backward: two phases per group of 16 records -- pixel lanes replay the recurrences and stage two scalars per
//     (record, pixel); record lanes sum them over their pixels serially and commit 12 gradient terms with two
//     vector reductions (REDG.ADD.F32x4) -- versus 9 scalar global atomics per (pixel, Gaussian) in the reference
//     (backward.cu:523-554);
//   * backward skips the tail of the list that no pixel of the warp/tile reached in the forward.
// Per-pixel arithmetic, skip rules and termination are the reference's (SURVEY 8a' T10-T18) and are
// spelled with rounding-explicit intrinsics so alpha decisions match the reference build bit for bit.
#include <math_constants.h>
#include <cstdlib>

#include <atomic>

#include "common.cuh"

namespace asb {

namespace {

constexpr int BL_THREADS = TILE_PIX;  // 256
constexpr int BL_BATCH = 128;
constexpr int BL_STAGES = 4;  // bulk-copy pipeline depth (power of two): batches b..b+3 are in flight
#ifndef ASB_FW_UNROLL
#define ASB_FW_UNROLL 4
#endif
constexpr int FW_UNROLL = ASB_FW_UNROLL;  // records a lane evaluates together in the forward (ILP)

__device__ __forceinline__ float eval_power(const float4& co, float dx, float dy) {
    // forward.cu:342: -0.5f*(A*dx*dx + C*dy*dy) - B*dx*dy, contracted as nvcc contracts it
    const float q = __fmaf_rn(__fmul_rn(co.x, dx), dx, __fmul_rn(__fmul_rn(co.z, dy), dy));
    return __fmaf_rn(q, -0.5f, -__fmul_rn(__fmul_rn(co.y, dx), dy));
}

template <int C, int STAGES = BL_STAGES, int BATCH_ = BL_BATCH>
struct BlendSmem {
    static constexpr int FV = (C == 3) ? 1 : 2;
    static constexpr int BATCH = BATCH_;
    float4 conic[STAGES][BATCH_];
    float4 geom[STAGES][BATCH_];
    float4 feat[STAGES][BATCH_ * FV];
    uint64_t bar[STAGES];
};

template <class SM>
__device__ __forceinline__ void issue_batch(SM& sm, int stage, const float4* __restrict__ inst_conic,
                                            const float4* __restrict__ inst_geom,
                                            const float4* __restrict__ inst_feat, uint32_t start, int cnt) {
    constexpr int FV = SM::FV;
    const uint32_t bytes = (uint32_t)cnt * 16u;
    mbar_expect_tx(&sm.bar[stage], bytes * (2 + FV));
    bulk_g2s(sm.conic[stage], inst_conic + start, bytes, &sm.bar[stage]);
    bulk_g2s(sm.geom[stage], inst_geom + start, bytes, &sm.bar[stage]);
    bulk_g2s(sm.feat[stage], inst_feat + (size_t)start * FV, bytes * FV, &sm.bar[stage]);
}

// ---- per-warp culling ---------------------------------------------------------------------------
// A pair (pixel, Gaussian) can only pass the skip rules (power <= 0, alpha >= 1/255; forward.cu:343-353)
// inside the ellipse  A dx^2 + 2B dx dy + C dy^2 <= L,  L = 2 ln(255 * opacity); ranges_gather stores L (with
// margin) per instance.  For one pixel row (dy fixed) that is an interval of columns: a warp computes, per record,
// the columns of its two rows inside the ellipse, skips the record when there are none, and otherwise keeps the
// 32-bit lane mask.  Pairs outside the mask would have been skipped by the rules anyway, so images, n_contrib and
// gradients are unchanged.
template <class SM>
__device__ __forceinline__ float inst_level(const SM& sm, int s, int j) {
    constexpr int FV = SM::FV;
    const float4 f = sm.feat[s][j * FV + FV - 1];
    return FV == 1 ? f.w : f.z;
}

// single-instruction reciprocal / square root (MUFU.RCP, MUFU.SQRT; ~1 ulp).  Used where the argument's range is known
// (1 - alpha in [0.01, 1] in the backward) or where the result only feeds the conservative culling intervals, whose
// 0.02-pixel margin is orders of magnitude above the approximation error.
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float sqrt_approx(float x) {   // NaN for x < 0, like sqrtf
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// columns (bit k = column k of the tile) with |column - c| <= hw
__device__ __forceinline__ uint32_t column_bits(float c, float hw) {
    const int lo = max(0, __float2int_ru(c - hw));             // -inf saturates to INT_MIN
    const int hi = min(TILE_X - 1, __float2int_rd(c + hw));    // +inf saturates to INT_MAX
    return (hi >= lo) ? ((2u << hi) - (1u << lo)) : 0u;
}

// Compacts the batch's records that can touch this warp's pixel strip into the warp's private list
// (ascending record order) and returns how many there are.  For every kept record it also stores the
// LANE MASK of the strip's pixels inside the record's ellipse (bit l = pixel of lane l: column l & 15,
// row l >> 4).  Two steps: a cheap row test over all records (does the ellipse reach either row at all?),
// then the column intervals for the survivors only.
template <class SM>
__device__ __forceinline__ int cull_batch(const SM& sm, int s, int limit, int lane, float strip_x0,
                                          float strip_y0, uint8_t* list, uint32_t* masks) {
    int n = 0;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < SM::BATCH / 32; k++) {
        const int j = lane + 32 * k;
        bool ok = false;
        if (j < limit) {
            const float4 co = sm.conic[s][j];
            const float gy = sm.geom[s][j].y;
            const float L = inst_level(sm, s, j);
            // row r is reached iff D_r = A L - det dy_r^2 >= 0; anything not finite is kept (never cull the unknown)
            const float det = co.x * co.z - co.y * co.y, al = co.x * L;
            const float dy0 = gy - strip_y0, dy1 = dy0 - 1.0f;
            const float D0 = al - det * dy0 * dy0, D1 = al - det * dy1 * dy1;
            ok = (L >= 0.f) && !(D0 < 0.f && D1 < 0.f);
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, ok);
        if (ok) list[n + __popc(bal & lt)] = (uint8_t)j;
        n += __popc(bal);
    }
    __syncwarp();
    // column intervals of the survivors; records whose interval misses the tile's columns drop out here
    int n2 = 0;
#pragma unroll 1
    for (int i0 = 0; i0 < n; i0 += 32) {
        uint32_t m = 0;
        int j = 0;
        if (i0 + lane < n) {
            j = list[i0 + lane];
            const float4 co = sm.conic[s][j];
            const float4 ge = sm.geom[s][j];
            const float L = inst_level(sm, s, j);
            const float det = co.x * co.z - co.y * co.y;
            const float inv_a = rcp_approx(co.x);
            const float rel = ge.x - strip_x0;
            const float dy0 = ge.y - strip_y0, dy1 = dy0 - 1.0f;
            // row r: A dx^2 + 2 B dy dx + (C dy^2 - L) <= 0  =>  pixel column in  rel + (B dy -+ sqrt(D)) / A
            const float D0 = co.x * L - det * dy0 * dy0, D1 = co.x * L - det * dy1 * dy1;
            const float c0 = rel + co.y * dy0 * inv_a, c1 = rel + co.y * dy1 * inv_a;
            const float h0 = sqrt_approx(D0) * inv_a + 0.02f, h1 = sqrt_approx(D1) * inv_a + 0.02f;  // NaN when D < 0
            uint32_t b0 = (D0 >= 0.f) ? column_bits(c0, h0) : 0u;
            uint32_t b1 = (D1 >= 0.f) ? column_bits(c1, h1) : 0u;
            if (!(L < CUDART_INF_F) || !(c0 == c0) || !(D0 == D0) || !(h0 == h0 || D0 < 0.f)) b0 = 0xffffu;
            if (!(L < CUDART_INF_F) || !(c1 == c1) || !(D1 == D1) || !(h1 == h1 || D1 < 0.f)) b1 = 0xffffu;
            m = b0 | (b1 << 16);
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, m != 0u);  // also orders this round's reads before its writes
        if (m != 0u) {
            const int slot = n2 + __popc(bal & lt);  // slot <= i0 + lane: in-place compaction never overtakes the reads
            list[slot] = (uint8_t)j;
            masks[slot] = m;
        }
        n2 += __popc(bal);
        __syncwarp();
    }
    return n2;
}

// 32x32 bit-matrix transpose across the warp: lane r passes row r, lane p receives column p
// (bit r of the result = bit p of lane r's input).  Five butterfly stages.
__device__ __forceinline__ uint32_t warp_bit_transpose(uint32_t x, int lane) {
#pragma unroll
    for (int st = 0; st < 5; st++) {
        const int j = 16 >> st;
        const uint32_t m = (st == 0) ? 0x0000ffffu : (st == 1) ? 0x00ff00ffu : (st == 2) ? 0x0f0f0f0fu
                         : (st == 3) ? 0x33333333u : 0x55555555u;
        const uint32_t t = __shfl_xor_sync(0xffffffffu, x, j);
        x = (lane & j) ? (((t >> j) & m) | (x & ~m)) : ((x & m) | ((t & m) << j));
    }
    return x;
}

// WARPS warps per CTA: a CTA blends 2*WARPS pixel rows of one tile, so 8/WARPS CTAs share a tile (each
// stages the tile's records for itself).  Small images have few tiles; splitting them keeps all SMs busy
// and spreads a heavy tile over several SMs.  Per-pixel results do not depend on WARPS.
template <int C, int WARPS, bool FRAME>
__global__ void __launch_bounds__(32 * WARPS)
blend_forward_kernel(const uint2* __restrict__ ranges, const float4* __restrict__ inst_conic,
                     const float4* __restrict__ inst_geom, const float4* __restrict__ inst_feat, int W, int H,
                     int gx, const float* __restrict__ bg, float* __restrict__ out_color,
                     float* __restrict__ out_depth, float* __restrict__ out_opacity, float* __restrict__ final_T,
                     uint32_t* __restrict__ n_contrib, const uint32_t* __restrict__ tile_order) {
    constexpr int FV = BlendSmem<C>::FV;
    constexpr int NT = 32 * WARPS, PARTS = (BL_THREADS / 32) / WARPS;
    __shared__ __align__(128) BlendSmem<C> sm;
    __shared__ uint8_t s_list[WARPS][BL_BATCH];
    __shared__ uint32_t s_mask[WARPS][BL_BATCH];
    const int tile = tile_order ? (int)tile_order[blockIdx.x / PARTS] : (int)(blockIdx.x / PARTS);
    const int tx = tile % gx, ty = tile / gx;
    const int lane = threadIdx.x & 31;
    const int strip = (blockIdx.x % PARTS) * WARPS + (threadIdx.x >> 5);  // which 16x2 strip of the tile
    const int px = tx * TILE_X + (lane & (TILE_X - 1));
    const int py = ty * TILE_Y + 2 * strip + (lane >> 4);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    // this warp's pixels: columns [tile x0, +15], rows 2*strip and 2*strip+1
    const float strip_x0 = (float)(tx * TILE_X);
    const float strip_y0 = (float)(ty * TILE_Y + 2 * strip);
    const uint2 range = ranges[tile];
    const int len = (int)(range.y - range.x);
    const int nb = (len + BL_BATCH - 1) / BL_BATCH;

    // Warps are not held together by a CTA-wide barrier per batch: a warp that is through with a stage moves on to the
    // next (already staged) batch, and the LAST warp to leave a stage refills it -- or, once every warp has reported
    // all of its pixels finished (forward.cu:315-317), sets s_stop and completes the stage's barrier without data so
    // that nobody waits for a batch that will never come.
    __shared__ int s_released[BL_STAGES];   // warps that are through with the stage's current batch
    __shared__ int s_done_warps;            // warps whose 32 pixels are all finished
    __shared__ volatile int s_stop, s_last_issued;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int q = 0; q < BL_STAGES; q++) {
            mbar_init(&sm.bar[q], 1);
            s_released[q] = 0;
        }
        s_done_warps = 0;
        s_stop = 0;
        s_last_issued = min(BL_STAGES, nb) - 1;
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int q = 0; q < BL_STAGES && q < nb; q++)
            issue_batch(sm, q, inst_conic, inst_geom, inst_feat, range.x + q * BL_BATCH,
                           min(BL_BATCH, len - q * BL_BATCH));
    }

    bool done = !inside;
    bool reported = false;   // this warp is counted in s_done_warps
    float T = 1.0f;
    float D = 15.0f;  // median depth default (forward.cu:309)
    uint32_t last_contributor = 0;
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ch++) acc[ch] = 0.f;

    int waited = 0;   // batches whose barrier this warp has waited on
    for (int b = 0; b < nb; b++) {
        const int s = b & (BL_STAGES - 1);
        mbar_wait(&sm.bar[s], (uint32_t)((b / BL_STAGES) & 1));
        waited = b + 1;
        if (__shfl_sync(0xffffffffu, (int)s_stop, 0)) break;   // every pixel of the tile is finished (one read per warp: the decision is uniform)
        const int cnt = min(BL_BATCH, len - b * BL_BATCH);
        // The per-pixel loop of forward.cu:332-381, restructured so that every lane works on a record that can
        // actually reach ITS pixel:
        // (0) the warp drops the records whose box cannot reach its strip and keeps, per surviving record, the
        //     mask of the strip's pixels inside the box (lane = record);
        // (1) 32 surviving records at a time, the 32x32 bit matrix (record x pixel) is transposed with five
        //     shuffles, so each lane holds the set of records that can touch its own pixel;
        // (2) each lane walks ITS OWN set in list order with the reference's arithmetic and skip rules.  Lanes are
        //     on different records at the same time; nothing in the forward needs them to agree.
        bool warp_done = __all_sync(0xffffffffu, done);
        if (!warp_done) {
            const int w = threadIdx.x >> 5;
            uint8_t* list = s_list[w];
            uint32_t* masks = s_mask[w];
            const int n = cull_batch(sm, s, cnt, lane, strip_x0, strip_y0, list, masks);
#pragma unroll 1
            for (int g0 = 0; g0 < n && !warp_done; g0 += 32) {
                const bool have = g0 + lane < n;
                const int jmine = have ? (int)list[g0 + lane] : 0;
                uint32_t hits = warp_bit_transpose(have ? masks[g0 + lane] : 0u, lane);
                if (done) hits = 0u;
#pragma unroll 1
                while (__any_sync(0xffffffffu, hits != 0u)) {
                    // pop up to FW_UNROLL of this lane's records (ascending = list order); their alphas are
                    // independent chains, so the exp latency of one hides behind the others
                    int jj[FW_UNROLL];
                    uint32_t valid = 0u;
#pragma unroll
                    for (int u = 0; u < FW_UNROLL; u++) {
                        const bool has = hits != 0u;
                        const int r = has ? (__ffs((int)hits) - 1) : 0;
                        hits &= hits - 1u;  // 0 stays 0
                        jj[u] = __shfl_sync(0xffffffffu, jmine, r);
                        if (has) valid |= 1u << u;
                    }
                    float al[FW_UNROLL];
                    uint32_t hit = 0u;
#pragma unroll
                    for (int u = 0; u < FW_UNROLL; u++) {
                        const float4 co = sm.conic[s][jj[u]];
                        const float4 ge = sm.geom[s][jj[u]];
                        const float dx = __fsub_rn(ge.x, pxf), dy = __fsub_rn(ge.y, pyf);
                        const float power = eval_power(co, dx, dy);
                        const float alpha = fminf(0.99f, __fmul_rn(co.w, expf(power)));
                        al[u] = alpha;
                        // skip rules of forward.cu:343-353 (written so that NaN behaves as in the reference)
                        if (((valid >> u) & 1u) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f)) hit |= 1u << u;
                    }
                    // order-dependent part: the only serial chain is T -> test_T -> T (a branch per record measured
                    // faster than selects: the selects became a multiply-add into a temporary plus a move, per channel)
#pragma unroll
                    for (int u = 0; u < FW_UNROLL; u++) {
                        const int j = jj[u];
                        const float alpha = al[u];
                        const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
                        const bool h = ((hit >> u) & 1u) && !done;
                        const bool stop = h && (test_T < 0.0001f);  // forward.cu:355-359: stop BEFORE blending it
                        const bool blend = h && !stop;
                        done = done || stop;
                        if (blend) {
                            const float4 f0 = sm.feat[s][j * FV];
                            acc[0] = __fmaf_rn(__fmul_rn(f0.x, alpha), T, acc[0]);
                            acc[1] = __fmaf_rn(__fmul_rn(f0.y, alpha), T, acc[1]);
                            acc[2] = __fmaf_rn(__fmul_rn(f0.z, alpha), T, acc[2]);
                            if (C == 6) {
                                acc[3 % C] = __fmaf_rn(__fmul_rn(f0.w, alpha), T, acc[3 % C]);
                                if (FRAME) {
                                    // fused frame path: the second triple is (z, 1, z^2) by construction (fused.cu
                                    // gathers): 1 * alpha = alpha, z^2 is the same product the gather stored
                                    acc[4 % C] = __fmaf_rn(alpha, T, acc[4 % C]);
                                    acc[5 % C] = __fmaf_rn(__fmul_rn(__fmul_rn(f0.w, f0.w), alpha), T, acc[5 % C]);
                                } else {
                                    const float4 f1 = sm.feat[s][j * FV + FV - 1];
                                    acc[4 % C] = __fmaf_rn(__fmul_rn(f1.x, alpha), T, acc[4 % C]);
                                    acc[5 % C] = __fmaf_rn(__fmul_rn(f1.y, alpha), T, acc[5 % C]);
                                }
                            }
                            if (T > 0.5f && test_T < 0.5f) D = sm.geom[s][j].z;                  // forward.cu:370-374
                            last_contributor = (uint32_t)(b * BL_BATCH + j + 1);                 // :335,380
                            T = test_T;
                        }
                    }
                    if (done) hits = 0u;
                }
                warp_done = __all_sync(0xffffffffu, done);
            }
            __syncwarp();  // the list is rewritten by the next batch
        }
        __syncwarp();
        if (lane == 0) {
            if (warp_done && !reported) atomicAdd(&s_done_warps, 1);   // before the release: the last one out sees it
            __threadfence_block();
            if (atomicAdd(&s_released[s], 1) == WARPS - 1) {
                s_released[s] = 0;
                if (b + BL_STAGES < nb) {
                    if (*(volatile int*)&s_done_warps == WARPS) {
                        s_stop = 1;
                        __threadfence_block();
                        mbar_arrive(&sm.bar[s]);
                    } else {
                        const int nstart = (b + BL_STAGES) * BL_BATCH;
                        issue_batch(sm, s, inst_conic, inst_geom, inst_feat, range.x + nstart, min(BL_BATCH, len - nstart));
                        s_last_issued = b + BL_STAGES;
                    }
                }
            }
        }
        reported = reported || warp_done;
    }
    // copies that were issued but never consumed must land before this CTA's shared memory is released (after an
    // early stop the batches up to s_last_issued are still in flight; s_last_issued is final once s_stop is visible)
    for (int q = waited; q <= s_last_issued; q++)
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

// ---- forward, split along the list ("chunked") ------------------------------------------------------
// A low-resolution wide-angle candidate view puts 10^4-10^5 records on each of a few dozen tiles; one warp per strip
// then walks them serially (1.4 of a 1.9 ms view, with 640 warps on the whole GPU).  Here a tile's list is cut into
// chunks of BL_CHUNK records and every chunk gets its own CTA, in two passes:
//   pass A  transmittance only: T_k = product of (1 - alpha) over the chunk's accepted records, from T = 1
//           (a pixel stops once its chunk-local T drops below 1e-4: everything behind it is dead anyway);
//   pass B  the reference's loop, unchanged, from the TRUE transmittance at the chunk's start, prefix_k = T_0 ... T_{k-1}:
//           the 1e-4 stop rule and the T = 0.5 median-depth rule fire inside whichever chunk they belong to; a chunk
//           whose prefix is already below 1e-4 does nothing;
//   combine sums the chunks' colour contributions in list order, takes the median depth from the chunk that set it and
//           the final T from the last live chunk.
// Differences from the single walk: T at a chunk start is a product of chunk products (rounding ~1e-7), and
// n_contrib is not produced -- this path is for renders that nobody differentiates (GetOpacity sweeps).
constexpr int BL_CHUNK = 1024;       // records per chunk (8 staged batches)
constexpr int CHUNK_FIELDS = 9;      // per (chunk, pixel): T_k | colour sums [6] | median depth or -1 | T after the chunk

// per tile: number of chunks and the index of its first chunk; per chunk: its tile.  One block.
__global__ void __launch_bounds__(1024)
blend_chunk_table_kernel(const uint2* __restrict__ ranges, int tiles, int max_chunks, int* __restrict__ chunk_base,
                         int* __restrict__ chunk_tile) {
    __shared__ int s_scan[1024];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < tiles; base += 1024) {
        const int t = base + threadIdx.x;
        int n = 0;
        if (t < tiles) {
            const uint2 r = ranges[t];
            n = (int)((r.y - r.x + BL_CHUNK - 1) / BL_CHUNK);
        }
        s_scan[threadIdx.x] = n;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {
            const int v = (threadIdx.x >= off) ? s_scan[threadIdx.x - off] : 0;
            __syncthreads();
            s_scan[threadIdx.x] += v;
            __syncthreads();
        }
        const int first = s_carry + s_scan[threadIdx.x] - n;
        if (t < tiles) {
            chunk_base[t] = first;
            for (int k = 0; k < n && first + k < max_chunks; k++) chunk_tile[first + k] = t;
        }
        __syncthreads();
        if (threadIdx.x == 1023) s_carry += s_scan[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) chunk_base[tiles] = min(s_carry, max_chunks);
}

template <int C, int PASS>
__global__ void __launch_bounds__(BL_THREADS)
blend_forward_chunk_kernel(const uint2* __restrict__ ranges, const float4* __restrict__ inst_conic,
                           const float4* __restrict__ inst_geom, const float4* __restrict__ inst_feat, int W, int H,
                           int gx, int tiles, const int* __restrict__ chunk_base, const int* __restrict__ chunk_tile,
                           float* __restrict__ scratch) {
    constexpr int FV = BlendSmem<C>::FV;
    const int chunk = blockIdx.x;
    if (chunk >= chunk_base[tiles]) return;
    __shared__ __align__(128) BlendSmem<C> sm;
    __shared__ uint8_t s_list[BL_THREADS / 32][BL_BATCH];
    __shared__ uint32_t s_mask[BL_THREADS / 32][BL_BATCH];
    const int tile = chunk_tile[chunk];
    const int k = chunk - chunk_base[tile];
    const int tx = tile % gx, ty = tile / gx;
    const int lane = threadIdx.x & 31, strip = threadIdx.x >> 5;
    const int px = tx * TILE_X + (lane & (TILE_X - 1));
    const int py = ty * TILE_Y + 2 * strip + (lane >> 4);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const float strip_x0 = (float)(tx * TILE_X), strip_y0 = (float)(ty * TILE_Y + 2 * strip);
    const uint2 tr = ranges[tile];
    const uint32_t begin = tr.x + (uint32_t)k * BL_CHUNK;
    const int len = (int)min((uint32_t)BL_CHUNK, tr.y - begin);
    const int nb = (len + BL_BATCH - 1) / BL_BATCH;
    const int pix = strip * 32 + lane;                                   // pixel slot of the tile
    float* mine = scratch + (size_t)chunk * CHUNK_FIELDS * TILE_PIX + pix;   // field f at mine[f * TILE_PIX]

    if (threadIdx.x == 0) {
#pragma unroll
        for (int q = 0; q < BL_STAGES; q++) mbar_init(&sm.bar[q], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int q = 0; q < BL_STAGES && q < nb; q++)
            issue_batch(sm, q, inst_conic, inst_geom, inst_feat, begin + q * BL_BATCH, min(BL_BATCH, len - q * BL_BATCH));

    float T = 1.0f;
    if (PASS == 2) {   // the true transmittance at the start of this chunk
        const float* tk = scratch + (size_t)chunk_base[tile] * CHUNK_FIELDS * TILE_PIX + pix;
        for (int j = 0; j < k; j++) T *= tk[(size_t)j * CHUNK_FIELDS * TILE_PIX];
    }
    bool done = !inside || (PASS == 2 && T < 0.0001f);
    float D = -1.0f;
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ch++) acc[ch] = 0.f;

    int consumed = 0;
    for (int b = 0; b < nb; b++) {
        const int s = b & (BL_STAGES - 1);
        mbar_wait(&sm.bar[s], (uint32_t)((b / BL_STAGES) & 1));
        const int cnt = min(BL_BATCH, len - b * BL_BATCH);
        bool warp_done = __all_sync(0xffffffffu, done);
        if (!warp_done) {
            uint8_t* list = s_list[strip];
            uint32_t* masks = s_mask[strip];
            const int n = cull_batch(sm, s, cnt, lane, strip_x0, strip_y0, list, masks);
#pragma unroll 1
            for (int g0 = 0; g0 < n && !warp_done; g0 += 32) {
                const bool have = g0 + lane < n;
                const int jmine = have ? (int)list[g0 + lane] : 0;
                uint32_t hits = warp_bit_transpose(have ? masks[g0 + lane] : 0u, lane);
                if (done) hits = 0u;
#pragma unroll 1
                while (__any_sync(0xffffffffu, hits != 0u)) {
                    int jj[FW_UNROLL];
                    uint32_t valid = 0u;
#pragma unroll
                    for (int u = 0; u < FW_UNROLL; u++) {
                        const bool has = hits != 0u;
                        const int r = has ? (__ffs((int)hits) - 1) : 0;
                        hits &= hits - 1u;
                        jj[u] = __shfl_sync(0xffffffffu, jmine, r);
                        if (has) valid |= 1u << u;
                    }
                    float al[FW_UNROLL];
                    uint32_t hit = 0u;
#pragma unroll
                    for (int u = 0; u < FW_UNROLL; u++) {
                        const float4 co = sm.conic[s][jj[u]];
                        const float4 ge = sm.geom[s][jj[u]];
                        const float dx = __fsub_rn(ge.x, pxf), dy = __fsub_rn(ge.y, pyf);
                        const float power = eval_power(co, dx, dy);
                        const float alpha = fminf(0.99f, __fmul_rn(co.w, expf(power)));
                        al[u] = alpha;
                        if (((valid >> u) & 1u) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f)) hit |= 1u << u;
                    }
#pragma unroll
                    for (int u = 0; u < FW_UNROLL; u++) {
                        const int j = jj[u];
                        const float alpha = al[u];
                        const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
                        const bool h = ((hit >> u) & 1u) && !done;
                        const bool stop = h && (test_T < 0.0001f);
                        if (PASS == 1) {
                            T = h ? test_T : T;       // pass A keeps the product (below 1e-4 = "dead from here on")
                            done = done || stop;
                        } else {
                            const bool blend = h && !stop;   // forward.cu:355-359: stop BEFORE blending it
                            done = done || stop;
                            if (blend) {   // as in blend_forward_kernel<C, WARPS, true>: this path only serves the frame entry
                                const float4 f0 = sm.feat[s][j * FV];
                                acc[0] = __fmaf_rn(__fmul_rn(f0.x, alpha), T, acc[0]);
                                acc[1] = __fmaf_rn(__fmul_rn(f0.y, alpha), T, acc[1]);
                                acc[2] = __fmaf_rn(__fmul_rn(f0.z, alpha), T, acc[2]);
                                if (C == 6) {   // the second triple is (z, 1, z^2) by construction
                                    acc[3 % C] = __fmaf_rn(__fmul_rn(f0.w, alpha), T, acc[3 % C]);
                                    acc[4 % C] = __fmaf_rn(alpha, T, acc[4 % C]);
                                    acc[5 % C] = __fmaf_rn(__fmul_rn(__fmul_rn(f0.w, f0.w), alpha), T, acc[5 % C]);
                                }
                                if (T > 0.5f && test_T < 0.5f) D = sm.geom[s][j].z;   // forward.cu:370-374
                                T = test_T;
                            }
                        }
                    }
                    if (done) hits = 0u;
                }
                warp_done = __all_sync(0xffffffffu, done);
            }
            __syncwarp();
        }
        consumed = b + 1;
        if (__syncthreads_count(done) == BL_THREADS) break;
        if (threadIdx.x == 0 && b + BL_STAGES < nb) {
            const int nstart = (b + BL_STAGES) * BL_BATCH;
            issue_batch(sm, s, inst_conic, inst_geom, inst_feat, begin + nstart, min(BL_BATCH, len - nstart));
        }
    }
    if (consumed >= 1)
        for (int q = consumed; q < nb && q < consumed - 1 + BL_STAGES; q++)
            mbar_wait(&sm.bar[q & (BL_STAGES - 1)], (uint32_t)((q / BL_STAGES) & 1));

    if (PASS == 1) {
        mine[0] = T;
    } else {
#pragma unroll
        for (int ch = 0; ch < C; ch++) mine[(1 + ch) * TILE_PIX] = acc[ch];
        mine[7 * TILE_PIX] = D;
        mine[8 * TILE_PIX] = T;
    }
}

// one CTA per tile: the chunks' results in list order -> the images
template <int C>
__global__ void __launch_bounds__(BL_THREADS)
blend_chunk_combine_kernel(int W, int H, int gx, const int* __restrict__ chunk_base, const float* __restrict__ scratch,
                           const float* __restrict__ bg, float* __restrict__ out_color, float* __restrict__ out_depth,
                           float* __restrict__ out_opacity, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib) {
    const int tile = blockIdx.x;
    const int tx = tile % gx, ty = tile / gx;
    const int lane = threadIdx.x & 31, strip = threadIdx.x >> 5;
    const int px = tx * TILE_X + (lane & (TILE_X - 1));
    const int py = ty * TILE_Y + 2 * strip + (lane >> 4);
    if (px >= W || py >= H) return;
    const int c0 = chunk_base[tile], c1 = chunk_base[tile + 1];
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ch++) acc[ch] = 0.f;
    float T = 1.0f, prefix = 1.0f, D = 15.0f;   // median depth default (forward.cu:309)
    bool d_set = false;
    for (int c = c0; c < c1 && !(prefix < 0.0001f); c++) {
        const float* f = scratch + (size_t)c * CHUNK_FIELDS * TILE_PIX + threadIdx.x;
#pragma unroll
        for (int ch = 0; ch < C; ch++) acc[ch] += f[(1 + ch) * TILE_PIX];
        const float d = f[7 * TILE_PIX];
        if (!d_set && d >= 0.f) { D = d; d_set = true; }
        T = f[8 * TILE_PIX];
        prefix *= f[0];
    }
    const size_t pix_id = (size_t)W * py + px;
    final_T[pix_id] = T;
    n_contrib[pix_id] = 0u;   // not produced on this path (no backward follows it)
#pragma unroll
    for (int ch = 0; ch < C; ch++) out_color[(size_t)ch * H * W + pix_id] = __fmaf_rn(T, __ldg(bg + ch), acc[ch]);
    out_depth[pix_id] = D;
    out_opacity[pix_id] = __fsub_rn(1.f, T);
}

// ---- backward -----------------------------------------------------------------------------------
// The reference walks every pixel back to front and issues 9 global atomics per (pixel, Gaussian)
// (backward.cu:486-554).  Here a warp handles its 16x2 strip in groups of BW_GROUP surviving records:
//   phase 1 (lane = pixel): each lane walks, back to front, only the records whose box contains ITS pixel
//     (same bit-matrix transpose as the forward), replays T / accum_rec / dL_dalpha with the reference's
//     arithmetic, and leaves two numbers per (record, pixel) in a per-warp staging tile:
//       q = G * dL_dalpha   and   w = alpha * T.
//     Every geometric term of backward.cu:505-554 is q times a polynomial in (dx, dy) and the record's own
//     conic/opacity, and the colour terms are w * dL_dpixel, so nothing else has to cross lanes.
//   phase 2 (lane = record x pixel row): each lane sums ITS record's q, q*dx, q*dx^2 and w*dL_dpixel over the
//     pixels of one row serially in registers -- no butterfly, every lane busy on its own record -- forms the
//     12 gradient terms, and adds them to the CTA's per-batch accumulator.
// The CTA flushes the accumulator once per (tile, instance) with 128-bit vector reductions (REDG.ADD.F32x4).
#ifndef ASB_BW_GROUP
#define ASB_BW_GROUP 16
#endif
constexpr int BW_GROUP = ASB_BW_GROUP;   // records per group (16 or 8): BW_GROUP x 32 staging slots per warp
constexpr int BW_SPLIT = 16 / BW_GROUP;  // phase-2 lanes per (record, pixel row): each sums 16 / BW_SPLIT pixels of the row
static_assert(BW_GROUP == 16 || BW_GROUP == 8, "phase 2 maps 32 lanes to BW_GROUP records x 2 rows x BW_SPLIT segments");
#ifndef ASB_BW_BATCH
#define ASB_BW_BATCH 128
#endif
constexpr int BW_BATCH = ASB_BW_BATCH;  // records per staged batch in the backward (64 gives 4 CTAs per SM but twice the barriers: measured slower)
#ifndef ASB_BW_STAGES
#define ASB_BW_STAGES 2
#endif
constexpr int BW_STAGES = ASB_BW_STAGES;   // bulk-copy pipeline depth of the backward (a batch takes long enough to hide one copy)
#ifndef ASB_BW_UNROLL
#define ASB_BW_UNROLL 2
#endif
constexpr int BW_UNROLL = ASB_BW_UNROLL;  // records a lane evaluates together in phase 1 (ILP)

// 16x16 bit-matrix transposes in both half-warps at once (lane r of a half passes row r, lane k receives column k).
__device__ __forceinline__ uint32_t half_warp_bit_transpose(uint32_t x, int lane) {
#pragma unroll
    for (int st = 0; st < 4; st++) {
        const int j = 8 >> st;
        const uint32_t m = (st == 0) ? 0x00ffu : (st == 1) ? 0x0f0fu : (st == 2) ? 0x3333u : 0x5555u;
        const uint32_t t = __shfl_xor_sync(0xffffffffu, x, j);
        x = (lane & j) ? (((t >> j) & m) | (x & ~m & 0xffffu)) : ((x & m) | ((t & m) << j));
    }
    return x;
}

// staging slot of (record r of the group, pixel lane p): rows are padded to 33 slots, so lanes reading the same
// column of different records (phase 2) hit distinct banks, lanes writing one record (phase 1) do too, and a lane
// walks its row with immediate offsets (an XOR swizzle cost three address instructions per pixel)
constexpr int STG_STRIDE = 33;
__device__ __forceinline__ int stage_slot(int r, int p) { return r * STG_STRIDE + p; }

// 32-bit shared-window address of the staging tile + raw vector accesses: the generic-pointer form made the compiler
// rebuild the window base (S2UR/ULEA chain) at every store of the inner loop
__device__ __forceinline__ void sts_float2(uint32_t addr, float x, float y) {
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ float2 lds_float2(uint32_t addr) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr) : "memory");
    return v;
}
// index of the highest set bit, -1 (all ones) for 0: one FLO instead of the clz / subtract / xor sequence
__device__ __forceinline__ int find_msb(uint32_t x) {
    int r;
    asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x));
    return r;
}
// Which of the 12 per-Gaussian terms the caller consumes (conic x,y,w | opacity | mean2D x,y | colour 0..5):
//   BWD_ALL       everything (rasterizer API)
//   BWD_MAPPING   fused frame path with Gaussian gradients: colour 4 is the constant 1 of [z,1,z^2], never read
//   BWD_TRACKING  fused frame path, pose only: conic, mean2D and the z / z^2 colour terms (T20)
enum BwdMode { BWD_ALL = 0, BWD_MAPPING = 1, BWD_TRACKING = 2 };
template <int C, int MODE>
__device__ __forceinline__ constexpr bool need_colour(int ch) {
    return MODE == BWD_ALL ? true : MODE == BWD_MAPPING ? (ch != 4) : (ch == 3 || ch == 5);
}

template <int C, int WARPS>
struct BwdLayout {
    using SM = BlendSmem<C, BW_STAGES, BW_BATCH>;
    static constexpr size_t off_qw = (sizeof(SM) + 127) / 128 * 128;
    static constexpr size_t off_dpx = off_qw + sizeof(float2) * WARPS * BW_GROUP * STG_STRIDE;
    static constexpr size_t off_mask = off_dpx + sizeof(float) * WARPS * C * 32;
    static constexpr size_t off_list = off_mask + sizeof(uint32_t) * WARPS * BW_BATCH;
    static constexpr size_t bytes = off_list + WARPS * BW_BATCH;
};

template <int C, int WARPS, int MODE>
__global__ void __launch_bounds__(32 * WARPS)
blend_backward_kernel(const uint2* __restrict__ ranges, const float4* __restrict__ inst_conic,
                      const float4* __restrict__ inst_geom, const float4* __restrict__ inst_feat, int W, int H,
                      int gx, const float* __restrict__ bg, const float* __restrict__ final_Ts,
                      const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                      float* __restrict__ gacc, const uint32_t* __restrict__ tile_order) {
    using LY = BwdLayout<C, WARPS>;
    using SM = typename LY::SM;
    constexpr int FV = SM::FV;
    constexpr int PARTS = (BL_THREADS / 32) / WARPS;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    SM& sm = *reinterpret_cast<SM*>(smem_raw);
    const int wid = threadIdx.x >> 5;
    float2* stg = reinterpret_cast<float2*>(smem_raw + LY::off_qw) + wid * BW_GROUP * STG_STRIDE;
    const uint32_t stg_s = (uint32_t)__cvta_generic_to_shared(stg);
    float* dpx = reinterpret_cast<float*>(smem_raw + LY::off_dpx) + wid * C * 32;   // [C][32]
    uint32_t* masks = reinterpret_cast<uint32_t*>(smem_raw + LY::off_mask) + wid * BW_BATCH;
    uint8_t* list = reinterpret_cast<uint8_t*>(smem_raw + LY::off_list) + wid * BW_BATCH;
    __shared__ int s_max_contrib;
    __shared__ int s_released[BW_STAGES];   // warps that are through with the stage's current batch

    const int tile = tile_order ? (int)tile_order[blockIdx.x / PARTS] : (int)(blockIdx.x / PARTS);
    const int tx = tile % gx, ty = tile / gx;
    const int lane = threadIdx.x & 31;
    const int strip = (blockIdx.x % PARTS) * WARPS + wid;
    const int px = tx * TILE_X + (lane & (TILE_X - 1));
    const int py = ty * TILE_Y + 2 * strip + (lane >> 4);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix_id = (size_t)W * py + px;
    const float strip_x0 = (float)(tx * TILE_X);
    const float strip_y0 = (float)(ty * TILE_Y + 2 * strip);
    const uint2 range = ranges[tile];

    const float T_final = inside ? final_Ts[pix_id] : 0.f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;

    if (threadIdx.x == 0) {
        s_max_contrib = 0;
#pragma unroll
        for (int q = 0; q < BW_STAGES; q++) {
            mbar_init(&sm.bar[q], 1);
            s_released[q] = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();
    const int warp_max = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0 && warp_max > 0) atomicMax(&s_max_contrib, warp_max);
    __syncthreads();
    // Entries at list index >= len are skipped by every pixel of the tile (backward.cu:489-490).
    const int len = s_max_contrib;
    const int nb = (len + BW_BATCH - 1) / BW_BATCH;
    // batch k covers list indices [hi_k - cnt_k, hi_k), hi_k = len - k*BATCH, walked back to front
    if (threadIdx.x == 0) {
        for (int q = 0; q < BW_STAGES && q < nb; q++) {
            const int hi = len - q * BW_BATCH, cnt = min(BW_BATCH, hi);
            issue_batch(sm, q, inst_conic, inst_geom, inst_feat, range.x + (uint32_t)(hi - cnt), cnt);
        }
    }

    float dL_dpixel[C];
    float adot = 0.f, last_cdot = 0.f;
    float bg_dot_dpixel = 0.f;
#pragma unroll
    for (int ch = 0; ch < C; ch++) {
        dL_dpixel[ch] = inside ? __ldg(dL_dpixels + (size_t)ch * H * W + pix_id) : 0.f;
        bg_dot_dpixel += __ldg(bg + ch) * dL_dpixel[ch];
        dpx[ch * 32 + lane] = dL_dpixel[ch];
    }
    float last_alpha = 0.f;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;  // backward.cu:460-461
    // phase-2 identity of this lane: (record rl, pixel row, segment of the row)
    const int rl = lane & (BW_GROUP - 1), row = lane >> 4, seg = (lane & 15) / BW_GROUP;
    const uint32_t seg_bits = (BW_SPLIT == 1) ? 0xffffu : (0xffu << (8 * seg));
    const float row_y = strip_y0 + (float)row;

    for (int b = 0; b < nb; b++) {
        const int s = b % BW_STAGES;
        const int hi = len - b * BW_BATCH;
        const int cnt = min(BW_BATCH, hi);
        const int lo = hi - cnt;  // list index of record 0 of this batch
        const int last_in_batch = last_contributor - lo;   // the pixel's last contributor relative to this batch
        mbar_wait(&sm.bar[s], (uint32_t)((b / BW_STAGES) & 1));

        // records at list index >= warp_max are skipped by the whole warp; so are records whose box cannot
        // reach this warp's strip (see cull_batch)
        const int n = cull_batch(sm, s, min(cnt, warp_max - lo), lane, strip_x0, strip_y0, list, masks);
#pragma unroll 1
        for (int t = n; t > 0; t -= BW_GROUP) {
            const int gbase = max(t - BW_GROUP, 0), gs = t - gbase;
            const bool have = rl < gs;
            const int jr = have ? (int)list[gbase + rl] : 0;           // both half-warps hold the group's records
            const uint32_t mrow = have ? ((masks[gbase + rl] >> (16 * row)) & 0xffffu) : 0u;  // my record, my row
            // records of the group that reach MY pixel (with 8-record groups only the first 8 lanes of a half-warp pass a row)
            uint32_t hits = half_warp_bit_transpose((BW_SPLIT == 1 || seg == 0) ? mrow : 0u, lane);

            // ---- phase 1: lane = pixel, back to front ----
#pragma unroll 1
            while (__any_sync(0xffffffffu, hits != 0u)) {
                int jj[BW_UNROLL], rr[BW_UNROLL];
                uint32_t valid = 0u;
#pragma unroll
                for (int u = 0; u < BW_UNROLL; u++) {
                    const bool has = hits != 0u;
                    const int r = find_msb(hits) & 31;   // 31 when hits == 0: clearing bit 31 of 0 is a no-op
                    hits &= ~(1u << r);
                    rr[u] = r;
                    jj[u] = __shfl_sync(0xffffffffu, jr, r);
                    if (has) valid |= 1u << u;
                }
                float al[BW_UNROLL], Gs[BW_UNROLL];
                uint32_t act = 0u;
#pragma unroll
                for (int u = 0; u < BW_UNROLL; u++) {
                    const int j = jj[u];
                    const float4 co = sm.conic[s][j];
                    const float4 ge = sm.geom[s][j];
                    const float dx = __fsub_rn(ge.x, pxf), dy = __fsub_rn(ge.y, pyf);
                    const float power = eval_power(co, dx, dy);
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, __fmul_rn(co.w, G));
                    al[u] = alpha;
                    Gs[u] = G;
                    // the reference's `contributor` after its decrement is lo + j (backward.cu:486-490)
                    if (((valid >> u) & 1u) && j < last_in_batch && !(power > 0.0f) &&
                        !(alpha < 1.0f / 255.0f))
                        act |= 1u << u;
                }
#pragma unroll
                for (int u = 0; u < BW_UNROLL; u++) {
                    float q = 0.f, w = 0.f;
                    if ((act >> u) & 1u) {
                        const int j = jj[u];
                        const float alpha = al[u];
                        const float inv_1ma = rcp_approx(1.f - alpha);    // 1 - alpha is in [0.01, 1]: the 1-ulp approximate reciprocal; shared by T / (1 - alpha) and the background term
                        T = T * inv_1ma;
                        const float4 f0 = sm.feat[s][j * FV];
                        float col[C];
                        col[0] = f0.x; col[1] = f0.y; col[2] = f0.z;
                        if (C == 6) {
                            col[3 % C] = f0.w;
                            if (MODE != BWD_ALL) {
                                // fused frame path: the second triple is (z, 1, z^2) by construction (fused.cu gathers),
                                // so the second feature vector need not be read
                                col[4 % C] = 1.f;
                                col[5 % C] = __fmul_rn(f0.w, f0.w);
                            } else {
                                const float4 f1 = sm.feat[s][j * FV + FV - 1];
                                col[4 % C] = f1.x; col[5 % C] = f1.y;
                            }
                        }
                        // backward.cu:516-528 keeps accum_rec[ch] and sums (col[ch] - accum_rec[ch]) * dL_dpixel[ch];
                        // only that sum is ever used, so the recurrence is carried on the two dot products
                        //   cdot = sum col[ch] dL_dpixel[ch]   and   adot = sum accum_rec[ch] dL_dpixel[ch]
                        float cdot = 0.f;
#pragma unroll
                        for (int ch = 0; ch < C; ch++) cdot = fmaf(col[ch], dL_dpixel[ch], cdot);
                        adot = last_alpha * last_cdot + (1.f - last_alpha) * adot;
                        last_cdot = cdot;
                        float dL_dalpha = (cdot - adot) * T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;  // backward.cu:531-534
                        q = Gs[u] * dL_dalpha;
                        w = alpha * T;                                      // dchannel_dcolor (:513)
                    }
                    if ((valid >> u) & 1u) sts_float2(stg_s + 8u * (uint32_t)stage_slot(rr[u], lane), q, w);
                }
            }
            __syncwarp();

            // ---- phase 2: lane = (record rl, pixel row): serial sums over the row's pixels inside the box ----
            {
                float Sq = 0.f, Sx = 0.f, Sxx = 0.f, cc[C];
#pragma unroll
                for (int ch = 0; ch < C; ch++) cc[ch] = 0.f;
                const float4 ge = sm.geom[s][jr];
                const float2* my_stg = stg + stage_slot(rl, 16 * row);
                const float* my_dpx = dpx + 16 * row;
                // a row's reachable columns are one interval (column_bits): walk it with a counter
                const uint32_t bits = mrow & seg_bits;
                const int k_end = find_msb(bits);                            // -1 for an empty row
                const int k_begin = __ffs((int)(bits | 0x10000u)) - 1;          // 16 for an empty row
                float xk = strip_x0 + (float)k_begin;
                for (int k = k_begin; k <= k_end; k++, xk += 1.0f) {
                    const float2 qw = my_stg[k];
                    const float dx = __fsub_rn(ge.x, xk);                    // same dx the pixel lane used
                    const float qdx = qw.x * dx;
                    Sq += qw.x;
                    Sx += qdx;
                    Sxx = fmaf(qdx, dx, Sxx);
#pragma unroll
                    for (int ch = 0; ch < C; ch++)
                        if (need_colour<C, MODE>(ch)) cc[ch] = fmaf(qw.y, my_dpx[ch * 32 + k], cc[ch]);
                }
                // the 12 gradient terms of backward.cu:505-554 for (record, row); dy is constant along the row
                const float4 co = sm.conic[s][jr];
                const float dy = __fsub_rn(ge.y, row_y);
                const float Sy = dy * Sq, Sxy = dy * Sx, Syy = dy * Sy;
                float v[12];
                v[0] = -0.5f * co.w * Sxx;                                  // dL_dconic.x : -0.5 gdx dx dL_dG
                v[1] = -0.5f * co.w * Sxy;                                  // dL_dconic.y
                v[2] = -0.5f * co.w * Syy;                                  // dL_dconic.w
                v[3] = Sq;                                                  // dL_dopacity : G dL_dalpha
                v[4] = -co.w * ddelx_dx * (co.x * Sx + co.y * Sy);          // dL_dmean2D.x : dL_dG dG_ddelx ddelx_dx
                v[5] = -co.w * ddely_dy * (co.z * Sy + co.y * Sx);          // dL_dmean2D.y
#pragma unroll
                for (int ch = 0; ch < 6; ch++) v[6 + ch] = (ch < C && need_colour<C, MODE>(ch)) ? cc[ch % C] : 0.f;
                // add the two rows of the strip; then each row-lane commits half of the accumulator row
                // (conic x,y,w | opacity | mean2D x,y || colour 0..5) with vector reductions to global memory
#pragma unroll
                for (int i = 0; i < 12; i++) {
                    const bool used = i < 6 ? (i != 3 || MODE != BWD_TRACKING) : (i - 6 < C && need_colour<C, MODE>(i - 6));
                    if (used && BW_SPLIT == 2) v[i] += __shfl_xor_sync(0xffffffffu, v[i], 8);   // the two halves of the row
                    if (used) v[i] += __shfl_xor_sync(0xffffffffu, v[i], 16);
                }
                if (have && seg == 0) {
                    float* dst = gacc + (size_t)__float_as_uint(ge.w) * GACC_STRIDE;
                    if (row == 0) {
                        red_add_v4(dst, v[0], v[1], v[2], v[3]);
                        red_add_v2(dst + 4, v[4], v[5]);
                    } else {
                        if (MODE != BWD_TRACKING) red_add_v2(dst + 6, v[6], v[7]);
                        if (C == 6 || MODE == BWD_ALL) red_add_v4(dst + 8, v[8], v[9], v[10], v[11]);
                    }
                }
            }
            __syncwarp();  // the staging tile is rewritten by the next group
        }
        // No CTA-wide barrier per batch: a warp that is through with stage s moves on to the next batch (already
        // staged) while slower strips finish; the LAST warp to leave the stage refills it.  The copy into stage s
        // is issued only after every warp's reads of it (ordered by the fence + the counter), and a warp reaches the
        // refilled stage only through the copy's mbarrier.
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(&s_released[s], 1) == WARPS - 1) {
                s_released[s] = 0;
                if (b + BW_STAGES < nb) {
                    const int nhi = len - (b + BW_STAGES) * BW_BATCH, ncnt = min(BW_BATCH, nhi);
                    issue_batch(sm, s, inst_conic, inst_geom, inst_feat, range.x + (uint32_t)(nhi - ncnt), ncnt);
                }
            }
        }
    }
}

}  // namespace

// Warps per CTA: whole tiles (8 warps) when there are plenty of tiles, finer splits for small images.
static int pick_warps(int tiles, bool forward) {
    if (tiles >= (forward ? 16 : 3) * NUM_SMS) return 8;  // forward: half-tile CTAs pack the SMs better (measured at 1200 tiles)
    if (tiles * 2 >= 3 * NUM_SMS) return 4;
    if (tiles * 4 >= 3 * NUM_SMS) return 2;
    return 1;
}

template <int C, bool FRAME>
static void launch_fwd_c(int warps, int tiles, cudaStream_t st, const uint2* ranges, const BinningState& b, int W, int H,
                         int gx, const float* bg, float* out_color, float* out_depth, float* out_opacity,
                         float* final_T, uint32_t* n_contrib, const uint32_t* tile_order) {
#define ASB_FWD(WP)                                                                                              \
    blend_forward_kernel<C, WP, FRAME><<<tiles*(8 / WP), 32 * WP, 0, st>>>(ranges, b.inst_conic, b.inst_geom, b.inst_feat, W, \
                                                                    H, gx, bg, out_color, out_depth, out_opacity,    \
                                                                    final_T, n_contrib, tile_order)
    if (warps == 8) ASB_FWD(8);
    else if (warps == 4) ASB_FWD(4);
    else if (warps == 2) ASB_FWD(2);
    else ASB_FWD(1);
#undef ASB_FWD
}

template <int C, int MODE>
static void launch_bwd_c(int warps, int tiles, cudaStream_t st, const uint2* ranges, const BinningState& b, int W, int H,
                         int gx, const float* bg, const float* final_T, const uint32_t* n_contrib,
                         const float* dL_dpix, float* gacc, const uint32_t* tile_order) {
#define ASB_BWD(WP)                                                                                                \
    do {                                                                                                           \
        constexpr size_t smem = BwdLayout<C, WP>::bytes;                                                           \
        if (smem > 48 * 1024) {                                                                                    \
            /* opt in to the large dynamic shared-memory carve-out once per (kernel instance, device) */           \
            static std::atomic<unsigned> done_mask{0};                                                             \
            int dev = 0;                                                                                           \
            cudaGetDevice(&dev);                                                                                   \
            const unsigned bit = 1u << (dev & 31);                                                                 \
            if (!(done_mask.load(std::memory_order_relaxed) & bit)) {                                              \
                cudaFuncSetAttribute(blend_backward_kernel<C, WP, MODE>,                                           \
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                      \
                done_mask.fetch_or(bit, std::memory_order_relaxed);                                                \
            }                                                                                                      \
        }                                                                                                          \
        blend_backward_kernel<C, WP, MODE><<<tiles*(8 / WP), 32 * WP, smem, st>>>(                                 \
            ranges, b.inst_conic, b.inst_geom, b.inst_feat, W, H, gx, bg, final_T, n_contrib, dL_dpix, gacc,        \
            tile_order);                                                                                           \
    } while (0)
    if (warps == 8) ASB_BWD(8);
    else if (warps == 4) ASB_BWD(4);
    else if (warps == 2) ASB_BWD(2);
    else ASB_BWD(1);
#undef ASB_BWD
}

void launch_blend_forward(int channels, int W, int H, const uint2* ranges, const BinningState& b, const float* bg,
                          float* out_color, float* out_depth, float* out_opacity, float* final_T,
                          uint32_t* n_contrib, cudaStream_t st, const uint32_t* tile_order, bool frame_features) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;
    if (tiles <= 0) return;
    const int warps = pick_warps(tiles, true);
    if (channels == 3)
        launch_fwd_c<3, false>(warps, tiles, st, ranges, b, W, H, gx, bg, out_color, out_depth, out_opacity, final_T,
                               n_contrib, tile_order);
    else if (frame_features)
        launch_fwd_c<6, true>(warps, tiles, st, ranges, b, W, H, gx, bg, out_color, out_depth, out_opacity, final_T,
                              n_contrib, tile_order);
    else
        launch_fwd_c<6, false>(warps, tiles, st, ranges, b, W, H, gx, bg, out_color, out_depth, out_opacity, final_T,
                               n_contrib, tile_order);
}

// scratch of the chunked forward: chunk_base [tiles + 1] | chunk_tile [max_chunks] | per-chunk fields
int blend_chunk_count_bound(int capacity, int tiles) { return capacity / BL_CHUNK + tiles + 1; }
size_t blend_chunk_scratch_bytes(int capacity, int tiles) {
    const size_t mc = (size_t)blend_chunk_count_bound(capacity, tiles);
    const size_t ints = ((size_t)tiles + 1 + mc + 3) / 4 * 4;
    return ints * sizeof(int) + mc * CHUNK_FIELDS * TILE_PIX * sizeof(float);
}

void launch_blend_forward_chunked(int W, int H, const uint2* ranges, const BinningState& b, int capacity, const float* bg,
                                  float* out_color, float* out_depth, float* out_opacity, float* final_T,
                                  uint32_t* n_contrib, void* scratch, cudaStream_t st) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;
    if (tiles <= 0) return;
    const int mc = blend_chunk_count_bound(capacity, tiles);
    int* chunk_base = reinterpret_cast<int*>(scratch);
    int* chunk_tile = chunk_base + tiles + 1;
    const size_t ints = ((size_t)tiles + 1 + mc + 3) / 4 * 4;
    float* fields = reinterpret_cast<float*>(chunk_base + ints);
    blend_chunk_table_kernel<<<1, 1024, 0, st>>>(ranges, tiles, mc, chunk_base, chunk_tile);
    blend_forward_chunk_kernel<6, 1><<<mc, BL_THREADS, 0, st>>>(ranges, b.inst_conic, b.inst_geom, b.inst_feat, W, H, gx,
                                                                tiles, chunk_base, chunk_tile, fields);
    blend_forward_chunk_kernel<6, 2><<<mc, BL_THREADS, 0, st>>>(ranges, b.inst_conic, b.inst_geom, b.inst_feat, W, H, gx,
                                                                tiles, chunk_base, chunk_tile, fields);
    blend_chunk_combine_kernel<6><<<tiles, BL_THREADS, 0, st>>>(W, H, gx, chunk_base, fields, bg, out_color, out_depth,
                                                               out_opacity, final_T, n_contrib);
}

void launch_blend_backward(int channels, int W, int H, const uint2* ranges, const BinningState& b, const float* bg,
                           const float* final_T, const uint32_t* n_contrib, const float* dL_dpix, float* gacc,
                           cudaStream_t st, int mode, const uint32_t* tile_order) {
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;
    if (tiles <= 0) return;
    const int warps = pick_warps(tiles, false);
    if (channels == 3)
        launch_bwd_c<3, BWD_ALL>(warps, tiles, st, ranges, b, W, H, gx, bg, final_T, n_contrib, dL_dpix, gacc,
                                  tile_order);
    else if (mode == BWD_MAPPING)
        launch_bwd_c<6, BWD_MAPPING>(warps, tiles, st, ranges, b, W, H, gx, bg, final_T, n_contrib, dL_dpix, gacc,
                                  tile_order);
    else if (mode == BWD_TRACKING)
        launch_bwd_c<6, BWD_TRACKING>(warps, tiles, st, ranges, b, W, H, gx, bg, final_T, n_contrib, dL_dpix, gacc,
                                  tile_order);
    else
        launch_bwd_c<6, BWD_ALL>(warps, tiles, st, ranges, b, W, H, gx, bg, final_T, n_contrib, dL_dpix, gacc,
                                  tile_order);
}

}  // namespace asb
