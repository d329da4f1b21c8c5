// densify.cu -- silhouette-driven densification on the device (sm_100a).
//
//   as_densify  replaces  add_new_gaussians + get_pointcloud + initialize_new_params
//               src/mapper/splatam/splatam.py:332-379, 25-75, 304-329
//
// The reference does this with ~40 torch kernels, three host synchronisations (torch.sum(mask) > 0, two boolean-index
// gathers) and a torch.cat that re-allocates and copies all five parameter tensors.  Here the new Gaussians are written
// straight behind the existing rows of tensors that were allocated with head-room:
//   1. depth_error = |gt_depth - depth| * (gt_depth > 0) and its exact MEDIAN (torch.median: the lower median, the
//      element of rank (n-1)/2) by a two-pass radix select on the float bits (the values are non-negative, so their
//      bit patterns order like the values): histogram of the high 16 bits, pick the bin, histogram of the low 16 bits
//      inside that bin, pick again;
//   2. per pixel  mask = (sil < thres) | (depth > gt_depth & depth_error > 2 median & sil > thres & gt_depth < 5),
//      then & (gt_depth > 0)                                                          (splatam.py:339-364);
//   3. ordered stream compaction (torch's boolean indexing keeps row-major order): per-block counts, one-block scan,
//      then every block writes its pixels' Gaussians at  P + offset + rank:
//         mean   = c2w * ((x - cx) / fx * z, (y - cy) / fy * z, z, 1)               (splatam.py:32-49)
//         colour = gt_rgb[:, y, x]          rotation = (1, 0, 0, 0)          logit opacity = 0
//         log scale = log(sqrt((z / ((fx + fy) / 2))^2))   (x1 isotropic, x3 anisotropic)   (splatam.py:55-57, 309-312)
// Nothing synchronises with the host; the caller reads result[0] (the number of new Gaussians) once, because the
// launches that follow need the new P.
#include <math_constants.h>

#include "common.cuh"

namespace asb {

namespace {

constexpr int DN_THREADS = 1024;
constexpr int DN_BINS = 1 << 16;

struct DensifyPix {
    float depth, sil, gt;
};

__device__ __forceinline__ DensifyPix densify_load(const float* __restrict__ c6, const float* __restrict__ gt_depth,
                                                  size_t N, size_t p) {
    DensifyPix q;
    q.depth = __ldg(c6 + 3 * N + p);   // [z, 1, z^2] channels of the bundle: depth, silhouette
    q.sil = __ldg(c6 + 4 * N + p);
    q.gt = __ldg(gt_depth + p);
    return q;
}

__device__ __forceinline__ float densify_error(const DensifyPix& q) {
    return fabsf(q.gt - q.depth) * ((q.gt > 0.f) ? 1.f : 0.f);   // splatam.py:345
}

// result words: 0 n_new, 1 overflow, 2 median bits, 3 mask count before the valid-depth test, 4 rank still to find,
// 5 selected high half, 6 number of NaN errors
template <bool LOW>
__global__ void __launch_bounds__(256)
densify_hist_kernel(size_t N, const float* __restrict__ c6, const float* __restrict__ gt_depth,
                    uint32_t* __restrict__ hist, uint32_t* __restrict__ result) {
    const uint32_t want_hi = LOW ? result[5] : 0u;
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < N; p += (size_t)gridDim.x * blockDim.x) {
        const uint32_t bits = __float_as_uint(densify_error(densify_load(c6, gt_depth, N, p)));
        if (!LOW) {
            if (bits > 0x7f800000u) atomicAdd(result + 6, 1u);   // NaN: torch.median then returns NaN
            atomicAdd(hist + (bits >> 16), 1u);
        } else if ((bits >> 16) == want_hi) {
            atomicAdd(hist + (bits & 0xffffu), 1u);
        }
    }
}

// One block: finds the bin that holds the element of rank result[4] (counting from 0), leaves the rank inside that bin
// in result[4] and clears the histogram for the next pass.
template <bool LOW>
__global__ void __launch_bounds__(DN_THREADS)
densify_select_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ result) {
    __shared__ uint32_t s_sum[DN_THREADS];
    __shared__ uint32_t s_found[2];
    constexpr int PER = DN_BINS / DN_THREADS;   // 64 consecutive bins per thread
    const uint32_t rank = result[4];
    uint32_t mine = 0;
    for (int i = 0; i < PER; i++) mine += hist[threadIdx.x * PER + i];
    s_sum[threadIdx.x] = mine;
    if (threadIdx.x == 0) { s_found[0] = 0xffffffffu; s_found[1] = 0; }
    __syncthreads();
    // inclusive scan of the per-thread sums (Hillis-Steele over 1024 entries)
    for (int off = 1; off < DN_THREADS; off <<= 1) {
        const uint32_t v = (threadIdx.x >= off) ? s_sum[threadIdx.x - off] : 0u;
        __syncthreads();
        s_sum[threadIdx.x] += v;
        __syncthreads();
    }
    const uint32_t before = s_sum[threadIdx.x] - mine;
    if (rank >= before && rank < before + mine) {   // exactly one thread
        uint32_t acc = before;
        for (int i = 0; i < PER; i++) {
            const uint32_t h = hist[threadIdx.x * PER + i];
            if (rank < acc + h) { s_found[0] = threadIdx.x * PER + i; s_found[1] = rank - acc; break; }
            acc += h;
        }
    }
    __syncthreads();
    for (int i = 0; i < PER; i++) hist[threadIdx.x * PER + i] = 0u;
    if (threadIdx.x == 0) {
        result[4] = s_found[1];
        if (!LOW) result[5] = s_found[0];
        else result[2] = result[6] ? 0x7fc00000u : ((result[5] << 16) | s_found[0]);
    }
}

__device__ __forceinline__ bool densify_mask(const DensifyPix& q, float sil_thres, float median, bool& before_valid) {
    const float err = densify_error(q);
    const bool non_presence_sil = q.sil < sil_thres;                                              // :340
    const bool non_presence_depth = (q.depth > q.gt) && (err > 2.f * median) && (q.sil > sil_thres) && (q.gt < 5.f);
    before_valid = non_presence_sil || non_presence_depth;                                          // :352
    return before_valid && (q.gt > 0.f);                                                            // :363-364
}

__global__ void __launch_bounds__(DN_THREADS)
densify_count_kernel(size_t N, const float* __restrict__ c6, const float* __restrict__ gt_depth, float sil_thres,
                     uint32_t* __restrict__ block_counts, uint32_t* __restrict__ result) {
    __shared__ uint32_t s_cnt[2];
    if (threadIdx.x == 0) { s_cnt[0] = 0; s_cnt[1] = 0; }
    __syncthreads();
    const size_t p = (size_t)blockIdx.x * DN_THREADS + threadIdx.x;
    bool before = false, m = false;
    if (p < N) m = densify_mask(densify_load(c6, gt_depth, N, p), sil_thres, __uint_as_float(result[2]), before);
    const uint32_t bm = __ballot_sync(0xffffffffu, m), bb = __ballot_sync(0xffffffffu, before);
    if ((threadIdx.x & 31) == 0) {
        if (bm) atomicAdd(&s_cnt[0], (uint32_t)__popc(bm));
        if (bb) atomicAdd(&s_cnt[1], (uint32_t)__popc(bb));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        block_counts[blockIdx.x] = s_cnt[0];
        if (s_cnt[1]) atomicAdd(result + 3, s_cnt[1]);
    }
}

// exclusive scan of the block counts in place (one block), total -> result[0], overflow flag -> result[1]
__global__ void __launch_bounds__(DN_THREADS)
densify_scan_kernel(int nblocks, uint32_t* __restrict__ block_counts, int room, uint32_t* __restrict__ result) {
    __shared__ uint32_t s[DN_THREADS];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += DN_THREADS) {
        const int i = base + threadIdx.x;
        const uint32_t v = (i < nblocks) ? block_counts[i] : 0u;
        s[threadIdx.x] = v;
        __syncthreads();
        for (int off = 1; off < DN_THREADS; off <<= 1) {
            const uint32_t t = (threadIdx.x >= off) ? s[threadIdx.x - off] : 0u;
            __syncthreads();
            s[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < nblocks) block_counts[i] = s_carry + s[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == DN_THREADS - 1) s_carry += s[threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        result[0] = s_carry;
        result[1] = (s_carry > (uint32_t)max(room, 0)) ? 1u : 0u;
    }
}

struct DensifyOut {
    float *means3D, *rgb, *rot, *logit, *log_scales, *timestep;
};

__global__ void __launch_bounds__(DN_THREADS)
densify_write_kernel(int W, int H, const float* __restrict__ c6, const float* __restrict__ gt_rgb,
                     const float* __restrict__ gt_depth, float sil_thres, float fx, float fy, float cx, float cy,
                     const float* __restrict__ c2w, int iso, float time_idx, int P, DensifyOut o,
                     const uint32_t* __restrict__ block_offsets, const uint32_t* __restrict__ result) {
    __shared__ uint32_t s_warp[DN_THREADS / 32];
    if (result[1]) return;   // not enough room: nothing is written, the caller grows the tensors and calls again
    const size_t N = (size_t)W * H;
    const size_t p = (size_t)blockIdx.x * DN_THREADS + threadIdx.x;
    bool before = false, m = false;
    DensifyPix q = {0.f, 0.f, 0.f};
    if (p < N) {
        q = densify_load(c6, gt_depth, N, p);
        m = densify_mask(q, sil_thres, __uint_as_float(result[2]), before);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t bm = __ballot_sync(0xffffffffu, m);
    if (lane == 0) s_warp[warp] = (uint32_t)__popc(bm);
    __syncthreads();
    if (!m) return;
    uint32_t rank = (uint32_t)__popc(bm & ((1u << lane) - 1u));
    for (int w = 0; w < warp; w++) rank += s_warp[w];
    const size_t row = (size_t)P + block_offsets[blockIdx.x] + rank;
    const int y = (int)(p / W), x = (int)(p - (size_t)y * W);
    const float z = q.gt;
    const float xc = ((float)x - cx) / fx * z, yc = ((float)y - cy) / fy * z;                       // splatam.py:37-44
    o.means3D[3 * row + 0] = fmaf(c2w[0], xc, fmaf(c2w[1], yc, fmaf(c2w[2], z, c2w[3])));           // :47-49
    o.means3D[3 * row + 1] = fmaf(c2w[4], xc, fmaf(c2w[5], yc, fmaf(c2w[6], z, c2w[7])));
    o.means3D[3 * row + 2] = fmaf(c2w[8], xc, fmaf(c2w[9], yc, fmaf(c2w[10], z, c2w[11])));
#pragma unroll
    for (int ch = 0; ch < 3; ch++) o.rgb[3 * row + ch] = __ldg(gt_rgb + ch * N + p);                // :62-63
    o.rot[4 * row + 0] = 1.f; o.rot[4 * row + 1] = 0.f; o.rot[4 * row + 2] = 0.f; o.rot[4 * row + 3] = 0.f;
    o.logit[row] = 0.f;
    const float sg = z / ((fx + fy) / 2.f);                                                         // :56
    const float ls = logf(sqrtf(sg * sg));                                                          // :57, :310-312
    if (iso) {
        o.log_scales[row] = ls;
    } else {
        o.log_scales[3 * row + 0] = ls; o.log_scales[3 * row + 1] = ls; o.log_scales[3 * row + 2] = ls;
    }
    if (o.timestep) o.timestep[row] = time_idx;                                                     // :376-377
}

}  // namespace

size_t densify_scratch_words(int W, int H) {
    const size_t N = (size_t)W * H;
    return DN_BINS + (N + DN_THREADS - 1) / DN_THREADS + 16;
}

void launch_densify(int W, int H, const float* c6, const float* gt_rgb, const float* gt_depth, float sil_thres, float fx,
                    float fy, float cx, float cy, const float* c2w, int iso, float time_idx, int P, int capacity,
                    float* means3D, float* rgb, float* rot, float* logit, float* log_scales, float* timestep,
                    uint32_t* scratch, uint32_t* result, cudaStream_t st) {
    const size_t N = (size_t)W * H;
    const int nblocks = (int)((N + DN_THREADS - 1) / DN_THREADS);
    uint32_t* hist = scratch;
    uint32_t* block_counts = scratch + DN_BINS;
    cudaMemsetAsync(hist, 0, sizeof(uint32_t) * DN_BINS, st);
    cudaMemsetAsync(result, 0, sizeof(uint32_t) * 8, st);
    const uint32_t rank = (uint32_t)((N - 1) / 2);   // torch.median of n values: the element of rank (n - 1) / 2
    cudaMemcpyAsync(result + 4, &rank, sizeof(uint32_t), cudaMemcpyHostToDevice, st);   // pageable 4 bytes: staged by the driver
    const int hb = (int)min((size_t)NUM_SMS * 8, (N + 255) / 256);
    densify_hist_kernel<false><<<hb, 256, 0, st>>>(N, c6, gt_depth, hist, result);
    densify_select_kernel<false><<<1, DN_THREADS, 0, st>>>(hist, result);
    densify_hist_kernel<true><<<hb, 256, 0, st>>>(N, c6, gt_depth, hist, result);
    densify_select_kernel<true><<<1, DN_THREADS, 0, st>>>(hist, result);
    densify_count_kernel<<<nblocks, DN_THREADS, 0, st>>>(N, c6, gt_depth, sil_thres, block_counts, result);
    densify_scan_kernel<<<1, DN_THREADS, 0, st>>>(nblocks, block_counts, capacity - P, result);
    DensifyOut o = {means3D, rgb, rot, logit, log_scales, timestep};
    densify_write_kernel<<<nblocks, DN_THREADS, 0, st>>>(W, H, c6, gt_rgb, gt_depth, sil_thres, fx, fy, cx, cy, c2w, iso,
                                                         time_idx, P, o, block_counts, result);
}

}  // namespace asb
