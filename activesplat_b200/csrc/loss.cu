// loss.cu -- the mapping loss of get_loss, forward and backward, fused (sm_100a).
//
//   reference: src/mapper/splatam/splatam.py:213-249 (mapping branch, use_l1) with
//              calc_ssim / _ssim                    src/mapper/splatam/utils/slam_external.py:54-97
//   loss = w_im * (0.8 * mean|im - gt| + 0.2 * (1 - mean SSIM(im, gt))) + w_depth * mean_mask|gt_d - depth|
//   mask = gt_depth > 0 & !isnan(depth) & !isnan(depth_sq - depth^2)
//
// The reference runs this as ~40 torch kernels (5 grouped 11x11 convolutions forward, their backward,
// boolean-index copies, reductions).  Here: two sweeps over the image.
//   mapping_loss_stats_kernel : separable 11-tap Gaussian window in shared memory (zero padding, sigma 1.5)
//       -> SSIM map, its partial derivatives w.r.t. the three window averages that depend on `im`
//       (dS/dmu1, dS/dE[x^2], dS/dE[xy]) and per-block partial sums (|d im|, SSIM, |d depth|, mask count);
//   mapping_loss_grad_kernel  : the same window applied to the three derivative maps (the adjoint of a
//       symmetric zero-padded window is the window itself) + the L1 terms -> dL/d(colour6).
#include <math_constants.h>

#include "common.cuh"

namespace asb {

namespace {

constexpr int ST = 16;            // output tile edge
constexpr int HALO = 5;           // window_size // 2
constexpr int SW = ST + 2 * HALO; // 26
constexpr int SSIM_THREADS = ST * ST;
constexpr float SSIM_C1 = 0.0001f;  // 0.01^2
constexpr float SSIM_C2 = 0.0009f;  // 0.03^2

struct Window {
    float g[11];
};

__device__ __forceinline__ float sgnf(float x) { return (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f); }

// Loads a (ST+10)^2 window of `img` (one channel, zero outside the image) into smem.
__device__ __forceinline__ void load_halo(const float* __restrict__ img, int W, int H, int x0, int y0, float* sm) {
    for (int i = threadIdx.x; i < SW * SW; i += SSIM_THREADS) {
        const int ly = i / SW, lx = i % SW;
        const int x = x0 + lx - HALO, y = y0 + ly - HALO;
        sm[i] = (x >= 0 && x < W && y >= 0 && y < H) ? __ldg(img + (size_t)y * W + x) : 0.f;
    }
}

// stats: per channel window averages -> SSIM and derivative maps.  partials[block][8]:
// 0 sum|im-gt|, 1 sum SSIM, 2 sum_mask|gt_d-depth|, 3 mask count.
__global__ void __launch_bounds__(SSIM_THREADS)
mapping_loss_stats_kernel(int W, int H, const float* __restrict__ c6, const float* __restrict__ gt_rgb,
                          const float* __restrict__ gt_depth, Window win, float* __restrict__ dmaps,
                          float* __restrict__ partials) {
    __shared__ float sx[SW * SW], sy[SW * SW];
    __shared__ float sh[5][SW * ST];  // horizontally filtered: x, y, xx, yy, xy  (26 rows x 16 cols)
    __shared__ float s_red[4 * 8];
    const int x0 = blockIdx.x * ST, y0 = blockIdx.y * ST;
    const int lx = threadIdx.x % ST, ly = threadIdx.x / ST;
    const int px = x0 + lx, py = y0 + ly;
    const bool inside = px < W && py < H;
    const size_t N = (size_t)W * H;
    float s_l1 = 0.f, s_ssim = 0.f, s_d = 0.f, s_cnt = 0.f;
    for (int ch = 0; ch < 3; ch++) {
        __syncthreads();
        load_halo(c6 + ch * N, W, H, x0, y0, sx);
        load_halo(gt_rgb + ch * N, W, H, x0, y0, sy);
        __syncthreads();
        for (int i = threadIdx.x; i < SW * ST; i += SSIM_THREADS) {
            const int r = i / ST, c = i % ST;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll
            for (int k = 0; k < 11; k++) {
                const float x = sx[r * SW + c + k], y = sy[r * SW + c + k], g = win.g[k];
                a0 = fmaf(g, x, a0);
                a1 = fmaf(g, y, a1);
                a2 = fmaf(g, x * x, a2);
                a3 = fmaf(g, y * y, a3);
                a4 = fmaf(g, x * y, a4);
            }
            sh[0][i] = a0; sh[1][i] = a1; sh[2][i] = a2; sh[3][i] = a3; sh[4][i] = a4;
        }
        __syncthreads();
        float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const int i = (ly + k) * ST + lx;
            const float g = win.g[k];
            mu1 = fmaf(g, sh[0][i], mu1);
            mu2 = fmaf(g, sh[1][i], mu2);
            e11 = fmaf(g, sh[2][i], e11);
            e22 = fmaf(g, sh[3][i], e22);
            e12 = fmaf(g, sh[4][i], e12);
        }
        if (inside) {
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
            const float sig1 = e11 - mu1_sq, sig2 = e22 - mu2_sq, sig12 = e12 - mu12;
            const float A1 = 2.f * mu12 + SSIM_C1, A2 = 2.f * sig12 + SSIM_C2;
            const float B1 = mu1_sq + mu2_sq + SSIM_C1, B2 = sig1 + sig2 + SSIM_C2;
            const float inv = 1.f / (B1 * B2);
            const float S = A1 * A2 * inv;
            s_ssim += S;
            const size_t o = ch * N + (size_t)py * W + px;
            dmaps[o] = (2.f * mu2 * (A2 - A1)) * inv - S * (2.f * mu1 / B1 - 2.f * mu1 / B2);  // dS/dmu1
            dmaps[3 * N + o] = -S / B2;                                                          // dS/dE[x^2]
            dmaps[6 * N + o] = 2.f * A1 * inv;                                                   // dS/dE[xy]
            s_l1 += fabsf(sx[(ly + HALO) * SW + lx + HALO] - sy[(ly + HALO) * SW + lx + HALO]);
        }
    }
    if (inside) {
        const size_t p = (size_t)py * W + px;
        const float depth = c6[3 * N + p], dsq = c6[5 * N + p], gd = gt_depth[p];
        const float invis = dsq - depth * depth;
        if ((gd > 0.f) && !isnan(depth) && !isnan(invis)) {
            s_d = fabsf(gd - depth);
            s_cnt = 1.f;
        }
    }
    float v[4] = {s_l1, s_ssim, s_d, s_cnt};
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) s_red[k * 8 + warp] = x;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        float x = 0.f;
#pragma unroll
        for (int w = 0; w < SSIM_THREADS / 32; w++) x += s_red[threadIdx.x * 8 + w];
        partials[(size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 4 + threadIdx.x] = x;
    }
}

// losses[0] weighted total, [1] weighted im, [2] weighted depth, [3] mask count, [4] R, [5] overflow, [6] mean SSIM
__global__ void mapping_loss_finalize_kernel(int nblocks, const float* __restrict__ partials, float n3, float w_im,
                                             float w_depth, float* __restrict__ losses,
                                             const uint32_t* __restrict__ status) {
    __shared__ float s[4][32];
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    for (int b = threadIdx.x; b < nblocks; b += 32)
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] += partials[(size_t)b * 4 + k];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        s[k][0] = x;
    }
    if (threadIdx.x == 0) {
        const float l1 = s[0][0] / n3, ssim = s[1][0] / n3;
        const float cnt = s[3][0];
        const float li = w_im * (0.8f * l1 + 0.2f * (1.0f - ssim));
        const float ld = w_depth * (s[2][0] / cnt);  // mean over an empty mask is NaN in torch as well
        losses[0] = li + ld;
        losses[1] = li;
        losses[2] = ld;
        losses[3] = cnt;
        if (status) {
            losses[4] = (float)status[0];
            losses[5] = (float)status[1];
            if (status[1]) {  // binning buffer overflowed: the frame is invalid -- make that impossible to miss
                losses[0] = CUDART_NAN_F; losses[1] = CUDART_NAN_F; losses[2] = CUDART_NAN_F;
            }
        }
        losses[6] = ssim;
    }
}

__global__ void __launch_bounds__(SSIM_THREADS)
mapping_loss_grad_kernel(int W, int H, const float* __restrict__ c6, const float* __restrict__ gt_rgb,
                         const float* __restrict__ gt_depth, const float* __restrict__ dmaps, Window win, float w_im,
                         float w_depth, const float* __restrict__ losses, float* __restrict__ dL) {
    __shared__ float sm[3][SW * SW];
    __shared__ float sh[3][SW * ST];
    const int x0 = blockIdx.x * ST, y0 = blockIdx.y * ST;
    const int lx = threadIdx.x % ST, ly = threadIdx.x / ST;
    const int px = x0 + lx, py = y0 + ly;
    const bool inside = px < W && py < H;
    const size_t N = (size_t)W * H;
    const float n3 = 3.f * (float)N;
    for (int ch = 0; ch < 3; ch++) {
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 3; q++) load_halo(dmaps + (3 * q + ch) * N, W, H, x0, y0, sm[q]);
        __syncthreads();
        for (int i = threadIdx.x; i < SW * ST; i += SSIM_THREADS) {
            const int r = i / ST, c = i % ST;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
            for (int k = 0; k < 11; k++) {
                const float g = win.g[k];
                a0 = fmaf(g, sm[0][r * SW + c + k], a0);
                a1 = fmaf(g, sm[1][r * SW + c + k], a1);
                a2 = fmaf(g, sm[2][r * SW + c + k], a2);
            }
            sh[0][i] = a0; sh[1][i] = a1; sh[2][i] = a2;
        }
        __syncthreads();
        float c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const int i = (ly + k) * ST + lx;
            const float g = win.g[k];
            c0 = fmaf(g, sh[0][i], c0);
            c1 = fmaf(g, sh[1][i], c1);
            c2 = fmaf(g, sh[2][i], c2);
        }
        if (inside) {
            const size_t o = ch * N + (size_t)py * W + px;
            const float x = c6[o], y = gt_rgb[o];
            const float dssim = c0 + 2.f * x * c1 + y * c2;  // d(sum SSIM)/d im
            dL[o] = w_im * (0.8f * sgnf(x - y) - 0.2f * dssim) / n3;
        }
    }
    if (inside) {
        const size_t p = (size_t)py * W + px;
        const float depth = c6[3 * N + p], dsq = c6[5 * N + p], gd = gt_depth[p];
        const float invis = dsq - depth * depth;
        float g = 0.f;
        if ((gd > 0.f) && !isnan(depth) && !isnan(invis)) g = -w_depth * sgnf(gd - depth) / losses[3];
        dL[3 * N + p] = g;
        dL[4 * N + p] = 0.f;
        dL[5 * N + p] = 0.f;
    }
}

Window make_window() {
    // gaussian(11, 1.5) of slam_external.py:54-56, normalised in fp32 like torch does
    Window w;
    float sum = 0.f;
    for (int i = 0; i < 11; i++) {
        w.g[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5));
        sum += w.g[i];
    }
    for (int i = 0; i < 11; i++) w.g[i] /= sum;
    return w;
}

}  // namespace

size_t mapping_loss_scratch_floats(int W, int H) {
    const size_t bx = (W + ST - 1) / ST, by = (H + ST - 1) / ST;
    return 9 * (size_t)W * H + bx * by * 4 + 16;
}

void launch_mapping_loss(int W, int H, const float* c6, const float* gt_rgb, const float* gt_depth, float w_im,
                         float w_depth, float* dL, float* losses, float* scratch, const uint32_t* status,
                         cudaStream_t st) {
    const dim3 grid((W + ST - 1) / ST, (H + ST - 1) / ST);
    const int nblocks = grid.x * grid.y;
    float* dmaps = scratch;
    float* partials = scratch + 9 * (size_t)W * H;
    const Window win = make_window();
    mapping_loss_stats_kernel<<<grid, SSIM_THREADS, 0, st>>>(W, H, c6, gt_rgb, gt_depth, win, dmaps, partials);
    mapping_loss_finalize_kernel<<<1, 32, 0, st>>>(nblocks, partials, 3.f * (float)W * (float)H, w_im, w_depth, losses,
                                                   status);
    mapping_loss_grad_kernel<<<grid, SSIM_THREADS, 0, st>>>(W, H, c6, gt_rgb, gt_depth, dmaps, win, w_im, w_depth,
                                                           losses, dL);
}

}  // namespace asb
