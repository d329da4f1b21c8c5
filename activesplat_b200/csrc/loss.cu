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

constexpr int TW = 32, TH = 16;   // output tile (columns x rows) of one block
constexpr int HALO = 5;           // window_size // 2
constexpr int XOFF = 8;             // staged column 0 is image column x0 - XOFF: a multiple of 4, so that rows can be
                                    // fetched with aligned 128-bit loads (the window needs x0 - HALO and up)
constexpr int SWP = TW + 2 * XOFF;  // 48 staged columns = 12 float4 per row
constexpr int SHH = TH + 2 * HALO;  // 26 staged rows
constexpr int SSIM_THREADS = 128;
constexpr int SEG = 4;              // outputs per thread along the filter direction (sliding window kept in registers)
constexpr float SSIM_C1 = 0.0001f;  // 0.01^2
constexpr float SSIM_C2 = 0.0009f;  // 0.03^2

struct Window {
    float g[11];
};

__device__ __forceinline__ float sgnf(float x) { return (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f); }

// Loads the neighbourhood of a tile of `img` (one channel, zero outside the image) into smem: TH + 10 rows of SWP columns
// starting at image column x0 - XOFF.  Rows of an image whose width is a multiple of 4 come in as aligned float4s (312
// loads per tile instead of 1144 scalar ones with their index arithmetic: the loads were half of the stall samples).
__device__ __forceinline__ void load_halo(const float* __restrict__ img, int W, int H, int x0, int y0, float* sm) {
    if ((W & 3) == 0 && (reinterpret_cast<uintptr_t>(img) & 15) == 0) {
        for (int i = threadIdx.x; i < SHH * (SWP / 4); i += SSIM_THREADS) {
            const int ly = i / (SWP / 4), q = i - ly * (SWP / 4);
            const int x = x0 - XOFF + 4 * q, y = y0 + ly - HALO;   // x is a multiple of 4: the four are in or out together
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (x >= 0 && x < W && y >= 0 && y < H) v = __ldg(reinterpret_cast<const float4*>(img + (size_t)y * W + x));
            reinterpret_cast<float4*>(sm)[i] = v;
        }
    } else {
        for (int i = threadIdx.x; i < SHH * SWP; i += SSIM_THREADS) {
            const int ly = i / SWP, lx = i - ly * SWP;
            const int x = x0 - XOFF + lx, y = y0 + ly - HALO;
            sm[i] = (x >= 0 && x < W && y >= 0 && y < H) ? __ldg(img + (size_t)y * W + x) : 0.f;
        }
    }
}

// The 14 inputs of outputs c .. c+3 of a staged row (`row` points at staged column c, 16-byte aligned): image columns
// x0 + c - HALO and up are staged columns c + XOFF - HALO and up, so five LDS.128 cover them
__device__ __forceinline__ void load_run(const float* row, float (&v)[16]) {
    float t[20];
#pragma unroll
    for (int q = 0; q < 5; q++) {
        const float4 u = *reinterpret_cast<const float4*>(row + 4 * q);
        t[4 * q] = u.x; t[4 * q + 1] = u.y; t[4 * q + 2] = u.z; t[4 * q + 3] = u.w;
    }
#pragma unroll
    for (int i = 0; i < 14; i++) v[i] = t[i + XOFF - HALO];
    v[14] = v[15] = 0.f;
}

// SEG outputs of the 11-tap window over a run of SEG + 10 inputs, taps accumulated in ascending order (the order the
// first version of these kernels used, so the results are unchanged bit for bit)
__device__ __forceinline__ void window_run(const float (&v)[16], const Window& win, float (&out)[SEG]) {
#pragma unroll
    for (int o = 0; o < SEG; o++) {
        float a = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) a = fmaf(win.g[k], v[o + k], a);
        out[o] = a;
    }
}

// stats: per channel window averages -> SSIM and derivative maps.  partials[block][4]:
// 0 sum|im-gt|, 1 sum SSIM, 2 sum_mask|gt_d-depth|, 3 mask count.
// Separable window with register-resident sliding runs: a thread filters SEG = 4 neighbouring outputs from one run of
// 14 inputs (horizontal pass: 4 x 128-bit shared loads per input map instead of 44 scalar ones; vertical pass: 14
// conflict-free column reads per map for 4 output rows).
__global__ void __launch_bounds__(SSIM_THREADS)
mapping_loss_stats_kernel(int W, int H, const float* __restrict__ c6, const float* __restrict__ gt_rgb,
                          const float* __restrict__ gt_depth, Window win, float* __restrict__ dmaps,
                          float* __restrict__ partials) {
    __shared__ __align__(16) float sx[SHH * SWP], sy[SHH * SWP];
    __shared__ __align__(16) float sh[5][SHH * TW];  // horizontally filtered: x, y, xx, yy, xy  (26 rows x 32 cols)
    __shared__ float s_red[4 * (SSIM_THREADS / 32)];
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const size_t N = (size_t)W * H;
    // vertical-pass ownership: column vc, rows vr0 .. vr0+3
    const int vc = threadIdx.x % TW, vr0 = (threadIdx.x / TW) * SEG;
    const int px = x0 + vc;
    float s_l1 = 0.f, s_ssim = 0.f, s_d = 0.f, s_cnt = 0.f;
    {
        const int ch = blockIdx.z;   // one colour channel per block: three times the blocks, a third of the serial chain
        load_halo(c6 + ch * N, W, H, x0, y0, sx);
        load_halo(gt_rgb + ch * N, W, H, x0, y0, sy);
        __syncthreads();
        for (int it = threadIdx.x; it < SHH * (TW / SEG); it += SSIM_THREADS) {
            const int r = it / (TW / SEG), c = (it - r * (TW / SEG)) * SEG;
            float x[16], y[16], p[16], o[SEG];
            load_run(sx + r * SWP + c, x);
            load_run(sy + r * SWP + c, y);
            window_run(x, win, o);
            *reinterpret_cast<float4*>(&sh[0][r * TW + c]) = make_float4(o[0], o[1], o[2], o[3]);
            window_run(y, win, o);
            *reinterpret_cast<float4*>(&sh[1][r * TW + c]) = make_float4(o[0], o[1], o[2], o[3]);
#pragma unroll
            for (int i = 0; i < 14; i++) p[i] = x[i] * x[i];
            window_run(p, win, o);
            *reinterpret_cast<float4*>(&sh[2][r * TW + c]) = make_float4(o[0], o[1], o[2], o[3]);
#pragma unroll
            for (int i = 0; i < 14; i++) p[i] = y[i] * y[i];
            window_run(p, win, o);
            *reinterpret_cast<float4*>(&sh[3][r * TW + c]) = make_float4(o[0], o[1], o[2], o[3]);
#pragma unroll
            for (int i = 0; i < 14; i++) p[i] = x[i] * y[i];
            window_run(p, win, o);
            *reinterpret_cast<float4*>(&sh[4][r * TW + c]) = make_float4(o[0], o[1], o[2], o[3]);
        }
        __syncthreads();
        float f[5][SEG];
#pragma unroll
        for (int q = 0; q < 5; q++) {
            float v[16];
#pragma unroll
            for (int i = 0; i < 14; i++) v[i] = sh[q][(vr0 + i) * TW + vc];
            window_run(v, win, f[q]);
        }
#pragma unroll
        for (int o = 0; o < SEG; o++) {
            const int py = y0 + vr0 + o;
            if (px < W && py < H) {
                const float mu1 = f[0][o], mu2 = f[1][o], e11 = f[2][o], e22 = f[3][o], e12 = f[4][o];
                const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
                const float sig1 = e11 - mu1_sq, sig2 = e22 - mu2_sq, sig12 = e12 - mu12;
                const float A1 = 2.f * mu12 + SSIM_C1, A2 = 2.f * sig12 + SSIM_C2;
                const float B1 = mu1_sq + mu2_sq + SSIM_C1, B2 = sig1 + sig2 + SSIM_C2;
                const float inv = 1.f / (B1 * B2);
                const float S = A1 * A2 * inv;
                s_ssim += S;
                const size_t off = ch * N + (size_t)py * W + px;
                dmaps[off] = (2.f * mu2 * (A2 - A1)) * inv - S * (2.f * mu1 / B1 - 2.f * mu1 / B2);  // dS/dmu1
                dmaps[3 * N + off] = -S / B2;                                                          // dS/dE[x^2]
                dmaps[6 * N + off] = 2.f * A1 * inv;                                                   // dS/dE[xy]
                const int ci = (vr0 + o + HALO) * SWP + vc + XOFF;
                s_l1 += fabsf(sx[ci] - sy[ci]);
            }
        }
    }
    if (blockIdx.z == 0) {   // the depth term rides with channel 0
#pragma unroll
        for (int o = 0; o < SEG; o++) {
            const int py = y0 + vr0 + o;
            if (px < W && py < H) {
                const size_t p = (size_t)py * W + px;
                const float depth = c6[3 * N + p], dsq = c6[5 * N + p], gd = gt_depth[p];
                const float invis = dsq - depth * depth;
                if ((gd > 0.f) && !isnan(depth) && !isnan(invis)) {
                    s_d += fabsf(gd - depth);
                    s_cnt += 1.f;
                }
            }
        }
    }
    float v[4] = {s_l1, s_ssim, s_d, s_cnt};
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) s_red[k * (SSIM_THREADS / 32) + warp] = x;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        float x = 0.f;
#pragma unroll
        for (int w = 0; w < SSIM_THREADS / 32; w++) x += s_red[threadIdx.x * (SSIM_THREADS / 32) + w];
        partials[(size_t)((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 4 + threadIdx.x] = x;
    }
}

// Totals of the per-block partial sums -> the loss record.  Every block of the gradient kernel calls this (it needs the
// mask count for its own pixels; the partials are a few KB that stay in L2), block 0 writes the record -- a separate
// one-warp finalize launch cost 9 us of serial latency.
// losses[0] weighted total, [1] weighted im, [2] weighted depth, [3] mask count, [4] R, [5] overflow, [6] mean SSIM
__device__ __forceinline__ float mapping_loss_totals(int nblocks, const float* __restrict__ partials, float n3, float w_im,
                                                     float w_depth, float* __restrict__ losses,
                                                     const uint32_t* __restrict__ status, float* s_red /*[4 * warps]*/) {
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    for (int b = threadIdx.x; b < nblocks; b += SSIM_THREADS) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(partials) + b);
        v[0] += q.x; v[1] += q.y; v[2] += q.z; v[3] += q.w;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) s_red[k * (SSIM_THREADS / 32) + warp] = x;
    }
    __syncthreads();
    float tot[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float x = 0.f;
#pragma unroll
        for (int w = 0; w < SSIM_THREADS / 32; w++) x += s_red[k * (SSIM_THREADS / 32) + w];
        tot[k] = x;
    }
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
        const float l1 = tot[0] / n3, ssim = tot[1] / n3;
        const float cnt = tot[3];
        const float li = w_im * (0.8f * l1 + 0.2f * (1.0f - ssim));
        const float ld = w_depth * (tot[2] / cnt);  // mean over an empty mask is NaN in torch as well
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
    return tot[3];
}

__global__ void __launch_bounds__(SSIM_THREADS)
mapping_loss_grad_kernel(int W, int H, const float* __restrict__ c6, const float* __restrict__ gt_rgb,
                         const float* __restrict__ gt_depth, const float* __restrict__ dmaps, Window win, float w_im,
                         float w_depth, int nblocks, const float* __restrict__ partials, float* __restrict__ losses,
                         const uint32_t* __restrict__ status, float* __restrict__ dL) {
    __shared__ __align__(16) float sm[3][SHH * SWP];
    __shared__ __align__(16) float sh[3][SHH * TW];
    __shared__ float s_red[4 * (SSIM_THREADS / 32)];
    // only the blocks that also write the depth gradient (z == 0) need the mask count; block (0,0,0) writes the record
    float mask_count = 1.f;
    if (blockIdx.z == 0)
        mask_count = mapping_loss_totals(nblocks, partials, 3.f * (float)W * (float)H, w_im, w_depth, losses, status, s_red);
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const int vc = threadIdx.x % TW, vr0 = (threadIdx.x / TW) * SEG;
    const int px = x0 + vc;
    const size_t N = (size_t)W * H;
    const float n3 = 3.f * (float)N;
    {
        const int ch = blockIdx.z;
#pragma unroll
        for (int q = 0; q < 3; q++) load_halo(dmaps + (3 * q + ch) * N, W, H, x0, y0, sm[q]);
        __syncthreads();
        for (int it = threadIdx.x; it < SHH * (TW / SEG); it += SSIM_THREADS) {
            const int r = it / (TW / SEG), c = (it - r * (TW / SEG)) * SEG;
#pragma unroll
            for (int q = 0; q < 3; q++) {
                float v[16], o[SEG];
                load_run(sm[q] + r * SWP + c, v);
                window_run(v, win, o);
                *reinterpret_cast<float4*>(&sh[q][r * TW + c]) = make_float4(o[0], o[1], o[2], o[3]);
            }
        }
        __syncthreads();
        float f[3][SEG];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            float v[16];
#pragma unroll
            for (int i = 0; i < 14; i++) v[i] = sh[q][(vr0 + i) * TW + vc];
            window_run(v, win, f[q]);
        }
#pragma unroll
        for (int o = 0; o < SEG; o++) {
            const int py = y0 + vr0 + o;
            if (px < W && py < H) {
                const size_t off = ch * N + (size_t)py * W + px;
                const float x = c6[off], y = gt_rgb[off];
                const float dssim = f[0][o] + 2.f * x * f[1][o] + y * f[2][o];  // d(sum SSIM)/d im
                dL[off] = w_im * (0.8f * sgnf(x - y) - 0.2f * dssim) / n3;
            }
        }
    }
    if (blockIdx.z != 0) return;
#pragma unroll
    for (int o = 0; o < SEG; o++) {
        const int py = y0 + vr0 + o;
        if (px < W && py < H) {
            const size_t p = (size_t)py * W + px;
            const float depth = c6[3 * N + p], dsq = c6[5 * N + p], gd = gt_depth[p];
            const float invis = dsq - depth * depth;
            float g = 0.f;
            if ((gd > 0.f) && !isnan(depth) && !isnan(invis)) g = -w_depth * sgnf(gd - depth) / mask_count;
            dL[3 * N + p] = g;
            dL[4 * N + p] = 0.f;
            dL[5 * N + p] = 0.f;
        }
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
    const size_t bx = (W + TW - 1) / TW, by = (H + TH - 1) / TH;
    return 9 * (size_t)W * H + 3 * bx * by * 4 + 16;
}

void launch_mapping_loss(int W, int H, const float* c6, const float* gt_rgb, const float* gt_depth, float w_im,
                         float w_depth, float* dL, float* losses, float* scratch, const uint32_t* status,
                         cudaStream_t st) {
    const dim3 grid((W + TW - 1) / TW, (H + TH - 1) / TH, 3);
    const int nblocks = grid.x * grid.y * grid.z;
    float* dmaps = scratch;
    float* partials = scratch + ((9 * (size_t)W * H + 3) & ~(size_t)3);   // float4 reads in the gradient kernel
    const Window win = make_window();
    mapping_loss_stats_kernel<<<grid, SSIM_THREADS, 0, st>>>(W, H, c6, gt_rgb, gt_depth, win, dmaps, partials);
    mapping_loss_grad_kernel<<<grid, SSIM_THREADS, 0, st>>>(W, H, c6, gt_rgb, gt_depth, dmaps, win, w_im, w_depth,
                                                           nblocks, partials, losses, status, dL);
}

}  // namespace asb
