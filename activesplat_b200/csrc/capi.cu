// capi.cu -- host orchestration and the C ABI of libactivesplat_b200.so (see include/activesplat_b200.h).
//
//   as_forward   follows CudaRasterizer::Rasterizer::forward   dgr/cuda_rasterizer/rasterizer_impl.cu:198-342
//   as_backward  follows CudaRasterizer::Rasterizer::backward  rasterizer_impl.cu:346-441
//   scratch      follows GeometryState/ImageState/BinningState::fromChunk + required<T>
//                rasterizer_impl.cu:155-194, rasterizer_impl.h:21-72
#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>

#include "common.cuh"

namespace asb {

static thread_local int g_last_cuda_error = 0;

GeometryState GeometryState::fromChunk(char*& chunk, size_t P) {
    GeometryState g;
    Bump b(chunk);
    g.depths = b.take<float>(P);
    g.means2D = b.take<float2>(P);
    g.conic_opacity = b.take<float4>(P);
    g.tiles_touched = b.take<uint32_t>(P);
    g.point_offsets = b.take<uint32_t>(P);
    g.radii = b.take<int>(P);
    g.gacc = b.take<float>(P * GACC_STRIDE);
    g.zsil = b.take<float>(P);
    g.vis_list = b.take<uint32_t>(P);
    g.scan_temp = b.take<uint32_t>((scan_temp_bytes((int)P) + 3) / 4 + 2 * ((P + 255) / 256) + 8);
    g.num_rendered = b.take<uint32_t>(64);
    chunk = b.p;
    return g;
}

ImageState ImageState::fromChunk(char*& chunk, size_t N, size_t tiles) {
    ImageState img;
    Bump b(chunk);
    img.final_T = b.take<float>(N);
    img.n_contrib = b.take<uint32_t>(N);
    img.ranges = b.take<uint2>(tiles);
    img.tile_count = b.take<uint32_t>(tiles * TB_SUB);   // sub-counters per tile (fused.cu)
    img.tile_cursor = b.take<uint32_t>(tiles * TB_SUB);
    img.tile_order = b.take<uint32_t>(tiles);
    chunk = b.p;
    return img;
}

BinningState BinningState::fromChunk(char*& chunk, size_t R, int channels) {
    BinningState s;
    Bump b(chunk);
    const size_t fv = channels == 3 ? 1 : 2;
    s.keys_unsorted = b.take<uint64_t>(R);
    s.keys_sorted = b.take<uint64_t>(R);
    s.values_unsorted = b.take<uint32_t>(R);
    s.point_list = b.take<uint32_t>(R);
    s.inst_conic = b.take<float4>(R);
    s.inst_geom = b.take<float4>(R);
    s.inst_feat = b.take<float4>(R * fv);
    s.sort_temp_bytes = asb::sort_temp_bytes((int)R);
    s.sort_temp = b.take<char>(s.sort_temp_bytes);
    chunk = b.p;
    return s;
}

template <typename F>
static size_t required_bytes(F make) {
    char* p = nullptr;
    make(p);
    return reinterpret_cast<size_t>(p) + 256;
}

static int check(cudaError_t e) {
    if (e == cudaSuccess) return AS_OK;
    g_last_cuda_error = (int)e;
    return AS_ERR_CUDA;
}

// auxiliary.h:166-173 CHECK_CUDA: in debug mode synchronise after the stage and surface failures
static int stage_check(cudaStream_t st, int debug) {
    if (debug) {
        int r = check(cudaStreamSynchronize(st));
        if (r) return r;
    }
    return check(cudaGetLastError());
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---- optional per-stage timing (bench.py's roofline attribution) and launch counting ----------
// Off by default.  When on, every stage is bracketed by two CUDA events recorded on the launching
// stream; as_profile_read() synchronises and sums elapsed times per stage.  Process-wide (torch runs
// backward passes on its own thread); the mutex is only taken while profiling is on -- with it off a
// stage costs two relaxed atomic operations and the library calls stay independent of each other.
struct Profiler {
    std::mutex mu;
    std::atomic<bool> on{false};
    static constexpr int MAX_EV = 8192;
    cudaEvent_t ev[MAX_EV];
    int n_created = 0;
    int n_used = 0;                   // events used since the last read (pairs: 2k, 2k+1)
    int stage_of[MAX_EV / 2];
    std::atomic<long long> launches{0};  // kernels launched by this library (always counted)
};
static Profiler g_prof;

struct StageScope {
    cudaStream_t st;
    int slot = -1;
    StageScope(int stage, cudaStream_t s, int kernels) : st(s) {
        g_prof.launches.fetch_add(kernels, std::memory_order_relaxed);
        if (!g_prof.on.load(std::memory_order_relaxed)) return;
        std::lock_guard<std::mutex> lock(g_prof.mu);
        if (g_prof.n_used + 2 > Profiler::MAX_EV) return;
        while (g_prof.n_created < g_prof.n_used + 2) {
            if (cudaEventCreate(&g_prof.ev[g_prof.n_created]) != cudaSuccess) return;
            g_prof.n_created++;
        }
        slot = g_prof.n_used;
        g_prof.n_used += 2;
        g_prof.stage_of[slot / 2] = stage;
        cudaEventRecord(g_prof.ev[slot], st);
    }
    ~StageScope() {
        if (slot >= 0) cudaEventRecord(g_prof.ev[slot + 1], st);
    }
};

}  // namespace asb

using namespace asb;

extern "C" {

size_t as_geometry_bytes(int P) {
    return required_bytes([&](char*& p) { GeometryState::fromChunk(p, (size_t)(P > 0 ? P : 0)); });
}
size_t as_image_bytes(int width, int height) {
    const size_t N = (size_t)width * height;
    const size_t tiles = (size_t)((width + TILE_X - 1) / TILE_X) * ((height + TILE_Y - 1) / TILE_Y);
    return required_bytes([&](char*& p) { ImageState::fromChunk(p, N, tiles); });
}
size_t as_binning_bytes(int num_rendered, int channels) {
    return required_bytes([&](char*& p) { BinningState::fromChunk(p, (size_t)(num_rendered > 0 ? num_rendered : 0), channels); });
}

int as_scratch_layout(int P, int width, int height, int num_rendered, int channels, AsScratchLayout* out) {
    if (!out || P < 0 || width <= 0 || height <= 0 || num_rendered < 0 || (channels != 3 && channels != 6))
        return AS_ERR_INVALID_ARG;
    char* p = nullptr;
    GeometryState g = GeometryState::fromChunk(p, (size_t)P);
    out->depths = (size_t)g.depths;
    out->means2D = (size_t)g.means2D;
    out->conic_opacity = (size_t)g.conic_opacity;
    out->cov3D = (size_t)-1;  // not stored: the backward recomputes it from scale/rotation
    out->tiles_touched = (size_t)g.tiles_touched;
    out->point_offsets = (size_t)g.point_offsets;
    out->radii = (size_t)g.radii;
    p = nullptr;
    BinningState b = BinningState::fromChunk(p, (size_t)num_rendered, channels);
    out->keys_unsorted = (size_t)b.keys_unsorted;
    out->values_unsorted = (size_t)b.values_unsorted;
    out->keys_sorted = (size_t)b.keys_sorted;
    out->point_list = (size_t)b.point_list;
    p = nullptr;
    const size_t tiles = (size_t)((width + TILE_X - 1) / TILE_X) * ((height + TILE_Y - 1) / TILE_Y);
    ImageState img = ImageState::fromChunk(p, (size_t)width * height, tiles);
    out->final_T = (size_t)img.final_T;
    out->n_contrib = (size_t)img.n_contrib;
    out->ranges = (size_t)img.ranges;
    return AS_OK;
}

uint32_t as_higher_msb(uint32_t n) {  // rasterizer_impl.cu:35-50
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step;
        else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

int as_forward(const AsForwardArgs* a, AsAllocFn geometryBuffer, AsAllocFn binningBuffer, AsAllocFn imageBuffer,
               void* alloc_ctx, void* cuda_stream) {
    if (!a || !geometryBuffer || !binningBuffer || !imageBuffer) return AS_ERR_INVALID_ARG;
    if (a->P < 0 || a->width <= 0 || a->height <= 0 || (a->channels != 3 && a->channels != 6)) return AS_ERR_INVALID_ARG;
    if (!a->out_color || !a->out_depth || !a->out_opacity || !a->background) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    const int P = a->P, W = a->width, H = a->height, C = a->channels;
    const size_t N = (size_t)W * H;
    if (P == 0) {  // rasterize_points.cu:83: nothing is launched, outputs stay at their fill value 0
        int r = check(cudaMemsetAsync(a->out_color, 0, sizeof(float) * C * N, st));
        if (!r) r = check(cudaMemsetAsync(a->out_depth, 0, sizeof(float) * N, st));
        if (!r) r = check(cudaMemsetAsync(a->out_opacity, 0, sizeof(float) * N, st));
        return r;
    }
    if (!a->colors_precomp) return AS_ERR_NO_COLORS;  // rasterizer_impl.cu:244-247
    if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix) return AS_ERR_INVALID_ARG;
    if (!a->cov3D_precomp && (!a->scales || !a->rotations)) return AS_ERR_INVALID_ARG;
    if (a->rotations && !aligned16(a->rotations)) return AS_ERR_INVALID_ARG;

    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;

    char* gchunk = geometryBuffer(alloc_ctx, as_geometry_bytes(P));
    if (!gchunk) return AS_ERR_ALLOC;
    GeometryState g = GeometryState::fromChunk(gchunk, (size_t)P);
    char* ichunk = imageBuffer(alloc_ctx, as_image_bytes(W, H));
    if (!ichunk) return AS_ERR_ALLOC;
    ImageState img = ImageState::fromChunk(ichunk, N, (size_t)tiles);

    int r;
    {
    StageScope _s(AS_STAGE_PREPROCESS, st, 1);
    launch_preprocess_forward(P, a->means3D, a->scales, a->rotations, a->opacities, a->cov3D_precomp, a->viewmatrix,
                              a->projmatrix, W, H, a->tan_fovx, a->tan_fovy, a->scale_modifier, a->radii, g, st);
    }
    if ((r = stage_check(st, a->debug))) return r;
    {
    StageScope _s(AS_STAGE_SCAN, st, 1);
    launch_inclusive_scan(g.tiles_touched, g.point_offsets, P, g.num_rendered, g.scan_temp, st);
    }
    if ((r = stage_check(st, a->debug))) return r;

    // The one host round trip of the reference (rasterizer_impl.cu:283): R sizes the binning buffer.
    uint32_t R_host = 0;
    if ((r = check(cudaMemcpyAsync(&R_host, g.num_rendered, sizeof(uint32_t), cudaMemcpyDeviceToHost, st)))) return r;
    if ((r = check(cudaStreamSynchronize(st)))) return r;
    const int R = (int)R_host;

    char* bchunk = binningBuffer(alloc_ctx, as_binning_bytes(R, C));
    if (!bchunk) return AS_ERR_ALLOC;
    BinningState b = BinningState::fromChunk(bchunk, (size_t)R, C);

    if (R > 0) {
        {
        StageScope _s(AS_STAGE_DUPLICATE, st, 1);
        launch_duplicate_with_keys(P, g, g.radii, W, H, b.keys_unsorted, b.values_unsorted, st);
        }
        if ((r = stage_check(st, a->debug))) return r;
        const int bit = (int)as_higher_msb((uint32_t)tiles);
        {
        StageScope _s(AS_STAGE_SORT, st, 4 + (32 + bit + 7) / 8);
        launch_sort_pairs(b.keys_unsorted, b.values_unsorted, b.keys_sorted, b.point_list, R, 32 + bit, b.sort_temp, st);
        }
        if ((r = stage_check(st, a->debug))) return r;
    }
    {
    StageScope _s(AS_STAGE_RANGES_GATHER, st, R > 0 ? 1 : 0);
    launch_ranges_and_gather(R, tiles, C, b.keys_sorted, b.point_list, g, a->colors_precomp, b, img.ranges, st);
    }
    if ((r = stage_check(st, a->debug))) return r;
    {
    StageScope _s(AS_STAGE_BLEND_FWD, st, 1);
    launch_blend_forward(C, W, H, img.ranges, b, a->background, a->out_color, a->out_depth, a->out_opacity,
                         img.final_T, img.n_contrib, st);
    }
    if ((r = stage_check(st, a->debug))) return r;
    return R;
}

int as_backward(const AsBackwardArgs* a, void* cuda_stream) {
    if (!a) return AS_ERR_INVALID_ARG;
    if (a->P < 0 || a->R < 0 || a->width <= 0 || a->height <= 0 || (a->channels != 3 && a->channels != 6))
        return AS_ERR_INVALID_ARG;
    if (a->P == 0) return AS_OK;
    if (!a->geom_buffer || !a->image_buffer || !a->dL_dpix || !a->means3D || !a->viewmatrix || !a->projmatrix ||
        !a->background)
        return AS_ERR_INVALID_ARG;
    if (a->R > 0 && !a->binning_buffer) return AS_ERR_INVALID_ARG;
    if (!a->cov3D_precomp && (!a->scales || !a->rotations)) return AS_ERR_INVALID_ARG;
    if ((a->rotations && !aligned16(a->rotations)) || (a->dL_drot && !aligned16(a->dL_drot)) ||
        (a->dL_dconic && !aligned16(a->dL_dconic)))
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    const int P = a->P, W = a->width, H = a->height, C = a->channels, R = a->R;
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;

    char* gchunk = a->geom_buffer;
    GeometryState g = GeometryState::fromChunk(gchunk, (size_t)P);
    char* ichunk = a->image_buffer;
    ImageState img = ImageState::fromChunk(ichunk, (size_t)W * H, (size_t)gx * gy);
    const int* radii = a->radii ? a->radii : g.radii;

    int r;
    if ((r = check(cudaMemsetAsync(g.gacc, 0, sizeof(float) * GACC_STRIDE * (size_t)P, st)))) return r;
    if (R > 0) {
        char* bchunk = a->binning_buffer;
        BinningState b = BinningState::fromChunk(bchunk, (size_t)R, C);
        {
        StageScope _s(AS_STAGE_BLEND_BWD, st, 1);
        launch_blend_backward(C, W, H, img.ranges, b, a->background, img.final_T, img.n_contrib, a->dL_dpix, g.gacc, st);
        }
        if ((r = stage_check(st, a->debug))) return r;
    }
    PreprocessBackwardOut o;
    o.dL_dmean2D = a->dL_dmean2D;
    o.dL_dconic = a->dL_dconic;
    o.dL_dopacity = a->dL_dopacity;
    o.dL_dcolor = a->dL_dcolor;
    o.dL_dmean3D = a->dL_dmean3D;
    o.dL_dcov3D = a->dL_dcov3D;
    o.dL_dscale = a->dL_dscale;
    o.dL_drot = a->dL_drot;
    {
    StageScope _s(AS_STAGE_PREPROCESS_BWD, st, 1);
    launch_preprocess_backward(P, C, a->means3D, radii, a->scales, a->rotations, a->cov3D_precomp, a->scale_modifier,
                               a->viewmatrix, a->projmatrix, W, H, a->tan_fovx, a->tan_fovy, nullptr, g.gacc, o, st);
    }
    return stage_check(st, a->debug);
}

// ---- fused frame path: no host synchronisation, caller-sized binning buffer ---------------------
int as_frame_forward(const AsFrameArgs* a, void* cuda_stream) {
    if (!a || a->P <= 0 || a->width <= 0 || a->height <= 0 || a->binning_capacity < 0) return AS_ERR_INVALID_ARG;
    if (!a->means3D || !a->rgb_colors || !a->unnorm_rotations || !a->logit_opacities || !a->log_scales ||
        !a->cam_unnorm_rot || !a->cam_trans || !a->viewmatrix || !a->projmatrix || !a->background ||
        !a->geom_buffer || !a->image_buffer || !a->binning_buffer || !a->status || !a->out_color || !a->out_depth ||
        !a->out_opacity)
        return AS_ERR_INVALID_ARG;
    if (!aligned16(a->unnorm_rotations)) return AS_ERR_INVALID_ARG;
    // the preprocess look-back word carries the visible count in 24 bits next to its flag bits
    if (a->P >= (1 << 24)) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    const int P = a->P, W = a->width, H = a->height;
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y, tiles = gx * gy;
    char* gchunk = a->geom_buffer;
    GeometryState g = GeometryState::fromChunk(gchunk, (size_t)P);
    char* ichunk = a->image_buffer;
    ImageState img = ImageState::fromChunk(ichunk, (size_t)W * H, (size_t)tiles);
    char* bchunk = a->binning_buffer;
    const int cap = a->binning_capacity;
    BinningState b = BinningState::fromChunk(bchunk, (size_t)cap, 6);
    int r;
    const bool buckets = a->binning_mode == AS_BINNING_TILE_BUCKETS;
    if (buckets) cudaMemsetAsync(img.tile_count, 0, sizeof(uint32_t) * (size_t)tiles * TB_SUB, st);
    {
        StageScope _s(AS_STAGE_PREPROCESS, st, 1);
        launch_frame_preprocess(*a, g, a->status, buckets ? img.tile_count : nullptr, st);
    }
    if ((r = stage_check(st, a->debug))) return r;
    if (buckets) {
        // tile-bucket binning: scan of per-tile counts, scatter, per-tile sort fused with the record gather
        StageScope _s(AS_STAGE_SORT, st, cap > 0 ? 3 : 1);
        launch_frame_tile_binning(P, g, W, H, (uint32_t)cap, a->rgb_colors, b, img, a->status, st);
    } else {
        if (cap > 0) {
            {
                StageScope _s(AS_STAGE_DUPLICATE, st, 1);
                launch_frame_duplicate(P, g, W, H, (uint32_t)cap, b.keys_unsorted, b.values_unsorted, a->status, st);
            }
            if ((r = stage_check(st, a->debug))) return r;
            const int bit = (int)as_higher_msb((uint32_t)tiles);
            {
                StageScope _s(AS_STAGE_SORT, st, 4 + (32 + bit + 7) / 8);
                launch_sort_pairs(b.keys_unsorted, b.values_unsorted, b.keys_sorted, b.point_list, cap, 32 + bit,
                                  b.sort_temp, st, a->status);
            }
            if ((r = stage_check(st, a->debug))) return r;
        }
        {
            StageScope _s(AS_STAGE_RANGES_GATHER, st, 1);
            launch_frame_ranges_gather(a->status, (uint32_t)cap, tiles, g, a->rgb_colors, b, img.ranges, st);
        }
    }
    if ((r = stage_check(st, a->debug))) return r;
    if (a->chunk_scratch) {   // crowded tiles: the lists are blended in chunks, one CTA per chunk (forward-only renders)
        StageScope _s(AS_STAGE_BLEND_FWD, st, 4);
        launch_blend_forward_chunked(W, H, img.ranges, b, cap, a->background, a->out_color, a->out_depth, a->out_opacity,
                                     img.final_T, img.n_contrib, a->chunk_scratch, st);
    } else {
        StageScope _s(AS_STAGE_BLEND_FWD, st, 1);
        launch_blend_forward(6, W, H, img.ranges, b, a->background, a->out_color, a->out_depth, a->out_opacity,
                             img.final_T, img.n_contrib, st, buckets ? img.tile_order : nullptr, true);
    }
    return stage_check(st, a->debug);
}

size_t as_chunk_scratch_bytes(int binning_capacity, int width, int height) {
    const int tiles = ((width + TILE_X - 1) / TILE_X) * ((height + TILE_Y - 1) / TILE_Y);
    return blend_chunk_scratch_bytes(binning_capacity > 0 ? binning_capacity : 0, tiles);
}

int as_frame_backward(const AsFrameBackwardArgs* ba, void* cuda_stream) {
    if (!ba || !ba->frame || !ba->dL_dpix) return AS_ERR_INVALID_ARG;
    const AsFrameArgs* a = ba->frame;
    if (a->P <= 0 || !a->geom_buffer || !a->image_buffer || !a->binning_buffer) return AS_ERR_INVALID_ARG;
    if (ba->camera_grad && (!ba->grad_cam || !ba->pose_scratch)) return AS_ERR_INVALID_ARG;
    if (ba->gaussians_grad && !ba->grad_params) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    const int P = a->P, W = a->width, H = a->height;
    const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y, tiles = gx * gy;
    char* gchunk = a->geom_buffer;
    GeometryState g = GeometryState::fromChunk(gchunk, (size_t)P);
    char* ichunk = a->image_buffer;
    ImageState img = ImageState::fromChunk(ichunk, (size_t)W * H, (size_t)tiles);
    char* bchunk = a->binning_buffer;
    BinningState b = BinningState::fromChunk(bchunk, (size_t)a->binning_capacity, 6);
    int r;
    {
        StageScope _s(AS_STAGE_BLEND_BWD, st, 1);
        launch_blend_backward(6, W, H, img.ranges, b, a->background, img.final_T, img.n_contrib, ba->dL_dpix, g.gacc, st,
                              ba->gaussians_grad ? 1 : 2,
                              a->binning_mode == AS_BINNING_TILE_BUCKETS ? img.tile_order : nullptr);
    }
    if ((r = stage_check(st, a->debug))) return r;
    {
        StageScope _s(AS_STAGE_PREPROCESS_BWD, st, ba->camera_grad ? 2 : 1);
        launch_frame_preprocess_backward(*a, g, ba->grad_scale, ba->gaussians_grad, ba->camera_grad, ba->accumulate,
                                         ba->grad_params, ba->grad_means2D, ba->pose_scratch, ba->grad_cam,
                                         ba->grad_cam_stride, st);
    }
    return stage_check(st, a->debug);
}

size_t as_pose_scratch_floats(int P) { return pose_scratch_floats(P > 0 ? P : 0); }
int as_tile_sort_capacity(void) { return tile_sort_capacity(); }

size_t as_loss_scratch_floats(int width, int height) { return (size_t)loss_partial_blocks(width * height) * 4 + 16; }

int as_tracking_loss(int width, int height, const float* color6, const float* gt_rgb, const float* gt_depth,
                     float w_im, float w_depth, float sil_thres, int use_sil_for_loss, int mask_color,
                     float* dL_dcolor6, float* losses, float* scratch, const uint32_t* frame_status,
                     void* cuda_stream) {
    if (width <= 0 || height <= 0 || !color6 || !gt_rgb || !gt_depth || !dL_dcolor6 || !losses || !scratch)
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    StageScope _s(AS_STAGE_LOSS, st, 2);
    launch_tracking_loss(width * height, color6, gt_rgb, gt_depth, w_im, w_depth, sil_thres, use_sil_for_loss,
                         mask_color, dL_dcolor6, scratch, losses, frame_status, st);
    return check(cudaGetLastError());
}

size_t as_mapping_loss_scratch_floats(int width, int height) { return mapping_loss_scratch_floats(width, height); }

int as_mapping_loss(int width, int height, const float* color6, const float* gt_rgb, const float* gt_depth,
                    float w_im, float w_depth, float* dL_dcolor6, float* losses, float* scratch,
                    const uint32_t* frame_status, void* cuda_stream) {
    if (width <= 0 || height <= 0 || !color6 || !gt_rgb || !gt_depth || !dL_dcolor6 || !losses || !scratch)
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    StageScope _s(AS_STAGE_LOSS, st, 2);
    launch_mapping_loss(width, height, color6, gt_rgb, gt_depth, w_im, w_depth, dL_dcolor6, losses, scratch,
                        frame_status, st);
    return check(cudaGetLastError());
}

int as_adam_step(int P, int isotropic, float* means3D, float* rgb_colors, float* unnorm_rotations,
                 float* logit_opacities, float* log_scales, const float* grad_params, float* exp_avg,
                 float* exp_avg_sq, const float* lrs, float beta1, float beta2, float eps, int step,
                 void* cuda_stream) {
    if (P < 0 || step < 1 || !lrs) return AS_ERR_INVALID_ARG;
    if (P == 0) return AS_OK;
    if (!means3D || !rgb_colors || !unnorm_rotations || !logit_opacities || !log_scales || !grad_params || !exp_avg ||
        !exp_avg_sq)
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    float* params[5] = {means3D, rgb_colors, unnorm_rotations, logit_opacities, log_scales};
    StageScope _s(AS_STAGE_ADAM, st, 1);
    launch_adam(P, isotropic, params, lrs, grad_params, exp_avg, exp_avg_sq, beta1, beta2, eps, step, st);
    return check(cudaGetLastError());
}

int as_adam_step_counted(int P, int isotropic, float* means3D, float* rgb_colors, float* unnorm_rotations,
                         float* logit_opacities, float* log_scales, const float* grad_params, float* exp_avg,
                         float* exp_avg_sq, const float* lrs, float beta1, float beta2, float eps, float* step_state,
                         void* cuda_stream) {
    if (P < 0 || !lrs || !step_state) return AS_ERR_INVALID_ARG;
    if (P > 0 && (!means3D || !rgb_colors || !unnorm_rotations || !logit_opacities || !log_scales || !grad_params ||
                  !exp_avg || !exp_avg_sq))
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    float* params[5] = {means3D, rgb_colors, unnorm_rotations, logit_opacities, log_scales};
    StageScope _s(AS_STAGE_ADAM, st, 1);
    launch_adam_counted(P, isotropic, params, lrs, grad_params, exp_avg, exp_avg_sq, beta1, beta2, eps, step_state, 0,
                        ~(size_t)0, 1, st);
    return check(cudaGetLastError());
}

int as_adam_step_range(int P, int isotropic, float* means3D, float* rgb_colors, float* unnorm_rotations,
                       float* logit_opacities, float* log_scales, const float* grad_params, float* exp_avg,
                       float* exp_avg_sq, const float* lrs, float beta1, float beta2, float eps, float* step_state,
                       long long e_begin, long long e_end, int advance_step, void* cuda_stream) {
    if (P < 0 || !lrs || !step_state || e_begin < 0 || e_end < e_begin || (e_begin & 3)) return AS_ERR_INVALID_ARG;
    if (P > 0 && (!means3D || !rgb_colors || !unnorm_rotations || !logit_opacities || !log_scales || !grad_params ||
                  !exp_avg || !exp_avg_sq))
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    float* params[5] = {means3D, rgb_colors, unnorm_rotations, logit_opacities, log_scales};
    StageScope _s(AS_STAGE_ADAM, st, 1);
    launch_adam_counted(P, isotropic, params, lrs, grad_params, exp_avg, exp_avg_sq, beta1, beta2, eps, step_state,
                        (size_t)e_begin, (size_t)e_end, advance_step, st);
    return check(cudaGetLastError());
}

int as_pose_adam_step(float* cam_unnorm_rots, const float* grad_rots, int n_rots, float lr_rots, float* cam_trans,
                      const float* grad_trans, int n_trans, float lr_trans, float* exp_avg, float* exp_avg_sq,
                      float beta1, float beta2, float eps, float* step_state, void* cuda_stream) {
    if (n_rots < 0 || n_trans < 0 || !step_state) return AS_ERR_INVALID_ARG;
    if ((n_rots > 0 && (!cam_unnorm_rots || !grad_rots)) || (n_trans > 0 && (!cam_trans || !grad_trans)) ||
        (n_rots + n_trans > 0 && (!exp_avg || !exp_avg_sq)))
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    StageScope _s(AS_STAGE_ADAM, st, 2);
    launch_pose_adam(cam_unnorm_rots, grad_rots, n_rots, lr_rots, cam_trans, grad_trans, n_trans, lr_trans, exp_avg,
                     exp_avg_sq, beta1, beta2, eps, step_state, st);
    return check(cudaGetLastError());
}

int as_enable_peer_access(int peer_device) {
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess) return AS_ERR_CUDA;
    if (cur == peer_device) return AS_OK;
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, cur, peer_device) != cudaSuccess || !can) return AS_ERR_INVALID_ARG;
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();  // clear the sticky-looking (but harmless) error
        return AS_OK;
    }
    return check(e);
}

int as_ipc_open(const void* handle64, void** out_ptr) {
    if (!handle64 || !out_ptr) return AS_ERR_INVALID_ARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    return check(cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
}

int as_ipc_close(void* ptr) { return ptr ? check(cudaIpcCloseMemHandle(ptr)) : AS_OK; }

int as_adam_step_peer(int P, int isotropic, int world, int rank, int e_begin, int e_end, float* const* params,
                      const float* const* grads, float* exp_avg, float* exp_avg_sq, const float* lrs, float beta1,
                      float beta2, float eps, int step, void* cuda_stream) {
    if (P < 0 || world < 1 || world > 8 || rank < 0 || rank >= world || step < 1 || !lrs || !params || !grads)
        return AS_ERR_INVALID_ARG;
    const long long n_flat = (long long)(isotropic ? 12 : 14) * P;
    if (n_flat > 0x7fffffffLL || e_begin < 0 || e_end < e_begin || e_end > n_flat) return AS_ERR_INVALID_ARG;
    if (e_end == e_begin) return AS_OK;
    if (!exp_avg || !exp_avg_sq) return AS_ERR_INVALID_ARG;
    for (int i = 0; i < 5 * world; i++)
        if (!params[i]) return AS_ERR_INVALID_ARG;
    for (int r = 0; r < world; r++)
        if (!grads[r]) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    StageScope _s(AS_STAGE_ADAM, st, 1);
    launch_adam_peer(P, e_begin, e_end, isotropic, world, rank, params, grads, lrs, exp_avg, exp_avg_sq, beta1, beta2,
                     eps, step, st);
    return check(cudaGetLastError());
}

int as_dbscan_grid(int batch, int height, int width, const float* values, float value_threshold, float eps,
                   int min_samples, int* labels, int* scratch, int max_clusters, int* n_clusters, int* stats_int,
                   double* stats_value, void* cuda_stream) {
    if (batch < 0 || height <= 0 || width <= 0 || max_clusters < 0 || min_samples < 1 || !(eps >= 0.f) || eps > 8.f)
        return AS_ERR_INVALID_ARG;
    if (batch == 0) return AS_OK;
    if (!values || !labels || !scratch || !n_clusters || (max_clusters > 0 && (!stats_int || !stats_value)))
        return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    StageScope _s(AS_STAGE_OTHER, st, 1);
    launch_dbscan_grid(batch, height, width, values, value_threshold, eps, min_samples, labels, scratch, max_clusters,
                       n_clusters, stats_int, stats_value, st);
    return check(cudaGetLastError());
}

size_t as_densify_scratch_words(int width, int height) { return densify_scratch_words(width, height); }

int as_densify(const AsDensifyArgs* a, void* cuda_stream) {
    if (!a || a->width <= 0 || a->height <= 0 || a->P < 0 || a->capacity < a->P) return AS_ERR_INVALID_ARG;
    if (!a->color6 || !a->gt_rgb || !a->gt_depth || !a->c2w || !a->means3D || !a->rgb_colors || !a->unnorm_rotations ||
        !a->logit_opacities || !a->log_scales || !a->scratch || !a->result)
        return AS_ERR_INVALID_ARG;
    if (!(a->fx > 0.f) || !(a->fy > 0.f)) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    StageScope _s(AS_STAGE_OTHER, st, 7);
    launch_densify(a->width, a->height, a->color6, a->gt_rgb, a->gt_depth, a->sil_thres, a->fx, a->fy, a->cx, a->cy,
                   a->c2w, a->isotropic, a->time_idx, a->P, a->capacity, a->means3D, a->rgb_colors, a->unnorm_rotations,
                   a->logit_opacities, a->log_scales, a->timestep, a->scratch, a->result, st);
    return check(cudaGetLastError());
}

int as_cluster_hull_volumes(int batch, int height, int width, const int* labels, const int* n_clusters,
                            const float* depth, double vfov_deg, int max_clusters, double* volumes, int* n_points,
                            void* cuda_stream) {
    if (batch < 0 || height <= 0 || width <= 0 || width > 32000 || height > 32000 || max_clusters < 0 || !(vfov_deg > 0))
        return AS_ERR_INVALID_ARG;
    if (batch == 0 || max_clusters == 0) return AS_OK;
    if (!labels || !n_clusters || !depth || !volumes || !n_points) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    StageScope _s(AS_STAGE_OTHER, st, 1);
    if (launch_cluster_hull(batch, height, width, labels, n_clusters, depth, vfov_deg, max_clusters, volumes, n_points, st))
        return AS_ERR_INVALID_ARG;   // the image's two bit planes do not fit in shared memory
    return check(cudaGetLastError());
}

int as_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* /*projmatrix*/,
                    uint8_t* present, void* cuda_stream) {
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    launch_mark_visible(P, means3D, viewmatrix, present, st);
    return check(cudaGetLastError());
}

size_t as_sort_temp_bytes(int n) { return sort_temp_bytes(n); }

int as_sort_pairs(const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out, uint32_t* values_out,
                  int n, int end_bit, void* temp, size_t temp_bytes, void* cuda_stream) {
    if (n < 0 || end_bit < 0 || end_bit > 64) return AS_ERR_INVALID_ARG;
    if (n == 0) return AS_OK;
    if (!keys_in || !values_in || !keys_out || !values_out || !temp) return AS_ERR_INVALID_ARG;
    if (temp_bytes < sort_temp_bytes(n)) return AS_ERR_CAPACITY;
    if (reinterpret_cast<uintptr_t>(temp) & 255) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    launch_sort_pairs(keys_in, values_in, keys_out, values_out, n, end_bit, reinterpret_cast<char*>(temp), st);
    return check(cudaGetLastError());
}

size_t as_scan_temp_bytes(int n) { return scan_temp_bytes(n); }

int as_inclusive_scan_u32(const uint32_t* in, uint32_t* out, int n, uint32_t* total_out, void* temp,
                          size_t temp_bytes, void* cuda_stream) {
    if (n < 0) return AS_ERR_INVALID_ARG;
    if (n > 0 && (!in || !out || !temp)) return AS_ERR_INVALID_ARG;
    if (temp_bytes < scan_temp_bytes(n)) return AS_ERR_CAPACITY;
    if (reinterpret_cast<uintptr_t>(temp) & 15) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    launch_inclusive_scan(in, out, n, total_out, reinterpret_cast<uint32_t*>(temp), st);
    return check(cudaGetLastError());
}

void as_profile_enable(int on) {
    std::lock_guard<std::mutex> lock(g_prof.mu);
    g_prof.on.store(on != 0);
}

int as_profile_read(double* ms_per_stage, int* calls_per_stage, int n_stages) {
    if (!ms_per_stage || !calls_per_stage || n_stages < AS_NUM_STAGES) return AS_ERR_INVALID_ARG;
    for (int i = 0; i < n_stages; i++) { ms_per_stage[i] = 0.0; calls_per_stage[i] = 0; }
    std::lock_guard<std::mutex> lock(g_prof.mu);
    int r = AS_OK;
    for (int s = 0; s + 1 < g_prof.n_used; s += 2) {
        if (cudaEventSynchronize(g_prof.ev[s + 1]) != cudaSuccess) { r = AS_ERR_CUDA; break; }
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, g_prof.ev[s], g_prof.ev[s + 1]) != cudaSuccess) { r = AS_ERR_CUDA; break; }
        const int stage = g_prof.stage_of[s / 2];
        ms_per_stage[stage] += ms;
        calls_per_stage[stage] += 1;
    }
    g_prof.n_used = 0;
    if (r) g_last_cuda_error = (int)cudaGetLastError();
    return r;
}

long long as_launch_count(void) { return g_prof.launches.load(std::memory_order_relaxed); }

int as_static_scan_max_n(void) { return static_scan_max_n(); }

int as_inclusive_scan_u32_static(const uint32_t* in, uint32_t* out, int n, uint32_t* total_out, void* cuda_stream) {
    if (n < 0) return AS_ERR_INVALID_ARG;
    if (n > static_scan_max_n()) return AS_ERR_STATIC_LIMIT;
    if (n > 0 && (!in || !out)) return AS_ERR_INVALID_ARG;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    void* scratch = nullptr;
    int r = check(static_scan_scratch(&scratch));  // first call on a device: symbol lookup, then cached
    if (r) return r;
    // The array still holds the previous call's ticket and tile states; launch_inclusive_scan clears
    // the part this call uses (cudaMemsetAsync on `st`) before the kernel reads it.
    launch_inclusive_scan(in, out, n, total_out, reinterpret_cast<uint32_t*>(scratch), st);
    return check(cudaGetLastError());
}

const char* as_error_string(int code) {
    switch (code) {
        case AS_ERR_STATIC_LIMIT:
            return "input larger than the built-in scratch allows: as_inclusive_scan_u32_static handles at most "
                   "4194304 elements (see as_static_scan_max_n); use as_inclusive_scan_u32 with caller scratch";
        case AS_OK: return "ok";
        case AS_ERR_INVALID_ARG: return "invalid argument";
        case AS_ERR_NO_COLORS: return "For non-RGB, provide precomputed Gaussian colors!";
        case AS_ERR_ALLOC: return "scratch allocator returned NULL";
        case AS_ERR_CUDA: return cudaGetErrorString((cudaError_t)g_last_cuda_error);
        case AS_ERR_CAPACITY: return "buffer capacity too small";
        default: return code > 0 ? "ok" : "unknown error";
    }
}

int as_last_cuda_error(void) { return g_last_cuda_error; }

const char* as_version(void) { return "activesplat_b200 0.1 (sm_100a)"; }

}  // extern "C"
