// scan_sort.cu -- hand-written single-pass prefix sum and onesweep LSD radix sort (sm_100a).
//
//   inclusive_scan_kernel   replaces cub::DeviceScan::InclusiveSum       dgr/cuda_rasterizer/rasterizer_impl.cu:279
//   radix_*_kernel          replace  cub::DeviceRadixSort::SortPairs     rasterizer_impl.cu:305-310
//
// Sort: keys are u64 (tile << 32 | depth bits), values u32 (Gaussian id).  Stable, ascending, on
// bits [0, end_bit), 8 bits per pass.  One upfront histogram kernel counts every pass's digits in a
// single read of the keys; each pass is then ONE kernel ("onesweep"): a tile of keys is ranked in
// shared memory, its per-digit counts are published, and the exclusive prefix over earlier tiles is
// obtained by decoupled look-back, so keys are read once and written once per pass.
// Passes whose digit is identical for all keys (one histogram bin holds everything) are detected
// on the device and skipped by a pointer swap that is also resolved on the device, so no host
// synchronisation is needed anywhere.
#include <mutex>

#include "common.cuh"

namespace asb {

namespace {

// ------------------------------------------------------------------------------------------------
// inclusive scan, decoupled look-back
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;
constexpr uint64_t FLAG_AGG = 1ull << 32;
constexpr uint64_t FLAG_PREFIX = 2ull << 32;

// temp layout: [0] ticket (u32), [2..] u64 status per tile (flag << 32 | value)
__global__ void __launch_bounds__(SCAN_THREADS)
inclusive_scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n,
                      uint32_t* __restrict__ total_out, uint32_t* ticket, uint64_t* status) {
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    __shared__ uint32_t s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    if (base + SCAN_ITEMS <= n && (reinterpret_cast<uintptr_t>(in) & 15) == 0) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(in + base));
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(in + base) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < n) ? __ldg(in + base + i) : 0u;
    }
    uint32_t run = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { run += v[i]; v[i] = run; }
    // block-wide exclusive scan of the per-thread totals
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) {
        const uint32_t c = s_warp[w];
        if (w < warp) warp_off += c;
        block_total += c;
    }
    const uint32_t thread_excl = warp_off + incl - run;
    if (threadIdx.x == 0) {
        uint32_t prefix = 0;
        if (tile == 0) {
            st_volatile_u64(status, FLAG_PREFIX | block_total);
        } else {
            st_volatile_u64(status + tile, FLAG_AGG | block_total);
            int p = (int)tile - 1;
            while (true) {
                uint64_t s;
                do { s = ld_volatile_u64(status + p); } while ((s >> 32) == 0);
                prefix += (uint32_t)s;
                if ((s >> 32) == 2) break;
                p--;
            }
            st_volatile_u64(status + tile, FLAG_PREFIX | (uint64_t)(prefix + block_total));
        }
        s_prefix = prefix;
    }
    __syncthreads();
    const uint32_t off = s_prefix + thread_excl;
    if (base + SCAN_ITEMS <= n && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
        reinterpret_cast<uint4*>(out + base)[0] = make_uint4(v[0] + off, v[1] + off, v[2] + off, v[3] + off);
        reinterpret_cast<uint4*>(out + base)[1] = make_uint4(v[4] + off, v[5] + off, v[6] + off, v[7] + off);
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++)
            if (base + i < n) out[base + i] = v[i] + off;
    }
    if (total_out && base <= n - 1 && n - 1 < base + SCAN_ITEMS) *total_out = v[n - 1 - base] + off;
}

// ------------------------------------------------------------------------------------------------
// onesweep radix sort
// ------------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_KPT = 8;  // keys per thread
constexpr int RS_TILE = RS_THREADS * RS_KPT;
constexpr int RS_BINS = 256;
constexpr int RS_MAX_PASSES = 8;
constexpr uint32_t ST_AGG = 1u << 30;
constexpr uint32_t ST_PREFIX = 2u << 30;
constexpr uint32_t ST_MASK = (1u << 30) - 1;

// temp layout (u32 words):
//   hist   [RS_MAX_PASSES][256]   digit counts, later exclusive digit offsets
//   ticket [RS_MAX_PASSES]
//   skip   [RS_MAX_PASSES]        1 = every key has the same digit in this pass
//   status [passes][tiles][256]
struct SortTemp {
    uint32_t* hist;
    uint32_t* ticket;
    uint32_t* skip;
    uint32_t* status;
};
__host__ __device__ inline size_t sort_header_words() { return RS_MAX_PASSES * RS_BINS + 2 * RS_MAX_PASSES; }

__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(const uint64_t* __restrict__ keys, int n, const uint32_t* __restrict__ n_dev, int passes,
                  uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[RS_MAX_PASSES * RS_BINS];
    if (n_dev) n = min(n, (int)*n_dev);  // device-side count (no-sync path); n is then the capacity
    for (int i = threadIdx.x; i < passes * RS_BINS; i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    const int stride = gridDim.x * RS_THREADS;
    for (int i = blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
        const uint64_t k = __ldg(keys + i);
        for (int p = 0; p < passes; p++) atomicAdd(&sh[p * RS_BINS + (uint32_t)((k >> (8 * p)) & 0xff)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RS_BINS; i += RS_THREADS) {
        const uint32_t c = sh[i];
        if (c) atomicAdd(hist + i, c);
    }
}

// one block per pass: counts -> exclusive offsets; flag passes that do not move anything
__global__ void __launch_bounds__(RS_BINS)
radix_scan_kernel(uint32_t* hist, uint32_t* skip, int n, const uint32_t* __restrict__ n_dev) {
    __shared__ uint32_t s_warp[RS_BINS / 32];
    if (n_dev) n = min(n, (int)*n_dev);
    const int p = blockIdx.x;
    const uint32_t c = hist[p * RS_BINS + threadIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    const int any_full = __syncthreads_or(c == (uint32_t)n);
    uint32_t off = 0;
    for (int w = 0; w < warp; w++) off += s_warp[w];
    hist[p * RS_BINS + threadIdx.x] = off + incl - c;
    if (threadIdx.x == 0) skip[p] = any_full ? 1u : 0u;
}

// Which of the two ping-pong buffers holds the input of pass `pass`: the caller's input for the
// first non-skipped pass, then alternating.  Evaluated identically by every thread from skip[].
struct PingPong {
    const uint64_t* kin; const uint32_t* vin; uint64_t* kout; uint32_t* vout;
};
__device__ __forceinline__ PingPong resolve_buffers(const uint32_t* __restrict__ skip, int pass,
                                                    const uint64_t* k_src, const uint32_t* v_src,
                                                    uint64_t* k_a, uint32_t* v_a, uint64_t* k_b, uint32_t* v_b) {
    // k_a/v_a is the final destination; the number of real passes from `pass` (inclusive) to the end decides
    // which buffer this pass must write so that the last real pass lands in k_a.
    PingPong r;
    int done_before = 0;
    for (int q = 0; q < pass; q++) done_before += skip[q] ? 0 : 1;
    int remaining_after = 0;
    for (int q = pass + 1; q < RS_MAX_PASSES; q++) remaining_after += (skip[q] == 0) ? 1 : 0;  // skip[] beyond `passes` is preset to 1
    // output of this pass: k_a if an even number of real passes follow, else k_b
    r.kout = (remaining_after & 1) ? k_b : k_a;
    r.vout = (remaining_after & 1) ? v_b : v_a;
    if (done_before == 0) { r.kin = k_src; r.vin = v_src; }
    else { r.kin = (remaining_after & 1) ? k_a : k_b; r.vin = (remaining_after & 1) ? v_a : v_b; }
    return r;
}

__global__ void __launch_bounds__(RS_THREADS)
radix_onesweep_kernel(const uint64_t* k_src, const uint32_t* v_src, uint64_t* k_a, uint32_t* v_a, uint64_t* k_b,
                      uint32_t* v_b, int n, const uint32_t* __restrict__ n_dev, int pass,
                      const uint32_t* __restrict__ hist, const uint32_t* __restrict__ skip, uint32_t* ticket,
                      uint32_t* status) {
    __shared__ uint32_t s_warp_hist[RS_WARPS][RS_BINS];  // per-warp digit counts -> per-warp exclusive offsets
    __shared__ uint32_t s_loc_off[RS_BINS];              // start of the digit's run inside the reordered tile
    __shared__ uint32_t s_dst_base[RS_BINS];             // global index of the run's first element minus s_loc_off
    __shared__ uint64_t s_keys[RS_TILE];
    __shared__ uint32_t s_vals[RS_TILE];
    __shared__ uint32_t s_scan[RS_WARPS];
    __shared__ uint32_t s_tile;

    if (skip[pass]) return;  // uniform: this pass would be the identity permutation
    if (n_dev) n = min(n, (int)*n_dev);
    const PingPong pp = resolve_buffers(skip, pass, k_src, v_src, k_a, v_a, k_b, v_b);

    if (threadIdx.x == 0) s_tile = atomicAdd(ticket + pass, 1u);
    for (int i = threadIdx.x; i < RS_WARPS * RS_BINS; i += RS_THREADS) (&s_warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int shift = 8 * pass;
    const int tile_base = tile * RS_TILE;
    if (tile_base >= n) return;  // launched for the capacity; nobody looks back at tiles past the end
    const int valid = min(RS_TILE, n - tile_base);

    // warp-striped load: item i of lane l of warp w is element w*32*KPT + i*32 + l of the tile
    uint64_t key[RS_KPT];
    uint32_t rank[RS_KPT];
    const int wbase = warp * 32 * RS_KPT + lane;
#pragma unroll
    for (int i = 0; i < RS_KPT; i++) {
        const int e = wbase + i * 32;
        key[i] = (e < valid) ? pp.kin[tile_base + e] : ~0ull;
    }
    // rank keys inside the warp, digit by digit (stable: earlier item / lower lane first)
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < RS_KPT; i++) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & 0xffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader) {
            old = s_warp_hist[warp][d];
            s_warp_hist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();

    // thread t owns digit t: exclusive scan over warps, publish, look back
    {
        const int d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            const uint32_t c = s_warp_hist[w][d];
            s_warp_hist[w][d] = run;
            run += c;
        }
        // padded keys (last tile only) all carry digit 0xff and rank last; do not publish them
        uint32_t count = run;
        if (d == 0xff) count -= (uint32_t)(RS_TILE - valid);
        uint32_t* my_status = status + (size_t)tile * RS_BINS + d;
        uint32_t excl = 0;
        if (tile == 0) {
            st_volatile_u32(my_status, ST_PREFIX | count);
        } else {
            st_volatile_u32(my_status, ST_AGG | count);
            int p = (int)tile - 1;
            while (true) {
                uint32_t s;
                do { s = ld_volatile_u32(status + (size_t)p * RS_BINS + d); } while ((s >> 30) == 0);
                excl += s & ST_MASK;
                if ((s >> 30) == 2) break;
                p--;
            }
            st_volatile_u32(my_status, ST_PREFIX | (excl + count));
        }
        // exclusive scan of the tile's digit counts (including padding: it sits at the very end)
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_scan[warp] = incl;
        __syncthreads();
        uint32_t woff = 0;
        for (int w = 0; w < warp; w++) woff += s_scan[w];
        const uint32_t loc = woff + incl - run;
        s_loc_off[d] = loc;
        s_dst_base[d] = hist[pass * RS_BINS + d] + excl - loc;
    }
    __syncthreads();

    // reorder through shared memory so each digit's run is contiguous
#pragma unroll
    for (int i = 0; i < RS_KPT; i++) {
        const int e = wbase + i * 32;
        const uint32_t d = (uint32_t)(key[i] >> shift) & 0xffu;
        const uint32_t pos = s_loc_off[d] + s_warp_hist[warp][d] + rank[i];
        s_keys[pos] = key[i];
        s_vals[pos] = (e < valid) ? pp.vin[tile_base + e] : 0u;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < valid; j += RS_THREADS) {
        const uint64_t k = s_keys[j];
        const uint32_t d = (uint32_t)(k >> shift) & 0xffu;
        const uint32_t dst = s_dst_base[d] + j;
        pp.kout[dst] = k;
        pp.vout[dst] = s_vals[j];
    }
}

// All passes skipped (n <= 1 or all keys equal): the output is the input.
__global__ void radix_copy_if_all_skipped(const uint64_t* k_src, const uint32_t* v_src, uint64_t* k_a, uint32_t* v_a,
                                          int n, const uint32_t* __restrict__ n_dev,
                                          const uint32_t* __restrict__ skip) {
    for (int q = 0; q < RS_MAX_PASSES; q++)
        if (!skip[q]) return;
    if (n_dev) n = min(n, (int)*n_dev);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        k_a[i] = k_src[i];
        v_a[i] = v_src[i];
    }
}

__global__ void radix_init_kernel(uint32_t* words, size_t nwords, int passes) {
    // zero everything, then preset skip[] = 1 for passes that do not exist
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += stride) {
        uint32_t v = 0;
        const size_t skip0 = RS_MAX_PASSES * RS_BINS + RS_MAX_PASSES;
        if (i >= skip0 && i < skip0 + RS_MAX_PASSES && (int)(i - skip0) >= passes) v = 1;
        words[i] = v;
    }
}

}  // namespace

size_t scan_temp_bytes(int n) {
    const size_t tiles = ((size_t)(n > 0 ? n : 1) + SCAN_TILE - 1) / SCAN_TILE;
    return 16 + tiles * sizeof(uint64_t);
}

void launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, uint32_t* total_out, uint32_t* temp,
                           cudaStream_t st) {
    if (n <= 0) {
        if (total_out) cudaMemsetAsync(total_out, 0, sizeof(uint32_t), st);
        return;
    }
    const int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    cudaMemsetAsync(temp, 0, scan_temp_bytes(n), st);
    inclusive_scan_kernel<<<tiles, SCAN_THREADS, 0, st>>>(in, out, n, total_out, temp,
                                                         reinterpret_cast<uint64_t*>(temp + 4));
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// temp: [header words][status words][pad][keys B: n u64][vals B: n u32]
size_t sort_temp_bytes(int n) {
    const size_t nn = n > 0 ? (size_t)n : 1;
    const size_t tiles = (nn + RS_TILE - 1) / RS_TILE;
    size_t b = (sort_header_words() + (size_t)RS_MAX_PASSES * tiles * RS_BINS) * sizeof(uint32_t);
    b = align_up(b, 256) + align_up(nn * sizeof(uint64_t), 256) + align_up(nn * sizeof(uint32_t), 256);
    return b;
}

void launch_sort_pairs(const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out,
                       uint32_t* values_out, int n, int end_bit, char* temp, cudaStream_t st,
                       const uint32_t* n_dev) {
    if (n <= 0) return;
    int passes = (end_bit + 7) / 8;
    if (passes < 1) passes = 1;
    if (passes > RS_MAX_PASSES) passes = RS_MAX_PASSES;
    const size_t tiles = ((size_t)n + RS_TILE - 1) / RS_TILE;
    uint32_t* words = reinterpret_cast<uint32_t*>(temp);
    SortTemp t;
    t.hist = words;
    t.ticket = words + RS_MAX_PASSES * RS_BINS;
    t.skip = t.ticket + RS_MAX_PASSES;
    t.status = words + sort_header_words();
    const size_t nwords = sort_header_words() + (size_t)passes * tiles * RS_BINS;
    char* p = temp + align_up((sort_header_words() + (size_t)RS_MAX_PASSES * tiles * RS_BINS) * sizeof(uint32_t), 256);
    uint64_t* k_b = reinterpret_cast<uint64_t*>(p);
    uint32_t* v_b = reinterpret_cast<uint32_t*>(p + align_up((size_t)n * sizeof(uint64_t), 256));

    size_t iblocks = (nwords + 1023) / 1024;
    if (iblocks > (size_t)NUM_SMS * 8) iblocks = (size_t)NUM_SMS * 8;
    radix_init_kernel<<<(int)iblocks, 1024, 0, st>>>(words, nwords, passes);
    int hblocks = (n + RS_THREADS * 16 - 1) / (RS_THREADS * 16);
    if (hblocks > NUM_SMS * 8) hblocks = NUM_SMS * 8;
    radix_hist_kernel<<<hblocks, RS_THREADS, 0, st>>>(keys_in, n, n_dev, passes, t.hist);
    radix_scan_kernel<<<passes, RS_BINS, 0, st>>>(t.hist, t.skip, n, n_dev);
    for (int pass = 0; pass < passes; pass++)
        radix_onesweep_kernel<<<(int)tiles, RS_THREADS, 0, st>>>(keys_in, values_in, keys_out, values_out, k_b, v_b, n,
                                                                 n_dev, pass, t.hist, t.skip, t.ticket,
                                                                 t.status + (size_t)pass * tiles * RS_BINS);
    int cblocks = (n + 255) / 256;
    if (cblocks > NUM_SMS * 4) cblocks = NUM_SMS * 4;
    radix_copy_if_all_skipped<<<cblocks, 256, 0, st>>>(keys_in, values_in, keys_out, values_out, n, n_dev, t.skip);
}


// ------------------------------------------------------------------------------------------------
// Prefix sum without any caller- or library-side allocation (as_inclusive_scan_u32_static).
//
// Some users of this library cannot hand it scratch memory: a mapping process on a robot that
// carved up all of its device memory into fixed arenas at start-up has nothing left to allocate
// from, and a control loop with a latency budget cannot afford cudaMalloc/cudaFree, which may
// synchronise the whole device.  For them the look-back state of the scan lives in the array below,
// which is part of the library image: every GPU that loads the library carries these 16 KiB until
// the process exits.
//
// In host code `g_static_scan_scratch` is only a handle for the symbol, not an address the GPU can
// use; the device address is looked up once per device with cudaGetSymbolAddress and cached.
// ------------------------------------------------------------------------------------------------
constexpr int STATIC_SCAN_MAX_TILES = 2048;
constexpr int STATIC_SCAN_MAX_N = STATIC_SCAN_MAX_TILES * SCAN_TILE;  // 4,194,304 elements
// [0..1]: ticket word + padding (16 bytes), [2..]: one status word per tile -- the layout of scan_temp
static __device__ __align__(256) unsigned long long g_static_scan_scratch[2 + STATIC_SCAN_MAX_TILES];

constexpr int MAX_DEVICES = 64;
static std::mutex g_static_scan_mu;
static void* g_static_scan_ptr[MAX_DEVICES] = {nullptr};

int static_scan_max_n() { return STATIC_SCAN_MAX_N; }

// Returns cudaSuccess and the device address of the built-in scratch for the CURRENT device.
// The first call on a device performs the symbol lookup (a host-side runtime query, no allocation).
cudaError_t static_scan_scratch(void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_scan_mu);
    if (!g_static_scan_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scan_scratch);
        if (e != cudaSuccess) return e;
        g_static_scan_ptr[dev] = p;
    }
    *out = g_static_scan_ptr[dev];
    return cudaSuccess;
}

}  // namespace asb
