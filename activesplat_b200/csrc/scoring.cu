// scoring.cu -- the clustering step that follows the GetOpacity renders (sm_100a).
//
//   dbscan_grid_kernel  replaces  sklearn.cluster.DBSCAN(eps, min_samples).fit_predict(np.column_stack(np.where(img > thr)))
//                       as called by get_invisibility_clusters   src/mapper/__init__.py:92-117
//                       (and by get_convexhull_volume            src/mapper/__init__.py:8-21)
//
// The samples are pixels of a regular grid, so a radius query is a fixed disc of offsets and DBSCAN reduces to
// image operations.  sklearn's labelling (cluster/_dbscan_inner.pyx) is reproduced exactly, not just the partition:
//   * a sample is a core point iff at least min_samples samples (itself included) lie within eps (<=);
//   * clusters are the connected components of core points under "within eps"; they are numbered in the order their
//     first core point appears in the sample list -- np.where order = row-major -- so a component's label is the
//     rank of its smallest pixel index;
//   * a non-core sample within eps of core points takes the label of the FIRST cluster expanded that reaches it,
//     i.e. the smallest label among those core points; otherwise it is noise (-1).
// One CTA per image (images of a batch in parallel): min-label propagation with pointer jumping over an int32
// parent image until nothing changes, a block-wide scan to rank the roots, then per-cluster statistics
// (count, sum of rows, sum of columns as exact integers; sum of values in double).
#include "common.cuh"

namespace asb {

namespace {

constexpr int DB_THREADS = 1024;
constexpr int DB_MAX_OFFSETS = 256;  // discs up to eps = 8

__global__ void __launch_bounds__(DB_THREADS)
dbscan_grid_kernel(int H, int W, const float* __restrict__ values, float thr, float eps, int min_samples,
                   int* __restrict__ labels, int* __restrict__ parent, int max_clusters, int* __restrict__ n_clusters,
                   int* __restrict__ stats_i, double* __restrict__ stats_v) {
    __shared__ short2 s_off[DB_MAX_OFFSETS];
    __shared__ int s_noff;
    __shared__ int s_scan[DB_THREADS / 32];
    __shared__ int s_carry;
    const int N = H * W;
    const int b = blockIdx.x;
    values += (size_t)b * N;
    labels += (size_t)b * N;
    parent += (size_t)b * N;
    stats_i += (size_t)b * max_clusters * 3;
    stats_v += (size_t)b * max_clusters;
    if (threadIdx.x == 0) {
        const int r = (int)floorf(eps);
        int n = 0;
        for (int dy = -r; dy <= r; dy++)
            for (int dx = -r; dx <= r; dx++)
                if ((float)(dx * dx + dy * dy) <= eps * eps && n < DB_MAX_OFFSETS) s_off[n++] = make_short2((short)dx, (short)dy);
        s_noff = n;
        s_carry = 0;
    }
    for (int c = threadIdx.x; c < max_clusters; c += DB_THREADS) {
        stats_i[3 * c] = 0; stats_i[3 * c + 1] = 0; stats_i[3 * c + 2] = 0;
        stats_v[c] = 0.0;
    }
    // samples: labels = -1 (a sample, not yet labelled) / -2 (not a sample)
    for (int p = threadIdx.x; p < N; p += DB_THREADS) labels[p] = (values[p] > thr) ? -1 : -2;
    __syncthreads();
    const int noff = s_noff;
    // core points; parent = own index for core points, INT_MAX otherwise
    for (int p = threadIdx.x; p < N; p += DB_THREADS) {
        int par = INT_MAX;
        if (labels[p] == -1) {
            const int y = p / W, x = p - y * W;
            int cnt = 0;
            for (int k = 0; k < noff; k++) {
                const int xx = x + s_off[k].x, yy = y + s_off[k].y;
                if (xx >= 0 && xx < W && yy >= 0 && yy < H && labels[yy * W + xx] != -2) cnt++;
            }
            if (cnt >= min_samples) par = p;
        }
        parent[p] = par;
    }
    __syncthreads();
    // connected components of the core points: parent <- min over core neighbours, then pointer jumping
    for (;;) {
        int changed = 0;
        for (int p = threadIdx.x; p < N; p += DB_THREADS) {
            const int mine = parent[p];
            if (mine == INT_MAX) continue;
            const int y = p / W, x = p - y * W;
            int m = mine;
            for (int k = 0; k < noff; k++) {
                const int xx = x + s_off[k].x, yy = y + s_off[k].y;
                if (xx >= 0 && xx < W && yy >= 0 && yy < H) m = min(m, parent[yy * W + xx]);  // INT_MAX for non-core
            }
            if (m < mine) {
                // hook the old root as well, so whole trees merge at once
                atomicMin(parent + mine, m);
                atomicMin(parent + p, m);
                changed = 1;
            }
        }
        __syncthreads();
        for (int j = 0; j < 4; j++) {  // values are indices of core pixels with parent <= index: safe to chase
            for (int p = threadIdx.x; p < N; p += DB_THREADS) {
                const int q = parent[p];
                if (q != INT_MAX) {
                    const int g = parent[q];
                    if (g < q) parent[p] = g;
                }
            }
            __syncthreads();
        }
        if (!__syncthreads_or(changed)) break;
    }
    // a sweep without changes means neighbouring core pixels agree, so every core pixel now holds its component's
    // smallest pixel index (that pixel's parent can never drop below itself)
    // rank the roots (row-major order of their index = order in which sklearn discovers the clusters)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < N; base += DB_THREADS) {
        const int p = base + threadIdx.x;
        const int is_root = (p < N && parent[p] == p) ? 1 : 0;
        int incl = is_root;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_scan[warp] = incl;
        __syncthreads();
        int woff = 0;
        for (int w = 0; w < warp; w++) woff += s_scan[w];
        if (is_root) labels[p] = s_carry + woff + incl - 1;
        __syncthreads();
        if (threadIdx.x == DB_THREADS - 1) s_carry += woff + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) n_clusters[b] = s_carry;
    // core pixels take their root's label
    for (int p = threadIdx.x; p < N; p += DB_THREADS) {
        const int q = parent[p];
        if (q != INT_MAX && q != p) labels[p] = labels[q];
    }
    __syncthreads();
    // border samples: smallest label among the core points within eps; the rest stays noise
    for (int p = threadIdx.x; p < N; p += DB_THREADS) {
        if (labels[p] != -1 || parent[p] != INT_MAX) continue;
        const int y = p / W, x = p - y * W;
        int best = INT_MAX;
        for (int k = 0; k < noff; k++) {
            const int xx = x + s_off[k].x, yy = y + s_off[k].y;
            if (xx >= 0 && xx < W && yy >= 0 && yy < H) {
                const int q = yy * W + xx;
                if (parent[q] != INT_MAX) best = min(best, labels[q]);
            }
        }
        // only labels of CORE pixels are read in this pass and only labels of border pixels are written: no race
        labels[p] = (best == INT_MAX) ? -1 : best;
    }
    __syncthreads();
    for (int p = threadIdx.x; p < N; p += DB_THREADS) {
        const int l = labels[p];
        if (l >= 0 && l < max_clusters) {
            const int y = p / W, x = p - y * W;
            atomicAdd(stats_i + 3 * l, 1);
            atomicAdd(stats_i + 3 * l + 1, y);
            atomicAdd(stats_i + 3 * l + 2, x);
            atomicAdd(stats_v + l, (double)values[p]);
        }
    }
}

}  // namespace

void launch_dbscan_grid(int batch, int H, int W, const float* values, float thr, float eps, int min_samples, int* labels,
                        int* scratch, int max_clusters, int* n_clusters, int* stats_i, double* stats_v, cudaStream_t st) {
    dbscan_grid_kernel<<<batch, DB_THREADS, 0, st>>>(H, W, values, thr, eps, min_samples, labels, scratch, max_clusters,
                                                     n_clusters, stats_i, stats_v);
}

}  // namespace asb
