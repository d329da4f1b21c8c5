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
#include <atomic>

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


// ---------------------------------------------------------------------------------------------------------------
// cluster_hull_kernel  replaces, per DBSCAN cluster of a GLOBAL GetOpacity panorama (src/mapper/__init__.py:29-86):
//     cv2.dilate(mask, getStructuringElement(MORPH_ELLIPSE, (15, 15)))                      :37-38
//     cv2.findContours(dilated, RETR_EXTERNAL, CHAIN_APPROX_SIMPLE), the largest contour      :41-42
//     back-projection of the contour pixels with the median depth, skipping z == 15           :49-66
//     scipy.spatial.ConvexHull(points).volume                                                 :69-78
// One CTA per (cluster, panorama).  The cluster mask and its dilation live in shared memory as bit images inside a
// one-pixel frame of zeros; the dilation tests, per element row, a run of at most 15 bits of the mask; one thread
// follows the outer border (Suzuki-Abe, as OpenCV does: first pixel in raster order, 8-connected foreground,
// counter-clockwise neighbour search) and keeps the points where the chain code changes (CHAIN_APPROX_SIMPLE); the
// convex hull is built incrementally over triangles in double precision -- visibility tests and horizon search spread
// over the CTA -- and its volume is the sum of the signed tetrahedra against an interior point.  A DBSCAN cluster
// dilated by a disc of radius 7 > eps is a single connected component, so "the largest contour" is the only one.
// ---------------------------------------------------------------------------------------------------------------
constexpr int HULL_THREADS = 256;
constexpr int HULL_MAX_POINTS = 1024;            // contour vertices kept per cluster
constexpr int HULL_MAX_FACES = 2 * HULL_MAX_POINTS;
constexpr float HULL_EMPTY_DEPTH = 15.0f;        // median depth of a pixel whose transmittance never crossed 0.5

struct HullSmem {
    int first;            // smallest frame bit index of the dilated mask
    int n_contour, n_pts, n_faces, n_vis, n_horizon, overflow;
    int bb[4];            // x0, y0, x1, y1 of the cluster (image coordinates)
    double red_val[HULL_THREADS / 32];
    int red_idx[HULL_THREADS / 32];
};

// index of the largest v over the block (ties: any); lanes with idx < 0 do not take part
__device__ __forceinline__ int hull_block_argmax(double v, int idx, HullSmem& sm) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (idx < 0) v = -1.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > v) { v = ov; idx = oi; }
    }
    __syncthreads();
    if (lane == 0) { sm.red_val[warp] = v; sm.red_idx[warp] = idx; }
    __syncthreads();
    double bv = sm.red_val[0];
    int bi = sm.red_idx[0];
    for (int w = 1; w < HULL_THREADS / 32; w++)
        if (sm.red_val[w] > bv) { bv = sm.red_val[w]; bi = sm.red_idx[w]; }
    return bi;
}

__global__ void __launch_bounds__(HULL_THREADS)
cluster_hull_kernel(int H, int W, const int* __restrict__ labels, const int* __restrict__ n_clusters,
                    const float* __restrict__ depth, double h_angle, double v_angle, int max_clusters,
                    double* __restrict__ volumes, int* __restrict__ n_points) {
    extern __shared__ __align__(16) unsigned char hull_raw[];
    const int b = blockIdx.y, c = blockIdx.x;
    const int N = H * W;
    const size_t out = (size_t)b * max_clusters + c;
    if (c >= min(n_clusters[b], max_clusters)) {
        if (threadIdx.x == 0) { volumes[out] = 0.0; n_points[out] = 0; }
        return;
    }
    labels += (size_t)b * N;
    depth += (size_t)b * N;
    const int RW = (W + 2 + 31) >> 5;            // words per frame row
    const int FH = H + 2;
    // carve the dynamic shared memory
    HullSmem& sm = *reinterpret_cast<HullSmem*>(hull_raw);
    double* pts = reinterpret_cast<double*>(hull_raw + ((sizeof(HullSmem) + 15) & ~(size_t)15));   // [HULL_MAX_POINTS][3]
    ushort4* faces = reinterpret_cast<ushort4*>(pts + 3 * HULL_MAX_POINTS);                // a, b, c, visible
    uint32_t* horizon = reinterpret_cast<uint32_t*>(faces + HULL_MAX_FACES);               // a | b << 16
    uint32_t* vis_list = horizon + 3 * HULL_MAX_FACES;
    short2* contour = reinterpret_cast<short2*>(vis_list + HULL_MAX_FACES);                // image coordinates
    uint32_t* mask = reinterpret_cast<uint32_t*>(contour + HULL_MAX_POINTS);
    uint32_t* dil = mask + (size_t)FH * RW;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NW = HULL_THREADS / 32;
    if (threadIdx.x == 0) {
        sm.first = INT_MAX; sm.n_contour = 0; sm.n_pts = 0; sm.n_faces = 0; sm.overflow = 0;
        sm.bb[0] = INT_MAX; sm.bb[1] = INT_MAX; sm.bb[2] = -1; sm.bb[3] = -1;
    }
    __syncthreads();
    // ---- the cluster's mask as a bit image (frame coordinates = image coordinates + 1) ----
    for (int w = warp; w < FH * RW; w += NW) {
        const int fy = w / RW, fx = (w - fy * RW) * 32 + lane;
        const int y = fy - 1, x = fx - 1;
        const bool in = (y >= 0 && y < H && x >= 0 && x < W) && labels[y * W + x] == c;
        const uint32_t bits = __ballot_sync(0xffffffffu, in);
        if (lane == 0) { mask[w] = bits; dil[w] = 0u; }
        if (in) { atomicMin(&sm.bb[0], x); atomicMin(&sm.bb[1], y); atomicMax(&sm.bb[2], x); atomicMax(&sm.bb[3], y); }
    }
    __syncthreads();
    // ---- dilation with OpenCV's 15x15 elliptic element ----
    const int R = 7;
    const int bx0 = max(sm.bb[0] - R, 0), by0 = max(sm.bb[1] - R, 0), bx1 = min(sm.bb[2] + R, W - 1), by1 = min(sm.bb[3] + R, H - 1);
    const int wlo = (bx0 + 1) >> 5, whi = (bx1 + 1) >> 5;
    const int wcount = whi - wlo + 1, rows = by1 - by0 + 1;
    for (int t = warp; t < rows * wcount; t += NW) {
        const int y = by0 + t / wcount, wi = wlo + t % wcount;
        const int fx = wi * 32 + lane, x = fx - 1;
        bool on = false;
        if (x >= 0 && x < W) {
#pragma unroll
            for (int dy = -R; dy <= R; dy++) {
                const int yy = y + dy;
                if (yy < 0 || yy >= H) continue;
                // half-width of the element row: round(7 * sqrt((49 - dy^2) / 49)) = 0 4 5 6 6 7 7 7 7 7 6 6 5 4 0
                const int ady = dy < 0 ? -dy : dy;
                const int dx = ady == 7 ? 0 : ady == 6 ? 4 : ady == 5 ? 5 : ady >= 3 ? 6 : 7;
                const int lo = max(fx - dx, 0), hi = min(fx + dx, W + 1);
                const uint32_t* row = mask + (size_t)(yy + 1) * RW;
                const int w0 = lo >> 5;
                uint64_t win = row[w0];
                if (w0 + 1 < RW) win |= (uint64_t)row[w0 + 1] << 32;
                win >>= (lo & 31);
                if (win & ((1ull << (hi - lo + 1)) - 1ull)) { on = true; break; }
            }
        }
        const uint32_t bits = __ballot_sync(0xffffffffu, on);
        if (lane == 0) {
            dil[(size_t)(y + 1) * RW + wi] = bits;
            if (bits) atomicMin(&sm.first, ((y + 1) * RW + wi) * 32 + __ffs(bits) - 1);
        }
    }
    __syncthreads();
    // ---- outer border of the dilated component, CHAIN_APPROX_SIMPLE; back-projection ----
    if (threadIdx.x == 0 && sm.first != INT_MAX) {
        auto at = [&](int fx, int fy) -> bool { return (dil[(size_t)fy * RW + (fx >> 5)] >> (fx & 31)) & 1u; };
        const int ddx[8] = {1, 1, 0, -1, -1, -1, 0, 1}, ddy[8] = {0, -1, -1, -1, 0, 1, 1, 1};
        const int w0 = sm.first >> 5;
        const int y0 = w0 / RW, x0 = (w0 - y0 * RW) * 32 + (sm.first & 31);
        int n = 0;
        int s = 4, x1 = x0, y1 = y0;
        const int s_end = 4;
        do {                                                    // first neighbour, clockwise from west
            s = (s - 1) & 7;
            x1 = x0 + ddx[s]; y1 = y0 + ddy[s];
        } while (!at(x1, y1) && s != s_end);
        if (s == s_end) {
            contour[n++] = make_short2((short)(x0 - 1), (short)(y0 - 1));       // an isolated pixel
        } else {
            int x3 = x0, y3 = y0, prev_s = s ^ 4;
            for (int guard = 0; guard < 8 * (H + 2) * (W + 2); guard++) {
                int x4, y4;
                for (;;) {                                      // next border pixel, counter-clockwise
                    s++;
                    x4 = x3 + ddx[s & 7]; y4 = y3 + ddy[s & 7];
                    if (at(x4, y4)) break;
                }
                s &= 7;
                if (s != prev_s) {                              // the chain turns here: the point is kept
                    if (n < HULL_MAX_POINTS) contour[n] = make_short2((short)(x3 - 1), (short)(y3 - 1));
                    n++;
                }
                prev_s = s;
                if (x4 == x0 && y4 == y0 && x3 == x1 && y3 == y1) break;
                x3 = x4; y3 = y4;
                s = (s + 4) & 7;
            }
        }
        if (n > HULL_MAX_POINTS) { sm.overflow = 1; n = HULL_MAX_POINTS; }
        sm.n_contour = n;
        // (x * h_angle, y * v_angle, median depth); pixels without a surface are skipped
        int m = 0;
        for (int i = 0; i < n; i++) {
            const int x = contour[i].x, y = contour[i].y;
            if (x < 0 || x >= W || y < 0 || y >= H) continue;
            const float z = depth[y * W + x];
            if (z == HULL_EMPTY_DEPTH) continue;
            pts[3 * m] = x * h_angle; pts[3 * m + 1] = y * v_angle; pts[3 * m + 2] = (double)z;
            m++;
        }
        sm.n_pts = m;
    }
    __syncthreads();
    const int n = sm.n_pts;
    if (threadIdx.x == 0) n_points[out] = sm.overflow ? -n : n;
    if (n < 4) { if (threadIdx.x == 0) volumes[out] = 0.0; return; }
    // ---- incremental convex hull ----
    auto P = [&](int i, int k) -> double { return pts[3 * i + k]; };
    double cx = 0, cy = 0, cz = 0;
    for (int i = 0; i < n; i++) { cx += P(i, 0); cy += P(i, 1); cz += P(i, 2); }   // every thread: n <= 1024 adds
    cx /= n; cy /= n; cz /= n;
    double sp = 0;
    for (int i = threadIdx.x; i < n; i += HULL_THREADS)
        sp = fmax(sp, fmax(fabs(P(i, 0) - cx), fmax(fabs(P(i, 1) - cy), fabs(P(i, 2) - cz))));
    const int isp = hull_block_argmax(sp, threadIdx.x, sm);
    __syncthreads();
    if (threadIdx.x == isp) sm.red_val[0] = sp;
    __syncthreads();
    const double span = sm.red_val[0];            // scale of the tolerances
    __syncthreads();
    if (span == 0.0) { if (threadIdx.x == 0) volumes[out] = 0.0; return; }
    // initial tetrahedron: farthest from the centroid, farthest from it, farthest from that line, farthest from that plane
    int i0, i1, i2, i3;
    {
        double best = -1; int bi = -1;
        for (int i = threadIdx.x; i < n; i += HULL_THREADS) {
            const double dx = P(i, 0) - cx, dy = P(i, 1) - cy, dz = P(i, 2) - cz;
            const double v = dx * dx + dy * dy + dz * dz;
            if (v > best) { best = v; bi = i; }
        }
        i0 = hull_block_argmax(best, bi, sm);
        best = -1; bi = -1;
        for (int i = threadIdx.x; i < n; i += HULL_THREADS) {
            const double dx = P(i, 0) - P(i0, 0), dy = P(i, 1) - P(i0, 1), dz = P(i, 2) - P(i0, 2);
            const double v = dx * dx + dy * dy + dz * dz;
            if (v > best) { best = v; bi = i; }
        }
        i1 = hull_block_argmax(best, bi, sm);
        const double ux = P(i1, 0) - P(i0, 0), uy = P(i1, 1) - P(i0, 1), uz = P(i1, 2) - P(i0, 2);
        best = -1; bi = -1;
        for (int i = threadIdx.x; i < n; i += HULL_THREADS) {
            const double dx = P(i, 0) - P(i0, 0), dy = P(i, 1) - P(i0, 1), dz = P(i, 2) - P(i0, 2);
            const double kx = dy * uz - dz * uy, ky = dz * ux - dx * uz, kz = dx * uy - dy * ux;
            const double v = kx * kx + ky * ky + kz * kz;
            if (v > best) { best = v; bi = i; }
        }
        i2 = hull_block_argmax(best, bi, sm);
        const double vx = P(i2, 0) - P(i0, 0), vy = P(i2, 1) - P(i0, 1), vz = P(i2, 2) - P(i0, 2);
        const double nx = uy * vz - uz * vy, ny = uz * vx - ux * vz, nz = ux * vy - uy * vx;
        best = -1; bi = -1;
        for (int i = threadIdx.x; i < n; i += HULL_THREADS) {
            const double v = fabs((P(i, 0) - P(i0, 0)) * nx + (P(i, 1) - P(i0, 1)) * ny + (P(i, 2) - P(i0, 2)) * nz);
            if (v > best) { best = v; bi = i; }
        }
        i3 = hull_block_argmax(best, bi, sm);
        const double d3 = (P(i3, 0) - P(i0, 0)) * nx + (P(i3, 1) - P(i0, 1)) * ny + (P(i3, 2) - P(i0, 2)) * nz;
        if (fabs(d3) <= 1e-12 * span * span * span) { if (threadIdx.x == 0) volumes[out] = 0.0; return; }   // coplanar
    }
    const double ix = 0.25 * (P(i0, 0) + P(i1, 0) + P(i2, 0) + P(i3, 0)), iy = 0.25 * (P(i0, 1) + P(i1, 1) + P(i2, 1) + P(i3, 1)),
                 iz = 0.25 * (P(i0, 2) + P(i1, 2) + P(i2, 2) + P(i3, 2));
    // signed plane test of point q against face f (outward normal): returns n.(q - a), and |n| in nn
    auto side = [&](const ushort4 f, double qx, double qy, double qz, double& nn) -> double {
        const double ax = P(f.x, 0), ay = P(f.x, 1), az = P(f.x, 2);
        const double e1x = P(f.y, 0) - ax, e1y = P(f.y, 1) - ay, e1z = P(f.y, 2) - az;
        const double e2x = P(f.z, 0) - ax, e2y = P(f.z, 1) - ay, e2z = P(f.z, 2) - az;
        const double nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;
        nn = sqrt(nx * nx + ny * ny + nz * nz);
        return nx * (qx - ax) + ny * (qy - ay) + nz * (qz - az);
    };
    if (threadIdx.x == 0) {
        const int tet[4][3] = {{i0, i1, i2}, {i0, i3, i1}, {i1, i3, i2}, {i2, i3, i0}};
        for (int f = 0; f < 4; f++) {
            ushort4 t = make_ushort4((unsigned short)tet[f][0], (unsigned short)tet[f][1], (unsigned short)tet[f][2], 0);
            double nn;
            if (side(t, ix, iy, iz, nn) > 0) { const unsigned short u = t.y; t.y = t.z; t.z = u; }   // outward w.r.t. the interior
            faces[f] = t;
        }
        sm.n_faces = 4;
    }
    __syncthreads();
    for (int p = 0; p < n; p++) {
        if (p == i0 || p == i1 || p == i2 || p == i3) continue;
        const int F = sm.n_faces;
        const double qx = P(p, 0), qy = P(p, 1), qz = P(p, 2);
        if (threadIdx.x == 0) { sm.n_vis = 0; sm.n_horizon = 0; }
        __syncthreads();
        for (int f = threadIdx.x; f < F; f += HULL_THREADS) {
            double nn;
            const double d = side(faces[f], qx, qy, qz, nn);
            const bool v = d > 1e-12 * span * nn;
            faces[f].w = v ? 1 : 0;
            if (v) vis_list[atomicAdd(&sm.n_vis, 1)] = f;
        }
        __syncthreads();
        const int V = sm.n_vis;
        if (V == 0) continue;                                   // inside the current hull (uniform over the block)
        // horizon: directed edges of visible faces whose reverse is not an edge of a visible face
        for (int e = threadIdx.x; e < 3 * V; e += HULL_THREADS) {
            const ushort4 f = faces[vis_list[e / 3]];
            const int k = e % 3;
            const unsigned short ea = k == 0 ? f.x : k == 1 ? f.y : f.z, eb = k == 0 ? f.y : k == 1 ? f.z : f.x;
            bool twin = false;
            for (int j = 0; j < V && !twin; j++) {
                const ushort4 g = faces[vis_list[j]];
                twin = (g.x == eb && g.y == ea) || (g.y == eb && g.z == ea) || (g.z == eb && g.x == ea);
            }
            if (!twin) horizon[atomicAdd(&sm.n_horizon, 1)] = (uint32_t)ea | ((uint32_t)eb << 16);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int m = 0;
            for (int f = 0; f < F; f++)
                if (!faces[f].w) faces[m++] = faces[f];
            const int nh = sm.n_horizon;
            for (int h = 0; h < nh && m < HULL_MAX_FACES; h++)
                faces[m++] = make_ushort4((unsigned short)(horizon[h] & 0xffffu), (unsigned short)(horizon[h] >> 16),
                                          (unsigned short)p, 0);
            sm.n_faces = m;
        }
        __syncthreads();
    }
    // ---- volume: signed tetrahedra of the faces against the interior point ----
    double vol = 0.0;
    for (int f = threadIdx.x; f < sm.n_faces; f += HULL_THREADS) {
        const ushort4 t = faces[f];
        const double ax = P(t.x, 0) - ix, ay = P(t.x, 1) - iy, az = P(t.x, 2) - iz;
        const double bx = P(t.y, 0) - ix, by = P(t.y, 1) - iy, bz = P(t.y, 2) - iz;
        const double cx2 = P(t.z, 0) - ix, cy2 = P(t.z, 1) - iy, cz2 = P(t.z, 2) - iz;
        vol += ax * (by * cz2 - bz * cy2) + ay * (bz * cx2 - bx * cz2) + az * (bx * cy2 - by * cx2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vol += __shfl_xor_sync(0xffffffffu, vol, o);
    __syncthreads();
    if (lane == 0) sm.red_val[warp] = vol;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int w = 0; w < NW; w++) t += sm.red_val[w];
        volumes[out] = fabs(t) / 6.0;
    }
}

size_t cluster_hull_smem_bytes(int H, int W) {
    const size_t RW = (size_t)(W + 2 + 31) >> 5;
    return ((sizeof(HullSmem) + 15) & ~(size_t)15) + sizeof(double) * 3 * HULL_MAX_POINTS + sizeof(ushort4) * HULL_MAX_FACES +
           4 * sizeof(uint32_t) * HULL_MAX_FACES + sizeof(short2) * HULL_MAX_POINTS + 2 * sizeof(uint32_t) * (size_t)(H + 2) * RW;
}

}  // namespace

int launch_cluster_hull(int batch, int H, int W, const int* labels, const int* n_clusters, const float* depth,
                        double vfov_deg, int max_clusters, double* volumes, int* n_points, cudaStream_t st) {
    const size_t smem = cluster_hull_smem_bytes(H, W);
    if (smem > 200 * 1024) return -1;
    static std::atomic<unsigned> done_mask{0};
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned bit = 1u << (dev & 31);
    if (!(done_mask.load(std::memory_order_relaxed) & bit)) {
        cudaFuncSetAttribute(cluster_hull_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        done_mask.fetch_or(bit, std::memory_order_relaxed);
    }
    const double pi = 3.14159265358979323846;
    const double h_angle = (360.0 / W) * (pi / 180.0), v_angle = (vfov_deg / H) * (pi / 180.0);   // src/mapper/__init__.py:63-64
    cluster_hull_kernel<<<dim3(max_clusters, batch), HULL_THREADS, smem, st>>>(H, W, labels, n_clusters, depth, h_angle,
                                                                              v_angle, max_clusters, volumes, n_points);
    return 0;
}

void launch_dbscan_grid(int batch, int H, int W, const float* values, float thr, float eps, int min_samples, int* labels,
                        int* scratch, int max_clusters, int* n_clusters, int* stats_i, double* stats_v, cudaStream_t st) {
    dbscan_grid_kernel<<<batch, DB_THREADS, 0, st>>>(H, W, values, thr, eps, min_samples, labels, scratch, max_clusters,
                                                     n_clusters, stats_i, stats_v);
}

}  // namespace asb
