"""CPU restatement of get_convexhull_volume (src/mapper/__init__.py:8-90), the post-processing of a GLOBAL GetOpacity
panorama: DBSCAN clusters of low-visibility pixels, each dilated, its outline back-projected with the median depth, and
the volume of the 3-D convex hull of that outline.

TEST INFRASTRUCTURE ONLY.  The reference calls three third-party routines; their algorithms are restated here so that
the CUDA kernels (activesplat_b200/csrc/scoring.cu) have something to be compared with line by line:

  cv2.getStructuringElement(MORPH_ELLIPSE, (15, 15)) + cv2.dilate   -> ellipse_rows(), dilate()
      OpenCV 4.x: row i of the element spans [c - dx, c + dx] with dx = round(c * sqrt((r^2 - dy^2) / r^2)), r = c = 7;
      dilation ignores pixels outside the image.
  cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_SIMPLE)               -> outer_contour_simple()
      Suzuki-Abe border following of the outer border of a component (8-connected foreground), started at its first
      pixel in raster order; CHAIN_APPROX_SIMPLE keeps a border point only where the chain changes direction.
  scipy.spatial.ConvexHull(points).volume (qhull)                     -> convex_hull_volume()
      the volume of a convex hull is a geometric quantity: an incremental hull in double precision gives it to rounding.

Pinned by tests/golden/ref_hull.npz: outputs of the reference's own function (with OpenCV, scikit-learn and SciPy as
installed in this image) on the seeded panoramas of `panorama()` (tests/golden/make_hull_golden.py).
"""
import numpy as np

from .scoring_oracle import dbscan_labels_image

KSIZE = 15
EMPTY_DEPTH = 15.0  # the rasterizer's median depth where T never crossed 0.5 (dgr forward.cu:309); skipped at :54


def ellipse_rows(ksize=KSIZE):
    """Half-widths dx[i] of the rows of OpenCV's elliptic structuring element (negative: empty row)."""
    r = c = ksize // 2
    inv_r2 = 1.0 / (r * r) if r else 0.0
    out = []
    for i in range(ksize):
        dy = i - r
        out.append(int(np.rint(c * np.sqrt((r * r - dy * dy) * inv_r2))) if abs(dy) <= r else -1)
    return out


def dilate(mask, ksize=KSIZE):
    """cv2.dilate(mask, ellipse) for a boolean mask; pixels outside the image do not contribute."""
    H, W = mask.shape
    r = ksize // 2
    out = np.zeros_like(mask, dtype=bool)
    ys, xs = np.nonzero(mask)
    for i, dx in enumerate(ellipse_rows(ksize)):
        if dx < 0:
            continue
        y = ys + (i - r)          # the element is symmetric: a set pixel reaches the rows ys + dy
        ok = (y >= 0) & (y < H)
        for yy, xx in zip(y[ok], xs[ok]):
            out[yy, max(xx - dx, 0):min(xx + dx + 1, W)] = True
    return out


# chain codes as OpenCV numbers them: 0 = east, then counter-clockwise on the screen (y grows downwards)
_DELTAS = [(1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1), (1, 1)]


def outer_contour_simple(mask):
    """Points (x, y) that cv2.findContours(mask, RETR_EXTERNAL, CHAIN_APPROX_SIMPLE) returns for the component that
    contains the first set pixel in raster order (the caller guarantees a single component)."""
    H, W = mask.shape
    m = np.zeros((H + 2, W + 2), dtype=bool)   # the one-pixel frame of zeros findContours works in
    m[1:-1, 1:-1] = mask
    ys, xs = np.nonzero(m)
    if len(ys) == 0:
        return []
    y0, x0 = int(ys[0]), int(xs[0])            # np.nonzero is row-major: top-most row, left-most pixel in it

    def at(x, y):
        return m[y, x]

    # first neighbour, clockwise from west
    s_end = s = 4
    while True:
        s = (s - 1) & 7
        x1, y1 = x0 + _DELTAS[s][0], y0 + _DELTAS[s][1]
        if at(x1, y1) or s == s_end:
            break
    if s == s_end:                              # an isolated pixel
        return [(x0 - 1, y0 - 1)]
    pts = []
    x3, y3 = x0, y0
    prev_s = s ^ 4
    while True:
        while True:                             # next border pixel, counter-clockwise from the one we came from
            s += 1
            x4, y4 = x3 + _DELTAS[s & 7][0], y3 + _DELTAS[s & 7][1]
            if at(x4, y4):
                break
        s &= 7
        if s != prev_s:                         # CHAIN_APPROX_SIMPLE: keep the point only where the direction changes
            pts.append((x3 - 1, y3 - 1))
        prev_s = s
        if (x4, y4) == (x0, y0) and (x3, y3) == (x1, y1):
            break
        x3, y3 = x4, y4
        s = (s + 4) & 7
    return pts


def convex_hull_volume(points, eps_rel=1e-12):
    """Volume of the convex hull of `points` [n,3] (float64): incremental hull over triangles with outward normals.
    Returns 0.0 when the points are (numerically) coplanar."""
    P = np.asarray(points, dtype=np.float64)
    n = len(P)
    if n < 4:
        return 0.0
    span = float(np.abs(P - P.mean(0)).max())
    if span == 0.0:
        return 0.0
    # initial tetrahedron: extreme point, farthest from it, farthest from the line, farthest from the plane
    i0 = int(np.argmin(P[:, 0]))
    i1 = int(np.argmax(((P - P[i0]) ** 2).sum(1)))
    d = P[i1] - P[i0]
    cr = np.cross(P - P[i0], d)
    i2 = int(np.argmax((cr ** 2).sum(1)))
    nrm = np.cross(P[i1] - P[i0], P[i2] - P[i0])
    dist = (P - P[i0]) @ nrm
    i3 = int(np.argmax(np.abs(dist)))
    if abs(dist[i3]) <= eps_rel * span ** 3:
        return 0.0
    faces = [(i0, i1, i2), (i0, i3, i1), (i1, i3, i2), (i2, i3, i0)]
    interior = P[[i0, i1, i2, i3]].mean(0)

    def plane(f):
        a, b, c = P[f[0]], P[f[1]], P[f[2]]
        nn = np.cross(b - a, c - a)
        return nn, float(nn @ a)

    # orient every face outwards w.r.t. the interior point
    fixed = []
    for f in faces:
        nn, off = plane(f)
        fixed.append(f if nn @ interior - off < 0 else (f[0], f[2], f[1]))
    faces = fixed
    used = {i0, i1, i2, i3}
    for p in range(n):
        if p in used:
            continue
        vis = []
        for fi, f in enumerate(faces):
            nn, off = plane(f)
            if nn @ P[p] - off > eps_rel * span * np.sqrt(nn @ nn):
                vis.append(fi)
        if not vis:
            continue
        vset = set(vis)
        edges = {}
        for fi in vis:
            a, b, c = faces[fi]
            for e in ((a, b), (b, c), (c, a)):
                edges[e] = fi
        horizon = [e for e in edges if (e[1], e[0]) not in edges]
        faces = [f for fi, f in enumerate(faces) if fi not in vset] + [(a, b, p) for a, b in horizon]
    vol = 0.0
    for a, b, c in faces:
        vol += np.dot(P[a] - interior, np.cross(P[b] - interior, P[c] - interior))
    return abs(vol) / 6.0


def get_convexhull_volume(depth_np, invisibility_np, vfov=150):
    """src/mapper/__init__.py:8-90.  Returns (sum over clusters of sum(invisibility) * volume, sum of volumes)."""
    if depth_np.ndim == 3:
        depth_np = np.squeeze(depth_np, axis=-1)
    labels = dbscan_labels_image(invisibility_np, 0.8, 5, 25)                          # :12-18
    if (labels >= -1).sum() == 0:
        return 0, 0
    H, W = depth_np.shape
    h_angle = np.deg2rad(360 / W)                                                    # :63-64
    v_angle = np.deg2rad(vfov / H)
    inv_sums, vol_sums = [], []
    for c in range(int(labels.max()) + 1):
        member = labels == c
        dil = dilate(member)                                                           # :37-38
        pts = []
        for x, y in outer_contour_simple(dil):                                         # :41-42, :49-57
            if 0 <= y < H and 0 <= x < W:
                z = depth_np[y, x]
                if z == EMPTY_DEPTH:
                    continue
                pts.append([x * h_angle, y * v_angle, float(z)])
        if not pts:
            continue                                                                   # :60-61
        vol = convex_hull_volume(np.array(pts)) if len(pts) >= 4 else 0.0              # :69-78
        cluster_invisibility = invisibility_np[member]
        inv_sums.append(np.sum(cluster_invisibility * vol))                            # :81
        vol_sums.append(vol)
    return np.sum(np.array(inv_sums)), np.sum(np.array(vol_sums))


def panorama(seed, H=150, W=360):
    """Seeded (depth [H,W,1], invisibility [H,W]) pair shaped like get_global_invisibility's panorama
    (src/mapper/splatam/__init__.py:721-738): a few unexplored regions (invisibility near 1) of different size, some
    touching the image border, speckle below the cluster size, depth varying smoothly with patches of the
    rasterizer's empty-pixel depth 15."""
    rng = np.random.default_rng(1000 + seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    inv = np.zeros((H, W), np.float32)
    for _ in range(int(rng.integers(2, 7))):
        cy, cx = rng.uniform(-5, H + 5), rng.uniform(-5, W + 5)
        sy, sx = rng.uniform(3.0, 16.0), rng.uniform(3.0, 30.0)
        inv += rng.uniform(0.8, 1.6) * np.exp(-0.5 * (((yy - cy) / sy) ** 4 + ((xx - cx) / sx) ** 4)).astype(np.float32)
    inv += (rng.uniform(0, 1, (H, W)) < 0.01).astype(np.float32)
    inv = np.clip(inv, 0.0, 1.0).astype(np.float32)
    depth = (2.0 + 1.5 * np.sin(xx / 37.0 + rng.uniform(0, 6)) + 0.8 * np.cos(yy / 23.0 + rng.uniform(0, 6))
             + 0.05 * rng.standard_normal((H, W))).astype(np.float32)
    for _ in range(int(rng.integers(0, 4))):
        cy, cx = int(rng.integers(0, H)), int(rng.integers(0, W))
        depth[max(cy - 12, 0):cy + 12, max(cx - 20, 0):cx + 20] = EMPTY_DEPTH
    return depth[..., None], inv
