"""CPU restatement of the clustering that follows the GetOpacity renders.

TEST INFRASTRUCTURE ONLY (imported by tests/ and by the golden-vector script).  The product path
(activesplat_b200/scoring.py -> as_dbscan_grid) never imports this module.

Follows src/mapper/__init__.py:92-117 (`get_invisibility_clusters`) line by line, with the same third-party call
(sklearn.cluster.DBSCAN; scikit-learn is installed in this image).  Pinned by tests/golden/ref_clusters.npz, which
holds outputs of the reference's own function on seeded images (tests/golden/make_clusters_golden.py).
"""
import numpy as np


def dbscan_labels_image(values, value_threshold, eps, min_samples):
    """Labels of sklearn's DBSCAN over np.where(values > thr) as an image: -2 = not a sample, -1 = noise."""
    from sklearn.cluster import DBSCAN
    pts = np.column_stack(np.where(values > value_threshold))  # src/mapper/__init__.py:93
    img = np.full(values.shape, -2, dtype=np.int32)
    if len(pts) == 0:
        return img
    lab = DBSCAN(eps=eps, min_samples=min_samples).fit_predict(pts)  # :99-100
    img[pts[:, 0], pts[:, 1]] = lab
    return img


def get_invisibility_clusters(invisibility_np, cluster_invisibility_threshold=30):
    """src/mapper/__init__.py:92-117."""
    from sklearn.cluster import DBSCAN
    low_visibility_points = np.column_stack(np.where(invisibility_np > 0.3))
    if len(low_visibility_points) == 0:
        return [], []
    clusters = DBSCAN(eps=5, min_samples=10).fit_predict(low_visibility_points)
    cluster_centers, cluster_invisibilities = [], []
    for cluster in sorted(set(clusters)):  # the reference iterates set(clusters): ascending for these small ints
        if cluster != -1:
            points = low_visibility_points[clusters == cluster]
            center = points.mean(axis=0)
            invisibility_sum = np.sum(invisibility_np[points[:, 0], points[:, 1]])
            if invisibility_sum > cluster_invisibility_threshold:
                cluster_centers.append(center)
                cluster_invisibilities.append(invisibility_sum)
    return cluster_centers, cluster_invisibilities


def blob_image(seed, H=75, W=180, n_blobs=6, noise=0.25):
    """Seeded invisibility-like test image in [0,1]: a few soft blobs of different size plus speckle, so that
    core/border/noise pixels, touching clusters and sub-threshold clusters all occur."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    img = np.zeros((H, W), np.float32)
    for _ in range(n_blobs):
        cy, cx = rng.uniform(0, H), rng.uniform(0, W)
        sy, sx = rng.uniform(1.5, 9.0), rng.uniform(1.5, 14.0)
        img += rng.uniform(0.3, 1.0) * np.exp(-0.5 * (((yy - cy) / sy) ** 2 + ((xx - cx) / sx) ** 2)).astype(np.float32)
    img += (rng.uniform(0, 1, (H, W)) < 0.04).astype(np.float32) * rng.uniform(0.3, 1.0, (H, W)).astype(np.float32) * noise * 4
    return np.clip(img, 0.0, 1.0).astype(np.float32)
