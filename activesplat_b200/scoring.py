"""What ActiveSplat does with a GetOpacity render, on the device.

Mirrors, by name and argument meaning:
  get_invisibility_clusters      src/mapper/__init__.py:92-117   (DBSCAN eps 5 / min_samples 10 on invisibility > 0.3)
  the LOCAL branch of GetOpacity src/mapper/splatam/__init__.py:762-838 (`get_local_invisibility`): panorama of three
      120x150 views (hfov 120, vfov 150: one pixel per degree), sum of invisibility, 2x area down-sampling, clusters,
      and the camera rotation towards the heaviest cluster.
The clustering runs in libactivesplat_b200.so (`as_dbscan_grid`); there is no CPU fallback.  The GLOBAL branch's
convex-hull volume (get_convexhull_volume: OpenCV contour tracing + qhull) is not ported; its DBSCAN step is this
same kernel with (eps 5, min_samples 25, threshold 0.8).
"""
import ctypes

import numpy as np
import torch

from . import capi


def dbscan_grid(values, value_threshold, eps=5.0, min_samples=10, max_clusters=256):
    """DBSCAN over the pixels of `values` ([H,W] or [B,H,W] CUDA float tensor) that exceed value_threshold.

    Returns (labels [B,H,W] int32: sklearn's labels, -2 where the pixel is not a sample,
             n_clusters [B] int32, stats_int [B,max_clusters,3] int32 = count / sum rows / sum cols,
             stats_value [B,max_clusters] float64 = sum of values)."""
    if not values.is_cuda:
        raise RuntimeError("dbscan_grid needs a CUDA tensor (there is no CPU path)")
    v = values.detach()
    if v.dim() == 2:
        v = v.unsqueeze(0)
    if v.dtype != torch.float32 or not v.is_contiguous():
        v = v.contiguous().float()
    B, H, W = v.shape
    dev = v.device
    with torch.cuda.device(dev):
        labels = torch.empty((B, H, W), dtype=torch.int32, device=dev)
        scratch = torch.empty((B, H, W), dtype=torch.int32, device=dev)
        n_clusters = torch.zeros((B,), dtype=torch.int32, device=dev)
        stats_i = torch.zeros((B, max_clusters, 3), dtype=torch.int32, device=dev)
        stats_v = torch.zeros((B, max_clusters), dtype=torch.float64, device=dev)
        capi.check(capi.lib().as_dbscan_grid(B, H, W, v.data_ptr(), float(value_threshold), float(eps), int(min_samples),
                                             labels.data_ptr(), scratch.data_ptr(), int(max_clusters),
                                             n_clusters.data_ptr(), stats_i.data_ptr(), stats_v.data_ptr(),
                                             torch.cuda.current_stream(dev).cuda_stream))
    return labels, n_clusters, stats_i, stats_v


def get_invisibility_clusters(invisibility, cluster_invisibility_threshold=30, max_clusters=256):
    """src/mapper/__init__.py:92-117 for one image (CUDA tensor [H,W]).  Returns (cluster_centers,
    cluster_invisibilities): lists, clusters in ascending label order, centre = mean (row, col) of the members,
    kept when the summed invisibility exceeds the threshold.  One device->host read of the per-cluster statistics."""
    _, n, si, sv = dbscan_grid(invisibility, 0.3, eps=5.0, min_samples=10, max_clusters=max_clusters)
    n = int(n[0])
    if n > max_clusters:
        return get_invisibility_clusters(invisibility, cluster_invisibility_threshold, max_clusters=2 * n)
    si = si[0, :n].cpu().numpy().astype(np.float64)
    sv = sv[0, :n].cpu().numpy()
    centers, sums = [], []
    for c in range(n):
        if sv[c] > cluster_invisibility_threshold:
            centers.append(np.array([si[c, 1] / si[c, 0], si[c, 2] / si[c, 0]]))
            sums.append(np.float32(sv[c]))
    return centers, sums


def area_downsample_2x(img):
    """cv2.resize(img, (W//2, H//2), interpolation=cv2.INTER_AREA) for even H, W: the mean of each 2x2 block
    (src/mapper/splatam/__init__.py:810-813 uses factor 0.5 on the 150x360 panorama)."""
    H, W = img.shape[-2:]
    x = img[..., : H // 2 * 2, : W // 2 * 2]
    return 0.25 * (x[..., 0::2, 0::2] + x[..., 0::2, 1::2] + x[..., 1::2, 0::2] + x[..., 1::2, 1::2])


def rot_axis(c2w, axis, angle):
    """Rotate a camera-to-world pose about one of its own axes (src/utils/pose_utils.py:13-44 `rot_axis` as used at
    src/mapper/splatam/__init__.py:717,781,829-830): c2w @ R_axis(angle)."""
    c, s = np.cos(angle), np.sin(angle)
    R = np.eye(4)
    if axis == 'x':
        R[1, 1], R[1, 2], R[2, 1], R[2, 2] = c, -s, s, c
    elif axis == 'y':
        R[0, 0], R[0, 2], R[2, 0], R[2, 2] = c, s, -s, c
    elif axis == 'z':
        R[0, 0], R[0, 1], R[1, 0], R[1, 1] = c, -s, s, c
    else:
        raise ValueError(axis)
    return np.asarray(c2w, dtype=np.float64) @ R


def local_invisibility_from_panorama(opacity_pano, view_c2w, cluster_invisibility_threshold=30, width=120, height=150):
    """The part of get_local_invisibility after the renders (src/mapper/splatam/__init__.py:793-832).
    opacity_pano: [150, 360] CUDA tensor (the three views side by side).  Returns (sum_invisibility, best_pose_c2w or
    None)."""
    inv = 1.0 - opacity_pano
    sum_invisibility = float(inv.sum())
    best_pose_c2w = None
    if sum_invisibility > 100:
        factor = 0.5
        small = area_downsample_2x(inv)
        centers, sums = get_invisibility_clusters(small, cluster_invisibility_threshold)
        if len(sums) > 0:
            c = centers[int(np.argmax(sums))]
            center_vec = np.array([c[1] / factor - width / 2, c[0] / factor - height / 2])
            horizontal_angle, vertical_angle = np.deg2rad(center_vec[0]), np.deg2rad(center_vec[1])
            # the middle of the field of view has been seen already
            if np.abs(horizontal_angle) > np.deg2rad(15) or np.abs(vertical_angle) > np.deg2rad(15):
                best_pose_c2w = rot_axis(rot_axis(view_c2w, 'y', horizontal_angle), 'x', vertical_angle)
    return sum_invisibility, best_pose_c2w


def look_around_views(view_c2w, position=None, hfov=120, vfov=150, width=120, height=150):
    """The camera sweep of get_local_invisibility / get_global_invisibility (src/mapper/splatam/__init__.py:707-719,
    771-783): int(360 / hfov) views rotated about the camera's own y axis, one pixel per degree.  `position` (x, -, z)
    moves the camera to a Voronoi node at the agent's camera height (:703-704).  Returns CameraSpec-like views."""
    from .scenes import CameraSpec, compute_intrinsics
    c2w = np.array(view_c2w, dtype=np.float64)
    if position is not None:
        c2w[0, 3] = position[0]
        c2w[2, 3] = position[2]
    fx, fy, cx, cy = compute_intrinsics(width, height, np.deg2rad(hfov), np.deg2rad(vfov))
    views = []
    for i in range(int(360 / hfov)):
        c2w_i = rot_axis(c2w, 'y', np.deg2rad(hfov * i))
        views.append(CameraSpec(width, height, fx, fy, cx, cy, np.linalg.inv(c2w_i)))
    return views


def get_local_invisibility(params, view_c2w, cluster_invisibility_threshold=30, scale_modifier=1.0):
    """SplaTAM.get_local_invisibility (src/mapper/splatam/__init__.py:762-838) without the visualisation: renders the
    three look-around views of the current camera, stacks their opacity side by side, and returns
    (sum_invisibility, best_pose_c2w or None) -- what the LOCAL branch of the GetOpacity service puts into
    targets_frustums_invisibility / targets_frustums (src/visualizer/visualizer.py:2198-2211)."""
    from . import dist as asd
    views = look_around_views(view_c2w)
    _, ops = asd.score_candidate_views(params, views, n_streams=len(views), scale_modifier=scale_modifier,
                                       return_opacity=True)
    pano = torch.cat([ops[i] for i in range(len(views))], dim=1)  # np.hstack of the per-view opacity images (:789)
    return local_invisibility_from_panorama(pano, view_c2w, cluster_invisibility_threshold,
                                            width=views[0].width, height=views[0].height)
