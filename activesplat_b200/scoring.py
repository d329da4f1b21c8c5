"""What ActiveSplat does with a GetOpacity render, on the device.

Mirrors, by name and argument meaning:
  get_invisibility_clusters      src/mapper/__init__.py:92-117   (DBSCAN eps 5 / min_samples 10 on invisibility > 0.3)
  the LOCAL branch of GetOpacity src/mapper/splatam/__init__.py:762-838 (`get_local_invisibility`): panorama of three
      120x150 views (hfov 120, vfov 150: one pixel per degree), sum of invisibility, 2x area down-sampling, clusters,
      and the camera rotation towards the heaviest cluster.
  get_convexhull_volume          src/mapper/__init__.py:8-90     (GLOBAL branch: DBSCAN eps 5 / min_samples 25 on
      invisibility > 0.8, per cluster dilation + outer contour + back-projection + 3-D convex-hull volume)
  the GLOBAL branch of GetOpacity src/mapper/splatam/__init__.py:698-759 (`get_global_invisibility`), all nodes at once
The clustering and the hull volumes run in libactivesplat_b200.so (`as_dbscan_grid`, `as_cluster_hull_volumes`); there
is no CPU fallback.
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


def get_convexhull_volume(depth, invisibility, vfov=150.0, max_clusters=32):
    """src/mapper/__init__.py:8-90 for a batch of panoramas on the device.
    depth [B,H,W] (median depth, 15 where no surface), invisibility [B,H,W] (1 - opacity), CUDA float tensors.
    Returns (last_invisibility [B] float64, last_volume [B] float64, n_clusters [B] int32, n_points [B,max_clusters]
    int32): per panorama the sum over its clusters of sum(invisibility of the members) * hull volume and the sum of
    the hull volumes, plus the cluster count (a count above max_clusters means clusters were dropped: call again with
    more) and the number of outline points each hull was built from.  No host synchronisation."""
    if not invisibility.is_cuda:
        raise RuntimeError("get_convexhull_volume needs CUDA tensors (there is no CPU path)")
    inv = invisibility.detach()
    dep = depth.detach()
    if inv.dim() == 2:
        inv = inv.unsqueeze(0)
    dep = dep.reshape(inv.shape)
    if dep.dtype != torch.float32 or not dep.is_contiguous():
        dep = dep.contiguous().float()
    B, H, W = inv.shape
    labels, n_clusters, _, stats_v = dbscan_grid(inv, 0.8, eps=5.0, min_samples=25, max_clusters=max_clusters)
    dev = inv.device
    with torch.cuda.device(dev):
        volumes = torch.empty((B, max_clusters), dtype=torch.float64, device=dev)
        n_points = torch.empty((B, max_clusters), dtype=torch.int32, device=dev)
        capi.check(capi.lib().as_cluster_hull_volumes(B, H, W, labels.data_ptr(), n_clusters.data_ptr(), dep.data_ptr(),
                                                      float(vfov), int(max_clusters), volumes.data_ptr(),
                                                      n_points.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
    used = n_points != 0     # a cluster whose whole outline lies on "no surface" pixels is skipped (:60-61)
    last_invisibility = (stats_v * volumes * used).sum(1)
    last_volume = (volumes * used).sum(1)
    return last_invisibility, last_volume, n_clusters, n_points


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


def get_global_invisibility(params, view_c2w, positions, group=None, scale_modifier=1.0, max_clusters=32):
    """SplaTAM.get_global_invisibility (src/mapper/splatam/__init__.py:698-759) for ALL Voronoi nodes of a GLOBAL
    GetOpacity query at once (the reference loops over them, src/visualizer/visualizer.py:991-1001): per node the three
    look-around renders at the node's position and the agent's camera height, the 150x360 panoramas of invisibility
    and median depth, DBSCAN clusters and convex-hull volumes -- all on the device, nodes sharded over the ranks of
    `group`.  Returns (invisibility [n_nodes], volume [n_nodes]) float64 tensors on every rank; a node at the origin
    is skipped like the reference does (:701-702) and scores (0, 0)."""
    from . import dist as asd
    positions = np.asarray(positions, dtype=np.float64).reshape(-1, 3)
    n_nodes = len(positions)
    dev = params["means3D"].device
    rank, world = asd.world_info(group)
    live = [i for i in range(n_nodes) if not (positions[i] == 0).all()]
    mine = live[rank::world]
    local = torch.zeros((len(mine), 2), dtype=torch.float64, device=dev)
    if mine:
        views = []
        for i in mine:
            views += look_around_views(view_c2w, position=positions[i])
        n_yaw = len(views) // len(mine)
        _, ops, depths = asd.score_candidate_views(params, views, n_streams=min(32, len(views)), return_opacity=True,
                                                   scale_modifier=scale_modifier, return_depth=True, shard=False)
        H, W = ops.shape[1:]
        # np.hstack of the per-yaw images (:731-736): view j of a node fills columns [j*W, (j+1)*W)
        inv = (1.0 - ops).view(len(mine), n_yaw, H, W).permute(0, 2, 1, 3).reshape(len(mine), H, n_yaw * W).contiguous()
        dep = depths.view(len(mine), n_yaw, H, W).permute(0, 2, 1, 3).reshape(len(mine), H, n_yaw * W).contiguous()
        li, lv, n_cl, _ = get_convexhull_volume(dep, inv, vfov=150.0, max_clusters=max_clusters)
        if int(n_cl.max()) > max_clusters:   # one read per query; the scores go to the host right after anyway
            li, lv, _, _ = get_convexhull_volume(dep, inv, vfov=150.0, max_clusters=int(n_cl.max()))
        local = torch.stack([li, lv], 1)
    out = torch.zeros((n_nodes, 2), dtype=torch.float64, device=dev)
    if world == 1:
        out[mine] = local
    else:
        n_max = (len(live) + world - 1) // world
        pad = torch.zeros((n_max, 2), dtype=torch.float64, device=dev)
        pad[: len(mine)] = local
        allr = torch.empty((world, n_max, 2), dtype=torch.float64, device=dev)
        torch.distributed.all_gather_into_tensor(allr.view(world * n_max, 2), pad, group=group)
        for r in range(world):
            idx = live[r::world]
            out[idx] = allr[r, : len(idx)]
    return out[:, 0], out[:, 1]
