"""The GetOpacity service, as data in / data out (no ROS).

Wire format: srv/GetOpacity.srv -- request {bool arrived_flag, geometry_msgs/Point[] nodes, int32[] nodes_id},
response {geometry_msgs/Pose[] targets_frustums, float32[] targets_frustums_invisibility,
float32[] targets_frustums_volume}.  Handler: src/visualizer/visualizer.py:2180-2221.

* LOCAL query (arrived_flag false): complete.  One invisibility value (the sum over the look-around panorama), volume
  [0], and one frustum pose: the camera turned towards the heaviest cluster of unseen pixels, converted with
  OPENCV_TO_OPENGL and matrix_to_pose like the handler does (an identity Pose when there is no such cluster), plus the
  high-loss-sample pose when the mapper has one.
* GLOBAL query (arrived_flag true): one invisibility and one volume per requested node
  (SplaTAM.get_global_invisibility -> get_convexhull_volume, src/mapper/__init__.py:8-90), all nodes in one pass on
  the device -- renders, panoramas, DBSCAN, dilation, contour, convex-hull volume -- sharded over the GPUs of the
  process group; the reference answers the same request with a sequential loop over the nodes
  (src/visualizer/visualizer.py:991-1001).
"""
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np
import torch

from . import scoring

OPENCV_TO_OPENGL = np.array([[1, 0, 0, 0], [0, -1, 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]], dtype=np.float64)  # src/utils/__init__.py:10-17


@dataclass
class Pose:
    """geometry_msgs/Pose: position (x, y, z) and orientation quaternion (x, y, z, w); defaults are ROS's zeros."""
    position: np.ndarray = field(default_factory=lambda: np.zeros(3))
    orientation_xyzw: np.ndarray = field(default_factory=lambda: np.zeros(4))


@dataclass
class GetOpacityRequest:
    arrived_flag: bool
    nodes: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))   # geometry_msgs/Point[]
    nodes_id: List[int] = field(default_factory=list)


@dataclass
class GetOpacityResponse:
    targets_frustums: List[Pose] = field(default_factory=list)
    targets_frustums_invisibility: List[float] = field(default_factory=list)
    targets_frustums_volume: List[float] = field(default_factory=list)


def matrix_to_pose(matrix):
    """src/utils/gui_utils.py:138-152 (quaternion.from_rotation_matrix on the 3x3 block: the unit quaternion with
    w >= 0 of a proper rotation)."""
    matrix = np.asarray(matrix, dtype=np.float64)
    if matrix.shape != (4, 4):
        raise ValueError("Expected a 4x4 transformation matrix")
    R = matrix[:3, :3]
    # Bar-Itzhack / eigenvector form used by the quaternion package for orthogonal input
    K = np.array([[R[0, 0] - R[1, 1] - R[2, 2], 0, 0, 0],
                  [R[1, 0] + R[0, 1], R[1, 1] - R[0, 0] - R[2, 2], 0, 0],
                  [R[2, 0] + R[0, 2], R[2, 1] + R[1, 2], R[2, 2] - R[0, 0] - R[1, 1], 0],
                  [R[1, 2] - R[2, 1], R[2, 0] - R[0, 2], R[0, 1] - R[1, 0], R[0, 0] + R[1, 1] + R[2, 2]]]) / 3.0
    w_, v_ = np.linalg.eigh(K)
    q = v_[:, np.argmax(w_)]            # (x, y, z, w) up to sign
    xyzw = np.array([-q[0], -q[1], -q[2], q[3]])
    if xyzw[3] < 0:
        xyzw = -xyzw
    return Pose(position=matrix[:3, 3].copy(), orientation_xyzw=xyzw)


def handle_local(params, view_c2w, high_loss_samples_pose_c2w=None, cluster_invisibility_threshold=30):
    """The LOCAL branch (visualizer.py:2198-2213): SplaTAM.get_local_invisibility on the device, then the response."""
    invisibility, best_pose = scoring.get_local_invisibility(params, view_c2w, cluster_invisibility_threshold)
    resp = GetOpacityResponse()
    if best_pose is None:
        resp.targets_frustums.append(Pose())
    else:
        resp.targets_frustums.append(matrix_to_pose(best_pose @ OPENCV_TO_OPENGL))   # z axis facing backward
    if high_loss_samples_pose_c2w is not None:
        resp.targets_frustums.append(matrix_to_pose(np.asarray(high_loss_samples_pose_c2w) @ OPENCV_TO_OPENGL))
    resp.targets_frustums_invisibility = [float(invisibility)]   # just one node
    resp.targets_frustums_volume = [0.0]
    return resp


def handle_global(params, view_c2w, nodes, group=None, scale_modifier=1.0):
    """The GLOBAL branch (visualizer.py:2183-2195 + the node loop :991-1001): every node's invisibility and volume in
    one pass on the device (scoring.get_global_invisibility), nodes sharded over the ranks of `group`."""
    invisibility, volume = scoring.get_global_invisibility(params, view_c2w, nodes, group=group,
                                                           scale_modifier=scale_modifier)
    resp = GetOpacityResponse()
    resp.targets_frustums_invisibility = [float(x) for x in invisibility.cpu()]
    resp.targets_frustums_volume = [float(x) for x in volume.cpu()]
    return resp


def handle_get_opacity(req: GetOpacityRequest, params, view_c2w, high_loss_samples_pose_c2w=None,
                       cluster_invisibility_threshold=30, group=None) -> Optional[GetOpacityResponse]:
    """visualizer.py:2180-2221: `arrived_flag` selects the GLOBAL branch (one invisibility + one volume per requested
    node, no frustum poses) or the LOCAL branch (the current camera's look-around, best pose towards the heaviest
    cluster)."""
    if not req.arrived_flag:
        return handle_local(params, view_c2w, high_loss_samples_pose_c2w, cluster_invisibility_threshold)
    nodes = np.array([[p.x, p.y, p.z] if hasattr(p, "x") else list(p) for p in req.nodes], dtype=np.float64).reshape(-1, 3)
    return handle_global(params, view_c2w, nodes, group=group)
