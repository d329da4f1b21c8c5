"""Seeded synthetic scenes and cameras for parity tests and benchmarks.

Everything is generated on the CPU with a seeded torch.Generator so that the same bytes reach the
CPU oracle, the reference CUDA build and the sm_100a kernels.  No dataset, no simulator.

Camera conventions follow the reference:
  * intrinsics: compute_intrinsics (src/dataloader/__init__.py:275-284): cx = W/2 - 1, cy = H/2 - 1
  * raster settings: setup_camera (src/mapper/splatam/utils/recon_helpers.py:4-28)
"""
import math
from typing import NamedTuple

import numpy as np
import torch


class CameraSpec(NamedTuple):
    """Plain-number camera: what setup_camera (recon_helpers.py:4-28) derives from (w, h, k, w2c)."""
    width: int
    height: int
    fx: float
    fy: float
    cx: float
    cy: float
    w2c: np.ndarray  # [4,4] float64 world->camera
    near: float = 0.01
    far: float = 100.0

    @property
    def tanfovx(self):
        return self.width / (2 * self.fx)

    @property
    def tanfovy(self):
        return self.height / (2 * self.fy)

    def k(self):
        return np.array([[self.fx, 0, self.cx], [0, self.fy, self.cy], [0, 0, 1]], dtype=np.float64)

    def matrices(self):
        """(viewmatrix [4,4] = w2c^T, projmatrix [4,4] = (P_gl w2c)^T, campos [3]) as float32 numpy,
        computed the way recon_helpers.py:6-14 does (fp32 matmul)."""
        w2c = torch.tensor(self.w2c).float()
        cam_center = torch.inverse(w2c)[:3, 3]
        w, h, fx, fy, cx, cy, near, far = (self.width, self.height, self.fx, self.fy, self.cx, self.cy,
                                           self.near, self.far)
        opengl_proj = torch.tensor([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                                    [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                                    [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                                    [0.0, 0.0, 1.0, 0.0]]).float().unsqueeze(0).transpose(1, 2)
        view = w2c.unsqueeze(0).transpose(1, 2)
        full = view.bmm(opengl_proj)
        return view[0].contiguous().numpy(), full[0].contiguous().numpy(), cam_center.numpy()


def compute_intrinsics(width, height, hfov_rad, vfov_rad=None):
    """src/dataloader/__init__.py:275-284."""
    fx = 0.5 * width / np.tan(hfov_rad / 2.)
    fy = fx if vfov_rad is None else 0.5 * height / np.tan(vfov_rad / 2.)
    return fx, fy, width / 2 - 1, height / 2 - 1


def frustum_camera(W, H):
    """Camera at the origin looking down +z, fx = fy = W/2 (tanfovx = 1)."""
    return CameraSpec(W, H, W / 2.0, W / 2.0, W / 2 - 1, H / 2 - 1, np.eye(4))


def look_around_camera(position, yaw_rad, width=120, height=150, hfov_deg=120.0, vfov_deg=150.0):
    """The GetOpacity candidate-view camera (src/mapper/splatam/__init__.py:707-716):
    120x150 px, hfov 120 deg, vfov 150 deg, rotated about the camera y axis."""
    fx, fy, cx, cy = compute_intrinsics(width, height, math.radians(hfov_deg), math.radians(vfov_deg))
    c, s = math.cos(yaw_rad), math.sin(yaw_rad)
    c2w = np.eye(4)
    c2w[:3, :3] = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    c2w[:3, 3] = position
    return CameraSpec(width, height, fx, fy, cx, cy, np.linalg.inv(c2w))


def _gen(seed):
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def _finish_params(pts_world, z_cam, fx, g, rgb=None):
    """SplaTAM-style parameters (splatam.py:77-95): 1-pixel projective initial scale with spread."""
    P = pts_world.shape[0]
    log_scales = torch.log(z_cam.clamp_min(0.05) / fx).unsqueeze(1) + 0.5 * torch.randn(P, 3, generator=g)
    rots = torch.randn(P, 4, generator=g)
    rots = rots / rots.norm(dim=1, keepdim=True)
    logit_op = 2.0 * torch.randn(P, 1, generator=g)
    if rgb is None:
        rgb = torch.rand(P, 3, generator=g)
    return {
        "means3D": pts_world.float().contiguous(),
        "rgb_colors": rgb.float().contiguous(),
        "unnorm_rotations": rots.float().contiguous(),
        "logit_opacities": logit_op.float().contiguous(),
        "log_scales": log_scales.float().contiguous(),
    }


def s_frustum(P, W, H, seed):
    """S-frustum(P, W, H, seed): points back-projected from pixels ~U over the image inflated by 10 %,
    depth ~U[0.5, 10] m, 5 % of the points at z in [0, 0.3] to exercise the near cull.
    Returns (params dict, CameraSpec)."""
    g = _gen(seed)
    cam = frustum_camera(W, H)
    u = (torch.rand(P, generator=g) * 1.2 - 0.1) * W
    v = (torch.rand(P, generator=g) * 1.2 - 0.1) * H
    z = 0.5 + 9.5 * torch.rand(P, generator=g)
    near = torch.rand(P, generator=g) < 0.05
    z = torch.where(near, 0.3 * torch.rand(P, generator=g), z)
    x = (u - cam.cx) / cam.fx * z
    y = (v - cam.cy) / cam.fy * z
    pts = torch.stack([x, y, z], dim=1)
    return _finish_params(pts, z, cam.fx, g), cam


def room_pose(k, K, radius=1.5, height=0.0):
    """k-th of K camera poses on a circle inside the room, looking outward (+ a little jitter-free)."""
    ang = 2 * math.pi * k / max(K, 1)
    pos = np.array([radius * math.cos(ang), height, radius * math.sin(ang)])
    # camera z axis points outward along the radius, y down
    yaw = math.atan2(math.cos(ang), math.sin(ang))
    c, s = math.cos(yaw), math.sin(yaw)
    c2w = np.eye(4)
    c2w[:3, :3] = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    c2w[:3, 3] = pos
    return c2w


def s_room(P, K, seed, W=640, H=480, hfov_deg=90.0, scale_mult=1.0):
    """S-room(P, K, seed): points on the 6 faces of a 10 x 3 x 8 m box seen from K poses on a circle.
    Returns (params dict, [CameraSpec] * K).  Scales follow the 1-pixel rule w.r.t. camera 0's focal
    length at the point's distance from the room centre."""
    g = _gen(seed)
    half = torch.tensor([5.0, 1.5, 4.0])
    face = torch.randint(0, 6, (P,), generator=g)
    uvw = (torch.rand(P, 3, generator=g) * 2 - 1) * half
    axis = face // 2
    sign = (face % 2).float() * 2 - 1
    pts = uvw.clone()
    pts[torch.arange(P), axis] = sign * half[axis]
    fx, fy, cx, cy = compute_intrinsics(W, H, math.radians(hfov_deg))
    dist = pts.norm(dim=1).clamp_min(0.5)
    params = _finish_params(pts, dist * scale_mult, fx, g)
    cams = [CameraSpec(W, H, fx, fy, cx, cy, np.linalg.inv(room_pose(k, K))) for k in range(K)]
    return params, cams


def candidate_views(n_views, seed=0):
    """GetOpacity candidate poses: node positions on a jittered grid inside the room x 3 yaws
    (0/120/240 deg), truncated to n_views."""
    g = _gen(seed)
    n_nodes = (n_views + 2) // 3
    pos = (torch.rand(n_nodes, 3, generator=g) * 2 - 1) * torch.tensor([3.5, 0.0, 2.5])
    cams = []
    for n in range(n_nodes):
        for i in range(3):
            if len(cams) < n_views:
                cams.append(look_around_camera(pos[n].numpy().astype(np.float64), math.radians(120.0 * i)))
    return cams


def perturbed(params, seed, sigma_means=0.01, sigma_rgb=0.05):
    """A perturbed copy used to render 'ground truth' so that losses have non-trivial gradients."""
    g = _gen(seed)
    out = {k: v.clone() for k, v in params.items()}
    out["means3D"] = out["means3D"] + sigma_means * torch.randn(out["means3D"].shape, generator=g)
    out["rgb_colors"] = (out["rgb_colors"] + sigma_rgb * torch.randn(out["rgb_colors"].shape, generator=g)).clamp(0, 1)
    return out


def activated(params):
    """get_rendervars-style activations (splatam.py:436-468) on CPU tensors: what the rasterizer is fed."""
    return {
        "means3D": params["means3D"],
        "colors_precomp": params["rgb_colors"],
        "rotations": torch.nn.functional.normalize(params["unnorm_rotations"]),
        "opacities": torch.sigmoid(params["logit_opacities"]),
        "scales": torch.exp(params["log_scales"] if params["log_scales"].shape[1] == 3
                            else torch.tile(params["log_scales"], (1, 3))),
    }


def rotmat_to_quat(Rm):
    """(r, x, y, z) unit quaternion of a 3x3 rotation (numpy, float64)."""
    Rm = np.asarray(Rm, dtype=np.float64)
    t = np.trace(Rm)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = [0.25 * s, (Rm[2, 1] - Rm[1, 2]) / s, (Rm[0, 2] - Rm[2, 0]) / s, (Rm[1, 0] - Rm[0, 1]) / s]
    else:
        i = int(np.argmax(np.diag(Rm)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = math.sqrt(1.0 + Rm[i, i] - Rm[j, j] - Rm[k, k]) * 2
        q = [0.0, 0.0, 0.0, 0.0]
        q[0] = (Rm[k, j] - Rm[j, k]) / s
        q[1 + i] = 0.25 * s
        q[1 + j] = (Rm[j, i] + Rm[i, j]) / s
        q[1 + k] = (Rm[k, i] + Rm[i, k]) / s
    q = np.array(q)
    return q / np.linalg.norm(q)


def camera_track(cams):
    """SplaTAM camera parameters for a list of CameraSpec: cam_unnorm_rots [1,4,T], cam_trans [1,3,T]
    (world == first-frame convention of splatam.py:97-103: w2c_t = [R(q_t) | t_t])."""
    T = len(cams)
    rots = np.zeros((1, 4, T), np.float32)
    trans = np.zeros((1, 3, T), np.float32)
    for t, c in enumerate(cams):
        rots[0, :, t] = rotmat_to_quat(c.w2c[:3, :3])
        trans[0, :, t] = c.w2c[:3, 3]
    return torch.from_numpy(rots), torch.from_numpy(trans)
