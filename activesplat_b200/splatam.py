"""Host-side mirror of the SplaTAM functions that wrap the rasterizer in ActiveSplat.

Same names, arguments and return values as the reference, so the mapper's call sites read the same:
  setup_camera                 src/mapper/splatam/utils/recon_helpers.py:4-28
  build_rotation               src/mapper/splatam/utils/slam_external.py:25-42
  quat_mult                    src/mapper/splatam/utils/slam_helpers.py:21-28
  transform_to_frame           src/mapper/splatam/utils/slam_helpers.py:252-304
  get_depth_and_silhouette     src/mapper/splatam/utils/slam_helpers.py:196-213
  transformed_params2rendervar / transformed_params2depthplussilhouette   slam_helpers.py:124-139, 234-249
  get_rendervars, render       src/mapper/splatam/splatam.py:436-468, 413-434
  get_loss                     src/mapper/splatam/splatam.py:172-301
  calc_ssim                    src/mapper/splatam/utils/slam_external.py:54-97
  initialize_optimizer         src/mapper/splatam/splatam.py:118-124
  get_pointcloud, initialize_new_params, add_new_gaussians   src/mapper/splatam/splatam.py:25-75, 304-379

What is different underneath: the reference renders every frame twice (RGB pass, then a [z,1,z^2]
pass over identical geometry, splatam.py:208,212); here both feature sets go through ONE 6-channel
rasterizer call (preprocess, sort and alpha evaluation shared), which is arithmetically identical.
All tensors live on the device of `params`; nothing here touches the CPU oracle.
"""
import numpy as np
import torch
import torch.nn.functional as F

from .rasterizer import GaussianRasterizationSettings as Camera
from .rasterizer import GaussianRasterizer as Renderer


def setup_camera(w, h, k, w2c, near=0.01, far=100, scale_modifier=1.0, device="cuda", bg=(0.0, 0.0, 0.0)):
    fx, fy, cx, cy = k[0][0], k[1][1], k[0][2], k[1][2]
    w2c = torch.as_tensor(np.asarray(w2c)).to(device).float()
    cam_center = torch.inverse(w2c)[:3, 3]
    w2c = w2c.unsqueeze(0).transpose(1, 2)
    opengl_proj = torch.tensor([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                                [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                                [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                                [0.0, 0.0, 1.0, 0.0]]).to(device).float().unsqueeze(0).transpose(1, 2)
    full_proj = w2c.bmm(opengl_proj)
    return Camera(image_height=h, image_width=w, tanfovx=w / (2 * fx), tanfovy=h / (2 * fy),
                  bg=torch.tensor(bg, dtype=torch.float32, device=device), scale_modifier=scale_modifier,
                  viewmatrix=w2c, projmatrix=full_proj, sh_degree=0, campos=cam_center, prefiltered=False,
                  debug=False)


def with_background(cam, bg):
    """A copy of `cam` with another background (the white-bg camera of splatam.py:416-429)."""
    return cam._replace(bg=torch.as_tensor(bg, dtype=torch.float32, device=cam.bg.device))


def build_rotation(q):
    norm = torch.sqrt(q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2] + q[:, 3] * q[:, 3])
    q = q / norm[:, None]
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rot = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=-1)
    return rot.reshape(-1, 3, 3)


def quat_mult(q1, q2):
    w1, x1, y1, z1 = q1.T
    w2, x2, y2, z2 = q2.T
    w = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2
    x = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2
    y = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2
    z = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2
    return torch.stack([w, x, y, z]).T


def l1_loss_v1(x, y):
    return torch.abs((x - y)).mean()


def transform_to_frame(params, time_idx, gaussians_grad, camera_grad):
    if camera_grad:
        cam_rot = F.normalize(params['cam_unnorm_rots'][..., time_idx])
        cam_tran = params['cam_trans'][..., time_idx]
    else:
        cam_rot = F.normalize(params['cam_unnorm_rots'][..., time_idx].detach())
        cam_tran = params['cam_trans'][..., time_idx].detach()
    dev = params['means3D'].device
    rot = build_rotation(cam_rot)[0]
    top = torch.cat([rot, cam_tran.reshape(3, 1)], dim=1)
    rel_w2c = torch.cat([top, torch.tensor([[0.0, 0.0, 0.0, 1.0]], device=dev)], dim=0)
    transform_rots = params['log_scales'].shape[1] != 1  # anisotropic Gaussians carry their own rotation
    if gaussians_grad:
        pts, unnorm_rots = params['means3D'], params['unnorm_rotations']
    else:
        pts, unnorm_rots = params['means3D'].detach(), params['unnorm_rotations'].detach()
    pts4 = torch.cat((pts, torch.ones(pts.shape[0], 1, device=dev)), dim=1)
    out = {'means3D': (rel_w2c @ pts4.T).T[:, :3]}
    if transform_rots:
        out['unnorm_rotations'] = quat_mult(cam_rot, F.normalize(unnorm_rots))
    else:
        out['unnorm_rotations'] = unnorm_rots
    return out


def get_depth_and_silhouette(pts_3D, w2c):
    pts4 = torch.cat((pts_3D, torch.ones_like(pts_3D[:, :1])), dim=-1)
    pts_in_cam = (w2c @ pts4.transpose(0, 1)).transpose(0, 1)
    depth_z = pts_in_cam[:, 2].unsqueeze(-1)
    return torch.cat((depth_z, torch.ones_like(depth_z), torch.square(depth_z)), dim=-1)


def _log_scales3(params):
    ls = params['log_scales']
    return torch.tile(ls, (1, 3)) if ls.shape[1] == 1 else ls


def transformed_params2rendervar(params, transformed_gaussians):
    return {
        'means3D': transformed_gaussians['means3D'],
        'colors_precomp': params['rgb_colors'],
        'rotations': F.normalize(transformed_gaussians['unnorm_rotations']),
        'opacities': torch.sigmoid(params['logit_opacities']),
        'scales': torch.exp(_log_scales3(params)),
        'means2D': torch.zeros_like(params['means3D'], requires_grad=True) + 0,
    }


def transformed_params2depthplussilhouette(params, w2c, transformed_gaussians):
    rv = transformed_params2rendervar(params, transformed_gaussians)
    rv['colors_precomp'] = get_depth_and_silhouette(transformed_gaussians['means3D'], w2c)
    return rv


def get_rendervars(params, w2c, curr_timestep=None):
    dev = params['means3D'].device
    w2c = torch.as_tensor(np.asarray(w2c)).to(dev).float()
    base = {
        'means3D': params['means3D'],
        'rotations': F.normalize(params['unnorm_rotations']),
        'opacities': torch.sigmoid(params['logit_opacities']),
        'scales': torch.exp(_log_scales3(params)),
    }
    rendervar = dict(base, colors_precomp=params['rgb_colors'], means2D=torch.zeros_like(params['means3D']))
    depth_rendervar = dict(base, colors_precomp=get_depth_and_silhouette(params['means3D'], w2c),
                           means2D=torch.zeros_like(params['means3D']))
    return rendervar, depth_rendervar


def render_bundle(cam, rendervar, depth_sil_colors, bg_rgb=None):
    """ONE 6-channel rasterizer call in place of the reference's two passes.
    Returns (im[3,H,W], radii, median_depth[1,H,W], opacity[1,H,W], depth_sil[3,H,W])."""
    bg6 = torch.cat([cam.bg if bg_rgb is None else torch.as_tensor(bg_rgb, dtype=torch.float32, device=cam.bg.device),
                     torch.zeros(3, device=cam.bg.device)])
    cam6 = cam._replace(bg=bg6)
    colors6 = torch.cat([rendervar['colors_precomp'], depth_sil_colors], dim=1)
    out, radii, depth, opacity = Renderer(raster_settings=cam6)(
        means3D=rendervar['means3D'], means2D=rendervar['means2D'], opacities=rendervar['opacities'],
        colors_precomp=colors6, scales=rendervar['scales'], rotations=rendervar['rotations'])
    return out[:3], radii, depth, opacity, out[3:]


def render(w2c, k, timestep_data, timestep_depth_data, cfg, scale_modifier=1.0):
    """splatam.py:413-434: white-background RGB + median depth + opacity, and the silhouette."""
    with torch.no_grad():
        dev = timestep_data['means3D'].device
        cam = setup_camera(cfg['viz_w'], cfg['viz_h'], k, w2c, cfg.get('viz_near', 0.01), cfg.get('viz_far', 100),
                           scale_modifier=scale_modifier, device=dev)
        im, _, depth, opacity, depth_sil = render_bundle(cam, timestep_data, timestep_depth_data['colors_precomp'],
                                                         bg_rgb=[1.0, 1.0, 1.0])
        return im, depth, opacity, depth_sil[1, :, :].unsqueeze(0)


def _gaussian_window(window_size, sigma, channel, like):
    x = torch.arange(window_size, dtype=torch.float32)
    g = torch.exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2))
    g = (g / g.sum()).unsqueeze(1)
    w2d = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)
    return w2d.expand(channel, 1, window_size, window_size).contiguous().to(like.device).type_as(like)


def calc_ssim(img1, img2, window_size=11, size_average=True):
    channel = img1.size(-3)
    window = _gaussian_window(window_size, 1.5, channel, img1)
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + c1) * (2 * sigma12 + c2)) / ((mu1_sq + mu2_sq + c1) * (sigma1_sq + sigma2_sq + c2))
    return ssim_map.mean() if size_average else ssim_map.mean(1).mean(1).mean(1)


def get_loss(params, curr_data, variables, iter_time_idx, loss_weights, use_sil_for_loss, sil_thres, use_l1,
             ignore_outlier_depth_loss, tracking=False, mapping=False, do_ba=False, plot_dir=None,
             visualize_tracking_loss=False, tracking_iteration=None):
    if (tracking and use_l1 and not ignore_outlier_depth_loss and not visualize_tracking_loss
            and params['means3D'].is_cuda):
        return _get_loss_tracking_fused(params, curr_data, variables, iter_time_idx, loss_weights, use_sil_for_loss,
                                        sil_thres)
    if (not tracking and use_l1 and not ignore_outlier_depth_loss and params['means3D'].is_cuda):
        return _get_loss_mapping_fused(params, curr_data, variables, iter_time_idx, loss_weights, mapping and do_ba)
    # The remaining combinations (outlier-rejecting depth mask, or no L1 depth term) are not on ActiveSplat's path
    # (config/splatam/online_habitat_sim.py:26-29,52-54): the bundle still comes from the fused frame node, differentiable
    # down to the raw parameters and the pose; only the masking and the reductions are torch ops.
    if not params['means3D'].is_cuda:
        raise RuntimeError("activesplat_b200.splatam.get_loss needs CUDA tensors (there is no CPU path)")
    from .frame import fused_render
    frozen = set()
    if tracking:
        frozen |= {'means3D', 'unnorm_rotations'}          # transform_to_frame(gaussians_grad=False)
    if not (tracking or (mapping and do_ba)):
        frozen |= {'cam_unnorm_rots', 'cam_trans'}         # transform_to_frame(camera_grad=False)
    p = {k: (v.detach() if k in frozen else v) for k, v in params.items()}
    bundle, radius, _, _, _ = fused_render(p, curr_data['cam'], int(iter_time_idx), curr_data['w2c'])
    variables['means2D'] = _NO_MEANS2D
    rgb, z, sil, z_sq = bundle[0:3], bundle[3:4], bundle[4], bundle[5:6]
    gt_rgb, gt_z = curr_data['im'], curr_data['depth']
    z_err = (gt_z - z).abs()
    valid = (gt_z > 0) & ~torch.isnan(z) & ~torch.isnan((z_sq - z * z).detach())
    if ignore_outlier_depth_loss:                          # splatam.py:223-226: ten times the median error
        masked_err = z_err * (gt_z > 0)
        valid = valid & (masked_err < 10 * masked_err.median())
    if tracking and use_sil_for_loss:
        valid = valid & (sil > sil_thres)
    valid = valid.detach()
    terms = {}
    if use_l1:
        picked = z_err[valid]
        terms['depth'] = picked.sum() if tracking else picked.mean()
    if tracking:
        rgb_err = (gt_rgb - rgb).abs()
        if use_sil_for_loss or ignore_outlier_depth_loss:
            rgb_err = rgb_err[valid.expand_as(rgb_err)]
        terms['im'] = rgb_err.sum()
    else:
        terms['im'] = 0.8 * l1_loss_v1(rgb, gt_rgb) + 0.2 * (1.0 - calc_ssim(rgb, gt_rgb))
    weighted_losses = {k: v * loss_weights[k] for k, v in terms.items()}
    loss = sum(weighted_losses.values())
    # max_2D_radius[seen] = max(radius[seen], ...) without a boolean-index host sync: radius is 0 where unseen
    variables['max_2D_radius'] = torch.maximum(variables['max_2D_radius'], radius.to(variables['max_2D_radius'].dtype))
    variables['seen'] = radius > 0
    weighted_losses['loss'] = loss
    return loss, variables, weighted_losses


class _ScreenSpaceGradientNotKept:
    """What variables['means2D'] holds after a fused get_loss.  The reference keeps the dummy means2D tensor of the
    COLOUR pass there so that its (disabled by default, online_habitat_sim.py:70) Gaussian-splatting densification can
    read `.grad` (utils/slam_external.py:100-108).  The fused path renders colour and depth/silhouette in one pass and
    does not keep a colour-only screen-space gradient, so asking for it is an error rather than a wrong number."""

    @property
    def grad(self):
        raise NotImplementedError(
            "activesplat_b200: the fused get_loss path does not keep the colour-pass means2D gradient "
            "(use_gaussian_splatting_densification); render through diff_gaussian_rasterization.GaussianRasterizer "
            "(activesplat_b200.rasterizer) if that gradient is needed")


_NO_MEANS2D = _ScreenSpaceGradientNotKept()


def _get_loss_tracking_fused(params, curr_data, variables, iter_time_idx, loss_weights, use_sil_for_loss, sil_thres):
    """get_loss(tracking=True) (splatam.py:178-182, 230-247) as one fused autograd node: frame transform,
    activations, the 6-channel bundle render, the masked L1 sums and the whole backward run as a dozen
    kernels on the current stream with no host synchronisation."""
    from .frame import fused_tracking_loss
    # transform_to_frame(gaussians_grad=False, camera_grad=True): means and rotations are detached,
    # colours / opacities / scales are not (slam_helpers.py:124-139 reads them from `params` directly)
    params = dict(params, means3D=params['means3D'].detach(), unnorm_rotations=params['unnorm_rotations'].detach())
    seen = _seen_buffer(variables, params)
    out, ws = fused_tracking_loss(params, curr_data['cam'], int(iter_time_idx), curr_data['w2c'], curr_data['im'],
                                  curr_data['depth'], loss_weights['im'], loss_weights['depth'], sil_thres,
                                  use_sil_for_loss, use_sil_for_loss, max_2D_radius=_radius_buffer(variables),
                                  seen=seen)
    loss = out[0]
    weighted_losses = {'depth': out[2], 'im': out[1], 'loss': loss}
    # max_2D_radius / seen (splatam.py:296-298) were updated in place by the preprocess kernel
    variables['seen'] = seen
    variables['means2D'] = _NO_MEANS2D
    return loss, variables, weighted_losses


def _get_loss_mapping_fused(params, curr_data, variables, iter_time_idx, loss_weights, camera_grad):
    """get_loss(mapping=True) (splatam.py:183-196, 213-249) as one fused autograd node, SSIM included."""
    from .frame import fused_mapping_loss
    p = dict(params)
    if not camera_grad:
        p['cam_unnorm_rots'], p['cam_trans'] = p['cam_unnorm_rots'].detach(), p['cam_trans'].detach()
    seen = _seen_buffer(variables, params)
    out, ws = fused_mapping_loss(p, curr_data['cam'], int(iter_time_idx), curr_data['w2c'], curr_data['im'],
                                 curr_data['depth'], loss_weights['im'], loss_weights['depth'],
                                 max_2D_radius=_radius_buffer(variables), seen=seen)
    loss = out[0]
    weighted_losses = {'depth': out[2], 'im': out[1], 'loss': loss}
    variables['seen'] = seen
    variables['means2D'] = _NO_MEANS2D
    return loss, variables, weighted_losses


def get_pointcloud(color, depth, intrinsics, w2c, transform_pts=True, mask=None, compute_mean_sq_dist=False,
                   mean_sq_dist_method="projective"):
    """Back-projection of an RGB-D frame into a coloured point cloud [n,6] (+ the squared 1-pixel footprint of each
    point, the initial Gaussian scale): the reference's get_pointcloud (src/mapper/splatam/splatam.py:25-75), used for
    map initialisation.  Densification does the same arithmetic inside as_densify."""
    if mean_sq_dist_method != "projective":
        raise ValueError(f"Unknown mean_sq_dist_method {mean_sq_dist_method}")
    _, H, W = color.shape
    fx, fy, cx, cy = intrinsics[0][0], intrinsics[1][1], intrinsics[0][2], intrinsics[1][2]
    v, u = torch.meshgrid(torch.arange(H, device=color.device, dtype=torch.float32),
                          torch.arange(W, device=color.device, dtype=torch.float32), indexing='ij')
    z = depth[0].reshape(-1)
    pts = torch.stack((((u - cx) / fx).reshape(-1) * z, ((v - cy) / fy).reshape(-1) * z, z), dim=-1)
    if transform_pts:
        c2w = torch.inverse(w2c)
        pts = pts @ c2w[:3, :3].T + c2w[:3, 3]
    cloud = torch.cat((pts, color.permute(1, 2, 0).reshape(-1, 3)), dim=-1)
    msd = (z / ((fx + fy) / 2)) ** 2
    if mask is not None:
        cloud, msd = cloud[mask], msd[mask]
    return (cloud, msd) if compute_mean_sq_dist else cloud


def initialize_new_params(new_pt_cld, mean3_sq_dist, gaussian_distribution):
    """Parameters of Gaussians created from a point cloud (splatam.py:304-329): identity rotation, logit opacity 0,
    log scale = log of the 1-pixel footprint, one scale per Gaussian ("isotropic") or three equal ones."""
    reps = {"isotropic": 1, "anisotropic": 3}.get(gaussian_distribution)
    if reps is None:
        raise ValueError(f"Unknown gaussian_distribution {gaussian_distribution}")
    n, dev = new_pt_cld.shape[0], new_pt_cld.device
    identity = torch.tensor([1.0, 0.0, 0.0, 0.0], device=dev).expand(n, 4)
    fresh = {'means3D': new_pt_cld[:, :3], 'rgb_colors': new_pt_cld[:, 3:6], 'unnorm_rotations': identity,
             'logit_opacities': torch.zeros((n, 1), device=dev),
             'log_scales': torch.log(torch.sqrt(mean3_sq_dist)).reshape(n, 1).expand(n, reps)}
    return {k: torch.nn.Parameter(v.float().contiguous().requires_grad_(True)) for k, v in fresh.items()}


class _MapStore:
    """Capacity behind a params dict: the five Gaussian tensors (and variables['timestep']) as views of buffers that
    hold more rows than the map has, so that densification appends in place.  Found again through the identity of the
    dict; if the caller swapped a tensor for one of its own, the store is rebuilt from what the dict holds now."""
    KEYS = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")

    def __init__(self, params, variables, room):
        self.P = int(params['means3D'].shape[0])
        self.capacity = 0
        self.buf, self.timestep = {}, None
        self._adopt(params, variables, self.P + int(room))

    def _adopt(self, params, variables, capacity):
        dev = params['means3D'].device
        capacity = max(int(capacity), self.P)
        new = {k: torch.empty((capacity,) + tuple(params[k].shape[1:]), dtype=torch.float32, device=dev) for k in self.KEYS}
        for k in self.KEYS:
            new[k][:self.P].copy_(params[k].detach()[:self.P])
        ts = torch.zeros(capacity, dtype=torch.float32, device=dev)
        t_old = variables.get('timestep')
        if t_old is not None:
            ts[:min(self.P, t_old.shape[0])].copy_(t_old[:self.P])
        self.buf, self.timestep, self.capacity = new, ts, capacity

    def matches(self, params, variables):
        if int(params['means3D'].shape[0]) != self.P:
            return False
        if any(params[k].data_ptr() != self.buf[k].data_ptr() for k in self.KEYS):
            return False
        t = variables.get('timestep')
        return t is not None and t.data_ptr() == self.timestep.data_ptr()

    def publish(self, params, variables, P):
        """Point the dict at rows [0, P) of the buffers (new leaf Parameters, like the reference's re-wrapping)."""
        self.P = int(P)
        for k in self.KEYS:
            params[k] = torch.nn.Parameter(self.buf[k][:self.P], requires_grad=True)
        variables['timestep'] = self.timestep[:self.P]


_MAP_STORES = {}


def _map_store(params, variables, room):
    st = _MAP_STORES.get(id(params))
    if st is None or not st.matches(params, variables):
        st = _MapStore(params, variables, room)
        st.publish(params, variables, st.P)
        _MAP_STORES[id(params)] = st
        if len(_MAP_STORES) > 8:   # dicts that went away
            _MAP_STORES.pop(next(iter(_MAP_STORES)))
    return st


def add_new_gaussians(params, variables, curr_data, sil_thres, time_idx, mean_sq_dist_method, gaussian_distribution):
    """Silhouette-driven densification (src/mapper/splatam/splatam.py:332-379): render depth and silhouette of the
    current map at the frame's pose, take the pixels the map does not explain, back-project them as new Gaussians.
    One fused forward from the raw parameters, then as_densify: median, mask, ordered compaction and the new rows are
    produced on the device and written behind the existing rows of tensors that were allocated with head-room (no
    torch.cat of the whole map; the map grows by at most one frame of pixels per call).  One host read: the count."""
    import ctypes
    from . import capi
    from .frame import fused_render
    if mean_sq_dist_method != "projective":
        raise ValueError(f"Unknown mean_sq_dist_method {mean_sq_dist_method}")
    if gaussian_distribution not in ("isotropic", "anisotropic"):
        raise ValueError(f"Unknown gaussian_distribution {gaussian_distribution}")
    dev = params['means3D'].device
    H, W = curr_data['im'].shape[1], curr_data['im'].shape[2]
    store = _map_store(params, variables, room=W * H)
    with torch.no_grad():
        p = {k: v.detach() for k, v in params.items()}
        c6, _, _, _, _ = fused_render(p, curr_data['cam'], int(time_idx), curr_data['w2c'])
        # the frame's pose (splatam.py:357-362) and its inverse, on the device
        w2c = torch.eye(4, device=dev)
        w2c[:3, :3] = build_rotation(F.normalize(params['cam_unnorm_rots'][..., time_idx].detach()))[0]
        w2c[:3, 3] = params['cam_trans'][..., time_idx].detach().reshape(3)
        c2w = torch.inverse(w2c).contiguous()
        k = curr_data['intrinsics']
        k = k.detach().cpu().numpy() if torch.is_tensor(k) else np.asarray(k)
        L = capi.lib()
        scratch = torch.empty(L.as_densify_scratch_words(W, H), dtype=torch.int32, device=dev)
        result = torch.zeros(8, dtype=torch.int32, device=dev)
        gt_rgb, gt_depth = curr_data['im'].contiguous().float(), curr_data['depth'].contiguous().float()
        for _ in range(2):
            a = capi.AsDensifyArgs()
            a.width, a.height = W, H
            a.color6, a.gt_rgb, a.gt_depth = c6.data_ptr(), gt_rgb.data_ptr(), gt_depth.data_ptr()
            a.sil_thres = float(sil_thres)
            a.fx, a.fy, a.cx, a.cy = float(k[0][0]), float(k[1][1]), float(k[0][2]), float(k[1][2])
            a.c2w = c2w.data_ptr()
            a.isotropic = 1 if gaussian_distribution == "isotropic" else 0
            a.time_idx = float(time_idx)
            a.P, a.capacity = store.P, store.capacity
            a.means3D, a.rgb_colors = store.buf['means3D'].data_ptr(), store.buf['rgb_colors'].data_ptr()
            a.unnorm_rotations = store.buf['unnorm_rotations'].data_ptr()
            a.logit_opacities, a.log_scales = store.buf['logit_opacities'].data_ptr(), store.buf['log_scales'].data_ptr()
            a.timestep = store.timestep.data_ptr()
            a.scratch, a.result = scratch.data_ptr(), result.data_ptr()
            with torch.cuda.device(dev):
                capi.check(L.as_densify(ctypes.byref(a), torch.cuda.current_stream(dev).cuda_stream))
            n_new, overflow, _, n_mask = result[:4].tolist()   # the one synchronisation: what follows needs the new P
            if not overflow:
                break
            store._adopt(params, variables, int(1.5 * (store.P + n_new)) + W * H)   # grow, then write again
            store.publish(params, variables, store.P)
    if n_mask > 0:                                   # splatam.py:355: otherwise nothing is touched
        num_pts = store.P + n_new
        store.publish(params, variables, num_pts)
        variables['means2D_gradient_accum'] = torch.zeros(num_pts, device=dev)
        variables['denom'] = torch.zeros(num_pts, device=dev)
        variables['max_2D_radius'] = torch.zeros(num_pts, device=dev)
    return params, variables


def _radius_buffer(variables):
    r = variables.get('max_2D_radius')
    return r if (r is not None and r.dtype == torch.float32 and r.is_contiguous()) else None


def _seen_buffer(variables, params):
    s = variables.get('seen')
    n = params['means3D'].shape[0]
    if s is None or s.dtype != torch.bool or s.shape[0] != n or s.device != params['means3D'].device:
        s = torch.zeros(n, dtype=torch.bool, device=params['means3D'].device)
    return s


def initialize_optimizer(params, lrs_dict, tracking):
    param_groups = [{'params': [v], 'name': k, 'lr': lrs_dict[k]} for k, v in params.items()]
    if tracking:
        return torch.optim.Adam(param_groups)
    return torch.optim.Adam(param_groups, lr=0.0, eps=1e-15)


def render_view(params, w2c, k, width, height, near=0.01, far=100, scale_modifier=1.0, ws=None):
    """get_rendervars + render (splatam.py:436-468, 413-434) for one candidate view, straight from the raw
    parameters: white-background RGB, median depth, opacity and silhouette, one fused forward, no gradients.
    This is the per-view body of the GetOpacity sweeps (src/mapper/splatam/__init__.py:716-737)."""
    from .frame import fused_render, identity_pose
    with torch.no_grad():
        dev = params['means3D'].device
        cam, sil_w2c = _camera_one_upload(width, height, k, w2c, near, far, scale_modifier, dev)
        rot, tr = identity_pose(dev)
        p = dict(params, cam_unnorm_rots=rot, cam_trans=tr)
        c6, _, depth, opacity, ws = fused_render(p, cam, 0, sil_w2c, bg_rgb=_white_bg(dev), ws=ws)
        return c6[:3], depth, opacity, c6[4:5]


def _camera_one_upload(w, h, k, w2c, near, far, scale_modifier, dev):
    """setup_camera's matrices (recon_helpers.py:6-14) computed on the host and sent to the device in ONE
    copy (the reference builds them with ~8 tiny device ops + a matrix inverse per candidate view)."""
    fx, fy, cx, cy = k[0][0], k[1][1], k[0][2], k[1][2]
    w2c_t = torch.as_tensor(np.asarray(w2c)).float()
    cam_center = torch.inverse(w2c_t)[:3, 3]
    view = w2c_t.unsqueeze(0).transpose(1, 2)
    opengl_proj = torch.tensor([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                                [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                                [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                                [0.0, 0.0, 1.0, 0.0]]).float().unsqueeze(0).transpose(1, 2)
    full_proj = view.bmm(opengl_proj)
    packed = torch.cat([view.reshape(-1), full_proj.reshape(-1), w2c_t.reshape(-1), cam_center, torch.zeros(3)])
    d = packed.to(dev, non_blocking=True)
    cam = Camera(image_height=h, image_width=w, tanfovx=w / (2 * fx), tanfovy=h / (2 * fy), bg=d[51:54],
                 scale_modifier=scale_modifier, viewmatrix=d[0:16].view(1, 4, 4), projmatrix=d[16:32].view(1, 4, 4),
                 sh_degree=0, campos=d[48:51], prefiltered=False, debug=False)
    return cam, d[32:48].view(4, 4)


_WHITE = {}


def _white_bg(dev):
    key = torch.device(dev)
    if key not in _WHITE:
        _WHITE[key] = torch.ones(3, device=dev)
    return _WHITE[key]
