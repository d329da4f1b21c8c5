"""CPU restatement of the SplaTAM functions around the rasterizer (TEST INFRASTRUCTURE ONLY).

Pure torch on CPU tensors with the C-oracle rasterizer underneath, following the reference line by
line -- including its TWO raster passes per frame (src/mapper/splatam/splatam.py:208,212), which the
product collapses into one 6-channel pass:
  build_rotation          src/mapper/splatam/utils/slam_external.py:25-42
  quat_mult               src/mapper/splatam/utils/slam_helpers.py:21-28
  transform_to_frame      slam_helpers.py:252-304
  get_depth_and_silhouette slam_helpers.py:196-213
  rendervars              slam_helpers.py:124-139, 234-249
  calc_ssim               slam_external.py:54-97
  get_loss                splatam.py:172-301
  adam_step               torch.optim.Adam as configured at splatam.py:118-124
"""
from math import exp

import torch
import torch.nn.functional as F

from .raster_oracle import GaussianRasterizer as Renderer  # default: the CPU oracle rasterizer

# Every function creates its temporaries on the device of its inputs, and get_loss / render take the
# rasterizer class as an argument, so the same restatement also drives the UNMODIFIED reference CUDA
# rasterizer (oracle/_ref) on a GPU: that combination is the "reference pipeline" timed by the benchmarks.


def build_rotation(q):
    norm = torch.sqrt(q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1] + q[:, 2] * q[:, 2] + q[:, 3] * q[:, 3])
    q = q / norm[:, None]
    rot = torch.zeros((q.size(0), 3, 3), device=q.device)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rot[:, 0, 0] = 1 - 2 * (y * y + z * z)
    rot[:, 0, 1] = 2 * (x * y - r * z)
    rot[:, 0, 2] = 2 * (x * z + r * y)
    rot[:, 1, 0] = 2 * (x * y + r * z)
    rot[:, 1, 1] = 1 - 2 * (x * x + z * z)
    rot[:, 1, 2] = 2 * (y * z - r * x)
    rot[:, 2, 0] = 2 * (x * z - r * y)
    rot[:, 2, 1] = 2 * (y * z + r * x)
    rot[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return rot


def quat_mult(q1, q2):
    w1, x1, y1, z1 = q1.T
    w2, x2, y2, z2 = q2.T
    w = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2
    x = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2
    y = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2
    z = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2
    return torch.stack([w, x, y, z]).T


def transform_to_frame(params, time_idx, gaussians_grad, camera_grad):
    if camera_grad:
        cam_rot = F.normalize(params['cam_unnorm_rots'][..., time_idx])
        cam_tran = params['cam_trans'][..., time_idx]
    else:
        cam_rot = F.normalize(params['cam_unnorm_rots'][..., time_idx].detach())
        cam_tran = params['cam_trans'][..., time_idx].detach()
    rel_w2c = torch.eye(4, device=params['means3D'].device)
    rel_w2c[:3, :3] = build_rotation(cam_rot)
    rel_w2c[:3, 3] = cam_tran
    transform_rots = params['log_scales'].shape[1] != 1
    if gaussians_grad:
        pts, unnorm_rots = params['means3D'], params['unnorm_rotations']
    else:
        pts, unnorm_rots = params['means3D'].detach(), params['unnorm_rotations'].detach()
    pts4 = torch.cat((pts, torch.ones(pts.shape[0], 1, device=pts.device)), dim=1)
    out = {'means3D': (rel_w2c @ pts4.T).T[:, :3]}
    out['unnorm_rotations'] = quat_mult(cam_rot, F.normalize(unnorm_rots)) if transform_rots else unnorm_rots
    return out


def get_depth_and_silhouette(pts_3D, w2c):
    pts4 = torch.cat((pts_3D, torch.ones_like(pts_3D[:, :1])), dim=-1)
    pts_in_cam = (w2c @ pts4.transpose(0, 1)).transpose(0, 1)
    depth_z = pts_in_cam[:, 2].unsqueeze(-1)
    ds = torch.zeros((pts_3D.shape[0], 3), device=pts_3D.device)
    ds[:, 0] = depth_z.squeeze(-1)
    ds[:, 1] = 1.0
    ds[:, 2] = torch.square(depth_z).squeeze(-1)
    return ds


def _log_scales3(params):
    ls = params['log_scales']
    return torch.tile(ls, (1, 3)) if ls.shape[1] == 1 else ls


def transformed_params2rendervar(params, tg):
    return {'means3D': tg['means3D'], 'colors_precomp': params['rgb_colors'],
            'rotations': F.normalize(tg['unnorm_rotations']), 'opacities': torch.sigmoid(params['logit_opacities']),
            'scales': torch.exp(_log_scales3(params)),
            'means2D': torch.zeros_like(params['means3D'], requires_grad=True) + 0}


def transformed_params2depthplussilhouette(params, w2c, tg):
    rv = transformed_params2rendervar(params, tg)
    rv['colors_precomp'] = get_depth_and_silhouette(tg['means3D'], w2c)
    return rv


def gaussian(window_size, sigma):
    g = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    return g / g.sum()


def calc_ssim(img1, img2, window_size=11):
    channel = img1.size(-3)
    w1 = gaussian(window_size, 1.5).unsqueeze(1)
    window = w1.mm(w1.t()).float().unsqueeze(0).unsqueeze(0).expand(channel, 1, window_size, window_size).contiguous()
    window = window.to(img1.device).type_as(img1)
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + c1) * (2 * sigma12 + c2)) / ((mu1_sq + mu2_sq + c1) * (sigma1_sq + sigma2_sq + c2))
    return ssim_map.mean()


def get_rendervars(params, w2c):
    """splatam.py:436-468 (no timestep filtering)."""
    base = {'means3D': params['means3D'], 'rotations': F.normalize(params['unnorm_rotations']),
            'opacities': torch.sigmoid(params['logit_opacities']), 'scales': torch.exp(_log_scales3(params))}
    rv = dict(base, colors_precomp=params['rgb_colors'], means2D=torch.zeros_like(params['means3D']))
    drv = dict(base, colors_precomp=get_depth_and_silhouette(params['means3D'], w2c),
               means2D=torch.zeros_like(params['means3D']))
    return rv, drv


def render(cam, white_bg_cam, timestep_data, timestep_depth_data, Renderer=Renderer):
    """splatam.py:413-434: white-background RGB pass, then the depth/silhouette pass."""
    with torch.no_grad():
        im, _, depth, opacity = Renderer(white_bg_cam)(**timestep_data)
        depth_sil, _, _, _ = Renderer(cam)(**timestep_depth_data)
        return im, depth, opacity, depth_sil[1, :, :].unsqueeze(0)


def get_loss(params, curr_data, iter_time_idx, loss_weights, use_sil_for_loss, sil_thres, use_l1,
             ignore_outlier_depth_loss, tracking=False, mapping=False, do_ba=False, Renderer=Renderer):
    losses = {}
    if tracking:
        tg = transform_to_frame(params, iter_time_idx, gaussians_grad=False, camera_grad=True)
    elif mapping and do_ba:
        tg = transform_to_frame(params, iter_time_idx, gaussians_grad=True, camera_grad=True)
    else:
        tg = transform_to_frame(params, iter_time_idx, gaussians_grad=True, camera_grad=False)
    rendervar = transformed_params2rendervar(params, tg)
    depth_sil_rendervar = transformed_params2depthplussilhouette(params, curr_data['w2c'], tg)
    im, radius, _, _ = Renderer(curr_data['cam'])(**rendervar)                       # splatam.py:208
    depth_sil, _, _, _ = Renderer(curr_data['cam'])(**depth_sil_rendervar)           # splatam.py:212
    depth = depth_sil[0, :, :].unsqueeze(0)
    silhouette = depth_sil[1, :, :]
    presence_sil_mask = (silhouette > sil_thres)
    depth_sq = depth_sil[2, :, :].unsqueeze(0)
    invisibility = (depth_sq - depth ** 2).detach()
    nan_mask = (~torch.isnan(depth)) & (~torch.isnan(invisibility))
    if ignore_outlier_depth_loss:
        depth_error = torch.abs(curr_data['depth'] - depth) * (curr_data['depth'] > 0)
        mask = (depth_error < 10 * depth_error.median()) & (curr_data['depth'] > 0)
    else:
        mask = (curr_data['depth'] > 0)
    mask = mask & nan_mask
    if tracking and use_sil_for_loss:
        mask = mask & presence_sil_mask
    if use_l1:
        mask = mask.detach()
        if tracking:
            losses['depth'] = torch.abs(curr_data['depth'] - depth)[mask].sum()
        else:
            losses['depth'] = torch.abs(curr_data['depth'] - depth)[mask].mean()
    if tracking and (use_sil_for_loss or ignore_outlier_depth_loss):
        color_mask = torch.tile(mask, (3, 1, 1)).detach()
        losses['im'] = torch.abs(curr_data['im'] - im)[color_mask].sum()
    elif tracking:
        losses['im'] = torch.abs(curr_data['im'] - im).sum()
    else:
        losses['im'] = 0.8 * torch.abs(im - curr_data['im']).mean() + 0.2 * (1.0 - calc_ssim(im, curr_data['im']))
    weighted = {k: v * loss_weights[k] for k, v in losses.items()}
    loss = sum(weighted.values())
    weighted['loss'] = loss
    return loss, weighted, dict(im=im, depth_sil=depth_sil, radius=radius)


def get_pointcloud(color, depth, intrinsics, w2c, mask=None):
    """splatam.py:25-75 (transform_pts=True, projective mean-squared distance)."""
    width, height = color.shape[2], color.shape[1]
    CX, CY, FX, FY = intrinsics[0][2], intrinsics[1][2], intrinsics[0][0], intrinsics[1][1]
    x_grid, y_grid = torch.meshgrid(torch.arange(width).float(), torch.arange(height).float(), indexing='xy')
    xx = ((x_grid - CX) / FX).reshape(-1)
    yy = ((y_grid - CY) / FY).reshape(-1)
    depth_z = depth[0].reshape(-1)
    pts_cam = torch.stack((xx * depth_z, yy * depth_z, depth_z), dim=-1)
    pts4 = torch.cat((pts_cam, torch.ones(height * width, 1)), dim=1)
    pts = (torch.inverse(w2c) @ pts4.T).T[:, :3]
    mean3_sq_dist = (depth_z / ((FX + FY) / 2)) ** 2
    point_cld = torch.cat((pts, torch.permute(color, (1, 2, 0)).reshape(-1, 3)), -1)
    if mask is not None:
        point_cld, mean3_sq_dist = point_cld[mask], mean3_sq_dist[mask]
    return point_cld, mean3_sq_dist


def add_new_gaussians(params, curr_data, sil_thres, time_idx, gaussian_distribution, Renderer=Renderer):
    """splatam.py:332-379.  Returns (non_presence_mask [H*W] bool, new parameter dict or None)."""
    tg = transform_to_frame(params, time_idx, gaussians_grad=False, camera_grad=False)
    drv = transformed_params2depthplussilhouette(params, curr_data['w2c'], tg)
    with torch.no_grad():
        depth_sil, _, _, _ = Renderer(curr_data['cam'])(**{k: v.detach() for k, v in drv.items()})
    silhouette = depth_sil[1, :, :]
    gt_depth = curr_data['depth'][0, :, :]
    render_depth = depth_sil[0, :, :]
    depth_error = torch.abs(gt_depth - render_depth) * (gt_depth > 0)
    non_presence_depth_mask = (render_depth > gt_depth) * (depth_error > 2 * depth_error.median())
    good_depth_mask = (silhouette > sil_thres) * (gt_depth < 5)
    non_presence_mask = ((silhouette < sil_thres) | (non_presence_depth_mask & good_depth_mask)).reshape(-1)
    if torch.sum(non_presence_mask) == 0:
        return non_presence_mask, None
    cam_rot = F.normalize(params['cam_unnorm_rots'][..., time_idx].detach())
    w2c = torch.eye(4)
    w2c[:3, :3] = build_rotation(cam_rot)
    w2c[:3, 3] = params['cam_trans'][..., time_idx].detach()
    non_presence_mask = non_presence_mask & (gt_depth > 0).reshape(-1)
    pt, msd = get_pointcloud(curr_data['im'], curr_data['depth'], curr_data['intrinsics'], w2c, mask=non_presence_mask)
    n = pt.shape[0]
    rots = torch.zeros(n, 4)
    rots[:, 0] = 1.0
    reps = 1 if gaussian_distribution == "isotropic" else 3
    new = {'means3D': pt[:, :3], 'rgb_colors': pt[:, 3:6], 'unnorm_rotations': rots,
           'logit_opacities': torch.zeros(n, 1), 'log_scales': torch.tile(torch.log(torch.sqrt(msd))[..., None], (1, reps))}
    return non_presence_mask, new
