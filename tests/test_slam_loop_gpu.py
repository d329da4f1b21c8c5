"""The SplaTAM loop of ActiveSplat's mapper as a caller drives it (src/mapper/splatam/__init__.py:380-520): for every
frame, pose-only tracking iterations, silhouette-driven densification, mapping iterations with a fresh optimizer --
on the fused path, with the map growing from frame to frame.  Parity of each piece against the reference is pinned
elsewhere (test_splatam_golden.py); this test is about the pieces working IN SEQUENCE: workspaces that follow a
growing map, optimizer state created per round, gradients that stay in the workspace, no stale buffers."""
import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu

TRACK_LRS = dict(cam_unnorm_rots=1e-3, cam_trans=4e-3)     # config/splatam/online_habitat_sim.py:37-45


def _pose_error(rots, trans, t, cam):
    """(degrees, millimetres) between the estimated pose of frame t and the camera the frame was rendered from."""
    from activesplat_b200 import splatam
    with torch.no_grad():
        R = splatam.build_rotation(torch.nn.functional.normalize(rots[..., t]))[0].cpu().numpy()
        tr = trans[0, :, t].cpu().numpy()
    Rg, tg = cam.w2c[:3, :3], cam.w2c[:3, 3]
    c = (np.trace(R @ Rg.T) - 1.0) / 2.0
    return float(np.degrees(np.arccos(np.clip(c, -1, 1)))), float(1e3 * np.linalg.norm(tr - tg))


@pytest.mark.parametrize("iso", [True, False])
def test_track_densify_map_over_frames(iso):
    from activesplat_b200 import frame, scenes, splatam
    dev = torch.device("cuda:0")
    W, H, P_full, K = 160, 128, 30_000, 4
    full_cpu, c0 = scenes.s_room(P_full, 1, seed=21, W=W, H=H, scale_mult=2.0)
    c0 = c0[0]
    # a slow turn on the circle, 1 degree per frame: what a tracker starting from the previous pose can follow
    cams = [scenes.CameraSpec(W, H, c0.fx, c0.fy, c0.cx, c0.cy, np.linalg.inv(scenes.room_pose(k, 360))) for k in range(K)]
    if iso:
        full_cpu = dict(full_cpu, log_scales=full_cpu["log_scales"][:, :1].contiguous())
    full = {k: v.to(dev) for k, v in full_cpu.items()}
    # ground-truth frames from the full map; the SLAM map starts from a third of it, slightly off
    frames = []
    with torch.no_grad():
        for c in cams:
            im, depth, opacity, _ = splatam.render_view(full, c.w2c, c.k(), W, H)
            frames.append((im.clone(), torch.where(opacity > 0.5, depth, torch.zeros_like(depth)).clone()))
    keep = torch.arange(P_full) % 3 == 0
    start = scenes.perturbed({k: v[keep] for k, v in full_cpu.items()}, seed=3, sigma_means=0.002, sigma_rgb=0.02)
    params = {k: torch.nn.Parameter(v.to(dev).contiguous()) for k, v in start.items()}
    rots_gt, trans_gt = scenes.camera_track(cams)
    rots, trans = rots_gt.clone(), trans_gt.clone()
    rots[..., 1:], trans[..., 1:] = 0, 0                       # later poses are found by tracking
    params["cam_unnorm_rots"] = torch.nn.Parameter(rots.to(dev))
    params["cam_trans"] = torch.nn.Parameter(trans.to(dev))
    P0 = int(params["means3D"].shape[0])
    variables = {"max_2D_radius": torch.zeros(P0, device=dev), "means2D_gradient_accum": torch.zeros(P0, device=dev),
                 "denom": torch.zeros(P0, device=dev), "timestep": torch.zeros(P0, device=dev)}
    intr = torch.tensor(cams[0].k(), dtype=torch.float32, device=dev)
    first_w2c = torch.eye(4, device=dev)          # the poses of camera_track are absolute, so the fixed camera is the identity
    cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    sizes, map_losses = [P0], []
    for t in range(K):
        im, depth = frames[t]
        curr = {"cam": cam, "im": im, "depth": depth, "id": t, "intrinsics": intr, "w2c": first_w2c}
        if t > 0:
            # ---- tracking: start from the previous pose (constant-position prior), pose-only iterations ----
            with torch.no_grad():
                params["cam_unnorm_rots"][..., t] = params["cam_unnorm_rots"][..., t - 1]
                params["cam_trans"][..., t] = params["cam_trans"][..., t - 1]
            before = _pose_error(params["cam_unnorm_rots"], params["cam_trans"], t, cams[t])
            opt = frame.FusedPoseAdam(params, TRACK_LRS)
            track_losses = []
            for it in range(40):
                for p in params.values():
                    p.grad = None
                loss, variables, _ = splatam.get_loss(params, curr, variables, t, helpers.LOSS_WEIGHTS, True, 0.99, True,
                                                      False, tracking=True)
                loss.backward()
                opt.step()
                track_losses.append(float(loss.detach()))
            after = _pose_error(params["cam_unnorm_rots"], params["cam_trans"], t, cams[t])
            assert all(np.isfinite(track_losses))
            # on this trajectory (a turn on a circle, looking outward) the world->camera translation is constant: the start
            # is off by the turn only.  40 iterations take most of the turn back (part of it may be traded for a few
            # centimetres of translation, which the loss cannot tell apart at this distance from the walls)
            assert track_losses[-1] < track_losses[0], (t, track_losses[0], track_losses[-1])
            assert after[0] < 0.6 * before[0] and after[1] < 50.0, (t, before, after)
        # ---- densification: the map does not explain this frame yet ----
        with torch.no_grad():
            params, variables = splatam.add_new_gaussians(params, variables, curr, 0.5, t, "projective",
                                                          "isotropic" if iso else "anisotropic")
        P = int(params["means3D"].shape[0])
        assert P >= sizes[-1] and variables["timestep"].shape[0] == P and variables["max_2D_radius"].shape[0] == P
        sizes.append(P)
        # ---- mapping: a fresh optimizer per round (src/mapper/splatam/__init__.py:440), gradients stay in the workspace ----
        opt = frame.FusedAdam(params, helpers.MAPPING_LRS)
        ws = frame.workspace_for(params, cam, "train")
        losses = []
        for it in range(30):
            for p in params.values():
                p.grad = None
            loss, variables, _ = splatam.get_loss(params, curr, variables, t, helpers.LOSS_WEIGHTS, False, 0.5, True,
                                                  False, mapping=True)
            with frame.gradients_stay_in_workspace():
                loss.backward()
            assert ws.P == P
            opt.step(ws.grad_params)
            losses.append(float(loss.detach()))
        assert all(np.isfinite(losses)), losses
        assert losses[-1] < losses[0], (t, losses[0], losses[-1])
        map_losses.append((losses[0], losses[-1]))
        for k in frame.PARAM_ORDER:
            assert torch.isfinite(params[k]).all()
    assert sizes[1] > sizes[0], sizes      # the first frame sees holes in a map that holds a third of the scene
    # the frames seen so far render close to their ground truth from the final map
    with torch.no_grad():
        p = {k: v.detach() for k, v in params.items()}
        c6, _, _, _, _ = frame.fused_render(p, cam, K - 1, first_w2c)
        mse = float(((c6[:3] - frames[K - 1][0]) ** 2).mean())
    assert 10 * np.log10(1.0 / max(mse, 1e-12)) > 10.0, mse
