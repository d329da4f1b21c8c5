"""Diagnostic: fused-path gradients vs the CPU oracle and vs the reference pipeline for tiny Gaussians."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from activesplat_b200 import scenes, splatam
from activesplat_b200.frame import PARAM_ORDER
from benchmarks.configs import load_ref_renderer, LW
from oracle import splatam_oracle as so
dev = torch.device("cuda:0")
for sm in (3.0, 1.0):
    W, H, P = 128, 96, 20000
    params, cams = scenes.s_room(P, 3, 7, W=W, H=H, scale_mult=sm)
    rots, trans = scenes.camera_track(cams)
    params["cam_unnorm_rots"], params["cam_trans"] = rots, trans
    g = torch.Generator().manual_seed(8)
    gt_rgb, gt_depth = torch.rand(3, H, W, generator=g), torch.rand(1, H, W, generator=g) * 6
    cam_c = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device="cpu")
    cam_g = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
    pc = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    lc, _, _ = so.get_loss(pc, {"cam": cam_c, "im": gt_rgb, "depth": gt_depth, "w2c": torch.eye(4)}, 1, LW, False, 0.98, True, False, mapping=True)
    lc.backward()
    cur = {"cam": cam_g, "im": gt_rgb.to(dev), "depth": gt_depth.to(dev), "w2c": torch.eye(4, device=dev)}
    pg = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    lg, _, _ = splatam.get_loss(pg, cur, {"max_2D_radius": torch.zeros(P, device=dev)}, 1, LW, False, 0.98, True, False, mapping=True)
    lg.backward()
    pr = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    lr_, _, _ = so.get_loss(pr, cur, 1, LW, False, 0.98, True, False, mapping=True, Renderer=load_ref_renderer())
    lr_.backward()
    pf = {k: v.clone().to(dev).requires_grad_(True) for k, v in params.items()}
    lf, _, _ = so.get_loss(pf, cur, 1, LW, False, 0.98, True, False, mapping=True, Renderer=load_ref_renderer(True))
    lf.backward()
    print("scale_mult", sm, "loss", float(lc), float(lg), float(lr_))
    for k in PARAM_ORDER:
        a, c = pg[k].grad.cpu(), pf[k].grad.cpu()
        print("  %-18s fused-vs-refpipe(contiguous rotations) %.2e" % (k, float((a - c).norm() / c.norm())))
    for k in PARAM_ORDER:
        a, b, c = pg[k].grad.cpu(), pc[k].grad, pr[k].grad.cpu()
        print("  %-18s fused-vs-oracle %.2e  refpipe-vs-oracle %.2e  fused-vs-refpipe %.2e  |oracle| %.3e" % (
            k, float((a - b).norm() / b.norm()), float((c - b).norm() / b.norm()), float((a - c).norm() / c.norm()), float(b.norm())))
