"""Diagnostic: gradient agreement and parameter drift, ours vs the reference pipeline, at cfg3 scale."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from activesplat_b200 import scenes, splatam
from activesplat_b200.frame import PARAM_ORDER, FusedAdam, workspace_for
from benchmarks.configs import load_ref_renderer, gt_frames, LW, MAP_LRS
from oracle import splatam_oracle as so
dev = torch.device("cuda:0")
W, H, P = 640, 480, int(os.environ.get("DIAG_P", "1000000"))
params_cpu, cams = scenes.s_room(P, 4, seed=3, W=W, H=H)
rots, trans = scenes.camera_track(cams)
gt = gt_frames(params_cpu, cams[:1], dev, seed=13)[0]
cam = splatam.setup_camera(W, H, cams[0].k(), np.eye(4), device=dev)
cur = {"cam": cam, "im": gt[:3], "depth": gt[3:4], "w2c": torch.eye(4, device=dev)}
Ref = load_ref_renderer()
def fresh():
    p = {k: v.clone().to(dev).requires_grad_(True) for k, v in params_cpu.items()}
    p["cam_unnorm_rots"] = rots.clone().to(dev).requires_grad_(True)
    p["cam_trans"] = trans.clone().to(dev).requires_grad_(True)
    return p
p1, p2 = fresh(), fresh()
opt1 = FusedAdam(p1, MAP_LRS)
opt2 = torch.optim.Adam([{"params": [v], "name": k, "lr": MAP_LRS[k]} for k, v in p2.items()], lr=0.0, eps=1e-15)
variables = {"max_2D_radius": torch.zeros(P, device=dev)}
ws = workspace_for(p1, cam)
for it in range(21):
    l1, _, wl1 = splatam.get_loss(p1, cur, variables, 0, LW, False, 0.98, True, False, mapping=True)
    for k in PARAM_ORDER: p1[k].grad = None
    l1.backward()
    l2, wl2, _ = so.get_loss(p2, cur, 0, LW, False, 0.98, True, False, mapping=True, Renderer=Ref)
    opt2.zero_grad(set_to_none=True)
    l2.backward()
    if it in (0, 1, 2, 5, 10, 20):
        gd = {k: float((p1[k].grad - p2[k].grad).norm() / (p2[k].grad.norm() + 1e-30)) for k in PARAM_ORDER}
        pd = {k: float((p1[k].detach() - p2[k].detach()).norm() / ((p2[k].detach() - params_cpu[k].to(dev)).norm() + 1e-30)) for k in PARAM_ORDER}
        print(it, "loss", float(l1), float(l2), "grad rel", {k: round(v, 6) for k, v in gd.items()})
        print("   param drift / update", {k: round(v, 5) for k, v in pd.items()})
    opt1.step(ws.grad_params)
    opt2.step()
