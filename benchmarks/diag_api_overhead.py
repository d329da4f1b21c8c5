"""Diagnostic: host-side cost of the rasterizer API path (cfg1 call pattern) under cProfile."""
import cProfile, pstats, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from activesplat_b200 import scenes, splatam
from activesplat_b200.rasterizer import GaussianRasterizer
dev = torch.device("cuda:0")
W, H, P = 640, 480, 100_000
params, cam = scenes.s_frustum(P, W, H, seed=0)
a = {k: v.to(dev).requires_grad_(True) for k, v in scenes.activated(params).items()}
view, proj, campos = cam.matrices()
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
feats = torch.cat([a["means3D"][:, 2:3], torch.ones_like(a["means3D"][:, :1]), a["means3D"][:, 2:3] ** 2], 1).detach().requires_grad_(True)
rs6 = splatam.Camera(H, W, cam.tanfovx, cam.tanfovy, torch.zeros(6, device=dev), 1.0, t(view).unsqueeze(0),
                     t(proj).unsqueeze(0), 0, t(campos), False, False)
dL = torch.rand(6, H, W, device=dev)
m2d = torch.zeros_like(a["means3D"])
leaves = [a["means3D"], a["colors_precomp"], feats, a["opacities"], a["scales"], a["rotations"]]
def ours(i):
    c6, _, _, _ = GaussianRasterizer(rs6)(means3D=a["means3D"], means2D=m2d, opacities=a["opacities"],
                                          colors_precomp=torch.cat([a["colors_precomp"], feats], 1),
                                          scales=a["scales"], rotations=a["rotations"])
    torch.autograd.grad((c6 * dL).sum(), leaves)
for i in range(10): ours(i)
torch.cuda.synchronize()
t0 = time.time()
for i in range(100): ours(i)
torch.cuda.synchronize()
print("ms per call (wall):", (time.time() - t0) * 10)
pr = cProfile.Profile(); pr.enable()
for i in range(100): ours(i)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
