"""Diagnostic: per-view cost of the GetOpacity render (cfg4 camera) with per-stage timings."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from activesplat_b200 import scenes, splatam, capi
dev = torch.device('cuda:0')
P = int(os.environ.get("DIAG_P", "1000000"))
NV = int(os.environ.get("DIAG_VIEWS", "64"))
params_cpu, _ = scenes.s_room(P, 1, seed=4)
params = {k: v.to(dev) for k, v in params_cpu.items()}
views = scenes.candidate_views(NV, seed=4)
L = capi.lib()
for c in views[:4]:
    splatam.render_view(params, c.w2c, c.k(), c.width, c.height)
torch.cuda.synchronize()
per_view = []
tot = {}
from activesplat_b200.frame import _WORKSPACES
for c in views:
    L.as_profile_enable(1); capi.profile_read()
    t0 = time.time()
    splatam.render_view(params, c.w2c, c.k(), c.width, c.height)
    torch.cuda.synchronize()
    wall = (time.time() - t0) * 1e3
    st = capi.profile_read()
    for k, v in st.items():
        tot[k] = tot.get(k, 0.0) + v[0]
    tot["wall"] = tot.get("wall", 0.0) + wall
    ws = list(_WORKSPACES.values())[0]
    R, _ = ws.last_status()
    per_view.append((round(st['blend_fwd'][0], 3), R, int(ws.radii.max()), int((ws.radii > 0).sum()), int((ws.radii > 200).sum())))
L.as_profile_enable(0)
print("mean stage ms per view:", {k: round(v / len(views), 4) for k, v in tot.items()})
print("blend_ms, R, max radius, visible, radius>200")
for p in per_view[:24]:
    print(p)
