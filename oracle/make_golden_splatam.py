"""Golden vectors for the SplaTAM layer, produced by the reference's OWN Python on top of its OWN CUDA build
(TEST INFRASTRUCTURE ONLY).

Runs on a GPU box:   python oracle/make_golden_splatam.py [outdir]
It imports the unmodified reference modules byte-compiled by oracle/build_ref_py.py (src/mapper/splatam/
splatam.py, utils/slam_helpers.py, utils/slam_external.py, utils/recon_helpers.py and the rasterizer's
diff_gaussian_rasterization/__init__.py) with `_C` = oracle/_ref/ref_dgr_C.so, feeds them the seeded cases of
tests/helpers.SPLATAM_CASES, and stores what they return as tests/golden/ref_splatam_<case>.npz:

  loss cases     get_loss (splatam.py:172-301): loss + weighted components, gradients of every parameter that gets
                 one, the screen-space means2D gradient, seen / max_2D_radius; plus transform_to_frame's output and
                 the two renders (im, depth_sil, radius) of the same frame, and calc_ssim of the frame.
  adam case      N iterations of get_loss(mapping) + backward + the reference's optimizer (initialize_optimizer).
  densify cases  add_new_gaussians (splatam.py:332-379): the appended rows of the five tensors + timestep.
  render case    get_rendervars + render (splatam.py:413-468).

Anisotropic maps: SplaTAM hands the rasterizer column-major rotations (quat_mult ends in `.T`), which the stock
backward misreads.  For those cases the gradients are stored twice: "stock_*" as the stock reference
returns them and the plain names from the same reference code with quat_mult's result made contiguous (a wrapper
around the reference function; nothing in the reference is edited).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402
from oracle import build_ref_py  # noqa: E402

DEV = "cuda:0"
PARAM_KEYS = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales", "cam_unnorm_rots",
              "cam_trans")


def npy(t):
    return t.detach().contiguous().cpu().numpy()


def setup(ns, spec):
    params, cams, gt_rgb, gt_depth = helpers.frame_case(**spec["frame"])
    W, H = cams[0].width, cams[0].height
    p = {k: torch.nn.Parameter(v.clone().to(DEV).float().contiguous()) for k, v in params.items()}
    P = p["means3D"].shape[0]
    k = torch.tensor(cams[0].k()).float().to(DEV)
    cam = ns.recon_helpers.setup_camera(W, H, cams[0].k(), np.eye(4))  # the reference's own camera builder
    w2c = torch.tensor(spec.get("w2c", np.eye(4).tolist())).float().to(DEV)
    curr = {"cam": cam, "im": gt_rgb.to(DEV), "depth": gt_depth.to(DEV), "id": spec["t"], "intrinsics": k, "w2c": w2c}
    variables = {"max_2D_radius": torch.zeros(P, device=DEV), "means2D_gradient_accum": torch.zeros(P, device=DEV),
                 "denom": torch.zeros(P, device=DEV), "timestep": torch.zeros(P, device=DEV)}
    return p, cams, curr, variables


def loss_once(ns, spec):
    p, cams, curr, variables = setup(ns, spec)
    loss, variables, wl = ns.splatam.get_loss(
        p, curr, variables, spec["t"], helpers.LOSS_WEIGHTS, spec["use_sil"], spec["sil_thres"], True,
        spec.get("outlier", False), tracking=spec.get("tracking", False), mapping=spec.get("mapping", False),
        do_ba=spec.get("do_ba", False))
    loss.backward()
    out = {"loss": npy(loss), "loss_im": npy(wl["im"]), "loss_depth": npy(wl["depth"])}
    for k in PARAM_KEYS:
        if p[k].grad is not None:
            out["grad_" + k] = npy(p[k].grad)
    if variables["means2D"].grad is not None:
        out["grad_means2D"] = npy(variables["means2D"].grad)
    out["seen"] = npy(variables["seen"])
    out["max_2D_radius"] = npy(variables["max_2D_radius"])
    # the pieces get_loss is made of, on the same frame
    with torch.no_grad():
        gg = not spec.get("tracking", False)
        tg = ns.slam_helpers.transform_to_frame(p, spec["t"], gaussians_grad=gg, camera_grad=False)
        out["tf_means3D"] = npy(tg["means3D"])
        out["tf_unnorm_rotations"] = npy(tg["unnorm_rotations"])
        rv = ns.slam_helpers.transformed_params2rendervar(p, tg)
        drv = ns.slam_helpers.transformed_params2depthplussilhouette(p, curr["w2c"], tg)
        out["rv_rotations"] = npy(rv["rotations"])
        out["rv_scales"] = npy(rv["scales"])
        out["rv_opacities"] = npy(rv["opacities"])
        out["depth_sil_colors"] = npy(drv["colors_precomp"])
        im, radius, mdepth, opacity = ns.splatam.Renderer(raster_settings=curr["cam"])(**rv)
        ds, _, _, _ = ns.splatam.Renderer(raster_settings=curr["cam"])(**drv)
        out.update(im=npy(im), radius=npy(radius), median_depth=npy(mdepth), opacity=npy(opacity), depth_sil=npy(ds))
        out["ssim"] = npy(ns.slam_external.calc_ssim(im, curr["im"]))
        out["cam_viewmatrix"] = npy(curr["cam"].viewmatrix)
        out["cam_projmatrix"] = npy(curr["cam"].projmatrix)
        out["cam_tanfov"] = np.array([curr["cam"].tanfovx, curr["cam"].tanfovy], np.float64)
    return out


class contiguous_quat_mult:
    """Wraps (does not edit) the reference's quat_mult so that its result is row-major."""

    def __init__(self, ns):
        self.mod, self.orig = ns.slam_helpers, ns.slam_helpers.quat_mult

    def __enter__(self):
        orig = self.orig
        self.mod.quat_mult = lambda a, b: orig(a, b).contiguous()

    def __exit__(self, *a):
        self.mod.quat_mult = self.orig


def run_loss(ns, spec):
    stock = loss_once(ns, spec)
    # the reference against a second run of itself (its float atomics commit in arbitrary order): the noise floor
    # any comparison with it inherits, per tensor, as max |a - b| / max |a|
    again = loss_once(ns, spec)
    for k in list(stock):
        if k.startswith("grad_") or k == "loss":
            a, b = stock[k].astype(np.float64), again[k].astype(np.float64)
            stock["selfnoise_" + k] = np.float64(np.abs(a - b).max() / max(np.abs(a).max(), 1e-30))
    if spec["frame"].get("iso", False):
        return stock
    with contiguous_quat_mult(ns):
        out = loss_once(ns, spec)
    for k in ("grad_log_scales", "grad_unnorm_rotations", "grad_cam_unnorm_rots", "grad_cam_trans"):
        if k in stock:
            out["stock_" + k] = stock[k]
    for k in stock:
        if k.startswith("selfnoise_"):
            out[k] = stock[k]
    # everything that does not pass through the rotation/scale backward must not depend on the memory order beyond
    # rounding (torch normalises a column-major tensor with a differently ordered reduction: rotations differ by an ulp)
    assert np.array_equal(stock["radius"], out["radius"])
    for k in ("loss", "im", "depth_sil"):
        d = float(np.abs(stock[k].astype(np.float64) - out[k]).max() / max(np.abs(out[k]).max(), 1e-30))
        out["stock_vs_rowmajor_" + k] = np.float64(d)
        assert d < 1e-5, (k, d)
    return out


def run_adam(ns, spec):
    p, cams, curr, variables = setup(ns, spec)
    opt = ns.splatam.initialize_optimizer(p, helpers.MAPPING_LRS, tracking=False)
    out = {}
    for it in range(spec["steps"]):
        loss, variables, wl = ns.splatam.get_loss(p, curr, variables, spec["t"], helpers.LOSS_WEIGHTS, False, 0.98, True,
                                                  False, mapping=True)
        loss.backward()
        out[f"loss_{it}"] = npy(loss)
        with torch.no_grad():
            opt.step()
            opt.zero_grad(set_to_none=True)
    for k in PARAM_KEYS[:5]:
        out["after_" + k] = npy(p[k])
    return out


def run_densify(ns, spec):
    p, cams, curr, variables = setup(ns, spec)
    P0 = p["means3D"].shape[0]
    dist = "isotropic" if spec["frame"].get("iso", False) else "anisotropic"
    before = {k: p[k].detach().clone() for k in PARAM_KEYS[:5]}
    with torch.no_grad():
        p2, v2 = ns.splatam.add_new_gaussians(p, variables, curr, spec["sil_thres"], spec["t"], "projective", dist)
    out = {"P0": np.int64(P0), "P1": np.int64(p2["means3D"].shape[0])}
    for k in PARAM_KEYS[:5]:
        out["new_" + k] = npy(p2[k][P0:])
        assert torch.equal(p2[k][:P0], before[k])
    out["timestep"] = npy(v2["timestep"])
    out["max_2D_radius_len"] = np.int64(v2["max_2D_radius"].shape[0])
    return out


def run_render(ns, spec):
    from activesplat_b200 import scenes
    params, _, _, _ = helpers.frame_case(**spec["frame"])
    cam = scenes.candidate_views(3, seed=spec["frame"]["seed"])[spec["view"]]
    p = {k: v.clone().to(DEV) for k, v in params.items()}
    rv, drv = ns.splatam.get_rendervars(p, cam.w2c)
    cfg = dict(viz_w=cam.width, viz_h=cam.height, viz_near=cam.near, viz_far=cam.far)
    im, depth, opacity, sil = ns.splatam.render(cam.w2c, cam.k(), rv, drv, cfg)
    return dict(im=npy(im), depth=npy(depth), opacity=npy(opacity), sil=npy(sil))


RUNNERS = {"loss": run_loss, "adam": run_adam, "densify": run_densify, "render": run_render}


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    outdir = args[0] if args else os.path.join(ROOT, "gpurun_out", "golden_splatam")
    os.makedirs(outdir, exist_ok=True)
    ns = build_ref_py.load(build_ref_py.load_ref_extension())
    for name, spec in helpers.SPLATAM_CASES.items():
        out = RUNNERS[spec["kind"]](ns, spec)
        np.savez_compressed(os.path.join(outdir, f"ref_splatam_{name}.npz"), **out)
        brief = {k: (float(v) if np.ndim(v) == 0 else list(np.shape(v))) for k, v in out.items()}
        print(f"[{name}] {brief}", flush=True)


if __name__ == "__main__":
    main()
