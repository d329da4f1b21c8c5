"""Generate golden vectors from the UNMODIFIED reference CUDA rasterizer (TEST INFRASTRUCTURE ONLY).

Runs on a GPU box:   python oracle/make_golden.py [outdir]
It loads oracle/_ref/ref_dgr_C.so (built by oracle/build_ref.py from the sources under
/root/reference, never copied), feeds it the seeded cases of tests/helpers.py, and stores what the
reference produced -- outputs, gradients and the integer intermediates pulled out of its scratch
buffers (layout: dgr/cuda_rasterizer/rasterizer_impl.cu:155-194) -- as tests/golden/ref_<case>.npz.
The reference ships no golden vectors of its own; these pin the oracle.

It also prints a three-way report (reference CUDA / CPU oracle / sm_100a kernels).
"""
import importlib.util
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402


def load_ref():
    path = os.path.join(HERE, "_ref", "ref_dgr_C.so")
    spec = importlib.util.spec_from_file_location("ref_dgr_C", path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _al(x, a=128):
    return (x + a - 1) // a * a


def ref_geometry_offsets(P):
    """GeometryState::fromChunk (rasterizer_impl.cu:155-170) up to tiles_touched."""
    o = {}
    p = 0
    for name, size in (("depths", 4 * P), ("clamped", 3 * P), ("internal_radii", 4 * P), ("means2D", 8 * P),
                       ("cov3D", 24 * P), ("conic_opacity", 16 * P), ("rgb", 12 * P), ("tiles_touched", 4 * P)):
        p = _al(p)
        o[name] = p
        p += size
    return o


def ref_binning_offsets(R):
    """BinningState::fromChunk (rasterizer_impl.cu:181-194) up to keys_unsorted."""
    o = {}
    p = 0
    for name, size in (("point_list", 4 * R), ("point_list_unsorted", 4 * R), ("keys", 8 * R), ("keys_unsorted", 8 * R)):
        p = _al(p)
        o[name] = p
        p += size
    return o


def ref_image_offsets(N):
    """ImageState::fromChunk (rasterizer_impl.cu:172-179)."""
    o = {}
    p = 0
    for name, size in (("accum_alpha", 4 * N), ("n_contrib", 4 * N), ("ranges", 8 * N)):
        p = _al(p)
        o[name] = p
        p += size
    return o


def view_of(buf, off, dtype, count):
    base = buf.data_ptr()
    pad = (-base) % 128
    raw = buf.cpu().numpy()
    o = pad + off
    return np.frombuffer(raw[o:o + count * np.dtype(dtype).itemsize].tobytes(), dtype=dtype).copy()


def run_reference(ref, case, device="cuda:0"):
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
    e = torch.Tensor([])
    P, H, W = case["means3D"].shape[0], case["H"], case["W"]
    ins = dict(bg=t(case["bg"]), means3D=t(case["means3D"]), colors=t(case["colors"]), op=t(case["opacities"]),
               scales=t(case["scales"]), rots=t(case["rotations"]), view=t(case["view"]), proj=t(case["proj"]),
               campos=t(case["campos"]))
    R, color, radii, geom, binning, img, depth, opacity = ref.rasterize_gaussians(
        ins["bg"], ins["means3D"], ins["colors"], ins["op"], ins["scales"], ins["rots"], case["scale_modifier"], e,
        ins["view"], ins["proj"], case["tanfovx"], case["tanfovy"], H, W, e, 0, ins["campos"], False, False)
    torch.cuda.synchronize()
    go, bo, io = ref_geometry_offsets(P), ref_binning_offsets(R), ref_image_offsets(H * W)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    vis = radii.cpu().numpy() > 0
    out = dict(num_rendered=np.int64(R), color=color.cpu().numpy(), radii=radii.cpu().numpy(),
               depth=depth.cpu().numpy(), opacity=opacity.cpu().numpy(),
               means2D=view_of(geom, go["means2D"], np.float32, 2 * P).reshape(P, 2),
               depths=view_of(geom, go["depths"], np.float32, P),
               conic_opacity=view_of(geom, go["conic_opacity"], np.float32, 4 * P).reshape(P, 4),
               cov3D=view_of(geom, go["cov3D"], np.float32, 6 * P).reshape(P, 6),
               tiles_touched=view_of(geom, go["tiles_touched"], np.uint32, P),
               final_T=view_of(img, io["accum_alpha"], np.float32, H * W).reshape(H, W),
               n_contrib=view_of(img, io["n_contrib"], np.uint32, H * W).reshape(H, W),
               ranges=view_of(img, io["ranges"], np.uint32, 2 * tiles).reshape(tiles, 2),
               keys=view_of(binning, bo["keys"], np.uint64, R),
               point_list=view_of(binning, bo["point_list"], np.uint32, R),
               keys_unsorted=view_of(binning, bo["keys_unsorted"], np.uint64, R),
               vals_unsorted=view_of(binning, bo["point_list_unsorted"], np.uint32, R))
    # the reference leaves per-Gaussian scratch of culled Gaussians uninitialised: blank it for storage
    for k in ("means2D", "depths", "conic_opacity", "cov3D"):
        out[k][~vis] = 0
    g = ref.rasterize_gaussians_backward(
        ins["bg"], ins["means3D"], radii, ins["colors"], ins["scales"], ins["rots"], case["scale_modifier"], e,
        ins["view"], ins["proj"], case["tanfovx"], case["tanfovy"], t(case["dL_dpix"]), e, 0, ins["campos"], geom, R,
        binning, img, False)
    torch.cuda.synchronize()
    names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"]
    for n, x in zip(names, g):
        if n != "sh":
            out["grad_" + n] = x.cpu().numpy()
    return out


def summarize(tag, a, b, int_keys, float_keys, vis=None):
    rep = {}
    for k in int_keys:
        if k in a and k in b and a[k].shape == b[k].shape:
            rep[k] = int((a[k] != b[k]).sum())
        else:
            rep[k] = "shape %s vs %s" % (getattr(a.get(k), "shape", None), getattr(b.get(k), "shape", None))
    for k in float_keys:
        if k in a and k in b and a[k].shape == b[k].shape:
            x, y = a[k].astype(np.float64), b[k].astype(np.float64)
            if vis is not None and x.shape[0] == vis.shape[0] and k in ("means2D", "depths", "conic_opacity"):
                x, y = x[vis], y[vis]
            scale = max(np.abs(y).max(), 1e-12)
            rep[k] = dict(max_rel=float(np.abs(x - y).max() / scale), bitexact=bool((x == y).all()),
                          n_gt_1e4=int((np.abs(x - y) / scale > 1e-4).sum()))
    print(tag, json.dumps(rep))
    return rep


INT_KEYS = ["radii", "tiles_touched", "keys_unsorted", "vals_unsorted", "keys", "point_list", "ranges", "n_contrib"]
FLOAT_KEYS = ["means2D", "depths", "conic_opacity", "color", "depth", "opacity", "final_T"]
GRAD_KEYS = ["grad_means2D", "grad_colors", "grad_opacity", "grad_means3D", "grad_cov3D", "grad_scales", "grad_rotations"]


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    ref_only = "--ref-only" in sys.argv
    outdir = args[0] if args else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(outdir, exist_ok=True)
    ref = load_ref()
    report = {}
    for name in helpers.CASES:
        case = helpers.make_case(name)
        r = run_reference(ref, case)
        np.savez_compressed(os.path.join(outdir, f"ref_{name}.npz"), **r)
        print(f"[{name}] P={case['means3D'].shape[0]} {case['W']}x{case['H']} R={int(r['num_rendered'])} "
              f"visible={(r['radii'] > 0).sum()}")
        vis = r["radii"] > 0
        st, g = helpers.run_oracle(case)
        orc = dict(st)
        orc.update({"grad_" + k: v for k, v in g.items()})
        report[name] = {"oracle_vs_ref": summarize(f"[{name}] oracle vs ref:", orc, r, INT_KEYS, FLOAT_KEYS + GRAD_KEYS, vis)}
        if ref_only:
            continue
        try:
            mine, mg = helpers.run_cuda(case)
            mine.update({"grad_" + k: v for k, v in mg.items()})
            report[name]["cuda_vs_ref"] = summarize(f"[{name}] cuda vs ref:  ", mine, r, INT_KEYS, FLOAT_KEYS + GRAD_KEYS, vis)
            report[name]["cuda_vs_oracle"] = summarize(f"[{name}] cuda vs oracle:", mine, orc, INT_KEYS, FLOAT_KEYS + GRAD_KEYS, vis)
        except Exception as ex:  # the golden files are still useful if our kernels are broken
            print(f"[{name}] cuda path failed: {type(ex).__name__}: {ex}")
            report[name]["cuda_error"] = str(ex)
    with open(os.path.join(outdir, "report.json"), "w") as f:
        json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
