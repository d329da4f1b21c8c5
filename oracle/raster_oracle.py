"""ctypes/numpy front end of the CPU oracle (oracle/raster_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.

`forward()` follows CudaRasterizer::Rasterizer::forward (dgr/cuda_rasterizer/rasterizer_impl.cu:198-342)
step by step and returns every intermediate the parity tests compare (keys, sorted list, ranges...).
`backward()` follows Rasterizer::backward (rasterizer_impl.cu:346-441).
`OracleRasterize` is the autograd wrapper with the gradient routing of
dgr/diff_gaussian_rasterization/__init__.py:101-155 (only grad_out_color is used).
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "raster_oracle.c")):
            import importlib.util
            spec = importlib.util.spec_from_file_location("_orc_build", os.path.join(_HERE, "build_oracle.py"))
            m = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(m)
            m.build(force=True)
        _lib = ctypes.CDLL(path)
        _lib.orc_inclusive_scan.restype = ctypes.c_int64
        _lib.orc_higher_msb.restype = ctypes.c_uint32
        _lib.orc_num_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    """Use n OpenMP threads (launchers such as torchrun export OMP_NUM_THREADS=1)."""
    lib().orc_set_num_threads(ctypes.c_int(int(n)))


def higher_msb(n):
    return int(lib().orc_higher_msb(ctypes.c_uint32(n)))


def mark_visible(means3D, viewmatrix):
    means3D = _f32(means3D)
    view = _f32(viewmatrix).reshape(-1)
    P = means3D.shape[0]
    out = np.zeros(P, dtype=np.uint8)
    lib().orc_mark_visible(ctypes.c_int(P), _p(means3D), _p(view), _p(out))
    return out.astype(bool)


def forward(means3D, colors, opacities, scales, rotations, scale_modifier, cov3D_precomp,
            viewmatrix, projmatrix, tanfovx, tanfovy, H, W, bg):
    """Returns a dict with outputs and all intermediates. colors is [P,C], C in {3,6}."""
    L = lib()
    means3D = _f32(means3D)
    colors = _f32(colors)
    opacities = _f32(opacities).reshape(-1)
    P, C = means3D.shape[0], colors.shape[1]
    scales = _f32(scales) if scales is not None and np.size(scales) else None
    rotations = _f32(rotations) if rotations is not None and np.size(rotations) else None
    cov_pre = _f32(cov3D_precomp) if cov3D_precomp is not None and np.size(cov3D_precomp) else None
    view = _f32(viewmatrix).reshape(-1)
    proj = _f32(projmatrix).reshape(-1)
    bg = _f32(bg).reshape(-1)
    assert bg.shape[0] == C
    gx, gy = (W + 15) // 16, (H + 15) // 16
    tiles = gx * gy

    st = dict(P=P, C=C, W=W, H=H, tanfovx=float(tanfovx), tanfovy=float(tanfovy),
              scale_modifier=float(scale_modifier), view=view, proj=proj, bg=bg,
              means3D=means3D, colors=colors, scales=scales, rotations=rotations, cov3D_precomp=cov_pre)
    radii = np.zeros(P, np.int32)
    means2D = np.zeros((P, 2), np.float32)
    depths = np.zeros(P, np.float32)
    cov3D = np.zeros((P, 6), np.float32)
    conic_op = np.zeros((P, 4), np.float32)
    tiles_touched = np.zeros(P, np.uint32)
    L.orc_preprocess(ctypes.c_int(P), _p(means3D), _p(scales), ctypes.c_float(scale_modifier), _p(rotations),
                     _p(opacities), _p(cov_pre), _p(view), _p(proj), ctypes.c_int(W), ctypes.c_int(H),
                     ctypes.c_float(tanfovx), ctypes.c_float(tanfovy), _p(radii), _p(means2D), _p(depths),
                     _p(cov3D), _p(conic_op), _p(tiles_touched))
    offsets = np.zeros(P, np.uint32)
    R = int(L.orc_inclusive_scan(ctypes.c_int(P), _p(tiles_touched), _p(offsets)))
    keys_unsorted = np.zeros(R, np.uint64)
    vals_unsorted = np.zeros(R, np.uint32)
    L.orc_duplicate_with_keys(ctypes.c_int(P), _p(means2D), _p(depths), _p(offsets), _p(radii),
                              ctypes.c_int(W), ctypes.c_int(H), _p(keys_unsorted), _p(vals_unsorted))
    bit = higher_msb(tiles)
    keys = np.zeros(R, np.uint64)
    point_list = np.zeros(R, np.uint32)
    L.orc_sort_pairs(ctypes.c_int64(R), _p(keys_unsorted), _p(vals_unsorted), _p(keys), _p(point_list),
                     ctypes.c_int(32 + bit))
    ranges = np.zeros((tiles, 2), np.uint32)
    L.orc_tile_ranges(ctypes.c_int64(R), _p(keys), ctypes.c_int(tiles), _p(ranges))
    out_color = np.zeros((C, H, W), np.float32)
    out_depth = np.zeros((1, H, W), np.float32)
    out_opacity = np.zeros((1, H, W), np.float32)
    final_T = np.zeros((H, W), np.float32)
    n_contrib = np.zeros((H, W), np.uint32)
    L.orc_blend_forward(ctypes.c_int(W), ctypes.c_int(H), ctypes.c_int(C), _p(ranges), _p(point_list),
                        _p(means2D), _p(colors), _p(conic_op), _p(depths), _p(bg), _p(out_color),
                        _p(out_depth), _p(out_opacity), _p(final_T), _p(n_contrib))
    st.update(radii=radii, means2D=means2D, depths=depths, cov3D=cov3D, conic_opacity=conic_op,
              tiles_touched=tiles_touched, offsets=offsets, num_rendered=R, sort_bits=32 + bit,
              keys_unsorted=keys_unsorted, vals_unsorted=vals_unsorted, keys=keys, point_list=point_list,
              ranges=ranges, color=out_color, depth=out_depth, opacity=out_opacity,
              final_T=final_T, n_contrib=n_contrib)
    return st


def backward(st, dL_dpix):
    """Rasterizer::backward. Returns dict of grads shaped like rasterize_points.cu:155-163."""
    L = lib()
    P, C, W, H = st["P"], st["C"], st["W"], st["H"]
    dL_dpix = _f32(dL_dpix).reshape(C, H, W)
    g = dict(means2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 4), np.float32),
             opacity=np.zeros((P, 1), np.float32), colors=np.zeros((P, C), np.float32),
             means3D=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32),
             scales=np.zeros((P, 3), np.float32), rotations=np.zeros((P, 4), np.float32))
    L.orc_blend_backward(ctypes.c_int(W), ctypes.c_int(H), ctypes.c_int(C), _p(st["ranges"]), _p(st["point_list"]),
                         _p(st["bg"]), _p(st["means2D"]), _p(st["conic_opacity"]), _p(st["colors"]),
                         _p(st["final_T"]), _p(st["n_contrib"]), _p(dL_dpix), _p(g["means2D"]), _p(g["conic"]),
                         _p(g["opacity"]), _p(g["colors"]))
    cov3D = st["cov3D_precomp"] if st["cov3D_precomp"] is not None else st["cov3D"]
    L.orc_cov2d_backward(ctypes.c_int(P), _p(st["means3D"]), _p(st["radii"]), _p(cov3D), ctypes.c_int(W),
                         ctypes.c_int(H), ctypes.c_float(st["tanfovx"]), ctypes.c_float(st["tanfovy"]),
                         _p(st["view"]), _p(g["conic"]), _p(g["means3D"]), _p(g["cov3D"]))
    L.orc_preprocess_backward(ctypes.c_int(P), _p(st["means3D"]), _p(st["radii"]), _p(st["scales"]),
                              _p(st["rotations"]), ctypes.c_float(st["scale_modifier"]), _p(st["proj"]),
                              _p(g["means2D"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["scales"]), _p(g["rotations"]))
    return g


# ----------------------------------------------------------------------------------------------
# torch surface (CPU tensors) mirroring dgr/diff_gaussian_rasterization/__init__.py
# ----------------------------------------------------------------------------------------------
def _torch():
    import torch
    return torch


class _OracleRasterizeFn:
    """Built lazily so that importing this module does not import torch."""
    fn = None


def _get_fn():
    if _OracleRasterizeFn.fn is not None:
        return _OracleRasterizeFn.fn
    torch = _torch()

    class OracleRasterize(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs):
            npy = lambda t: t.detach().cpu().numpy()
            st = forward(npy(means3D), npy(colors_precomp), npy(opacities), npy(scales), npy(rotations),
                         rs.scale_modifier, npy(cov3Ds_precomp), npy(rs.viewmatrix), npy(rs.projmatrix),
                         rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, npy(rs.bg))
            ctx.st = st
            ctx.has_scales = scales.numel() > 0
            out = (torch.from_numpy(st["color"]), torch.from_numpy(st["radii"]),
                   torch.from_numpy(st["depth"]), torch.from_numpy(st["opacity"]))
            ctx.mark_non_differentiable(out[1])
            return out

        @staticmethod
        def backward(ctx, grad_out_color, _r, _d, _o):
            # __init__.py:101: only the colour gradient is consumed
            g = backward(ctx.st, grad_out_color.detach().cpu().numpy())
            t = torch.from_numpy
            return (t(g["means3D"]), t(g["means2D"]), None, t(g["colors"]), t(g["opacity"]),
                    t(g["scales"]) if ctx.has_scales else None, t(g["rotations"]) if ctx.has_scales else None,
                    None if ctx.has_scales else t(g["cov3D"]), None)

    _OracleRasterizeFn.fn = OracleRasterize
    return OracleRasterize


class GaussianRasterizer:
    """CPU-oracle twin of GaussianRasterizer (dgr/diff_gaussian_rasterization/__init__.py:171-220)."""

    def __init__(self, raster_settings):
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        torch = _torch()
        return torch.from_numpy(mark_visible(positions.detach().cpu().numpy(),
                                             self.raster_settings.viewmatrix.detach().cpu().numpy()))

    def __call__(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                 rotations=None, cov3D_precomp=None):
        torch = _torch()
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is not None:
            raise NotImplementedError("oracle: SH colours are not on the ActiveSplat path")
        e = torch.Tensor([])
        return _get_fn().apply(means3D, means2D, e, colors_precomp, opacities,
                               scales if scales is not None else e, rotations if rotations is not None else e,
                               cov3D_precomp if cov3D_precomp is not None else e, self.raster_settings)
