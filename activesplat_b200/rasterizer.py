"""Python surface of the rasterizer: a drop-in for `diff_gaussian_rasterization`.

Mirrors dgr/diff_gaussian_rasterization/__init__.py of the reference:
  GaussianRasterizationSettings  (:157-169)  same 12 fields, same order
  GaussianRasterizer             (:171-220)  .forward(...) -> (color, radii, depth, opacity), .markVisible
  rasterize_gaussians            (:21-42)
  _RasterizeGaussians            (:44-155)   only grad_out_color is back-propagated (:101)
backed by activesplat_b200._C (C ABI -> sm_100a kernels).  `colors_precomp` may carry 3 channels
(reference behaviour) or 6 (rgb | z, 1, z^2 in one sweep).
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _snapshot(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _call_with_debug_dump(fn, args, debug, dump_name, note):
    """debug=True keeps a CPU copy of the arguments and writes it out if the call throws (__init__.py:82-92)."""
    if not debug:
        return fn(*args)
    saved = _snapshot(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print(f"\nAn error occured in {note}. Writing {dump_name} for debugging.\n")
        raise


def make_rasterize_function(backend):
    """The autograd.Function of the reference (__init__.py:44-155) bound to a `_C`-shaped backend:
    any object with rasterize_gaussians / rasterize_gaussians_backward in the reference's signature."""

    class _Fn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs):
            return _forward_impl(backend, ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                 cov3Ds_precomp, rs)

        @staticmethod
        def backward(ctx, grad_out_color, _grad_radii, _grad_depth, _grad_opacity):
            return _backward_impl(backend, ctx, grad_out_color)

    return _Fn


def _forward_impl(backend, ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                  raster_settings):
    rs = raster_settings
    args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
            rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
    num_rendered, color, radii, geom, binning, img, depth, opacity = _call_with_debug_dump(
        backend.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
    ctx.raster_settings = rs
    ctx.num_rendered = num_rendered
    ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img)
    ctx.mark_non_differentiable(radii)
    return color, radii, depth, opacity


def _backward_impl(backend, ctx, grad_out_color):
    rs = ctx.raster_settings
    colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img = ctx.saved_tensors
    args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos,
            geom, ctx.num_rendered, binning, img, rs.debug)
    (g_means2D, g_colors, g_opacities, g_means3D, g_cov3D, g_sh, g_scales, g_rotations) = _call_with_debug_dump(
        backend.rasterize_gaussians_backward, args, rs.debug, "snapshot_bw.dump", "backward")
    return (g_means3D, g_means2D, g_sh, g_colors, g_opacities, g_scales, g_rotations, g_cov3D, None)


_RasterizeGaussians = make_rasterize_function(_C)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None or rotations is not None
        full_sr = scales is not None and rotations is not None
        if (not full_sr and cov3D_precomp is None) or (have_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        absent = torch.Tensor([])  # empty CPU tensor == "not provided" (null data pointer in the reference)
        return rasterize_gaussians(
            means3D, means2D,
            absent if shs is None else shs,
            absent if colors_precomp is None else colors_precomp,
            opacities,
            absent if scales is None else scales,
            absent if rotations is None else rotations,
            absent if cov3D_precomp is None else cov3D_precomp,
            self.raster_settings)
