"""Import-compatible alias: `from diff_gaussian_rasterization import GaussianRasterizer` keeps working
(reference call sites: src/mapper/splatam/splatam.py:22-23, utils/recon_helpers.py:2)."""
from activesplat_b200 import _C  # noqa: F401
from activesplat_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                         rasterize_gaussians)
