"""Drop-in for the reference's `diff_gaussian_rasterization` package
(submodules_fsgs/diff-gaussian-rasterization-confidence/diff_gaussian_rasterization/__init__.py): same names, served by
the sm_100a kernels of dynamic3dgaussians_b200 through the C ABI in include/gsr_b200.h.

    from diff_gaussian_rasterization import GaussianRasterizer as Renderer            # REF/train.py:9
    from diff_gaussian_rasterization import GaussianRasterizationSettings as Camera   # REF/helpers.py:5
"""
from dynamic3dgaussians_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    _RasterizeGaussians,
    rasterize_gaussians,
    rasterize_views,
)
import dynamic3dgaussians_b200.rasterizer as _impl


def set_default_returns(n: int) -> None:
    """Selects which tuple GaussianRasterizer(...)() returns by default: 3, 4 or 5 (see GaussianRasterizer)."""
    if n not in (3, 4, 5):
        raise ValueError("returns must be 3, 4 or 5")
    _impl.DEFAULT_RETURNS = n
