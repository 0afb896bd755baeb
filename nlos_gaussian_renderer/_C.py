"""`from nlos_gaussian_renderer import _C` -> the C-ABI backed mirror of the reference's pybind module."""
try:
    from nlos_gaussian_renderer_b200 import _lib as _l
    _l.lib()                                            # load libnlosb200.so now: a missing library must surface as ImportError
except (OSError, RuntimeError) as e:                    # pragma: no cover
    raise ImportError(f"libnlosb200.so is not available: {e}")
from nlos_gaussian_renderer_b200._C import (MAX_GAUSSIANS_PER_RAY, MAX_SECTIONS_PER_RAY, filter_gaussians_per_ray,  # noqa: F401,E402
                                            render_rays, render_rays_analytic, render_rays_backward)
