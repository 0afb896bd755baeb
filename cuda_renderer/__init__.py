"""Import-name shim for the reference package `cuda_renderer` (submodules/cuda_renderer/__init__.py; imported by
gaussian_model/rendering_section.py:12 and test_cuda_renderer.py:15)."""
from nlos_gaussian_renderer import _C  # noqa: F401
from nlos_gaussian_renderer_b200.dropin.cuda_renderer_pkg import (CUDA_AVAILABLE, NLOSGaussianRenderer,  # noqa: F401
                                                                  SectionGaussianRendererCUDA, create_renderer,
                                                                  create_section_renderer)
