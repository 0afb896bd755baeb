"""Host-side mirror of the reference's renderer facade for the hot path: the same module-level names,
class names, call signatures, return shapes and error behaviour as

    gaussian_model/cuda_autograd.py       CUDARenderFunction, CUDARenderModule, create_cuda_render_module
    gaussian_model/rendering_cuda.py      GaussianRendererCUDA, create_cuda_renderer, CUDA_AVAILABLE
    gaussian_model/rendering_section.py   GaussianSectionRenderer, create_section_renderer
    submodules/cuda_renderer/__init__.py  NLOSGaussianRenderer, create_renderer
    submodules/cuda_renderer/section_renderer.py  SectionGaussianRendererCUDA, create_section_renderer

so that main.py / nlos_helpers.py / the two test scripts of the reference run unchanged once these files replace
theirs.  All arithmetic happens in libnlosb200.so; this layer only builds ray lattices and reshapes."""
from .lattice import ray_lattice, attenuate_and_integrate  # noqa: F401
