"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/nlosb200.h declares, the
Python mirror has the reference's names and call signatures (checked against the reference sources with `ast` when
/root/reference is present), and the error convention holds without a GPU (no compute call is made here)."""
import ast
import ctypes
import inspect
import os

import pytest
import torch

from nlos_gaussian_renderer_b200 import _lib

REF = os.environ.get("NLOS_REFERENCE_ROOT", "/root/reference")
has_ref = os.path.isdir(os.path.join(REF, "gaussian_model"))


def test_library_exports_every_declared_symbol():
    from nlos_gaussian_renderer_b200 import build
    L = ctypes.CDLL(build.build())
    names = _lib.declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), n
    assert _lib.lib().nlosb200_record_floats() == 32 and _lib.lib().nlosb200_grad_floats() == 16
    assert _lib.lib().nlosb200_version() >= 100


def test_shim_modules_have_the_reference_names():
    from nlos_gaussian_renderer import _C
    import cuda_renderer
    from nlos_gaussian_renderer_b200.dropin import cuda_autograd, rendering_cuda, rendering_section
    for f in ("render_rays", "filter_gaussians_per_ray", "render_rays_analytic"):      # bindings.cpp:26-35
        assert callable(getattr(_C, f))
    for n in ("NLOSGaussianRenderer", "SectionGaussianRendererCUDA", "create_renderer", "create_section_renderer", "CUDA_AVAILABLE"):
        assert hasattr(cuda_renderer, n)
    for m, names in ((cuda_autograd, ("CUDARenderFunction", "CUDARenderModule", "create_cuda_render_module", "CUDA_AVAILABLE")),
                     (rendering_cuda, ("GaussianRendererCUDA", "create_cuda_renderer", "CUDA_AVAILABLE")),
                     (rendering_section, ("GaussianSectionRenderer", "create_section_renderer", "CUDA_AVAILABLE"))):
        for n in names:
            assert hasattr(m, n), (m.__name__, n)
    assert cuda_autograd.CUDARenderModule(sigma_threshold=2.5).sigma_threshold == 2.5


def _ref_signatures(path):
    """{qualified name: [(arg, has_default)]} of every def in a reference source file (never imported: it does not import at HEAD)."""
    tree = ast.parse(open(os.path.join(REF, path)).read())
    out = {}

    def visit(node, prefix):
        for ch in ast.iter_child_nodes(node):
            if isinstance(ch, ast.ClassDef):
                visit(ch, prefix + ch.name + ".")
            elif isinstance(ch, ast.FunctionDef):
                a = ch.args
                nd = len(a.args) - len(a.defaults)
                out[prefix + ch.name] = [(x.arg, i >= nd) for i, x in enumerate(a.args)]
    visit(tree, "")
    return out


def _our_signature(obj):
    sig = inspect.signature(obj)
    return [(n, p.default is not inspect.Parameter.empty) for n, p in sig.parameters.items()
            if p.kind in (p.POSITIONAL_OR_KEYWORD,)]


@pytest.mark.skipif(not has_ref, reason="/root/reference not present (GPU box)")
def test_signatures_match_the_reference_sources():
    import cuda_renderer
    from nlos_gaussian_renderer_b200.dropin import cuda_autograd, rendering_cuda, rendering_section
    cases = [
        ("gaussian_model/cuda_autograd.py", cuda_autograd, ["CUDARenderFunction.forward", "CUDARenderFunction.backward",
                                                             "CUDARenderModule.__init__", "CUDARenderModule.forward",
                                                             "create_cuda_render_module"]),
        ("gaussian_model/rendering_cuda.py", rendering_cuda, ["GaussianRendererCUDA.__init__", "GaussianRendererCUDA.render_transient",
                                                              "GaussianRendererCUDA.render_from_spherical_samples", "create_cuda_renderer"]),
        ("gaussian_model/rendering_section.py", rendering_section, ["GaussianSectionRenderer.__init__", "GaussianSectionRenderer.render_transient",
                                                                    "GaussianSectionRenderer.render_from_spherical_samples",
                                                                    "create_section_renderer"]),
        ("submodules/cuda_renderer/__init__.py", cuda_renderer, ["NLOSGaussianRenderer.__init__", "NLOSGaussianRenderer.render",
                                                                 "NLOSGaussianRenderer.filter_gaussians", "create_renderer"]),
        ("submodules/cuda_renderer/section_renderer.py", cuda_renderer, ["SectionGaussianRendererCUDA.__init__",
                                                                         "SectionGaussianRendererCUDA.render_transient",
                                                                         "SectionGaussianRendererCUDA.render_from_spherical_samples",
                                                                         "SectionGaussianRendererCUDA.filter_gaussians",
                                                                         "create_section_renderer"]),
    ]
    for path, mod, names in cases:
        ref = _ref_signatures(path)
        for q in names:
            obj = mod
            for part in q.split("."):
                obj = getattr(obj, part)
            assert _our_signature(obj) == ref[q], (path, q, _our_signature(obj), ref[q])


@pytest.mark.skipif(not has_ref, reason="/root/reference not present (GPU box)")
def test_C_positional_arguments_match_the_pybind_headers():
    """include/volume_renderer.h:8-24, include/ray_aabb.h:7-13, src/bindings.cpp:6-24 -- argument ORDER of the three functions."""
    import re
    from nlos_gaussian_renderer import _C

    def c_args(path, fn):
        text = open(os.path.join(REF, "submodules/cuda_renderer", path)).read()
        text = re.sub(r"//[^\n]*", "", text)
        m = re.search(fn + r"\s*\(([^)]*)\)", text, flags=re.S)
        return [re.split(r"[\s&*]+", a.strip())[-1] for a in m.group(1).split(",")]
    for path, fn in (("include/volume_renderer.h", "render_rays"), ("include/ray_aabb.h", "filter_gaussians_per_ray"),
                     ("src/bindings.cpp", "render_rays_analytic")):
        ours = [n for n, p in inspect.signature(getattr(_C, fn)).parameters.items() if p.default is inspect.Parameter.empty]
        assert ours == c_args(path, fn), (fn, ours, c_args(path, fn))


def test_error_convention_without_gpu():
    """CHECK_INPUT (cuda_utils.cuh:7-9): CPU tensors raise RuntimeError -- and nothing silently runs on the CPU."""
    from nlos_gaussian_renderer import _C
    z = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        _C.filter_gaussians_per_ray(z, z, z, torch.zeros(4, 6), 3.0)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        _C.render_rays(z, z, torch.zeros(5), z, z, torch.zeros(4, 4), torch.zeros(4, 1), torch.zeros(4, 1), torch.zeros(3), 0, 1.0, 0.01,
                       1.0, True, "netf")
    from nlos_gaussian_renderer_b200.transient import _f32c
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _f32c(z)
    rc = _lib.lib().nlosb200_bin_window_plan(None, 0, 0, None, 0, 0.0, 1.0, 0, 128, 8, 1, 0.0, 5.3, 0, 0, 0, None, None, None, None, None, None)
    assert rc < 0 and b"bad sizes" in _lib.lib().nlosb200_last_error()
