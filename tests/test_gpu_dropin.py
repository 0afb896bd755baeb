"""The reference-facing facade on the GPU: what the reference's own scripts (test_cuda_autograd.py:47-179,
test_cuda_renderer.py:8-106) exercise -- imports, construction, shapes, backward -- but with asserts, plus the
identity that ties the batched [scan_y, scan_x, T] renderer to the per-scan-point operator."""
import math

import numpy as np
import pytest
import torch

from nlos_gaussian_renderer_b200 import scenes
from nlos_gaussian_renderer_b200.geometry import ScanGeometry
from nlos_gaussian_renderer_b200.transient import TransientRenderer
from tests import util
from tests.fake_model import FakeGaussianModel

pytestmark = pytest.mark.gpu


def test_reference_smoke_shapes_and_real_gradients():
    """test_cuda_autograd.py: 10 Gaussians, 4x4 rays, 10 samples, r in [0.5, 2], c=1, deltaT=0.01, occlusion on."""
    from nlos_gaussian_renderer_b200.dropin.cuda_autograd import CUDARenderModule, create_cuda_render_module
    torch.manual_seed(0)
    gm = FakeGaussianModel.random(10, device="cuda", spread=0.3, centre=(0.0, 1.0, 0.0), scale=0.15)
    module = create_cuda_render_module(sigma_threshold=3.0)
    assert isinstance(module, CUDARenderModule)
    cam = torch.zeros(3, device="cuda")
    result, hist = module(gaussian_model=gm, camera_pos=cam, theta_range=(math.pi / 4, 3 * math.pi / 4),
                          phi_range=(math.pi / 4, 3 * math.pi / 4), r_range=(0.5, 2.0), num_theta=4, num_phi=4, num_r=10, c=1.0,
                          deltaT=0.01, scaling_modifier=1.0, use_occlusion=True, rendering_type="netf")
    assert tuple(result.shape) == (10, 4, 4) and tuple(hist.shape) == (10,)
    assert torch.isfinite(result).all() and float(hist.abs().max()) > 0
    hist.sum().backward()
    for p in (gm._mu, gm._scaling, gm._rotation, gm._opacity, gm._features_dc):
        assert p.grad is not None and torch.isfinite(p.grad).all()
        assert float(p.grad.abs().max()) > 0          # the reference returns zeros here (cuda_autograd.py:147-156)


def test_factories_and_facades():
    import cuda_renderer
    from nlos_gaussian_renderer_b200.dropin import rendering_cuda, rendering_section
    assert cuda_renderer.CUDA_AVAILABLE and rendering_cuda.CUDA_AVAILABLE
    gm = FakeGaussianModel.random(300, device="cuda", spread=0.2, centre=(0.0, 0.75, 0.0), scale=0.04)
    cam = torch.tensor([0.1, 0.0, -0.05], device="cuda")
    kw = dict(gaussian_model=gm, camera_pos=cam, theta_range=(0.9, 2.2), phi_range=(0.8, 2.3), r_range=(0.35, 1.65), num_theta=6,
              num_phi=6, num_r=32, c=1.0, deltaT=1.3 / 32, scaling_modifier=1.0, use_occlusion=False, rendering_type="netf")
    r1 = rendering_cuda.create_cuda_renderer()
    res1, h1 = r1.render_transient(**kw)
    r2 = cuda_renderer.create_renderer()
    res2, h2 = r2.render(**kw)
    assert torch.allclose(res1, res2) and torch.allclose(h1, h2)
    # render_from_spherical_samples recovers the same ranges from a [Na,5] lattice (rendering_cuda.py:265-336)
    th = torch.linspace(0.9, 2.2, 6, device="cuda")
    ph = torch.linspace(0.8, 2.3, 6, device="cuda")
    pts = torch.zeros(36, 5, device="cuda")
    pts[:, 3] = th.repeat_interleave(6)
    pts[:, 4] = ph.repeat(6)
    start = int(round(0.35 / (1.3 / 32)))
    res3, h3 = r1.render_from_spherical_samples(gm, pts, cam, start, start + 32, 32, 6, 0.0, 0.0, 1.0, 1.3 / 32, 1.0, False, "netf")
    assert tuple(res3.shape) == (32, 36) and tuple(h3.shape) == (32,)
    # section renderer: per-ray scalar in bin num_r//2, histogram broadcast (section_renderer.py:171-184)
    sec = rendering_section.create_section_renderer()
    res4, h4 = sec.render_transient(**kw)
    assert tuple(res4.shape) == (32, 6, 6) and tuple(h4.shape) == (32,)
    assert float(res4[:16].abs().max()) == 0 and float(res4[17:].abs().max()) == 0
    assert torch.allclose(h4, h4[0].expand(32))
    flt = r2.filter_gaussians(cam.expand(36, 3).contiguous(), torch.nn.functional.normalize(torch.rand(36, 3, device="cuda"), dim=1),
                              gm.get_mu.detach(), gm.get_bboxes())
    assert tuple(flt.shape) == (36, 257) and flt.dtype == torch.int32


@pytest.mark.parametrize("mode_occl", [False])
def test_batched_slice_equals_per_scan_point_operator(mode_occl):
    """H[y, x, :] of the batched renderer == pred_histogram * vp_y^2 of one CUDARenderModule call for that scan point
    (nlos_helpers.py:255-276), with the 256-per-ray cap not binding."""
    from nlos_gaussian_renderer_b200.dropin.cuda_autograd import CUDARenderModule
    N, Sy, Sx, T, A = 200, 3, 5, 64, 8
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=2.0, device="cuda")
    H = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode="numerical").forward(*params)
    gm = FakeGaussianModel(*params)
    module = CUDARenderModule()
    r0, r1 = ta["start"] * ta["c"] * ta["deltaT"], ta["end"] * ta["c"] * ta["deltaT"]
    for p in (0, 7, 14):
        th, ph = geom.theta_range[p].tolist(), geom.phi_range[p].tolist()
        _, hist = module(gm, geom.scan_o[p], tuple(th), tuple(ph), (r0, r1), A, A, T, ta["c"], ta["deltaT"], 1.0, mode_occl, "netf")
        ref = (hist * geom.vp_y ** 2).detach().cpu().numpy()
        got = H.reshape(Sy * Sx, T)[p].cpu().numpy()
        assert util.rel_to_max(got, ref) < 1e-4


@pytest.mark.parametrize("occl", [True, False])
def test_occluded_batch_equals_per_scan_point_loop(occl):
    """render_transient_batch_occluded (native batched kernels, csrc/occluded.cu) == the reference's loop of CUDARenderModule.forward over
    the scan points (main.py:198-203) times vp_y^2 (nlos_helpers.py:275-276), values and parameter gradients."""
    from nlos_gaussian_renderer_b200.dropin.cuda_autograd import CUDARenderModule
    from nlos_gaussian_renderer_b200.occluded import render_transient_batch_occluded
    N, Sy, Sx, T, A = 120, 2, 3, 48, 6
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=3.0, device="cuda")
    params = list(params)
    params[3] = params[3] + 3.0                                      # opaque enough for the transmittance to matter
    r0, r1 = ta["start"] * ta["c"] * ta["deltaT"], ta["end"] * ta["c"] * ta["deltaT"]
    t_samples = torch.linspace(r0, r1, T, device="cuda")
    pa = [p.clone().requires_grad_() for p in params]
    H = render_transient_batch_occluded(geom, t_samples, *pa, ta["c"], ta["deltaT"], use_occlusion=occl, chunk_scan_points=32)
    assert tuple(H.shape) == (Sy, Sx, T)
    wgt = torch.linspace(0.5, 1.5, Sy * Sx * T, device="cuda").reshape(Sy, Sx, T)
    (H * wgt).sum().backward()
    gm = FakeGaussianModel(*params)
    module = CUDARenderModule()
    rows = []
    for p in range(Sy * Sx):
        th, ph = geom.theta_range[p].tolist(), geom.phi_range[p].tolist()
        _, hist = module(gm, geom.scan_o[p], tuple(th), tuple(ph), (r0, r1), A, A, T, ta["c"], ta["deltaT"], 1.0, occl, "netf")
        rows.append(hist * geom.vp_y ** 2)
    Href = torch.stack(rows).reshape(Sy, Sx, T)
    (Href * wgt).sum().backward()
    assert util.rel_to_max(H.detach().cpu().numpy(), Href.detach().cpu().numpy()) < 1e-5
    if occl:                                                          # occlusion changes the result on this scene
        Hn = render_transient_batch_occluded(geom, t_samples, *params, ta["c"], ta["deltaT"], use_occlusion=False)
        assert util.rel_to_max(Hn.cpu().numpy(), H.detach().cpu().numpy()) > 1e-3
    for a, b in zip(pa, (gm._mu, gm._scaling, gm._rotation, gm._opacity, gm._features_dc)):
        assert float(b.grad.abs().max()) > 0
        assert util.rel_to_max(a.grad.reshape(-1).cpu().numpy(), b.grad.reshape(-1).cpu().numpy()) < 1e-4
