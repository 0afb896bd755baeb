"""Dense point evaluation (csrc/evaluate.cu; SURVEY 8 f4: gaussian2volume nlos_helpers.py:40-69 ->
estimate_rho_w_no_occlusion(out_separately=True) gaussian_model.py:346-364) against the fp64 restatement."""
import pytest
import torch

from nlos_gaussian_renderer_b200 import _C, scenes
from tests import ref_torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,Na,scale_mul", [(1, 1, 3.0), (37, 1000, 3.0), (700, 5001, 2.0), (3000, 300, 1.0), (0, 10, 1.0)])
def test_evaluate_points_matches_fp64(N, Na, scale_mul):
    params = scenes.make_gaussians(max(N, 1), 7, scale_mul=scale_mul)
    params = [p[:N].contiguous() for p in params]
    g = torch.Generator().manual_seed(N + Na)
    x = torch.rand(Na, 3, generator=g) * 0.5 + torch.tensor([-0.25, 0.5, -0.25])
    if N:                                                                   # half of the points near Gaussian centres: values of O(1)
        idx = torch.randint(0, N, (Na // 2,), generator=g)
        x[: Na // 2] = params[0][idx] + 0.3 * torch.exp(params[1][idx]) * torch.randn(Na // 2, 3, generator=g)
    rho, dens = _C.evaluate_points(x.cuda(), *[p.cuda() for p in params])
    want_rho, want_dens = ref_torch.evaluate_points(x.double(), *[p.double() for p in params]) if N else (torch.zeros(Na), torch.zeros(Na))
    for got, want in ((rho, want_rho), (dens, want_dens)):
        err = float((got.cpu().double() - want.double()).abs().max())
        assert err <= 2e-5 * max(float(want.abs().max()), 1e-3), err
    if N:
        assert float(dens.max()) > 0.1


def test_evaluate_points_equals_one_sample_of_render_rays():
    """Gaussians so wide that every 3-sigma box contains the sample: rho_density / (c deltaT) and density of an uncapped render_rays at
    x = o + t d are the dense sums evaluate_points returns at x (same per-camera albedo)."""
    params = [p.cuda() for p in scenes.make_gaussians(50, 3, scale_mul=20.0)]
    cam = torch.tensor([0.1, 0.0, -0.2]).cuda()
    d = torch.tensor([[0.05, 1.0, 0.1]])
    d = (d / d.norm()).cuda()
    t = torch.tensor([0.8]).cuda()
    rho_r, dens_r, _ = _C.render_rays(cam[None].contiguous(), d, t, *params, cam, 0, 1.0, 0.01, 1.0, False, "netf", max_per_ray=0)
    x = (cam[None] + t * d).contiguous()
    rho_e, dens_e = _C.evaluate_points(x, *params, camera_pos=cam, active_sh_degree=0)
    assert float(rho_e[0]) > 0.1
    assert abs(float(rho_r[0, 0]) / 0.01 - float(rho_e[0])) <= 1e-4 * float(rho_e[0])
    assert abs(float(dens_r[0, 0]) - float(dens_e[0])) <= 1e-4 * float(dens_e[0])


def test_evaluate_points_dense_semantics_matches_the_reference_dense_port():
    """semantics="dense" == the reference's dense PyTorch arithmetic (estimate_gaussian_pdf / estimate_rho_w_no_occlusion,
    gaussian_model.py:253-294,346-364: scale = exp(exp(s) mod), R not R^T) -- checked against oracle/dense_torch.py, which
    tests/test_ref_python.py pins to the reference's own output."""
    from oracle import dense_torch
    N, Na = 300, 2000
    g = torch.Generator().manual_seed(4)
    means = torch.tensor([0.0, 0.75, 0.0]) + (torch.rand(N, 3, generator=g) - 0.5) * 0.4
    ls = torch.log(torch.log(1.0 + 0.02 + 0.06 * torch.rand(N, 3, generator=g)))       # exp(exp(ls)) - 1 ~ 2-8 cm ... in the dense semantics s ~ 1.02-1.08
    rots = torch.randn(N, 4, generator=g)
    opac = torch.randn(N, 1, generator=g)
    fdc = (0.2 * torch.rand(N, 1, generator=g) - 0.5) / 0.28209479177387814
    x = torch.tensor([0.0, 0.75, 0.0]) + (torch.rand(Na, 3, generator=g) - 0.5) * 3.0
    rho, dens = _C.evaluate_points(x.cuda(), means.cuda(), ls.cuda(), rots.cuda(), opac.cuda(), fdc.cuda(), semantics="dense", scaling_modifier=0.9)
    # the port evaluates sum_g pdf op alb on a lattice: feed the points as a [Na,1,1] lattice with unit attenuation
    s = torch.exp(torch.exp(ls.double()) * 0.9)
    R = dense_torch.build_rotation(rots.double())
    loc = torch.einsum("gij,gnj->gni", R, x.double()[None] - means.double()[:, None]) / s[:, None, :]
    pdf = torch.exp(-0.5 * (loc * loc).sum(-1))
    op = torch.sigmoid(opac.double().reshape(-1))
    alb = torch.clamp_min(0.28209479177387814 * fdc.double().reshape(-1) + 0.5, 0.0)
    want_d, want_r = (pdf * op[:, None]).sum(0), (pdf * (op * alb)[:, None]).sum(0)
    assert float(want_d.max()) > 1.0
    assert float((dens.cpu().double() - want_d).abs().max()) <= 2e-5 * float(want_d.max())
    assert float((rho.cpu().double() - want_r).abs().max()) <= 2e-5 * float(want_r.max())
    # and the two semantics differ: same inputs, CUDA pdf
    rho_c, _ = _C.evaluate_points(x.cuda(), means.cuda(), ls.cuda(), rots.cuda(), opac.cuda(), fdc.cuda(), scaling_modifier=0.9)
    assert float((rho_c - rho).abs().max()) > 0.1 * float(rho.max())
