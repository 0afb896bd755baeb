"""View-dependent albedo (active_sh_degree 1..3): albedo(g, p) = max(SH(dir(mu_g - o_p)) + 0.5, 0) per (Gaussian, scan point)
(volume_renderer.cu:107-111, spherical_harmonics.cuh:20-80; configs/default.py:31 sets sh_degree = 3 and main.py:240-241 raises the
active degree during training) in the batched renderer and in BOTH backward passes (the reference's backward is zeros)."""
import numpy as np
import pytest
import torch

from nlos_gaussian_renderer_b200.dropin.cuda_autograd import CUDARenderFunction
from nlos_gaussian_renderer_b200.dropin.lattice import ray_lattice
from nlos_gaussian_renderer_b200.transient import TransientRenderer, render_transient_batch
from tests import ref_torch, util
from tests.test_gpu_transient import scenes_cpu

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _scene(N, Sy, Sx, T, A, deg, scale_mul=2.0, seed=3):
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=seed, scale_mul=scale_mul, device="cuda")
    g = torch.Generator().manual_seed(100 + deg)
    feats = torch.randn(N, 16, generator=g) * 0.4
    feats[:, 0] = params[4].cpu()[:, 0]
    params = list(params[:4]) + [feats.cuda()]
    return geom, ta, params


@pytest.mark.parametrize("variant", ["wide", "narrow", "auto"])
@pytest.mark.parametrize("mode", ["numerical", "analytic"])
@pytest.mark.parametrize("deg", [1, 2, 3])
def test_batched_forward_matches_oracle(deg, mode, variant):
    N, Sy, Sx, T, A = 200, 5, 9, 48, 6
    geom, ta, params = _scene(N, Sy, Sx, T, A, deg)
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, variant=variant, active_sh_degree=deg, one_pass=True)
    H = r.forward(*params).cpu().numpy().reshape(Sy * Sx, T)
    H64 = util.oracle_H(scenes_cpu(geom), ta, [p.cpu() for p in params], mode=0 if mode == "numerical" else 1, precision="f64", sh_degree=deg)
    H0 = util.oracle_H(scenes_cpu(geom), ta, [p.cpu() for p in params], mode=0 if mode == "numerical" else 1, precision="f64", sh_degree=0)
    assert util.rel_to_max(H0, H64) > 1e-2            # the view dependence matters in this scene
    assert util.rel_to_max(H, H64) < TOL


@pytest.mark.parametrize("variant", ["wide", "narrow"])
@pytest.mark.parametrize("mode", ["numerical", "analytic"])
@pytest.mark.parametrize("deg", [1, 3])
def test_batched_backward_matches_autograd(deg, mode, variant):
    N, Sy, Sx, T, A = 80, 4, 8, 40, 5
    geom, ta, params = _scene(N, Sy, Sx, T, A, deg)
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(5))
    ps = [p.clone().requires_grad_() for p in params]
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, variant=variant, active_sh_degree=deg, one_pass=True)
    H = render_transient_batch(r, *ps)
    (H * gH.cuda()).sum().backward()
    cg = scenes_cpu(geom)
    masks = torch.from_numpy(util.ray_masks(cg, [p.cpu() for p in params]))
    pd = [p.detach().cpu().double().requires_grad_() for p in params]
    dd = lambda x: x.cpu().double()
    Ht = ref_torch.transient_batch(dd(geom.scan_o), dd(geom.sin_th), dd(geom.cos_th), dd(geom.sin_ph), dd(geom.cos_ph), dd(ta["t_samples"]),
                                   dd(geom.hscale), masks, *pd, ta["c"], ta["deltaT"], mode=0 if mode == "numerical" else 1, sh_degree=deg)
    (Ht * gH.reshape(Sy * Sx, T).double()).sum().backward()
    assert util.rel_to_max(H.detach().cpu().numpy().reshape(-1, T), Ht.detach().numpy()) < TOL
    K = (deg + 1) ** 2
    for name, a, b in zip(("means", "log_scales", "rotations", "opacities", "features"), ps, pd):
        ga, gb = a.grad.cpu().numpy(), b.grad.numpy()
        assert util.rel_to_max(ga, gb) < TOL, name
        if name == "features":
            assert np.abs(gb[:, 1:K]).max() > 0 and (K == 16 or np.abs(ga[:, K:]).max() == 0)


@pytest.mark.parametrize("deg", [1, 2, 3])
def test_batched_slice_equals_per_scan_point_operator(deg):
    """One scan point of the batched transient == the drop-in one-scan-point operator (CUDARenderFunction) with the same degree and camera,
    values and gradients."""
    N, Sy, Sx, T, A = 150, 2, 3, 40, 6
    geom, ta, params = _scene(N, Sy, Sx, T, A, deg, scale_mul=1.5)
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode="numerical", active_sh_degree=deg)
    ps = [p.clone().requires_grad_() for p in params]
    H = render_transient_batch(r, *ps)
    w = torch.linspace(0.5, 1.5, T, device="cuda")
    p_idx = (1, 2)
    (H[p_idx] * w).sum().backward()
    flat = p_idx[0] * Sx + p_idx[1]
    cam = geom.scan_o[flat].contiguous()
    st, ct, sp, cp = geom.sin_th[flat], geom.cos_th[flat], geom.sin_ph[flat], geom.cos_ph[flat]
    d = torch.stack([(st[:, None] * cp[None, :]).reshape(-1), (st[:, None] * sp[None, :]).reshape(-1), ct[:, None].expand(A, A).reshape(-1)], 1).contiguous()
    o = cam[None].expand(A * A, 3).contiguous()
    qs = [p.clone().requires_grad_() for p in params]
    rho, _, _ = CUDARenderFunction.apply(o, d, ta["t_samples"].cuda(), qs[0], qs[1], qs[2], qs[3], qs[4], cam, deg, ta["c"], ta["deltaT"], 1.0, False, "netf")
    Hp = (rho * st[:, None].expand(A, A).reshape(-1, 1)).sum(0) / (ta["t_samples"].cuda() ** 2 + 1e-8) * geom.hscale[flat]
    (Hp * w).sum().backward()
    assert util.rel_to_max(H[p_idx].detach().cpu().numpy(), Hp.detach().cpu().numpy()) < 2e-5
    for a, b in zip(ps, qs):
        assert util.rel_to_max(a.grad.cpu().numpy(), b.grad.cpu().numpy()) < 5e-5


@pytest.mark.parametrize("occ", [False, True])
@pytest.mark.parametrize("deg", [1, 2, 3])
def test_one_scan_point_backward_with_sh_matches_autograd(deg, occ):
    """CUDARenderFunction.backward at active_sh_degree >= 1 (round 1 raised here): fp64 autograd through the restatement with the view-dependent albedo."""
    N, T, A = 60, 32, 5
    geom, ta, params = _scene(N, 1, 1, T, A, deg, scale_mul=2.5)
    cam = geom.scan_o[0].contiguous()
    st, ct, sp, cp = geom.sin_th[0], geom.cos_th[0], geom.sin_ph[0], geom.cos_ph[0]
    d = torch.stack([(st[:, None] * cp[None, :]).reshape(-1), (st[:, None] * sp[None, :]).reshape(-1), ct[:, None].expand(A, A).reshape(-1)], 1).contiguous()
    o = cam[None].expand(A * A, 3).contiguous()
    ts = ta["t_samples"].cuda()
    g = torch.rand(A * A, T, generator=torch.Generator().manual_seed(2)).cuda()
    qs = [p.clone().requires_grad_() for p in params]
    rho, _, _ = CUDARenderFunction.apply(o, d, ts, qs[0], qs[1], qs[2], qs[3], qs[4], cam, deg, ta["c"], 0.5 if occ else ta["deltaT"], 1.0, occ, "netf")
    (rho * g).sum().backward()
    # truth
    import tests.ref_torch as R
    pd = [p.detach().cpu().double().requires_grad_() for p in params]
    cg = scenes_cpu(geom)
    masks = torch.from_numpy(util.ray_masks(cg, [p.cpu() for p in params]))[0]
    rho_t, _, _ = R.render_rays(o.cpu().double(), d.cpu().double(), ts.cpu().double(), masks, *pd, ta["c"], 0.5 if occ else ta["deltaT"],
                                use_occlusion=occ, sh_degree=deg, cam=cam.cpu().double())
    (rho_t * g.cpu().double()).sum().backward()
    assert util.rel_to_max(rho.detach().cpu().numpy(), rho_t.detach().numpy()) < TOL
    for name, a, b in zip(("means", "log_scales", "rotations", "opacities", "features"), qs, pd):
        assert util.rel_to_max(a.grad.cpu().numpy(), b.grad.numpy()) < TOL, name
