"""Pins against the reference's OWN Python, imported from /root/reference with the two shims of SURVEY.md 0.7
(tests/golden/ref_python_shims.py) -- live here, through the committed fixture tests/golden/ref_python.npz on the GPU box:
  * the lattice of ScanGeometry.from_volume == nlos_helpers.spherical_sample_histogram:124-188 (row f1);
  * oracle/dense_torch.py (the CPU baseline bench.py times) == nlos_helpers.gaussian_transient_rendering:206-232 ->
    GaussianModel.estimate_rho_w_no_occlusion, values and gradients (row a12);
  * the reference's unmodified gaussian_model/cuda_autograd.py / rendering_cuda.py / nlos_helpers.py bind to THIS repo's
    `nlos_gaussian_renderer._C` when the repo root is on the path (row a10; the GPU half is tests/test_gpu_reference_python.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from nlos_gaussian_renderer_b200.geometry import ScanGeometry
from oracle import dense_torch
from tests.golden import make_ref_python_golden as gen
from tests.golden import ref_python_shims as shims

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(ROOT, "tests", "golden", "ref_python.npz"))


@pytest.mark.skipif(not shims.available(), reason="/root/reference not present (GPU box): the committed fixture stands in")
def test_fixture_is_what_the_reference_produces_here():
    live = gen.make()
    assert set(live) == set(G.files)
    for k in G.files:
        assert np.allclose(live[k], G[k], rtol=1e-6, atol=1e-9), k


@pytest.mark.parametrize("n", range(len(gen.cases())))
def test_lattice_matches_spherical_sample_histogram(n):
    pos, vp, vs, A, start, end, c, dT = gen.cases()[n]
    geom = ScanGeometry.from_volume(torch.tensor([[pos]], dtype=torch.float32), vp, vs, A)
    I1, I2, num_r, dth, dph, tmin, tmax, pmin, pmax = G[f"lat{n}_scalars"]
    assert int(num_r) == end - start and int(I1) == start and int(I2) == end
    # ranges: the reference takes .item() of fp32 min / max over the 8 box corners
    assert float(geom.theta_range[0, 0]) == np.float32(tmin) and float(geom.theta_range[0, 1]) == np.float32(tmax)
    assert float(geom.phi_range[0, 0]) == np.float32(pmin) and float(geom.phi_range[0, 1]) == np.float32(pmax)
    # linspaces, bit for bit
    assert np.array_equal(geom.theta[0].numpy(), G[f"lat{n}_theta"])
    assert np.array_equal(geom.phi[0].numpy(), G[f"lat{n}_phi"])
    # quadrature weight dtheta * dphi * vp_y^2 with the reference's /A (nlos_helpers.py:162-163, 226-231)
    assert abs(float(geom.hscale[0]) - dth * dph * vp[1] ** 2) <= 2e-7 * dth * dph * vp[1] ** 2
    # radii: torch.linspace(start c dT, end c dT, num_r) (nlos_helpers.py:165-172), as scenes.time_axis builds them
    r = torch.linspace(start * c * dT, end * c * dT, end - start, dtype=torch.float32).numpy()
    assert np.allclose(r, G[f"lat{n}_r"], rtol=3e-6)
    # ray directions: (sin th cos ph, sin th sin ph, cos th) * r + o == the reference's lattice points (first and last shell)
    d = np.stack([geom.sin_th[0].numpy()[:, None] * geom.cos_ph[0].numpy()[None, :], geom.sin_th[0].numpy()[:, None] * geom.sin_ph[0].numpy()[None, :],
                  np.repeat(geom.cos_th[0].numpy()[:, None], A, 1)], -1)
    for k, rk in ((0, r[0]), (1, r[-1])):
        assert np.allclose(d * rk + np.array(pos, np.float32), G[f"lat{n}_xyz_first_last"][k], atol=3e-6)


def test_dense_port_matches_the_reference_dense_path():
    N, A, start, end = [int(v) for v in G["dense_shape"]]
    c, dT, vs, vx, vy, vz, px, py, pz = [float(v) for v in G["dense_scalars"]]
    flat = torch.from_numpy(G["dense_params"].astype(np.float32))
    sizes = [3 * N, 3 * N, 4 * N, N, N]
    means, ls, rots, opac, fdc = [t.clone().requires_grad_() for t in torch.split(flat, sizes)]
    geom = ScanGeometry.from_volume(torch.tensor([[[px, py, pz]]]), (vx, vy, vz), vs, A)
    T = end - start
    r = torch.linspace(start * c * dT, end * c * dT, T, dtype=torch.float32)
    th, ph = geom.theta[0], geom.phi[0]
    pts = torch.stack([r[:, None, None] * torch.sin(th)[None, :, None] * torch.cos(ph)[None, None, :],
                       r[:, None, None] * torch.sin(th)[None, :, None] * torch.sin(ph)[None, None, :],
                       (r[:, None, None] * torch.cos(th)[None, :, None]).expand(T, A, A)], -1) + torch.tensor([px, py, pz])
    dth = float(geom.theta_range[0, 1] - geom.theta_range[0, 0]) / A
    dph = float(geom.phi_range[0, 1] - geom.phi_range[0, 0]) / A
    hist = dense_torch.dense_histogram(pts, th, r, means.reshape(N, 3), ls.reshape(N, 3), rots.reshape(N, 4), opac, fdc, dth, dph, vy)
    w = torch.linspace(0.5, 1.5, T)
    (hist * w).sum().backward()
    ref = G["dense_hist"]
    assert np.abs(hist.detach().numpy() - ref).max() <= 2e-5 * np.abs(ref).max()
    for name, t, shape in (("mu", means, (N, 3)), ("scaling", ls, (N, 3)), ("rotation", rots, (N, 4)), ("opacity", opac, (N, 1)), ("fdc", fdc, (N, 1))):
        g = G["dense_grad_" + name]
        assert np.abs(t.grad.numpy().reshape(shape) - g).max() <= 2e-4 * np.abs(g).max(), name


@pytest.mark.skipif(not shims.available(), reason="/root/reference not present (GPU box)")
def test_reference_python_binds_to_this_repos_extension():
    """Unmodified reference files + this repo's root on the path: `from nlos_gaussian_renderer import _C` resolves to our C-ABI shim,
    rendering_cuda.create_cuda_renderer() returns a renderer and nlos_helpers builds its global CUDA_RENDERER from it
    (nlos_helpers.py:21-27) -- in a fresh interpreter so that nothing of this test session leaks in."""
    code = r"""
import sys
sys.path.insert(0, %r)
from tests.golden import ref_python_shims as shims
ref = shims.import_reference()
NH = ref["nlos_helpers"]
import nlos_gaussian_renderer._C as ours
import gaussian_model.cuda_autograd as ca, gaussian_model.rendering_cuda as rc
assert ca.__file__.startswith(shims.REF) and rc.__file__.startswith(shims.REF), (ca.__file__, rc.__file__)
assert ca.CUDA_AVAILABLE and rc.CUDA_AVAILABLE
assert ca._C is ours and rc._C is ours
assert ours.__file__.startswith(%r)
assert NH.CUDA_RENDERER is not None and type(NH.CUDA_RENDERER).__name__ == "GaussianRendererCUDA"
assert type(NH.CUDA_RENDERER.renderer).__name__ == "CUDARenderModule"
for f in ("render_rays", "filter_gaussians_per_ray", "render_rays_analytic"):
    assert callable(getattr(ours, f))
# the reference's section renderer imports the package `cuda_renderer` (rendering_section.py:12): ours
import gaussian_model.rendering_section as rs
assert rs.CUDA_AVAILABLE and rs.SectionGaussianRendererCUDA.__module__.startswith("nlos_gaussian_renderer_b200")
print("bound")
""" % (ROOT, ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd="/tmp")
    assert out.returncode == 0 and "bound" in out.stdout, out.stderr[-2000:]
