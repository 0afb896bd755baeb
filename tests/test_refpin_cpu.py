"""CPU pins of the two truths the GPU parity tests lean on where the reference defines nothing (its backward is zeros, its analytic kernel
has no time resolution), against the reference's OWN kernel source compiled for the host (oracle/_ref):
  * the oracle's analytic-exact closed form == 64-point fp64 quadrature, per bin, of the reference's numerical kernel (its pdf);
  * the gradient truth (torch.autograd through tests/ref_torch.py, fp64) == central finite differences of the reference-source forward with
    frozen ray membership, all five parameter tensors, both integration modes.
The same comparisons with the CUDA path on the other side are tests/test_gpu_refpin.py."""
import numpy as np
import pytest
import torch

import oracle.ref as oref
from tests import ref_torch, util
from tests.test_gpu_refpin import FD_STEPS, _ref_histogram

pytestmark = pytest.mark.skipif(not oref.available(), reason="oracle/_ref not built and /root/reference absent")


@pytest.mark.parametrize("scale_mul", [0.5, 1.0, 3.0])
def test_closed_form_is_the_bin_integral_of_the_reference_pdf(scale_mul):
    N, Sy, Sx, T, A = 40, 2, 3, 48, 6
    cg, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=21, scale_mul=scale_mul)
    x, w = np.polynomial.legendre.leggauss(64)
    rk = ta["t_samples"].double().numpy()
    nodes = (rk[:, None] + 0.5 * ta["dt"] * x[None, :]).reshape(-1)
    Hn = _ref_histogram(cg, ta, [util.np32(p) for p in params], nodes.astype(np.float32))
    I = (Hn.reshape(cg.S, T, 64) * (0.5 * ta["dt"] * w)[None, None, :]).sum(-1) / (ta["c"] * ta["deltaT"])
    Hq = I / (rk[None, :] ** 2 + 1e-8)
    Ho = util.oracle_H(cg, ta, params, mode=1, precision="f64")
    assert Hq.max() > 0 and util.rel_to_max(Ho, Hq) < 5e-6


@pytest.mark.parametrize("mode,m", [("numerical", 0), ("analytic", 1)])
def test_autograd_truth_matches_finite_differences_of_the_reference_forward(mode, m):
    N, Sy, Sx, T, A = 8, 2, 2, 32, 6
    cg, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=33, scale_mul=0.8)
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(8), dtype=torch.float64).reshape(-1, T)
    rk = ta["t_samples"].double().numpy()
    masks = torch.from_numpy(util.ray_masks(cg, params))
    pd = [p.detach().double().requires_grad_() for p in params]
    dd = lambda x: x.double()
    Ht = ref_torch.transient_batch(dd(cg.scan_o), dd(cg.sin_th), dd(cg.cos_th), dd(cg.sin_ph), dd(cg.cos_ph), dd(ta["t_samples"]), dd(cg.hscale),
                                   masks, *pd, ta["c"], ta["deltaT"], mode=m)
    (Ht * gH).sum().backward()
    if mode == "numerical":
        t_eval, fold = ta["t_samples"].numpy(), None
    else:
        x, w = np.polynomial.legendre.leggauss(32)
        t_eval = (rk[:, None] + 0.5 * ta["dt"] * x[None, :]).reshape(-1).astype(np.float32)
        fold = 0.5 * ta["dt"] * w / (ta["c"] * ta["deltaT"])
    base = [util.np32(p).astype(np.float64) for p in params]
    _, filters = _ref_histogram(cg, ta, [b.astype(np.float32) for b in base], t_eval, want_filters=True)

    def loss(pp):
        H = _ref_histogram(cg, ta, [q.astype(np.float32) for q in pp], t_eval, filters=filters)
        if fold is not None:
            H = (H.reshape(cg.S, T, -1) * fold[None, None, :]).sum(-1)
        return float((H / (rk[None, :] ** 2 + 1e-8) * gH.numpy()).sum())

    assert abs(loss(base) - float((Ht.detach() * gH).sum())) <= 2e-6 * abs(loss(base))
    for ti, h in enumerate(FD_STEPS):
        g_fd = np.zeros_like(base[ti])
        it = np.nditer(base[ti], flags=["multi_index"])
        for _ in it:
            i = it.multi_index
            pp = [b.copy() for b in base]; pm = [b.copy() for b in base]
            pp[ti][i] += h; pm[ti][i] -= h
            hp = float(pp[ti].astype(np.float32)[i]) - float(pm[ti].astype(np.float32)[i])
            g_fd[i] = (loss(pp) - loss(pm)) / hp
        g = pd[ti].grad.numpy().reshape(base[ti].shape)
        assert np.abs(g - g_fd).max() <= 2e-3 * np.abs(g_fd).max(), (mode, ti, np.abs(g - g_fd).max() / np.abs(g_fd).max())
