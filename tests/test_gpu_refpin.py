"""Pins of the parts of the path the reference does NOT define, against the reference's own source compiled for the host (oracle/_ref):

  * analytic-exact mode  ==  fine fp64 Gauss-Legendre quadrature, over every bin, of the reference's numerical kernel evaluated at the quadrature
    nodes (volume_renderer.cu:139-182 -> eval_gaussian_pdf, include/cuda_utils.cuh:124-151): the closed form of the oracle and the GPU's
    moment-matched mixtures both have to agree with the integral of the reference's own pdf;
  * gradients  ==  central finite differences of the reference-source forward with FROZEN ray membership, for means, scales,
    rotations, opacities and features on a 8-Gaussian scene;
  * the timed configurations at their REAL shapes: 8 - 16 scan points of BASELINE configs[1..3] and of a 16-row slab of the north-star scene
    (tile corners, grid corners, last tile) against the fp64 oracle, both integration modes.
"""
import numpy as np
import pytest
import torch

import oracle
import oracle.ref as oref
from nlos_gaussian_renderer_b200 import scenes
from nlos_gaussian_renderer_b200.geometry import ScanGeometry
from nlos_gaussian_renderer_b200.transient import TransientRenderer, render_transient_batch
from tests import util

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not oref.available(), reason="oracle/_ref (the reference's kernels compiled for the host) not built")
TOL = 1e-4
FD_STEPS = [3e-5, 2e-3, 2e-3, 1e-2, 1e-2]          # means [m], log-scales, quaternion components, opacity logits, features


def _rays(geom_cpu, p):
    st, ct, sp, cp = [util.np32(x)[p] for x in (geom_cpu.sin_th, geom_cpu.cos_th, geom_cpu.sin_ph, geom_cpu.cos_ph)]
    A = st.shape[0]
    d = np.stack([(st[:, None] * cp[None, :]).reshape(-1), (st[:, None] * sp[None, :]).reshape(-1), np.repeat(ct, A)], 1).astype(np.float32)
    o = np.repeat(util.np32(geom_cpu.scan_o)[p][None], A * A, 0)
    return o, d, st


def _ref_histogram(geom_cpu, ta, params_np, t_eval, weights=None, filters=None, want_filters=False):
    """H[p, k] built from the reference-source numerical kernel (no occlusion): rho[ray, s] at the radii t_eval, then the post-processing of
    cuda_autograd.py:302-314 in fp64.  weights [len(t_eval)] folds quadrature nodes into bins (None: one sample per bin)."""
    S, A = geom_cpu.S, geom_cpu.A
    hs = geom_cpu.hscale.double().numpy()
    out, fl = [], []
    for p in range(S):
        o, d, st = _rays(geom_cpu, p)
        r = oref.render_rays(o, d, t_eval, *params_np, util.np32(geom_cpu.scan_o)[p], 0, ta["c"], ta["deltaT"], 1.0, False,
                             return_filter=want_filters, filter_lists=None if filters is None else filters[p])
        if want_filters:
            fl.append(r[3])
        rho = r[0].astype(np.float64)                                           # [A*A, n_eval]
        out.append((rho * np.repeat(st.astype(np.float64), A)[:, None]).sum(0) * hs[p])
    H = np.stack(out)
    return (H, fl) if want_filters else H


@needs_ref
@pytest.mark.parametrize("scale_mul", [0.5, 1.0, 3.0])
def test_analytic_exact_is_the_bin_integral_of_the_reference_pdf(scale_mul):
    N, Sy, Sx, T, A = 40, 2, 3, 48, 6
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=21, scale_mul=scale_mul, device="cuda")
    cg = ScanGeometry.wall_grid(Sy, Sx, A=A)
    pn = [util.np32(p) for p in params]
    # 64-point Gauss-Legendre nodes inside every bin [r_k - dt/2, r_k + dt/2]
    x, w = np.polynomial.legendre.leggauss(64)
    rk = ta["t_samples"].double().numpy()
    nodes = (rk[:, None] + 0.5 * ta["dt"] * x[None, :]).reshape(-1)
    Hn = _ref_histogram(cg, ta, pn, nodes.astype(np.float32))                   # [S, T*64], includes the kernel's factor c*deltaT
    # the float32 rounding of the node radii moves a node by <= 6e-8 relative: evaluate the weights at the rounded nodes' own positions
    I = (Hn.reshape(cg.S, T, 64) * (0.5 * ta["dt"] * w)[None, None, :]).sum(-1) / (ta["c"] * ta["deltaT"])
    Hq = I / (rk[None, :] ** 2 + 1e-8)                                          # attenuation at the bin centre (cuda_autograd.py:308)
    assert Hq.max() > 0
    Ho = util.oracle_H(cg, ta, params, mode=1, precision="f64")                 # the oracle's closed form
    e_o = util.rel_to_max(Ho, Hq)
    assert e_o < 2e-5
    for one_pass in (False, True):                                              # the mixture plan and the one-pass (1-node rows + 7-tap filter) plan
        Hg = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode="analytic", one_pass=one_pass).forward(*params)
        e_g = util.rel_to_max(Hg.cpu().numpy().reshape(cg.S, T), Hq)
        print(f"scale x{scale_mul}: closed form vs quadrature of the reference pdf {e_o:.2e}; GPU (one_pass={one_pass}) vs quadrature {e_g:.2e}")
        assert e_g < TOL


@needs_ref
@pytest.mark.parametrize("mode", ["numerical", "analytic"])
def test_gradients_match_finite_differences_of_the_reference_forward(mode):
    """L = sum H * gH.  Truth: central differences of the reference-source forward, per-ray lists frozen at the base point.  For the analytic
    mode the forward is the bin integral of the reference's pdf (64-point quadrature as above)."""
    N, Sy, Sx, T, A = 8, 2, 2, 32, 6
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=33, scale_mul=0.8, device="cuda")
    cg = ScanGeometry.wall_grid(Sy, Sx, A=A)
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(8), dtype=torch.float64)
    rk = ta["t_samples"].double().numpy()
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
        H = H / (rk[None, :] ** 2 + 1e-8)
        return float((H * gH.reshape(cg.S, T).numpy()).sum())

    # GPU gradients
    ps = [p.clone().requires_grad_() for p in params]
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, one_pass=True)
    (render_transient_batch(r, *ps) * gH.float().cuda()).sum().backward()
    # the reference's per-ray lists at the base point == the membership the GPU used (cap not binding: 8 Gaussians)
    steps = FD_STEPS
    names = ("means", "log_scales", "rotations", "opacities", "features")
    for ti, (name, h) in enumerate(zip(names, steps)):
        g_gpu = ps[ti].grad.detach().cpu().double().numpy().reshape(base[ti].shape)
        g_fd = np.zeros_like(base[ti])
        it = np.nditer(base[ti], flags=["multi_index"])
        for _ in it:
            i = it.multi_index
            pp = [b.copy() for b in base]; pm = [b.copy() for b in base]
            pp[ti][i] += h; pm[ti][i] -= h
            # (the forward rounds the parameters to fp32: use the step that was actually taken)
            hp = float(pp[ti].astype(np.float32)[i]) - float(pm[ti].astype(np.float32)[i])
            g_fd[i] = (loss(pp) - loss(pm)) / hp
        err = np.abs(g_gpu - g_fd).max() / np.abs(g_fd).max()
        print(f"{mode} d/d{name}: GPU vs finite differences of the reference forward {err:.2e} (max |g| {np.abs(g_fd).max():.3e})")
        assert err < 3e-3, name


# ----------------------------------------------------------------------------------------------------------------------------------------
def _sample_points(geom):
    """flat scan indices: the four grid corners, the four corners of the first and of the last tile, an interior tile corner."""
    Sy, Sx, ty, tx = geom.Sy, geom.Sx, geom.tile_y, geom.tile_x
    pts = {(0, 0), (0, Sx - 1), (Sy - 1, 0), (Sy - 1, Sx - 1),
           (0, tx - 1), (ty - 1, 0), (ty - 1, tx - 1), (ty, tx),
           (Sy - ty, Sx - tx), (Sy - 1, Sx - tx), (Sy - ty, Sx - 1),
           (Sy // 2, Sx // 2), (Sy // 2 - 1, Sx // 2 - 1), (Sy // 3, 2 * Sx // 3), (2 * Sy // 3, Sx // 3), (Sy // 2, 0)}
    return sorted(y * Sx + x for y, x in pts if 0 <= y < Sy and 0 <= x < Sx)


@pytest.mark.parametrize("cfg,rows,n_pts", [("cfg2", None, 16), ("cfg3", None, 12), ("cfg4", None, 8), ("north_star", 16, 8)])
def test_sampled_scan_points_of_the_timed_configurations_match_the_oracle(cfg, rows, n_pts):
    """The GPU renders the WHOLE configuration (A = 32, real Gaussian counts: footprints of ~1 - 9 rays, the class mix of the real scenes);
    the fp64 oracle renders a sample of its scan points.  Both integration modes, default (auto) kernel choice."""
    N, Sy, Sx, T, A = scenes.CONFIGS[cfg]
    full = ScanGeometry.wall_grid(Sy, Sx, A=A)
    if rows:
        full = full.select_rows(range(rows))
    geom = ScanGeometry(full.scan_o.reshape(full.Sy, Sx, 3), full.theta_range.reshape(full.Sy, Sx, 2), full.phi_range.reshape(full.Sy, Sx, 2), A,
                        vp_y=full.vp_y, device="cuda")
    ta = scenes.time_axis(T)
    seed = 1000 + list(scenes.CONFIGS).index(cfg) + 1
    params = scenes.make_gaussians(N, seed)
    pts = _sample_points(full)
    pts = pts[:: max(1, len(pts) // n_pts)][:n_pts]
    f = util.np32
    sub = lambda x: f(x)[pts]
    oracle.set_num_threads(__import__("os").cpu_count() or 1)
    for mode, m in (("numerical", 0), ("analytic", 1)):
        H = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, one_pass=True).forward(*[p.cuda() for p in params])
        H = H.reshape(full.S, T)[pts].cpu().numpy()
        torch.cuda.empty_cache()
        H64 = oracle.transient_batch(sub(full.scan_o), sub(full.sin_th), sub(full.cos_th), sub(full.sin_ph), sub(full.cos_ph), f(ta["t_samples"]),
                                     sub(full.hscale), *[f(p) for p in params], sh_degree=0, c=ta["c"], deltaT=ta["deltaT"], mode=m, precision="f64")
        worst = 0.0
        for i in range(len(pts)):
            e = np.abs(H[i] - H64[i]).max() / H64[i].max()
            worst = max(worst, e)
            assert e <= TOL, (cfg, mode, pts[i], e)
            big = H64[i] > 1e-3 * H64[i].max()          # (rtol on the large bins + the bound of the default 5.3-sigma cut, see test_forward_matches_oracle)
            assert (np.abs(H[i] - H64[i])[big] <= TOL * H64[i][big] + (1.6e-6 if m == 0 else 5e-6) * H64[i].max()).all(), (cfg, mode, pts[i])
        print(f"{cfg} {mode}: {len(pts)} scan points, worst rel-to-max error {worst:.2e}")
