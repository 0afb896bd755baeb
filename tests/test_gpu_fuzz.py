"""Randomised parity sweep of the batched transient path against the CPU oracle: odd grid shapes (ragged tiles, single rows), lattice sizes
from 1 to 13, bin counts that are no multiple of anything, narrow to very wide Gaussians, every window / bucket / slice decomposition the
host accepts.  Index lists bit-exact, histograms and gradients within 1e-4 of the tensor maximum (north_star)."""
import os

import numpy as np
import pytest
import torch

import oracle
from nlos_gaussian_renderer_b200.transient import TransientRenderer, render_transient_batch
from tests import ref_torch, util
from tests.test_gpu_transient import scenes_cpu

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _case(seed):
    rng = np.random.default_rng(seed)
    N = int(rng.integers(1, 400))
    Sy, Sx = int(rng.integers(1, 10)), int(rng.integers(1, 13))
    T = int(rng.integers(8, 300))
    A = int(rng.integers(1, 14))
    scale_mul = float(np.exp(rng.uniform(np.log(0.15), np.log(8.0))))
    bucket = int(rng.choice([4, 8, 16, 32]))
    while -(-T // bucket) > 64:                               # at most 64 buckets per tile (binning.cu BIN_MAX_RANGES)
        bucket *= 2
    window = int(8 * rng.integers(4, 15))                     # the host accepts multiples of 8 >= 32 with 2 * bucket <= window (shared memory caps it near 120)
    window = max(window, 2 * bucket)
    slice_len = int(rng.choice([1, 2, 3, 4, 7, 16, 64, 4096]))
    return N, Sy, Sx, T, A, scale_mul, window, bucket, slice_len


@pytest.mark.parametrize("seed", list(range(int(os.environ.get("NLOS_FUZZ_SEEDS", "24")))))     # NLOS_FUZZ_SEEDS=400 for a soak run
def test_random_scene_forward_and_lists(seed):
    N, Sy, Sx, T, A, scale_mul, window, bucket, slice_len = _case(seed)
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=100 + seed, scale_mul=scale_mul, device="cuda")
    r = None
    for mode, m in (("numerical", 0), ("analytic", 1)):
        kw = dict(mode=mode, window=window, bucket=bucket, slice_len=slice_len, one_pass=True)
        r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], **kw)
        H = r.forward(*params).cpu().numpy().reshape(Sy * Sx, T)
        H64 = util.oracle_H(geom, ta, params, mode=m, precision="f64")
        assert np.isfinite(H).all()
        if H64.max() == 0:
            assert np.abs(H).max() == 0
            continue
        if util.rel_to_max(H, H64) < TOL:
            continue
        # Not within 1e-4 of the maximum: legitimate only in a scene whose histogram is made of far tails (A = 1 or 2 rays that miss every
        # Gaussian's core), where the 5.3-sigma radial cut of SURVEY 8(d) -- samples with pdf < exp(-5.3^2/2) = 7.9e-7 are dropped -- is
        # visible against a maximum that is itself a tail.  Then (i) the error obeys the rigorous bound of the cut and (ii) parity holds
        # once the cut is moved out (mcut = 8: 1.3e-14).
        case = (mode, N, Sy, Sx, T, A, scale_mul, window, bucket, slice_len)
        op = torch.sigmoid(params[3].reshape(-1))
        alb = torch.clamp_min(ref_torch.C0 * params[4].reshape(N, -1)[:, 0] + 0.5, 0.0)
        bound = float(np.exp(-0.5 * 5.3 ** 2) * (op * alb).sum() * geom.hscale.max() * A * A / float(ta["t_samples"].min()) ** 2)
        assert np.abs(H - H64).max() <= 1.05 * bound + TOL * H64.max(), case
        H8 = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mcut=8.0, **kw).forward(*params).cpu().numpy().reshape(Sy * Sx, T)
        assert util.rel_to_max(H8, H64) < TOL, case
    cnt, off, lst = r.lists()
    (ocnt, ooff, olst), _ = util.oracle_window_lists(geom, ta, params, r)
    assert np.array_equal(cnt, ocnt) and np.array_equal(off, ooff) and np.array_equal(lst, olst)


@pytest.mark.parametrize("seed", [1, 4, 9, 13, 17, 22] + list(range(1000, 1000 + int(os.environ.get("NLOS_FUZZ_BWD_SEEDS", "0")))))
@pytest.mark.parametrize("mode", ["numerical", "analytic"])
def test_random_scene_backward(seed, mode):
    N, Sy, Sx, T, A, scale_mul, window, bucket, slice_len = _case(seed)
    N, T, Sy, Sx = min(N, 120), min(T, 96), min(Sy, 5), min(Sx, 8)             # the fp64 autograd truth is O(N S A^2 T)
    scale_mul = max(scale_mul, 0.6)
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=200 + seed, scale_mul=scale_mul, device="cuda")
    m = 0 if mode == "numerical" else 1
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(seed))
    ps = [p.clone().requires_grad_() for p in params]
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, window=window, bucket=bucket, slice_len=slice_len, one_pass=True)
    H = render_transient_batch(r, *ps)
    (H * gH.cuda()).sum().backward()
    cg = scenes_cpu(geom)
    masks = torch.from_numpy(util.ray_masks(cg, [p.cpu() for p in params]))
    pd = [p.detach().cpu().double().requires_grad_() for p in params]
    dd = lambda x: x.cpu().double()
    Ht = ref_torch.transient_batch(dd(geom.scan_o), dd(geom.sin_th), dd(geom.cos_th), dd(geom.sin_ph), dd(geom.cos_ph),
                                   dd(ta["t_samples"]), dd(geom.hscale), masks, *pd, ta["c"], ta["deltaT"], mode=m)
    (Ht * gH.reshape(Sy * Sx, T).double()).sum().backward()
    if float(Ht.max()) == 0:
        return
    assert util.rel_to_max(H.detach().cpu().numpy().reshape(-1, T), Ht.detach().numpy()) < TOL
    for name, a, b in zip(("means", "log_scales", "rotations", "opacities", "features"), ps, pd):
        if float(b.grad.abs().max()) > 0:
            assert util.rel_to_max(a.grad.cpu().numpy(), b.grad.numpy()) < TOL, (name, N, Sy, Sx, T, A, scale_mul, window, bucket, slice_len)


def test_invalid_decompositions_raise():
    geom, ta, params = util.small_scene(20, 2, 2, 32, 3, device="cuda")
    for window, bucket in [(16, 4), (36, 4), (32, 32), (64, 6)]:
        with pytest.raises(RuntimeError):
            TransientRenderer(geom, 32, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], window=window, bucket=bucket).forward(*params)
