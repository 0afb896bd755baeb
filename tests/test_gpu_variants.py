"""The two scan-tile kernels (transient.cu: register accumulators, transient_narrow.cu: direct to window) and the per-entry choice between
them ("auto": entry classes decided at binning time) must be interchangeable: same hit masks as the reference's per-ray filter
(src/ray_aabb.cu:10-61, include/cuda_utils.cuh:97-121) BIT FOR BIT, same histograms and gradients within the fp32 tolerance."""
import numpy as np
import pytest
import torch

from nlos_gaussian_renderer_b200.transient import TransientRenderer, render_transient_batch
from tests import util

pytestmark = pytest.mark.gpu


def _r(geom, ta, mode, **kw):
    return TransientRenderer(geom, ta["t_samples"].numel(), ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, **{"one_pass": True, **kw})


@pytest.mark.parametrize("variant", ["wide", "narrow", "auto"])
@pytest.mark.parametrize("shape,scale_mul", [((150, 5, 9, 48, 6), 1.0), ((300, 6, 7, 64, 12), 0.4), ((200, 4, 8, 64, 9), 2.5), ((64, 3, 5, 40, 3), 6.0),
                                             ((500, 8, 8, 128, 16), 0.25)])
def test_hit_masks_equal_the_reference_filter_bit_for_bit(variant, shape, scale_mul):
    """Every (scan point, ray, Gaussian) triple the render kernel evaluates == the reference's slab test with cap = N
    (oracle.filter_gaussians, itself pinned bit-exact against the reference source in tests/test_oracle.py)."""
    N, Sy, Sx, T, A = shape
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=scale_mul, device="cuda")
    # mcut = 60: no radial reject and every ray of a hit inside the cut, so the dump holds the bare slab-test decisions
    r = _r(geom, ta, "numerical", variant=variant, mcut=60.0, window=120 if T > 100 else 96)
    got = util.hit_masks_gpu(r, params)
    from tests.test_gpu_transient import scenes_cpu
    want = util.ray_masks(scenes_cpu(geom), [p.cpu() for p in params])
    # Gaussians the time window does not reach at all are not listed (radial cull of the binning): compare where listed
    listed = np.zeros((geom.S, N), bool)
    cnt, off, lst = r.lists()
    perm = geom.perm.cpu().numpy()
    cpt = r.n_classes * r.n_buckets
    for t in range(geom.n_tiles):
        gs = lst[off[t * cpt]: off[(t + 1) * cpt]] & ((1 << 26) - 1)
        pts = perm[t * 32:(t + 1) * 32]
        listed[np.ix_(pts[pts >= 0], gs)] = True
    assert listed.mean() > 0.5
    diff = (got != want) & listed[:, None, :]
    assert want.sum() > 0
    assert diff.sum() == 0, f"{diff.sum()} of {want.sum()} slab decisions differ from the reference"
    assert not (got & ~listed[:, None, :]).any()


@pytest.mark.parametrize("mode", ["numerical", "analytic"])
@pytest.mark.parametrize("scale_mul", [0.3, 1.0, 4.0])
def test_variants_agree(mode, scale_mul):
    N, Sy, Sx, T, A = 400, 6, 10, 160, 10
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=scale_mul, device="cuda")
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(2)).cuda()
    outs = {}
    for variant in ("wide", "narrow", "auto"):
        r = _r(geom, ta, mode, variant=variant, narrow_rays=1.5)
        ps = [p.clone().requires_grad_() for p in params]
        H = render_transient_batch(r, *ps)
        (H * gH).sum().backward()
        outs[variant] = (H.detach().cpu().numpy(), [p.grad.cpu().numpy() for p in ps])
    for v in ("narrow", "auto"):
        assert util.rel_to_max(outs[v][0], outs["wide"][0]) < 2e-5
        for a, b in zip(outs[v][1], outs["wide"][1]):
            if np.abs(b).max() > 0:
                assert util.rel_to_max(a, b) < 3e-5


@pytest.mark.parametrize("scale_mul", [0.5, 1.0, 3.0])
def test_analytic_plans_agree(scale_mul):
    """analytic-exact mode: the mixture plan (one_pass=False: the default below 64 tiles) and the one-pass plan (time-wide entries as 1-node
    rows + the fixed 7-tap filter, merged two-class launches) give the same histograms and gradients; the deterministic mode (one class per
    launch) as well."""
    N, Sy, Sx, T, A = 400, 6, 10, 160, 10
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=scale_mul, device="cuda")
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(3)).cuda()
    outs = []
    for kw in (dict(one_pass=False), dict(one_pass=True), dict(one_pass=True, deterministic=True), dict(one_pass=None)):
        r = _r(geom, ta, "analytic", **kw)
        if kw["one_pass"]:
            assert r.n_classes == 4 and len(r._launch_groups()) == (4 if r.deterministic else 2)
        else:
            assert not r.one_pass and r.n_classes == 2                   # 2 tiles < 64: the plan is off by default
        ps = [p.clone().requires_grad_() for p in params]
        H = render_transient_batch(r, *ps)
        (H * gH).sum().backward()
        outs.append((H.detach().cpu().numpy(), [p.grad.cpu().numpy() for p in ps]))
    for H, g in outs[1:]:
        assert util.rel_to_max(H, outs[0][0]) < 4e-5                     # 2 x the 1.6e-5-of-a-ray's-peak bound of the filter
        for a, b in zip(g, outs[0][1]):
            if np.abs(b).max() > 0:
                assert util.rel_to_max(a, b) < 1e-4


@pytest.mark.parametrize("scale_mul", [0.3, 1.0])
@pytest.mark.parametrize("mode", ["numerical", "analytic"])
def test_culling_words(mode, scale_mul):
    """The backward pass that reads the forward's culling words (narrow kernel: anchor + hit mask of the <= 2 x 2 candidate rays per (entry, scan
    point)) uses exactly the hit masks the forward decided -- and, in the deterministic mode, produces bit-identical gradients to the backward
    that decides again; a backward without a forward of the same prepare() must not read them."""
    N, Sy, Sx, T, A = 300, 6, 10, 96, 12
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=scale_mul, device="cuda")
    r_on = _r(geom, ta, mode, culling_words=True, narrow_rays=2.5)
    fwd_masks = util.hit_masks_gpu(r_on, params)
    assert r_on._stash is not None and r_on._stash_gen == r_on.generation
    words = r_on._stash[: r_on.check_lists() * 32].cpu().numpy().view(np.uint16)
    assert (words != 0xFFFF).any() and ((words >> 12) != 0).any()                  # the words carry hits
    bwd_masks = util.hit_masks_gpu(r_on, params, backward=True)
    assert fwd_masks.sum() > 0 and np.array_equal(fwd_masks, bwd_masks)
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(5)).cuda()
    grads = {}
    for cw in (True, False):
        r = _r(geom, ta, mode, culling_words=cw, deterministic=True)
        assert (r._stash is None) or cw
        ps = [p.clone().requires_grad_() for p in params]
        (render_transient_batch(r, *ps) * gH).sum().backward()
        assert (r._stash is not None) == cw
        grads[cw] = [p.grad.clone() for p in ps]
    for a, b in zip(grads[True], grads[False]):
        assert torch.equal(a, b)
    # prepare() without a forward: the words of the previous scene are stale and must be ignored
    r = _r(geom, ta, mode, culling_words=True, deterministic=True)
    r.forward(*[p[:100].contiguous() for p in params])
    r.prepare(*params)
    assert r._stash_gen != r.generation
    r.render_backward(geom.tile_grad(gH))
    want = _r(geom, ta, mode, culling_words=False, deterministic=True)
    want.prepare(*params)
    want.render_backward(geom.tile_grad(gH))
    assert torch.equal(r.dacc, want.dacc)


def test_backward_after_another_forward_uses_its_own_lists():
    """a second forward() on the same renderer before the first graph's backward() must not corrupt the first gradients."""
    geom, ta, params = util.small_scene(200, 5, 9, 48, 6, device="cuda")
    r = _r(geom, ta, "numerical")
    ps = [p.clone().requires_grad_() for p in params]
    H = render_transient_batch(r, *ps)
    with torch.no_grad():
        other = [p[:50].contiguous() for p in params]
        r.forward(*other)                                   # e.g. a validation render
    H.sum().backward()
    ps2 = [p.clone().requires_grad_() for p in params]
    render_transient_batch(_r(geom, ta, "numerical"), *ps2).sum().backward()
    for a, b in zip(ps, ps2):
        assert util.rel_to_max(a.grad.cpu().numpy(), b.grad.cpu().numpy()) < 1e-5


def test_list_overflow_is_detected_one_step_later():
    from nlos_gaussian_renderer_b200.transient import ListOverflow
    geom, ta, params = util.small_scene(100, 5, 9, 48, 6, device="cuda")
    r = _r(geom, ta, "numerical")
    H0 = r.forward(*[p[:10].contiguous() for p in params])          # sizes the list buffer for 10 Gaussians
    r._N = None                                                      # (new N: workspaces are rebuilt, the list buffer is sized again)
    H1 = r.forward(*params)
    assert float(H1.abs().max()) > 0
    # same N, many more entries than the buffer holds: shrink the buffer by hand
    r._list = r._list[:8]
    r._list_bins = r._list_bins[:8]
    H2 = r.forward(*params)
    assert float(H2.abs().max()) == 0.0                              # the launch rendered nothing ...
    with pytest.raises(ListOverflow):
        r.check_lists()                                              # ... and says so
    H3 = r.forward(*params)                                          # the buffer has been regrown
    assert torch.equal(H3, H1) or util.rel_to_max(H3.cpu().numpy(), H1.cpu().numpy()) < 1e-6


@pytest.mark.parametrize("mode", ["numerical", "analytic"])
def test_deterministic_mode_is_bit_reproducible(mode):
    """deterministic=True: H and every gradient are bit-identical run to run (the default mode accumulates with float REDs in scheduling order and
    is only reproducible to rounding); and it agrees with the default mode to rounding."""
    N, Sy, Sx, T, A = 500, 8, 16, 128, 8
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=1.5, device="cuda")
    feats = (torch.randn(N, 16, generator=torch.Generator().manual_seed(3)) * 0.3).cuda()
    params = list(params[:4]) + [feats]
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(2)).cuda()
    runs = []
    for rep in range(3):
        r = _r(geom, ta, mode, deterministic=rep < 2, active_sh_degree=2, window=64, bucket=8)
        ps = [p.clone().requires_grad_() for p in params]
        H = render_transient_batch(r, *ps)
        (H * gH).sum().backward()
        runs.append((H.detach().clone(), [p.grad.clone() for p in ps]))
    assert torch.equal(runs[0][0], runs[1][0])
    for a, b in zip(runs[0][1], runs[1][1]):
        assert torch.equal(a, b)
    assert util.rel_to_max(runs[2][0].cpu().numpy(), runs[0][0].cpu().numpy()) < 1e-5
    for a, b in zip(runs[2][1], runs[0][1]):
        assert util.rel_to_max(a.cpu().numpy(), b.cpu().numpy()) < 2e-5
