"""Native batched occlusion renderer (csrc/occluded.cu): the reference's per-ray semantics (src/ray_aabb.cu:10-61 lists with the 256 cap in
index order, src/volume_renderer.cu:80-137 shared transmittance with the early stop) for a whole scan grid, with gradients."""
import numpy as np
import pytest
import torch

import oracle
from nlos_gaussian_renderer_b200 import _lib
from nlos_gaussian_renderer_b200._C import _aligned
from nlos_gaussian_renderer_b200.occluded import render_transient_batch_occluded
from tests import ref_torch, util
from tests.test_gpu_transient import scenes_cpu

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _gpu_lists(geom, params, cap):
    L = _lib.lib()
    N = params[0].shape[0]
    rec = _aligned(N * 32, torch.float32, "cuda", 128)
    st = _lib.stream()
    m, ls, rt, op, ft = [p.contiguous() for p in params]
    _lib.call("nlosb200_preprocess", m.data_ptr(), ls.data_ptr(), rt.data_ptr(), op.reshape(-1).data_ptr(), ft.data_ptr(), N, ft.reshape(N, -1).shape[1], 1.0,
              3.0, 5.3, 0.0, rec.data_ptr(), None, None, st)
    n_rows = geom.n_tiles * 32
    lists = torch.full((n_rows, geom.A * geom.A, cap + 1), -1, dtype=torch.int32, device="cuda")
    _lib.call("nlosb200_occluded_lists", geom.tiles().data_ptr(), geom.n_tiles, geom.A, rec.data_ptr(), N, cap, 0, n_rows, lists.data_ptr(), st)
    out = torch.full((geom.S, geom.A * geom.A, cap + 1), -1, dtype=torch.int32, device="cuda")
    out[geom.perm[geom.valid]] = lists[geom.valid]
    return out.cpu().numpy()


@pytest.mark.parametrize("shape,scale_mul,cap", [((300, 3, 5, 6), 1.0, 256), ((900, 2, 3, 8), 4.0, 256), ((700, 2, 2, 5), 6.0, 40), ((64, 5, 9, 12), 0.4, 256)])
def test_lists_equal_the_reference_filter_bit_for_bit(shape, scale_mul, cap):
    """per-ray lists == the reference's filter (first `cap` AABB hits in index order, -1 padding aside), including scenes where the cap binds."""
    N, Sy, Sx, A = shape
    geom, ta, params = util.small_scene(N, Sy, Sx, 32, A, scale_mul=scale_mul, device="cuda")
    got = _gpu_lists(geom, params, cap)
    cg = scenes_cpu(geom)
    means, ls, rots = [util.np32(p) for p in params[:3]]
    bb = oracle.bboxes(means, ls, rots, 1.0, 3.0, 1)
    so, st, ct, sp, cp = [util.np32(x) for x in (cg.scan_o, cg.sin_th, cg.cos_th, cg.sin_ph, cg.cos_ph)]
    binds = False
    for p in range(cg.S):
        d = np.stack([(st[p][:, None] * cp[p][None, :]).reshape(-1), (st[p][:, None] * sp[p][None, :]).reshape(-1), np.repeat(ct[p], A)], 1).astype(np.float32)
        o = np.repeat(so[p][None], A * A, 0)
        want = oracle.filter_gaussians(o, d, bb, cap=cap)
        assert np.array_equal(got[p][:, 0], want[:, 0]), p
        for r in range(A * A):
            n = want[r, 0]
            assert np.array_equal(got[p][r, 1:1 + n], want[r, 1:1 + n]), (p, r)
        binds = binds or (want[:, 0] == cap).any()
    if scale_mul >= 4.0:
        assert binds                                         # the cap is exercised


@pytest.mark.parametrize("occ", [True, False])
@pytest.mark.parametrize("deg", [0, 2])
def test_forward_matches_oracle_and_backward_matches_autograd(occ, deg):
    N, Sy, Sx, T, A = 150, 3, 5, 48, 6
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=3.0, device="cuda")
    params = list(params)
    params[3] = params[3] + 3.0                                                   # opaque enough for the transmittance to matter
    if deg:
        f = torch.randn(N, 16, generator=torch.Generator().manual_seed(9)) * 0.4
        f[:, 0] = params[4].cpu()[:, 0]
        params[4] = f.cuda()
    deltaT = 0.6 if occ else ta["deltaT"]
    ts = ta["t_samples"].cuda()
    ps = [p.clone().requires_grad_() for p in params]
    H = render_transient_batch_occluded(geom, ts, *ps, ta["c"], deltaT, use_occlusion=occ, active_sh_degree=deg, chunk_scan_points=32)
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(4)).cuda()
    (H * gH).sum().backward()
    cg = scenes_cpu(geom)
    tb = dict(ta); tb["deltaT"] = deltaT
    H64 = util.oracle_H(cg, tb, [p.cpu() for p in params], mode=0, precision="f64", use_occlusion=occ, cap=256, sh_degree=deg)
    if occ:
        Hno = util.oracle_H(cg, tb, [p.cpu() for p in params], mode=0, precision="f64", use_occlusion=False, cap=256, sh_degree=deg)
        assert util.rel_to_max(Hno, H64) > 1e-2                                   # occlusion matters in this scene
    assert util.rel_to_max(H.detach().cpu().numpy().reshape(-1, T), H64) < TOL
    # gradient truth: fp64 autograd through the per-ray restatement (membership from the oracle; the cap does not bind with 150 Gaussians)
    masks = torch.from_numpy(util.ray_masks(cg, [p.cpu() for p in params]))
    pd = [p.detach().cpu().double().requires_grad_() for p in params]
    tsd = ts.cpu().double()
    rows = []
    for p in range(cg.S):
        o = cg.scan_o[p].double()
        d = torch.stack([(cg.sin_th[p][:, None] * cg.cos_ph[p][None, :]).reshape(-1), (cg.sin_th[p][:, None] * cg.sin_ph[p][None, :]).reshape(-1),
                         cg.cos_th[p][:, None].expand(A, A).reshape(-1)], 1).double()
        rho, _, _ = ref_torch.render_rays(o[None].expand(A * A, 3), d, tsd, masks[p], *pd, ta["c"], deltaT, use_occlusion=occ, sh_degree=deg, cam=o)
        w = cg.sin_th[p][:, None].expand(A, A).reshape(-1).double()
        rows.append(cg.hscale[p].double() * (rho * w[:, None]).sum(0) / (tsd * tsd + 1e-8))
    Ht = torch.stack(rows)
    (Ht * gH.cpu().double().reshape(-1, T)).sum().backward()
    assert util.rel_to_max(H.detach().cpu().numpy().reshape(-1, T), Ht.detach().numpy()) < TOL
    for name, a, b in zip(("means", "log_scales", "rotations", "opacities", "features"), ps, pd):
        assert util.rel_to_max(a.grad.cpu().numpy(), b.grad.numpy()) < 2e-4, name


def test_early_stop_and_cap_in_a_dense_opaque_scene():
    """many opaque Gaussians: the 256 cap binds on most rays and the transmittance falls below 1e-4 (zero-filled tail, volume_renderer.cu:128-136)."""
    N, Sy, Sx, T, A = 3000, 2, 3, 64, 6
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=2.5, device="cuda")
    params = list(params)
    params[3] = params[3] * 0 + 6.0
    ts = ta["t_samples"].cuda()
    H = render_transient_batch_occluded(geom, ts, *params, ta["c"], 2.0, use_occlusion=True)
    tb = dict(ta); tb["deltaT"] = 2.0
    H64 = util.oracle_H(scenes_cpu(geom), tb, [p.cpu() for p in params], mode=0, precision="f64", use_occlusion=True, cap=256)
    Hu = util.oracle_H(scenes_cpu(geom), tb, [p.cpu() for p in params], mode=0, precision="f64", use_occlusion=True, cap=0)
    assert util.rel_to_max(Hu, H64) > 1e-3                                        # the cap changes the result here
    assert H64[:, -8:].max() < 1e-9 * H64.max() and H64.max() > 0                 # occluded tail
    assert util.rel_to_max(H.cpu().numpy().reshape(-1, T), H64) < 2e-4


def test_workspace_blocking_does_not_change_the_result():
    N, Sy, Sx, T, A = 200, 9, 9, 40, 5
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=2.0, device="cuda")
    ts = ta["t_samples"].cuda()
    a = render_transient_batch_occluded(geom, ts, *params, ta["c"], 0.3, chunk_scan_points=32)
    b = render_transient_batch_occluded(geom, ts, *params, ta["c"], 0.3)
    assert util.rel_to_max(a.cpu().numpy(), b.cpu().numpy()) < 1e-6
