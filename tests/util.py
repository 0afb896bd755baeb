"""Shared helpers for the tests: small seeded scenes, oracle calls, error metrics."""
import numpy as np
import torch

import oracle
from nlos_gaussian_renderer_b200 import scenes
from nlos_gaussian_renderer_b200.geometry import ScanGeometry  # noqa: F401 (re-exported for the tests)


def small_scene(N=150, Sy=5, Sx=9, T=48, A=6, seed=7, scale_mul=1.0, device="cpu"):
    geom = ScanGeometry.wall_grid(Sy, Sx, A=A, device=device)
    ta = scenes.time_axis(T)
    params = scenes.make_gaussians(N, seed, device=device, scale_mul=scale_mul)
    return geom, ta, params


def np32(t):
    return t.detach().cpu().numpy().astype(np.float32)


def oracle_H(geom, ta, params, mode=0, precision="f64", use_occlusion=False, cap=0, count=False, sh_degree=0):
    means, ls, rots, opac, feats = [np32(p) for p in params]
    return oracle.transient_batch(np32(geom.scan_o), np32(geom.sin_th), np32(geom.cos_th), np32(geom.sin_ph), np32(geom.cos_ph),
                                  np32(ta["t_samples"]), np32(geom.hscale), means, ls, rots, opac, feats, sh_degree=sh_degree,
                                  c=ta["c"], deltaT=ta["deltaT"], scaling_modifier=1.0, use_occlusion=use_occlusion, mode=mode,
                                  cap=cap, exp_mode=1, precision=precision, count=count)


def ray_masks(geom, params):
    """[S, A*A, N] bool: slab-test membership of every (scan point, ray, Gaussian) from the C oracle."""
    means, ls, rots, _, _ = [np32(p) for p in params]
    N = means.shape[0]
    bb = oracle.bboxes(means, ls, rots, 1.0, 3.0, 1)
    S, A = geom.S, geom.A
    out = np.zeros((S, A * A, N), bool)
    so, st, ct, sp, cp = [np32(x) for x in (geom.scan_o, geom.sin_th, geom.cos_th, geom.sin_ph, geom.cos_ph)]
    for p in range(S):
        d = np.stack([(st[p][:, None] * cp[p][None, :]).reshape(-1), (st[p][:, None] * sp[p][None, :]).reshape(-1),
                      np.repeat(ct[p], A)], 1).astype(np.float32)
        o = np.repeat(so[p][None], A * A, 0)
        lst = oracle.filter_gaussians(o, d, bb, cap=N)
        for r in range(A * A):
            out[p, r, lst[r, 1:1 + lst[r, 0]]] = True
    return out


def rel_to_max(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def oracle_window_lists(geom, ta, params, r):
    """The oracle's definition of the renderer r's entry lists (counts, offsets, list), with r's entry classes."""
    T = r.T
    klo, khi = oracle.shell_bins(np32(geom.scan_o), geom.Sy, geom.Sx, geom.tile_y, geom.tile_x, np32(params[0]), np32(params[1]),
                                 r.mod, r.mcut, ta["t0"], ta["dt"], T, exp_mode=1)
    cls = None
    if r.n_classes > 1:
        lat = np32(geom.lattice)                         # (theta0, dtheta, phi0, dphi) of the linspaces
        A, PI = geom.A, np.float32(3.14159265358979)
        fast = (A > 1) & (lat[:, 1] > 1e-9) & (lat[:, 3] > 1e-9) & (lat[:, 0] >= 0) & (lat[:, 0] + lat[:, 1] * (A - 1) <= PI) & \
               (lat[:, 2] >= -PI) & (lat[:, 2] + lat[:, 3] * (A - 1) <= PI)             # csrc/preprocess.cu:build_scan_tiles_kernel `fast`
        step = np.where(fast, np.minimum(lat[:, 1], lat[:, 3]), np.float32(0)).astype(np.float32)
        cls = oracle.entry_classes(np32(geom.scan_o), None, step, geom.Sy, geom.Sx, geom.tile_y, geom.tile_x, np32(params[0]), np32(params[1]),
                                   r.mod, r.mcut, klo, khi, r.window, r.bucket, r.narrow_rays if r.n_fp == 2 else 0.0, exp_mode=1,
                                   time_wide_min=r.tw_min, T=T)
    return oracle.window_lists(klo, khi, T, r.window, r.bucket, cls, n_classes=r.n_classes), cls


def hit_masks_gpu(r, params, backward=False):
    """[S, A*A, N] bool: the slab-test membership the render kernels of renderer r actually used (debug dump of the forward launch; of the
    backward launch after an undumped forward with backward=True -- the masks that pass then reads from the forward's culling words)."""
    from nlos_gaussian_renderer_b200 import _lib
    g = r.geom
    N = params[0].shape[0]
    dump = torch.zeros(g.n_tiles * 32, g.A * g.A, N, dtype=torch.uint8, device="cuda")
    r.prepare(*params)
    if backward:
        H = r.forward_tiled()
    _lib.call("nlosb200_debug_hit_dump", dump.data_ptr())
    try:
        if backward:
            r.render_backward(torch.ones_like(H))
        else:
            r.forward_tiled()
        torch.cuda.synchronize()
    finally:
        _lib.call("nlosb200_debug_hit_dump", None)
    out = torch.zeros(g.S, g.A * g.A, N, dtype=torch.uint8, device="cuda")
    out[g.perm[g.valid]] = dump[g.valid]
    return out.cpu().numpy().astype(bool)
