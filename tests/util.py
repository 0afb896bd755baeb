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


def oracle_H(geom, ta, params, mode=0, precision="f64", use_occlusion=False, cap=0, count=False):
    means, ls, rots, opac, feats = [np32(p) for p in params]
    return oracle.transient_batch(np32(geom.scan_o), np32(geom.sin_th), np32(geom.cos_th), np32(geom.sin_ph), np32(geom.cos_ph),
                                  np32(ta["t_samples"]), np32(geom.hscale), means, ls, rots, opac, feats, sh_degree=0,
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
