"""Hit-count statistics of (Gaussian, scan point) pairs per scan tile -- design input for the small-footprint kernel variant.
CPU only (numpy fp32 slab tests, the reference predicate of cuda_utils.cuh:97-121).   python tools/footprint_stats.py cfg4"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from nlos_gaussian_renderer_b200 import scenes

name = sys.argv[1] if len(sys.argv) > 1 else "north_star"
n_g = int(sys.argv[2]) if len(sys.argv) > 2 else 400
N, Sy, Sx, T, A = scenes.CONFIGS[name]
geom, ta, params = scenes.make_scene(name)
means, ls, rots, opac, feats = [p.numpy() for p in params]
rng = np.random.default_rng(0)
gs = rng.choice(N, n_g, replace=False)
# AABB (3 sigma) of the sampled Gaussians
q = rots[gs] / np.linalg.norm(rots[gs], axis=1, keepdims=True)
w, x, y, z = q.T
R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y), 2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
              2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)
s = np.exp(ls[gs])
ext = 3 * np.sqrt(((R * s[:, None, :]) ** 2).sum(-1))
bmin, bmax = means[gs] - ext, means[gs] + ext
dt = ta["dt"]
tiles = rng.choice(geom.n_tiles, 6, replace=False)
perm = geom.perm.numpy()
so, st, ct, sp, cp = [t.numpy() for t in (geom.scan_o, geom.sin_th, geom.cos_th, geom.sin_ph, geom.cos_ph)]
stats = {"hits": [], "maxhits": [], "pairit": [], "maxpairit": [], "nbins": []}
for t in tiles:
    pts = perm[t * 32:(t + 1) * 32]
    hits = np.zeros((32, n_g), int)
    for l, p in enumerate(pts):
        d = np.stack([st[p][:, None] * cp[p][None, :], st[p][:, None] * sp[p][None, :], np.repeat(ct[p][:, None], A, 1)], -1).reshape(-1, 3).astype(np.float32)
        inv = (1.0 / (d + np.float32(1e-8))).astype(np.float32)
        o = so[p]
        t0 = (bmin[None, :, :] - o) * inv[:, None, :]
        t1 = (bmax[None, :, :] - o) * inv[:, None, :]
        tmin = np.minimum(t0, t1).max(-1)
        tmax = np.maximum(t0, t1).min(-1)
        hit = (tmax >= tmin) & (tmax >= 0)
        hits[l] = hit.sum(0)
    stats["hits"].append(hits)
h = np.concatenate([x.reshape(-1) for x in stats["hits"]])
print(name, "dt", dt, "mean hits per pair", h.mean(), "hist", np.bincount(h, minlength=10)[:12] / h.size)
mx = np.concatenate([x.max(0) for x in stats["hits"]])
print("max over 32 lanes: mean", mx.mean(), "hist", np.bincount(mx, minlength=10)[:16] / mx.size)
pit = np.concatenate([np.ceil(x / 2).max(0) for x in stats["hits"]])
print("ray-pair iterations per entry (max over lanes of ceil(h/2)): mean", pit.mean())
tot = np.concatenate([x.sum(0) for x in stats["hits"]])
print("hits per entry (sum over lanes): mean", tot.mean(), " -> lane utilisation single-ray loop", tot.mean() / (32 * mx.mean()),
      " two-ray loop", tot.mean() / (64 * pit.mean()))
print("sigma/dt quantiles", np.quantile(s / dt, [0.05, 0.5, 0.95]))
