#!/usr/bin/env python
"""The tensor-core question of SURVEY.md 7, answered with data (CPU, numpy): the Mahalanobis form m2(x) = w_g . phi(x), phi = [x2,y2,z2,xy,xz,yz,x,y,z,1],
IS a dense [N_g x 10] . [10 x N_s] contraction -- what precision does it need?  exp(-m2/2) to 1e-4 needs |delta m2| < 2e-4.  The sample points of the
real path lie ~1 m from the scan point and the Gaussians are millimetres wide, so the ten terms are ~1e4 .. 1e5 each and cancel to m2 <= 28.
Operands are rounded to TF32 (10-bit mantissa, what tcgen05 kind::tf32 multiplies), to a 2-term and a 3-term TF32 split (3 / 6 MMAs per product) and
to fp32; products are accumulated in fp32 as the tensor core does.  The re-centred fp32 form the kernels use (m2min + C (r - r*)^2) is the last row.

    python tools/tf32_contraction_error.py [cfg4]"""
import sys

import numpy as np

sys.path.insert(0, ".")
from nlos_gaussian_renderer_b200 import scenes  # noqa: E402


def tf32(a):
    b = np.asarray(a, np.float32).view(np.uint32)
    b = (b + np.uint32(0x1000)) & np.uint32(0xFFFFE000)          # round to 10 mantissa bits
    return b.view(np.float32)


def split(a, n):
    parts, r = [], np.asarray(a, np.float32).copy()
    for _ in range(n):
        p = tf32(r)
        parts.append(p)
        r = (r - p).astype(np.float32)
    return parts


def contract(W, PHI, n):
    """sum_k W[:,k] PHI[:,k] with n-term TF32 splits of both operands (all cross products of the leading terms), fp32 accumulation."""
    if n == 0:
        return (W.astype(np.float32) * PHI.astype(np.float32)).sum(1, dtype=np.float32)
    Ws, Ps = split(W, n), split(PHI, n)
    acc = np.zeros(W.shape[0], np.float32)
    for i in range(n):
        for j in range(n - i):
            acc = (acc + (Ws[i].astype(np.float32) * Ps[j].astype(np.float32)).sum(1, dtype=np.float32)).astype(np.float32)
    return acc


cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg4"
N, Sy, Sx, T, A = scenes.CONFIGS[cfg]
rng = np.random.default_rng(0)
means, ls, rots, _, _ = [p.numpy().astype(np.float64) for p in scenes.make_gaussians(N, 1005)]
g = rng.choice(N, 20000)
q = rots[g] / np.linalg.norm(rots[g], axis=1, keepdims=True)
w, x, y, z = q.T
R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y), 2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
              2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)
s = np.exp(ls[g])
M = R.transpose(0, 2, 1) / s[:, :, None]
Sinv = np.einsum("nji,njk->nik", M, M)                                            # Sigma^-1
o = np.stack([rng.uniform(-0.5, 0.5, len(g)), np.zeros(len(g)), rng.uniform(-0.5, 0.5, len(g))], 1)     # scan point (origin of the lattice)
# sample point within ~2 sigma of the mean, coordinates RELATIVE TO THE SCAN POINT (what a [10 x N_s] operand shared by all Gaussians would hold)
xs = means[g] + np.einsum("nij,nj->ni", R, s * rng.normal(size=(len(g), 3))) - o
mu = means[g] - o
d = xs - mu
m2_true = np.einsum("ni,nij,nj->n", d, Sinv, d)
b = -2 * np.einsum("nij,nj->ni", Sinv, mu)
c0 = np.einsum("ni,nij,nj->n", mu, Sinv, mu)
W = np.stack([Sinv[:, 0, 0], Sinv[:, 1, 1], Sinv[:, 2, 2], 2 * Sinv[:, 0, 1], 2 * Sinv[:, 0, 2], 2 * Sinv[:, 1, 2], b[:, 0], b[:, 1], b[:, 2], c0], 1)
PHI = np.stack([xs[:, 0] ** 2, xs[:, 1] ** 2, xs[:, 2] ** 2, xs[:, 0] * xs[:, 1], xs[:, 0] * xs[:, 2], xs[:, 1] * xs[:, 2], xs[:, 0], xs[:, 1], xs[:, 2],
                np.ones(len(g))], 1)
print(f"{cfg}: {len(g)} (Gaussian, sample) pairs, m2 in [{m2_true.min():.2f}, {m2_true.max():.2f}], largest term of the contraction {np.abs(W * PHI).max():.3g}")
print("requirement: |delta m2| < 2e-4 (exp(-m2/2) to 1e-4)")
for name, n, mmas in (("TF32 operands (1 MMA)", 1, 1), ("2-term TF32 split (3 MMAs)", 2, 3), ("3-term TF32 split (6 MMAs)", 3, 6), ("fp32 operands, fp32 accumulate", 0, 0)):
    err = np.abs(contract(W, PHI, n).astype(np.float64) - m2_true)
    print(f"  {name:34s} |delta m2|: median {np.median(err):.3g}  99% {np.quantile(err, 0.99):.3g}  max {err.max():.3g}   within 2e-4: {100 * (err < 2e-4).mean():.1f} %")
# the kernels' form: a' = M (x0 - mu) at the point of the ray nearest the mean, m2 = |a' + tau b|^2, all fp32
a32 = np.einsum("nij,nj->ni", M.astype(np.float32), d.astype(np.float32)).astype(np.float32)
err = np.abs((a32 * a32).sum(1, dtype=np.float32).astype(np.float64) - m2_true)
print(f"  {'re-centred fp32 (the kernels)':34s} |delta m2|: median {np.median(err):.3g}  99% {np.quantile(err, 0.99):.3g}  max {err.max():.3g}   within 2e-4: {100 * (err < 2e-4).mean():.1f} %")
