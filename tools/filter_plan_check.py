#!/usr/bin/env python
"""Histogram-level check (CPU, fp64) of the plan in DESIGN.md section 8 item 4: many rays of random width >= 1.3 bins and random centre summed into one
bin row with 3 halo bins, 1-node mixture per ray, ONE 7-tap filter over the summed row, the 1/(r^2+1e-8) weighting applied after the filter.
Compared with the exact bin integrals, relative to the largest bin of the row; also the adjoint identity the backward pass relies on.

    python tools/filter_plan_check.py
"""
import numpy as np
from scipy.special import erf

C4, C6, HALO = 2.4e-4, -1.7e-4, 3


def d2(a):                        # zero-padded second difference: a symmetric Toeplitz matrix, so F below is its own adjoint
    return np.convolve(a, [1.0, -2.0, 1.0], mode="same")


def filt(x):                      # F = I - C4 d4 - C6 d6   (symmetric)
    d4 = d2(d2(x))
    return x - C4 * d4 - C6 * d2(d4)


rng = np.random.default_rng(0)
T, dt, t0 = 128, 1.3 / 128, 0.35
k = np.arange(-HALO, T + HALO)
r = t0 + k * dt
worst = 0.0
for trial in range(20):
    n = 400
    sig = rng.uniform(1.3, 6.0, n) * dt                    # ray widths
    mu = rng.uniform(r[0] - 4 * dt, r[-1] + 4 * dt, n)     # some rays straddle the ends of the row
    amp = rng.uniform(0.1, 1.0, n)
    exact = (amp[:, None] * sig[:, None] * np.sqrt(np.pi / 2) *
             (erf((r[None] + dt / 2 - mu[:, None]) / (sig[:, None] * np.sqrt(2))) - erf((r[None] - dt / 2 - mu[:, None]) / (sig[:, None] * np.sqrt(2))))).sum(0)
    u = 1 + (dt * dt / 12) / sig ** 2
    one = (amp[:, None] * dt * u[:, None] ** -0.5 * np.exp(-(r[None] - mu[:, None]) ** 2 / (2 * sig[:, None] ** 2 * u[:, None]))).sum(0)
    w = 1.0 / (r * r + 1e-8)
    H_exact = (exact * w)[HALO:-HALO]
    H_plan = (filt(one) * w)[HALO:-HALO]
    H_raw = (one * w)[HALO:-HALO]
    e_plan = np.abs(H_plan - H_exact).max() / np.abs(H_exact).max()
    e_raw = np.abs(H_raw - H_exact).max() / np.abs(H_exact).max()
    worst = max(worst, e_plan)
    if trial < 3:
        print(f"trial {trial}: 1-node only {e_raw:.2e}   1-node + filter {e_plan:.2e}")
print(f"worst of 20 rows (400 rays each, widths 1.3-6 bins): {worst:.2e}")
# adjoint: <g, crop(w F x)> == <F pad(w g), x>
x = rng.normal(size=T + 2 * HALO)
g = rng.normal(size=T)
w = 1.0 / (r * r + 1e-8)
lhs = (g * (filt(x) * w)[HALO:-HALO]).sum()
gp = np.zeros(T + 2 * HALO)
gp[HALO:-HALO] = g * w[HALO:-HALO]
print(f"adjoint identity: {lhs:.12e} vs {(filt(gp) * x).sum():.12e}")
