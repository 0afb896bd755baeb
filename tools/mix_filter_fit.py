#!/usr/bin/env python
"""Design study for DESIGN.md section 8 item 4 (CPU, fp64): how accurate is the bin integral of a ray's Gaussian when it is replaced by
  (a) the 1-node / 2-node moment-matched mixtures the analytic mode uses today, and
  (b) the 1-node mixture followed by ONE fixed linear filter of the bin row,  I' = I - c4 d4(I) - c6 d6(I)   (d4, d6 = central differences),
as a function of the ray's width in bins.  (b) is what would make every ray wider than ~1.3 bins a one-pass ray: the filter is the same for all
rays, so it can be applied once per window flush (forward) / refill (backward) instead of per ray.  Errors are relative to the ray's peak bin.

    python tools/mix_filter_fit.py
"""
import numpy as np
from scipy.special import erf


def exact(sig, off):
    k = np.arange(-60, 61) + off
    return k, sig * np.sqrt(np.pi / 2) * (erf((k + .5) / (sig * np.sqrt(2))) - erf((k - .5) / (sig * np.sqrt(2))))


def one_node(k, sig):
    u = 1 + (1 / 12) / sig ** 2
    return u ** -0.5 * np.exp(-k ** 2 / (2 * sig ** 2 * u))


def two_node(k, sig):
    x = (1 / 240) ** 0.25
    u = 1 + (1 / 12 - x * x) / sig ** 2
    return 0.5 * u ** -0.5 * (np.exp(-(k + x) ** 2 / (2 * sig ** 2 * u)) + np.exp(-(k - x) ** 2 / (2 * sig ** 2 * u)))


def diff(a, n):
    for _ in range(n // 2):
        b = np.zeros_like(a)
        b[1:-1] = a[:-2] - 2 * a[1:-1] + a[2:]
        a = b
    return a


def worst(fn, sigs):
    w = 0.0
    for sig in sigs:
        for off in np.linspace(0, 0.5, 6):
            k, ex = exact(sig, off)
            w = max(w, np.abs(fn(k, sig) - ex).max() / ex.max())
    return w


if __name__ == "__main__":
    for s in (5.0, 3.0, 2.5, 2.0, 1.8, 1.5, 1.3, 1.0, 0.8):
        f4 = lambda k, sig: one_node(k, sig) - diff(one_node(k, sig), 4) / 2880
        f6 = lambda k, sig: one_node(k, sig) - 2.4e-4 * diff(one_node(k, sig), 4) + 1.7e-4 * diff(one_node(k, sig), 6)
        print(f"width {s:4.2f} bins: 1 node {worst(one_node, [s]):.2e}   2 nodes {worst(two_node, [s]):.2e}   "
              f"1 node + d4/2880 {worst(f4, [s]):.2e}   1 node + (2.4e-4 d4 - 1.7e-4 d6) {worst(f6, [s]):.2e}")
