"""TEST INFRASTRUCTURE (not product code): CPU restatement of the MCMC relocation helper.

GaussianModel._update_params (/root/reference/gaussian_model/gaussian_model.py:471-483) calls `compute_relocation_cuda(opacity_old,
scale_old, N)`; the reference neither defines nor imports that name.  It is the helper of the public 3DGS-MCMC
code (Kheradmand et al., "3D Gaussian Splatting as Markov Chain Monte Carlo", NeurIPS 2024), which is not vendored under
/root/reference, so this restatement follows the publication (eq. 9) -- PARITY UNPINNED:

    o_new = 1 - (1 - o_old)^(1/N)
    s_new = s_old * o_old / sum_{i=1..N} sum_{k=0..i-1} C(i-1, k) (-1)^k o_new^(k+1) / sqrt(k+1)

The double sum is kept in the published order with a float32 accumulator (each term formed in float64): the series alternates with
binomial-sized terms, so for large N its rounding is part of the function.  `relocation_closed_form` is the same quantity by the
hockey-stick identity (sum_i C(i-1,k) = C(N,k+1)) in exact rational-free float64 -- the cross-check for small N."""
import math

import numpy as np

N_MAX = 51


def binoms(n_max=N_MAX):
    b = np.zeros((n_max, n_max), np.float32)
    for n in range(n_max):
        for k in range(n + 1):
            b[n, k] = math.comb(n, k)
    return b


def compute_relocation(opacity_old, scale_old, counts, n_max=N_MAX):
    opacity_old = np.asarray(opacity_old, np.float32)
    scale_old = np.asarray(scale_old, np.float32).reshape(-1, 3)
    counts = np.clip(np.asarray(counts, np.int64), 1, n_max - 1)
    B = binoms(n_max)
    o_new = np.empty_like(opacity_old)
    s_new = np.empty_like(scale_old)
    for idx in range(opacity_old.shape[0]):
        N = int(counts[idx])
        o = np.float32(1.0) - np.float32(np.power(np.float32(1.0) - opacity_old[idx], np.float32(1.0) / np.float32(N), dtype=np.float32))
        o_new[idx] = o
        term = [(-1.0 if k & 1 else 1.0) / math.sqrt(k + 1) * float(o) ** (k + 1) for k in range(N)]
        denom = np.float32(0.0)
        for i in range(1, N + 1):
            for k in range(i):
                denom = np.float32(float(denom) + float(B[i - 1, k]) * term[k])
        s_new[idx] = (opacity_old[idx] / denom) * scale_old[idx]
    return o_new, s_new


def relocation_closed_form(opacity_old, counts):
    """-> (o_new, scale factor) in float64 through sum_k C(N, k+1) (-1)^k o_new^(k+1) / sqrt(k+1)."""
    opacity_old = np.asarray(opacity_old, np.float64)
    counts = np.asarray(counts, np.int64)
    o_new = 1.0 - (1.0 - opacity_old) ** (1.0 / counts)
    fac = np.empty_like(o_new)
    for idx in range(o_new.shape[0]):
        N = int(counts[idx])
        d = sum(math.comb(N, k + 1) * (-1.0) ** k * o_new[idx] ** (k + 1) / math.sqrt(k + 1) for k in range(N))
        fac[idx] = opacity_old[idx] / d
    return o_new, fac
