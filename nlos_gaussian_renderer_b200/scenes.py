"""Synthetic scenes of SURVEY.md 8(d) (there is no network for datasets): relay wall y=0, hidden cube at
(0,0.75,0) side 0.5, c=1, bins on r in [0.35,1.65].  Used by tests/ and bench.py."""
import math

import torch

from .geometry import ScanGeometry

C0 = 0.28209479177387814

CONFIGS = {
    # name: (N, Sy, Sx, T, A)   -- BASELINE.json configs[0..4]
    "cfg1": (1_000, 16, 16, 128, 32),
    "cfg2": (50_000, 64, 64, 512, 32),
    "cfg3": (200_000, 64, 64, 512, 32),
    "cfg4": (500_000, 128, 128, 1024, 32),
    "cfg5": (1_000_000, 256, 256, 1024, 32),
    "north_star": (1_000_000, 128, 128, 1024, 32),
}


def time_axis(T, r_lo=0.35, r_hi=1.65, c=1.0):
    """c*deltaT = (r_hi-r_lo)/T, start = round(r_lo/(c dT)), end = start+T, radii = linspace(start cdT, end cdT, T)
    (nlos_helpers.py:165-172).  Returns dict(c, deltaT, start, end, t0, dt, t_samples)."""
    cdt = (r_hi - r_lo) / T
    start = int(round(r_lo / cdt))
    end = start + T
    deltaT = cdt / c
    r_min, r_max = start * c * deltaT, end * c * deltaT
    t = torch.linspace(r_min, r_max, T, dtype=torch.float32)
    dt = (r_max - r_min) / max(T - 1, 1)
    return {"c": c, "deltaT": deltaT, "start": start, "end": end, "t0": float(r_min), "dt": float(dt), "t_samples": t}


def make_gaussians(N, seed, volume_position=(0.0, 0.75, 0.0), volume_size=0.5, device="cpu", scale_mul=1.0):
    g = torch.Generator().manual_seed(seed)
    vp = torch.tensor(volume_position, dtype=torch.float32)
    half = 0.5 * volume_size * 0.9
    means = vp + (torch.rand(N, 3, generator=g) * 2 - 1) * half
    s0 = 0.5 * (volume_size ** 3 / N) ** (1.0 / 3.0) * scale_mul
    log_scales = torch.log(s0 * (0.5 + 1.5 * torch.rand(N, 3, generator=g)))
    rots = torch.randn(N, 4, generator=g) * (0.5 + 1.5 * torch.rand(N, 1, generator=g))
    opac = torch.randn(N, 1, generator=g)
    feats = (0.2 * torch.rand(N, 1, generator=g) - 0.5) / C0          # RHO2SH(U(0,0.2)), sh_utils.py:114
    return [t.to(device).contiguous() for t in (means, log_scales, rots, opac, feats)]


def make_scene(name_or_tuple, seed=None, device="cpu", A=None, tile=(4, 8)):
    """-> (geom, time axis dict, [means, log_scales, rots, opac, feats])"""
    if isinstance(name_or_tuple, str):
        N, Sy, Sx, T, A0 = CONFIGS[name_or_tuple]
        if seed is None:
            seed = 1000 + list(CONFIGS).index(name_or_tuple) + 1
    else:
        N, Sy, Sx, T, A0 = name_or_tuple
        seed = 1234 if seed is None else seed
    A = A or A0
    geom = ScanGeometry.wall_grid(Sy, Sx, A=A, tile=tile, device=device)
    return geom, time_axis(T), make_gaussians(N, seed, device=device)
