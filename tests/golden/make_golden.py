"""Generate tests/golden/ref_small.npz from the REFERENCE's own kernel source compiled for the host
(oracle/_ref; needs /root/reference).  Run once here:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle.ref as oref  # noqa: E402
from nlos_gaussian_renderer_b200 import scenes  # noqa: E402

A = 6
th = torch.linspace(0.9, 2.2, A)
ph = torch.linspace(0.8, 2.3, A)
tg, pg = torch.meshgrid(th, ph, indexing="ij")
d = torch.stack([torch.sin(tg) * torch.cos(pg), torch.sin(tg) * torch.sin(pg), torch.cos(tg)], -1).reshape(-1, 3).numpy().astype(np.float32)
o = np.repeat(np.array([[0.1, 0.0, -0.05]], np.float32), A * A, 0)
t = np.linspace(0.35, 1.65, 32).astype(np.float32)
means, ls, rots, opac, feats = [p.numpy().astype(np.float32) for p in scenes.make_gaussians(120, 11, scale_mul=2.0)]
out = dict(o=o, d=d, t=t, means=means, ls=ls, rots=rots, opac=opac, feats=feats, deltaT=np.float32(0.04))
for occl in (0, 1):
    r = oref.render_rays(o, d, t, means, ls, rots, opac, feats, o[0], 0, 1.0, 0.04, 1.0, bool(occl), return_filter=True)
    out[f"rho{occl}"], out[f"den{occl}"], out[f"tr{occl}"], out["filter"] = r
out["analytic"] = oref.render_rays_analytic(o, d, 0.35, 1.65, out["filter"], means, ls, rots, opac, feats, o[0])
np.savez_compressed(os.path.join(os.path.dirname(__file__), "ref_small.npz"), **out)
print({k: v.shape for k, v in out.items()})
