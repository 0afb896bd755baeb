#!/usr/bin/env python
"""One small forward + backward of the batched renderer (both integration modes, both scan-tile kernels, entry classes, SH degree 2) and of the
batched occlusion renderer -- the program run under `compute-sanitizer --tool racecheck|memcheck` (one tool per gpurun call):
    compute-sanitizer --tool racecheck python tools/sanitize_one.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from nlos_gaussian_renderer_b200 import scenes  # noqa: E402
from nlos_gaussian_renderer_b200.geometry import ScanGeometry  # noqa: E402
from nlos_gaussian_renderer_b200.occluded import render_transient_batch_occluded  # noqa: E402
from nlos_gaussian_renderer_b200.transient import TransientRenderer, render_transient_batch  # noqa: E402

dev = torch.device("cuda", 0)
N, Sy, Sx, T, A = 600, 8, 16, 128, 8
geom = ScanGeometry.wall_grid(Sy, Sx, A=A, device=dev)
ta = scenes.time_axis(T)
base = scenes.make_gaussians(N, 1001, device=dev, scale_mul=float(sys.argv[1]) if len(sys.argv) > 1 else 1.0)
for mode in ("numerical", "analytic"):
    for variant, deg in (("wide", 0), ("narrow", 0), ("auto", 2)):
        params = [p.clone().requires_grad_() for p in base]
        if deg:
            params[4] = (torch.randn(N, 16, device=dev) * 0.3).requires_grad_()
        r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, variant=variant, active_sh_degree=deg, window=64, bucket=8)
        H = render_transient_batch(r, *params)
        (H * H).sum().backward()
        torch.cuda.synchronize()
        print(mode, variant, deg, float(H.sum()), float(params[0].grad.abs().max()))
params = [p.clone().requires_grad_() for p in base]
H = render_transient_batch_occluded(geom, ta["t_samples"].to(dev), *params, ta["c"], 0.3)
(H * H).sum().backward()
torch.cuda.synchronize()
print("occluded", float(H.sum()), float(params[0].grad.abs().max()))
