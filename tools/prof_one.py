#!/usr/bin/env python
"""One prepare + forward + backward of a named config, for ncu captures:  python tools/prof_one.py cfg2 numerical [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from nlos_gaussian_renderer_b200 import scenes  # noqa: E402
from nlos_gaussian_renderer_b200.geometry import ScanGeometry  # noqa: E402
from nlos_gaussian_renderer_b200.transient import TransientRenderer  # noqa: E402

cfg, mode = sys.argv[1], sys.argv[2]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
kw = {}
rows = None
for a in sys.argv[4:]:
    k, v = a.split("=")
    if k == "rows":                       # rows=16: render only the first 16 scan rows of the grid (a slab, as one GPU of a sharded run would)
        rows = int(v)
    elif k == "variant":                  # variant=wide|narrow|auto
        kw[k] = v
    elif k == "narrow_rays":
        kw[k] = float(v)
    else:
        kw[k] = int(v)
N, Sy, Sx, T, A = scenes.CONFIGS[cfg]
dev = torch.device("cuda", 0)
geom = ScanGeometry.wall_grid(Sy, Sx, A=A, device=dev)
if rows:
    geom = geom.rows(0, rows)
ta = scenes.time_axis(T)
params = scenes.make_gaussians(N, 1000 + list(scenes.CONFIGS).index(cfg) + 1, device=dev)
r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, **kw)
for _ in range(reps):
    r.prepare(*params)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record()
    H = r.forward_tiled()
    ev[1].record()
    g = r.backward_tiled(torch.ones_like(H))
    ev[2].record()
    torch.cuda.synchronize()
    print(f"{cfg} {mode} {kw}: fwd {ev[0].elapsed_time(ev[1]):.2f} ms  bwd {ev[1].elapsed_time(ev[2]):.2f} ms  slice {r._slice} list {r.check_lists()}")
