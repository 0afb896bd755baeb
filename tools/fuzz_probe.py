#!/usr/bin/env python
"""Details of randomised parity cases (tests/test_gpu_fuzz.py):  python tools/fuzz_probe.py 91 182 ..."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
from nlos_gaussian_renderer_b200.transient import TransientRenderer  # noqa: E402
from tests import util  # noqa: E402
from tests.test_gpu_fuzz import _case  # noqa: E402

for seed in map(int, sys.argv[1:]):
    N, Sy, Sx, T, A, scale_mul, window, bucket, slice_len = _case(seed)
    print("seed", seed, dict(N=N, Sy=Sy, Sx=Sx, T=T, A=A, scale_mul=round(scale_mul, 3), window=window, bucket=bucket, slice_len=slice_len))
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, seed=100 + seed, scale_mul=scale_mul, device="cuda")
    for mode, m in (("numerical", 0), ("analytic", 1)):
        H64 = util.oracle_H(geom, ta, params, mode=m, precision="f64")
        H32 = util.oracle_H(geom, ta, params, mode=m, precision="f32")
        for kw in (dict(window=window, bucket=bucket, slice_len=slice_len), dict(), dict(mcut=6.5), dict(mcut=8.0)):
            r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, **kw)
            H = r.forward(*params).cpu().numpy().reshape(Sy * Sx, T)
            d = np.abs(H - H64)
            p, k = np.unravel_index(d.argmax(), d.shape)
            print(f"  {mode:9s} {('mcut=%g' % kw['mcut']) if 'mcut' in kw else ('case' if kw else 'default'):8s} err/max {d.max() / H64.max():.3e} at p={p} k={k}: gpu {H[p, k]:.6e} f64 {H64[p, k]:.6e} f32 {H32[p, k]:.6e}"
                  f"  max {H64.max():.3e}  oracle32 err {np.abs(H32 - H64).max() / H64.max():.2e}")
