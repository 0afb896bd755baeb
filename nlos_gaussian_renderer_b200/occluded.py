"""Batched [scan_y, scan_x, time_bin] transient WITH occlusion (transmittance along each ray) -- native kernels (csrc/occluded.cu).

With occlusion the reference's semantics are per RAY, not per scan point: per-ray list of at most 256 Gaussians in index order
(src/ray_aabb.cu:10-61), alpha_g = 1 - exp(-pdf_g sigma(o_g) c dT), running transmittance over the samples of the ray with the early stop at
1e-4 (src/volume_renderer.cu:80-137), then /(r^2+1e-8) * sin(theta) and the angular sum (gaussian_model/cuda_autograd.py:302-314) times
volume_position[1]^2 (nlos_helpers.py:275-276).  The reference loops CUDARenderModule.forward over the scan points (main.py:198-203) and its
backward is zeros.  Here: per-ray lists for a block of scan points (O(N) per scan point, bit-identical to the reference's lists), one
forward launch that writes H[p, :] directly, one backward launch that recomputes every ray -- no [rays, T] tensor exists at any point.
Scan points are processed in blocks only to bound the list workspace (n_block * A^2 * 257 int32).
(The no-occlusion transient has its own, much faster, scan-tile kernels without the 256 cap: transient.TransientRenderer.)"""
import torch

from . import _lib
from ._C import _aligned, effective_sh_degree
from .geometry import ScanGeometry, TILE
from .transient import _f32c

MAX_GAUSSIANS_PER_RAY = 256     # src/ray_aabb.cu:6


def _blocks(n_rows, A, cap, workspace_bytes):
    per = A * A * (cap + 1) * 4
    n = max(TILE, int(workspace_bytes // per) // TILE * TILE)
    return [(s, min(n, n_rows - s)) for s in range(0, n_rows, n)]


class _OccludedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, geom, t_samples, means, log_scales, rotations, opacities, features, c, deltaT, mod, occ, cap, sh_degree, mcut, ws_bytes):
        L = _lib.lib()
        m, ls, rt = _f32c(means), _f32c(log_scales), _f32c(rotations)
        N = m.shape[0]
        op, ft = _f32c(opacities).reshape(-1), _f32c(features).reshape(N, -1)
        ts = _f32c(t_samples)
        T = ts.shape[0]
        dev = m.device
        deg = effective_sh_degree(sh_degree, ft.shape[1])
        records = _aligned(N * L.nlosb200_record_floats(), torch.float32, dev, 128)
        st = _lib.stream()
        _lib.call("nlosb200_preprocess", m.data_ptr(), ls.data_ptr(), rt.data_ptr(), op.data_ptr(), ft.data_ptr(), N, ft.shape[1], float(mod), 3.0,
                  float(mcut), 0.0, records.data_ptr(), None, None, st)
        n_rows = geom.n_tiles * TILE
        H = torch.zeros(n_rows, T, dtype=torch.float32, device=dev)
        t0 = float(ts[0])
        dt = float((ts[-1] - ts[0]) / max(T - 1, 1)) if T > 1 else 1.0
        blocks = _blocks(n_rows, geom.A, cap, ws_bytes)
        lists = torch.empty(L.nlosb200_occluded_list_ints(max(b[1] for b in blocks), geom.A, cap), dtype=torch.int32, device=dev)
        sh = (ft.data_ptr() if deg > 0 else None, ft.shape[1], deg)
        common = lambda sp0, n: (geom.tiles().data_ptr(), geom.n_tiles, geom.A, records.data_ptr(), N, cap, lists.data_ptr(), sp0, n)
        for sp0, n in blocks:
            _lib.call("nlosb200_occluded_lists", geom.tiles().data_ptr(), geom.n_tiles, geom.A, records.data_ptr(), N, cap, sp0, n, lists.data_ptr(), st)
            _lib.call("nlosb200_occluded_forward", *common(sp0, n), ts.data_ptr(), T, t0, dt, float(c) * float(deltaT), float(mcut), int(occ), *sh,
                      H.data_ptr(), st)
        ctx.save_for_backward(m, ls, rt, op, ft, ts)
        ctx.misc = (geom, float(c) * float(deltaT), float(mod), int(occ), cap, deg, float(mcut), t0, dt, blocks, opacities.shape, features.shape, ws_bytes)
        return geom.untile(H)

    @staticmethod
    def backward(ctx, gH):
        m, ls, rt, op, ft, ts = ctx.saved_tensors
        geom, cdt, mod, occ, cap, deg, mcut, t0, dt, blocks, oshape, fshape, ws_bytes = ctx.misc
        L = _lib.lib()
        N, T, dev = m.shape[0], ts.shape[0], m.device
        st = _lib.stream()
        records = _aligned(N * L.nlosb200_record_floats(), torch.float32, dev, 128)
        _lib.call("nlosb200_preprocess", m.data_ptr(), ls.data_ptr(), rt.data_ptr(), op.data_ptr(), ft.data_ptr(), N, ft.shape[1], mod, 3.0, mcut,
                  0.0, records.data_ptr(), None, None, st)
        g = geom.tile_grad(gH.to(torch.float32)).contiguous()
        dacc = torch.zeros(N, L.nlosb200_grad_floats(), dtype=torch.float32, device=dev)
        dsh = torch.zeros(N, 16, dtype=torch.float32, device=dev) if deg > 0 else None
        lists = torch.empty(L.nlosb200_occluded_list_ints(max(b[1] for b in blocks), geom.A, cap), dtype=torch.int32, device=dev)
        sh = (ft.data_ptr() if deg > 0 else None, ft.shape[1], deg)
        for sp0, n in blocks:
            _lib.call("nlosb200_occluded_lists", geom.tiles().data_ptr(), geom.n_tiles, geom.A, records.data_ptr(), N, cap, sp0, n, lists.data_ptr(), st)
            _lib.call("nlosb200_occluded_backward", geom.tiles().data_ptr(), geom.n_tiles, geom.A, records.data_ptr(), N, cap, lists.data_ptr(), sp0, n,
                      ts.data_ptr(), T, t0, dt, cdt, mcut, occ, *sh, g.data_ptr(), dacc.data_ptr(), dsh.data_ptr() if deg > 0 else None, st)
        outs = [torch.empty_like(m), torch.empty_like(ls), torch.empty_like(rt), torch.empty_like(op), torch.empty_like(ft)]
        if deg > 0:
            _lib.call("nlosb200_preprocess_backward_sh", ls.data_ptr(), rt.data_ptr(), op.data_ptr(), ft.data_ptr(), N, ft.shape[1], deg, mod,
                      dacc.data_ptr(), dsh.data_ptr(), *[o.data_ptr() for o in outs], 0, st)
        else:
            _lib.call("nlosb200_preprocess_backward", ls.data_ptr(), rt.data_ptr(), op.data_ptr(), ft.data_ptr(), N, ft.shape[1], mod, dacc.data_ptr(),
                      *[o.data_ptr() for o in outs], 0, st)
        return (None, None, outs[0], outs[1], outs[2], outs[3].reshape(oshape), outs[4].reshape(fshape)) + (None,) * 8


def render_transient_batch_occluded(geom: ScanGeometry, t_samples, means, log_scales, rotations, opacities, features, c, deltaT,
                                    scaling_modifier=1.0, use_occlusion=True, chunk_scan_points=None, max_per_ray=None, active_sh_degree=0,
                                    mcut=5.3, workspace_bytes=4 << 30):
    """-> H [Sy, Sx, T] fp32, differentiable w.r.t. the five parameter tensors.  t_samples must be the uniform radii (linspace) of the path.
    max_per_ray: the reference's 256-per-ray cap (None = 256); chunk_scan_points (multiple of 32) / workspace_bytes bound the list workspace.
    active_sh_degree 1..3: albedo per (Gaussian, scan point), features [N, >= (deg+1)^2]."""
    cap = MAX_GAUSSIANS_PER_RAY if max_per_ray is None else int(max_per_ray)
    if cap <= 0:
        cap = int(means.shape[0])
    ws = workspace_bytes if chunk_scan_points is None else max(1, int(chunk_scan_points)) * geom.A * geom.A * (cap + 1) * 4
    return _OccludedFn.apply(geom, t_samples, means, log_scales, rotations, opacities, features, float(c), float(deltaT), float(scaling_modifier),
                             bool(use_occlusion), cap, int(active_sh_degree), float(mcut), ws)
