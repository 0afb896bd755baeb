"""CPU oracle for the NLOS transient renderer -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (nlos_gaussian_renderer_b200) never does.

numpy in / numpy out wrappers over oracle/_build/libnlos_oracle.so (nlos_oracle.c).
`precision` is "f32" (faithful fp32 restatement of the reference CUDA kernels) or
"f64" (same formulas evaluated in double; culling stays fp32).
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(_build.build())
        _LIB.orc_det_expf.restype = C.c_float
        _LIB.orc_det_expf.argtypes = [C.c_float]
        _LIB.orc_num_threads.restype = C.c_int
    return _LIB


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(int(n)))


def det_expf(x):
    x = _f(x)
    out = np.empty_like(x)
    l = lib()
    flat, oflat = x.reshape(-1), out.reshape(-1)
    for i in range(flat.size):
        oflat[i] = l.orc_det_expf(C.c_float(float(flat[i])))
    return out


def bboxes(means, log_scales, rots, scaling_modifier=1.0, sigma=3.0, exp_mode=1):
    """bbox_compute.cuh:23-120 -> [N,6] (min xyz, max xyz)."""
    means, log_scales, rots = _f(means), _f(log_scales), _f(rots)
    N = means.shape[0]
    out = np.empty((N, 6), np.float32)
    lib().orc_bboxes(_p(means), _p(log_scales), _p(rots), C.c_int(N), C.c_float(scaling_modifier),
                     C.c_float(sigma), C.c_int(exp_mode), _p(out))
    return out


def filter_gaussians(ray_o, ray_d, bbox, cap=256):
    """ray_aabb.cu:10-102 -> int32 [n_rays, cap+1] (col 0 = count, unused = -1)."""
    ray_o, ray_d, bbox = _f(ray_o), _f(ray_d), _f(bbox).reshape(-1, 6)
    n_rays, N = ray_o.shape[0], bbox.shape[0]
    out = np.empty((n_rays, cap + 1), np.int32)
    lib().orc_filter(_p(ray_o), _p(ray_d), _p(bbox), C.c_int(n_rays), C.c_int(N), C.c_int(cap), _p(out))
    return out


def _rt(precision):
    return (np.float32, "orc32_") if precision == "f32" else (np.float64, "orc64_")


def render_rays(ray_o, ray_d, t_samples, means, log_scales, rots, opac, feats, cam,
                sh_degree=0, c=1.0, deltaT=0.01, scaling_modifier=1.0, use_occlusion=True,
                cap=256, mode=0, exp_mode=1, precision="f32", filter_lists=None):
    """volume_renderer.cu:189-305.  cap<=0: every Gaussian passing the slab test (no 256 cap).
    Returns rho_density, density, transmittance each [n_rays, T]."""
    ray_o, ray_d, t_samples = _f(ray_o), _f(ray_d), _f(t_samples)
    means, log_scales, rots, opac, feats, cam = _f(means), _f(log_scales), _f(rots), _f(opac).reshape(-1), _f(feats), _f(cam)
    n_rays, T, N = ray_o.shape[0], t_samples.shape[0], means.shape[0]
    feats = feats.reshape(N, -1)
    dt, pre = _rt(precision)
    bb = bboxes(means, log_scales, rots, scaling_modifier, 3.0, exp_mode)
    if cap > 0 and filter_lists is None:
        filter_lists = filter_gaussians(ray_o, ray_d, bb, cap)
    if filter_lists is not None:
        filter_lists = np.ascontiguousarray(filter_lists, np.int32)
        cap = filter_lists.shape[1] - 1
    outs = [np.zeros((n_rays, T), dt) for _ in range(3)]
    getattr(lib(), pre + "render_rays")(
        _p(ray_o), _p(ray_d), _p(t_samples), _p(filter_lists), C.c_int(cap), _p(bb),
        _p(means), _p(log_scales), _p(rots), _p(opac), _p(feats), _p(cam),
        C.c_int(n_rays), C.c_int(T), C.c_int(N), C.c_int(sh_degree), C.c_int(feats.shape[1]),
        C.c_float(c), C.c_float(deltaT), C.c_float(scaling_modifier), C.c_int(int(use_occlusion)),
        C.c_int(mode), C.c_int(exp_mode), _p(outs[0]), _p(outs[1]), _p(outs[2]))
    return tuple(outs)


def render_rays_analytic(ray_o, ray_d, t_min, t_max, filter_lists, means, log_scales, rots, opac, feats, cam,
                         sh_degree=0, scaling_modifier=1.0, sigma_threshold=3.0, max_sections=128,
                         exp_mode=1, precision="f32"):
    """volume_renderer_analytic.cu:23-241 -> [n_rays]."""
    ray_o, ray_d = _f(ray_o), _f(ray_d)
    means, log_scales, rots, opac, feats, cam = _f(means), _f(log_scales), _f(rots), _f(opac).reshape(-1), _f(feats), _f(cam)
    n_rays, N = ray_o.shape[0], means.shape[0]
    feats = feats.reshape(N, -1)
    filter_lists = np.ascontiguousarray(filter_lists, np.int32)
    cap = filter_lists.shape[1] - 1
    dt, pre = _rt(precision)
    out = np.zeros((n_rays,), dt)
    getattr(lib(), pre + "render_rays_analytic")(
        _p(ray_o), _p(ray_d), C.c_float(t_min), C.c_float(t_max), _p(filter_lists), C.c_int(cap),
        _p(means), _p(log_scales), _p(rots), _p(opac), _p(feats), _p(cam),
        C.c_int(n_rays), C.c_int(N), C.c_int(sh_degree), C.c_int(feats.shape[1]),
        C.c_float(scaling_modifier), C.c_float(sigma_threshold), C.c_int(max_sections), C.c_int(exp_mode), _p(out))
    return out


def transient_batch(scan_o, sin_th, cos_th, sin_ph, cos_ph, t_samples, hscale,
                    means, log_scales, rots, opac, feats, sh_degree=0, c=1.0, deltaT=0.01,
                    scaling_modifier=1.0, use_occlusion=False, mode=0, cap=0, exp_mode=1,
                    precision="f32", count=False):
    """Batched transient H[S,T]; per-scan-point slice == cuda_autograd.py:254-314 x vp_y^2.
    Returns H (and counters dict if count)."""
    scan_o, t_samples, hscale = _f(scan_o).reshape(-1, 3), _f(t_samples), _f(hscale).reshape(-1)
    S, T = scan_o.shape[0], t_samples.shape[0]
    sin_th, cos_th, sin_ph, cos_ph = [_f(a).reshape(S, -1) for a in (sin_th, cos_th, sin_ph, cos_ph)]
    A = sin_th.shape[1]
    means, log_scales, rots, opac, feats = _f(means), _f(log_scales), _f(rots), _f(opac).reshape(-1), _f(feats)
    N = means.shape[0]
    feats = feats.reshape(N, -1)
    dt, pre = _rt(precision)
    H = np.zeros((S, T), dt)
    cnt = np.zeros(4, np.int64) if count else None
    getattr(lib(), pre + "transient_batch")(
        _p(scan_o), _p(sin_th), _p(cos_th), _p(sin_ph), _p(cos_ph), _p(t_samples), _p(hscale),
        _p(means), _p(log_scales), _p(rots), _p(opac), _p(feats),
        C.c_int(S), C.c_int(A), C.c_int(T), C.c_int(N), C.c_int(sh_degree), C.c_int(feats.shape[1]),
        C.c_float(c), C.c_float(deltaT), C.c_float(scaling_modifier), C.c_int(int(use_occlusion)),
        C.c_int(mode), C.c_int(cap), C.c_int(exp_mode), _p(H), _p(cnt))
    if count:
        return H, {"hits": int(cnt[0]), "ref_sample_evals": int(cnt[1]), "W_s": int(cnt[2]), "W_b": int(cnt[3])}
    return H


def shell_intervals(scan_o, Sy, Sx, tile_y, tile_x, means, log_scales, scaling_modifier, mcut,
                    t0, dt, T, bins_per_range, exp_mode=1):
    scan_o, means, log_scales = _f(scan_o).reshape(-1, 3), _f(means), _f(log_scales)
    N = means.shape[0]
    nt = ((Sy + tile_y - 1) // tile_y) * ((Sx + tile_x - 1) // tile_x)
    lo = np.empty((nt, N), np.int32)
    hi = np.empty((nt, N), np.int32)
    lib().orc_shell_intervals(_p(scan_o), C.c_int(Sy), C.c_int(Sx), C.c_int(tile_y), C.c_int(tile_x),
                              _p(means), _p(log_scales), C.c_int(N), C.c_float(scaling_modifier), C.c_float(mcut),
                              C.c_int(exp_mode), C.c_float(t0), C.c_float(dt), C.c_int(T), C.c_int(bins_per_range),
                              _p(lo), _p(hi))
    return lo, hi


def shell_lists(lo, hi, n_ranges):
    """-> counts [n_tiles*n_ranges], offsets (exclusive scan, int64, len cells+1), list (ascending g per cell)."""
    lo = np.ascontiguousarray(lo, np.int32)
    hi = np.ascontiguousarray(hi, np.int32)
    nt, N = lo.shape
    counts = np.zeros(nt * n_ranges, np.int32)
    lib().orc_shell_lists(_p(lo), _p(hi), C.c_int(nt), C.c_int(n_ranges), C.c_int(N), _p(counts), None, None)
    offsets = np.zeros(nt * n_ranges + 1, np.int64)
    np.cumsum(counts, out=offsets[1:])
    lst = np.empty(int(offsets[-1]), np.int32)
    c2 = np.zeros_like(counts)
    lib().orc_shell_lists(_p(lo), _p(hi), C.c_int(nt), C.c_int(n_ranges), C.c_int(N), _p(c2), _p(offsets), _p(lst))
    return counts, offsets, lst


def shell_bins(scan_o, Sy, Sx, tile_y, tile_x, means, log_scales, scaling_modifier, mcut, t0, dt, T, exp_mode=1):
    """Tile-level shell of every Gaussian in bin units -> (klo, khi) int32 [n_tiles, N]; klo > khi = empty."""
    scan_o, means, log_scales = _f(scan_o).reshape(-1, 3), _f(means), _f(log_scales)
    N = means.shape[0]
    nt = ((Sy + tile_y - 1) // tile_y) * ((Sx + tile_x - 1) // tile_x)
    lo = np.empty((nt, N), np.int32)
    hi = np.empty((nt, N), np.int32)
    lib().orc_shell_bins(_p(scan_o), C.c_int(Sy), C.c_int(Sx), C.c_int(tile_y), C.c_int(tile_x), _p(means), _p(log_scales),
                         C.c_int(N), C.c_float(scaling_modifier), C.c_float(mcut), C.c_int(exp_mode), C.c_float(t0), C.c_float(dt),
                         C.c_int(T), _p(lo), _p(hi))
    return lo, hi


ENTRY_SHIFT = 26      # list entry = gaussian id | (sub-interval index << 26)


def entry_classes(scan_o, valid, lattice_step, Sy, Sx, tile_y, tile_x, means, log_scales, scaling_modifier, mcut, klo, khi, window, bucket,
                  narrow_rays, exp_mode=1, time_wide_min=0.0, T=None):
    """DEFINITION of the entry classes of the window binning (csrc/binning.cu header): class 0 ("narrow") when the Gaussian is listed once
    in the tile (khi - klo + 1 <= L) and  (6 / mcut) rad <= (narrow_rays * pitch) * sqrt(d2min)  in single-rounded fp32, with
    rad = mcut * max(scale), pitch = min over the tile's valid scan points of lattice_step (= min(dtheta, dphi), 0 where the lattice is
    unusable) and d2min = min over the tile's valid scan points of |mu - o|^2 (summed as ((dx^2 + dy^2) + dz^2)); else class 1.
    narrow_rays = 0: no footprint classes (bit 0 everywhere).  time_wide_min > 0 adds the time-wide bit (class += n_fp): smallest scale >=
    time_wide_min, listed once in the tile and klo >= 3, khi <= T - 4.  scan_o [Sy*Sx,3], lattice_step [Sy*Sx] -> int32 [n_tiles, N]."""
    f = np.float32
    scan_o, means, log_scales = _f(scan_o).reshape(Sy, Sx, 3), _f(means), _f(log_scales)
    step = _f(lattice_step).reshape(Sy, Sx)
    N = means.shape[0]
    nty, ntx = -(-Sy // tile_y), -(-Sx // tile_x)
    L = window - bucket
    s = det_expf(log_scales) if exp_mode == 1 else np.exp(log_scales).astype(f)
    s = (s * f(scaling_modifier)).astype(f)
    rad = (f(mcut) * s.max(1)).astype(f)
    lhs = (rad * (f(6.0) / f(mcut))).astype(f)
    out = np.ones((nty * ntx, N), np.int32)
    for ty in range(nty):
        for tx in range(ntx):
            o = scan_o[ty * tile_y:(ty + 1) * tile_y, tx * tile_x:(tx + 1) * tile_x].reshape(-1, 3)
            st = step[ty * tile_y:(ty + 1) * tile_y, tx * tile_x:(tx + 1) * tile_x].reshape(-1)
            d = (means[:, None, :] - o[None, :, :]).astype(f)
            d2 = ((d[..., 0] * d[..., 0]).astype(f) + (d[..., 1] * d[..., 1]).astype(f)).astype(f)
            d2 = (d2 + (d[..., 2] * d[..., 2]).astype(f)).astype(f)
            rhs = ((f(narrow_rays) * st.min()).astype(f) * np.sqrt(d2.min(1)).astype(f)).astype(f)
            t = ty * ntx + tx
            out[t] = np.where((khi[t] - klo[t] + 1 <= L) & (lhs <= rhs), 0, 1) if narrow_rays > 0 else 0
    n_fp = 2 if narrow_rays > 0 else 1
    if time_wide_min > 0:
        tw = s.min(1) >= f(time_wide_min)
        out = out + n_fp * (tw[None, :] & (klo >= 3) & (khi <= T - 4) & (khi - klo + 1 <= L)).astype(np.int32)
    return out


def window_lists(klo, khi, T, window, bucket, cls=None, n_classes=None):
    """DEFINITION of the window-binning lists the batched renderer consumes.  A Gaussian whose tile-level shell is
    [klo, khi] contributes entries e = 0 .. ceil(len/L)-1 with L = window - bucket; entry e covers bins
    [klo + e L, min(khi, klo + (e+1) L - 1)] and lives in cell (tile, (klo + e L) // bucket).  Per cell the entries are
    ascending in the Gaussian id.  With entry classes (cls int [n_tiles, N] of 0 / 1, entry_classes()) the cells of a tile are class-major:
    cell = tile * n_classes n_buckets + cls * n_buckets + bucket.  -> counts [cells] int32, offsets int64 [cells+1], list uint32."""
    klo = np.asarray(klo, np.int64)
    khi = np.asarray(khi, np.int64)
    nt, N = klo.shape
    L = window - bucket
    nb = -(-T // bucket)
    ncls = 1 if cls is None else (2 if n_classes is None else int(n_classes))
    n_e = np.where(klo <= khi, -(-(khi - klo + 1) // L), 0)
    cells, vals = [], []
    for e in range(int(n_e.max()) if n_e.size else 0):
        t, g = np.nonzero(n_e > e)
        s = klo[t, g] + e * L
        c = 0 if cls is None else np.asarray(cls, np.int64)[t, g]
        cells.append(t * (ncls * nb) + c * nb + s // bucket)
        vals.append(g | (e << ENTRY_SHIFT))
    if cells:
        cells, vals = np.concatenate(cells), np.concatenate(vals)
        order = np.lexsort((vals & ((1 << ENTRY_SHIFT) - 1), cells))
        cells, vals = cells[order], vals[order]
    else:
        cells, vals = np.zeros(0, np.int64), np.zeros(0, np.int64)
    counts = np.bincount(cells, minlength=nt * nb * ncls).astype(np.int32)
    offsets = np.zeros(nt * nb * ncls + 1, np.int64)
    np.cumsum(counts, out=offsets[1:])
    return counts, offsets, vals.astype(np.uint32)


def knn_dist2(points):
    points = _f(points).reshape(-1, 3)
    out = np.empty(points.shape[0], np.float32)
    lib().orc_knn_dist2(_p(points), C.c_int(points.shape[0]), _p(out))
    return out
