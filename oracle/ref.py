"""ctypes wrapper over oracle/_ref/libnlos_ref.so = the reference's OWN kernel source compiled for the
host (see build_ref.py).  TEST INFRASTRUCTURE ONLY.  `available()` is False when neither
/root/reference nor a prebuilt library exists."""
import ctypes as C

import numpy as np

from . import build_ref as _b

_LIB = None


def available() -> bool:
    return _b.build() is not None


def lib():
    global _LIB
    if _LIB is None:
        path = _b.build()
        if path is None:
            raise RuntimeError("oracle/_ref not built and /root/reference not present")
        _LIB = C.CDLL(path)
    return _LIB


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def cap():
    return int(lib().ref_max_gaussians_per_ray())


def num_threads():
    return int(lib().ref_num_threads())


def set_num_threads(n):
    lib().ref_set_num_threads(C.c_int(int(n)))


def bboxes(means, log_scales, rots, scaling_modifier=1.0, sigma=3.0):
    means, log_scales, rots = _f(means), _f(log_scales), _f(rots)
    N = means.shape[0]
    out = np.empty((N, 6), np.float32)
    lib().ref_bboxes(_p(means), _p(log_scales), _p(rots), C.c_int(N), C.c_float(scaling_modifier), C.c_float(sigma), _p(out))
    return out


def filter_gaussians(ray_o, ray_d, bbox):
    ray_o, ray_d, bbox = _f(ray_o), _f(ray_d), _f(bbox).reshape(-1, 6)
    out = np.empty((ray_o.shape[0], cap() + 1), np.int32)
    lib().ref_filter(_p(ray_o), _p(ray_d), _p(bbox), C.c_int(ray_o.shape[0]), C.c_int(bbox.shape[0]), _p(out))
    return out


def render_rays(ray_o, ray_d, t_samples, means, log_scales, rots, opac, feats, cam,
                sh_degree=0, c=1.0, deltaT=0.01, scaling_modifier=1.0, use_occlusion=True,
                rendering_type="netf", return_filter=False, filter_lists=None):
    """filter_lists int32 [n_rays, 257]: render with these per-ray lists instead of filtering (frozen membership)."""
    ray_o, ray_d, t_samples = _f(ray_o), _f(ray_d), _f(t_samples)
    means, log_scales, rots, opac, feats, cam = _f(means), _f(log_scales), _f(rots), _f(opac).reshape(-1), _f(feats), _f(cam)
    n_rays, T, N = ray_o.shape[0], t_samples.shape[0], means.shape[0]
    feats = feats.reshape(N, -1)
    outs = [np.empty((n_rays, T), np.float32) for _ in range(3)]
    if filter_lists is not None:
        fl = np.ascontiguousarray(filter_lists, np.int32)
        assert fl.shape == (n_rays, cap() + 1)
        lib().ref_render_rays_with_filter(_p(ray_o), _p(ray_d), _p(t_samples), _p(means), _p(log_scales), _p(rots), _p(opac),
                                          _p(feats), _p(cam), C.c_int(n_rays), C.c_int(T), C.c_int(N), C.c_int(sh_degree),
                                          C.c_int(feats.shape[1]), C.c_float(c), C.c_float(deltaT), C.c_float(scaling_modifier),
                                          C.c_int(int(use_occlusion)), C.c_int(0 if rendering_type == "netf" else 1), _p(fl),
                                          _p(outs[0]), _p(outs[1]), _p(outs[2]))
        return tuple(outs)
    filt = np.empty((n_rays, cap() + 1), np.int32) if return_filter else None
    lib().ref_render_rays(_p(ray_o), _p(ray_d), _p(t_samples), _p(means), _p(log_scales), _p(rots), _p(opac),
                          _p(feats), _p(cam), C.c_int(n_rays), C.c_int(T), C.c_int(N), C.c_int(sh_degree),
                          C.c_int(feats.shape[1]), C.c_float(c), C.c_float(deltaT), C.c_float(scaling_modifier),
                          C.c_int(int(use_occlusion)), C.c_int(0 if rendering_type == "netf" else 1),
                          _p(outs[0]), _p(outs[1]), _p(outs[2]), _p(filt))
    return tuple(outs) + ((filt,) if return_filter else ())


def render_rays_analytic(ray_o, ray_d, t_min, t_max, filter_lists, means, log_scales, rots, opac, feats, cam,
                         sh_degree=0, c=1.0, deltaT=0.01, scaling_modifier=1.0, sigma_threshold=3.0,
                         rendering_type="netf"):
    ray_o, ray_d = _f(ray_o), _f(ray_d)
    means, log_scales, rots, opac, feats, cam = _f(means), _f(log_scales), _f(rots), _f(opac).reshape(-1), _f(feats), _f(cam)
    n_rays, N = ray_o.shape[0], means.shape[0]
    feats = feats.reshape(N, -1)
    filter_lists = np.ascontiguousarray(filter_lists, np.int32)
    assert filter_lists.shape[1] == cap() + 1
    out = np.empty((n_rays,), np.float32)
    lib().ref_render_rays_analytic(_p(ray_o), _p(ray_d), C.c_float(t_min), C.c_float(t_max), _p(filter_lists),
                                   _p(means), _p(log_scales), _p(rots), _p(opac), _p(feats), _p(cam),
                                   C.c_int(n_rays), C.c_int(N), C.c_int(sh_degree), C.c_int(feats.shape[1]),
                                   C.c_float(c), C.c_float(deltaT), C.c_float(scaling_modifier),
                                   C.c_float(sigma_threshold), C.c_int(0 if rendering_type == "netf" else 1), _p(out))
    return out
