"""Mirror of the reference's pybind module `nlos_gaussian_renderer._C`
(submodules/cuda_renderer/src/bindings.cpp:26-35): same three function names, positional arguments,
return types and error behaviour (RuntimeError for non-CUDA / non-contiguous tensors, cuda_utils.cuh:7-9),
implemented by libnlosb200.so through the C ABI (include/nlosb200.h).  torch only allocates the outputs.

Differences from the reference, all documented in DESIGN.md:
  * no cudaDeviceSynchronize, launches go to torch's current stream;
  * active_sh_degree > 0 with too few feature columns evaluates the highest degree the columns support, with ONE warning (the reference
    reads out of bounds there; its own caller passes the dc column only, cuda_autograd.py:280);
  * `max_per_ray` / `max_sections` keyword extras (defaults = the reference's 256 / 128).
"""
import torch

from . import _lib

MAX_GAUSSIANS_PER_RAY = 256     # src/ray_aabb.cu:6
MAX_SECTIONS_PER_RAY = 128      # src/volume_renderer_analytic.cu:11


def _check(t, name, dtype=torch.float32):
    """CHECK_INPUT of include/cuda_utils.cuh:7-9 + the dtype test data_ptr<float>() / data_ptr<int>() performs in the reference."""
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if t.dtype != dtype:
        raise RuntimeError(f"expected scalar type {'Float' if dtype == torch.float32 else 'Int'} but found {t.dtype} ({name})")
    return t


_warned_sh = False


def effective_sh_degree(active_sh_degree, n_coeffs):
    """The reference's own caller passes ONLY the dc coefficient (cuda_autograd.py:280) while main.py:240-241 raises active_sh_degree almost
    every iteration, so its kernel reads past the feature row (volume_renderer.cu:110, spherical_harmonics.cuh:72-77).  Documented choice here:
    evaluate the highest degree the passed coefficients support (degree 0 for a dc-only row) and say so once."""
    global _warned_sh
    deg = int(active_sh_degree)
    sup = 0
    while sup < 3 and (sup + 2) ** 2 <= n_coeffs:
        sup += 1
    if deg > sup:
        if not _warned_sh:
            import warnings
            warnings.warn(f"active_sh_degree={deg} but the features hold {n_coeffs} SH coefficient(s) per Gaussian: evaluating degree {sup} "
                          "(the reference reads out of bounds here, volume_renderer.cu:110); pass get_features ([N,(deg+1)^2]) for view-dependent albedo",
                          RuntimeWarning, stacklevel=3)
            _warned_sh = True
        return sup
    return max(deg, 0)


def _aligned(numel, dtype, device, align=128):
    """1-D tensor whose data_ptr is `align`-byte aligned (TMA bulk copies need 16, records 128)."""
    item = torch.empty((), dtype=dtype).element_size()
    buf = torch.empty(numel + align // item, dtype=dtype, device=device)
    off = ((-buf.data_ptr()) % align) // item
    return buf[off: off + numel]


def filter_gaussians_per_ray(ray_origins, ray_directions, gaussian_means, gaussian_bboxes, sigma_threshold, max_per_ray=None):
    """-> int32 [N_rays, 257]; [:,0] = count, then ascending Gaussian ids, unused = -1 (src/ray_aabb.cu:63-102).
    gaussian_means and sigma_threshold are accepted and unused, exactly like the reference."""
    _check(ray_origins, "ray_origins"); _check(ray_directions, "ray_directions"); _check(gaussian_bboxes, "gaussian_bboxes")
    cap = MAX_GAUSSIANS_PER_RAY if max_per_ray is None else int(max_per_ray)
    n_rays, N = ray_origins.shape[0], gaussian_bboxes.reshape(-1, 6).shape[0]
    bb = gaussian_bboxes.reshape(-1)
    if bb.data_ptr() % 16:
        a = _aligned(bb.numel(), torch.float32, bb.device, 16)
        a.copy_(bb)
        bb = a
    out = torch.empty(n_rays, cap + 1, dtype=torch.int32, device=ray_origins.device)
    _lib.call("nlosb200_filter_gaussians_per_ray", ray_origins.data_ptr(), ray_directions.data_ptr(), bb.data_ptr(), n_rays, N, cap,
              out.data_ptr(), _lib.stream())
    return out


class _Prepared:
    """records / bboxes / filter lists of one render_rays call (kept by the autograd function for the backward)."""
    __slots__ = ("records", "bboxes", "filter", "cap", "N", "sh_dim", "sh_degree", "camera_pos", "means")


def _prepare(ray_origins, ray_directions, means, scales, rotations, opacities, features, camera_pos, sh_degree, scaling_modifier, cap):
    N = means.shape[0]
    dev = means.device
    L = _lib.lib()
    P = _Prepared()
    P.N, P.cap = N, cap
    feats = features.reshape(N, -1) if N else features.reshape(0, 1)
    P.sh_dim = feats.shape[1]
    sh_degree = P.sh_degree = effective_sh_degree(sh_degree, P.sh_dim)
    P.records = _aligned(max(N, 1) * L.nlosb200_record_floats(), torch.float32, dev, 128)
    P.bboxes = _aligned(max(N, 1) * 6, torch.float32, dev, 16)
    st = _lib.stream()
    if N:
        # sigma is hard-coded 3.0 inside render_rays (volume_renderer.cu:232,244); mcut unused here
        _lib.call("nlosb200_preprocess", means.data_ptr(), scales.data_ptr(), rotations.data_ptr(), opacities.data_ptr(), feats.data_ptr(),
                  N, P.sh_dim, float(scaling_modifier), 3.0, 0.0, 0.0, P.records.data_ptr(), None, P.bboxes.data_ptr(), st)
        _lib.call("nlosb200_compat_albedo", means.data_ptr(), feats.data_ptr(), camera_pos.data_ptr(), N, P.sh_dim, int(sh_degree),
                  P.records.data_ptr(), st)
    P.camera_pos, P.means = camera_pos, means
    n_rays = ray_origins.shape[0]
    P.filter = torch.empty(n_rays, cap + 1, dtype=torch.int32, device=dev)
    _lib.call("nlosb200_filter_gaussians_per_ray", ray_origins.data_ptr(), ray_directions.data_ptr(), P.bboxes.data_ptr(), n_rays, N, cap,
              P.filter.data_ptr(), st)
    return P


def render_rays(ray_origins, ray_directions, t_samples, gaussian_means, gaussian_scales, gaussian_rotations, gaussian_opacities,
                gaussian_features, camera_pos, active_sh_degree, c, deltaT, scaling_modifier, use_occlusion, rendering_type,
                max_per_ray=None, _return_prepared=False):
    """-> (rho_density, density, transmittance), each [N_rays, N_samples] (src/volume_renderer.cu:189-305).
    `rendering_type` is accepted and never read, like the reference kernel."""
    for t, n in ((ray_origins, "ray_origins"), (ray_directions, "ray_directions"), (t_samples, "t_samples"),
                 (gaussian_means, "gaussian_means"), (gaussian_scales, "gaussian_scales"), (gaussian_rotations, "gaussian_rotations"),
                 (gaussian_opacities, "gaussian_opacities"), (gaussian_features, "gaussian_features"), (camera_pos, "camera_pos")):
        _check(t, n)
    cap = MAX_GAUSSIANS_PER_RAY if max_per_ray is None else int(max_per_ray)
    if cap <= 0:
        cap = max(int(gaussian_means.shape[0]), 1)          # "no cap": every Gaussian passing the slab test
    P = _prepare(ray_origins, ray_directions, gaussian_means, gaussian_scales, gaussian_rotations, gaussian_opacities,
                 gaussian_features, camera_pos, active_sh_degree, scaling_modifier, cap)
    n_rays, T = ray_origins.shape[0], t_samples.shape[0]
    outs = [torch.empty(n_rays, T, dtype=torch.float32, device=ray_origins.device) for _ in range(3)]
    _lib.call("nlosb200_render_rays", ray_origins.data_ptr(), ray_directions.data_ptr(), t_samples.data_ptr(), P.records.data_ptr(),
              P.filter.data_ptr(), cap, n_rays, T, P.N, float(c), float(deltaT), int(bool(use_occlusion)),
              outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(), _lib.stream())
    if _return_prepared:
        return tuple(outs), P
    return tuple(outs)


def render_rays_backward(P, ray_origins, ray_directions, t_samples, gaussian_scales, gaussian_rotations, gaussian_opacities,
                         gaussian_features, c, deltaT, scaling_modifier, use_occlusion, rho_density, transmittance, grad_rho, grad_density,
                         grad_transmittance):
    """Gradients of (rho_density, density, transmittance) w.r.t. (means, scales, rotations, opacities, features)
    -- the kernel call commented out in gaussian_model/cuda_autograd.py:139-172.  active_sh_degree >= 1: the albedo's dependence on the SH
    coefficients and on the mean (through the view direction) is chained by nlosb200_compat_albedo_backward."""
    N, dev = P.N, ray_origins.device
    n_rays, T = ray_origins.shape[0], t_samples.shape[0]
    L = _lib.lib()
    dacc = torch.empty(max(N, 1), L.nlosb200_grad_floats(), dtype=torch.float32, device=dev)
    gp = lambda g: None if g is None else g.contiguous().data_ptr()
    keep = [g.contiguous() for g in (grad_rho, grad_density, grad_transmittance) if g is not None]   # noqa: F841 (lifetime)
    st = _lib.stream()
    _lib.call("nlosb200_render_rays_backward", ray_origins.data_ptr(), ray_directions.data_ptr(), t_samples.data_ptr(),
              P.records.data_ptr(), P.filter.data_ptr(), P.cap, n_rays, T, N, float(c), float(deltaT), int(bool(use_occlusion)),
              rho_density.data_ptr(), transmittance.data_ptr(), gp(grad_rho), gp(grad_density), gp(grad_transmittance),
              dacc.data_ptr(), st)
    feats = gaussian_features.reshape(N, -1)
    opac = gaussian_opacities.reshape(-1)
    outs = [torch.empty(N, 3, dtype=torch.float32, device=dev), torch.empty(N, 3, dtype=torch.float32, device=dev),
            torch.empty(N, 4, dtype=torch.float32, device=dev), torch.empty(N, dtype=torch.float32, device=dev),
            torch.empty(N, feats.shape[1], dtype=torch.float32, device=dev)]
    if N:
        _lib.call("nlosb200_preprocess_backward", gaussian_scales.data_ptr(), gaussian_rotations.data_ptr(), opac.data_ptr(),
                  feats.data_ptr(), N, feats.shape[1], float(scaling_modifier), dacc.data_ptr(), *[o.data_ptr() for o in outs], 0, st)
        if P.sh_degree > 0:
            _lib.call("nlosb200_compat_albedo_backward", P.means.data_ptr(), feats.data_ptr(), P.camera_pos.data_ptr(), N, feats.shape[1],
                      P.sh_degree, dacc.data_ptr(), outs[0].data_ptr(), outs[4].data_ptr(), st)
    return outs[0], outs[1], outs[2], outs[3].reshape(gaussian_opacities.shape), outs[4].reshape(gaussian_features.shape)


def render_rays_analytic(ray_origins, ray_directions, t_min, t_max, gaussian_filter, gaussian_means, gaussian_scales, gaussian_rotations,
                         gaussian_opacities, gaussian_features, camera_pos, active_sh_degree, c, deltaT, scaling_modifier,
                         sigma_threshold, rendering_type, max_sections=None):
    """-> [N_rays] (src/volume_renderer_analytic.cu:178-241).  c, deltaT and rendering_type are accepted and unused,
    like the reference kernel."""
    _check(ray_origins, "ray_origins"); _check(ray_directions, "ray_directions"); _check(gaussian_means, "gaussian_means")
    _check(gaussian_filter, "gaussian_filter", torch.int32)
    for t, n in ((gaussian_scales, "gaussian_scales"), (gaussian_rotations, "gaussian_rotations"),
                 (gaussian_opacities, "gaussian_opacities"), (gaussian_features, "gaussian_features"), (camera_pos, "camera_pos")):
        _check(t, n)

    n_rays, N = ray_origins.shape[0], gaussian_means.shape[0]
    cap = gaussian_filter.shape[1] - 1
    ms = MAX_SECTIONS_PER_RAY if max_sections is None else int(max_sections)
    feats = gaussian_features.reshape(N, -1) if N else gaussian_features.reshape(0, 1)
    out = torch.zeros(n_rays, dtype=torch.float32, device=ray_origins.device)
    _lib.call("nlosb200_render_rays_analytic", ray_origins.data_ptr(), ray_directions.data_ptr(), float(t_min), float(t_max),
              gaussian_filter.data_ptr(), cap, gaussian_means.data_ptr(), gaussian_scales.data_ptr(), gaussian_rotations.data_ptr(),
              gaussian_opacities.data_ptr(), feats.data_ptr(), camera_pos.data_ptr(), n_rays, N, feats.shape[1],
              effective_sh_degree(active_sh_degree, feats.shape[1]),
              float(scaling_modifier), float(sigma_threshold), ms, out.data_ptr(), _lib.stream())
    return out


def evaluate_points(points, gaussian_means, gaussian_scales, gaussian_rotations, gaussian_opacities, gaussian_features, camera_pos=None,
                    active_sh_degree=0, scaling_modifier=1.0, semantics="cuda"):
    """Dense evaluation at `points` [Na,3] -> (rho_density [Na], density [Na]): sums over ALL Gaussians, the quantities
    GaussianModel.estimate_rho_w_no_occlusion(..., out_separately=True) returns (gaussian_model/gaussian_model.py:346-364) and
    gaussian2volume samples (nlos_helpers.py:40-69), with the CUDA renderer's pdf arithmetic (include/cuda_utils.cuh:124-151).
    camera_pos None -> degree-0 albedo; else the view-dependent albedo of that position (volume_renderer.cu:107-111).
    semantics="dense": the arithmetic gaussian2volume really evaluates in the reference -- GaussianModel.estimate_gaussian_pdf
    (gaussian_model.py:253-294): scale = exp(exp(_scaling) * mod), R (x - mu), no epsilon (SURVEY App. A.1 column a), degree-0 albedo --
    so that a model fitted with the reference's dense path is sampled with ITS pdf; "cuda" (default) = the CUDA renderer's pdf, i.e.
    what this library's renderers fit."""
    if semantics not in ("cuda", "dense"):
        raise ValueError("semantics must be 'cuda' or 'dense'")
    for t, n in ((points, "points"), (gaussian_means, "gaussian_means"), (gaussian_scales, "gaussian_scales"),
                 (gaussian_rotations, "gaussian_rotations"), (gaussian_opacities, "gaussian_opacities"), (gaussian_features, "gaussian_features")):
        _check(t, n)
    N, dev = gaussian_means.shape[0], gaussian_means.device
    L = _lib.lib()
    feats = gaussian_features.reshape(N, -1) if N else gaussian_features.reshape(0, 1)
    records = _aligned(max(N, 1) * L.nlosb200_record_floats(), torch.float32, dev, 128)
    st = _lib.stream()
    if N and semantics == "dense":
        if camera_pos is not None and int(active_sh_degree) > 0:
            raise RuntimeError("evaluate_points(semantics='dense') implements the degree-0 albedo only")
        _lib.call("nlosb200_preprocess_dense", gaussian_means.data_ptr(), gaussian_scales.data_ptr(), gaussian_rotations.data_ptr(),
                  gaussian_opacities.data_ptr(), feats.data_ptr(), N, feats.shape[1], float(scaling_modifier), records.data_ptr(), st)
    elif N:
        _lib.call("nlosb200_preprocess", gaussian_means.data_ptr(), gaussian_scales.data_ptr(), gaussian_rotations.data_ptr(),
                  gaussian_opacities.data_ptr(), feats.data_ptr(), N, feats.shape[1], float(scaling_modifier), 3.0, 0.0, 0.0, records.data_ptr(),
                  None, None, st)
        if camera_pos is not None:
            _check(camera_pos, "camera_pos")
            _lib.call("nlosb200_compat_albedo", gaussian_means.data_ptr(), feats.data_ptr(), camera_pos.data_ptr(), N, feats.shape[1],
                      effective_sh_degree(active_sh_degree, feats.shape[1]), records.data_ptr(), st)
    Na = points.shape[0]
    density = torch.zeros(Na, dtype=torch.float32, device=dev)
    rho_density = torch.zeros(Na, dtype=torch.float32, device=dev)
    _lib.call("nlosb200_evaluate_points", points.data_ptr(), Na, records.data_ptr(), N, density.data_ptr(), rho_density.data_ptr(), st)
    return rho_density, density
