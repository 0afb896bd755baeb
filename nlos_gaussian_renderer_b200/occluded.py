"""Batched [scan_y, scan_x, time_bin] transient WITH occlusion (transmittance along each ray).

With occlusion the reference's semantics are per RAY, not per scan point: per-ray list of at most 256 Gaussians in index order
(src/ray_aabb.cu:10-61), alpha_g = 1 - exp(-pdf_g sigma(o_g) c dT), running transmittance over the samples of the ray with the early stop at
1e-4 (src/volume_renderer.cu:80-137), then /(r^2+1e-8) * sin(theta) and the angular sum (gaussian_model/cuda_autograd.py:302-314) times
volume_position[1]^2 (nlos_helpers.py:275-276).  The reference loops CUDARenderModule.forward over the scan points (main.py:198-203);
here the rays of a chunk of scan points go through ONE call of the per-ray operator (`_C.render_rays`: compat_filter_kernel +
compat_render_kernel<OCC>, gradients by compat_render_backward_kernel<OCC>), which is legal because every ray carries its own origin and
the degree-0 albedo does not depend on the scan point.  (The no-occlusion transient has its own, much faster, scan-tile kernel:
transient.TransientRenderer.)"""
import torch

from .dropin.cuda_autograd import CUDARenderFunction
from .geometry import ScanGeometry


def render_transient_batch_occluded(geom: ScanGeometry, t_samples, means, log_scales, rotations, opacities, features, c, deltaT,
                                    scaling_modifier=1.0, use_occlusion=True, chunk_scan_points=256):
    """-> H [Sy, Sx, T] fp32, differentiable w.r.t. the five parameter tensors (degree-0 SH)."""
    S, A, T = geom.S, geom.A, int(t_samples.shape[0])
    dev = means.device
    feats = features.reshape(features.shape[0], -1)[:, :1].contiguous()
    inv_r2 = 1.0 / (t_samples.to(torch.float32) ** 2 + 1e-8)                                   # [T]
    cam0 = geom.scan_o[0].contiguous()                                                          # unused by the degree-0 albedo
    out = []
    for s0 in range(0, S, chunk_scan_points):
        sl = slice(s0, min(S, s0 + chunk_scan_points))
        st, ct = geom.sin_th[sl], geom.cos_th[sl]                                               # [n,A]
        sp, cp = geom.sin_ph[sl], geom.cos_ph[sl]
        n = st.shape[0]
        d = torch.stack([st[:, :, None] * cp[:, None, :], st[:, :, None] * sp[:, None, :], ct[:, :, None].expand(n, A, A)], -1)
        d = d.reshape(-1, 3).contiguous()                                                       # ray (p, i_theta, j_phi)
        o = geom.scan_o[sl][:, None, :].expand(n, A * A, 3).reshape(-1, 3).contiguous()
        rho, _, _ = CUDARenderFunction.apply(o, d, t_samples, means, log_scales, rotations, opacities, feats, cam0, 0, c, deltaT,
                                             scaling_modifier, use_occlusion, "netf")
        w = (st[:, :, None].expand(n, A, A).reshape(n, A * A) * geom.hscale[sl, None])          # sin(theta) dtheta dphi vp_y^2
        H = torch.einsum("pr,prt->pt", w, rho.reshape(n, A * A, T)) * inv_r2[None, :]
        out.append(H)
    return torch.cat(out, 0).reshape(geom.Sy, geom.Sx, T)
