"""Differentiable PyTorch (fp64) restatement of the CUDA-kernel semantics -- TEST INFRASTRUCTURE ONLY.

The reference backward returns zeros (gaussian_model/cuda_autograd.py:139-172), so gradient truth is
torch.autograd through this restatement of
  volume_renderer.cu:139-182   (numerical, no occlusion)   and :80-137 (occlusion)
  cuda_utils.cuh:54-85,124-151 (quat_to_rotmat, eval_gaussian_pdf)
  cuda_autograd.py:302-314     (attenuation + angular sum), nlos_helpers.py:275-276 (x vp_y^2)
and of the analytic-exact bin integral.  Ray/Gaussian membership (the non-
differentiable 3-sigma AABB slab test) comes from the C oracle as a boolean mask.
"""
import math

import torch

C0 = 0.28209479177387814


def quat_to_rotmat(q):
    n = q.norm(dim=1, keepdim=True)
    w, x, y, z = (q / n).unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1)
    return R.reshape(-1, 3, 3)


def sh_basis(deg, v):
    """[...,3] unit vectors -> [..., (deg+1)^2]: real SH polynomials, the all-positive sign convention of spherical_harmonics.cuh:20-53."""
    x, y, z = v.unbind(-1)
    Y = [torch.full_like(x, C0)]
    if deg >= 1:
        Y += [0.4886025119029199 * y, 0.4886025119029199 * z, 0.4886025119029199 * x]
    if deg >= 2:
        Y += [1.0925484305920792 * x * y, 1.0925484305920792 * y * z, 0.31539156525252005 * (3 * z * z - 1), 1.0925484305920792 * x * z,
              0.5462742152960396 * (x * x - y * y)]
    if deg >= 3:
        Y += [0.5900435899266435 * y * (3 * x * x - y * y), 2.890611442640554 * x * y * z, 0.4570457994644658 * y * (5 * z * z - 1),
              0.3731763325901154 * z * (5 * z * z - 3), 0.4570457994644658 * x * (5 * z * z - 1), 1.445305721320277 * z * (x * x - y * y),
              0.5900435899266435 * x * (x * x - 3 * y * y)]
    return torch.stack(Y, -1)


def view_albedo(means, feats, cam, deg):
    """max(SH(dir(mu - cam)) + 0.5, 0), dir = c / (|c| + 1e-8)   (volume_renderer.cu:107-111, cuda_utils.cuh:48-51)."""
    c = means - cam[None, :]
    v = c / (c.norm(dim=1, keepdim=True) + 1e-8)
    K = (deg + 1) ** 2
    return torch.clamp_min((sh_basis(deg, v) * feats.reshape(feats.shape[0], -1)[:, :K]).sum(-1) + 0.5, 0.0)


def activate(means, log_scales, rots, opac, feats, mod=1.0):
    R = quat_to_rotmat(rots)
    s = torch.exp(log_scales) * mod
    M = R.transpose(1, 2) / (s + 1e-8)[:, :, None]          # M[j][i] = R[i][j] / (s_j + eps)
    op = torch.sigmoid(opac.reshape(-1))
    alb = torch.clamp_min(C0 * feats.reshape(feats.shape[0], -1)[:, 0] + 0.5, 0.0)
    return M, op, alb


def ray_values(o, d, t, mask, means, M, op, mode, dt_lat):
    """o [3], d [R,3], t [T], mask [R,N] -> per (ray, gaussian, sample) value v with
    numerical: pdf*op ; analytic: op * bin integral.   -> [R,N,T]"""
    a = torch.einsum("nij,nj->ni", M, (o[None, :] - means))             # [N,3]
    b = torch.einsum("nij,rj->rni", M, d)                                # [R,N,3]
    if mode == 0:
        q = a[None, :, None, :] + t[None, None, :, None] * b[:, :, None, :]      # [R,N,T,3]
        v = torch.exp(-0.5 * (q * q).sum(-1)) * op[None, :, None]
    else:
        C = (b * b).sum(-1)
        ab = (a[None] * b).sum(-1)
        rs = -ab / C
        perp = a[None] + rs[..., None] * b
        m2 = (perp * perp).sum(-1)
        sq = torch.sqrt(0.5 * C)
        pref = torch.sqrt(math.pi / (2 * C)) * torch.exp(-0.5 * m2)
        k = torch.arange(t.shape[0], dtype=t.dtype, device=t.device)
        e0 = t[0] + (k - 0.5) * dt_lat
        e1 = t[0] + (k + 0.5) * dt_lat
        I = pref[..., None] * (torch.erf(sq[..., None] * (e1 - rs[..., None])) - torch.erf(sq[..., None] * (e0 - rs[..., None])))
        v = I * op[None, :, None]
    return v * mask[:, :, None].to(v.dtype)


def transient_batch(scan_o, sin_th, cos_th, sin_ph, cos_ph, t, hscale, masks, means, log_scales, rots, opac, feats,
                    c, deltaT, mod=1.0, mode=0, sh_degree=0):
    """H [S,T]; masks [S, A*A, N] bool (ray r = i*A + j)."""
    M, op, alb = activate(means, log_scales, rots, opac, feats, mod)
    S, A = sin_th.shape
    T = t.shape[0]
    dt_lat = (t[-1] - t[0]) / (T - 1)
    out = []
    for p in range(S):
        d = torch.stack([(sin_th[p][:, None] * cos_ph[p][None, :]).reshape(-1),
                         (sin_th[p][:, None] * sin_ph[p][None, :]).reshape(-1),
                         cos_th[p][:, None].expand(A, A).reshape(-1)], 1)
        v = ray_values(scan_o[p], d, t, masks[p], means, M, op, mode, dt_lat)       # [R,N,T]
        a_p = alb if sh_degree == 0 else view_albedo(means, feats, scan_o[p], sh_degree)     # per (Gaussian, scan point)
        rho = (v * a_p[None, :, None]).sum(1)                                         # [R,T]
        if mode == 0:
            rho = rho * c * deltaT
        w = sin_th[p][:, None].expand(A, A).reshape(-1)
        out.append(hscale[p] * (rho * w[:, None]).sum(0) / (t * t + 1e-8))
    return torch.stack(out)


def render_rays(ray_o, ray_d, t, mask, means, log_scales, rots, opac, feats, c, deltaT, mod=1.0, use_occlusion=False, sh_degree=0, cam=None):
    """(rho, density, transmittance) [R,T] of volume_renderer.cu for rays sharing one origin; sh_degree > 0: albedo of camera `cam`."""
    M, op, alb = activate(means, log_scales, rots, opac, feats, mod)
    if sh_degree > 0:
        alb = view_albedo(means, feats, cam, sh_degree)
    v = ray_values(ray_o[0], ray_d, t, mask, means, M, op, 0, None)               # pdf*op  [R,N,T]
    density = v.sum(1)
    if not use_occlusion:
        rho = (v * alb[None, :, None]).sum(1) * c * deltaT
        return rho, density, torch.ones_like(density)
    alpha = 1 - torch.exp(-v * c * deltaT)
    wa = (alpha * alb[None, :, None]).sum(1)
    step = torch.exp(-density * c * deltaT)
    Tr = torch.cumprod(torch.cat([torch.ones_like(step[:, :1]), step[:, :-1]], 1), 1)
    after = Tr * step                                                              # T after the update at sample s
    alive = torch.cat([torch.ones_like(after[:, :1], dtype=torch.bool), (after[:, :-1] >= 1e-4)], 1)
    alive = torch.cumprod(alive.to(torch.int64), 1).bool()
    z = torch.zeros_like(density)
    return torch.where(alive, Tr * wa, z), torch.where(alive, density, z), torch.where(alive, Tr, z)


def evaluate_points(x, means, log_scales, rots, opac, feats, mod=1.0, chunk=512):
    """Dense sums over ALL Gaussians at points x [Na,3] -> (rho_density, density), the quantities of
    GaussianModel.estimate_rho_w_no_occlusion(out_separately=True) (gaussian_model/gaussian_model.py:346-364) with the CUDA renderer's
    pdf (include/cuda_utils.cuh:124-151) and degree-0 albedo."""
    M, op, alb = activate(means, log_scales, rots, opac, feats, mod)
    dens = torch.zeros(x.shape[0], dtype=x.dtype)
    rho = torch.zeros(x.shape[0], dtype=x.dtype)
    for g0 in range(0, means.shape[0], chunk):
        sl = slice(g0, g0 + chunk)
        loc = torch.einsum("gij,gnj->gni", M[sl], x[None] - means[sl][:, None])
        pdf = torch.exp(-0.5 * (loc * loc).sum(-1))
        dens += (pdf * op[sl, None]).sum(0)
        rho += (pdf * (op[sl] * alb[sl])[:, None]).sum(0)
    return rho, dens
