"""CPU restatement of the reference's DENSE PyTorch renderer -- TEST / BASELINE INFRASTRUCTURE ONLY.

This is semantics (a) of SURVEY.md App. A: the path that actually trains in the reference
(use_cuda_renderer=False, configs/default.py:42) and the only one with working gradients:
  gaussian_model/gaussian_model.py:253-294  estimate_gaussian_pdf   (scale = exp(exp(_scaling)*mod) double
                                            activation :265, R (x-mu) :281-282, no epsilon)
  gaussian_model/gaussian_model.py:346-364  estimate_rho_w_no_occlusion  (sum_g pdf * sigmoid(o) * albedo)
  gaussian_model/sh_utils.py:57-117         degree-0 SH:  C0 * dc ; albedo = clamp_min(. + 0.5, 0) (:310)
  nlos_helpers.py:206-232                   /r^2 * sin(theta), angular sum * dtheta*dphi, * vp_y^2
It cannot be imported from /root/reference on the GPU box (the tree is absent there, and at HEAD it
does not import at all: gaussian_model.py:4), so bench.py times this port as the "PyTorch CPU path".
It is NOT a parity oracle for the CUDA kernels (different semantics, App. A.1)."""
import torch

C0 = 0.28209479177387814


def build_rotation(q):
    """gaussian_utils.py:189-210 (no zero-norm guard)."""
    q = q / q.norm(dim=1, keepdim=True)
    r, x, y, z = q.unbind(1)
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)


def dense_histogram(points, theta, r, means, log_scales, rots, opac, feats_dc, dtheta, dphi, vp_y, mod=1.0, chunk=256):
    """points [T,A,A,3] lattice of ONE scan point, theta [A], r [T] -> pred_histogram [T] (autograd-able)."""
    T, A = points.shape[0], points.shape[1]
    x = points.reshape(-1, 3)
    rho = torch.zeros(x.shape[0], dtype=x.dtype)
    for g0 in range(0, means.shape[0], chunk):
        sl = slice(g0, g0 + chunk)
        s = torch.exp(torch.exp(log_scales[sl]) * mod)              # the reference's double activation
        R = build_rotation(rots[sl])
        diff = x[None, :, :] - means[sl][:, None, :]                 # [G,Na,3]
        loc = torch.einsum("gij,gnj->gni", R, diff) / s[:, None, :]
        pdf = torch.exp(-0.5 * (loc * loc).sum(-1))
        alb = torch.clamp_min(C0 * feats_dc[sl].reshape(-1) + 0.5, 0.0)
        rho = rho + (pdf * (torch.sigmoid(opac[sl].reshape(-1)) * alb)[:, None]).sum(0)
    res = rho.reshape(T, A, A) / (r.reshape(-1, 1, 1) ** 2) * torch.sin(theta).reshape(1, -1, 1)
    return res.sum((1, 2)) * dtheta * dphi * vp_y ** 2
