"""Drop-in for the reference's gaussian_model/cuda_autograd.py: same names and signatures, but the backward is REAL
(the reference returns zeros, gaussian_model/cuda_autograd.py:139-172)."""
from typing import Optional, Tuple

import torch
import torch.nn as nn

try:
    from nlos_gaussian_renderer import _C
    CUDA_AVAILABLE = True
except ImportError:                                   # same import-failure convention as cuda_autograd.py:10-15
    CUDA_AVAILABLE = False
    print("Warning: CUDA renderer not available")

from .lattice import attenuate_and_integrate, ray_lattice


class CUDARenderFunction(torch.autograd.Function):
    """apply(ray_origins, ray_directions, t_samples, means, scales(log), rotations(w,x,y,z raw), opacities(logit),
    features[N,K], camera_pos, active_sh_degree, c, deltaT, scaling_modifier, use_occlusion, rendering_type)
    -> rho_density, density, transmittance, each [N_rays, N_samples]   (cuda_autograd.py:27-108)."""

    @staticmethod
    def forward(ctx, ray_origins, ray_directions, t_samples, gaussian_means, gaussian_scales, gaussian_rotations,
                gaussian_opacities, gaussian_features, camera_pos, active_sh_degree, c, deltaT, scaling_modifier,
                use_occlusion, rendering_type):
        if not CUDA_AVAILABLE:
            raise RuntimeError("CUDA renderer not available")
        args = [t.detach().contiguous() for t in (ray_origins, ray_directions, t_samples, gaussian_means, gaussian_scales,
                                                  gaussian_rotations, gaussian_opacities, gaussian_features, camera_pos)]
        (rho, den, tr), prepared = _C.render_rays(*args, active_sh_degree, c, deltaT, scaling_modifier, use_occlusion, rendering_type,
                                                  _return_prepared=True)
        ctx.save_for_backward(*args, rho, den, tr)
        ctx.prepared = prepared
        ctx.consts = (int(active_sh_degree), float(c), float(deltaT), float(scaling_modifier), bool(use_occlusion))
        return rho, den, tr

    @staticmethod
    def backward(ctx, grad_rho_density, grad_density, grad_transmittance):
        """15-tuple with gradients at positions 3..7 (means, scales, rotations, opacities, features), None elsewhere
        (cuda_autograd.py:174-191)."""
        ro, rd, ts, means, scales, rots, opac, feats, cam, rho, den, tr = ctx.saved_tensors
        deg, c, deltaT, mod, occ = ctx.consts
        needs = ctx.needs_input_grad[3:8]
        out = [None] * 15
        if not any(needs):
            return tuple(out)
        g = _C.render_rays_backward(ctx.prepared, ro, rd, ts, scales, rots, opac, feats, c, deltaT, mod, occ, rho, tr,
                                    grad_rho_density, grad_density, grad_transmittance)
        for i, (need, gi) in enumerate(zip(needs, g)):
            if need:
                out[3 + i] = gi
        return tuple(out)


class CUDARenderModule(nn.Module):
    """forward(gaussian_model, camera_pos, theta_range, phi_range, r_range, num_theta, num_phi, num_r, c, deltaT,
    scaling_modifier=1.0, use_occlusion=True, rendering_type='netf') -> (result [num_r,num_theta,num_phi], pred_histogram [num_r])
    (cuda_autograd.py:194-316)."""

    def __init__(self, sigma_threshold: float = 3.0):
        super().__init__()
        if not CUDA_AVAILABLE:
            raise RuntimeError("CUDA renderer not available")
        self.sigma_threshold = sigma_threshold       # kept for API parity; render_rays hard-codes 3.0 (volume_renderer.cu:232)

    def forward(self, gaussian_model, camera_pos: torch.Tensor, theta_range: Tuple[float, float], phi_range: Tuple[float, float],
                r_range: Tuple[float, float], num_theta: int, num_phi: int, num_r: int, c: float, deltaT: float,
                scaling_modifier: float = 1.0, use_occlusion: bool = True, rendering_type: str = "netf"):
        origins, dirs, t_samples, theta_grid = ray_lattice(camera_pos, theta_range, phi_range, r_range, num_theta, num_phi, num_r)
        # cuda_autograd.py:280 passes the dc coefficient only; with active_sh_degree > 0 the kernel needs (deg+1)^2 coefficients per
        # Gaussian: take them from get_features ([N, K, 1] dc + rest) when the model has them, else _C evaluates degree 0 (documented)
        feats = gaussian_model.get_features_dc.squeeze(1)
        deg = int(gaussian_model.active_sh_degree)
        if deg > 0 and hasattr(gaussian_model, "get_features"):
            full = gaussian_model.get_features
            full = full.reshape(full.shape[0], -1)
            if full.shape[1] >= (deg + 1) ** 2:
                feats = full[:, :(deg + 1) ** 2].contiguous()
        rho, _, _ = CUDARenderFunction.apply(origins, dirs, t_samples, gaussian_model.get_mu, gaussian_model._scaling,
                                             gaussian_model._rotation, gaussian_model._opacity, feats, camera_pos,
                                             gaussian_model.active_sh_degree, c, deltaT, scaling_modifier, use_occlusion, rendering_type)
        return attenuate_and_integrate(rho, t_samples, theta_grid, theta_range, phi_range)


def create_cuda_render_module(sigma_threshold: float = 3.0) -> Optional[CUDARenderModule]:
    return CUDARenderModule(sigma_threshold=sigma_threshold) if CUDA_AVAILABLE else None
