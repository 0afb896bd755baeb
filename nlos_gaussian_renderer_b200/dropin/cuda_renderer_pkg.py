"""Drop-in for the reference package `cuda_renderer` (submodules/cuda_renderer/__init__.py:24-186 and
section_renderer.py:21-305): NLOSGaussianRenderer (numerical, no autograd) and SectionGaussianRendererCUDA (analytic)."""
from typing import Optional, Tuple

import torch

try:
    from nlos_gaussian_renderer import _C
    CUDA_AVAILABLE = True
except ImportError:
    print("Warning: CUDA extension not compiled. Please run 'python setup.py install'")
    CUDA_AVAILABLE = False

from .lattice import angular_range_from_points, attenuate_and_integrate, ray_lattice


def _model_tensors(gm):
    """(means, log-scales, raw quaternions, opacity logits, dc features [N,K]) as contiguous tensors."""
    return [t.contiguous() for t in (gm.get_mu, gm._scaling, gm._rotation, gm._opacity, gm.get_features_dc.squeeze(1))]


class _FilterMixin:
    def filter_gaussians(self, ray_origins, ray_directions, gaussian_means, gaussian_bboxes):
        """bboxes [N,2,3] (min,max) -> int32 [N_rays, 257], column 0 = count (cuda_renderer/__init__.py:154-180)."""
        return _C.filter_gaussians_per_ray(ray_origins.contiguous(), ray_directions.contiguous(), gaussian_means.contiguous(),
                                           gaussian_bboxes.contiguous().view(-1, 6), self.sigma_threshold)


class NLOSGaussianRenderer(_FilterMixin):
    def __init__(self, sigma_threshold=3.0):
        if not CUDA_AVAILABLE:
            raise RuntimeError("CUDA extension is not available. Cannot use NLOSGaussianRenderer.")
        self.sigma_threshold = sigma_threshold

    def render(self, gaussian_model, camera_pos, theta_range, phi_range, r_range, num_theta, num_phi, num_r, c, deltaT,
               scaling_modifier=1.0, use_occlusion=True, rendering_type="netf"):
        """-> result [num_r,num_theta,num_phi], pred_histogram [num_r]   (cuda_renderer/__init__.py:43-152)"""
        o, d, t, tg = ray_lattice(camera_pos, theta_range, phi_range, r_range, num_theta, num_phi, num_r)
        rho, _, _ = _C.render_rays(o, d, t, *[x.detach() for x in _model_tensors(gaussian_model)], camera_pos.contiguous(),
                                   gaussian_model.active_sh_degree, c, deltaT, scaling_modifier, use_occlusion, rendering_type)
        return attenuate_and_integrate(rho, t, tg, theta_range, phi_range)


class SectionGaussianRendererCUDA(_FilterMixin):
    def __init__(self, sigma_threshold=3.0):
        if not CUDA_AVAILABLE:
            raise RuntimeError("CUDA extension is not available. Cannot use SectionGaussianRendererCUDA.\n"
                               "Please run: cd cuda_renderer && python setup.py install")
        self.sigma_threshold = sigma_threshold

    def render_transient(self, gaussian_model, camera_pos: torch.Tensor, theta_range: Tuple[float, float],
                         phi_range: Tuple[float, float], r_range: Tuple[float, float], num_theta: int, num_phi: int, num_r: int,
                         c: float, deltaT: float, scaling_modifier: float = 1.0, use_occlusion: bool = True,
                         rendering_type: str = "netf"):
        """Reference-faithful section renderer (section_renderer.py:55-186): the per-ray scalar lands in bin num_r//2 of `result`
        and the angular sum is broadcast to every bin of pred_histogram; no sin(theta)/r^2, no autograd."""
        o, d, _, _ = ray_lattice(camera_pos, theta_range, phi_range, r_range, num_theta, num_phi, num_r)
        means, scales, rots, opac, feats = [x.detach() for x in _model_tensors(gaussian_model)]
        bboxes = gaussian_model.get_bboxes(scaling_modifier=scaling_modifier, sigma_scale=self.sigma_threshold)
        flt = _C.filter_gaussians_per_ray(o, d, means, bboxes.detach().contiguous().view(-1, 6), self.sigma_threshold)
        per_ray = _C.render_rays_analytic(o, d, r_range[0], r_range[1], flt, means, scales, rots, opac, feats, camera_pos.contiguous(),
                                          gaussian_model.active_sh_degree, c, deltaT, scaling_modifier, self.sigma_threshold,
                                          rendering_type)
        h2 = per_ray.reshape(num_theta, num_phi)
        result = torch.zeros(num_r, num_theta, num_phi, device=camera_pos.device)
        result[num_r // 2] = h2
        dth = (theta_range[1] - theta_range[0]) / num_theta
        dph = (phi_range[1] - phi_range[0]) / num_phi
        return result, (h2.sum() * dth * dph).expand(num_r)

    def render_from_spherical_samples(self, gaussian_model, input_points: torch.Tensor, camera_pos: torch.Tensor, I1: int, I2: int,
                                      num_r: int, num_angular: int, dtheta: float, dphi: float, c: float, deltaT: float,
                                      scaling_modifier: float = 1.0, use_occlusion: bool = True, rendering_type: str = "netf"):
        th, ph = angular_range_from_points(input_points)
        res, hist = self.render_transient(gaussian_model=gaussian_model, camera_pos=camera_pos, theta_range=th, phi_range=ph,
                                          r_range=(I1 * c * deltaT, I2 * c * deltaT), num_theta=num_angular, num_phi=num_angular,
                                          num_r=num_r, c=c, deltaT=deltaT, scaling_modifier=scaling_modifier,
                                          use_occlusion=use_occlusion, rendering_type=rendering_type)
        return res.reshape(num_r, num_angular * num_angular), hist


def create_renderer(sigma_threshold=3.0):
    return NLOSGaussianRenderer(sigma_threshold=sigma_threshold)


def create_section_renderer(sigma_threshold=3.0) -> Optional[SectionGaussianRendererCUDA]:
    if not CUDA_AVAILABLE:
        print("Warning: CUDA renderer not available")
        return None
    return SectionGaussianRendererCUDA(sigma_threshold=sigma_threshold)
