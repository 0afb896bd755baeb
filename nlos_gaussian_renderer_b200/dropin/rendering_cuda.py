"""Drop-in for the reference's gaussian_model/rendering_cuda.py:188-353."""
from typing import Optional, Tuple

import torch

from .cuda_autograd import CUDA_AVAILABLE, CUDARenderModule
from .lattice import angular_range_from_points


class GaussianRendererCUDA:
    """render_transient(...) / render_from_spherical_samples(...) with the reference's keyword signature; autograd-enabled."""

    def __init__(self, sigma_threshold=3.0):
        self.use_cuda = CUDA_AVAILABLE
        self.renderer = CUDARenderModule(sigma_threshold=sigma_threshold) if self.use_cuda else None
        if not self.use_cuda:
            print("CUDA renderer not available, will use fallback implementation")

    def render_transient(self, gaussian_model, camera_pos: torch.Tensor, theta_range: Tuple[float, float],
                         phi_range: Tuple[float, float], r_range: Tuple[float, float], num_theta: int, num_phi: int, num_r: int,
                         c: float, deltaT: float, scaling_modifier: float = 1.0, use_occlusion: bool = True,
                         rendering_type: str = "netf"):
        """-> result [num_r, num_theta, num_phi], pred_histogram [num_r]   (rendering_cuda.py:208-263)"""
        if not self.use_cuda or self.renderer is None:
            raise RuntimeError("CUDA renderer is not available")
        return self.renderer(gaussian_model=gaussian_model, camera_pos=camera_pos, theta_range=theta_range, phi_range=phi_range,
                             r_range=r_range, num_theta=num_theta, num_phi=num_phi, num_r=num_r, c=c, deltaT=deltaT,
                             scaling_modifier=scaling_modifier, use_occlusion=use_occlusion, rendering_type=rendering_type)

    def render_from_spherical_samples(self, gaussian_model, input_points: torch.Tensor, camera_pos: torch.Tensor, I1: int, I2: int,
                                      num_r: int, num_angular: int, dtheta: float, dphi: float, c: float, deltaT: float,
                                      scaling_modifier: float = 1.0, use_occlusion: bool = True, rendering_type: str = "netf"):
        """input_points [Na,5] (x,y,z,theta,phi) of spherical_sample_histogram -> result [num_r, num_angular^2], pred_histogram [num_r]
        (rendering_cuda.py:265-336; dtheta/dphi are accepted and recomputed from the ranges, as there)."""
        th, ph = angular_range_from_points(input_points)
        res, hist = self.render_transient(gaussian_model=gaussian_model, camera_pos=camera_pos, theta_range=th, phi_range=ph,
                                          r_range=(I1 * c * deltaT, I2 * c * deltaT), num_theta=num_angular, num_phi=num_angular,
                                          num_r=num_r, c=c, deltaT=deltaT, scaling_modifier=scaling_modifier,
                                          use_occlusion=use_occlusion, rendering_type=rendering_type)
        return res.reshape(num_r, num_angular * num_angular), hist


def create_cuda_renderer(sigma_threshold=3.0) -> Optional[GaussianRendererCUDA]:
    if not CUDA_AVAILABLE:
        print("Warning: CUDA renderer not available")
        return None
    return GaussianRendererCUDA(sigma_threshold=sigma_threshold)
