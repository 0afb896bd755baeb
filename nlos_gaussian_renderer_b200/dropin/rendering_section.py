"""Drop-in for the reference's gaussian_model/rendering_section.py:18-176 (thin forwarder to SectionGaussianRendererCUDA)."""
from typing import Optional

try:
    from cuda_renderer import SectionGaussianRendererCUDA, CUDA_AVAILABLE
except ImportError:
    CUDA_AVAILABLE = False
    print("Warning: cuda_renderer not available. Section rendering disabled.")


class GaussianSectionRenderer:
    def __init__(self, sigma_threshold=3.0):
        self.use_cuda = CUDA_AVAILABLE
        self.renderer = SectionGaussianRendererCUDA(sigma_threshold=sigma_threshold) if self.use_cuda else None
        if not self.use_cuda:
            print("CUDA renderer not available, section rendering disabled")

    def _need(self):
        if not self.use_cuda or self.renderer is None:
            raise RuntimeError("Section renderer is not available")
        return self.renderer

    def render_transient(self, gaussian_model, camera_pos, theta_range, phi_range, r_range, num_theta, num_phi, num_r, c, deltaT,
                         scaling_modifier=1.0, use_occlusion=True, rendering_type="netf"):
        return self._need().render_transient(gaussian_model=gaussian_model, camera_pos=camera_pos, theta_range=theta_range,
                                             phi_range=phi_range, r_range=r_range, num_theta=num_theta, num_phi=num_phi, num_r=num_r,
                                             c=c, deltaT=deltaT, scaling_modifier=scaling_modifier, use_occlusion=use_occlusion,
                                             rendering_type=rendering_type)

    def render_from_spherical_samples(self, gaussian_model, input_points, camera_pos, I1, I2, num_r, num_angular, dtheta, dphi, c, deltaT,
                                      scaling_modifier=1.0, use_occlusion=True, rendering_type="netf"):
        return self._need().render_from_spherical_samples(gaussian_model=gaussian_model, input_points=input_points, camera_pos=camera_pos,
                                                          I1=I1, I2=I2, num_r=num_r, num_angular=num_angular, dtheta=dtheta, dphi=dphi,
                                                          c=c, deltaT=deltaT, scaling_modifier=scaling_modifier,
                                                          use_occlusion=use_occlusion, rendering_type=rendering_type)


def create_section_renderer(sigma_threshold=3.0) -> Optional[GaussianSectionRenderer]:
    if not CUDA_AVAILABLE:
        print("Warning: Section renderer not available")
        return None
    return GaussianSectionRenderer(sigma_threshold=sigma_threshold)
