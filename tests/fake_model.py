"""Minimal stand-in for the attributes of the reference's GaussianModel that the renderer facade reads
(gaussian_model/gaussian_model.py:107-178: get_mu, _scaling, _rotation, _opacity, get_features_dc, active_sh_degree,
get_bboxes).  TEST INFRASTRUCTURE ONLY -- the real GaussianModel (host-side training logic) is out of scope."""
import torch


class FakeGaussianModel:
    def __init__(self, means, log_scales, rotations, opacities, features_dc, active_sh_degree=0):
        mk = lambda t: torch.nn.Parameter(t.detach().clone().float())
        self._mu, self._scaling, self._rotation = mk(means), mk(log_scales), mk(rotations)
        self._opacity = mk(opacities.reshape(-1, 1))
        self._features_dc = mk(features_dc.reshape(means.shape[0], 1, -1))          # [N,1,K]
        self.active_sh_degree = active_sh_degree

    @classmethod
    def random(cls, N, device="cpu", spread=0.3, centre=(0.0, 1.0, 0.0), scale=0.1):
        means = torch.tensor(centre, device=device) + (torch.rand(N, 3, device=device) * 2 - 1) * spread
        ls = torch.log(scale * (0.5 + torch.rand(N, 3, device=device)))
        return cls(means, ls, torch.randn(N, 4, device=device), torch.randn(N, 1, device=device),
                   torch.rand(N, 1, device=device) - 0.5)

    get_mu = property(lambda self: self._mu)
    get_features_dc = property(lambda self: self._features_dc)

    def get_bboxes(self, scaling_modifier=1.0, sigma_scale=3.0):
        """[N,2,3] = mean -+ sigma_scale*sqrt(diag(R S^2 R^T))   (semantics of gaussian_model.py:140-178)."""
        q = torch.nn.functional.normalize(self._rotation, dim=1)
        w, x, y, z = q.unbind(1)
        R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                         2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                         2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)
        s = torch.exp(self._scaling) * scaling_modifier
        ext = sigma_scale * torch.sqrt(torch.clamp_min(((R * s[:, None, :]) ** 2).sum(-1), 1e-8))
        return torch.stack([self._mu - ext, self._mu + ext], 1)
