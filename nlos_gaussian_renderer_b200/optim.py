"""Fused Adam for the Gaussian parameter groups: the optimiser step of the reference's fitting loop
(gaussian_model/gaussian_model.py:223-249 training_setup -> torch.optim.Adam(l, lr=0.0, eps=1e-15); main.py:161,213) as ONE kernel
launch over all groups (csrc/optim.cu).  Same update as torch.optim.Adam (no weight decay, no amsgrad); CUDA tensors only."""
import ctypes as C

import torch

from . import _lib


class FusedAdam:
    """param_groups: list of dicts {'params': [tensor], 'lr': float, 'name': str} exactly as the reference builds them; every group holds
    ONE fp32 CUDA tensor.  `param_groups[i]['lr']` may be changed between steps (GaussianModel.update_learning_rate does)."""

    def __init__(self, param_groups, betas=(0.9, 0.999), eps=1e-15):
        self.param_groups = [dict(g) for g in param_groups]
        assert 1 <= len(self.param_groups) <= 8
        for g in self.param_groups:
            assert len(g["params"]) == 1 and g["params"][0].is_cuda and g["params"][0].dtype == torch.float32 and g["params"][0].is_contiguous()
        self.betas, self.eps, self.t = betas, float(eps), 0
        self.state = [{"exp_avg": torch.zeros_like(g["params"][0]), "exp_avg_sq": torch.zeros_like(g["params"][0])} for g in self.param_groups]

    def zero_grad(self, set_to_none=True):
        for g in self.param_groups:
            p = g["params"][0]
            if p.grad is not None:
                p.grad = None if set_to_none else p.grad.zero_()

    @torch.no_grad()
    def step(self, grads=None):
        """grads: optional list of gradient tensors (one per group); default = each parameter's .grad (groups without one are skipped)."""
        ps, gs, ms, vs, lrs = [], [], [], [], []
        for i, g in enumerate(self.param_groups):
            p = g["params"][0]
            gr = grads[i] if grads is not None else p.grad
            if gr is None or p.numel() == 0:
                continue
            gr = gr.detach().to(torch.float32).contiguous().reshape(p.shape)
            ps.append(p); gs.append(gr); ms.append(self.state[i]["exp_avg"]); vs.append(self.state[i]["exp_avg_sq"]); lrs.append(float(g["lr"]))
        self.t += 1
        n = len(ps)
        if n == 0:
            return
        arr = lambda ts: (C.c_void_p * n)(*[t.data_ptr() for t in ts])
        sizes = (C.c_longlong * n)(*[p.numel() for p in ps])
        lr = (C.c_float * n)(*lrs)
        _lib.call("nlosb200_adam_step", n, arr(ps), arr(gs), arr(ms), arr(vs), sizes, lr, float(self.betas[0]), float(self.betas[1]), self.eps,
                  self.t, _lib.stream())

    @torch.no_grad()
    def reset_state(self, inds=None):
        """Zero both moments of every group at the rows `inds` (all rows when None) -- what GaussianModel.replace_tensors_to_optimizer does
        to the sources of relocated / cloned Gaussians (gaussian_model/gaussian_model.py:432-458)."""
        for st in self.state:
            if inds is None:
                st["exp_avg"].zero_(); st["exp_avg_sq"].zero_()
            else:
                st["exp_avg"][inds] = 0
                st["exp_avg_sq"][inds] = 0
