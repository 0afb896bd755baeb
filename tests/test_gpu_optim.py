"""Fused multi-tensor Adam (csrc/optim.cu) against torch.optim.Adam configured as the reference does
(gaussian_model/gaussian_model.py:230-240: six groups, per-group lr, eps=1e-15)."""
import pytest
import torch

from nlos_gaussian_renderer_b200.optim import FusedAdam

pytestmark = pytest.mark.gpu


def test_fused_adam_matches_torch():
    g = torch.Generator().manual_seed(0)
    shapes = {"mu": (1000, 3), "f_dc": (1000, 1, 1), "f_rest": (1000, 0, 1), "opacity": (1000, 1), "scaling": (1000, 3), "rotation": (1000, 4)}
    lrs = {"mu": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 5e-2, "scaling": 5e-3, "rotation": 1e-3}
    ref_p = {k: torch.randn(*s, generator=g).cuda().requires_grad_() for k, s in shapes.items()}
    our_p = {k: v.detach().clone() for k, v in ref_p.items()}
    ref = torch.optim.Adam([{"params": [ref_p[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    ours = FusedAdam([{"params": [our_p[k]], "lr": lrs[k], "name": k} for k in shapes], eps=1e-15)
    for it in range(1, 26):
        grads = {k: (torch.randn(*s, generator=g) * (10.0 ** float(torch.randint(-6, 2, (1,), generator=g)))).cuda() for k, s in shapes.items()}
        if it == 10:                                        # the reference changes the position lr every iteration (update_learning_rate)
            for opt in (ref, ours):
                opt.param_groups[0]["lr"] = 3e-5
        for k in shapes:
            ref_p[k].grad = grads[k].clone()
        ref.step()
        ours.step([grads[k] for k in shapes])
    for k in shapes:
        if ref_p[k].numel():
            err = float((ref_p[k].detach() - our_p[k]).abs().max())
            assert err <= 2e-6 * float(ref_p[k].detach().abs().max()), (k, err)
