"""The fitting step end to end on the device (fit.TransientFit: render -> fused loss -> backward -> fused Adam), the loop of
main.py:148-162 on the batched renderer: the loss against a transient rendered from known Gaussians must go down."""
import pytest
import torch

from nlos_gaussian_renderer_b200 import fit
from nlos_gaussian_renderer_b200.transient import TransientRenderer
from tests import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["numerical", "analytic"])
def test_fitting_loop_reduces_the_loss(mode):
    N, Sy, Sx, T, A = 300, 8, 8, 64, 8
    geom, ta, truth = util.small_scene(N, Sy, Sx, T, A, scale_mul=2.0, device="cuda")
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, one_pass=True)
    target = r.forward(*truth).clone()
    g = torch.Generator(device="cuda").manual_seed(5)
    start = [p.clone() for p in truth]
    start[3] = start[3] + 0.8 * torch.randn(start[3].shape, device="cuda", generator=g)          # opacity logits
    start[4] = start[4] + 0.5 * torch.randn(start[4].shape, device="cuda", generator=g)          # dc features
    start[0] = start[0] + 0.004 * torch.randn(start[0].shape, device="cuda", generator=g)        # means, a few mm
    f = fit.TransientFit(r, start, target, target.numel(), opacity_lr=5e-2, feature_lr=2e-2, position_lr_max_steps=100)
    losses = [float(f.step(it)) for it in range(60)]
    assert all(l == l for l in losses)
    assert losses[-1] < 0.2 * losses[0], (losses[0], losses[-1])
    assert f.optimizer.param_groups[0]["lr"] == pytest.approx(fit.expon_lr(59, 1.6e-4, 1.6e-6, lr_delay_mult=0.01, max_steps=100))
