"""The fitting step of the reference's training loop on the batched renderer: what main.py:148-162 / 200-214 do per iteration
(update_learning_rate -> zero_grad -> compute_loss -> backward -> optimizer.step) with every stage on the device and no host sync:

    preprocess + binning -> transient forward -> fused MSE loss + dL/dH -> transient backward -> ONE all-reduce of the packed [N,16]
    gradients (multi-GPU) -> per-Gaussian chain rule -> fused Adam over the five parameter tensors (csrc/optim.cu)

Optimiser set-up as GaussianModel.training_setup (gaussian_model/gaussian_model.py:223-249): Adam(eps=1e-15), one group per tensor, the
position learning rate on the log-linear schedule of get_expon_lr_func (gaussian_model/gaussian_utils.py:223-257), defaults of
configs/default.py:62-69."""
import math

from .optim import FusedAdam
from .parallel import ShardedFitter


def expon_lr(step, lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1_000_000):
    """lr_init at step 0 -> lr_final at max_steps, log-linear in between, optionally eased in over lr_delay_steps
    (gaussian_utils.py:241-255)."""
    if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
        return 0.0
    ease = 1.0
    if lr_delay_steps > 0:
        ease = lr_delay_mult + (1.0 - lr_delay_mult) * math.sin(0.5 * math.pi * min(max(step / lr_delay_steps, 0.0), 1.0))
    t = min(max(step / max_steps, 0.0), 1.0)
    return ease * math.exp(math.log(lr_init) * (1.0 - t) + math.log(lr_final) * t)


class TransientFit:
    """params: (means [N,3], log_scales [N,3], rotations [N,4], opacities [N] or [N,1], features [N,K]) fp32 CUDA tensors, updated IN PLACE.
    target_slab: this rank's rows of the measured transient [rows, Sx, T]; global_numel = Sy*Sx*T of the whole grid."""

    GROUPS = ("mu", "scaling", "rotation", "opacity", "f_dc")

    def __init__(self, renderer, params, target_slab, global_numel, group=None, position_lr_init=1.6e-4, position_lr_final=1.6e-6,
                 position_lr_delay_mult=0.01, position_lr_max_steps=50_000, spatial_lr_scale=1.0, feature_lr=2.5e-3, opacity_lr=2.5e-2,
                 scaling_lr=5e-3, rotation_lr=1e-3):
        self.params = [p.detach() for p in params]
        self.fitter = ShardedFitter(renderer, target_slab, global_numel, group)
        self.mu_schedule = lambda it: expon_lr(it, position_lr_init * spatial_lr_scale, position_lr_final * spatial_lr_scale,
                                               lr_delay_mult=position_lr_delay_mult, max_steps=position_lr_max_steps)
        lrs = (position_lr_init * spatial_lr_scale, scaling_lr, rotation_lr, opacity_lr, feature_lr)
        self.optimizer = FusedAdam([{"params": [p], "lr": lr, "name": n} for p, lr, n in zip(self.params, lrs, self.GROUPS)], eps=1e-15)

    def update_learning_rate(self, iteration):
        """gaussian_model.py:244-249: only the position group is scheduled."""
        lr = self.mu_schedule(iteration)
        self.optimizer.param_groups[0]["lr"] = lr
        return lr

    def step(self, iteration):
        """One iteration -> this rank's loss contribution (a device tensor; summing it over ranks gives the global mean squared error)."""
        self.update_learning_rate(iteration)
        loss, grads = self.fitter.step(*self.params)
        self.optimizer.step(list(grads))
        return loss
