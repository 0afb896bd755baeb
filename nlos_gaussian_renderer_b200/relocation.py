"""MCMC relocation / densification of the Gaussian set -- SURVEY.md 8 row f3 (gaussian_model/gaussian_model.py:471-549, main.py:243-247).

The reference's methods cannot run as written: `_update_params` calls `compute_relocation_cuda`, which it neither defines nor imports
(`:472`), misspells `unsqueeze` (`:477`), and `add_new_gs` omits an argument of `_sample_alives` (`:532`); `configs/default.py:78`
disables the branch.  This module provides the missing helper under the name the reference calls (`compute_relocation_cuda`, one CUDA
launch: csrc/optim.cu relocation_kernel) and the two operations with those three slips repaired, on a plain dict of the raw parameter
tensors {"mu", "f_dc", "f_rest", "opacity", "scaling", "rotation"} (the names of the reference's optimizer groups, `:233-240`).
The helper is the one of the public 3DGS-MCMC code (Kheradmand et al., NeurIPS 2024, eq. 9), which is not vendored in the reference:
PARITY UNPINNED (oracle/relocation.py restates the publication)."""
import math

import torch

from . import _lib

N_MAX = 51
_BINOMS = {}


def _binoms(device):
    key = str(device)
    if key not in _BINOMS:
        b = torch.zeros(N_MAX, N_MAX, dtype=torch.float32)
        for n in range(N_MAX):
            for k in range(n + 1):
                b[n, k] = math.comb(n, k)
        _BINOMS[key] = b.to(device)
    return _BINOMS[key]


def compute_relocation_cuda(opacity_old, scale_old, N):
    """(opacity_old [P] activated, scale_old [P,3] activated, N [P] copies per Gaussian) -> (new_opacity [P], new_scale [P,3]):
    N copies with new_opacity compose to opacity_old, the scale shrinks so that the rendered mass is kept."""
    if not opacity_old.is_cuda:
        raise RuntimeError("compute_relocation_cuda: CUDA tensors only (there is no CPU path)")
    opacity_old = opacity_old.detach().to(torch.float32).contiguous().reshape(-1)
    scale_old = scale_old.detach().to(torch.float32).contiguous().reshape(-1, 3)
    P = opacity_old.shape[0]
    if scale_old.shape[0] != P or N.numel() != P:
        raise RuntimeError(f"compute_relocation_cuda: {P} opacities, {scale_old.shape[0]} scales, {N.numel()} counts")
    counts = N.detach().reshape(-1).clamp(1, N_MAX - 1).to(torch.int32).contiguous()
    new_opacity, new_scale = torch.empty_like(opacity_old), torch.empty_like(scale_old)
    _lib.call("nlosb200_compute_relocation", opacity_old.data_ptr(), scale_old.data_ptr(), counts.data_ptr(), _binoms(opacity_old.device).data_ptr(),
              N_MAX, P, new_opacity.data_ptr(), new_scale.data_ptr(), _lib.stream())
    return new_opacity, new_scale


def _inverse_sigmoid(x):
    return torch.log(x / (1 - x))


def _update_params(params, idxs, ratio):
    """gaussian_model.py:471-483: the six tensors of the Gaussians `idxs`, opacity / scaling replaced by their relocated values."""
    new_opacity, new_scaling = compute_relocation_cuda(torch.sigmoid(params["opacity"][idxs, 0]), torch.exp(params["scaling"][idxs]), ratio[idxs, 0] + 1)
    new_opacity = torch.clamp(new_opacity.unsqueeze(-1), max=1.0 - torch.finfo(torch.float32).eps, min=0.005)
    return (params["mu"][idxs], params["f_dc"][idxs], params["f_rest"][idxs], _inverse_sigmoid(new_opacity), torch.log(new_scaling.reshape(-1, 3)),
            params["rotation"][idxs])


def _sample_alives(probs, num, alive_indices=None, generator=None):
    """gaussian_model.py:485-491: `num` draws (with replacement) proportional to probs -> (indices, how often every index was drawn)."""
    probs = probs / (probs.sum() + torch.finfo(torch.float32).eps)
    sampled = torch.multinomial(probs, num, replacement=True, generator=generator)
    if alive_indices is not None:
        sampled = alive_indices[sampled]
    return sampled, torch.bincount(sampled).unsqueeze(-1)


@torch.no_grad()
def relocate_gs(params, dead_mask, optimizer=None, generator=None):
    """gaussian_model.py:495-524, in place on `params`: every dead Gaussian becomes a copy of an alive one drawn proportionally to opacity;
    a Gaussian drawn n times is then present n + 1 times and all copies get the relocated opacity / scale.  `optimizer` (optim.FusedAdam whose
    groups carry the same names) has its moments zeroed at the re-initialised sources (replace_tensors_to_optimizer(inds), `:432-458`).
    -> the indices of the sources, one per dead Gaussian (empty when nothing moved)."""
    dead_mask = dead_mask.reshape(-1)
    dead = dead_mask.nonzero(as_tuple=True)[0]
    alive = (~dead_mask).nonzero(as_tuple=True)[0]
    if dead.numel() == 0 or alive.numel() == 0:
        return dead.new_empty(0)
    probs = torch.sigmoid(params["opacity"][alive, 0])
    src, ratio = _sample_alives(probs, dead.numel(), alive, generator)
    new = _update_params(params, src, ratio)
    for name, value in zip(("mu", "f_dc", "f_rest", "opacity", "scaling", "rotation"), new):
        params[name][dead] = value
    params["opacity"][src] = params["opacity"][dead]
    params["scaling"][src] = params["scaling"][dead]
    if optimizer is not None:
        optimizer.reset_state(src)
    return src


@torch.no_grad()
def add_new_gs(params, cap_max, optimizer=None, generator=None):
    """gaussian_model.py:526-549: grow the set by 5 % (up to cap_max) with copies of Gaussians drawn proportionally to opacity.
    -> (new params dict with the copies appended, number added).  The caller rebuilds its optimizer groups over the new tensors (the
    reference's densification_postfix does the same); `optimizer.reset_state(idx)` is applied to the sources first."""
    n = params["opacity"].shape[0]
    num = max(0, min(int(cap_max), int(1.05 * n)) - n)
    if num <= 0:
        return params, 0
    src, ratio = _sample_alives(torch.sigmoid(params["opacity"][:, 0]), num, None, generator)
    new = dict(zip(("mu", "f_dc", "f_rest", "opacity", "scaling", "rotation"), _update_params(params, src, ratio)))
    params["opacity"][src] = new["opacity"]
    params["scaling"][src] = new["scaling"]
    if optimizer is not None:
        optimizer.reset_state(src)
    return {k: torch.cat([params[k], new[k]], 0) for k in params}, num
