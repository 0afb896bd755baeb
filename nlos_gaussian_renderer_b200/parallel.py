"""Scan-point sharding across GPUs: each rank renders a slab of the relay-wall grid with a full replica
of the Gaussians; the forward needs no communication (the loss is separable per scan point,
nlos_helpers.py:323-325); the backward sums the packed per-Gaussian gradient buffer [N,16] with ONE
all-reduce per fitting step (NCCL over NVLink on GPUs, gloo in the CPU tests).  The reference has no
multi-GPU path (submodules/cuda_renderer/README.md:126 lists it as future work)."""
import torch
import torch.distributed as dist


def slab_rows(Sy, tile_y, rank, world):
    """Contiguous scan-row slab [y0,y1) of `rank`, split at tile_y granularity, near-equal sizes."""
    n_tr = -(-Sy // tile_y)
    base, rem = divmod(n_tr, world)
    t0 = rank * base + min(rank, rem)
    t1 = t0 + base + (1 if rank < rem else 0)
    return min(t0 * tile_y, Sy), min(t1 * tile_y, Sy)


def all_reduce_packed(dacc, group=None):
    """In-place sum of the packed gradient buffer over ranks (no-op for world size 1)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(dacc, op=dist.ReduceOp.SUM, group=group)
    return dacc


def gather_transient(H_slab, Sy, tile_y, group=None):
    """All-gather the per-rank slabs [rows_r, Sx, T] into the full [Sy, Sx, T] transient (optional; fitting never needs it)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    rows = [slab_rows(Sy, tile_y, r, world) for r in range(world)]
    mx = max(b - a for a, b in rows)
    pad = torch.zeros(mx, *H_slab.shape[1:], dtype=H_slab.dtype, device=H_slab.device)
    pad[: H_slab.shape[0]] = H_slab
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    return torch.cat([o[: b - a] for o, (a, b) in zip(outs, rows)], 0)


class ShardedFitter:
    """One fitting step on this rank's slab: loss = mean over the GLOBAL grid of (H - target)^2.
    With a TransientRenderer the step stays in the tile-major layout end to end: forward kernel -> fused loss + dL/dH kernel
    (csrc/loss.cu) -> backward kernel -> ONE all-reduce of the packed gradients -> per-Gaussian chain."""

    def __init__(self, renderer, target_slab, global_numel, group=None):
        self.r, self.target, self.numel, self.group = renderer, target_slab, float(global_numel), group
        self.fused = hasattr(renderer, "mse_loss_grad_tiled")
        if self.fused:
            self.target_tiled = renderer.geom.tile_grad(target_slab).contiguous()      # padding rows zero, as in H

    def step(self, means, log_scales, rotations, opacities, features):
        """-> (local loss contribution [tensor], grads tuple summed over all ranks)."""
        reduce_fn = lambda d: all_reduce_packed(d, self.group)
        if self.fused:
            self.r.prepare(means, log_scales, rotations, opacities, features)
            Ht = self.r.forward_tiled()
            loss, gH = self.r.mse_loss_grad_tiled(Ht, self.target_tiled, 1.0 / self.numel)
            return loss, self.r.backward_tiled(gH, reduce_fn=reduce_fn)
        H = self.r.forward(means, log_scales, rotations, opacities, features)
        diff = H - self.target
        loss = (diff * diff).sum() / self.numel
        grads = self.r.backward(diff * (2.0 / self.numel), reduce_fn=reduce_fn)
        return loss, grads
