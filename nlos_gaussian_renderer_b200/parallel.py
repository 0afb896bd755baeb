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


def rank_rows(Sy, tile_y, rank, world, interleave=True):
    """Scan rows of `rank` as a sorted index list.  interleave=True: tile rows (blocks of tile_y scan rows) are dealt round-robin,
    tile row j -> rank j mod world, so that every rank gets near and far rows of the relay wall alike (the work per row varies with its
    distance to the hidden volume: contiguous slabs left the 8-GPU step of BASELINE configs[3] waiting for the slowest slab);
    interleave=False: the contiguous slab of slab_rows()."""
    if not interleave:
        y0, y1 = slab_rows(Sy, tile_y, rank, world)
        return list(range(y0, y1))
    n_tr = -(-Sy // tile_y)
    rows = []
    for j in range(rank, n_tr, world):
        rows.extend(range(j * tile_y, min((j + 1) * tile_y, Sy)))
    return rows


def all_reduce_packed(dacc, group=None):
    """In-place sum of the packed gradient buffer over ranks (no-op for world size 1)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(dacc, op=dist.ReduceOp.SUM, group=group)
    return dacc


def gather_transient(H_slab, Sy, tile_y, group=None, interleave=True):
    """All-gather the per-rank row sets [rows_r, Sx, T] into the full [Sy, Sx, T] transient (optional; fitting never needs it).
    `interleave` must be the partition the ranks rendered with (rank_rows)."""
    world = dist.get_world_size(group)
    rows = [rank_rows(Sy, tile_y, r, world, interleave) for r in range(world)]
    mx = max(len(r) for r in rows)
    pad = torch.zeros(mx, *H_slab.shape[1:], dtype=H_slab.dtype, device=H_slab.device)
    pad[: H_slab.shape[0]] = H_slab
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    full = torch.empty(Sy, *H_slab.shape[1:], dtype=H_slab.dtype, device=H_slab.device)
    for o, r in zip(outs, rows):
        if r:
            full[torch.as_tensor(r, device=full.device)] = o[: len(r)]
    return full


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
        def reduce_fn(d):
            all_reduce_packed(d, self.group)
            if getattr(self.r, "_deg", 0) > 0:                       # view-dependent albedo: the SH-coefficient gradients live in a second buffer
                all_reduce_packed(self.r.dsh, self.group)
            return d
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
