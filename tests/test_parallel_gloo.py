"""world_size-2 `gloo` tests (CPU) of the scan-slab sharding host logic: slab partition, ONE packed-gradient all-reduce per
step, optional transient all-gather.  The renderer is replaced by a CPU stand-in with the same forward/backward(reduce_fn)
protocol (tests/ref_torch.py fp64 restatement), so that the identity
    sum over ranks of slab gradients (after the all-reduce)  ==  gradient of the full-grid loss
is checked exactly as the GPU path relies on it."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nlos_gaussian_renderer_b200 import parallel, scenes
from nlos_gaussian_renderer_b200.geometry import ScanGeometry
from tests import ref_torch, util


def test_slab_rows_partition():
    for Sy, ty, world in [(64, 4, 8), (66, 4, 4), (5, 4, 2), (128, 4, 3), (3, 4, 8)]:
        rows = [parallel.slab_rows(Sy, ty, r, world) for r in range(world)]
        assert rows[0][0] == 0 and rows[-1][1] == Sy
        for (a, b), (c, d) in zip(rows, rows[1:]):
            assert b == c and a <= b
        assert all(a % ty == 0 or a == Sy for a, _ in rows)
        tile_rows = [-(-(b - a) // ty) for a, b in rows]          # near-equal in units of tile rows (the last one may be ragged)
        assert max(tile_rows) - min(tile_rows) <= 1


def test_rank_rows_partition_interleaved():
    for Sy, ty, world in [(64, 4, 8), (66, 4, 4), (5, 4, 2), (128, 4, 3), (3, 4, 8), (128, 4, 8)]:
        parts = [parallel.rank_rows(Sy, ty, r, world) for r in range(world)]
        allr = sorted(sum(parts, []))
        assert allr == list(range(Sy))                               # a partition
        for r, rows in enumerate(parts):
            assert all((y // ty) % world == r for y in rows)          # tile row j -> rank j mod world
            assert rows == sorted(rows)
        n_tr = [len({y // ty for y in rows}) for rows in parts]
        assert max(n_tr) - min(n_tr) <= 1
        assert [parallel.rank_rows(Sy, ty, r, world, interleave=False) for r in range(world)] == \
            [list(range(*parallel.slab_rows(Sy, ty, r, world))) for r in range(world)]


class _CpuStandIn:
    """forward/backward(reduce_fn) protocol of TransientRenderer on the CPU restatement (test infrastructure)."""

    def __init__(self, geom, ta):
        self.g, self.ta = geom, ta
        self.masks = torch.from_numpy(util.ray_masks(geom, self.params0))

    def forward(self, *params):
        g, ta = self.g, self.ta
        self.p = [p.detach().double().requires_grad_() for p in params]
        dd = lambda x: x.double()
        self.H = ref_torch.transient_batch(dd(g.scan_o), dd(g.sin_th), dd(g.cos_th), dd(g.sin_ph), dd(g.cos_ph), dd(ta["t_samples"]),
                                           dd(g.hscale), self.masks, *self.p, ta["c"], ta["deltaT"]).reshape(g.Sy, g.Sx, -1)
        return self.H.detach()

    def backward(self, gH, reduce_fn=None):
        grads = torch.autograd.grad(self.H, self.p, gH)
        packed = torch.cat([x.reshape(x.shape[0], -1) for x in grads], 1).contiguous()      # [N, 3+3+4+1+1]
        if reduce_fn is not None:
            reduce_fn(packed)
        return packed


def _worker(rank, world, port, Sy, out, interleave):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = ScanGeometry.wall_grid(Sy, 3, A=4)
        ta = scenes.time_axis(24)
        params = scenes.make_gaussians(30, 5, scale_mul=3.0)
        target = torch.zeros(Sy, 3, 24, dtype=torch.float64)
        rows = parallel.rank_rows(Sy, full.tile_y, rank, world, interleave)
        slab = full.select_rows(rows)
        _CpuStandIn.params0 = params
        fit = parallel.ShardedFitter(_CpuStandIn(slab, ta), target[rows], Sy * 3 * 24)
        loss, packed = fit.step(*params)
        tot = loss.clone()
        dist.all_reduce(tot)
        Hfull = parallel.gather_transient(fit.r.H.detach(), Sy, full.tile_y, interleave=interleave)
        if rank == 0:
            torch.save({"packed": packed, "loss": tot, "H": Hfull}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("interleave", [True, False])
def test_two_rank_fit_step_equals_single_rank(tmp_path, interleave):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    Sy = 16
    out = str(tmp_path / "r.pt")
    mp.spawn(_worker, args=(2, port, Sy, out, interleave), nprocs=2, join=True)
    got = torch.load(out)
    # single process, full grid
    full = ScanGeometry.wall_grid(Sy, 3, A=4)
    ta = scenes.time_axis(24)
    params = scenes.make_gaussians(30, 5, scale_mul=3.0)
    _CpuStandIn.params0 = params
    fit = parallel.ShardedFitter(_CpuStandIn(full, ta), torch.zeros(Sy, 3, 24, dtype=torch.float64), Sy * 3 * 24)
    loss, packed = fit.step(*params)
    assert float(packed.abs().max()) > 0
    np.testing.assert_allclose(got["packed"].numpy(), packed.numpy(), rtol=1e-10, atol=1e-300)
    np.testing.assert_allclose(float(got["loss"]), float(loss), rtol=1e-12)
    np.testing.assert_allclose(got["H"].numpy(), fit.r.H.detach().numpy(), rtol=1e-12)
