"""Batched transient renderer  H[scan_y, scan_x, time_bin]  with a real backward pass.

Host-side mirror of the reference operator boundary (gaussian_model/cuda_autograd.py:27-191,
CUDARenderFunction / CUDARenderModule) generalised from ONE scan point per call
(main.py:198-203, nlos_helpers.py:302-306) to the whole relay-wall grid; per scan point the
result equals `pred_histogram * volume_position[1]**2` of nlos_helpers.py:255-276.

All compute happens in libnlosb200.so (hand-written sm_100a CUDA) through the C ABI; torch only
owns device memory and streams.  There is no CPU or eager fallback.
"""
import torch

from . import _lib
from .geometry import ScanGeometry, TILE

MODES = {"numerical": 0, "analytic": 1, "analytic_exact": 1}
N_SM = 148                    # B200


def _f32c(t):
    assert t.is_cuda, "libnlosb200 needs CUDA tensors (there is no CPU fallback)"
    return t.detach().to(torch.float32).contiguous()


class TransientRenderer:
    """Stateful renderer for one ScanGeometry: owns the preprocess / binning workspaces."""

    def __init__(self, geom: ScanGeometry, T, t0, dt, c_deltaT, mode="numerical", window=96, bucket=None, slice_len=None,
                 mcut=5.3, scaling_modifier=1.0, sigma_bbox=3.0):
        """window: bins of the per-warp sliding accumulator (multiple of 8; 96 bins x 12 warps + tables fill the 227 KB of one SM); bucket: bins the window advances by
        (default max(8, next pow2 of T/64): at most 64 buckets); slice_len: list entries a warp pulls at a time from its
        tile's counter (default: 16..256, sized so that every warp of the persistent grid gets >= 32 slices)."""
        self.geom = geom
        self.T, self.t0, self.dt, self.c_deltaT = int(T), float(t0), float(dt), float(c_deltaT)
        self.mode = MODES[mode] if isinstance(mode, str) else int(mode)
        self.window = int(window)
        if bucket is None:
            bucket = 8
            while -(-self.T // bucket) > 64:
                bucket *= 2
        self.bucket = int(bucket)
        self.n_buckets = -(-self.T // self.bucket)
        self.slice_len = slice_len
        self.mcut, self.mod, self.sigma = float(mcut), float(scaling_modifier), float(sigma_bbox)
        self.wpb = _lib.lib().nlosb200_transient_warps_per_block()
        self._N = None
        self._list = None
        self.last_total = 0
        self.n_items = 0

    # ------------------------------------------------------------------------------------------
    def _alloc(self, N, dev):
        if self._N == N:
            return
        L = _lib.lib()
        n_cells = self.geom.n_tiles * self.n_buckets
        self.records = torch.empty(N * L.nlosb200_record_floats() + 32, dtype=torch.float32, device=dev)
        off = ((-self.records.data_ptr()) % 128) // 4         # 128-byte alignment for the TMA gather
        self.records = self.records[off: off + N * L.nlosb200_record_floats()]
        self.shell = torch.empty(N, 4, dtype=torch.float32, device=dev)
        self.chunk_counts = torch.empty(n_cells * L.nlosb200_bin_num_chunks(N), dtype=torch.int32, device=dev)
        self.cell_counts = torch.empty(n_cells, dtype=torch.int32, device=dev)
        self.cell_offsets = torch.empty(n_cells + 1, dtype=torch.int64, device=dev)
        self.tile_items = torch.empty(self.geom.n_tiles + 1, dtype=torch.int32, device=dev)
        self.work = torch.empty(self.geom.n_tiles, dtype=torch.int32, device=dev)      # per-tile slice counters (work stealing)
        self.header = torch.zeros(4, dtype=torch.int64, device=dev)
        self.dacc = torch.empty(N, L.nlosb200_grad_floats(), dtype=torch.float32, device=dev)
        self._N = N
        if self.slice_len is None:
            est = self.geom.n_tiles * N * 1.05                    # ~one entry per (tile, Gaussian)
            # >= 32 slices per warp (fine-grained stealing); small scenes go down to 2 entries per slice so that every warp of the
            # grid gets work (cfg1, 8 tiles x 1000 entries: 16 -> 2 entries per slice took the kernels from 1.6 / 3.5 ms to 0.44 / 0.77 ms)
            self._slice = int(min(256, max(2, est / (N_SM * self.wpb * 32))))
            if self._slice >= 4:
                self._slice = self._slice // 4 * 4
        else:
            self._slice = int(self.slice_len)

    def _bin_args(self):
        g = self.geom
        return (g.tiles().data_ptr(), g.n_tiles, g.A, self.shell.data_ptr(), self._N, self.t0, self.dt, self.T, self.window, self.bucket)

    def prepare(self, means, log_scales, rotations, opacities, features):
        """preprocess + shell binning for the current parameters (stream ordered; ONE 16-byte readback
        to size the list buffer and the launch grid)."""
        means, log_scales, rotations = _f32c(means), _f32c(log_scales), _f32c(rotations)
        opacities, features = _f32c(opacities).reshape(-1), _f32c(features)
        N = means.shape[0]
        features = features.reshape(N, -1)
        self._params = (means, log_scales, rotations, opacities, features)
        self._alloc(N, means.device)
        st = _lib.stream()
        _lib.call("nlosb200_preprocess", means.data_ptr(), log_scales.data_ptr(), rotations.data_ptr(), opacities.data_ptr(),
                  features.data_ptr(), N, features.shape[1], self.mod, self.sigma, self.mcut, self.records.data_ptr(),
                  self.shell.data_ptr(), None, st)
        _lib.call("nlosb200_bin_window_plan", *self._bin_args(), self._slice, self.wpb, self.chunk_counts.data_ptr(),
                  self.cell_counts.data_ptr(), self.cell_offsets.data_ptr(), self.tile_items.data_ptr(), self.header.data_ptr(), st)
        hdr = self.header.cpu()
        total, self.n_items = int(hdr[0]), int(hdr[1])
        self.last_total = total
        if self._list is None or self._list.numel() < total:
            self._list = torch.empty(max(int(total * 1.25), 1024), dtype=torch.int32, device=means.device)
            self._list_bins = torch.empty_like(self._list)                # per entry: shell klo | khi << 16
        _lib.call("nlosb200_bin_window_fill", *self._bin_args(), self.chunk_counts.data_ptr(), self.cell_offsets.data_ptr(),
                  self._list.data_ptr(), self._list_bins.data_ptr(), self._list.numel(), self.header.data_ptr(), st)

    def _render_args(self):
        g = self.geom
        return (g.tiles().data_ptr(), g.n_tiles, g.A, self.records.data_ptr(), self._N, self._list.data_ptr(),
                self._list_bins.data_ptr(), self.cell_offsets.data_ptr(), self.work.data_ptr(), self.t0, self.dt, self.T, self.window, self.bucket, self._slice)

    def forward_tiled(self):
        """-> H [n_tiles*32, T] in tile-major scan order (after prepare())."""
        H = torch.empty(self.geom.n_tiles * TILE, self.T, dtype=torch.float32, device=self.records.device)
        _lib.call("nlosb200_transient_forward", *self._render_args(), self.c_deltaT, self.mcut, self.mode, H.data_ptr(), _lib.stream())
        return H

    def forward(self, means, log_scales, rotations, opacities, features):
        self.prepare(means, log_scales, rotations, opacities, features)
        return self.geom.untile(self.forward_tiled())

    def backward_tiled(self, gH_tiled, reduce_fn=None):
        """dL/dH (tile-major) -> (d_means, d_log_scales, d_rotations, d_opacities, d_features).
        reduce_fn(dacc) runs between the render backward and the per-Gaussian chain: the multi-GPU path
        all-reduces the packed [N,16] buffer there (ONE NCCL all-reduce per fitting step)."""
        means, log_scales, rotations, opacities, features = self._params
        N = self._N
        gH_tiled = _f32c(gH_tiled)
        st = _lib.stream()
        _lib.call("nlosb200_transient_backward", *self._render_args(), self.c_deltaT, self.mcut, self.mode, gH_tiled.data_ptr(),
                  self.dacc.data_ptr(), st)
        if reduce_fn is not None:
            reduce_fn(self.dacc)
        outs = [torch.empty_like(means), torch.empty_like(log_scales), torch.empty_like(rotations), torch.empty_like(opacities),
                torch.empty_like(features)]
        _lib.call("nlosb200_preprocess_backward", log_scales.data_ptr(), rotations.data_ptr(), opacities.data_ptr(), features.data_ptr(),
                  N, features.shape[1], self.mod, self.dacc.data_ptr(), *[o.data_ptr() for o in outs], 0, st)
        return tuple(outs)

    def backward(self, gH, reduce_fn=None):
        return self.backward_tiled(self.geom.tile_grad(gH), reduce_fn)

    def mse_loss_grad_tiled(self, H_tiled, target_tiled, inv_numel, want_grad=True):
        """Fused  loss = inv_numel * sum (H - target)^2  and  dL/dH = 2 inv_numel (H - target)  on the tile-major layout
        (csrc/loss.cu; the batched counterpart of nlos_helpers.py:280-346).  -> (loss 0-d float64 tensor, dL/dH tile-major or None)."""
        H_tiled, target_tiled = _f32c(H_tiled), _f32c(target_tiled)
        assert H_tiled.shape == target_tiled.shape
        loss = torch.zeros((), dtype=torch.float64, device=H_tiled.device)
        gH = torch.empty_like(H_tiled) if want_grad else None
        _lib.call("nlosb200_mse_loss_grad", H_tiled.data_ptr(), target_tiled.data_ptr(), H_tiled.numel(), float(inv_numel),
                  gH.data_ptr() if want_grad else None, loss.data_ptr(), _lib.stream())
        return loss, gH

    def count_work(self):
        """SURVEY.md 8(d) work counters for the prepared scene (dict)."""
        c = torch.zeros(4, dtype=torch.int64, device=self.records.device)
        _lib.call("nlosb200_transient_count", *self._render_args(), self.mcut, c.data_ptr(), _lib.stream())
        c = c.cpu().tolist()
        return {"hits": c[0], "W_s": c[1], "evals": c[2], "W_b": c[3]}

    def lists(self):
        """(cell_counts, cell_offsets, list) as CPU numpy arrays -- for the bit-exact binning tests."""
        total = self.last_total
        import numpy as np
        return (self.cell_counts.cpu().numpy(), self.cell_offsets.cpu().numpy(), self._list[:total].cpu().numpy().view(np.uint32))


class _TransientFn(torch.autograd.Function):
    """autograd boundary, same role as CUDARenderFunction (gaussian_model/cuda_autograd.py:18-191)."""

    @staticmethod
    def forward(ctx, renderer, means, log_scales, rotations, opacities, features):
        ctx.renderer = renderer
        ctx.shapes = (opacities.shape, features.shape)
        return renderer.forward(means, log_scales, rotations, opacities, features)

    @staticmethod
    def backward(ctx, gH):
        r = ctx.renderer
        dm, ds, dr, do, df = r.backward(gH)
        return None, dm, ds, dr, do.reshape(ctx.shapes[0]), df.reshape(ctx.shapes[1])


def render_transient_batch(renderer, means, log_scales, rotations, opacities, features):
    """Differentiable [Sy,Sx,T] transient of the whole scan grid."""
    return _TransientFn.apply(renderer, means, log_scales, rotations, opacities, features)
