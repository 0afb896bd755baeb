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
VARIANTS = {"wide": 0, "narrow": 1, "auto": 2}


def _f32c(t):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError("libnlosb200 needs CUDA tensors (there is no CPU fallback)")
    if not t.dtype.is_floating_point:
        raise RuntimeError(f"expected a floating-point tensor, got {t.dtype}")
    return t.detach().to(torch.float32).contiguous()


class ListOverflow(RuntimeError):
    """The entry lists of a step did not fit the list buffer (the step rendered nothing); the buffer has been regrown."""


class TransientRenderer:
    """Stateful renderer for one ScanGeometry: owns the preprocess / binning workspaces."""

    def __init__(self, geom: ScanGeometry, T, t0, dt, c_deltaT, mode="numerical", window=96, bucket=None, slice_len=None,
                 mcut=5.3, scaling_modifier=1.0, sigma_bbox=3.0, variant="auto", narrow_rays=2.5, active_sh_degree=0,
                 deterministic=False, one_pass=None, one_pass_min_bins=1.3, culling_words=None):
        """window: bins of the per-warp sliding accumulator (multiple of 8; 96 bins x 12 warps + tables fill the 227 KB of one SM); bucket: bins the window advances by
        (default max(8, next pow2 of T/64): at most 64 buckets); slice_len: list entries a warp pulls at a time from its
        tile's counter (default: 16..256, sized so that every warp of the persistent grid gets >= 32 slices).
        variant: which scan-tile kernel renders an entry -- "wide" (register accumulators, transient.cu), "narrow" (direct to window,
        transient_narrow.cu) or "auto": per (tile, Gaussian) entry at binning time, narrow when the Gaussian's 3-sigma AABB spans at most
        `narrow_rays` lattice pitches (the two kernels give the same result, they differ in speed only; measured crossover: BASELINE
        configs[3], 1.8 hit rays per pair, is 16 % / 7 % faster forward / backward in the narrow kernel, configs[2], 3.5 hit rays, 7 % slower
        forward)."""
        self.geom = geom
        self.T, self.t0, self.dt, self.c_deltaT = int(T), float(t0), float(dt), float(c_deltaT)
        self.mode = MODES[mode] if isinstance(mode, str) else int(mode)
        self.variant = VARIANTS[variant] if isinstance(variant, str) else int(variant)
        self.narrow_rays = float(narrow_rays)
        self.window = int(window)
        if bucket is None:
            bucket = 8
            while -(-self.T // bucket) > 64:
                bucket *= 2
        self.bucket = int(bucket)
        self.n_buckets = -(-self.T // self.bucket)
        self.slice_len = slice_len
        self.mcut, self.mod, self.sigma = float(mcut), float(scaling_modifier), float(sigma_bbox)
        # view-dependent albedo max(SH(dir(mu - o_p)) + 0.5, 0) per (Gaussian, scan point) (volume_renderer.cu:107-111); settable between steps
        # like GaussianModel.active_sh_degree (main.py:240-241).  Needs features [N, >= (deg+1)^2]; a dc-only row falls back to degree 0 (_C.effective_sh_degree)
        self.active_sh_degree = int(active_sh_degree)
        # deterministic=True: bit-identical H and gradients run to run ("deterministic mode, no float atomics, so that races show up
        # as non-reproducibility").  One warp walks a whole tile list in order (no work stealing inside a tile) and the backward accumulates
        # into per-tile partial buffers [n_tiles, N, 16] that are summed in tile order.  Slower and memory-hungry: a debugging mode.
        self.deterministic = bool(deterministic)
        self.wpb = _lib.lib().nlosb200_transient_warps_per_block()
        # analytic mode, one-pass plan: entries of Gaussians whose smallest scale is >= one_pass_min_bins bins render every ray with ONE mixture
        # node into a second buffer of unweighted rows; a fixed 7-tap filter (nlosb200_analytic_filter, <= 1.6e-5 of a ray's peak) turns those
        # rows into bin integrals.  Halves the recurrence work of the 0.8 - 2.5 bin rays that otherwise take two nodes.
        # (one_pass=None: on for the analytic mode unless the scan grid is tiny -- the plan costs two more launches per pass, which is what a
        #  launch-bound scene like BASELINE configs[0] (8 tiles, 1.4 ms per step) would then be made of)
        if one_pass is None:
            one_pass = geom.n_tiles >= 64
        self.one_pass = bool(one_pass) and self.mode == 1
        self.tw_min = float(one_pass_min_bins) * self.dt if self.one_pass else 0.0
        self.n_fp = 2 if self.variant == 2 else 1
        self.n_tw = 2 if self.one_pass else 1
        # culling words: the forward of the narrow kernel leaves, per (list entry, scan point), the bit-exact hit mask of its <= 2 x 2 candidate
        # rays; the backward of the same prepare() reads them instead of repeating the footprint rectangle and the slab tests (64 B per entry:
        # 21 GB at cfg4).  None = on when the lattice allows (A <= 64) and the buffer stays under a quarter of the device memory.
        self.culling_words = culling_words
        self._stash = None
        self._stash_gen = -1
        self.n_classes = self.n_fp * self.n_tw
        self._N = None
        self._list = None
        self._pending = None           # (pinned header copy, event) of the previous prepare(): checked lazily, no host sync per step
        self.generation = 0            # incremented by every prepare(): a backward must see the lists of ITS forward
        self.last_total = 0

    # ------------------------------------------------------------------------------------------
    def _alloc(self, N, dev):
        if self._N == N:
            return
        L = _lib.lib()
        n_cells = self.geom.n_tiles * self.n_buckets * self.n_classes
        self.records = torch.empty(N * L.nlosb200_record_floats() + 32, dtype=torch.float32, device=dev)
        off = ((-self.records.data_ptr()) % 128) // 4         # 128-byte alignment for the TMA gather
        self.records = self.records[off: off + N * L.nlosb200_record_floats()]
        self.shell = torch.empty(N, 4, dtype=torch.float32, device=dev)
        self.chunk_counts = torch.empty(n_cells * L.nlosb200_bin_num_chunks(N), dtype=torch.int32, device=dev)
        self.cell_counts = torch.empty(n_cells, dtype=torch.int32, device=dev)
        self.cell_offsets = torch.empty(n_cells + 1, dtype=torch.int64, device=dev)
        self.tile_items = torch.empty(self.geom.n_tiles + 1, dtype=torch.int32, device=dev)
        self.work = torch.empty(2 * self.geom.n_tiles, dtype=torch.int32, device=dev)  # per-(segment, tile) slice counters (work stealing)
        self.header = torch.zeros(4, dtype=torch.int64, device=dev)
        self.header_host = torch.zeros(4, dtype=torch.int64).pin_memory()
        self.dacc = torch.empty(N, L.nlosb200_grad_floats(), dtype=torch.float32, device=dev)
        self._N = N
        self._list = None
        self._pending = None
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        if self.slice_len is None:
            est = self.geom.n_tiles * N * 1.05                    # ~one entry per (tile, Gaussian)
            # >= 32 slices per warp (fine-grained stealing); small scenes go down to 2 entries per slice so that every warp of the
            # grid gets work (cfg1, 8 tiles x 1000 entries: 16 -> 2 entries per slice took the kernels from 1.6 / 3.5 ms to 0.44 / 0.77 ms)
            self._slice = int(min(256, max(2, est / (n_sm * self.wpb * 32))))
            if self._slice >= 4:
                self._slice = self._slice // 4 * 4
        else:
            self._slice = int(self.slice_len)
        if self.deterministic:
            self._slice = 1 << 30
            self.dacc_part = torch.empty(self.geom.n_tiles, N, L.nlosb200_grad_floats(), dtype=torch.float32, device=dev)

    def _bin_args(self):
        g = self.geom
        return (g.tiles().data_ptr(), g.n_tiles, g.A, self.shell.data_ptr(), self._N, self.t0, self.dt, self.T, self.window, self.bucket,
                self.n_classes, self.narrow_rays if self.n_fp == 2 else 0.0, self.mcut, int(self.n_tw == 2))

    def _check_pending(self):
        """Lazy overflow check of the PREVIOUS prepare(): its header copy finished long ago, so this does not stall the stream."""
        if self._pending is None:
            return
        ev, gen, capacity = self._pending
        self._pending = None
        ev.synchronize()
        total, overflow = int(self.header_host[0]), int(self.header_host[2])
        self.last_total = total
        if overflow or total > capacity:
            self._grow(total)
            raise ListOverflow(f"the entry lists of step {gen} ({total} entries) did not fit the list buffer: that step rendered nothing; "
                               "the buffer has been regrown, repeat the step")

    def _grow(self, total):
        dev = self.records.device
        self._list = torch.empty(max(int(total * 1.25) + 1024, 1024), dtype=torch.int32, device=dev)
        self._list_bins = torch.empty_like(self._list)                # per entry: shell klo | khi << 16
        self._stash, self._stash_gen = None, -1
        want = self.culling_words
        nbytes = self._list.numel() * TILE * 2
        if want is None:
            want = nbytes <= torch.cuda.get_device_properties(dev).total_memory // 4
        if want and self.geom.A <= 64 and any(v == 1 for v, _ in self._launch_groups()):
            self._stash = torch.empty(self._list.numel() * TILE, dtype=torch.int16, device=dev)

    def prepare(self, means, log_scales, rotations, opacities, features):
        """preprocess + shell binning for the current parameters.  Stream ordered: the first call sizes the list buffer with one 32-byte
        readback; later calls reuse the buffer (25 % head-room over the last known total) and read the header back asynchronously --
        an overflow is detected one call later (ListOverflow) and by check_lists()."""
        means, log_scales, rotations = _f32c(means), _f32c(log_scales), _f32c(rotations)
        opacities, features = _f32c(opacities).reshape(-1), _f32c(features)
        N = means.shape[0]
        features = features.reshape(N, -1)
        self._alloc(N, means.device)
        self._check_pending()
        self._params = (means, log_scales, rotations, opacities, features)
        from ._C import effective_sh_degree
        self._deg = effective_sh_degree(self.active_sh_degree, features.shape[1])
        if self._deg > 0 and (getattr(self, "dsh", None) is None or self.dsh.shape[0] != N):
            self.dsh = torch.empty(N, 16, dtype=torch.float32, device=means.device)
            if self.deterministic:
                self.dsh_part = torch.empty(self.geom.n_tiles, N, 16, dtype=torch.float32, device=means.device)
        self.generation += 1
        st = _lib.stream()
        _lib.call("nlosb200_preprocess", means.data_ptr(), log_scales.data_ptr(), rotations.data_ptr(), opacities.data_ptr(),
                  features.data_ptr(), N, features.shape[1], self.mod, self.sigma, self.mcut, self.tw_min, self.records.data_ptr(),
                  self.shell.data_ptr(), None, st)
        _lib.call("nlosb200_bin_window_plan", *self._bin_args(), self._slice, self.wpb,
                  self.chunk_counts.data_ptr(), self.cell_counts.data_ptr(), self.cell_offsets.data_ptr(), self.tile_items.data_ptr(),
                  self.header.data_ptr(), st)
        if self._list is None:
            total = int(self.header.cpu()[0])                 # first call only: size the list buffer
            self.last_total = total
            self._grow(total)
        _lib.call("nlosb200_bin_window_fill", *self._bin_args(), self.chunk_counts.data_ptr(),
                  self.cell_offsets.data_ptr(), self._list.data_ptr(), self._list_bins.data_ptr(), self._list.numel(), self.header.data_ptr(), st)
        self.header_host.copy_(self.header, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._pending = (ev, self.generation, self._list.numel())

    def check_lists(self):
        """Block until the last prepare() is known to have fitted its lists (raises ListOverflow otherwise) -> total entries."""
        self._check_pending()
        return self.last_total

    def _launches(self):
        """(cls, variant, one_pass) per entry class: class = footprint bit + n_fp * time-wide bit.  Footprint bit 0 = narrow entries ->
        direct-to-window kernel, 1 = wide -> register accumulators; time-wide entries render one-pass."""
        out = []
        for tw in range(self.n_tw):
            for fp in range(self.n_fp):
                var = (1 if fp == 0 else 0) if self.n_fp == 2 else self.variant
                out.append((fp + self.n_fp * tw, var, tw))
        return tuple(out)

    def _launch_groups(self):
        """Render launches: the classes of one kernel variant go into ONE launch as its (up to two) segments, so a pass is one launch per
        kernel and the CTAs that drain one class move on to the other instead of idling to the launch boundary.  The deterministic mode
        keeps one class per launch (its per-tile partial buffers are owned by one warp at a time)."""
        groups = []
        for cls, var, op in self._launches():
            for g in groups:
                if g[0] == var and len(g[1]) < 2 and not self.deterministic:
                    g[1].append((cls, op))
                    break
            else:
                groups.append((var, [(cls, op)]))
        return groups

    def _render_args(self, cls):
        g = self.geom
        return (g.tiles().data_ptr(), g.n_tiles, g.A, self.records.data_ptr(), self._N, self._list.data_ptr(),
                self._list_bins.data_ptr(), self._list.numel(), self.cell_offsets.data_ptr(), self.n_classes, cls, self.work.data_ptr(),
                self.t0, self.dt, self.T, self.window, self.bucket, self._slice)

    def _sh_args(self):
        f = self._params[4]
        return (f.data_ptr() if self._deg > 0 else None, f.shape[1], self._deg)

    def forward_tiled(self):
        """-> H [n_tiles*32, T] in tile-major scan order (after prepare())."""
        H = torch.zeros(self.geom.n_tiles * TILE, self.T, dtype=torch.float32, device=self.records.device)
        Ha = torch.zeros_like(H) if self.one_pass else None          # unweighted one-node rows of the time-wide entries
        for var, segs in self._launch_groups():
            (cls, op), (cls2, op2) = segs[0], (segs[1] if len(segs) > 1 else (-1, 0))
            _lib.call("nlosb200_transient_forward", *self._render_args(cls), self.c_deltaT, self.mcut, self.mode, var, op, 1,
                      *self._sh_args(), (Ha if op else H).data_ptr(), cls2, op2, (Ha if op2 else H).data_ptr(),
                      self._stash.data_ptr() if self._stash is not None else None, _lib.stream())
        self._stash_gen = self.generation if self._stash is not None else -1
        if self.one_pass:
            _lib.call("nlosb200_analytic_filter", Ha.data_ptr(), H.data_ptr(), H.shape[0], self.T, self.t0, self.dt, 0, _lib.stream())
        return H

    def forward(self, means, log_scales, rotations, opacities, features):
        self.prepare(means, log_scales, rotations, opacities, features)
        return self.geom.untile(self.forward_tiled())

    def render_backward(self, gH_tiled):
        """transient backward kernels only: dL/dH (tile-major) -> the packed [N,16] gradient buffer self.dacc."""
        gH_tiled = _f32c(gH_tiled)
        det = self.deterministic
        dacc = self.dacc_part if det else self.dacc
        dsh = (self.dsh_part if det else self.dsh) if self._deg > 0 else None
        gHa = None
        if self.one_pass:                                            # weights of the one-pass launches: F(w gH)
            gHa = torch.empty_like(gH_tiled)
            _lib.call("nlosb200_analytic_filter", gH_tiled.data_ptr(), gHa.data_ptr(), gH_tiled.shape[0], self.T, self.t0, self.dt, 1, _lib.stream())
        for n, (var, segs) in enumerate(self._launch_groups()):
            (cls, op), (cls2, op2) = segs[0], (segs[1] if len(segs) > 1 else (-1, 0))
            _lib.call("nlosb200_transient_backward", *self._render_args(cls), self.c_deltaT, self.mcut, self.mode, var, op, int(n > 0),
                      *self._sh_args(), (gHa if op else gH_tiled).data_ptr(), cls2, op2, (gHa if op2 else gH_tiled).data_ptr(),
                      dacc.data_ptr(), dsh.data_ptr() if dsh is not None else None, int(det),
                      self._stash.data_ptr() if (self._stash is not None and self._stash_gen == self.generation) else None, _lib.stream())
        if det:
            _lib.call("nlosb200_reduce_tile_partials", dacc.data_ptr(), self.geom.n_tiles, self.dacc.numel(), self.dacc.data_ptr(), _lib.stream())
            if dsh is not None:
                _lib.call("nlosb200_reduce_tile_partials", dsh.data_ptr(), self.geom.n_tiles, self.dsh.numel(), self.dsh.data_ptr(), _lib.stream())
        return self.dacc

    def chain(self, params=None):
        """per-Gaussian chain rule: self.dacc -> (d_means, d_log_scales, d_rotations, d_opacities, d_features)."""
        means, log_scales, rotations, opacities, features = params if params is not None else self._params
        outs = [torch.empty_like(means), torch.empty_like(log_scales), torch.empty_like(rotations), torch.empty_like(opacities),
                torch.empty_like(features)]
        if self._deg > 0:
            _lib.call("nlosb200_preprocess_backward_sh", log_scales.data_ptr(), rotations.data_ptr(), opacities.data_ptr(), features.data_ptr(),
                      self._N, features.shape[1], self._deg, self.mod, self.dacc.data_ptr(), self.dsh.data_ptr(), *[o.data_ptr() for o in outs], 0,
                      _lib.stream())
        else:
            _lib.call("nlosb200_preprocess_backward", log_scales.data_ptr(), rotations.data_ptr(), opacities.data_ptr(), features.data_ptr(),
                      self._N, features.shape[1], self.mod, self.dacc.data_ptr(), *[o.data_ptr() for o in outs], 0, _lib.stream())
        return tuple(outs)

    def backward_tiled(self, gH_tiled, reduce_fn=None):
        """dL/dH (tile-major) -> (d_means, d_log_scales, d_rotations, d_opacities, d_features).
        reduce_fn(dacc) runs between the render backward and the per-Gaussian chain: the multi-GPU path
        all-reduces the packed [N,16] buffer there (ONE NCCL all-reduce per fitting step)."""
        self.render_backward(gH_tiled)
        if reduce_fn is not None:
            reduce_fn(self.dacc)
        return self.chain()

    def backward(self, gH, reduce_fn=None):
        return self.backward_tiled(self.geom.tile_grad(gH), reduce_fn)

    def mse_loss_grad_tiled(self, H_tiled, target_tiled, inv_numel, want_grad=True):
        """Fused  loss = inv_numel * sum (H - target)^2  and  dL/dH = 2 inv_numel (H - target)  on the tile-major layout
        (csrc/loss.cu; the batched counterpart of nlos_helpers.py:280-346).  -> (loss 0-d float64 tensor, dL/dH tile-major or None)."""
        H_tiled, target_tiled = _f32c(H_tiled), _f32c(target_tiled)
        if H_tiled.shape != target_tiled.shape:
            raise RuntimeError(f"H {tuple(H_tiled.shape)} and target {tuple(target_tiled.shape)} differ in shape")
        loss = torch.zeros((), dtype=torch.float64, device=H_tiled.device)
        gH = torch.empty_like(H_tiled) if want_grad else None
        _lib.call("nlosb200_mse_loss_grad", H_tiled.data_ptr(), target_tiled.data_ptr(), H_tiled.numel(), float(inv_numel),
                  gH.data_ptr() if want_grad else None, loss.data_ptr(), _lib.stream())
        return loss, gH

    def count_work(self):
        """SURVEY.md 8(d) work counters for the prepared scene (dict)."""
        c = torch.zeros(4, dtype=torch.int64, device=self.records.device)
        for cls in range(self.n_classes):
            _lib.call("nlosb200_transient_count", *self._render_args(cls), self.mcut, int(cls > 0), c.data_ptr(), _lib.stream())
        c = c.cpu().tolist()
        return {"hits": c[0], "W_s": c[1], "evals": c[2], "W_b": c[3]}

    def lists(self):
        """(cell_counts, cell_offsets, list) as CPU numpy arrays -- for the bit-exact binning tests."""
        total = self.check_lists()
        import numpy as np
        return (self.cell_counts.cpu().numpy(), self.cell_offsets.cpu().numpy(), self._list[:total].cpu().numpy().view(np.uint32))


class _TransientFn(torch.autograd.Function):
    """autograd boundary, same role as CUDARenderFunction (gaussian_model/cuda_autograd.py:18-191).  The five parameter tensors are
    saved on ctx (an in-place update between forward and backward trips autograd's version check); the binning state lives on the
    renderer, so a backward whose renderer has since prepared another scene re-runs prepare() from the saved tensors first."""

    @staticmethod
    def forward(ctx, renderer, means, log_scales, rotations, opacities, features):
        out = renderer.forward(means, log_scales, rotations, opacities, features)
        ctx.renderer = renderer
        ctx.generation = renderer.generation
        ctx.shapes = (opacities.shape, features.shape)
        ctx.save_for_backward(means, log_scales, rotations, opacities, features)
        return out

    @staticmethod
    def backward(ctx, gH):
        r = ctx.renderer
        if r.generation != ctx.generation:          # another forward()/prepare() ran on this renderer since: rebuild this call's lists
            r.prepare(*ctx.saved_tensors)
            ctx.generation = r.generation
        dm, ds, dr, do, df = r.backward(gH)
        return None, dm, ds, dr, do.reshape(ctx.shapes[0]), df.reshape(ctx.shapes[1])


def render_transient_batch(renderer, means, log_scales, rotations, opacities, features):
    """Differentiable [Sy,Sx,T] transient of the whole scan grid."""
    return _TransientFn.apply(renderer, means, log_scales, rotations, opacities, features)
