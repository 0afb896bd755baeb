"""Scan geometry of a confocal NLOS capture: relay-wall scan points, the per-scan-point spherical
lattice that covers the hidden volume, and the tile-major packing the CUDA kernels consume.

Restates, batched over scan points, what the reference does once per optimiser step:
  nlos_helpers.py:107-118   volume_box_point
  nlos_helpers.py:86-94     cartesian2spherical_torch
  nlos_helpers.py:124-188   spherical_sample_histogram (theta/phi range = min/max over the 8 box corners,
                            linspace lattices, dtheta = (max-min)/A)
  gaussian_model/cuda_autograd.py:254-268,311-314  ray directions and the dtheta*dphi quadrature weight
  nlos_helpers.py:275-276   the x volume_position[1]^2 factor
This is host-side plumbing (torch); it runs once per scan grid, not per step.
"""
import math

import torch

from . import _lib

TILE = 32


def volume_box_point(volume_position, volume_size):
    """8 corners of the hidden-volume cube, same corner order as nlos_helpers.py:107-118."""
    xv, yv, zv = [float(v) for v in volume_position]
    h = float(volume_size) / 2
    pts = [[xv + sx * h, yv + sy * h, zv + sz * h] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)]
    return torch.tensor(pts, dtype=torch.float32)


def torch_linspace_rows(lo, hi, steps):
    """Row i = torch.linspace(lo[i], hi[i], steps, dtype=float32) on lo's device, through torch.linspace ITSELF, one call per row with the
    bounds as Python floats -- exactly what the reference does per scan point (nlos_helpers.py:159-160 after the .item() calls of :153-156).
    A closed-form restatement is not bit-stable: torch's CPU kernel evaluates (start + step * i0) + step * j per SIMD chunk and its CUDA
    kernel contracts start + step * i into one FMA, so the last bit of a lattice angle depends on the device the lattice is built on
    (tests/test_ref_python.py pins this function against the reference's own output).  Host-side, once per scan grid."""
    dev = lo.device
    lo_l, hi_l = lo.detach().to("cpu", torch.float32).tolist(), hi.detach().to("cpu", torch.float32).tolist()
    if not lo_l:
        return torch.empty(0, steps, dtype=torch.float32, device=dev)
    return torch.stack([torch.linspace(a, b, steps, dtype=torch.float32, device=dev) for a, b in zip(lo_l, hi_l)])


class ScanGeometry:
    """Scan points [Sy,Sx,3] + lattice tables, packed tile-major (tile = tile_y x tile_x block, 32 points)."""

    def __init__(self, scan_points, theta_range, phi_range, A, vp_y=1.0, tile=(4, 8), device=None):
        """scan_points [Sy,Sx,3]; theta_range / phi_range [Sy,Sx,2] (min,max) per scan point."""
        scan_points = torch.as_tensor(scan_points, dtype=torch.float32)
        device = torch.device(device) if device is not None else scan_points.device
        assert scan_points.dim() == 3 and scan_points.shape[-1] == 3
        self.Sy, self.Sx = int(scan_points.shape[0]), int(scan_points.shape[1])
        self.A = int(A)
        self.vp_y = float(vp_y)
        self.tile_y, self.tile_x = int(tile[0]), int(tile[1])
        assert self.tile_y * self.tile_x == TILE, "a tile holds exactly 32 scan points"
        self.device = device
        S = self.Sy * self.Sx
        self.S = S
        sp = scan_points.reshape(S, 3).to(device)
        th = torch.as_tensor(theta_range, dtype=torch.float32).reshape(S, 2).to(device)
        ph = torch.as_tensor(phi_range, dtype=torch.float32).reshape(S, 2).to(device)
        self.scan_o = sp.contiguous()
        self.theta_range, self.phi_range = th, ph
        self.theta = torch_linspace_rows(th[:, 0], th[:, 1], self.A).contiguous()          # [S,A]
        self.phi = torch_linspace_rows(ph[:, 0], ph[:, 1], self.A).contiguous()
        self.sin_th, self.cos_th = torch.sin(self.theta).contiguous(), torch.cos(self.theta).contiguous()
        self.sin_ph, self.cos_ph = torch.sin(self.phi).contiguous(), torch.cos(self.phi).contiguous()
        # quadrature weight: dtheta*dphi with the reference's /A (not /(A-1)), times vp_y^2
        dth = (th[:, 1].double() - th[:, 0].double()) / self.A
        dph = (ph[:, 1].double() - ph[:, 0].double()) / self.A
        self.hscale = (dth * dph * self.vp_y ** 2).to(torch.float32).contiguous()
        den = max(self.A - 1, 1)
        self.lattice = torch.stack([th[:, 0], (th[:, 1] - th[:, 0]) / den, ph[:, 0], (ph[:, 1] - ph[:, 0]) / den], 1).contiguous()
        # tile-major permutation (tile index = ty*ntx + tx, inside a tile y-major) with -1 padding
        nty, ntx = -(-self.Sy // self.tile_y), -(-self.Sx // self.tile_x)
        self.n_tiles = nty * ntx
        yy = torch.arange(nty * self.tile_y).reshape(nty, 1, self.tile_y, 1)
        xx = torch.arange(ntx * self.tile_x).reshape(1, ntx, 1, self.tile_x)
        flat = (yy * self.Sx + xx).expand(nty, ntx, self.tile_y, self.tile_x)
        ok = ((yy < self.Sy) & (xx < self.Sx)).expand_as(flat)
        perm = torch.where(ok, flat, torch.full_like(flat, -1)).reshape(-1)
        self.perm = perm.to(device)                                   # [n_tiles*32] -> flat scan index or -1
        self.valid = (self.perm >= 0)
        self._tiles = None

    # -- construction helpers -------------------------------------------------------------------
    @classmethod
    def from_volume(cls, scan_points, volume_position, volume_size, A, tile=(4, 8), device=None):
        """Lattice bounds exactly as spherical_sample_histogram (nlos_helpers.py:145-156)."""
        scan_points = torch.as_tensor(scan_points, dtype=torch.float32)
        box = volume_box_point(volume_position, volume_size).to(scan_points.device)            # [8,3]
        rel = box[None, None, :, :] - scan_points[:, :, None, :]                                # [Sy,Sx,8,3]
        r = torch.linalg.norm(rel, dim=-1)
        theta = torch.acos(rel[..., 2] / r)
        phi = torch.atan2(rel[..., 1], rel[..., 0])
        th = torch.stack([theta.min(-1).values, theta.max(-1).values], -1)
        ph = torch.stack([phi.min(-1).values, phi.max(-1).values], -1)
        return cls(scan_points, th, ph, A, vp_y=float(volume_position[1]), tile=tile, device=device)

    @classmethod
    def wall_grid(cls, Sy, Sx, volume_position=(0.0, 0.75, 0.0), volume_size=0.5, A=32, extent=1.0, tile=(4, 8), device=None):
        """Synthetic relay wall y=0, scan grid uniform on x,z in [-extent/2, extent/2]."""
        xs = torch.linspace(-extent / 2, extent / 2, Sx) if Sx > 1 else torch.zeros(1)
        zs = torch.linspace(-extent / 2, extent / 2, Sy) if Sy > 1 else torch.zeros(1)
        pts = torch.stack([xs[None, :].expand(Sy, Sx), torch.zeros(Sy, Sx), zs[:, None].expand(Sy, Sx)], -1)
        return cls.from_volume(pts, volume_position, volume_size, A, tile=tile, device=device)

    def rows(self, y0, y1):
        """Sub-geometry of scan rows [y0, y1) -- the slab one GPU renders."""
        sl = slice(y0 * self.Sx, y1 * self.Sx)
        n = y1 - y0
        return ScanGeometry(self.scan_o[sl].reshape(n, self.Sx, 3), self.theta_range[sl].reshape(n, self.Sx, 2),
                            self.phi_range[sl].reshape(n, self.Sx, 2), self.A, vp_y=self.vp_y, tile=(self.tile_y, self.tile_x),
                            device=self.device)

    def select_rows(self, rows):
        """Sub-geometry of the given scan rows (any subset, kept in the given order) -- what one GPU of a row-interleaved sharded run
        renders (parallel.rank_rows)."""
        idx = torch.as_tensor(list(rows), dtype=torch.long, device=self.scan_o.device)
        n = int(idx.numel())
        pick = lambda x, k: x.reshape(self.Sy, self.Sx, k)[idx]
        return ScanGeometry(pick(self.scan_o, 3), pick(self.theta_range, 2), pick(self.phi_range, 2), self.A, vp_y=self.vp_y,
                            tile=(self.tile_y, self.tile_x), device=self.device) if n else None

    # -- packing ---------------------------------------------------------------------------------
    def gather_tile_major(self, x):
        """[S,...] -> [n_tiles*32,...] in tile-major order (padding lanes repeat index 0)."""
        return x[self.perm.clamp_min(0)].contiguous()

    def tiles(self):
        """Device block [n_tiles, tile_floats(A)] built by the CUDA packing kernel (cached)."""
        if self._tiles is None:
            assert self.device.type == "cuda", "scan tiles live on the GPU"
            g = self.gather_tile_major
            n = self.n_tiles * TILE
            out = torch.empty(self.n_tiles * _lib.lib().nlosb200_tile_floats(self.A), dtype=torch.float32, device=self.device)
            valid = self.valid.to(torch.int32).contiguous()
            args = [g(self.scan_o), g(self.sin_th), g(self.cos_th), g(self.sin_ph), g(self.cos_ph), g(self.lattice), g(self.hscale)]
            _lib.call("nlosb200_build_scan_tiles", *[a.data_ptr() for a in args], valid.data_ptr(), n, self.A, out.data_ptr(),
                      _lib.stream())
            self._keep = args + [valid]
            self._tiles = out
        return self._tiles

    def untile(self, H_tiled):
        """[n_tiles*32, T] tile-major -> [Sy,Sx,T]."""
        T = H_tiled.shape[-1]
        out = torch.empty(self.S, T, dtype=H_tiled.dtype, device=H_tiled.device)
        out[self.perm[self.valid]] = H_tiled[self.valid]
        return out.reshape(self.Sy, self.Sx, T)

    def tile_grad(self, gH):
        """[Sy,Sx,T] -> [n_tiles*32, T] tile-major (padding rows zero)."""
        T = gH.shape[-1]
        flat = gH.reshape(self.S, T)
        out = flat[self.perm.clamp_min(0)] * self.valid[:, None].to(flat.dtype)
        return out.contiguous()

    def slab(self, rank, world):
        """Contiguous slab of scan rows [y0, y1) of one rank: parallel.slab_rows (the ONE definition bench.py, the tests and
        gather_transient share)."""
        from .parallel import slab_rows
        return slab_rows(self.Sy, self.tile_y, rank, world)
