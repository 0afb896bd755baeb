"""Host-side launch planning of TransientRenderer (no GPU needed: the library loads, nothing is launched): which entry classes the binning
produces for a mode / variant and how they are grouped into render launches (one launch per kernel variant, its classes as segments;
one class per launch in the deterministic mode)."""
import pytest

from nlos_gaussian_renderer_b200 import scenes
from nlos_gaussian_renderer_b200.geometry import ScanGeometry
from nlos_gaussian_renderer_b200.transient import TransientRenderer


def _r(**kw):
    geom = ScanGeometry.wall_grid(8, 16, A=8)
    ta = scenes.time_axis(128)
    return TransientRenderer(geom, 128, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], **kw)


def test_launch_groups():
    r = _r(mode="numerical")                                   # footprint classes only
    assert r.n_classes == 2 and r._launches() == ((0, 1, 0), (1, 0, 0))
    assert r._launch_groups() == [(1, [(0, 0)]), (0, [(1, 0)])]
    r = _r(mode="numerical", variant="wide")
    assert r.n_classes == 1 and r._launch_groups() == [(0, [(0, 0)])]
    r = _r(mode="analytic", one_pass=True)                     # footprint x time-wide: two segments per kernel
    assert r.n_classes == 4
    assert r._launch_groups() == [(1, [(0, 0), (2, 1)]), (0, [(1, 0), (3, 1)])]
    r = _r(mode="analytic", one_pass=True, deterministic=True)
    assert [len(s) for _, s in r._launch_groups()] == [1, 1, 1, 1]
    r = _r(mode="analytic", one_pass=True, variant="narrow")
    assert r.n_classes == 2 and r._launch_groups() == [(1, [(0, 0), (1, 1)])]
    r = _r(mode="analytic")                                    # 4 tiles < 64: the one-pass plan is off by default
    assert not r.one_pass and r.n_classes == 2
    assert not _r(mode="numerical", one_pass=True).one_pass    # a flag of the analytic mode only


def test_cpu_tensors_are_refused():
    import torch
    r = _r(mode="numerical")
    with pytest.raises(RuntimeError, match="CUDA"):
        r.prepare(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 4), torch.zeros(4, 1), torch.zeros(4, 1))
