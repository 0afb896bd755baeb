"""nlosb200_transient_forward_static: the batched forward for callers that may not allocate (no counterpart in the reference).
Its slice counters live in a static device array inside the library instead of a caller buffer.  Called
through the C ABI and compared with nlosb200_transient_forward on the same prepared lists, twice in a row (the array keeps the last call's
contents).  Collected last on purpose."""
import numpy as np
import pytest
import torch

from nlos_gaussian_renderer_b200 import _lib
from nlos_gaussian_renderer_b200.transient import TILE, TransientRenderer
from tests import util

pytestmark = pytest.mark.gpu
TOL = 1e-5          # the two entry points run the same kernels; only the order of the float REDs differs (tests/test_gpu_variants.py uses 2e-5)


def _static_forward(r, H, Ha):
    """One forward pass of the prepared renderer r through the no-alloc entry point, into the caller's H (and Ha for one-pass rows)."""
    H.zero_()
    if Ha is not None:
        Ha.zero_()
    for var, segs in r._launch_groups():
        (cls, op), (cls2, op2) = segs[0], (segs[1] if len(segs) > 1 else (-1, 0))
        args = list(r._render_args(cls))
        del args[11]                                              # `work`: the static entry point takes no scratch from the caller
        _lib.call("nlosb200_transient_forward_static", *args, r.c_deltaT, r.mcut, r.mode, var, op, 1, *r._sh_args(),
                  (Ha if op else H).data_ptr(), cls2, op2, (Ha if op2 else H).data_ptr(), _lib.stream())
    if r.one_pass:
        _lib.call("nlosb200_analytic_filter", Ha.data_ptr(), H.data_ptr(), H.shape[0], r.T, r.t0, r.dt, 0, _lib.stream())


@pytest.mark.parametrize("mode,kw", [("numerical", {}), ("analytic", {"one_pass": True}), ("analytic", {"one_pass": False, "variant": "wide"})])
def test_static_forward_matches_the_existing_entry_point(mode, kw):
    N, Sy, Sx, T, A = 600, 8, 16, 128, 8
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, device="cuda")
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, **kw)
    assert geom.n_tiles * 2 <= _lib.lib().nlosb200_transient_static_work_ints()
    r.prepare(*params)
    r.check_lists()
    want = r.forward_tiled().clone()                              # nlosb200_transient_forward with the renderer's own `work` buffer
    assert float(want.abs().max()) > 0
    H = torch.empty(geom.n_tiles * TILE, T, dtype=torch.float32, device="cuda")
    Ha = torch.empty_like(H) if r.one_pass else None
    r.work.fill_(0x7FFFFFFF)                                      # poison the caller-side counters: the static entry point must not touch them
    torch.cuda.synchronize()
    for call in range(2):                                         # two calls in a row: the second finds the first one's counters in the array
        if call == 1:
            held = torch.cuda.memory_allocated()
            free0 = torch.cuda.mem_get_info()[0]
        _static_forward(r, H, Ha)
        torch.cuda.synchronize()
        if call == 1:
            assert torch.cuda.memory_allocated() == held and torch.cuda.mem_get_info()[0] == free0      # nothing allocated by the call
        assert torch.isfinite(H).all()
        assert util.rel_to_max(H.cpu().numpy(), want.cpu().numpy()) < TOL, f"call {call}"
    assert int(r.work.min()) == 0x7FFFFFFF
    # and the existing entry point still agrees afterwards
    assert util.rel_to_max(r.forward_tiled().cpu().numpy(), want.cpu().numpy()) < TOL


def test_static_forward_rejects_grids_beyond_its_scratch():
    L = _lib.lib()
    cap = L.nlosb200_transient_static_work_ints()
    assert cap == 65536
    d = torch.zeros(64, dtype=torch.float32, device="cuda")
    for n_tiles, cls2 in ((cap + 1, -1), (cap // 2 + 1, 1)):
        with pytest.raises(RuntimeError, match="static scratch"):
            _lib.call("nlosb200_transient_forward_static", d.data_ptr(), n_tiles, 8, d.data_ptr(), 1, d.data_ptr(), d.data_ptr(), 1, d.data_ptr(), 2, 0,
                      0.1, 0.01, 32, 96, 8, 4, 1.0, 5.3, 0, 0, 0, 0, None, 1, 0, d.data_ptr(), cls2, 0, d.data_ptr(), _lib.stream())
    torch.cuda.synchronize()
    assert float(d.abs().max()) == 0                              # nothing was launched
