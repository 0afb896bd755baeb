"""distCUDA2 (simple-knn helper) against the brute-force oracle: bit-exact (same squared-distance expression and summation
order; the hash grid only changes the search order).  Parity is unpinned by the reference (absent submodule) -- oracle/nlos_oracle.c."""
import numpy as np
import pytest
import torch

import oracle
from nlos_gaussian_renderer_b200.knn import distCUDA2

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,kind", [(1, "uniform"), (2, "uniform"), (3, "uniform"), (4, "uniform"), (257, "uniform"), (5000, "uniform"),
                                    (4000, "clustered"), (3000, "plane"), (2000, "duplicates"), (1500, "line")])
def test_knn_matches_bruteforce(N, kind):
    g = torch.Generator().manual_seed(N)
    p = torch.rand(N, 3, generator=g)
    if kind == "clustered":                      # very uneven cell occupancy
        p = torch.cat([p[: N // 2] * 0.01 + 0.3, p[N // 2:] * torch.tensor([3.0, 0.2, 1.0])])
    elif kind == "plane":                        # degenerate bounding box (zero extent along y)
        p[:, 1] = 0.25
    elif kind == "line":
        p[:, 1:] = 0.5
    elif kind == "duplicates":                   # coincident points: distance 0 to another point is legal
        p[N // 2:] = p[: N - N // 2]
    p = p.float()
    got = distCUDA2(p.cuda()).cpu().numpy()
    ref = oracle.knn_dist2(p.numpy())
    assert np.array_equal(got, ref), np.abs(got - ref).max()


def test_knn_large_property():
    """1M points (the north-star size): every result is >= 0, finite, and equals the oracle on a random sample of queries
    (brute force over all points for the sampled rows only)."""
    N = 1_000_000
    p = torch.rand(N, 3, generator=torch.Generator().manual_seed(1)) * torch.tensor([0.5, 0.5, 0.5])
    out = distCUDA2(p.cuda())
    assert torch.isfinite(out).all() and (out >= 0).all()
    idx = torch.randint(0, N, (64,), generator=torch.Generator().manual_seed(2))
    pc = p.cuda()
    for i in idx.tolist():
        d = pc - pc[i]
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        d2[i] = float("inf")
        b = torch.topk(d2, 3, largest=False).values.sort().values
        ref = ((b[0] + b[1]) + b[2]) / 3.0
        assert abs(float(out[i]) - float(ref)) <= 1e-6 * float(ref) + 1e-12


def test_simple_knn_import_name():
    from simple_knn._C import distCUDA2 as f            # the reference's import (gaussian_model/gaussian_model.py:8)
    p = torch.rand(100, 3).cuda()
    assert f(p).shape == (100,)
    with pytest.raises(RuntimeError):
        f(p.cpu())
