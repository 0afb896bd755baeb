"""MCMC relocation helper (gaussian_model/gaussian_model.py:471-549).  Parity UNPINNED: the helper the reference calls is
not in its tree; the oracle restates the publication (oracle/relocation.py).  CPU tests pin the restatement to the closed form and to the
properties the construction guarantees; GPU tests compare the kernel with the oracle and run the two host operations."""
import numpy as np
import pytest
import torch

from oracle import relocation as orel


def test_oracle_properties():
    rng = np.random.default_rng(0)
    o = rng.uniform(0.01, 0.95, 200).astype(np.float32)
    s = rng.uniform(0.001, 0.1, (200, 3)).astype(np.float32)
    n = rng.integers(1, 9, 200)
    o_new, s_new = orel.compute_relocation(o, s, n)
    # N copies of the new opacity compose to the old one
    assert np.allclose(1.0 - (1.0 - o_new.astype(np.float64)) ** n, o, rtol=0, atol=3e-6)
    # N = 1 is the identity
    one = n == 1
    assert one.any() and np.allclose(o_new[one], o[one], atol=1e-7) and np.allclose(s_new[one], s[one], rtol=3e-6)
    # the published double sum == its closed form (hockey-stick identity) while the alternating series is well conditioned
    o64, fac = orel.relocation_closed_form(o, n)
    assert np.allclose(o_new, o64, atol=2e-7)
    assert np.allclose(s_new, s * fac[:, None], rtol=2e-5)
    assert (s_new <= s * (1 + 3e-6)).all()                      # copies never grow
    # counts are clamped to [1, 50]
    a = orel.compute_relocation(o[:4], s[:4], [0, 80, -3, 50])
    b = orel.compute_relocation(o[:4], s[:4], [1, 50, 1, 50])
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.gpu
def test_kernel_matches_oracle():
    from nlos_gaussian_renderer_b200.relocation import compute_relocation_cuda
    rng = np.random.default_rng(1)
    P = 3000
    o = rng.uniform(0.005, 0.99, P).astype(np.float32)
    s = np.exp(rng.normal(-4, 1, (P, 3))).astype(np.float32)
    n = np.concatenate([rng.integers(1, 12, P - 200), rng.integers(1, 70, 200)])
    want_o, want_s = orel.compute_relocation(o, s, n)
    got_o, got_s = compute_relocation_cuda(torch.from_numpy(o).cuda(), torch.from_numpy(s).cuda(), torch.from_numpy(n).cuda())
    got_o, got_s = got_o.cpu().numpy(), got_s.cpu().numpy()
    assert np.abs(got_o - want_o).max() <= 2e-7                 # powf: <= 2 ulp of a value < 1
    # where the alternating series is well conditioned the scale factor agrees to fp32 rounding; for large N * opacity both evaluations are
    # dominated by the rounding of the float accumulator in the published summation order, which is identical term by term
    ok = np.isfinite(want_s).all(1) & (n <= 12)
    assert ok.sum() > 2000
    assert np.allclose(got_s[ok], want_s[ok], rtol=1e-4)
    rest = np.isfinite(want_s).all(1) & ~ok
    assert np.allclose(got_s[rest], want_s[rest], rtol=5e-3)
    assert compute_relocation_cuda(torch.zeros(0).cuda(), torch.zeros(0, 3).cuda(), torch.zeros(0, dtype=torch.long).cuda())[0].numel() == 0


@pytest.mark.gpu
def test_relocate_and_add():
    from nlos_gaussian_renderer_b200 import relocation
    from nlos_gaussian_renderer_b200.optim import FusedAdam
    g = torch.Generator(device="cuda").manual_seed(3)
    N = 500
    mk = lambda *s: torch.randn(*s, device="cuda", generator=g)
    params = {"mu": mk(N, 3), "f_dc": mk(N, 1, 1), "f_rest": mk(N, 15, 1), "opacity": mk(N, 1), "scaling": mk(N, 3) - 4, "rotation": mk(N, 4)}
    params["opacity"][:60] = -8.0                                # dead: sigmoid(-8) = 3e-4 <= 0.005 (main.py:245)
    before = {k: v.clone() for k, v in params.items()}
    opt = FusedAdam([{"params": [v], "lr": 1e-3, "name": k} for k, v in params.items()])
    opt.step([torch.ones_like(v) for v in params.values()])
    dead = (torch.sigmoid(params["opacity"]) <= 0.005).squeeze(-1)
    assert int(dead.sum()) == 60
    params = {k: g_["params"][0] for k, g_ in zip(params, opt.param_groups)}
    stepped = {k: v.clone() for k, v in params.items()}
    src = relocation.relocate_gs(params, dead, optimizer=opt, generator=g)
    assert src.numel() == 60 and not dead[src].any()
    d = dead.nonzero(as_tuple=True)[0]
    for k in ("mu", "f_dc", "f_rest", "rotation"):                # copies of their sources
        assert torch.equal(params[k][d], stepped[k][src])
    assert torch.equal(params["opacity"][d], params["opacity"][src]) and torch.equal(params["scaling"][d], params["scaling"][src])
    # a source drawn n times: n + 1 copies compose to its old opacity
    cnt = torch.bincount(src, minlength=N)[src] + 1
    o_old = torch.sigmoid(stepped["opacity"][src, 0]).double()
    o_new = torch.sigmoid(params["opacity"][src, 0]).double()
    comp = 1 - (1 - o_new) ** cnt
    free = o_new > 0.0051                                        # (not clamped at the 0.005 floor)
    assert free.any() and torch.allclose(comp[free], o_old[free], atol=1e-5)
    assert (params["scaling"][src] <= stepped["scaling"][src] + 1e-6).all()
    untouched = torch.ones(N, dtype=torch.bool, device="cuda"); untouched[d] = False; untouched[src] = False
    for k in params:
        assert torch.equal(params[k][untouched], stepped[k][untouched])
    for st in opt.state:                                          # moments zeroed at the sources only
        assert float(st["exp_avg"][src].abs().max()) == 0 and float(st["exp_avg_sq"][src].abs().max()) == 0
        assert float(st["exp_avg"][untouched].abs().min()) > 0
    assert relocation.relocate_gs(params, torch.zeros(N, dtype=torch.bool, device="cuda")).numel() == 0
    grown, num = relocation.add_new_gs(params, cap_max=510, optimizer=opt, generator=g)
    assert num == 10 and all(v.shape[0] == N + 10 for v in grown.values())
    grown2, num2 = relocation.add_new_gs(grown, cap_max=510)
    assert num2 == 0 and grown2 is grown
    del before
