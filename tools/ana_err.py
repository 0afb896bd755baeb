"""Error report of the analytic-exact mode (forward vs the fp64 C oracle, gradients vs fp64 autograd) over Gaussian widths.
Run on a GPU box:  python tools/ana_err.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nlos_gaussian_renderer_b200.transient import TransientRenderer, render_transient_batch
from tests import ref_torch, util
from tests.test_gpu_transient import scenes_cpu

for sm in (0.5, 0.7, 1.0, 1.4, 2.0, 3.0, 5.0):
    N, Sy, Sx, T, A = 120, 5, 9, 40, 5
    geom, ta, params = util.small_scene(N, Sy, Sx, T, A, scale_mul=sm, device="cuda")
    r = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode="analytic")
    gH = torch.rand(Sy, Sx, T, generator=torch.Generator().manual_seed(5))
    ps = [p.clone().requires_grad_() for p in params]
    H = render_transient_batch(r, *ps)
    (H * gH.cuda()).sum().backward()
    cg = scenes_cpu(geom)
    masks = torch.from_numpy(util.ray_masks(cg, [p.cpu() for p in params]))
    pd = [p.detach().cpu().double().requires_grad_() for p in params]
    dd = lambda x: x.cpu().double()
    Ht = ref_torch.transient_batch(dd(geom.scan_o), dd(geom.sin_th), dd(geom.cos_th), dd(geom.sin_ph), dd(geom.cos_ph),
                                   dd(ta["t_samples"]), dd(geom.hscale), masks, *pd, ta["c"], ta["deltaT"], mode=1)
    (Ht * gH.reshape(Sy * Sx, T).double()).sum().backward()
    ef = util.rel_to_max(H.detach().cpu().numpy().reshape(-1, T), Ht.detach().numpy())
    eg = [util.rel_to_max(a.grad.cpu().numpy(), b.grad.numpy()) for a, b in zip(ps, pd)]
    s0 = 0.5 * (0.125 / N) ** (1 / 3) * sm / ta["dt"]
    print(f"scale_mul {sm}: widths {0.5*s0:.2f}..{2*s0:.2f} bins  fwd {ef:.2e}  grads " + " ".join(f"{e:.2e}" for e in eg), flush=True)
