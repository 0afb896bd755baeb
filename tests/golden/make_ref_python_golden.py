"""Golden vectors from the reference's OWN Python, imported from /root/reference in this container (it cannot travel to the GPU box):
  * the per-scan-point lattice of nlos_helpers.spherical_sample_histogram:124-188 (theta / phi ranges, linspaces, dtheta, radii);
  * the dense PyTorch path nlos_helpers.gaussian_transient_rendering:206-232 -> GaussianModel.estimate_rho_w_no_occlusion
    (gaussian_model.py:253-294,346-364): pred_histogram and its gradients for a small scene -- the pin of oracle/dense_torch.py, the CPU
    baseline bench.py times.
    python tests/golden/make_ref_python_golden.py   ->  tests/golden/ref_python.npz"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_python_shims as shims  # noqa: E402


def cases():
    # (scan point, volume position, volume size, A, start, end, c, deltaT)
    return [((0.0, 0.0, 0.0), (0.0, 0.75, 0.0), 0.5, 8, 28, 60, 1.0, 0.0125),
            ((-0.5, 0.0, 0.5), (0.0, 0.75, 0.0), 0.5, 32, 276, 1300, 1.0, 1.3 / 1024),
            ((0.31, 0.0, -0.17), (0.1, 0.6, -0.05), 0.4, 5, 10, 31, 1.0, 0.03),
            ((0.26, 0.0, 0.24), (0.0, 0.75, 0.0), 0.5, 16, 69, 325, 1.0, 1.3 / 256)]


def make():
    ref = shims.import_reference()
    NH, GM = ref["nlos_helpers"], ref["gaussian_model"]
    out = {}
    for n, (pos, vp, vs, A, start, end, c, dT) in enumerate(cases()):
        args = types.SimpleNamespace(num_sampling_points=A, start=start, end=end)
        dk = {"volume_box_point": torch.from_numpy(NH.volume_box_point(np.array(vp), vs)).float(), "c": c, "deltaT": dT}
        p = torch.tensor(pos, dtype=torch.float32)
        cart, I1, I2, num_r, dth, dph, tmin, tmax, pmin, pmax = NH.spherical_sample_histogram(args, dk, p)
        g = cart.reshape(num_r, A, A, 5)
        out[f"lat{n}_scalars"] = np.array([I1, I2, num_r, dth, dph, tmin, tmax, pmin, pmax], np.float64)
        out[f"lat{n}_theta"] = g[0, :, 0, 3].numpy()
        out[f"lat{n}_phi"] = g[0, 0, :, 4].numpy()
        out[f"lat{n}_xyz_first_last"] = torch.stack([g[0, :, :, :3], g[-1, :, :, :3]]).numpy()
        out[f"lat{n}_r"] = torch.linalg.norm(g[:, A // 2, A // 2, :3].double() - p.double(), dim=-1).numpy()
    # ---- dense path, small scene (fp32, as the reference runs it)
    torch.manual_seed(5)
    N, A, start, end, c, dT = 40, 6, 20, 52, 1.0, 0.02
    vp, vs = (0.0, 0.75, 0.0), 0.5
    margs = types.SimpleNamespace(sh_degree=0, num_sampling_points=A, start=start, end=end, occlusion=False, scaling_modifier=1.0,
                                  use_cuda_renderer=False, rendering_type="netf")
    model = GM.GaussianModel(margs, torch.device("cpu"))
    means = torch.tensor(vp) + (torch.rand(N, 3) - 0.5) * 0.4
    # (dense semantics: scale = exp(exp(_scaling)): log-scales around log(log(1.03)) give ~3 cm Gaussians)
    ls = torch.log(torch.log(1.0 + 0.02 + 0.04 * torch.rand(N, 3)))
    rots = torch.randn(N, 4)
    opac = torch.randn(N, 1)
    fdc = ((0.2 * torch.rand(N, 1) - 0.5) / 0.28209479177387814).reshape(N, 1, 1)
    model._mu = means.clone().requires_grad_()
    model._scaling = ls.clone().requires_grad_()
    model._rotation = rots.clone().requires_grad_()
    model._opacity = opac.clone().requires_grad_()
    model._features_dc = fdc.clone().requires_grad_()
    model._features_rest = torch.zeros(N, 0, 1)
    dk = {"volume_box_point": torch.from_numpy(NH.volume_box_point(np.array(vp), vs)).float(), "c": c, "deltaT": dT,
          "volume_position": torch.tensor(vp)}
    p = torch.tensor([0.2, 0.0, -0.1])
    cart, I1, I2, num_r, dth, dph, *_ = NH.spherical_sample_histogram(margs, dk, p)
    result, hist = NH.gaussian_transient_rendering(margs, model, dk, cart, p, I1, I2, num_r, dth, dph)
    w = torch.linspace(0.5, 1.5, num_r)
    (hist * w).sum().backward()
    out["dense_params"] = np.concatenate([means.numpy().ravel(), ls.numpy().ravel(), rots.numpy().ravel(), opac.numpy().ravel(), fdc.numpy().ravel()])
    out["dense_shape"] = np.array([N, A, start, end], np.int64)
    out["dense_scalars"] = np.array([c, dT, vs, *vp, *p.tolist()], np.float64)
    out["dense_hist"] = hist.detach().numpy()
    out["dense_result_sum"] = result.detach().sum(1).numpy()
    out["dense_grad_mu"] = model._mu.grad.numpy()
    out["dense_grad_scaling"] = model._scaling.grad.numpy()
    out["dense_grad_rotation"] = model._rotation.grad.numpy()
    out["dense_grad_opacity"] = model._opacity.grad.numpy()
    out["dense_grad_fdc"] = model._features_dc.grad.numpy().reshape(N, 1)
    return out


if __name__ == "__main__":
    o = make()
    np.savez_compressed(os.path.join(HERE, "ref_python.npz"), **o)
    print("wrote", os.path.join(HERE, "ref_python.npz"), {k: v.shape for k, v in o.items()})
