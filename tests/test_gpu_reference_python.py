"""The reference's OWN, UNMODIFIED Python driven over this repo's extension on the GPU (row a10 of SURVEY.md 8):
nlos_helpers.gaussian_transient_rendering -> gaussian_transient_rendering_cuda (nlos_helpers.py:192-278) -> GaussianRendererCUDA.render_transient
(gaussian_model/rendering_cuda.py:208-270) -> CUDARenderModule.forward -> CUDARenderFunction (gaussian_model/cuda_autograd.py:27-316) ->
`nlos_gaussian_renderer._C.render_rays` = libnlosb200.so.  The files are the copies oracle/build_ref.py stages into the git-ignored
oracle/_ref/py/ (the GPU box has no /root/reference), imported with the two shims of SURVEY.md 0.7."""
import os
import subprocess
import sys
import types

import numpy as np
import pytest
import torch

import oracle
import oracle.build_ref as build_ref
from nlos_gaussian_renderer_b200 import scenes
from nlos_gaussian_renderer_b200.geometry import ScanGeometry
from nlos_gaussian_renderer_b200.transient import TransientRenderer
from tests import util

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PY = build_ref.stage_python()
needs_ref = pytest.mark.skipif(PY is None, reason="no staged reference Python (oracle/_ref/py) and no /root/reference")

DRIVER = r"""
import sys, types, json
sys.path.insert(0, %(root)r)
import os
os.environ["NLOS_REFERENCE_ROOT"] = %(py)r
from tests.golden import ref_python_shims as shims
ref = shims.import_reference()
NH, GM = ref["nlos_helpers"], ref["gaussian_model"]
import numpy as np, torch
import gaussian_model.cuda_autograd as ca
assert ca.__file__.startswith(%(py)r) and NH.CUDA_RENDERER is not None
dev = torch.device("cuda")
d = np.load(%(inp)r)
N = d["means"].shape[0]
A, start, end, occ = int(d["A"]), int(d["start"]), int(d["end"]), bool(d["occ"])
args = types.SimpleNamespace(sh_degree=0, num_sampling_points=A, start=start, end=end, occlusion=occ, scaling_modifier=1.0,
                             use_cuda_renderer=True, rendering_type="netf")
model = GM.GaussianModel(args, dev)
P = lambda a: torch.nn.Parameter(torch.from_numpy(a).to(dev))
model._mu, model._scaling, model._rotation, model._opacity = P(d["means"]), P(d["ls"]), P(d["rots"]), P(d["opac"].reshape(N, 1))
model._features_dc = P(d["feats"].reshape(N, 1, 1))
model._features_rest = torch.zeros(N, 0, 1, device=dev)
vp = torch.tensor(d["vp"], device=dev)
dk = {"volume_box_point": torch.from_numpy(NH.volume_box_point(d["vp"], float(d["vs"]))).float().to(dev), "c": float(d["c"]), "deltaT": float(d["dT"]),
      "volume_position": vp}
hists = []
for pos in d["scan"]:
    p = torch.tensor(pos, device=dev)
    cart, I1, I2, num_r, dth, dph, *_ = NH.spherical_sample_histogram(args, dk, p)
    result, hist = NH.gaussian_transient_rendering(args, model, dk, cart, p, I1, I2, num_r, dth, dph)
    assert result.shape == (num_r, A * A) and hist.shape == (num_r,)
    hists.append(hist.detach().cpu().numpy())
np.save(%(out)r, np.stack(hists))
print("driven")
"""


@needs_ref
@pytest.mark.parametrize("occ", [False, True])
def test_reference_dispatch_over_the_shim_matches_oracle_and_batched_renderer(tmp_path, occ):
    N, A, T = 300, 8, 64
    vp, vs = (0.0, 0.75, 0.0), 0.5
    ta = scenes.time_axis(T)
    params = scenes.make_gaussians(N, 77, scale_mul=1.5)
    scan = np.array([[-0.4, 0.0, 0.3], [0.0, 0.0, 0.0], [0.27, 0.0, -0.45]], np.float32)
    inp, out = str(tmp_path / "in.npz"), str(tmp_path / "out.npy")
    np.savez(inp, means=params[0].numpy(), ls=params[1].numpy(), rots=params[2].numpy(), opac=params[3].numpy(), feats=params[4].numpy(),
             A=A, start=ta["start"], end=ta["end"], occ=occ, vp=np.array(vp, np.float32), vs=vs, c=ta["c"], dT=ta["deltaT"], scan=scan)
    r = subprocess.run([sys.executable, "-c", DRIVER % dict(root=ROOT, py=PY, inp=inp, out=out)], capture_output=True, text=True, cwd="/tmp")
    assert r.returncode == 0 and "driven" in r.stdout, r.stderr[-3000:]
    got = np.load(out)                                                             # [3, T]: pred_histogram * vp_y^2 of the reference's own dispatch
    geom = ScanGeometry.from_volume(torch.from_numpy(scan).reshape(1, 3, 3), vp, vs, A, device="cuda")
    cpu = ScanGeometry.from_volume(torch.from_numpy(scan).reshape(1, 3, 3), vp, vs, A)
    # the oracle with the reference's 256-per-ray cap (the one-scan-point operator keeps it)
    H64 = util.oracle_H(cpu, ta, params, mode=0, precision="f64", use_occlusion=occ, cap=256)
    for p in range(3):
        assert np.abs(got[p] - H64[p]).max() <= 1e-4 * H64[p].max(), p
    if not occ:
        Hb = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode="numerical").forward(*[q.cuda() for q in params])
        Hb = Hb.cpu().numpy().reshape(3, T)
        Hu = util.oracle_H(cpu, ta, params, mode=0, precision="f64", use_occlusion=False, cap=0)
        cap_binds = np.abs(Hu - H64).max() > 1e-6 * H64.max()
        if not cap_binds:                                                          # (the batched renderer has no 256 cap)
            assert np.abs(got - Hb).max() <= 1e-4 * Hb.max()


@needs_ref
def test_reference_test_scripts_run_over_the_shim():
    """test_cuda_renderer.py / test_cuda_autograd.py: print-only scripts; what they pin is the import surface.  Both run to the
    end; the forward test of test_cuda_autograd.py fails inside the reference (GaussianModel(sh_degree=0) TypeError, test_cuda_autograd.py:64
    vs gaussian_model.py:33) exactly as it does against the reference's own extension."""
    pre = ("import sys, os, runpy; sys.path.insert(0, %r); os.environ['NLOS_REFERENCE_ROOT'] = %r; "
           "from tests.golden import ref_python_shims as s; s.import_reference(); "
           "import configs.default; runpy.run_path(os.path.join(%r, %%r), run_name='__main__')") % (ROOT, PY, PY)
    for script, must in (("test_cuda_autograd.py", ["CUDA extension imported successfully", "CUDARenderModule created successfully"]),
                         ("test_cuda_renderer.py", ["CUDA renderer imported successfully", "Renderer instance created"])):
        r = subprocess.run([sys.executable, "-c", pre % script], capture_output=True, text=True, cwd=PY)
        assert r.returncode == 0, r.stderr[-2000:]
        for m in must:
            assert m in r.stdout, (script, r.stdout[-1500:])
