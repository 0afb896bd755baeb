"""Import the reference's own Python (nlos_helpers.py, gaussian_model/*) in a container that lacks its optional
dependencies -- the two shims of SURVEY.md 0.7:  (1) stub modules for open3d / trimesh / matplotlib / skimage / cv2 (plotting and mesh
export, never on the path we pin), (2) the name gaussian_model.py:4 imports but gaussian_utils.py does not define
(inverse_opacity_activation := inverse_sigmoid, gaussian_utils.py:171).  TEST INFRASTRUCTURE ONLY; nothing here is copied from
the reference -- the modules are imported from where they lie under /root/reference (absent on the GPU box)."""
import importlib
import os
import sys
import types

REF = os.environ.get("NLOS_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF, "nlos_helpers.py"))


def _stub(name):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__path__ = []
    m.__getattr__ = lambda attr: (_ for _ in ()).throw(AttributeError(f"stub module {name} has no {attr}")) if attr.startswith("__") else _Dummy()
    sys.modules[name] = m
    return m


class _Dummy:
    def __call__(self, *a, **k):
        return _Dummy()

    def __getattr__(self, a):
        return _Dummy()


def import_reference(extra_path=()):
    """-> dict(nlos_helpers=module, gaussian_model=module GaussianModel lives in, gaussian_utils=...)."""
    if not available():
        raise RuntimeError("/root/reference is not present")
    for name in ("open3d", "trimesh", "matplotlib", "matplotlib.pyplot", "skimage", "skimage.measure", "cv2"):
        try:
            importlib.import_module(name)
        except Exception:
            _stub(name)
            if "." in name:
                setattr(sys.modules[name.split(".")[0]], name.split(".")[1], sys.modules[name])
    for p in (REF,) + tuple(extra_path):
        if p not in sys.path:
            sys.path.insert(0, p)
    GU = importlib.import_module("gaussian_model.gaussian_utils")
    if not hasattr(GU, "inverse_opacity_activation"):
        GU.inverse_opacity_activation = GU.inverse_sigmoid
    GM = importlib.import_module("gaussian_model.gaussian_model")
    NH = importlib.import_module("nlos_helpers")
    return {"nlos_helpers": NH, "gaussian_model": GM, "gaussian_utils": GU}
