"""`from simple_knn._C import distCUDA2` (gaussian_model/gaussian_model.py:8) -> libnlosb200.so:nlosb200_knn_dist2."""
try:
    from nlos_gaussian_renderer_b200 import _lib as _l
    _l.lib()                                            # a missing library must surface as ImportError (the reference then sets KNN_FLAG = False)
except (OSError, RuntimeError) as e:                    # pragma: no cover
    raise ImportError(f"libnlosb200.so is not available: {e}")
from nlos_gaussian_renderer_b200.knn import distCUDA2  # noqa: F401,E402
