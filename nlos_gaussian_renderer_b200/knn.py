"""distCUDA2 of the reference's simple-knn dependency (gaussian_model/gaussian_model.py:7-11,202-203): mean squared distance
to the 3 nearest other points, used to initialise the Gaussian scales.  All compute is in libnlosb200.so (csrc/knn.cu);
CUDA tensors only -- there is no CPU fallback."""
import torch

from . import _lib


def distCUDA2(points):
    """points [N,3] float CUDA tensor -> [N] float32 (same device)."""
    if not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor (libnlosb200 has no CPU fallback)")
    pts = points.detach().to(torch.float32).contiguous().reshape(-1, 3)
    N = pts.shape[0]
    out = torch.empty(N, dtype=torch.float32, device=pts.device)
    if N == 0:
        return out
    L = _lib.lib()
    nbytes = int(L.nlosb200_knn_workspace_bytes(N))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=pts.device)
    with torch.cuda.device(pts.device):
        _lib.call("nlosb200_knn_dist2", pts.data_ptr(), N, out.data_ptr(), ws.data_ptr(), nbytes, _lib.stream())
    return out
