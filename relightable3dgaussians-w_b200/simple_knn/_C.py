"""`simple_knn._C` drop-in: distCUDA2(points[P,3]) -> mean squared distance to the 3 nearest neighbours [P].

Reference: submodules/simple-knn/spatial.cu:16-26 (torch glue) and simple_knn.cu:185-221 (SimpleKNN::knn);
used by scene/gaussian_model.py:189,249.  Backed by r3dgw_knn_dist2 in libr3dgw_b200.so (csrc/knn.cu).
"""
import torch

from r3dgw_b200 import native as _nat


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    if points.device.type != "cuda":
        raise RuntimeError("distCUDA2 (B200-native): points must be a CUDA tensor; there is no CPU path")
    P = int(points.size(0))
    pts = points.contiguous()
    if pts.dtype != torch.float32:
        pts = pts.float()
    means = torch.zeros(P, dtype=torch.float32, device=points.device)
    if P == 0:
        return means
    L = _nat.lib()
    ws = torch.empty(L.r3dgw_knn_workspace_bytes(P), dtype=torch.uint8, device=points.device)
    with _nat.on_device(points.device):
        rc = L.r3dgw_knn_dist2(P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(),
                               _nat.raw_stream())
    _nat.check(rc, "r3dgw_knn_dist2")
    return means
