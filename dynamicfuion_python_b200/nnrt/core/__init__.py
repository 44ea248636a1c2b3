"""nnrt.core — matmul3d (cpp/pybind/core/core.cpp:31-46) and the linalg submodule."""
import torch

from .. import _tensor as T
from . import linalg


def matmul3d(a, b):
    """Batched matrix product; 3x3 blocks run on the hand-written kernel (Matmul3D_CUDA.cpp used cuBLAS gemmBatched)."""
    a, b = T.cuda(a, torch.float32), T.cuda(b, torch.float32)
    if a.dim() != 3 or b.dim() != 3 or a.shape[0] != b.shape[0] or a.shape[2] != b.shape[1]:
        raise RuntimeError(f"matmul3d: incompatible shapes {tuple(a.shape)} and {tuple(b.shape)}")
    if tuple(a.shape[1:]) == (3, 3) and tuple(b.shape[1:]) == (3, 3):
        c = torch.empty_like(a)
        T.call("nnrt_b200_matmul3d_3x3", T.ptr(a), T.ptr(b), a.shape[0], T.ptr(c), T.stream())
        return c
    raise RuntimeError("matmul3d: only [N,3,3] x [N,3,3] is on the dense-depth path of this build")
