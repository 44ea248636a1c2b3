"""nnrt.core — matmul3d, KDTree (cpp/pybind/core/core.cpp:31-46) and the linalg submodule."""
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


class KDTree:
    """nnrt.core.KDTree (cpp/pybind/core/core.cpp:36-46; cpp/core/KdTree.cpp). The index is the point array itself: queries are a
    brute-force scan with the indexed points broadcast from shared memory, which at warp-field sizes (hundreds to thousands of nodes)
    is faster on a B200 than building and walking a tree, and fixes the slot order the tree leaves to its traversal."""

    def __init__(self, points):
        self.points = T.cuda(points, torch.float32)
        if self.points.dim() != 2 or self.points.shape[1] != 3:
            raise RuntimeError(f"KDTree (B200 build) indexes 3D points, got shape {tuple(self.points.shape)}")

    def find_k_nearest_to_points(self, query_points, k, sort_output=False):
        q = T.cuda(query_points, torch.float32)
        if q.dim() != 2 or q.shape[1] != self.points.shape[1]:  # KdTree.cpp:58-63
            raise RuntimeError(f"Query point array of shape {tuple(q.shape)} is incompatible to the set of points being indexed by the KD Tree "
                               f"(reference points), which has shape {tuple(self.points.shape)}.")
        indices, distances = T.empty((q.shape[0], k), torch.int32), T.empty((q.shape[0], k), torch.float32)
        T.call("nnrt_b200_find_k_nearest_to_points", T.ptr(self.points), self.points.shape[0], T.ptr(q), q.shape[0], int(k), int(bool(sort_output)),
               T.ptr(indices), T.ptr(distances), T.stream())
        return indices, distances
