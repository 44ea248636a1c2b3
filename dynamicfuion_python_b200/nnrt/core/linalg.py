"""nnrt.core.linalg — block-sparse arrowhead solver and friends (cpp/core/linalg/*). AxisAngleVectorsToMatricesRodrigues is bound by the
reference (cpp/pybind/core/linalg/linalg.cpp:30); the others are C++-only there and are added under their natural names."""
import torch

from .. import _tensor as T


def AxisAngleVectorsToMatricesRodrigues(vectors):
    v = T.cuda(vectors, torch.float32)
    if v.dim() != 2 or v.shape[1] != 3:
        raise RuntimeError(f"vectors must have shape [N, 3], got {tuple(v.shape)}")
    out = T.empty((v.shape[0], 3, 3), torch.float32)
    T.call("nnrt_b200_rodrigues", T.ptr(v), v.shape[0], T.ptr(out), T.stream())
    return out


axis_angle_vectors_to_matrices_rodrigues = AxisAngleVectorsToMatricesRodrigues


class BlockSparseArrowheadMatrix:
    """cpp/core/linalg/BlockSparseArrowheadMatrix.h:46-107 (public data members kept; the int16 breadboard is not needed: wing rows are
    grouped through a CSR built from upper_wing_block_coordinates)."""

    def __init__(self):
        self.upper_wing_blocks = None
        self.upper_wing_block_coordinates = None
        self.corner_upper_blocks = None
        self.corner_upper_block_coordinates = None
        self.arrow_base_block_index = 0
        self.diagonal_block_count = 0
        self._diagonal_blocks = None

    def set_diagonal_blocks(self, blocks):
        self._diagonal_blocks = T.cuda(blocks, torch.float32)

    def diagonal_blocks(self):
        return self._diagonal_blocks

    def stem_diagonal_blocks(self):
        return self._diagonal_blocks[:self.arrow_base_block_index]

    def corner_diagonal_blocks(self):
        return self._diagonal_blocks[self.arrow_base_block_index:]

    def get_block_size(self):
        return 6

    def _args(self):
        d = self._diagonal_blocks
        if d is None or d.shape[0] != self.diagonal_block_count:
            raise RuntimeError("diagonal blocks must be set after diagonal_block_count (SetDiagonalBlocks)")
        w = None if self.upper_wing_blocks is None else T.cuda(self.upper_wing_blocks, torch.float32)
        wc = None if self.upper_wing_block_coordinates is None else T.cuda(self.upper_wing_block_coordinates, torch.int32)
        cu = None if self.corner_upper_blocks is None else T.cuda(self.corner_upper_blocks, torch.float32)
        cuc = None if self.corner_upper_block_coordinates is None else T.cuda(self.corner_upper_block_coordinates, torch.int32)
        keep = (d, w, wc, cu, cuc)
        return keep, (T.ptr(d), int(self.diagonal_block_count), int(self.arrow_base_block_index), T.ptr(w), T.ptr(wc),
                      0 if w is None else w.shape[0], T.ptr(cu), T.ptr(cuc), 0 if cu is None else cu.shape[0])


def solve_block_sparse_arrowhead_cholesky(a, b):
    """SolveBlockSparseArrowheadCholesky(x, A, b) (cpp/core/linalg/SolveBlockSparseArrowheadCholesky.cpp:31-94); returns x."""
    b = T.cuda(b, torch.float32)
    if b.dim() != 1 or b.shape[0] != a.diagonal_block_count * 6:
        raise RuntimeError(f"Size of lhs tensor b must be {{{a.diagonal_block_count * 6}}}, matching the leading dimension of rhs matrix a, "
                           f"but is {tuple(b.shape)} instead.")
    keep, args = a._args()
    x = torch.empty_like(b)
    T.call("nnrt_b200_solve_block_sparse_arrowhead_cholesky", *args, T.ptr(b), T.ptr(x), T.stream())
    return x


def compute_schur_complement_of_arrow_stem_dense(a):
    """ComputeSchurComplementOfArrowStem_Dense (cpp/core/linalg/SchurComplement.cpp:32-77)."""
    keep, args = a._args()
    n = (a.diagonal_block_count - a.arrow_base_block_index) * 6
    out = T.empty((n, n), torch.float32)
    T.call("nnrt_b200_arrowhead_schur_complement", *args, T.ptr(out), T.stream())
    return out


def solve_block_diagonal_cholesky(a_blocks, b):
    """SolveBlockDiagonalCholesky(X, A_blocks, B) (cpp/core/linalg/SolveBlockDiagonalGeneric.h:27-90); single right-hand side."""
    a_blocks, b = T.cuda(a_blocks, torch.float32), T.cuda(b, torch.float32)
    n, s = a_blocks.shape[0], a_blocks.shape[1]
    if a_blocks.dim() != 3 or a_blocks.shape[2] != s or b.numel() != n * s:
        raise RuntimeError("A_blocks must be [N,S,S] and B must have N*S entries")
    x = torch.empty_like(b)
    T.call("nnrt_b200_solve_block_diagonal_cholesky", T.ptr(a_blocks), T.ptr(b), n, s, T.ptr(x), T.stream())
    return x


def precondition_diagonal_blocks(blocks, dampening_factor):
    """PreconditionDiagonalBlocks (cpp/core/linalg/PreconditionDiagonalBlocksImpl.h:29-51): in place, returns the tensor."""
    if not (isinstance(blocks, torch.Tensor) and blocks.is_cuda and blocks.dtype == torch.float32 and blocks.is_contiguous()):
        raise RuntimeError("blocks must be a contiguous float32 CUDA tensor (modified in place)")
    T.call("nnrt_b200_precondition_diagonal_blocks", T.ptr(blocks), blocks.shape[0], blocks.shape[1], float(dampening_factor), T.stream())
    return blocks
