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


# --- generic block-sparse routines (C++-only in the reference: cpp/core/linalg/*.h); names follow the reference functions ------------
class MatrixPreprocessingOperation:
    NONE = 0
    TRANSPOSE = 1


class UpLoTriangular:
    LOWER = 0
    UPPER = 1


def _blocks(x):
    b = T.cuda(x, torch.float32)
    if b.dim() != 3 or b.shape[1] != b.shape[2]:
        raise RuntimeError(f"blocks must have shape [count, size, size], got {tuple(b.shape)}")
    return b


def matmul_block_sparse_row_wise_padded(blocks_a, blocks_b, blocks_b_coordinates):
    """MatmulBlockSparseRowWisePadded (cpp/core/linalg/MatmulBlockSparse.h): one output block per block of B (zeros where A has no row)."""
    a, b, bc = _blocks(blocks_a), _blocks(blocks_b), T.cuda(blocks_b_coordinates, torch.int32)
    c = torch.empty_like(b)
    T.call("nnrt_b200_matmul_block_sparse_row_wise_padded", T.ptr(a), a.shape[0], T.ptr(b), T.ptr(bc), b.shape[0], a.shape[1], T.ptr(c), T.stream())
    return c


def matmul_block_sparse_row_wise(blocks_a, blocks_b, blocks_b_coordinates):
    """MatmulBlockSparseRowWise: (product blocks, their coordinates) for the rows of B that have an A block."""
    bc = T.cuda(blocks_b_coordinates, torch.int32)
    c = matmul_block_sparse_row_wise_padded(blocks_a, blocks_b, bc)
    keep = bc[:, 0] < T.to_torch(blocks_a).shape[0]
    return c[keep].contiguous(), bc[keep].contiguous()


def matmul_block_sparse(blocks_a, blocks_a_breadboard, matrix_a_preprocessing, blocks_b, blocks_b_breadboard, matrix_b_preprocessing):
    """MatmulBlockSparse: (product blocks, coordinates [K,2] int32) in row-major order of the output positions that received a product."""
    a, b = _blocks(blocks_a), _blocks(blocks_b)
    ab, bb = T.cuda(blocks_a_breadboard, torch.int16), T.cuda(blocks_b_breadboard, torch.int16)
    ta, tb = int(matrix_a_preprocessing), int(matrix_b_preprocessing)
    out_rows = ab.shape[1] if ta else ab.shape[0]
    out_cols = bb.shape[0] if tb else bb.shape[1]
    s = a.shape[1]
    blocks, mask = T.empty((out_rows * out_cols, s, s), torch.float32), T.empty((out_rows * out_cols,), torch.uint8)
    T.call("nnrt_b200_matmul_block_sparse", T.ptr(a), T.ptr(ab), ab.shape[0], ab.shape[1], ta, T.ptr(b), T.ptr(bb), bb.shape[0], bb.shape[1], tb, s,
           T.ptr(blocks), T.ptr(mask), T.stream())
    keep = mask.to(torch.bool)
    index = torch.arange(out_rows * out_cols, device=blocks.device)
    coordinates = torch.stack([index // out_cols, index % out_cols], 1).to(torch.int32)
    return blocks[keep].contiguous(), coordinates[keep].contiguous()


def block_sparse_and_vector_product(blocks_a, matrix_a_other_dimension, blocks_a_coordinates, block_coordinate_offset, matrix_a_preprocessing, vector_b):
    """BlockSparseAndVectorProduct: op(A) b with A given as blocks + block coordinates (+ offset)."""
    a, co, b = _blocks(blocks_a), T.cuda(blocks_a_coordinates, torch.int32), T.cuda(vector_b, torch.float32)
    offset = tuple(block_coordinate_offset) if block_coordinate_offset else (0, 0)
    c = T.empty((int(matrix_a_other_dimension),) + tuple(b.shape[1:]), torch.float32)
    T.call("nnrt_b200_block_sparse_and_vector_product", T.ptr(a), T.ptr(co), a.shape[0], a.shape[1], int(matrix_a_other_dimension), int(matrix_a_preprocessing),
           int(offset[0]), int(offset[1]), T.ptr(b), T.ptr(c), T.stream())
    return c


def diagonal_block_sparse_and_vector_product(blocks_d, vector_b):
    d, b = _blocks(blocks_d), T.cuda(vector_b, torch.float32)
    c = torch.empty_like(b)
    T.call("nnrt_b200_diagonal_block_sparse_and_vector_product", T.ptr(d), d.shape[0], d.shape[1], T.ptr(b), T.ptr(c), T.stream())
    return c


def invert_positive_semidefinite_blocks(blocks):
    """InvertPositiveSemidefiniteBlocks (cpp/core/linalg/InvertBlocks.cpp:82-127); raises if a block is not positive definite."""
    b = _blocks(blocks)
    out, failures = torch.empty_like(b), T.empty((1,), torch.int32)
    T.call("nnrt_b200_invert_positive_semidefinite_blocks", T.ptr(b), b.shape[0], b.shape[1], T.ptr(out), T.ptr(failures), T.stream())
    if int(failures.item()) != 0:
        raise RuntimeError(f"{int(failures.item())} block(s) are not positive definite")
    return out


def invert_triangular_blocks(blocks, uplo):
    b = _blocks(blocks)
    out = torch.empty_like(b)
    T.call("nnrt_b200_invert_triangular_blocks", T.ptr(b), b.shape[0], b.shape[1], int(uplo), T.ptr(out), T.stream())
    return out


def transpose_blocks_in_place(blocks):
    if not (isinstance(blocks, torch.Tensor) and blocks.is_cuda and blocks.dtype == torch.float32 and blocks.is_contiguous()):
        raise RuntimeError("blocks must be a contiguous float32 CUDA tensor (modified in place)")
    T.call("nnrt_b200_transpose_blocks_in_place", T.ptr(blocks), blocks.shape[0], blocks.shape[1], T.stream())
    return blocks


def _matrix_blocks(matrix, blocks, coordinates, offset, transpose, gather):
    if not (isinstance(matrix, torch.Tensor) and matrix.is_cuda and matrix.dtype == torch.float32 and matrix.is_contiguous() and matrix.dim() == 2):
        raise RuntimeError("matrix must be a contiguous two-dimensional float32 CUDA tensor")
    co = None if coordinates is None else T.cuda(coordinates, torch.int32)
    T.call("nnrt_b200_dense_matrix_blocks", T.ptr(matrix), matrix.shape[0], matrix.shape[1], T.ptr(blocks), T.ptr(co), blocks.shape[0], blocks.shape[1],
           int(offset[0]), int(offset[1]), int(bool(transpose)), int(gather), T.stream())


def fill_in_diagonal_blocks(matrix, blocks):
    _matrix_blocks(matrix, _blocks(blocks), None, (0, 0), False, 0)
    return matrix


def get_diagonal_blocks(matrix, block_size):
    out = T.empty((matrix.shape[0] // block_size, block_size, block_size), torch.float32)
    _matrix_blocks(matrix, out, None, (0, 0), False, 1)
    return out


def fill_in_sparse_blocks(matrix, blocks, coordinates, block_coordinate_offset=(0, 0), transpose=False):
    _matrix_blocks(matrix, _blocks(blocks), coordinates, block_coordinate_offset, transpose, 0)
    return matrix


def get_sparse_blocks(matrix, block_size, coordinates):
    co = T.cuda(coordinates, torch.int32)
    out = T.empty((co.shape[0], block_size, block_size), torch.float32)
    _matrix_blocks(matrix, out, co, (0, 0), False, 1)
    return out
