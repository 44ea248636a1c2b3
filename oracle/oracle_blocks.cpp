// ORACLE — TEST INFRASTRUCTURE ONLY. Generic block-sparse building blocks of the reference's arrowhead solve, restated on flat
// arrays with the reference's loop structure. Pinned by the inline known-answer tensors of
// cpp/tests/test_linalg_matmul_block_sparse.cpp and cpp/tests/test_linalg_block_routines.cpp (tests/golden/fitter_kats.npz).
#include "oracle_common.h"

using namespace oracle;

namespace {
// row-major s x s product C = op(A) op(B), plain fp32 accumulation in k order (CBLAS reference order for these sizes)
void block_matmul(const float* A, bool ta, const float* B, bool tb, int s, float* C) {
	for (int r = 0; r < s; r++)
		for (int c = 0; c < s; c++) {
			float acc = 0.f;
			for (int k = 0; k < s; k++) acc += (ta ? A[k * s + r] : A[r * s + k]) * (tb ? B[c * s + k] : B[k * s + c]);
			C[r * s + c] = acc;
		}
}
}  // namespace

// MatmulBlockSparseRowWise{,Padded} (cpp/core/linalg/MatmulBlockSparseImpl.h:48-150): C_k = A[row(k)] * B_k for every block k of B;
// rows >= a_count give zero blocks (padded form). The unpadded form keeps the blocks with row < a_count, in order.
ORACLE_API void oracle_matmul_block_sparse_row_wise_padded(const float* a_blocks, int a_count, const float* b_blocks,
                                                           const int32_t* b_coordinates, int b_count, int s, float* c_blocks) {
	for (int k = 0; k < b_count; k++) {
		const int row = b_coordinates[k * 2];
		float* C = c_blocks + (int64_t) k * s * s;
		if (row >= a_count) { std::fill(C, C + s * s, 0.f); continue; }
		block_matmul(a_blocks + (int64_t) row * s * s, false, b_blocks + (int64_t) k * s * s, false, s, C);
	}
}

// MatmulBlockSparse (MatmulBlockSparseImpl.h:171-420): C = op(A) op(B) with both factors given as block lists + int16 breadboards
// (block index at every block position, -1 = empty). Output: one s x s block per output position (row-major over
// out_rows x out_cols) and a mask of the positions that received at least one product; the reference returns the masked blocks and
// their coordinates in that row-major order. Sums run over the inner block index in ascending order.
ORACLE_API void oracle_matmul_block_sparse(const float* a_blocks, const int16_t* a_board, int a_rows, int a_cols, int transpose_a,
                                           const float* b_blocks, const int16_t* b_board, int b_rows, int b_cols, int transpose_b,
                                           int s, float* out_blocks, uint8_t* out_mask) {
	const int out_rows = transpose_a ? a_cols : a_rows, inner = transpose_a ? a_rows : a_cols, out_cols = transpose_b ? b_rows : b_cols;
	std::vector<float> P((size_t) s * s);
	for (int i = 0; i < out_rows; i++)
		for (int j = 0; j < out_cols; j++) {
			float* C = out_blocks + ((int64_t) i * out_cols + j) * s * s;
			std::fill(C, C + s * s, 0.f);
			uint8_t any = 0;
			for (int k = 0; k < inner; k++) {
				const int16_t ia = transpose_a ? a_board[k * a_cols + i] : a_board[i * a_cols + k];
				const int16_t ib = transpose_b ? b_board[j * b_cols + k] : b_board[k * b_cols + j];
				if (ia < 0 || ib < 0) continue;
				block_matmul(a_blocks + (int64_t) ia * s * s, transpose_a != 0, b_blocks + (int64_t) ib * s * s, transpose_b != 0, s, P.data());
				for (int q = 0; q < s * s; q++) C[q] += P[q];
				any = 1;
			}
			out_mask[i * out_cols + j] = any;
		}
}

// BlockSparseAndVectorProduct (MatmulBlockSparseImpl.h:441-600): c = op(A) b, A given by blocks + coordinates (+ offset), m = len(c)
ORACLE_API void oracle_block_sparse_and_vector_product(const float* blocks, const int32_t* coordinates, int count, int s, int m, int transpose,
                                                       int row_offset, int column_offset, const float* b, float* c) {
	std::fill(c, c + m, 0.f);
	for (int k = 0; k < count; k++) {
		const int bi = transpose ? coordinates[k * 2 + 1] + column_offset : coordinates[k * 2] + row_offset;
		const int bj = transpose ? coordinates[k * 2] + row_offset : coordinates[k * 2 + 1] + column_offset;
		const float* A = blocks + (int64_t) k * s * s;
		for (int r = 0; r < s; r++) {
			float acc = 0.f;
			for (int q = 0; q < s; q++) acc += (transpose ? A[q * s + r] : A[r * s + q]) * b[bj * s + q];
			c[bi * s + r] += acc;
		}
	}
}

// DiagonalBlockSparseAndVectorProduct (MatmulBlockSparseImpl.h:620-700): c_i = D_i b_i
ORACLE_API void oracle_diagonal_block_sparse_and_vector_product(const float* blocks, int count, int s, const float* b, float* c) {
	for (int k = 0; k < count; k++)
		for (int r = 0; r < s; r++) {
			float acc = 0.f;
			for (int q = 0; q < s; q++) acc += blocks[((int64_t) k * s + r) * s + q] * b[k * s + q];
			c[k * s + r] = acc;
		}
}

// InvertPositiveSemidefiniteBlocks (cpp/core/linalg/InvertBlocksCPU.cpp:80-112): potrf + potrs against the identity, per block.
ORACLE_API int oracle_invert_positive_semidefinite_blocks(const float* blocks, int count, int s, float* inverted) {
	int failures = 0;
	std::vector<float> L((size_t) s * s), x((size_t) s);
	for (int b = 0; b < count; b++) {
		const float* A = blocks + (int64_t) b * s * s;
		std::copy(A, A + s * s, L.begin());
		bool ok = true;
		for (int j = 0; j < s && ok; j++) {
			float d = L[j * s + j];
			for (int k = 0; k < j; k++) d -= L[j * s + k] * L[j * s + k];
			if (!(d > 0.f)) { ok = false; break; }
			d = sqrtf(d);
			L[j * s + j] = d;
			for (int i = j + 1; i < s; i++) {
				float v = L[i * s + j];
				for (int k = 0; k < j; k++) v -= L[i * s + k] * L[j * s + k];
				L[i * s + j] = v / d;
			}
		}
		if (!ok) { failures++; continue; }
		for (int c = 0; c < s; c++) {
			for (int i = 0; i < s; i++) {
				float v = i == c ? 1.f : 0.f;
				for (int k = 0; k < i; k++) v -= L[i * s + k] * x[k];
				x[i] = v / L[i * s + i];
			}
			for (int i = s - 1; i >= 0; i--) {
				float v = x[i];
				for (int k = i + 1; k < s; k++) v -= L[k * s + i] * x[k];
				x[i] = v / L[i * s + i];
			}
			for (int i = 0; i < s; i++) inverted[((int64_t) b * s + i) * s + c] = x[i];
		}
	}
	return failures;
}

// InvertTriangularBlocks (cpp/core/linalg/InvertBlocks.cpp:33-80, trtri): inverse of a lower / upper triangular block, the other
// triangle of the output is zero.
ORACLE_API void oracle_invert_triangular_blocks(const float* blocks, int count, int s, int upper, float* inverted) {
	for (int b = 0; b < count; b++) {
		const float* A = blocks + (int64_t) b * s * s;
		float* X = inverted + (int64_t) b * s * s;
		std::fill(X, X + s * s, 0.f);
		for (int c = 0; c < s; c++) {
			if (!upper) {
				for (int i = c; i < s; i++) {
					float v = i == c ? 1.f : 0.f;
					for (int k = c; k < i; k++) v -= A[i * s + k] * X[k * s + c];
					X[i * s + c] = v / A[i * s + i];
				}
			} else {
				for (int i = c; i >= 0; i--) {
					float v = i == c ? 1.f : 0.f;
					for (int k = i + 1; k <= c; k++) v -= A[i * s + k] * X[k * s + c];
					X[i * s + c] = v / A[i * s + i];
				}
			}
		}
	}
}

// TransposeBlocksInPlace (cpp/core/linalg/TransposeBlocks.cpp)
ORACLE_API void oracle_transpose_blocks_in_place(float* blocks, int count, int s) {
	for (int b = 0; b < count; b++)
		for (int r = 0; r < s; r++)
			for (int c = r + 1; c < s; c++) std::swap(blocks[((int64_t) b * s + r) * s + c], blocks[((int64_t) b * s + c) * s + r]);
}

// FillInDiagonalBlocks / GetDiagonalBlocks (cpp/core/linalg/DiagonalBlocksImpl.h)
ORACLE_API void oracle_fill_in_diagonal_blocks(float* matrix, int n, const float* blocks, int count, int s) {
	for (int b = 0; b < count; b++)
		for (int r = 0; r < s; r++)
			for (int c = 0; c < s; c++) matrix[(int64_t) (b * s + r) * n + b * s + c] = blocks[((int64_t) b * s + r) * s + c];
}
ORACLE_API void oracle_get_diagonal_blocks(const float* matrix, int n, int s, float* blocks) {
	for (int b = 0; b < n / s; b++)
		for (int r = 0; r < s; r++)
			for (int c = 0; c < s; c++) blocks[((int64_t) b * s + r) * s + c] = matrix[(int64_t) (b * s + r) * n + b * s + c];
}

// FillInSparseBlocks / GetSparseBlocks (cpp/core/linalg/SparseBlocksImpl.h:30-140): with `transpose` the block lands at (j, i)
// transposed.
ORACLE_API void oracle_fill_in_sparse_blocks(float* matrix, int rows, int cols, const float* blocks, const int32_t* coordinates, int count,
                                             int s, int row_offset, int column_offset, int transpose) {
	(void) rows;
	for (int b = 0; b < count; b++)
		for (int q = 0; q < s * s; q++) {
			const int64_t bi = (transpose ? coordinates[b * 2 + 1] : coordinates[b * 2]) + row_offset;
			const int64_t bj = (transpose ? coordinates[b * 2] : coordinates[b * 2 + 1]) + column_offset;
			const int64_t r = bi * s + (transpose ? q % s : q / s), c = bj * s + (transpose ? q / s : q % s);
			matrix[r * cols + c] = blocks[(int64_t) b * s * s + q];
		}
}
ORACLE_API void oracle_get_sparse_blocks(const float* matrix, int cols, int s, const int32_t* coordinates, int count, float* blocks) {
	for (int b = 0; b < count; b++)
		for (int q = 0; q < s * s; q++)
			blocks[(int64_t) b * s * s + q] = matrix[((int64_t) coordinates[b * 2] * s + q / s) * cols + (int64_t) coordinates[b * 2 + 1] * s + q % s];
}
