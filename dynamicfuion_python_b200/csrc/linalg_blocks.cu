// Generic block-sparse building blocks of the reference's arrowhead solve as stand-alone stage entry points (any block size up to 8).
// The fitter never calls them: its solve is fused (system_stem_kernel + corner kernels). They mirror the reference functions one to
// one so that the reference's own known-answer tests (cpp/tests/test_linalg_matmul_block_sparse.cpp,
// cpp/tests/test_linalg_block_routines.cpp) run against the CUDA library.
#include "host_util.h"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {

constexpr int kMaxBlockSize = 8;

__device__ __forceinline__ float block_product_entry(const float* __restrict__ A, bool ta, const float* __restrict__ B, bool tb, int s, int r, int c) {
	float acc = 0.f;
	for (int k = 0; k < s; k++) acc += (ta ? A[k * s + r] : A[r * s + k]) * (tb ? B[c * s + k] : B[k * s + c]);
	return acc;
}

// MatmulBlockSparseRowWisePadded (cpp/core/linalg/MatmulBlockSparseImpl.h:48-150): one thread per output coefficient
__global__ void __launch_bounds__(256) row_wise_kernel(const float* __restrict__ a_blocks, int a_count, const float* __restrict__ b_blocks,
                                                        const int32_t* __restrict__ b_coordinates, int64_t total, int s, float* __restrict__ c_blocks) {
	const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= total) return;
	const int64_t k = idx / (s * s);
	const int q = (int) (idx - k * s * s), row = b_coordinates[k * 2];
	c_blocks[idx] = row >= a_count ? 0.f : block_product_entry(a_blocks + (int64_t) row * s * s, false, b_blocks + k * s * s, false, s, q / s, q % s);
}

// MatmulBlockSparse (MatmulBlockSparseImpl.h:171-420): one thread per coefficient of every output block position; products summed
// over the inner block index in ascending order (the reference sums them with float atomics)
__global__ void __launch_bounds__(256) block_sparse_matmul_kernel(const float* __restrict__ a_blocks, const int16_t* __restrict__ a_board, int a_rows,
                                                                   int a_cols, int ta, const float* __restrict__ b_blocks,
                                                                   const int16_t* __restrict__ b_board, int b_rows, int b_cols, int tb, int s,
                                                                   float* __restrict__ out_blocks, uint8_t* __restrict__ out_mask) {
	const int out_rows = ta ? a_cols : a_rows, inner = ta ? a_rows : a_cols, out_cols = tb ? b_rows : b_cols;
	const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= (int64_t) out_rows * out_cols * s * s) return;
	const int position = (int) (idx / (s * s)), q = (int) (idx % (s * s)), i = position / out_cols, j = position % out_cols;
	float acc = 0.f;
	bool any = false;
	for (int k = 0; k < inner; k++) {
		const int ia = ta ? a_board[k * a_cols + i] : a_board[i * a_cols + k];
		const int ib = tb ? b_board[j * b_cols + k] : b_board[k * b_cols + j];
		if (ia < 0 || ib < 0) continue;
		acc += block_product_entry(a_blocks + (int64_t) ia * s * s, ta != 0, b_blocks + (int64_t) ib * s * s, tb != 0, s, q / s, q % s);
		any = true;
	}
	out_blocks[idx] = acc;
	if (q == 0) out_mask[position] = any ? 1 : 0;
}

// BlockSparseAndVectorProduct (MatmulBlockSparseImpl.h:441-600): one thread per output scalar, blocks visited in list order
__global__ void __launch_bounds__(128) block_sparse_vector_kernel(const float* __restrict__ blocks, const int32_t* __restrict__ coordinates, int count,
                                                                   int s, int m, int transpose, int row_offset, int column_offset,
                                                                   const float* __restrict__ b, float* __restrict__ c) {
	const int out = blockIdx.x * blockDim.x + threadIdx.x;
	if (out >= m) return;
	const int block_row = out / s, r = out % s;
	float total = 0.f;
	for (int k = 0; k < count; k++) {
		const int bi = transpose ? coordinates[k * 2 + 1] + column_offset : coordinates[k * 2] + row_offset;
		if (bi != block_row) continue;
		const int bj = transpose ? coordinates[k * 2] + row_offset : coordinates[k * 2 + 1] + column_offset;
		const float* A = blocks + (int64_t) k * s * s;
		float acc = 0.f;
		for (int q = 0; q < s; q++) acc += (transpose ? A[q * s + r] : A[r * s + q]) * b[bj * s + q];
		total += acc;
	}
	c[out] = total;
}

__global__ void __launch_bounds__(128) diagonal_block_vector_kernel(const float* __restrict__ blocks, int64_t total, int s, const float* __restrict__ b,
                                                                     float* __restrict__ c) {
	const int64_t out = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (out >= total) return;
	const int64_t k = out / s;
	const int r = (int) (out % s);
	float acc = 0.f;
	for (int q = 0; q < s; q++) acc += blocks[(k * s + r) * s + q] * b[k * s + q];
	c[out] = acc;
}

// InvertPositiveSemidefiniteBlocks (cpp/core/linalg/InvertBlocksCPU.cpp:80-112 potrf + potrs(I); InvertBlocksCUDA.cu:96-139): thread / block
__global__ void __launch_bounds__(64) invert_spd_blocks_kernel(const float* __restrict__ blocks, int count, int s, float* __restrict__ inverted,
                                                                int* __restrict__ failure_count) {
	const int b = blockIdx.x * blockDim.x + threadIdx.x;
	if (b >= count) return;
	float L[kMaxBlockSize * kMaxBlockSize], x[kMaxBlockSize];
	const float* A = blocks + (int64_t) b * s * s;
	for (int k = 0; k < s * s; k++) L[k] = A[k];
	for (int j = 0; j < s; j++) {
		float d = L[j * s + j];
		for (int k = 0; k < j; k++) d -= L[j * s + k] * L[j * s + k];
		if (!(d > 0.f)) { atomicAdd(failure_count, 1); return; }
		d = sqrtf(d);
		L[j * s + j] = d;
		for (int i = j + 1; i < s; i++) {
			float v = L[i * s + j];
			for (int k = 0; k < j; k++) v -= L[i * s + k] * L[j * s + k];
			L[i * s + j] = v / d;
		}
	}
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

// InvertTriangularBlocks (cpp/core/linalg/InvertBlocks.cpp:33-80): thread per (block, column)
__global__ void __launch_bounds__(64) invert_triangular_blocks_kernel(const float* __restrict__ blocks, int count, int s, int upper,
                                                                       float* __restrict__ inverted) {
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= count * s) return;
	const int b = idx / s, c = idx % s;
	const float* A = blocks + (int64_t) b * s * s;
	float* X = inverted + (int64_t) b * s * s;
	float x[kMaxBlockSize];
	for (int i = 0; i < s; i++) x[i] = 0.f;
	if (!upper) {
		for (int i = c; i < s; i++) {
			float v = i == c ? 1.f : 0.f;
			for (int k = c; k < i; k++) v -= A[i * s + k] * x[k];
			x[i] = v / A[i * s + i];
		}
	} else {
		for (int i = c; i >= 0; i--) {
			float v = i == c ? 1.f : 0.f;
			for (int k = i + 1; k <= c; k++) v -= A[i * s + k] * x[k];
			x[i] = v / A[i * s + i];
		}
	}
	for (int i = 0; i < s; i++) X[i * s + c] = x[i];
}

__global__ void __launch_bounds__(256) transpose_blocks_kernel(float* __restrict__ blocks, int64_t count, int s) {
	const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= count * s * s) return;
	const int64_t b = idx / (s * s);
	const int r = (int) (idx % (s * s)) / s, c = (int) (idx % s);
	if (c <= r) return;
	float* B = blocks + b * s * s;
	const float upper = B[r * s + c];
	B[r * s + c] = B[c * s + r];
	B[c * s + r] = upper;
}

// FillInDiagonalBlocks / GetDiagonalBlocks / FillInSparseBlocks / GetSparseBlocks (DiagonalBlocksImpl.h, SparseBlocksImpl.h:30-140)
__global__ void __launch_bounds__(256) sparse_blocks_kernel(float* __restrict__ matrix, int64_t cols, float* __restrict__ blocks,
                                                             const int32_t* __restrict__ coordinates /*nullptr: diagonal*/, int64_t count, int s,
                                                             int row_offset, int column_offset, int transpose, int gather) {
	const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= count * s * s) return;
	const int64_t b = idx / (s * s);
	const int q = (int) (idx % (s * s));
	int64_t bi = coordinates ? coordinates[b * 2] : b, bj = coordinates ? coordinates[b * 2 + 1] : b;
	if (transpose) { const int64_t t = bi; bi = bj; bj = t; }
	const int64_t r = (bi + row_offset) * s + (transpose ? q % s : q / s), c = (bj + column_offset) * s + (transpose ? q / s : q % s);
	if (gather) blocks[idx] = matrix[r * cols + c];
	else matrix[r * cols + c] = blocks[idx];
}

}  // namespace nnrt_b200

using namespace nnrt_b200;

#define NNRT_BLOCK_SIZE_CHECK(s) NNRT_CHECK_ARG((s) >= 1 && (s) <= kMaxBlockSize, "block size must be in [1, %d], got %d", kMaxBlockSize, (int) (s))

extern "C" int nnrt_b200_matmul_block_sparse_row_wise_padded(const float* blocks_a, int32_t a_block_count, const float* blocks_b,
                                                             const int32_t* blocks_b_coordinates, int32_t b_block_count, int32_t block_size,
                                                             float* blocks_c, void* stream) {
	NNRT_CHECK_ARG(blocks_a && blocks_b && blocks_b_coordinates && blocks_c, "matmul_block_sparse_row_wise: null pointer argument");
	NNRT_BLOCK_SIZE_CHECK(block_size);
	const int64_t total = (int64_t) b_block_count * block_size * block_size;
	if (total == 0) return kOk;
	row_wise_kernel<<<blocks_for(total, 256), 256, 0, (cudaStream_t) stream>>>(blocks_a, a_block_count, blocks_b, blocks_b_coordinates, total, block_size, blocks_c);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_matmul_block_sparse(const float* blocks_a, const int16_t* a_breadboard, int32_t a_rows, int32_t a_columns, int32_t transpose_a,
                                             const float* blocks_b, const int16_t* b_breadboard, int32_t b_rows, int32_t b_columns, int32_t transpose_b,
                                             int32_t block_size, float* out_blocks, uint8_t* out_block_mask, void* stream) {
	NNRT_CHECK_ARG(blocks_a && a_breadboard && blocks_b && b_breadboard && out_blocks && out_block_mask, "matmul_block_sparse: null pointer argument");
	NNRT_BLOCK_SIZE_CHECK(block_size);
	const int inner_a = transpose_a ? a_rows : a_columns, inner_b = transpose_b ? b_columns : b_rows;
	NNRT_CHECK_ARG(inner_a == inner_b, "matmul_block_sparse: inner block dimensions differ (%d vs %d)", inner_a, inner_b);
	const int64_t total = (int64_t) (transpose_a ? a_columns : a_rows) * (transpose_b ? b_rows : b_columns) * block_size * block_size;
	if (total == 0) return kOk;
	block_sparse_matmul_kernel<<<blocks_for(total, 256), 256, 0, (cudaStream_t) stream>>>(blocks_a, a_breadboard, a_rows, a_columns, transpose_a, blocks_b,
	                                                                                     b_breadboard, b_rows, b_columns, transpose_b, block_size,
	                                                                                     out_blocks, out_block_mask);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_block_sparse_and_vector_product(const float* blocks_a, const int32_t* blocks_a_coordinates, int32_t block_count,
                                                         int32_t block_size, int32_t output_length, int32_t transpose_a, int32_t block_row_offset,
                                                         int32_t block_column_offset, const float* vector_b, float* vector_c, void* stream) {
	NNRT_CHECK_ARG(blocks_a && blocks_a_coordinates && vector_b && vector_c, "block_sparse_and_vector_product: null pointer argument");
	NNRT_BLOCK_SIZE_CHECK(block_size);
	NNRT_CHECK_ARG(output_length % block_size == 0, "output length %d is not a multiple of the block size %d", output_length, block_size);
	if (output_length == 0) return kOk;
	block_sparse_vector_kernel<<<blocks_for(output_length, 128), 128, 0, (cudaStream_t) stream>>>(blocks_a, blocks_a_coordinates, block_count, block_size,
	                                                                                             output_length, transpose_a, block_row_offset,
	                                                                                             block_column_offset, vector_b, vector_c);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_diagonal_block_sparse_and_vector_product(const float* blocks_d, int32_t block_count, int32_t block_size, const float* vector_b,
                                                                  float* vector_c, void* stream) {
	NNRT_CHECK_ARG(blocks_d && vector_b && vector_c, "diagonal_block_sparse_and_vector_product: null pointer argument");
	NNRT_BLOCK_SIZE_CHECK(block_size);
	const int64_t total = (int64_t) block_count * block_size;
	if (total == 0) return kOk;
	diagonal_block_vector_kernel<<<blocks_for(total, 128), 128, 0, (cudaStream_t) stream>>>(blocks_d, total, block_size, vector_b, vector_c);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_invert_positive_semidefinite_blocks(const float* blocks, int32_t block_count, int32_t block_size, float* inverted_blocks,
                                                             int32_t* failure_count_device, void* stream) {
	NNRT_CHECK_ARG(blocks && inverted_blocks && failure_count_device, "invert_positive_semidefinite_blocks: null pointer argument");
	NNRT_BLOCK_SIZE_CHECK(block_size);
	NNRT_CUDA(cudaMemsetAsync(failure_count_device, 0, sizeof(int32_t), (cudaStream_t) stream));
	if (block_count == 0) return kOk;
	invert_spd_blocks_kernel<<<blocks_for(block_count, 64), 64, 0, (cudaStream_t) stream>>>(blocks, block_count, block_size, inverted_blocks, failure_count_device);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_invert_triangular_blocks(const float* blocks, int32_t block_count, int32_t block_size, int32_t upper, float* inverted_blocks,
                                                  void* stream) {
	NNRT_CHECK_ARG(blocks && inverted_blocks, "invert_triangular_blocks: null pointer argument");
	NNRT_BLOCK_SIZE_CHECK(block_size);
	if (block_count == 0) return kOk;
	invert_triangular_blocks_kernel<<<blocks_for((int64_t) block_count * block_size, 64), 64, 0, (cudaStream_t) stream>>>(blocks, block_count, block_size,
	                                                                                                                 upper, inverted_blocks);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_transpose_blocks_in_place(float* blocks, int64_t block_count, int32_t block_size, void* stream) {
	NNRT_CHECK_ARG(blocks, "transpose_blocks_in_place: null pointer argument");
	NNRT_CHECK_ARG(block_size >= 1, "block size must be positive");
	if (block_count == 0) return kOk;
	transpose_blocks_kernel<<<blocks_for(block_count * block_size * block_size, 256), 256, 0, (cudaStream_t) stream>>>(blocks, block_count, block_size);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

// gather == 0: FillInDiagonalBlocks / FillInSparseBlocks (coordinates == nullptr: diagonal); gather == 1: GetDiagonalBlocks / GetSparseBlocks
extern "C" int nnrt_b200_dense_matrix_blocks(float* matrix, int64_t matrix_rows, int64_t matrix_columns, float* blocks, const int32_t* coordinates,
                                             int64_t block_count, int32_t block_size, int32_t block_row_offset, int32_t block_column_offset,
                                             int32_t transpose, int32_t gather, void* stream) {
	NNRT_CHECK_ARG(matrix && blocks, "dense_matrix_blocks: null pointer argument");
	// SparseBlocksImpl.h:50-53
	NNRT_CHECK_ARG(block_size >= 1 && matrix_rows % block_size == 0 && matrix_columns % block_size == 0,
	               "Matrix row & column counts (presently, %lld and %lld) should be evenly divisible by block size, presently %d.",
	               (long long) matrix_rows, (long long) matrix_columns, block_size);
	if (block_count == 0) return kOk;
	sparse_blocks_kernel<<<blocks_for(block_count * block_size * block_size, 256), 256, 0, (cudaStream_t) stream>>>(
			matrix, matrix_columns, blocks, coordinates, block_count, block_size, block_row_offset, block_column_offset, transpose, gather);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
