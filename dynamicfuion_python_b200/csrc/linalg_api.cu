// Stage-level core::linalg entry points: block-sparse arrowhead solve / Schur complement on explicit blocks,
// block-diagonal Cholesky solve, LM damping, Rodrigues, batched 3x3 matmul.
#include <vector>
#include "solver_kernels.cuh"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {

__global__ void __launch_bounds__(128) block_diagonal_solve_kernel(const float* __restrict__ blocks, const float* __restrict__ b, int count,
                                                                    int bs, float* __restrict__ x, int* __restrict__ flag) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	float L[36], y[6];
	const float* A = blocks + (int64_t) i * bs * bs;
	for (int k = 0; k < bs * bs; k++) L[k] = A[k];
	bool ok = true;
	for (int j = 0; j < bs; j++) {  // potrf
		float dj = L[j * bs + j];
		for (int k = 0; k < j; k++) dj -= L[j * bs + k] * L[j * bs + k];
		if (!(dj > 0.f)) ok = false;
		dj = sqrtf(dj);
		L[j * bs + j] = dj;
		for (int r = j + 1; r < bs; r++) {
			float s = L[r * bs + j];
			for (int k = 0; k < j; k++) s -= L[r * bs + k] * L[j * bs + k];
			L[r * bs + j] = s / dj;
		}
	}
	for (int r = 0; r < bs; r++) {  // two trsm
		float s = b[i * bs + r];
		for (int k = 0; k < r; k++) s -= L[r * bs + k] * y[k];
		y[r] = s / L[r * bs + r];
	}
	for (int r = bs - 1; r >= 0; r--) {
		float s = y[r];
		for (int k = r + 1; k < bs; k++) s -= L[k * bs + r] * y[k];
		y[r] = s / L[r * bs + r];
	}
	for (int r = 0; r < bs; r++) x[i * bs + r] = y[r];
	if (!ok && flag) atomicExch(flag, 1);
}

__global__ void __launch_bounds__(256) precondition_kernel(float* blocks, int64_t count, int bs, float lambda) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count * bs) return;
	const int64_t blk = i / bs;
	const int d = (int) (i % bs);
	blocks[blk * bs * bs + d * bs + d] += lambda;
}


__global__ void __launch_bounds__(128) rodrigues_kernel(const float* __restrict__ v, int count, float* __restrict__ out) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	const float w[3] = {v[i * 3], v[i * 3 + 1], v[i * 3 + 2]};
	const float angle = sqrtf(sum3(w[0] * w[0], w[1] * w[1], w[2] * w[2]));
	const float kx = w[0] / angle, ky = w[1] / angle, kz = w[2] / angle;
	const float K[9] = {0.f, -kz, ky, kz, 0.f, -kx, -ky, kx, 0.f};
	const float s = sinf(angle), c1 = 1 - cosf(angle);
	for (int r = 0; r < 3; r++)
		for (int c = 0; c < 3; c++) {
			const float kk = sum3(K[r * 3] * K[c], K[r * 3 + 1] * K[3 + c], K[r * 3 + 2] * K[6 + c]);
			out[i * 9 + r * 3 + c] = ((r == c ? 1.f : 0.f) + s * K[r * 3 + c]) + c1 * kk;
		}
}

__global__ void __launch_bounds__(128) matmul3x3_kernel(const float* __restrict__ a, const float* __restrict__ b, int count, float* __restrict__ c) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	float A[9], B[9];
	for (int k = 0; k < 9; k++) { A[k] = a[i * 9 + k]; B[k] = b[i * 9 + k]; }
	for (int r = 0; r < 3; r++)
		for (int cc = 0; cc < 3; cc++) {
			float acc = 0.f;
			for (int k = 0; k < 3; k++) acc += A[r * 3 + k] * B[k * 3 + cc];
			c[i * 9 + r * 3 + cc] = acc;
		}
}

// Explicit-block arrowhead workspace: CSR over wing rows is built on the host from the coordinate list (the reference
// also synchronises with the host several times inside this call, see SURVEY.md §3.1 "syncs").
struct ExplicitArrowhead {
	explicit ExplicitArrowhead(cudaStream_t stream) : scratch(stream) {}
	ArrowheadSystem sys{};
	char* buffer = nullptr;
	Scratch scratch;  // returns `buffer` to the pool on every path
};

static int prepare_explicit(ExplicitArrowhead& ws, const float* diag, int N, int base, const float* wing, const int32_t* coords, int64_t W,
                            const float* cu, const int32_t* cuc, int64_t Cu, const float* b, cudaStream_t s) {
	const int nc = N - base;
	const size_t n = (size_t) nc * 6;
	std::vector<int32_t> h((size_t) W * 2);
	if (W > 0) {
		NNRT_CUDA(cudaMemcpyAsync(h.data(), coords, sizeof(int32_t) * W * 2, cudaMemcpyDeviceToHost, s));
		NNRT_CUDA(cudaStreamSynchronize(s));
	}
	std::vector<int> off(base + 2, 0), blocks((size_t) W + 1), cols((size_t) W + 1);
	for (int64_t w = 0; w < W; w++) {
		const int i = h[w * 2], j = h[w * 2 + 1];
		NNRT_CHECK_ARG(i >= 0 && i < base && j >= base && j < N, "upper wing block %lld at (%d, %d) is outside the wing (rows < %d, columns in [%d, %d))",
		               (long long) w, i, j, base, base, N);
		off[i + 1]++; cols[w] = j;
	}
	for (int i = 0; i < base; i++) off[i + 1] += off[i];
	std::vector<int> cur(off.begin(), off.end());
	for (int64_t w = 0; w < W; w++) blocks[cur[h[w * 2]]++] = (int) w;
	auto al = [](size_t v) { return (v + 255) & ~size_t(255); };
	const size_t sz_off = al(sizeof(int) * (base + 2)), sz_blk = al(sizeof(int) * (W + 1)), sz_dinv = al(sizeof(float) * (size_t) (base + 1) * 36),
	             sz_y = al(sizeof(float) * (size_t) (base + 1) * 6), sz_x = al(sizeof(float) * (size_t) (W + 1) * 36), sz_s = al(sizeof(float) * corner_schur_floats(nc)),
	             sz_r = al(sizeof(float) * (n + 1)), sz_flags = 256;
	const size_t total = sz_off + 2 * sz_blk + sz_dinv + sz_y + sz_x + sz_s + sz_r + sz_flags;
	NNRT_CUDA(ws.scratch.alloc(&ws.buffer, total));
	char* p = ws.buffer;
	int* d_off = (int*) p; p += sz_off;
	int* d_blk = (int*) p; p += sz_blk;
	int* d_col = (int*) p; p += sz_blk;
	ws.sys.inverted_stem = (float*) p; p += sz_dinv;
	ws.sys.stem_y = (float*) p; p += sz_y;
	ws.sys.wing_x = (float*) p; p += sz_x;
	ws.sys.schur = (float*) p; p += sz_s;
	ws.sys.corner_rhs = (float*) p; p += sz_r;
	ws.sys.flags = (int*) p; p += sz_flags;
	NNRT_CUDA(cudaMemsetAsync(ws.sys.schur, 0, sz_s + sz_r + sz_flags, s));
	NNRT_CUDA(cudaMemcpyAsync(d_off, off.data(), sizeof(int) * (base + 2), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(d_blk, blocks.data(), sizeof(int) * (W + 1), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(d_col, cols.data(), sizeof(int) * (W + 1), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaStreamSynchronize(s));
	ws.sys.node_count = N; ws.sys.arrow_base = base; ws.sys.diagonal_blocks = diag; ws.sys.rhs = b;
	ws.sys.wing_blocks = wing; ws.sys.wing_columns = d_col; ws.sys.row_offsets = d_off; ws.sys.row_blocks = d_blk;
	ws.sys.corner_upper_blocks = cu; ws.sys.corner_upper_coordinates = cuc; ws.sys.corner_upper_count = (int) Cu;
	return kOk;
}

__global__ void __launch_bounds__(256) zero_kernel(float* p, int64_t n) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) p[i] = 0.f;
}

}  // namespace nnrt_b200

using namespace nnrt_b200;

static int check_arrowhead_args(const float* diag, int N, int base, const float* wing, const int32_t* coords, int64_t W) {
	NNRT_CHECK_ARG(diag && N > 0, "arrowhead: diagonal blocks missing");
	NNRT_CHECK_ARG(base >= 0 && base <= N, "arrow_base_block_index (%d) must be in [0, %d]", base, N);
	NNRT_CHECK_ARG(W == 0 || (wing && coords), "arrowhead: wing blocks / coordinates missing");
	return kOk;
}

extern "C" int nnrt_b200_solve_block_sparse_arrowhead_cholesky(const float* diagonal_blocks, int32_t diagonal_block_count, int32_t arrow_base,
                                                               const float* upper_wing_blocks, const int32_t* upper_wing_coordinates,
                                                               int64_t upper_wing_count, const float* corner_upper_blocks,
                                                               const int32_t* corner_upper_coordinates, int64_t corner_upper_count,
                                                               const float* b, float* x, void* stream) {
	int rc = check_arrowhead_args(diagonal_blocks, diagonal_block_count, arrow_base, upper_wing_blocks, upper_wing_coordinates, upper_wing_count);
	if (rc != kOk) return rc;
	NNRT_CHECK_ARG(b && x, "Size of lhs tensor b must be {%d}, matching the leading dimension of rhs matrix a.", diagonal_block_count * 6);
	auto s = (cudaStream_t) stream;
	ExplicitArrowhead ws(s);
	rc = prepare_explicit(ws, diagonal_blocks, diagonal_block_count, arrow_base, upper_wing_blocks, upper_wing_coordinates, upper_wing_count,
	                      corner_upper_blocks, corner_upper_coordinates, corner_upper_count, b, s);
	if (rc != kOk) return rc;
	ws.sys.schur_tile_packed = corner_uses_tile_packed_schur(diagonal_block_count - arrow_base) ? 1 : 0;
	launch_stem_eliminate(ws.sys, s);
	rc = launch_corner_solve(ws.sys, x, nullptr, nullptr, false, s);
	int flags[3] = {0, 0, 0};
	cudaMemcpyAsync(flags, ws.sys.flags, sizeof(flags), cudaMemcpyDeviceToHost, s);
	cudaStreamSynchronize(s);
	if (rc != kOk) return rc;
	NNRT_LAUNCH_CHECK();
	if (flags[2]) { set_last_error("solve_block_sparse_arrowhead_cholesky: matrix is not positive definite"); return kNotPositiveDefinite; }
	return kOk;
}

extern "C" int nnrt_b200_arrowhead_schur_complement(const float* diagonal_blocks, int32_t diagonal_block_count, int32_t arrow_base,
                                                    const float* upper_wing_blocks, const int32_t* upper_wing_coordinates,
                                                    int64_t upper_wing_count, const float* corner_upper_blocks,
                                                    const int32_t* corner_upper_coordinates, int64_t corner_upper_count, float* schur,
                                                    void* stream) {
	int rc = check_arrowhead_args(diagonal_blocks, diagonal_block_count, arrow_base, upper_wing_blocks, upper_wing_coordinates, upper_wing_count);
	if (rc != kOk) return rc;
	NNRT_CHECK_ARG(schur, "arrowhead_schur_complement: null output");
	auto s = (cudaStream_t) stream;
	const int64_t n = (int64_t) (diagonal_block_count - arrow_base) * 6;
	ExplicitArrowhead ws(s);
	float* zero_rhs = nullptr;
	NNRT_CUDA(ws.scratch.alloc(&zero_rhs, sizeof(float) * (size_t) diagonal_block_count * 6));
	NNRT_CUDA(cudaMemsetAsync(zero_rhs, 0, sizeof(float) * (size_t) diagonal_block_count * 6, s));
	rc = prepare_explicit(ws, diagonal_blocks, diagonal_block_count, arrow_base, upper_wing_blocks, upper_wing_coordinates, upper_wing_count,
	                      corner_upper_blocks, corner_upper_coordinates, corner_upper_count, zero_rhs, s);
	if (rc == kOk) {
		launch_stem_eliminate(ws.sys, s);
		if (n > 0) cudaMemcpyAsync(schur, ws.sys.schur, sizeof(float) * n * n, cudaMemcpyDeviceToDevice, s);
	}
	cudaStreamSynchronize(s);
	if (rc != kOk) return rc;
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_solve_block_diagonal_cholesky(const float* blocks, const float* b, int32_t block_count, int32_t block_size, float* x,
                                                       void* stream) {
	NNRT_CHECK_ARG(blocks && b && x, "solve_block_diagonal_cholesky: null pointer argument");
	NNRT_CHECK_ARG(block_size >= 1 && block_size <= 6, "block size must be in [1, 6], got %d", block_size);
	if (block_count == 0) return kOk;
	block_diagonal_solve_kernel<<<blocks_for(block_count, 128), 128, 0, (cudaStream_t) stream>>>(blocks, b, block_count, block_size, x, nullptr);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_precondition_diagonal_blocks(float* blocks, int32_t block_count, int32_t block_size, float lambda, void* stream) {
	NNRT_CHECK_ARG(blocks, "precondition_diagonal_blocks: null pointer argument");
	if (block_count == 0) return kOk;
	precondition_kernel<<<blocks_for((int64_t) block_count * block_size, 256), 256, 0, (cudaStream_t) stream>>>(blocks, block_count, block_size, lambda);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_rodrigues(const float* axis_angles, int32_t count, float* matrices, void* stream) {
	NNRT_CHECK_ARG(axis_angles && matrices, "rodrigues: null pointer argument");
	if (count == 0) return kOk;
	rodrigues_kernel<<<blocks_for(count, 128), 128, 0, (cudaStream_t) stream>>>(axis_angles, count, matrices);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_matmul3d_3x3(const float* a, const float* b, int32_t count, float* c, void* stream) {
	NNRT_CHECK_ARG(a && b && c, "matmul3d: null pointer argument");
	if (count == 0) return kOk;
	matmul3x3_kernel<<<blocks_for(count, 128), 128, 0, (cudaStream_t) stream>>>(a, b, count, c);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
