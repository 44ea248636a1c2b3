#include "solver_kernels.cuh"

namespace nnrt_b200 {

// in-place lower Cholesky of a 6x6 (row-major) followed by inversion through two triangular solves per unit column
// (potrf + potrs(I), cpp/core/linalg/InvertBlocksCPU.cpp:80-112)
__device__ bool invert_spd6(const float* __restrict__ A, float* __restrict__ Ainv) {
	float L[36];
#pragma unroll
	for (int k = 0; k < 36; k++) L[k] = A[k];
	bool ok = true;
#pragma unroll
	for (int j = 0; j < 6; j++) {
		float dj = L[j * 6 + j];
#pragma unroll
		for (int k = 0; k < j; k++) dj -= L[j * 6 + k] * L[j * 6 + k];
		if (!(dj > 0.f)) ok = false;
		dj = sqrtf(dj);
		L[j * 6 + j] = dj;
#pragma unroll
		for (int i = j + 1; i < 6; i++) {
			float s = L[i * 6 + j];
#pragma unroll
			for (int k = 0; k < j; k++) s -= L[i * 6 + k] * L[j * 6 + k];
			L[i * 6 + j] = s / dj;
		}
	}
#pragma unroll
	for (int c = 0; c < 6; c++) {
		float x[6];
#pragma unroll
		for (int i = 0; i < 6; i++) {
			float s = (i == c) ? 1.f : 0.f;
#pragma unroll
			for (int k = 0; k < i; k++) s -= L[i * 6 + k] * x[k];
			x[i] = s / L[i * 6 + i];
		}
#pragma unroll
		for (int i = 5; i >= 0; i--) {
			float s = x[i];
#pragma unroll
			for (int k = i + 1; k < 6; k++) s -= L[k * 6 + i] * x[k];
			x[i] = s / L[i * 6 + i];
		}
#pragma unroll
		for (int i = 0; i < 6; i++) Ainv[i * 6 + c] = x[i];
	}
	return ok;
}

__global__ void __launch_bounds__(128) stem_eliminate_kernel(ArrowheadSystem sys) {
	__shared__ float s_dinv[4][36];
	__shared__ float s_b[4][6];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const int row = blockIdx.x * 4 + wib;
	if (row >= sys.node_count) return;
	const int nc = sys.node_count - sys.arrow_base, n = nc * 6;
	const float* D = sys.diagonal_blocks + (int64_t) row * 36;
	if (row >= sys.arrow_base) {  // corner row: scatter the diagonal block and rhs into the dense corner
		const int c = row - sys.arrow_base;
		for (int k = lane; k < 36; k += 32) atomicAdd(sys.schur + (int64_t) (c * 6 + k / 6) * n + c * 6 + k % 6, D[k]);
		if (lane < 6) atomicAdd(sys.corner_rhs + c * 6 + lane, sys.rhs[row * 6 + lane]);
		return;
	}
	if (lane == 0) {
		if (!invert_spd6(D, s_dinv[wib])) atomicExch(sys.flags + 2, 1);
	}
	if (lane < 6) s_b[wib][lane] = sys.rhs[row * 6 + lane];
	__syncwarp();
	for (int k = lane; k < 36; k += 32) sys.inverted_stem[(int64_t) row * 36 + k] = s_dinv[wib][k];
	if (lane < 6) {
		float acc = 0.f;
#pragma unroll
		for (int k = 0; k < 6; k++) acc += s_dinv[wib][lane * 6 + k] * s_b[wib][k];
		sys.stem_y[row * 6 + lane] = acc;
	}
	const int e0 = sys.row_offsets[row], e1 = sys.row_offsets[row + 1], deg = e1 - e0;
	// X_e = D^-1 B_e
	for (int idx = lane; idx < deg * 36; idx += 32) {
		const int w = sys.row_blocks[e0 + idx / 36], r = (idx % 36) / 6, c = idx % 6;
		const float* B = sys.wing_blocks + (int64_t) w * 36;
		float acc = 0.f;
#pragma unroll
		for (int k = 0; k < 6; k++) acc += s_dinv[wib][r * 6 + k] * B[k * 6 + c];
		sys.wing_x[(int64_t) w * 36 + idx % 36] = acc;
	}
	__syncwarp();
	__threadfence_block();
	// S[c1, c2] -= B_e1^T X_e2
	for (int idx = lane; idx < deg * deg * 36; idx += 32) {
		const int pair = idx / 36, r = (idx % 36) / 6, c = idx % 6;
		const int w1 = sys.row_blocks[e0 + pair / deg], w2 = sys.row_blocks[e0 + pair % deg];
		const float* B1 = sys.wing_blocks + (int64_t) w1 * 36;
		const float* X2 = sys.wing_x + (int64_t) w2 * 36;
		float acc = 0.f;
#pragma unroll
		for (int k = 0; k < 6; k++) acc += B1[k * 6 + r] * X2[k * 6 + c];
		const int c1 = sys.wing_columns[w1] - sys.arrow_base, c2 = sys.wing_columns[w2] - sys.arrow_base;
		atomicAdd(sys.schur + (int64_t) (c1 * 6 + r) * n + c2 * 6 + c, -acc);
	}
	// b_C[c] -= X_e^T b_row
	for (int idx = lane; idx < deg * 6; idx += 32) {
		const int w = sys.row_blocks[e0 + idx / 6], c = idx % 6;
		const float* X = sys.wing_x + (int64_t) w * 36;
		float acc = 0.f;
#pragma unroll
		for (int k = 0; k < 6; k++) acc += X[k * 6 + c] * s_b[wib][k];
		atomicAdd(sys.corner_rhs + (sys.wing_columns[w] - sys.arrow_base) * 6 + c, -acc);
	}
}

__global__ void __launch_bounds__(256) corner_upper_scatter_kernel(ArrowheadSystem sys) {
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= sys.corner_upper_count * 36) return;
	const int u = idx / 36, r = (idx % 36) / 6, c = idx % 6;
	const int n = (sys.node_count - sys.arrow_base) * 6;
	const int bi = sys.corner_upper_coordinates[u * 2] - sys.arrow_base, bj = sys.corner_upper_coordinates[u * 2 + 1] - sys.arrow_base;
	const float v = sys.corner_upper_blocks[idx];
	atomicAdd(sys.schur + (int64_t) (bi * 6 + r) * n + bj * 6 + c, v);
	atomicAdd(sys.schur + (int64_t) (bj * 6 + c) * n + bi * 6 + r, v);
}

__device__ __forceinline__ int64_t packed_index(int i, int k) { return (int64_t) i * (i + 1) / 2 + k; }

// core::linalg::AxisAngleVectorsToMatricesRodrigues (cpp/core/linalg/RodriguesImpl.h:66-100), no small-angle guard (as the reference)
__device__ void rodrigues(const float* __restrict__ w, float* __restrict__ R) {
	const float angle = sqrtf(sum3(w[0] * w[0], w[1] * w[1], w[2] * w[2]));
	const float kx = w[0] / angle, ky = w[1] / angle, kz = w[2] / angle;
	const float K[9] = {0.f, -kz, ky, kz, 0.f, -kx, -ky, kx, 0.f};
	const float s = sinf(angle), c1 = 1 - cosf(angle);
#pragma unroll
	for (int r = 0; r < 3; r++)
#pragma unroll
		for (int c = 0; c < 3; c++) {
			const float kk = sum3(K[r * 3] * K[c], K[r * 3 + 1] * K[3 + c], K[r * 3 + 2] * K[6 + c]);
			R[r * 3 + c] = ((r == c ? 1.f : 0.f) + s * K[r * 3 + c]) + c1 * kk;
		}
}

template<bool kSharedScratch>
__global__ void __launch_bounds__(1024) corner_solve_kernel(ArrowheadSystem sys, float* __restrict__ x, float* __restrict__ rotations,
                                                             float* __restrict__ translations, int clear_corner,
                                                             float* __restrict__ scratch_global) {
	extern __shared__ float smem[];
	const int tid = threadIdx.x, T = blockDim.x;
	const int nc = sys.node_count - sys.arrow_base, n = nc * 6;
	const int64_t packed_count = (int64_t) (n + 1) * (n + 2) / 2;
	float* L = kSharedScratch ? smem : scratch_global;                      // packed lower triangle, row n = rhs
	float* col = kSharedScratch ? smem + packed_count : scratch_global + packed_count;  // [n+1]
	if (n > 0) {
		for (int64_t q = tid; q < (int64_t) n * n; q += T) {
			const int i = (int) (q / n), k = (int) (q % n);
			if (k <= i) L[packed_index(i, k)] = sys.schur[q];
			if (clear_corner) sys.schur[q] = 0.f;
		}
		for (int k = tid; k < n; k += T) {
			L[packed_index(n, k)] = sys.corner_rhs[k];
			if (clear_corner) sys.corner_rhs[k] = 0.f;
		}
		__syncthreads();
		const int tx = tid & 31, ty = tid >> 5, TY = T >> 5;
		for (int j = 0; j < n; j++) {
			if (tid == 0) {
				float dj = L[packed_index(j, j)];
				if (!(dj > 0.f)) atomicExch(sys.flags + 2, 1);
				L[packed_index(j, j)] = sqrtf(dj);
			}
			__syncthreads();
			const float dj = L[packed_index(j, j)];
			for (int i = j + 1 + tid; i <= n; i += T) {
				const float v = L[packed_index(i, j)] / dj;
				L[packed_index(i, j)] = v;
				col[i] = v;
			}
			__syncthreads();
			for (int i = j + 1 + ty; i <= n; i += TY) {
				const float ci = col[i];
				const int k_end = min(i, n - 1);
				float* Li = L + packed_index(i, 0);
				for (int k = j + 1 + tx; k <= k_end; k += 32) Li[k] -= ci * col[k];
			}
			__syncthreads();
		}
		// back substitution L^T x = y with y = row n, one warp
		if (tid < 32) {
			float* y = L + packed_index(n, 0);
			for (int j = n - 1; j >= 0; j--) {
				const float xj = y[j] / L[packed_index(j, j)];
				__syncwarp();
				if (tid == 0) y[j] = xj;
				const float* Lj = L + packed_index(j, 0);
				for (int k = tid; k < j; k += 32) y[k] -= Lj[k] * xj;
				__syncwarp();
			}
		}
		__syncthreads();
		for (int k = tid; k < n; k += T) x[sys.arrow_base * 6 + k] = L[packed_index(n, k)];
	}
	// stem back-substitution: x_i = y_i - sum_e X_e x_C[col(e)]
	const float* xc = L + packed_index(n, 0);
	for (int idx = tid; idx < sys.arrow_base * 6; idx += T) {
		const int i = idx / 6, r = idx % 6;
		float acc = sys.stem_y[idx];
		for (int e = sys.row_offsets[i]; e < sys.row_offsets[i + 1]; e++) {
			const int w = sys.row_blocks[e];
			const float* X = sys.wing_x + (int64_t) w * 36 + r * 6;
			const float* xv = xc + (sys.wing_columns[w] - sys.arrow_base) * 6;
#pragma unroll
			for (int k = 0; k < 6; k++) acc -= X[k] * xv[k];
		}
		x[idx] = acc;
	}
	if (rotations == nullptr) return;
	__syncthreads();
	__threadfence_block();
	// node update (DeformableMeshToImageFitter.cpp:258-273, HierarchicalGraphWarpField.cpp:261-284): t += dt, R <- R * Rodrigues(dw)
	for (int node = tid; node < sys.node_count; node += T) {
		const float* dlt = x + node * 6;
		float dR[9], R[9];
		rodrigues(dlt, dR);
#pragma unroll
		for (int k = 0; k < 9; k++) R[k] = rotations[node * 9 + k];
#pragma unroll
		for (int r = 0; r < 3; r++)
#pragma unroll
			for (int c = 0; c < 3; c++) {
				float acc = 0.f;
#pragma unroll
				for (int k = 0; k < 3; k++) acc += R[r * 3 + k] * dR[k * 3 + c];
				rotations[node * 9 + r * 3 + c] = acc;
			}
#pragma unroll
		for (int c = 0; c < 3; c++) translations[node * 3 + c] = translations[node * 3 + c] + dlt[3 + c];
	}
}

size_t corner_scratch_bytes(int corner_blocks) {
	const int64_t n = (int64_t) corner_blocks * 6;
	return (size_t) (((n + 1) * (n + 2) / 2 + (n + 1)) * sizeof(float));
}

void launch_stem_eliminate(const ArrowheadSystem& sys, cudaStream_t s) {
	if (sys.node_count > 0) stem_eliminate_kernel<<<(sys.node_count + 3) / 4, 128, 0, s>>>(sys);
	if (sys.corner_upper_count > 0) corner_upper_scatter_kernel<<<blocks_for((int64_t) sys.corner_upper_count * 36, 256), 256, 0, s>>>(sys);
}

int launch_corner_solve(const ArrowheadSystem& sys, float* x, float* node_rotations, float* node_translations, bool clear_corner,
                        float* scratch_global, cudaStream_t s) {
	const size_t bytes = corner_scratch_bytes(sys.node_count - sys.arrow_base);
	static bool attribute_set = false;
	constexpr size_t kMaxShared = 227 * 1024;
	if (bytes <= kMaxShared) {
		if (!attribute_set) {
			if (cudaFuncSetAttribute(corner_solve_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kMaxShared) != cudaSuccess) return kCudaError;
			attribute_set = true;
		}
		corner_solve_kernel<true><<<1, 1024, bytes, s>>>(sys, x, node_rotations, node_translations, clear_corner ? 1 : 0, nullptr);
	} else {
		if (scratch_global == nullptr) return kUnsupported;
		corner_solve_kernel<false><<<1, 1024, 0, s>>>(sys, x, node_rotations, node_translations, clear_corner ? 1 : 0, scratch_global);
	}
	return kOk;
}

}  // namespace nnrt_b200
