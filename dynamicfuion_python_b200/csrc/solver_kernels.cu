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


// node update (DeformableMeshToImageFitter.cpp:258-273, HierarchicalGraphWarpField.cpp:261-284): t += dt, R <- R * Rodrigues(dw)
__device__ __forceinline__ void apply_node_update(const float* __restrict__ dlt, float* __restrict__ rotation, float* __restrict__ translation) {
	float dR[9], R[9];
	rodrigues(dlt, dR);
#pragma unroll
	for (int k = 0; k < 9; k++) R[k] = rotation[k];
#pragma unroll
	for (int r = 0; r < 3; r++)
#pragma unroll
		for (int c = 0; c < 3; c++) {
			float acc = 0.f;
#pragma unroll
			for (int k = 0; k < 3; k++) acc += R[r * 3 + k] * dR[k * 3 + c];
			rotation[r * 3 + c] = acc;
		}
#pragma unroll
	for (int c = 0; c < 3; c++) translation[c] = translation[c] + dlt[3 + c];
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


// ---- blocked single-CTA corner solve ------------------------------------------------------------------------------------
// Right-looking Cholesky in shared memory, panels of kPanel = 8 columns, the right-hand side carried as row n (forward
// substitution for free). Storage: lower triangle, rows padded to a multiple of 4 columns so every row starts 16-byte aligned
// (row i of 4-row group g = i / 4 has width 4 (g + 1)); the padding lies in the strict upper triangle and is never read as data.
//   panel phase   : one thread per row i >= j. Every row thread factors the 8 x 8 diagonal block redundantly in registers
//                   (one rsqrt per pivot), solves its own row strip against it and publishes the strip to PT[i][0..7]
//                   (row stride 12 floats: conflict-free 128-bit reads by 8 consecutive rows).
//   trailing phase: all warps; a warp owns a 4-row group, every lane a 4 x 4 tile (128-bit loads/stores of L, 128 FMAs per
//                   16 shared-memory instructions), rank-8 update.
// Two __syncthreads per panel; the back substitution is blocked the same way on a double-buffered right-hand side.
constexpr int kPanel = 8;
constexpr int kCornerThreads = 512;
constexpr int kPanelStride = 12;

__host__ __device__ __forceinline__ int aligned_row_start(int i) {
	const int g = i >> 2;
	return 8 * g * (g + 1) + (i - 4 * g) * 4 * (g + 1);
}
__host__ __device__ inline int padded_rows(int n) { return (n + 1 + 3) & ~3; }

__host__ __device__ inline size_t blocked_corner_floats(int n) {
	const int pr = padded_rows(n);
	return (size_t) aligned_row_start(pr) + (size_t) kPanelStride * pr + 4 * (size_t) pr;
}

__global__ void __launch_bounds__(kCornerThreads) corner_solve_blocked_kernel(ArrowheadSystem sys, float* __restrict__ x,
                                                                              float* __restrict__ rotations, float* __restrict__ translations,
                                                                              int clear_corner) {
	extern __shared__ __align__(16) float smem[];
	const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, warps = T >> 5;
	const int nc = sys.node_count - sys.arrow_base, n = nc * 6;
	const int rows = n + 1;                                   // + rhs row
	const int prow = padded_rows(n);
	float* L = smem;                                          // aligned-packed lower triangle, rows 0..prow-1
	float* PT = L + aligned_row_start(prow);                  // [prow][kPanelStride]
	float* inv_diag = PT + kPanelStride * prow;               // [prow]
	float* ybuf = inv_diag + prow;                            // [2][prow]
	float* xs = ybuf + 2 * prow;                              // [prow] solution of the corner
	if (n > 0) {
		// zero the padding rows/columns once (finite values everywhere the tiles may touch)
		for (int q = aligned_row_start(n) + tid; q < aligned_row_start(prow); q += T) L[q] = 0.f;
		for (int q = tid; q < kPanelStride * prow; q += T) PT[q] = 0.f;
		// stage the lower triangle of S: batches of independent 64-bit loads (n is even, rows are 8-byte aligned); zeroing
		// happens after the whole batch is in registers so the loads are not serialised behind the stores
		{
			const int half = n >> 1, total = n * half;
			constexpr int kBatch = 8;
			const float2* S2 = reinterpret_cast<const float2*>(sys.schur);
			float2* S2w = reinterpret_cast<float2*>(sys.schur);
			for (int base = 0; base < total; base += T * kBatch) {
				float2 v[kBatch];
#pragma unroll
				for (int b = 0; b < kBatch; b++) {
					const int q = base + b * T + tid;
					v[b] = make_float2(0.f, 0.f);
					if (q < total) {
						const int i = q / half, k2 = q - i * half;
						if (2 * k2 <= i) v[b] = __ldcg(S2 + q);
					}
				}
#pragma unroll
				for (int b = 0; b < kBatch; b++) {
					const int q = base + b * T + tid;
					if (q < total) {
						const int i = q / half, k = 2 * (q - i * half);
						if (k <= i) {
							// the element right of the diagonal (k + 1 == i + 1) lands in the row's padding
							*reinterpret_cast<float2*>(L + aligned_row_start(i) + k) = make_float2(v[b].x, k + 1 <= i ? v[b].y : 0.f);
							if (clear_corner) S2w[q] = make_float2(0.f, 0.f);
						}
					}
				}
			}
		}
		__syncthreads();  // padding of row n was zeroed above; now place the right-hand side
		for (int k = tid; k < n; k += T) {
			L[aligned_row_start(n) + k] = sys.corner_rhs[k];
			if (clear_corner) sys.corner_rhs[k] = 0.f;
		}
		__syncthreads();
		for (int j = 0; j < n; j += kPanel) {
			const int nb = min(kPanel, n - j);
			// ---- panel phase
			for (int i = j + tid; i < rows; i += T) {
				if (i > j && i < j + nb) continue;  // rows of the diagonal block are written by the thread of row j
				float d[kPanel][kPanel], inv[kPanel];
#pragma unroll
				for (int r = 0; r < kPanel; r++) {
					const float* Lr = L + aligned_row_start(min(j + r, rows - 1)) + j;
#pragma unroll
					for (int c = 0; c <= r; c++) d[r][c] = (r < nb) ? Lr[c] : (r == c ? 1.f : 0.f);
				}
				bool ok = true;
#pragma unroll
				for (int c = 0; c < kPanel; c++) {
					float v = d[c][c];
#pragma unroll
					for (int k = 0; k < c; k++) v = fmaf(-d[c][k], d[c][k], v);
					if (!(v > 0.f)) ok = false;
					inv[c] = rsqrtf(v);
					d[c][c] = v * inv[c];
#pragma unroll
					for (int r = c + 1; r < kPanel; r++) {
						float w = d[r][c];
#pragma unroll
						for (int k = 0; k < c; k++) w = fmaf(-d[r][k], d[c][k], w);
						d[r][c] = w * inv[c];
					}
				}
				if (i == j) {
					if (!ok) atomicExch(sys.flags + 2, 1);
#pragma unroll
					for (int r = 0; r < kPanel; r++) {
						if (r < nb) {
							inv_diag[j + r] = inv[r];
							float* Lr = L + aligned_row_start(j + r) + j;
#pragma unroll
							for (int c = 0; c <= r; c++) Lr[c] = d[r][c];
						}
					}
				} else {
					float* Li = L + aligned_row_start(i) + j;
					float a[kPanel];
					if (nb == kPanel) {
						const float4 lo = *reinterpret_cast<const float4*>(Li), hi = *reinterpret_cast<const float4*>(Li + 4);
						a[0] = lo.x; a[1] = lo.y; a[2] = lo.z; a[3] = lo.w; a[4] = hi.x; a[5] = hi.y; a[6] = hi.z; a[7] = hi.w;
					} else {
#pragma unroll
						for (int c = 0; c < kPanel; c++) a[c] = (c < nb) ? Li[c] : 0.f;
					}
#pragma unroll
					for (int c = 0; c < kPanel; c++) {
						float w = a[c];
#pragma unroll
						for (int k = 0; k < c; k++) w = fmaf(-a[k], d[c][k], w);
						a[c] = w * inv[c];
					}
					if (nb == kPanel) {
						*reinterpret_cast<float4*>(Li) = make_float4(a[0], a[1], a[2], a[3]);
						*reinterpret_cast<float4*>(Li + 4) = make_float4(a[4], a[5], a[6], a[7]);
						float* P = PT + i * kPanelStride;
						*reinterpret_cast<float4*>(P) = make_float4(a[0], a[1], a[2], a[3]);
						*reinterpret_cast<float4*>(P + 4) = make_float4(a[4], a[5], a[6], a[7]);
					} else {
#pragma unroll
						for (int c = 0; c < kPanel; c++)
							if (c < nb) Li[c] = a[c];
					}
				}
			}
			__syncthreads();
			if (nb < kPanel) break;  // a partial last panel leaves only the rhs row, which has no columns >= n to update
			// ---- trailing phase: 4-row groups [i0, i0+3] from j + 8, 4-column tiles k in [j + 8, i0]
			const int first = j + kPanel;
			for (int i0 = first + warp * 4; i0 < rows; i0 += warps * 4) {
				float4 pr[4][2];
				float* Lrow[4];
#pragma unroll
				for (int r = 0; r < 4; r++) {
					Lrow[r] = L + aligned_row_start(i0 + r);
					const float* P = PT + (i0 + r) * kPanelStride;
					pr[r][0] = *reinterpret_cast<const float4*>(P);
					pr[r][1] = *reinterpret_cast<const float4*>(P + 4);
				}
				for (int k = first + lane * 4; k <= i0; k += 128) {
					float4 pk[4][2], v[4];
#pragma unroll
					for (int c = 0; c < 4; c++) {
						const float* P = PT + (k + c) * kPanelStride;
						pk[c][0] = *reinterpret_cast<const float4*>(P);
						pk[c][1] = *reinterpret_cast<const float4*>(P + 4);
					}
#pragma unroll
					for (int r = 0; r < 4; r++) v[r] = *reinterpret_cast<const float4*>(Lrow[r] + k);
#pragma unroll
					for (int r = 0; r < 4; r++) {
#pragma unroll
						for (int h = 0; h < 2; h++) {
							const float4 a = pr[r][h];
							v[r].x = fmaf(-a.x, pk[0][h].x, v[r].x); v[r].y = fmaf(-a.x, pk[1][h].x, v[r].y);
							v[r].z = fmaf(-a.x, pk[2][h].x, v[r].z); v[r].w = fmaf(-a.x, pk[3][h].x, v[r].w);
							v[r].x = fmaf(-a.y, pk[0][h].y, v[r].x); v[r].y = fmaf(-a.y, pk[1][h].y, v[r].y);
							v[r].z = fmaf(-a.y, pk[2][h].y, v[r].z); v[r].w = fmaf(-a.y, pk[3][h].y, v[r].w);
							v[r].x = fmaf(-a.z, pk[0][h].z, v[r].x); v[r].y = fmaf(-a.z, pk[1][h].z, v[r].y);
							v[r].z = fmaf(-a.z, pk[2][h].z, v[r].z); v[r].w = fmaf(-a.z, pk[3][h].z, v[r].w);
							v[r].x = fmaf(-a.w, pk[0][h].w, v[r].x); v[r].y = fmaf(-a.w, pk[1][h].w, v[r].y);
							v[r].z = fmaf(-a.w, pk[2][h].w, v[r].z); v[r].w = fmaf(-a.w, pk[3][h].w, v[r].w);
						}
					}
#pragma unroll
					for (int r = 0; r < 4; r++) *reinterpret_cast<float4*>(Lrow[r] + k) = v[r];
				}
			}
			__syncthreads();
		}
		// ---- blocked back substitution L^T x = y, y = row n of L (forward substitution happened in the factorisation)
		for (int k = tid; k < n; k += T) ybuf[k] = L[aligned_row_start(n) + k];
		__syncthreads();
		int p = 0;
		for (int j = ((n - 1) / kPanel) * kPanel; j >= 0; j -= kPanel) {
			const int nb = min(kPanel, n - j);
			const float* src = ybuf + p * prow;
			float* dst = ybuf + (p ^ 1) * prow;
			for (int k = tid; k < j + nb; k += T) {
				float xb[kPanel];
#pragma unroll
				for (int c = kPanel - 1; c >= 0; c--) {
					xb[c] = 0.f;
					if (c < nb) {
						float v = src[j + c];
#pragma unroll
						for (int r = kPanel - 1; r > c; r--)
							if (r < nb) v = fmaf(-L[aligned_row_start(j + r) + j + c], xb[r], v);
						xb[c] = v * inv_diag[j + c];
					}
				}
				if (k < j) {
					float v = src[k];
#pragma unroll
					for (int c = 0; c < kPanel; c++)
						if (c < nb) v = fmaf(-L[aligned_row_start(j + c) + k], xb[c], v);
					dst[k] = v;
				} else {
#pragma unroll
					for (int c = 0; c < kPanel; c++)
						if (k == j + c) xs[k] = xb[c];
				}
			}
			__syncthreads();
			p ^= 1;
		}
		for (int k = tid; k < n; k += T) x[sys.arrow_base * 6 + k] = xs[k];
	}
	// stem back-substitution: x_i = y_i - sum_e X_e x_C[col(e)]
	for (int idx = tid; idx < sys.arrow_base * 6; idx += T) {
		const int i = idx / 6, r = idx - i * 6;
		float acc = sys.stem_y[idx];
		for (int e = sys.row_offsets[i]; e < sys.row_offsets[i + 1]; e++) {
			const int w = sys.row_blocks[e];
			const float* X = sys.wing_x + (int64_t) w * 36 + r * 6;
			const float* xv = xs + (sys.wing_columns[w] - sys.arrow_base) * 6;
#pragma unroll
			for (int k = 0; k < 6; k++) acc = fmaf(-X[k], xv[k], acc);
		}
		x[idx] = acc;
	}
	if (rotations == nullptr) return;
	__syncthreads();
	__threadfence_block();
	for (int node = tid; node < sys.node_count; node += T) apply_node_update(x + node * 6, rotations + node * 9, translations + node * 3);
}

size_t corner_scratch_bytes(int corner_blocks) {
	const int64_t n = (int64_t) corner_blocks * 6;
	return (size_t) (((n + 1) * (n + 2) / 2 + (n + 1)) * sizeof(float));
}
bool corner_needs_global_scratch(int corner_blocks) { return blocked_corner_floats(corner_blocks * 6) * sizeof(float) > 227 * 1024; }

void launch_stem_eliminate(const ArrowheadSystem& sys, cudaStream_t s) {
	if (sys.node_count > 0) stem_eliminate_kernel<<<(sys.node_count + 3) / 4, 128, 0, s>>>(sys);
	if (sys.corner_upper_count > 0) corner_upper_scatter_kernel<<<blocks_for((int64_t) sys.corner_upper_count * 36, 256), 256, 0, s>>>(sys);
}

int launch_corner_solve(const ArrowheadSystem& sys, float* x, float* node_rotations, float* node_translations, bool clear_corner,
                        float* scratch_global, cudaStream_t s) {
	const int n = (sys.node_count - sys.arrow_base) * 6;
	constexpr size_t kMaxShared = 227 * 1024;
	const size_t blocked_bytes = blocked_corner_floats(n) * sizeof(float);
	if (blocked_bytes <= kMaxShared) {
		static bool blocked_attribute_set = false;
		if (!blocked_attribute_set) {
			if (cudaFuncSetAttribute(corner_solve_blocked_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kMaxShared) != cudaSuccess) return kCudaError;
			blocked_attribute_set = true;
		}
		corner_solve_blocked_kernel<<<1, kCornerThreads, blocked_bytes, s>>>(sys, x, node_rotations, node_translations, clear_corner ? 1 : 0);
		return kOk;
	}
	// corners too large for one SM's shared memory: unblocked kernel on a global scratch
	if (scratch_global == nullptr) return kUnsupported;
	corner_solve_kernel<false><<<1, 1024, 0, s>>>(sys, x, node_rotations, node_translations, clear_corner ? 1 : 0, scratch_global);
	return kOk;
}

}  // namespace nnrt_b200
