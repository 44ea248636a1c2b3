#include "solver_kernels.cuh"
#include "corner_finish.cuh"

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
		for (int k = lane; k < 36; k += 32) {
			const int64_t slot = schur_slot(sys, n, c * 6 + k / 6, c * 6 + k % 6);
			if (slot >= 0) atomicAdd(sys.schur + slot, D[k]);
		}
		if (lane < 6) {
			float* dst = sys.schur_tile_packed ? sys.schur + corner_tile_index(n, n, c * 6 + lane) : sys.corner_rhs + c * 6 + lane;
			atomicAdd(dst, sys.rhs[row * 6 + lane]);
		}
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
		const int64_t slot = schur_slot(sys, n, c1 * 6 + r, c2 * 6 + c);
		if (slot >= 0) atomicAdd(sys.schur + slot, -acc);
	}
	// b_C[c] -= X_e^T b_row
	for (int idx = lane; idx < deg * 6; idx += 32) {
		const int w = sys.row_blocks[e0 + idx / 6], c = idx % 6;
		const float* X = sys.wing_x + (int64_t) w * 36;
		float acc = 0.f;
#pragma unroll
		for (int k = 0; k < 6; k++) acc += X[k * 6 + c] * s_b[wib][k];
		const int col = (sys.wing_columns[w] - sys.arrow_base) * 6 + c;
		atomicAdd(sys.schur_tile_packed ? sys.schur + corner_tile_index(n, n, col) : sys.corner_rhs + col, -acc);
	}
}

__global__ void __launch_bounds__(256) corner_upper_scatter_kernel(ArrowheadSystem sys) {
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= sys.corner_upper_count * 36) return;
	const int u = idx / 36, r = (idx % 36) / 6, c = idx % 6;
	const int n = (sys.node_count - sys.arrow_base) * 6;
	const int bi = sys.corner_upper_coordinates[u * 2] - sys.arrow_base, bj = sys.corner_upper_coordinates[u * 2 + 1] - sys.arrow_base;
	const float v = sys.corner_upper_blocks[idx];
	const int64_t upper = schur_slot(sys, n, bi * 6 + r, bj * 6 + c), lower = schur_slot(sys, n, bj * 6 + c, bi * 6 + r);
	if (upper >= 0) atomicAdd(sys.schur + upper, v);
	if (lower >= 0) atomicAdd(sys.schur + lower, v);
}

// ---- blocked single-CTA corner solve ------------------------------------------------------------------------------------
// Right-looking Cholesky in shared memory, panels of kPanel = 8 columns, the right-hand side carried as row n (forward
// substitution for free). Storage: lower triangle, rows padded to a multiple of 4 columns so every row starts 16-byte aligned
// (row i of 4-row group g = i / 4 has width 4 (g + 1)); the padding lies in the strict upper triangle and is never read as data.
//   panel phase   : one thread per row i >= j. Every row thread factors the 8 x 8 diagonal block redundantly in registers
//                   (one rsqrt per pivot), solves its own row strip against it and publishes the strip to the transposed panel
//                   buffer PT[c][i] (128-bit reads of 4 consecutive rows are contiguous across lanes: conflict-free).
//   trailing phase: all warps; a warp owns a 4-row group, every lane a 4 x 4 tile (128-bit loads/stores of L, 128 FMAs per
//                   16 shared-memory instructions), rank-8 update.
//   back substitution: one warp, right-hand side in registers (lane owns y[lane + 32 m]), no block-wide barriers.
constexpr int kPanel = 8;
constexpr int kMaxRhsPerLane = 10;  // n <= 320

__host__ __device__ __forceinline__ int aligned_row_start(int i) {
	const int g = i >> 2;
	return 8 * g * (g + 1) + (i - 4 * g) * 4 * (g + 1);
}
__host__ __device__ inline int padded_rows(int n) { return (n + 1 + 3) & ~3; }

__host__ __device__ inline size_t blocked_corner_floats(int n) {
	const int pr = padded_rows(n);
	return (size_t) aligned_row_start(pr) + (size_t) kPanel * pr + 2 * (size_t) pr + 64;
}

#ifdef NNRT_PHASE_TIMING
#define NNRT_STAMP(slot) do { if (threadIdx.x == 0) { const long long now = clock64(); g_phase_clock[slot] = now - t_phase; t_phase = now; } } while (0)
#define NNRT_ACC(var) do { if (threadIdx.x == 0) { const long long now = clock64(); var += now - t_phase; t_phase = now; } } while (0)
#else
#define NNRT_STAMP(slot)
#define NNRT_ACC(var)
#endif

// Cholesky of the (up to) 8x8 diagonal block at (j, j) in registers; d is overwritten with its factor, inv[c] = 1 / L_cc.
template<bool kFull>
__device__ __forceinline__ bool factor_diagonal_block(const float* __restrict__ L, int j, int nb, int rows, float (&d)[kPanel][kPanel],
                                                      float (&inv)[kPanel]) {
#pragma unroll
	for (int r = 0; r < kPanel; r++) {
		const float* Lr = L + aligned_row_start(kFull ? j + r : min(j + r, rows - 1)) + j;
#pragma unroll
		for (int c = 0; c <= r; c++) d[r][c] = (kFull || r < nb) ? Lr[c] : (r == c ? 1.f : 0.f);
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
	return ok;
}

template<bool kFull>
__device__ __forceinline__ void panel_phase(float* __restrict__ L, float* __restrict__ PT, float* __restrict__ inv_diag, int* flags, int j,
                                            int nb, int rows, int prow) {
	for (int i = j + (int) threadIdx.x; i < rows; i += blockDim.x) {
		if (i > j && i < j + nb) continue;  // rows of the diagonal block are written by the thread of row j
		float d[kPanel][kPanel], inv[kPanel];
		const bool ok = factor_diagonal_block<kFull>(L, j, nb, rows, d, inv);
		if (i == j) {
			if (!ok) atomicExch(flags + 2, 1);
#pragma unroll
			for (int r = 0; r < kPanel; r++) {
				if (kFull || r < nb) {
					inv_diag[j + r] = inv[r];
					float* Lr = L + aligned_row_start(j + r) + j;
#pragma unroll
					for (int c = 0; c <= r; c++) Lr[c] = d[r][c];
				}
			}
		} else {
			float* Li = L + aligned_row_start(i) + j;
			float a[kPanel];
			if (kFull) {
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
			if (kFull) {
				*reinterpret_cast<float4*>(Li) = make_float4(a[0], a[1], a[2], a[3]);
				*reinterpret_cast<float4*>(Li + 4) = make_float4(a[4], a[5], a[6], a[7]);
#pragma unroll
				for (int c = 0; c < kPanel; c++) PT[c * prow + i] = a[c];
			} else {
#pragma unroll
				for (int c = 0; c < kPanel; c++)
					if (c < nb) Li[c] = a[c];
			}
		}
	}
}

// rank-8 update of rows [first, rows) x columns [first, row]: 4-row groups per warp, 4 x 4 tile per lane
__device__ __forceinline__ void trailing_phase(float* __restrict__ L, const float* __restrict__ PT, int first, int rows, int prow) {
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
	for (int i0 = first + warp * 4; i0 < rows; i0 += warps * 4) {
		float4 pr[kPanel];
		float* Lrow[4];
#pragma unroll
		for (int c = 0; c < kPanel; c++) pr[c] = *reinterpret_cast<const float4*>(PT + c * prow + i0);
#pragma unroll
		for (int r = 0; r < 4; r++) Lrow[r] = L + aligned_row_start(i0 + r);
		for (int k = first + lane * 4; k <= i0; k += 128) {
			float4 pk[kPanel], v[4];
#pragma unroll
			for (int c = 0; c < kPanel; c++) pk[c] = *reinterpret_cast<const float4*>(PT + c * prow + k);
#pragma unroll
			for (int r = 0; r < 4; r++) v[r] = *reinterpret_cast<const float4*>(Lrow[r] + k);
#pragma unroll
			for (int c = 0; c < kPanel; c++) {
				const float4 a = pr[c], b = pk[c];
				v[0].x = fmaf(-a.x, b.x, v[0].x); v[0].y = fmaf(-a.x, b.y, v[0].y); v[0].z = fmaf(-a.x, b.z, v[0].z); v[0].w = fmaf(-a.x, b.w, v[0].w);
				v[1].x = fmaf(-a.y, b.x, v[1].x); v[1].y = fmaf(-a.y, b.y, v[1].y); v[1].z = fmaf(-a.y, b.z, v[1].z); v[1].w = fmaf(-a.y, b.w, v[1].w);
				v[2].x = fmaf(-a.z, b.x, v[2].x); v[2].y = fmaf(-a.z, b.y, v[2].y); v[2].z = fmaf(-a.z, b.z, v[2].z); v[2].w = fmaf(-a.z, b.w, v[2].w);
				v[3].x = fmaf(-a.w, b.x, v[3].x); v[3].y = fmaf(-a.w, b.y, v[3].y); v[3].z = fmaf(-a.w, b.z, v[3].z); v[3].w = fmaf(-a.w, b.w, v[3].w);
			}
#pragma unroll
			for (int r = 0; r < 4; r++) *reinterpret_cast<float4*>(Lrow[r] + k) = v[r];
		}
	}
}

// L^T x = y by one warp: lane owns y[lane + 32 m] in registers; blocks of 8 from the bottom.
__device__ __forceinline__ void back_substitution_warp(const float* __restrict__ L, const float* __restrict__ inv_diag, float* __restrict__ ysh,
                                                       float* __restrict__ xs, int n) {
	const int lane = threadIdx.x & 31;
	float y[kMaxRhsPerLane];
	const float* rhs_row = L + aligned_row_start(n);
#pragma unroll
	for (int m = 0; m < kMaxRhsPerLane; m++) y[m] = (lane + 32 * m < n) ? rhs_row[lane + 32 * m] : 0.f;
	for (int j = ((n - 1) / kPanel) * kPanel; j >= 0; j -= kPanel) {
		const int nb = min(kPanel, n - j);
		// publish y[j .. j+nb): owners are lanes (j & 31) .. +7 of register slot j >> 5
		const int slot = j >> 5, off = lane - (j & 31);
#pragma unroll
		for (int m = 0; m < kMaxRhsPerLane; m++)
			if (m == slot && off >= 0 && off < kPanel) ysh[off] = y[m];
		__syncwarp();
		float xb[kPanel];
#pragma unroll
		for (int c = kPanel - 1; c >= 0; c--) {
			xb[c] = 0.f;
			if (c < nb) {
				float v = ysh[c];
#pragma unroll
				for (int r = kPanel - 1; r > c; r--)
					if (r < nb) v = fmaf(-L[aligned_row_start(j + r) + j + c], xb[r], v);
				xb[c] = v * inv_diag[j + c];
			}
		}
		__syncwarp();
#pragma unroll
		for (int c = 0; c < kPanel; c++)
			if (lane == c && c < nb) xs[j + c] = xb[c];
		// y[k] -= sum_c L[j+c][k] xb[c] for k < j
#pragma unroll
		for (int m = 0; m < kMaxRhsPerLane; m++) {
			const int k = lane + 32 * m;
			if (k < j) {
				float v = y[m];
#pragma unroll
				for (int c = 0; c < kPanel; c++)
					if (c < nb) v = fmaf(-L[aligned_row_start(j + c) + k], xb[c], v);
				y[m] = v;
			}
		}
	}
}

__global__ void __launch_bounds__(kCornerThreads) corner_solve_blocked_kernel(ArrowheadSystem sys, float* __restrict__ x,
                                                                              float* __restrict__ rotations, float* __restrict__ translations,
                                                                              int clear_corner) {
	extern __shared__ __align__(16) float smem[];
	const int tid = threadIdx.x, T = blockDim.x;
	const int nc = sys.node_count - sys.arrow_base, n = nc * 6;
	const int rows = n + 1;                                   // + rhs row
	const int prow = padded_rows(n);
	float* L = smem;                                          // aligned-packed lower triangle, rows 0..prow-1
	float* PT = L + aligned_row_start(prow);                  // [kPanel][prow]
	float* inv_diag = PT + kPanel * prow;                     // [prow]
	float* xs = inv_diag + prow;                              // [prow] solution of the corner
	float* ysh = xs + prow;                                   // [8] back-substitution exchange
#ifdef NNRT_PHASE_TIMING
	long long t_phase = clock64(), acc_panel = 0, acc_trail = 0;
#endif
	if (n > 0) {
		// zero the padding rows/columns once (finite values everywhere the tiles may touch)
		for (int q = aligned_row_start(n) + tid; q < aligned_row_start(prow); q += T) L[q] = 0.f;
		for (int q = tid; q < kPanel * prow; q += T) PT[q] = 0.f;
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
		NNRT_STAMP(0);
		for (int j = 0; j < n; j += kPanel) {
			const int nb = min(kPanel, n - j);
			if (nb == kPanel) panel_phase<true>(L, PT, inv_diag, sys.flags, j, nb, rows, prow);
			else panel_phase<false>(L, PT, inv_diag, sys.flags, j, nb, rows, prow);
			__syncthreads();
			NNRT_ACC(acc_panel);
			if (nb < kPanel) break;  // a partial last panel leaves only the rhs row, which has no columns >= n to update
			trailing_phase(L, PT, j + kPanel, rows, prow);
			__syncthreads();
			NNRT_ACC(acc_trail);
		}
#ifdef NNRT_PHASE_TIMING
		if (tid == 0) { g_phase_clock[1] = acc_panel; g_phase_clock[2] = acc_trail; }
#endif
		if (tid < 32) back_substitution_warp(L, inv_diag, ysh, xs, n);
		__syncthreads();
		for (int k = tid; k < n; k += T) x[sys.arrow_base * 6 + k] = xs[k];
		NNRT_STAMP(3);
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
	NNRT_STAMP(4);
	__threadfence_block();
	for (int node = tid; node < sys.node_count; node += T) apply_node_update(x + node * 6, rotations + node * 9, translations + node * 3, sys.update_mode);
#ifdef NNRT_PHASE_TIMING
	__syncthreads();
	NNRT_STAMP(5);
#endif
}

// ---- register-tiled single-CTA corner solve (n + 1 <= 256) ---------------------------------------------------------------
// The trailing matrix never round-trips through shared memory: thread t owns ONE 8 x 8 tile (bi, bj), bi >= bj, of the
// augmented lower triangle (rhs = row n) in 64 registers for the whole factorisation. Tiles are handed out by descending
// block column, so at panel p the live tiles are exactly threads [0, live(p)) and finished warps fall through to the barriers.
//   panel p, step A: one thread per ROW of block column p (up to 8 warps) forward-substitutes its row, read from the raw strip
//                    RAW that the column's tile owners published at the end of the previous step B, against the factored
//                    diagonal tile; the row goes to PT[c][row] and to the packed factor Lf (read by the back substitution).
//                    The diagonal tile itself is factored one panel AHEAD by warp 0 (which owns the block-column-0 tiles
//                    only and is otherwise idle), concurrently with step B: see lookahead_factor.
//   panel p, step B: every live tile with bj > p does its rank-8 update from PT: 32 LDS.128 per 512 FMAs.
// Shared-memory traffic per panel is the panel strip only; the smem-resident variant above moves every trailing element twice.
__host__ __device__ inline size_t regtile_corner_floats(int n) {
	const int tr = (n + 1 + kTile - 1) / kTile, prow = tr * kTile;
	return (size_t) aligned_row_start(prow) + (size_t) kTile * prow /*PT*/ + (size_t) tr * 72 /*raw strip*/ + 128 + 144 /*look-ahead tiles*/ + (size_t) prow + 32 /*xs*/ +
	       (size_t) ((n + 31) / 32) * 1024 /*Dinv*/ + (size_t) prow /*inv_diag*/ + 64;
}
__host__ __device__ inline bool regtile_corner_fits(int n) {
	const int tr = (n + 1 + kTile - 1) / kTile, tc = (n + kTile - 1) / kTile;
	// tiles (bi >= bj): sum over bj of (tr - bj); block column 0 takes warp 0, the rest start at thread 32
	const int tiles = tc * tr - tc * (tc - 1) / 2;
	return n > 0 && n + 1 <= kRegTileMaxRows && 32 + tiles - tr <= kCornerThreads && regtile_corner_floats(n) * sizeof(float) <= 227 * 1024;
}

// Look-ahead factorisation of one 8 x 8 diagonal tile by a single warp, run while the other warps do the trailing update of the
// previous panel. Lane l holds entries (l / 8, l % 8) and (l / 8 + 4, l % 8); `current` is the tile after all updates but the last,
// `strip` (PT + 8 q, or nullptr for the first tile) supplies that last rank-8 update with the owner's exact operation order.
// Writes the Cholesky factor (strict lower triangle) to fac[0..63], the reciprocal diagonal to fac[64..71] and inv_diag.
__device__ __forceinline__ void lookahead_factor(const float* current, const float* strip, int prow, int nb, float* fac,
                                                 float* inv_diag, int* flags) {
	const int lane = threadIdx.x & 31, r = lane >> 3, c = lane & 7;
	float d0 = current[r * kTile + c], d1 = current[(r + 4) * kTile + c];
	if (strip != nullptr) {
#pragma unroll
		for (int k = 0; k < kTile; k++) {
			const float* col = strip + k * prow;
			const float pc = col[c];
			d0 = fmaf(-col[r], pc, d0);
			d1 = fmaf(-col[r + 4], pc, d1);
		}
	}
	// exchange through fac so that every lane holds the whole updated tile, then factor it redundantly in registers: the
	// dependent chain per column is FMA -> MUFU -> FMUL -> FMA, no shuffles (a lane-parallel factor was 2 x slower)
	fac[r * kTile + c] = d0;
	fac[(r + 4) * kTile + c] = d1;
	__syncwarp();
	float d[kTile][kTile], inv[kTile];
	{
		const float4* F = reinterpret_cast<const float4*>(fac);
#pragma unroll
		for (int i = 0; i < kTile; i++) {
			const float4 lo = F[i * 2], hi = (i >= 4) ? F[i * 2 + 1] : make_float4(0.f, 0.f, 0.f, 0.f);
			const float row[kTile] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
			for (int k = 0; k <= i; k++) d[i][k] = (i < nb) ? row[k] : (i == k ? 1.f : 0.f);   // identity below a partial last block
		}
	}
	__syncwarp();
	bool ok = true;
#pragma unroll
	for (int k = 0; k < kTile; k++) {
		const float pivot = d[k][k];
		if (!(pivot > 0.f)) ok = false;
		asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(inv[k]) : "f"(pivot));
#pragma unroll
		for (int i = k + 1; i < kTile; i++) d[i][k] *= inv[k];
#pragma unroll
		for (int i = k + 1; i < kTile; i++)
#pragma unroll
			for (int m = k + 1; m <= i; m++) d[i][m] = fmaf(-d[i][k], d[m][k], d[i][m]);
	}
	if (lane == 0) {
		float4* F = reinterpret_cast<float4*>(fac);
#pragma unroll
		for (int i = 1; i < kTile; i++) {
			F[i * 2] = make_float4(d[i][0], i > 1 ? d[i][1] : 0.f, i > 2 ? d[i][2] : 0.f, i > 3 ? d[i][3] : 0.f);
			if (i > 4) F[i * 2 + 1] = make_float4(d[i][4], i > 5 ? d[i][5] : 0.f, i > 6 ? d[i][6] : 0.f, 0.f);
		}
		F[16] = make_float4(inv[0], inv[1], inv[2], inv[3]);
		F[17] = make_float4(inv[4], inv[5], inv[6], inv[7]);
#pragma unroll
		for (int i = 0; i < kTile; i++)
			if (i < nb) inv_diag[i] = inv[i];
		if (!ok) atomicExch(flags + 2, 1);
	}
}

// kReduced: the fitter's deterministic path (ReducedSystem, solver_kernels.cuh): the matrix is the 3 Nc x 3 Nc system over the corner
// translations, gathered here in a fixed order from the stem kernel's per-node products instead of being read from an
// atomically accumulated buffer.
template<bool kReduced>
__global__ void __launch_bounds__(kCornerThreads) corner_solve_regtile_kernel(ArrowheadSystem sys, ReducedSystem red, float* __restrict__ x,
                                                                              float* __restrict__ rotations, float* __restrict__ translations,
                                                                              int clear_corner) {
	extern __shared__ __align__(16) float smem[];
	const int tid = threadIdx.x, T = blockDim.x;
	const int nc = kReduced ? red.node_count - red.arrow_base : sys.node_count - sys.arrow_base, n = nc * (kReduced ? 3 : 6);
	int* const flags = kReduced ? red.flags : sys.flags;
	if (kReduced && n == 0) {  // no corner (edge-free warp field): the stem kernel already solved every block
		reduced_back_substitution(red, smem, x, tid, T);
		if (rotations == nullptr) return;
		__syncthreads();
		__threadfence_block();
		for (int node = tid; node < red.node_count; node += T) apply_node_update(x + node * 6, rotations + node * 9, translations + node * 3, red.update_mode);
		return;
	}
	const int tr = (n + 1 + kTile - 1) / kTile, tc = (n + kTile - 1) / kTile, prow = tr * kTile;
	float* Lf = smem;                                         // final factor, aligned-packed rows (back substitution)
	float* PT = Lf + aligned_row_start(prow);                 // [8][prow] current panel strip, transposed
	const int raw_plane = tr * 9;                             // float4 slots per column half; row r of the strip sits at r + r / 8
	float4* RAW = reinterpret_cast<float4*>(PT + kTile * prow);  // [2][tr * 9] raw (updated, unfactored) strip of the next panel
	float* DN = PT + kTile * prow + 2 * raw_plane * 4;        // [2][64] diagonal tiles q = p + 1, p + 2 before their last update
	float* FAC = DN + 128;                                    // [2][72] factored diagonal tile + reciprocal diagonal, by panel parity
	float* xs = FAC + 144;                                    // [prow + 32] corner solution
	float* Dinv = xs + prow + 32;                             // [groups][32][32] inverses of the 32 x 32 diagonal blocks of L
	float* inv_diag = Dinv + ((n + 31) / 32) * 1024;          // [prow] 1 / L_jj
#ifdef NNRT_PHASE_TIMING
	long long t_phase = clock64(), acc_panel = 0, acc_trail = 0;
#endif
	// ---- tile assignment: block columns in descending order, rows ascending inside a column
	int bi = -1, bj = -1;
	if (tid < 32) {
		if (tid < tr) { bj = 0; bi = tid; }
	} else {
		int offset = 32;
		for (int b = tc - 1; b >= 1; b--) {
			const int count = tr - b;
			if (tid >= offset && tid < offset + count) { bj = b; bi = b + (tid - offset); }
			offset += count;
		}
	}
	const bool has_tile = bj >= 0;
	float a[kTile][kTile];
	// ---- load this thread's tile: the stem kernel accumulated S and the rhs row directly in warp-interleaved tile order, so the
	// tile is 16 fully coalesced 128-bit loads; the buffer is zeroed behind the load for the next iteration's accumulation
	if (has_tile) {
		float4* tile = reinterpret_cast<float4*>(kReduced ? red.packed : sys.schur) + (tid >> 5) * 512 + (tid & 31);
		float4 v[16];
#pragma unroll
		for (int q = 0; q < 16; q++) v[q] = __ldcg(tile + q * 32);
#pragma unroll
		for (int r = 0; r < kTile; r++) {
			a[r][0] = v[2 * r].x; a[r][1] = v[2 * r].y; a[r][2] = v[2 * r].z; a[r][3] = v[2 * r].w;
			a[r][4] = v[2 * r + 1].x; a[r][5] = v[2 * r + 1].y; a[r][6] = v[2 * r + 1].z; a[r][7] = v[2 * r + 1].w;
		}
		if (!kReduced && clear_corner) {  // the reduced system is rewritten entry by entry each iteration: nothing accumulates
#pragma unroll
			for (int q = 0; q < 16; q++) tile[q * 32] = make_float4(0.f, 0.f, 0.f, 0.f);
		}
	}
	if (has_tile && bj == 0) {
		float4* dst = RAW + bi * 9;
#pragma unroll
		for (int r = 0; r < kTile; r++) {
			dst[r] = make_float4(a[r][0], a[r][1], a[r][2], a[r][3]);
			dst[raw_plane + r] = make_float4(a[r][4], a[r][5], a[r][6], a[r][7]);
		}
	}
	if (has_tile && bi == bj && bj < 2) {
#pragma unroll
		for (int r = 0; r < kTile; r++)
#pragma unroll
			for (int c = 0; c < kTile; c++) DN[bj * 64 + r * kTile + c] = a[r][c];
	}
	const bool lookahead_warp = tid < 32;   // warp 0: oldest warp of the CTA, so the scheduler does not starve the critical chain
	__syncthreads();
	if (lookahead_warp) lookahead_factor(DN, nullptr, prow, min(kTile, n), FAC, inv_diag, flags);
	__syncthreads();
	NNRT_STAMP(0);
	for (int p = 0; p < tc; p++) {
		const int j = p * kTile, nb = min(kTile, n - j);
		if (tid < prow - j) {
			// ---- step A, one thread per row of block column p: forward-substitute the row against the diagonal factor the
			// look-ahead warp left in FAC. Rows of the diagonal block itself fall out of the same recurrence.
			float d[kTile][kTile], inv[kTile];
			{
				const float4* F = reinterpret_cast<const float4*>(FAC + (p & 1) * 72);
#pragma unroll
				for (int c = 1; c < kTile; c++) {
					const float4 lo = F[c * 2], hi = (c > 4) ? F[c * 2 + 1] : make_float4(0.f, 0.f, 0.f, 0.f);
					const float row[kTile] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
					for (int k = 0; k < c; k++) d[c][k] = row[k];
				}
				const float4 i0 = F[16], i1 = F[17];
				inv[0] = i0.x; inv[1] = i0.y; inv[2] = i0.z; inv[3] = i0.w; inv[4] = i1.x; inv[5] = i1.y; inv[6] = i1.z; inv[7] = i1.w;
			}
			const int lr = tid, pos = lr + (lr >> 3), row = j + lr;
			const float4 v0 = RAW[pos], v1 = RAW[raw_plane + pos];
			float w[kTile] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
			for (int c = 0; c < kTile; c++) {
				float acc = w[c];
#pragma unroll
				for (int k = 0; k < c; k++) acc = fmaf(-w[k], d[c][k], acc);
				w[c] = acc * inv[c];
			}
			float* Lr = Lf + aligned_row_start(row) + j;
			if (lr >= kTile && nb == kTile) {
#pragma unroll
				for (int c = 0; c < kTile; c++) PT[c * prow + row] = w[c];
				*reinterpret_cast<float4*>(Lr) = make_float4(w[0], w[1], w[2], w[3]);
				*reinterpret_cast<float4*>(Lr + 4) = make_float4(w[4], w[5], w[6], w[7]);
			} else {
				// a row of the diagonal block (columns up to the diagonal), or a row below a partial last block (its nb columns)
				const int last = lr < nb ? lr : nb - 1;
#pragma unroll
				for (int c = 0; c < kTile; c++)
					if (c <= last) Lr[c] = w[c];
			}
		}
		__syncthreads();
		NNRT_ACC(acc_panel);
		if (bj > p) {
			// ---- step B: rank-8 update of the register tile
			const int row0 = bi * kTile, col0 = bj * kTile;
			// c-loop kept rolled (two steps per trip, operands of the next step prefetched) so the body stays inside the L0
			// instruction cache; the fully unrolled form is fetch-bound (measured 2.3 cycles / instruction)
			const float* pr_base = PT + row0;
			const float* pk_base = PT + col0;
			float4 r0 = *reinterpret_cast<const float4*>(pr_base), r1 = *reinterpret_cast<const float4*>(pr_base + 4);
			float4 k0 = *reinterpret_cast<const float4*>(pk_base), k1 = *reinterpret_cast<const float4*>(pk_base + 4);
#pragma unroll 1
			for (int c = 0; c < kTile; c += 2) {
				const float* pr1 = pr_base + (c + 1) * prow;
				const float* pk1 = pk_base + (c + 1) * prow;
				const float4 s0 = *reinterpret_cast<const float4*>(pr1), s1 = *reinterpret_cast<const float4*>(pr1 + 4);
				const float4 q0 = *reinterpret_cast<const float4*>(pk1), q1 = *reinterpret_cast<const float4*>(pk1 + 4);
				{
					const float pr[kTile] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
					const float pk[kTile] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
#pragma unroll
					for (int r = 0; r < kTile; r++)
#pragma unroll
						for (int cc = 0; cc < kTile; cc++) a[r][cc] = fmaf(-pr[r], pk[cc], a[r][cc]);
				}
				const int cn = min(c + 2, kTile - 1);
				const float* pr2 = pr_base + cn * prow;
				const float* pk2 = pk_base + cn * prow;
				r0 = *reinterpret_cast<const float4*>(pr2); r1 = *reinterpret_cast<const float4*>(pr2 + 4);
				k0 = *reinterpret_cast<const float4*>(pk2); k1 = *reinterpret_cast<const float4*>(pk2 + 4);
				{
					const float pr[kTile] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
					const float pk[kTile] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
					for (int r = 0; r < kTile; r++)
#pragma unroll
						for (int cc = 0; cc < kTile; cc++) a[r][cc] = fmaf(-pr[r], pk[cc], a[r][cc]);
				}
			}
			if (bi == bj && bj == p + 2) {
				// diagonal tile two panels ahead: the look-ahead warp applies its last update and factors it during the next step B
				float4* dst = reinterpret_cast<float4*>(DN + (p & 1) * 64);
#pragma unroll
				for (int r = 0; r < kTile; r++) {
					dst[2 * r] = make_float4(a[r][0], a[r][1], a[r][2], a[r][3]);
					dst[2 * r + 1] = make_float4(a[r][4], a[r][5], a[r][6], a[r][7]);
				}
			}
			if (bj == p + 1) {
				// raw strip of the next panel: step A reads it back row by row
				float4* dst = RAW + (bi - bj) * 9;
#pragma unroll
				for (int r = 0; r < kTile; r++) {
					dst[r] = make_float4(a[r][0], a[r][1], a[r][2], a[r][3]);
					dst[raw_plane + r] = make_float4(a[r][4], a[r][5], a[r][6], a[r][7]);
				}
			}
		}
#ifdef NNRT_PHASE_TIMING
		const long long tl0 = clock64();
#endif
		if (lookahead_warp && p + 1 < tc)
			lookahead_factor(DN + ((p + 1) & 1) * 64, PT + (p + 1) * kTile, prow, min(kTile, n - (p + 1) * kTile), FAC + ((p + 1) & 1) * 72,
			                 inv_diag + (p + 1) * kTile, flags);
#ifdef NNRT_PHASE_TIMING
		if (lookahead_warp && (tid & 31) == 0 && (p == 0 || p == 12 || p == 24)) g_phase_clock[13 + p / 12] = clock64() - tl0 + (long long) (FAC[((p + 1) & 1) * 72] == 12345.f);
#endif
		__syncthreads();
		NNRT_ACC(acc_trail);
#ifdef NNRT_PHASE_TIMING
		if (tid == 64) g_warp_abs[p] = clock64() + (long long) (((volatile float*) PT)[0] == 12345.f);
#endif
	}
#ifdef NNRT_PHASE_TIMING
	if (tid == 0) { g_phase_clock[1] = acc_panel; g_phase_clock[2] = acc_trail; }
#endif
	// ---- back substitution L^T x = y (y = row n of Lf), rows in groups of 32.
	// (I) every 32 x 32 diagonal block of L is inverted explicitly, one thread per column (256 independent forward substitutions,
	//     all groups at once); (II) per group, top down: the owning warp forms x_g = Linv_gg^T y_g as a 32-term dot product per
	//     lane (y_g by shuffle), one block barrier publishes x_g, every thread above folds it into its y (32 FMAs).
	{
		const int lane = tid & 31, warp = tid >> 5;
		const int groups = (n + 31) >> 5;
		if (warp < groups) {
			const int base = warp << 5, jc = base + lane;
			float z[32];
#pragma unroll
			for (int r = 0; r < 32; r++) {
				float acc = 0.f;
				const bool row_ok = base + r < n;
				if (r > 0) {
					const float* Lr = Lf + aligned_row_start(row_ok ? base + r : base) + base;
#pragma unroll
					for (int k = 0; k < r; k++) acc = fmaf(Lr[k], z[k], acc);
				}
				// rows >= n read padding above: never let those values (possibly NaN bit patterns left in shared memory) through
				const float di = row_ok ? inv_diag[base + r] : 0.f;
				z[r] = !row_ok ? 0.f : (r == lane) ? di : (r > lane ? -acc * di : 0.f);
				Dinv[(warp * 32 + r) * 32 + lane] = (jc < n) ? z[r] : 0.f;
			}
		}
		float y = (tid < n) ? Lf[aligned_row_start(n) + tid] : 0.f;
		__syncthreads();
#ifdef NNRT_PHASE_TIMING
		if (tid == 0) { const long long now = clock64() + (long long) (((volatile float*) Dinv)[0] == 12345.f); g_phase_clock[6] = now - t_phase; t_phase = now; }
#endif
		for (int g = groups - 1; g >= 0; g--) {
			const int base = g << 5;
			if (warp == g) {
				const float* Dg = Dinv + g * 1024 + lane;
				float acc = 0.f;
#pragma unroll
				for (int r = 0; r < 32; r++) acc = fmaf(Dg[r * 32], __shfl_sync(0xFFFFFFFFu, y, r), acc);
				xs[base + lane] = acc;
			}
			__syncthreads();
#ifdef NNRT_PHASE_TIMING
			if (tid == 0) g_warp_abs[32 + g] = clock64() + (long long) (((volatile float*) xs)[base] == 12345.f);
#endif
			if (tid < base) {
				// unconditional, fully unrolled so the 64 loads are issued ahead of the FMA chain: columns >= n of the last group
				// have x = 0 (their Dinv columns are zero) and read the rhs row instead of running past the factor
				float4 xg[8];
#pragma unroll
				for (int q = 0; q < 8; q++) xg[q] = *reinterpret_cast<const float4*>(xs + base + 4 * q);
				float l[32];
#pragma unroll
				for (int c = 0; c < 32; c++) l[c] = Lf[aligned_row_start(min(base + c, n)) + tid];
#pragma unroll
				for (int q = 0; q < 8; q++) {
					y = fmaf(-l[4 * q], xg[q].x, y);
					y = fmaf(-l[4 * q + 1], xg[q].y, y);
					y = fmaf(-l[4 * q + 2], xg[q].z, y);
					y = fmaf(-l[4 * q + 3], xg[q].w, y);
				}
			}
#ifdef NNRT_PHASE_TIMING
			if (tid == 0) { long long now; asm volatile("mov.u64 %0, %%clock64;" : "=l"(now) : "f"(y)); g_warp_abs[48 + g] = now; }
#endif
		}
		__syncthreads();
		if (!kReduced)
			for (int k = tid; k < n; k += T) x[sys.arrow_base * 6 + k] = xs[k];
		NNRT_STAMP(3);
	}
	if (kReduced) {
		reduced_back_substitution(red, xs, x, tid, T);
		if (rotations == nullptr) return;
		__syncthreads();
		NNRT_STAMP(4);
		__threadfence_block();
		for (int node = tid; node < red.node_count; node += T) apply_node_update(x + node * 6, rotations + node * 9, translations + node * 3, red.update_mode);
#ifdef NNRT_PHASE_TIMING
		__syncthreads();
		NNRT_STAMP(5);
#endif
		return;
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
	NNRT_STAMP(4);
	__threadfence_block();
	for (int node = tid; node < sys.node_count; node += T) apply_node_update(x + node * 6, rotations + node * 9, translations + node * 3, sys.update_mode);
#ifdef NNRT_PHASE_TIMING
	__syncthreads();
	NNRT_STAMP(5);
#endif
}

// ---- multi-kernel corner solve for corners beyond one SM (n > 246 unknowns) -----------------------------------------------------
// Right-looking blocked Cholesky of the DENSE Schur buffer, in place (lower triangle), panels of 32 columns, the right-hand side
// carried as row n of the augmented system. Per panel: one launch factors the diagonal block (every CTA redundantly, in shared
// memory) and forward-substitutes 64 rows per CTA; one launch applies the rank-32 update with one CTA per 64 x 64 tile of the
// lower triangle. The matrix (<= tens of MB) stays in L2 between the launches; all launches are captured in the fit's graph.
constexpr int kWide = 32;

__device__ __forceinline__ float* augmented_row(const ArrowheadSystem& sys, int n, int r) {
	return r < n ? sys.schur + (int64_t) r * n : sys.corner_rhs;
}

__global__ void __launch_bounds__(64) corner_wide_panel_kernel(ArrowheadSystem sys, int n, int j) {
	__shared__ float D[kWide][kWide + 1];
	__shared__ float inv_diag[kWide];
	const int tid = threadIdx.x, nb = min(kWide, n - j);
	for (int idx = tid; idx < nb * nb; idx += 64) {
		const int r = idx / nb, c = idx - r * nb;
		if (c <= r) D[r][c] = sys.schur[(int64_t) (j + r) * n + j + c];
	}
	__syncthreads();
	if (tid < 32) {
		// lane r keeps row r of the block in registers; column c of the factor travels by shuffle
		const int lane = tid;
		float d[kWide];
#pragma unroll
		for (int k = 0; k < kWide; k++) d[k] = (lane < nb && k <= lane) ? D[lane][k] : (k == lane ? 1.f : 0.f);
		bool ok = true;
#pragma unroll
		for (int c = 0; c < kWide; c++) {
			const float pivot = __shfl_sync(0xFFFFFFFFu, d[c], c);
			if (c < nb && !(pivot > 0.f)) ok = false;
			const float inv = rsqrtf(pivot);
			if (lane == c) inv_diag[c] = inv;
			const float lc = d[c] * inv;  // L(lane, c) for lane >= c
			d[c] = lc;
#pragma unroll
			for (int k = c + 1; k < kWide; k++) {
				const float lk = __shfl_sync(0xFFFFFFFFu, lc, k);  // L(k, c)
				if (k <= lane) d[k] = fmaf(-lc, lk, d[k]);
			}
		}
		if (!ok && lane == 0 && blockIdx.x == 0) atomicExch(sys.flags + 2, 1);
#pragma unroll
		for (int k = 0; k < kWide; k++)
			if (lane < nb && k <= lane) D[lane][k] = d[k];
	}
	__syncthreads();
	if (blockIdx.x == 0)
		for (int idx = tid; idx < nb * nb; idx += 64) {
			const int r = idx / nb, c = idx - r * nb;
			if (c <= r) sys.schur[(int64_t) (j + r) * n + j + c] = D[r][c];
		}
	const int i = j + nb + blockIdx.x * 64 + tid;
	if (i > n) return;
	float* row = augmented_row(sys, n, i) + j;
	float w[kWide];
#pragma unroll
	for (int c = 0; c < kWide; c++) w[c] = c < nb ? row[c] : 0.f;
#pragma unroll
	for (int c = 0; c < kWide; c++) {
		if (c < nb) {
			float acc = w[c];
#pragma unroll
			for (int k = 0; k < c; k++) acc = fmaf(-w[k], D[c][k], acc);
			w[c] = acc * inv_diag[c];
		}
	}
#pragma unroll
	for (int c = 0; c < kWide; c++)
		if (c < nb) row[c] = w[c];
}

__global__ void __launch_bounds__(256) corner_wide_trailing_kernel(ArrowheadSystem sys, int n, int j) {
	if (blockIdx.x > blockIdx.y) return;  // lower triangle of tiles: blockIdx.y = row tile, blockIdx.x = column tile
	__shared__ float Lr[64][kWide + 1], Lc[64][kWide + 1];
	const int tid = threadIdx.x, nb = min(kWide, n - j), first = j + nb;
	const int row0 = first + blockIdx.y * 64, col0 = first + blockIdx.x * 64;
	for (int idx = tid; idx < 64 * kWide; idx += 256) {
		const int r = idx / kWide, c = idx - r * kWide;
		const int i = row0 + r, k = col0 + r;
		Lr[r][c] = (i <= n && c < nb) ? augmented_row(sys, n, i)[j + c] : 0.f;
		Lc[r][c] = (k < n && c < nb) ? sys.schur[(int64_t) k * n + j + c] : 0.f;
	}
	__syncthreads();
	const int tx = tid & 15, ty = tid >> 4;
	float acc[4][4] = {};
#pragma unroll 8
	for (int c = 0; c < kWide; c++) {
		float a[4], b[4];
#pragma unroll
		for (int q = 0; q < 4; q++) { a[q] = Lr[ty * 4 + q][c]; b[q] = Lc[tx * 4 + q][c]; }
#pragma unroll
		for (int q = 0; q < 4; q++)
#pragma unroll
			for (int m = 0; m < 4; m++) acc[q][m] = fmaf(a[q], b[m], acc[q][m]);
	}
#pragma unroll
	for (int q = 0; q < 4; q++) {
		const int i = row0 + ty * 4 + q;
		if (i > n) continue;
		float* row = augmented_row(sys, n, i);
#pragma unroll
		for (int m = 0; m < 4; m++) {
			const int k = col0 + tx * 4 + m;
			if (k < n && k <= i) row[k] -= acc[q][m];
		}
	}
}

// Back substitution L^T x = y (y = the forward-substituted right-hand side), stem back substitution and node update: one CTA.
template<bool kReduced>
__global__ void __launch_bounds__(1024) corner_wide_finish_kernel(ArrowheadSystem sys, ReducedSystem red, int n, float* __restrict__ x,
                                                                   float* __restrict__ rotations, float* __restrict__ translations) {
	extern __shared__ float y[];  // [n]
	const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31;
	for (int k = tid; k < n; k += T) y[k] = sys.corner_rhs[k];
	__syncthreads();
	for (int jb = ((n - 1) / kWide) * kWide; jb >= 0; jb -= kWide) {
		const int nb = min(kWide, n - jb);
		if (tid < 32) {
			// lane l holds column l of the diagonal block, L(jb + c, jb + l) for c >= l: all loads issued before the dependent chain
			float Lcol[kWide];
#pragma unroll
			for (int c = 0; c < kWide; c++) Lcol[c] = (c < nb && lane <= c) ? sys.schur[(int64_t) (jb + c) * n + jb + lane] : 1.f;
			float yl = lane < nb ? y[jb + lane] : 0.f;
#pragma unroll
			for (int c = kWide - 1; c >= 0; c--) {
				if (c >= nb) continue;
				const float xc = __shfl_sync(0xFFFFFFFFu, yl / Lcol[c], c);  // lane c divides by the diagonal entry L(c, c)
				if (lane < c) yl = fmaf(-Lcol[c], xc, yl);
				if (lane == c) yl = xc;
			}
			if (lane < nb) y[jb + lane] = yl;
		}
		__syncthreads();
		for (int k = tid; k < jb; k += T) {
			float acc = y[k];
			if (nb == kWide) {
				float l[kWide];
#pragma unroll
				for (int c = 0; c < kWide; c++) l[c] = sys.schur[(int64_t) (jb + c) * n + k];
#pragma unroll
				for (int c = 0; c < kWide; c++) acc = fmaf(-l[c], y[jb + c], acc);
			} else {
				for (int c = 0; c < nb; c++) acc = fmaf(-sys.schur[(int64_t) (jb + c) * n + k], y[jb + c], acc);
			}
			y[k] = acc;
		}
		__syncthreads();
	}
	if (kReduced) {
		reduced_back_substitution(red, y, x, tid, T);
		if (rotations == nullptr) return;
		__syncthreads();
		__threadfence_block();
		for (int node = tid; node < red.node_count; node += T) apply_node_update(x + node * 6, rotations + node * 9, translations + node * 3, red.update_mode);
		return;
	}
	for (int k = tid; k < n; k += T) x[sys.arrow_base * 6 + k] = y[k];
	for (int idx = tid; idx < sys.arrow_base * 6; idx += T) {
		const int i = idx / 6, r = idx - i * 6;
		float acc = sys.stem_y[idx];
		for (int e = sys.row_offsets[i]; e < sys.row_offsets[i + 1]; e++) {
			const int w = sys.row_blocks[e];
			const float* X = sys.wing_x + (int64_t) w * 36 + r * 6;
			const float* xv = y + (sys.wing_columns[w] - sys.arrow_base) * 6;
#pragma unroll
			for (int k = 0; k < 6; k++) acc = fmaf(-X[k], xv[k], acc);
		}
		x[idx] = acc;
	}
	if (rotations == nullptr) return;
	__syncthreads();
	__threadfence_block();
	for (int node = tid; node < sys.node_count; node += T) apply_node_update(x + node * 6, rotations + node * 9, translations + node * 3, sys.update_mode);
}

static int launch_corner_solve_wide(const ArrowheadSystem& sys, int n, float* x, float* node_rotations, float* node_translations, bool clear_corner,
                                    cudaStream_t s, const ReducedSystem* red = nullptr) {
	if ((size_t) n * sizeof(float) > 160 * 1024) return kUnsupported;
	// one thread-block-cluster kernel (corner_cluster.cu) where the device can co-schedule it; otherwise two launches per panel
	const int cluster_status = launch_corner_cluster_factor(sys.schur, sys.corner_rhs, n, sys.flags, s);
	if (cluster_status != kOk && cluster_status != kUnsupported) return cluster_status;
	if (cluster_status == kOk) {
		const int rc = launch_corner_cluster_finish(sys, red, n, x, node_rotations, node_translations, s);
		if (rc != kOk) return rc;
		if (!red && clear_corner) {
			if (cudaMemsetAsync(sys.schur, 0, (size_t) n * n * sizeof(float), s) != cudaSuccess) return kCudaError;
			if (cudaMemsetAsync(sys.corner_rhs, 0, (size_t) n * sizeof(float), s) != cudaSuccess) return kCudaError;
		}
		return kOk;
	}
	for (int j = 0; j < n; j += kWide) {
		const int nb = n - j < kWide ? n - j : kWide;
		const int rows_below = n - (j + nb) + 1;  // the right-hand side row is always there
		corner_wide_panel_kernel<<<(rows_below + 63) / 64, 64, 0, s>>>(sys, n, j);
		if (j + nb < n) {
			const int tiles = (rows_below + 63) / 64;
			corner_wide_trailing_kernel<<<dim3(tiles, tiles), 256, 0, s>>>(sys, n, j);
		}
	}
	static SharedMemoryOptIn finish_opt_in, finish_reduced_opt_in;
	if (red) {
		if (finish_reduced_opt_in.ensure(corner_wide_finish_kernel<true>, 160 * 1024) != kOk) return kCudaError;
		corner_wide_finish_kernel<true><<<1, 1024, (size_t) n * sizeof(float), s>>>(sys, *red, n, x, node_rotations, node_translations);
		return kOk;  // the dense buffer is rewritten by the next gather: nothing to clear
	}
	if (finish_opt_in.ensure(corner_wide_finish_kernel<false>, 160 * 1024) != kOk) return kCudaError;
	corner_wide_finish_kernel<false><<<1, 1024, (size_t) n * sizeof(float), s>>>(sys, ReducedSystem{}, n, x, node_rotations, node_translations);
	if (clear_corner) {
		if (cudaMemsetAsync(sys.schur, 0, (size_t) n * n * sizeof(float), s) != cudaSuccess) return kCudaError;
		if (cudaMemsetAsync(sys.corner_rhs, 0, (size_t) n * sizeof(float), s) != cudaSuccess) return kCudaError;
	}
	return kOk;
}

// Gather of the reduced corner system: ONE WARP per listed block pair (then one per corner node for the right-hand side). The lanes
// load the pair's contiguous run of 3 x 3 products (32 at a time), a fixed shuffle tree sums them, the diagonal block is added and
// the nine entries are written where the corner kernel expects them (tile-packed, or the dense buffer of the multi-kernel path).
// Nothing is accumulated atomically: the result does not depend on scheduling.
__global__ void __launch_bounds__(128) reduced_gather_kernel(ReducedSystem red) {
	const int lane = threadIdx.x & 31, item = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int nc = red.node_count - red.arrow_base, n = nc * 3;
	if (item >= red.block_pair_count + nc) return;
	float v[9];
#pragma unroll
	for (int k = 0; k < 9; k++) v[k] = 0.f;
	const bool is_pair = item < red.block_pair_count;
	int c1 = 0, c2 = 0, first, last;
	if (is_pair) {
		const int2 cc = red.block_pairs[item];
		c1 = cc.x; c2 = cc.y; first = red.list_offsets[item]; last = red.list_offsets[item + 1];
	} else {
		c1 = c2 = item - red.block_pair_count; first = red.in_offsets[c1]; last = red.in_offsets[c1 + 1];
	}
	for (int q0 = first; q0 < last; q0 += 32) {
		float w[9];
#pragma unroll
		for (int k = 0; k < 9; k++) w[k] = 0.f;
		if (q0 + lane < last) {
			if (is_pair) {
				const float4* src = reinterpret_cast<const float4*>(red.pair_blocks) + (int64_t) (q0 + lane) * 3;
				const float4 a = __ldcg(src), b = __ldcg(src + 1), c = __ldcg(src + 2);
				w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w; w[8] = c.x;
			} else {
				const float4 a = __ldcg(reinterpret_cast<const float4*>(red.edge_rhs) + (q0 + lane));
				w[0] = a.x; w[1] = a.y; w[2] = a.z;
			}
		}
#pragma unroll
		for (int off = 16; off >= 1; off >>= 1)
#pragma unroll
			for (int k = 0; k < 9; k++)
				if (is_pair || k < 3) w[k] += __shfl_xor_sync(0xFFFFFFFFu, w[k], off);
#pragma unroll
		for (int k = 0; k < 9; k++) v[k] += w[k];
	}
	if (is_pair) {
		if (lane < 9) {
			float total = 0.f;
#pragma unroll
			for (int k = 0; k < 9; k++) total = lane == k ? v[k] : total;
			const float value = (c1 == c2 ? red.corner_c[c1 * 9 + lane] : 0.f) - total;
			const int row = c1 * 3 + lane / 3, col = c2 * 3 + lane % 3;
			if (col <= row) {
				if (red.packed) red.packed[corner_tile_index(n, row, col)] = value;
				else red.dense[(int64_t) row * n + col] = value;
			}
		}
	} else if (lane < 3) {
		const float total = lane == 0 ? v[0] : lane == 1 ? v[1] : v[2];
		const float value = red.corner_r[c1 * 3 + lane] - total;
		if (red.packed) red.packed[corner_tile_index(n, n, c1 * 3 + lane)] = value;
		else red.dense_rhs[c1 * 3 + lane] = value;
	}
}

bool reduced_corner_fits_one_cta(int corner_blocks) { return corner_blocks == 0 || regtile_corner_fits(corner_blocks * 3); }
size_t reduced_packed_floats(int corner_blocks) { return corner_blocks == 0 ? 64 : (size_t) corner_tile_floats(corner_blocks * 3) + 64; }

int launch_reduced_corner_solve(const ReducedSystem& red, float* x, float* node_rotations, float* node_translations, cudaStream_t s) {
	const int nc = red.node_count - red.arrow_base, n = nc * 3;
	const int items = red.block_pair_count + nc;
	if (reduced_corner_fits_one_cta(nc)) {
		if (items > 0) reduced_gather_kernel<<<(items + 3) / 4, 128, 0, s>>>(red);
		static SharedMemoryOptIn opt_in;
		if (opt_in.ensure(corner_solve_regtile_kernel<true>, 227 * 1024) != kOk) return kCudaError;
		corner_solve_regtile_kernel<true><<<1, kCornerThreads, regtile_corner_floats(n > 0 ? n : 1) * sizeof(float), s>>>(ArrowheadSystem{}, red, x, node_rotations,
		                                                                                                             node_translations, 0);
		return kOk;
	}
	if (red.dense == nullptr) return kInvalidArgument;
	// the multi-kernel factorisation works in place: the dense buffer is rebuilt (zeros included) every iteration
	if (cudaMemsetAsync(red.dense, 0, (size_t) n * n * sizeof(float), s) != cudaSuccess) return kCudaError;
	reduced_gather_kernel<<<(items + 3) / 4, 128, 0, s>>>(red);
	ArrowheadSystem dense{};
	dense.schur = red.dense; dense.corner_rhs = red.dense_rhs; dense.flags = red.flags;
	return launch_corner_solve_wide(dense, n, x, node_rotations, node_translations, false, s, &red);
}

bool corner_uses_tile_packed_schur(int corner_blocks) { return regtile_corner_fits(corner_blocks * 6); }
size_t corner_schur_floats(int corner_blocks) {
	const size_t n = (size_t) corner_blocks * 6, dense = n * n, packed = (size_t) corner_tile_floats((int) n);
	return (dense > packed ? dense : packed) + 64;
}

void launch_stem_eliminate(const ArrowheadSystem& sys, cudaStream_t s) {
	if (sys.node_count > 0) stem_eliminate_kernel<<<(sys.node_count + 3) / 4, 128, 0, s>>>(sys);
	if (sys.corner_upper_count > 0) corner_upper_scatter_kernel<<<blocks_for((int64_t) sys.corner_upper_count * 36, 256), 256, 0, s>>>(sys);
}

int launch_corner_solve(const ArrowheadSystem& sys, float* x, float* node_rotations, float* node_translations, bool clear_corner,
                        cudaStream_t s) {
	const int n = (sys.node_count - sys.arrow_base) * 6;
	constexpr size_t kMaxShared = 227 * 1024;
	const size_t blocked_bytes = blocked_corner_floats(n) * sizeof(float);
	if (sys.schur_tile_packed) {
		if (!regtile_corner_fits(n)) return kInvalidArgument;
		static SharedMemoryOptIn regtile_opt_in;
		if (regtile_opt_in.ensure(corner_solve_regtile_kernel<false>, kMaxShared) != kOk) return kCudaError;
		corner_solve_regtile_kernel<false><<<1, kCornerThreads, regtile_corner_floats(n) * sizeof(float), s>>>(sys, ReducedSystem{}, x, node_rotations, node_translations, clear_corner ? 1 : 0);
		return kOk;
	}
	if (blocked_bytes <= kMaxShared) {
		static SharedMemoryOptIn blocked_opt_in;
		if (blocked_opt_in.ensure(corner_solve_blocked_kernel, kMaxShared) != kOk) return kCudaError;
		corner_solve_blocked_kernel<<<1, kCornerThreads, blocked_bytes, s>>>(sys, x, node_rotations, node_translations, clear_corner ? 1 : 0);
		return kOk;
	}
	// corners too large for one SM's shared memory: multi-kernel blocked factorisation of the dense Schur buffer, in place
	return launch_corner_solve_wide(sys, n, x, node_rotations, node_translations, clear_corner, s);
}

}  // namespace nnrt_b200
