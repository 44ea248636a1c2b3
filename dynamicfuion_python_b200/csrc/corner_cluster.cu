// Corner systems beyond one SM (n > 255 unknowns): ONE thread-block-cluster kernel factors the dense Schur buffer in place, a second
// single-CTA kernel back-substitutes and applies the node update.
//
// Right-looking blocked Cholesky, panels of 32 columns, the right-hand side carried as row n of the augmented system (forward
// substitution for free). The matrix stays in global memory: it is 3 - 13 MB and lives in the 126 MB L2 for the whole kernel, so the
// cluster is used for what the hardware gives it — co-scheduled CTAs and a ~0.2 us hardware barrier — not for distributed shared
// memory (which would cap n at what 16 SMs hold). Per panel j there are two cluster barriers:
//   A  every CTA takes 64-row chunks of the rows below the diagonal block and multiplies them by inv(L_jj)^T                 | barrier 1
//   B  the CTA that owns panel j + 1 updates ONLY the next 32 x 32 diagonal block, factors it and inverts the factor (the dependency
//      chain of the whole factorisation) while the other CTAs apply the rank-32 update to the 64 x 64 tiles of the trailing matrix | barrier 2
// so the chain factor -> substitute -> update-diagonal -> factor never waits for the bulk of the update.
// Both A and the tile update are small GEMMs and run on the tensor cores: mma.sync m16n8k8 TF32 with the 3-term hi/lo split
// (a_hi b_hi + a_hi b_lo + a_lo b_hi, fp32 accumulate; the dropped a_lo b_lo term is 2^-22 relative) — the one dense contraction of the
// iteration large enough to pay for it (108 MFMA at C3). Measured on the way here (profiles/r2_corner_cluster.md): thread-per-row
// substitution and a single-warp 32 x 32 factorisation are serial instruction chains that run at ~4.5 cycles per instruction and
// dominated the first version (15.8 us per diagonal block, 4.1 us per substitution); fully unrolled variants were bound by
// instruction fetch. Hence: the diagonal block is factored 8 columns at a time by the whole CTA (every thread factors and inverts the
// 8 x 8 sub-block redundantly in registers, the rest is data-parallel), its INVERSE is published in the unused upper triangle of the
// block, and substitutions (here and in the finish kernel) become products with that inverse.
// Data written by one CTA and read by another goes through L2 (st / ld.global.cg); barrier.cluster release / acquire orders it.
#include "solver_kernels.cuh"
#include "corner_finish.cuh"

namespace nnrt_b200 {
namespace {

constexpr int kPanel = 32, kTrail = 64, kClusterThreads = 256;  // threads of one GROUP; the factor kernel runs two groups per CTA
constexpr int kGroups = 2;
constexpr int kLd = kPanel + 4;  // shared rows of 36 floats: 16-byte aligned for 128-bit loads, conflict-free for the fragment reads

#ifdef NNRT_CLUSTER_TIMING  // developer harness (tools/corner_cluster_bench.cu): cycles per phase, per CTA rank
__shared__ long long s_cluster_clock[16];  // accumulated in shared memory (a global read-modify-write per stamp costs 0.4 us itself)
#define NNRT_CLUSTER_STAMP(slot)                                                     \
	do {                                                                             \
		if (threadIdx.x == 0) {                                                      \
			const long long now = clock64();                                         \
			s_cluster_clock[slot] += now - t_phase;                                  \
			t_phase = now;                                                           \
		}                                                                            \
	} while (0)
#else
#define NNRT_CLUSTER_STAMP(slot)
#endif

__device__ __forceinline__ unsigned cluster_rank() {
	unsigned r;
	asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
	return r;
}
__device__ __forceinline__ unsigned cluster_ctas() {
	unsigned r;
	asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
	return r;
}
// barrier of one 256-thread group of the CTA (named barrier 1 + group): the two groups run different tiles at the same time
__device__ __forceinline__ void group_sync(int group) { asm volatile("bar.sync %0, %1;" ::"r"(group + 1), "r"(kClusterThreads) : "memory"); }
// Every cross-CTA read in this kernel bypasses L1 (ld.global.cg), so the L1 invalidation of an acquiring cluster barrier is not needed
// (measured: no difference): the L2 traffic is ordered with explicit fences around a relaxed barrier. Arrive and wait are separate so the
// owner of the next panel can arrive early and keep working.
__device__ __forceinline__ void cluster_arrive() {
	__threadfence();
	asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
	asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
	__threadfence();
}
__device__ __forceinline__ void cluster_barrier() {
	cluster_arrive();
	cluster_wait();
}

// row r of the augmented system: rows < n live in S, row n is the right-hand side (n * n < 2^31: 32-bit offsets)
__device__ __forceinline__ float* augmented(float* S, float* rhs, int n, int r) { return r < n ? S + r * n : rhs; }

__device__ __forceinline__ float dot8(const float4& a0, const float4& a1, const float4& b0, const float4& b1) {
	float acc = a0.x * b0.x;
	acc = fmaf(a0.y, b0.y, acc); acc = fmaf(a0.z, b0.z, acc); acc = fmaf(a0.w, b0.w, acc);
	acc = fmaf(a1.x, b1.x, acc); acc = fmaf(a1.y, b1.y, acc); acc = fmaf(a1.z, b1.z, acc); acc = fmaf(a1.w, b1.w, acc);
	return acc;
}

// In-place Cholesky of the 32 x 32 block D (lower triangle; rows >= nb hold the identity) and the inverse of the factor into Linv
// (lower triangle, zero above), by all 256 threads of the CTA. T is an 8 x 32-column staging area for the inverse's off-diagonal blocks.
__device__ __noinline__ void factor_and_invert_block(float (*D)[kLd], float (*Linv)[kLd], float (*T)[kLd], int nb, int* flags, int tid) {
	for (int idx = tid; idx < kPanel * kLd; idx += kClusterThreads) (&Linv[0][0])[idx] = 0.f;
	bool ok = true;
#pragma unroll 1
	for (int b = 0; b < kPanel; b += 8) {
		group_sync(0);
		// 8 x 8 diagonal sub-block: factored and inverted redundantly by every thread, in registers
		float a[8][8], m[8][8], rinv[8];
#pragma unroll
		for (int i = 0; i < 8; i++)
#pragma unroll
			for (int j = 0; j <= i; j++) a[i][j] = D[b + i][b + j];
#pragma unroll
		for (int j = 0; j < 8; j++) {
			const float pivot = a[j][j];
			if (b + j < nb && !(pivot > 0.f)) ok = false;
			const float inv = rsqrtf(pivot);
			rinv[j] = inv;
			a[j][j] = pivot * inv;
#pragma unroll
			for (int i = j + 1; i < 8; i++) a[i][j] *= inv;
#pragma unroll
			for (int i = j + 1; i < 8; i++)
#pragma unroll
				for (int k = j + 1; k <= i; k++) a[i][k] = fmaf(-a[i][j], a[k][j], a[i][k]);
		}
#pragma unroll
		for (int j = 0; j < 8; j++) {
			m[j][j] = rinv[j];
#pragma unroll
			for (int i = j + 1; i < 8; i++) {
				float acc = 0.f;
#pragma unroll
				for (int k = j; k < i; k++) acc = fmaf(a[i][k], m[k][j], acc);
				m[i][j] = -rinv[i] * acc;
			}
		}
		group_sync(0);  // every thread has read the sub-block
		if (tid < 64) {
			const int i = tid >> 3, j = tid & 7;
			float lv = 0.f, mv = 0.f;
#pragma unroll
			for (int ii = 0; ii < 8; ii++)
#pragma unroll
				for (int jj = 0; jj <= ii; jj++)
					if (ii == i && jj == j) { lv = a[ii][jj]; mv = m[ii][jj]; }
			if (j <= i) { D[b + i][b + j] = lv; Linv[b + i][b + j] = mv; }
		} else if (tid < 64 + kPanel - 8 - b) {
			// rows below the sub-block: x <- x inv(L8)^T
			const int r = b + 8 + tid - 64;
			const float4 x0 = *reinterpret_cast<const float4*>(&D[r][b]), x1 = *reinterpret_cast<const float4*>(&D[r][b + 4]);
			const float x[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
			float y[8];
#pragma unroll
			for (int q = 0; q < 8; q++) {
				float acc = 0.f;
#pragma unroll
				for (int p = 0; p <= q; p++) acc = fmaf(x[p], m[q][p], acc);
				y[q] = acc;
			}
			*reinterpret_cast<float4*>(&D[r][b]) = make_float4(y[0], y[1], y[2], y[3]);
			*reinterpret_cast<float4*>(&D[r][b + 4]) = make_float4(y[4], y[5], y[6], y[7]);
		}
		group_sync(0);
		// rank-8 update of the rest of the block
		const int rest = kPanel - 8 - b;
		for (int e = tid; e < rest * rest; e += kClusterThreads) {
			const int r = b + 8 + e / rest, k = b + 8 + e % rest;
			if (k <= r)
				D[r][k] -= dot8(*reinterpret_cast<const float4*>(&D[r][b]), *reinterpret_cast<const float4*>(&D[r][b + 4]),
				                *reinterpret_cast<const float4*>(&D[k][b]), *reinterpret_cast<const float4*>(&D[k][b + 4]));
		}
	}
	group_sync(0);
	if (!ok && tid == 0) atomicExch(flags + 2, 1);
	// off-diagonal 8 x 8 blocks of the inverse, one block sub-diagonal at a time:
	// Linv(i, j) = -Linv(i, i) * sum_{m = j}^{i - 1} L(i, m) Linv(m, j)
#pragma unroll 1
	for (int d = 1; d < 4; d++) {
		const int pairs = 4 - d;
		if (tid < pairs * 64) {
			const int pair = tid >> 6, e = tid & 63, i = d + pair, j = pair, r = i * 8 + (e >> 3), c = j * 8 + (e & 7);
			float acc = 0.f;
			for (int k = j * 8; k < i * 8; k++) acc = fmaf(D[r][k], Linv[k][c], acc);
			T[r & 7][pair * 8 + (e & 7)] = acc;  // T: [8 rows of the block][pair * 8 + column]
		}
		group_sync(0);
		if (tid < pairs * 64) {
			const int pair = tid >> 6, e = tid & 63, i = d + pair, j = pair, r = i * 8 + (e >> 3), c = j * 8 + (e & 7);
			float acc = 0.f;
#pragma unroll
			for (int p = 0; p < 8; p++) acc = fmaf(Linv[r][i * 8 + p], T[p][pair * 8 + (e & 7)], acc);
			Linv[r][c] = -acc;
		}
		group_sync(0);
	}
}

__device__ __forceinline__ unsigned to_tf32(float x) {
	unsigned r;
	asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
	return r;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
	asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
	             : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
	             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// Operand staging: 64 rows x 32 k of the row operand A(i, k) and of the column operand B(k, j) stored as Bt[j][k], raw fp32 (filled with
// cp.async so the loop that issues the copies stays rolled; the TF32 hi / lo split happens when a fragment is loaded).
struct OperandStage {
	float a[kTrail][kLd], bt[kTrail][kLd];
};
__device__ __forceinline__ void cp_async4(float* shared_destination, const float* source, bool valid) {
	const unsigned destination = (unsigned) __cvta_generic_to_shared(shared_destination);
	const int bytes = valid ? 4 : 0;  // 0: the destination is zero-filled, the source is not read
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(destination), "l"(source), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void split_tf32(float v, unsigned& hi, unsigned& lo) {
	hi = to_tf32(v);
	lo = to_tf32(v - __uint_as_float(hi));
}

// acc(16 x 8 * kBlocks) += A[wr .. wr + 16, 0 .. 32] * Bt[wc .. wc + 8 kBlocks, 0 .. 32]^T for one warp; the k loop stays rolled
// (code that runs once per panel with a cold instruction cache is bound by instruction fetch: keep it short).
// Fragment layouts of mma.m16n8k8 (g = lane / 4, t = lane % 4): a0 (g, t) a1 (g + 8, t) a2 (g, t + 4) a3 (g + 8, t + 4);
// b0 (k = t, n = g) b1 (k = t + 4, n = g); c0 (g, 2t) c1 (g, 2t + 1) c2 (g + 8, 2t) c3 (g + 8, 2t + 1).
template<int kBlocks>
__device__ __forceinline__ void warp_mma_32(const OperandStage& st, int wr, int wc, int g, int t, float (&acc)[kBlocks][4]) {
#pragma unroll 1
	for (int ks = 0; ks < kPanel; ks += 8) {
		unsigned a_hi[4], a_lo[4];
		split_tf32(st.a[wr + g][ks + t], a_hi[0], a_lo[0]);
		split_tf32(st.a[wr + g + 8][ks + t], a_hi[1], a_lo[1]);
		split_tf32(st.a[wr + g][ks + t + 4], a_hi[2], a_lo[2]);
		split_tf32(st.a[wr + g + 8][ks + t + 4], a_hi[3], a_lo[3]);
#pragma unroll
		for (int nbk = 0; nbk < kBlocks; nbk++) {
			unsigned b_hi[2], b_lo[2];
			split_tf32(st.bt[wc + nbk * 8 + g][ks + t], b_hi[0], b_lo[0]);
			split_tf32(st.bt[wc + nbk * 8 + g][ks + t + 4], b_hi[1], b_lo[1]);
			mma_tf32(acc[nbk], a_lo, b_hi);
			mma_tf32(acc[nbk], a_hi, b_lo);
			mma_tf32(acc[nbk], a_hi, b_hi);
		}
	}
}

// One 64 x 64 tile of the trailing matrix: T(i, k) -= sum_c L(i, j0 + c) L(k, j0 + c), rows row0.., columns col0.. (col0 <= row0);
// entries with k > i or beyond the augmented system are left alone, and so is the next diagonal block (rows and columns below
// skip_end; 0 for every tile but the first), which belongs to the panel owner.
__device__ __noinline__ void update_tile(float* __restrict__ S, float* __restrict__ rhs, int n, int j0, int nb, int row0, int col0, int skip_end,
                                         OperandStage& st, int tid, int group) {
	const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
	{
		float va[kTrail / 8], vb[kTrail / 8];  // one k column per lane, 8 rows per pass; all loads in flight before the first store
#pragma unroll
		for (int pass = 0; pass < kTrail / 8; pass++) {
			const int r = pass * 8 + warp, i = row0 + r, k = col0 + r;
			va[pass] = (i <= n && lane < nb) ? __ldcg(augmented(S, rhs, n, i) + j0 + lane) : 0.f;
			vb[pass] = (k < n && lane < nb) ? __ldcg(S + k * n + j0 + lane) : 0.f;
		}
#pragma unroll
		for (int pass = 0; pass < kTrail / 8; pass++) {
			st.a[pass * 8 + warp][lane] = va[pass];
			st.bt[pass * 8 + warp][lane] = vb[pass];
		}
	}
	// 8 warps: warp w owns rows [16 (w % 4), +16) x columns [32 (w / 4), +32) of the tile; the old values are fetched before the math
	const int wr = (warp & 3) * 16, wc = (warp >> 2) * 32;
	float old[4][4];
	unsigned writable = 0;
#pragma unroll
	for (int half = 0; half < 2; half++) {
		const int i = row0 + wr + g + half * 8;
		const int last = i > n ? -1 : min(i, n - 1);  // last column of row i that exists in the lower triangle
		const float* row = augmented(S, rhs, n, min(i, n)) + col0 + wc + 2 * t;
#pragma unroll
		for (int nbk = 0; nbk < 4; nbk++)
#pragma unroll
			for (int e = 0; e < 2; e++) {
				const int k = col0 + wc + nbk * 8 + 2 * t + e, q = half * 2 + e;
				const bool valid = k <= last && !(i < skip_end && k < skip_end);
				old[nbk][q] = valid ? __ldcg(row + nbk * 8 + e) : 0.f;
				writable |= (valid ? 1u : 0u) << (nbk * 4 + q);
			}
	}
	group_sync(group);
	float acc[4][4] = {};
	warp_mma_32<4>(st, wr, wc, g, t, acc);
#pragma unroll
	for (int half = 0; half < 2; half++) {
		const int i = row0 + wr + g + half * 8;
		float* row = augmented(S, rhs, n, min(i, n)) + col0 + wc + 2 * t;
#pragma unroll
		for (int nbk = 0; nbk < 4; nbk++)
#pragma unroll
			for (int e = 0; e < 2; e++) {
				const int q = half * 2 + e;
				if (writable >> (nbk * 4 + q) & 1u) row[nbk * 8 + e] = old[nbk][q] - acc[nbk][q];
			}
	}
	group_sync(group);  // staging is reused by the next tile
}

struct alignas(16) ClusterShared {
	float D[kPanel][kLd];        // owner: the block being factored
	float Linv[kPanel][kLd];     // inverse of the current diagonal factor
	float T[8][kLd];
	OperandStage stage[kGroups];
};

// inv(L_jj) of the factored diagonal block at (j0, j0): strict lower part from the block's upper triangle (where the owner left its
// transpose), diagonal as the reciprocal of the factor's diagonal; identity beyond nb.
__device__ __forceinline__ float load_inverse_entry(const float* __restrict__ S, int n, int j0, int nb, int r, int c) {
	// one branch-free load: for r >= c the address (j0 + c, j0 + r) is the diagonal entry itself (r == c) or the transposed inverse
	// entry above the diagonal (divergent loads would serialise their L2 round trips: measured 1.6 us per call site)
	const bool stored = r < nb && c <= r;
	const float v = stored ? __ldcg(S + (j0 + c) * n + j0 + r) : 0.f;
	const bool diagonal = r == c;
	const float reciprocal = 1.f / (diagonal && stored ? v : 1.f);  // never divides by the zeros of the unused lanes (slow path of the IEEE division)
	return diagonal ? reciprocal : v;
}

// Rows [row_begin, row_end) x columns [j0, j0 + nb): Y = A inv(L_jj)^T, 64 rows per chunk; this caller takes chunks part, part + parts, ...
__device__ __noinline__ void substitute_rows(float* __restrict__ S, float* __restrict__ rhs, int n, int j0, int nb, int row_begin, int row_end, int part,
                                             int parts, ClusterShared& sh, int tid) {
	const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
	const int chunks = (row_end - row_begin + kTrail - 1) / kTrail;
	if (part >= chunks) return;
	// column operand: Bt[j][k] = Linv(j, k) (Y(i, j) = sum_k A(i, k) Linv(j, k))
	{
		float entry[kPanel / 8];  // all four loads in flight before the first is used
#pragma unroll
		for (int pass = 0; pass < kPanel / 8; pass++) entry[pass] = load_inverse_entry(S, n, j0, nb, lane, pass * 8 + warp);  // lane = row: contiguous in memory
#pragma unroll
		for (int pass = 0; pass < kPanel / 8; pass++) sh.stage[0].bt[lane][pass * 8 + warp] = entry[pass];
	}
#pragma unroll 1
	for (int chunk = part; chunk < chunks; chunk += parts) {
		const int row0 = row_begin + chunk * kTrail;
		{
			float va[kTrail / 8];
#pragma unroll
			for (int pass = 0; pass < kTrail / 8; pass++) {
				const int i = row0 + pass * 8 + warp;
				va[pass] = (i < row_end && lane < nb) ? __ldcg(augmented(S, rhs, n, i) + j0 + lane) : 0.f;
			}
#pragma unroll
			for (int pass = 0; pass < kTrail / 8; pass++) sh.stage[0].a[pass * 8 + warp][lane] = va[pass];
		}
		group_sync(0);
		// warp w: rows [16 (w % 4), +16) x columns [16 (w / 4), +16)
		const int wr = (warp & 3) * 16, wc = (warp >> 2) * 16;
		float acc[2][4] = {};
		warp_mma_32<2>(sh.stage[0], wr, wc, g, t, acc);
#pragma unroll
		for (int half = 0; half < 2; half++) {
			const int i = row0 + wr + g + half * 8;
			if (i >= row_end) continue;
			float* row = augmented(S, rhs, n, i) + j0 + wc + 2 * t;
#pragma unroll
			for (int nbk = 0; nbk < 2; nbk++)
#pragma unroll
				for (int e = 0; e < 2; e++)
					if (wc + nbk * 8 + 2 * t + e < nb) row[nbk * 8 + e] = acc[nbk][half * 2 + e];
		}
		group_sync(0);
	}
}

// Owner: factored block and its inverse to global memory (factor: lower triangle incl. diagonal; inverse: transposed, strictly above it)
__device__ __forceinline__ void publish_block(float* __restrict__ S, int n, int j0, int nb, const ClusterShared& sh, int tid) {
#pragma unroll 4
	for (int idx = tid; idx < kPanel * kPanel; idx += kClusterThreads) {
		const int r = idx >> 5, c = idx & 31;
		if (r < nb && c < nb) S[(j0 + r) * n + j0 + c] = c <= r ? sh.D[r][c] : sh.Linv[c][r];
	}
}

__global__ void __launch_bounds__(kClusterThreads * kGroups) corner_cluster_factor_kernel(float* __restrict__ S, float* __restrict__ rhs, int n, int* __restrict__ flags) {
	__shared__ ClusterShared sh;
	const int tid = threadIdx.x & (kClusterThreads - 1), group = threadIdx.x / kClusterThreads;  // group 1 only takes tiles
	const int rank = (int) cluster_rank(), ctas = (int) cluster_ctas();
#ifdef NNRT_CLUSTER_TIMING
	if (threadIdx.x < 16) s_cluster_clock[threadIdx.x] = 0;
	__syncthreads();
	long long t_phase = clock64();
#endif
	if (rank == 0 && group == 0) {
		const int nb0 = min(kPanel, n);
#pragma unroll 4
		for (int idx = tid; idx < kPanel * kPanel; idx += kClusterThreads) {
			const int r = idx >> 5, c = idx & 31;
			sh.D[r][c] = (r < nb0 && c <= r) ? __ldcg(S + r * n + c) : (r == c ? 1.f : 0.f);
		}
		factor_and_invert_block(sh.D, sh.Linv, sh.T, nb0, flags, tid);
		publish_block(S, n, 0, nb0, sh, tid);
	}
	cluster_barrier();
	NNRT_CLUSTER_STAMP(0);
	for (int j0 = 0; j0 < n; j0 += kPanel) {
		const int nb = min(kPanel, n - j0), first = j0 + nb, step = j0 / kPanel;
		const bool has_next = first < n;
		const int nb1 = has_next ? min(kPanel, n - first) : 0;
		const int owner = (step + 1) % ctas;                 // of panel j + 1
		const int worker = (rank - owner - 1 + ctas) % ctas;  // position among the other CTAs
		const bool lone = ctas == 1;
		// ---- A: rows below the diagonal block (and the right-hand side row) times inv(L_jj)^T. The owner of panel j + 1 takes the nb1 rows
		// of its diagonal block, arrives at barrier 1 as soon as they are written and runs ahead: next diagonal block, factor, inverse,
		// publish — the dependency chain of the factorisation — while the others substitute the remaining rows and update the tiles.
		if (group == 0 && rank == owner && has_next) {
			substitute_rows(S, rhs, n, j0, nb, first, first + nb1, 0, 1, sh, tid);
			if (lone) substitute_rows(S, rhs, n, j0, nb, first + nb1, n + 1, 0, 1, sh, tid);
			NNRT_CLUSTER_STAMP(1);
			cluster_arrive();
			// P = L[first .. first + nb1, j0 .. j0 + nb] (this CTA's own writes); D = A[next block] - P P^T on the lower triangle
			float (*P)[kLd] = sh.stage[0].a;
#pragma unroll 4
			for (int idx = tid; idx < kPanel * kPanel; idx += kClusterThreads) {
				const int r = idx >> 5, c = idx & 31;
				P[r][c] = (r < nb1 && c < nb) ? __ldcg(S + (first + r) * n + j0 + c) : 0.f;
			}
			const int r = tid >> 3, cl = tid & 7;  // entries (r, cl + 8 q): conflict-free reads of P
			float prior[4];
#pragma unroll
			for (int q = 0; q < 4; q++) {
				const int c = cl + 8 * q;
				prior[q] = (r < nb1 && c <= r) ? __ldcg(S + (first + r) * n + first + c) : (r == c ? 1.f : 0.f);
			}
			group_sync(0);
			float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
			for (int k = 0; k < kPanel; k++) {
				const float pr = P[r][k];
#pragma unroll
				for (int q = 0; q < 4; q++) acc[q] = fmaf(pr, P[cl + 8 * q][k], acc[q]);
			}
#pragma unroll
			for (int q = 0; q < 4; q++) sh.D[r][cl + 8 * q] = prior[q] - acc[q];
			NNRT_CLUSTER_STAMP(6);
			factor_and_invert_block(sh.D, sh.Linv, sh.T, nb1, flags, tid);
			NNRT_CLUSTER_STAMP(7);
			publish_block(S, n, first, nb1, sh, tid);
			group_sync(0);
			NNRT_CLUSTER_STAMP(3);
			cluster_wait();
		} else {
			if (group == 0 && !has_next) {
				if (rank == 0) substitute_rows(S, rhs, n, j0, nb, first, n + 1, 0, 1, sh, tid);  // only the right-hand side row is left
			} else if (group == 0) {
				substitute_rows(S, rhs, n, j0, nb, first + nb1, n + 1, worker, ctas - 1, sh, tid);
			}
			NNRT_CLUSTER_STAMP(1);
			cluster_arrive();
			cluster_wait();
		}
		NNRT_CLUSTER_STAMP(2);
		if (!has_next) break;
		// ---- B: rank-nb update of the trailing matrix (the next diagonal block is the owner's)
		{
			const int tile_rows = (n + 1 - first + kTrail - 1) / kTrail, tiles = tile_rows * (tile_rows + 1) / 2;
			// tiles are dealt to (CTA, group) units: two per worker CTA, one (group 1) on the owner; a lone CTA works them after the factor
			const int units = lone ? kGroups : kGroups * (ctas - 1) + 1;
			const int unit = lone ? group : (rank == owner ? (group == 1 ? kGroups * (ctas - 1) : -1) : worker * kGroups + group);
#pragma unroll 1
			for (int tile = unit < 0 ? tiles : unit; tile < tiles; tile += units) {
				int a = (int) ((sqrtf(8.f * (float) tile + 1.f) - 1.f) * 0.5f);
				while (a * (a + 1) / 2 > tile) a--;
				while ((a + 1) * (a + 2) / 2 <= tile) a++;
				const int b = tile - a * (a + 1) / 2;
				update_tile(S, rhs, n, j0, nb, first + a * kTrail, first + b * kTrail, tile == 0 ? first + nb1 : 0, sh.stage[group], tid, group);
			}
			NNRT_CLUSTER_STAMP(4);
		}
		cluster_barrier();
		NNRT_CLUSTER_STAMP(5);
	}
#ifdef NNRT_CLUSTER_TIMING
	if (threadIdx.x == 0)
		for (int k = 0; k < 16; k++) g_cluster_clock[rank * 16 + k] += s_cluster_clock[k];
#endif
}

// Back substitution L^T x = y with the published block inverses, then the stem back substitution and the node update: one CTA of 1024
// threads. Per block (last to first): x_j = inv(L_jj)^T y_j (warp 0, inverse in shared memory), then y_k -= sum_c L(j0 + c, k) x(j0 + c)
// for every k < j0 (thread per k; the 32 loads are issued before the block solve, so their L2 latency hides behind it).
constexpr int kFinishThreads = 1024;

template<bool kReduced>
__global__ void __launch_bounds__(kFinishThreads) corner_cluster_finish_kernel(ArrowheadSystem sys, ReducedSystem red, const float* __restrict__ S,
                                                                               const float* __restrict__ rhs, int n, float* __restrict__ x,
                                                                               float* __restrict__ rotations, float* __restrict__ translations) {
	extern __shared__ float y[];  // [n]
	__shared__ float Linv[kPanel][kPanel + 1];
	const int tid = threadIdx.x, T = blockDim.x;
	for (int k = tid; k < n; k += T) y[k] = rhs[k];
	for (int j0 = ((n - 1) / kPanel) * kPanel; j0 >= 0; j0 -= kPanel) {
		const int nb = min(kPanel, n - j0);
		{
			const int c = tid >> 5, r = tid & 31;  // 1024 threads: one entry of the inverse each (lane = row: contiguous in memory)
			const bool stored = r < nb && c <= r;
			const float v = stored ? S[(j0 + c) * n + j0 + r] : 0.f;  // diagonal of L for r == c, transposed inverse above it otherwise
			Linv[r][c] = r == c ? 1.f / (stored ? v : 1.f) : v;
		}
		// this thread's column k of the rows of the block, for every pass over k < j0
		float l[kPanel];
		const int k0 = tid;
		if (k0 < j0) {
#pragma unroll
			for (int c = 0; c < kPanel; c++) l[c] = c < nb ? S[(j0 + c) * n + k0] : 0.f;
		}
		__syncthreads();
		if (tid < kPanel) {
			float acc = 0.f;
#pragma unroll 8
			for (int r = 0; r < kPanel; r++) acc = fmaf(Linv[r][tid], r < nb ? y[j0 + r] : 0.f, acc);  // x(c) = sum_r Linv(r, c) y(r)
			if (tid < nb) y[j0 + tid] = acc;
		}
		__syncthreads();
		if (k0 < j0) {
			float acc = y[k0];
#pragma unroll
			for (int c = 0; c < kPanel; c++) acc = fmaf(-l[c], c < nb ? y[j0 + c] : 0.f, acc);
			y[k0] = acc;
		}
		for (int k = tid + T; k < j0; k += T) {  // corners beyond 1024 unknowns
			float acc = y[k];
			for (int c = 0; c < nb; c++) acc = fmaf(-S[(j0 + c) * n + k], y[j0 + c], acc);
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

// cluster size per device ordinal: 0 = not probed yet, -1 = the device cannot co-schedule a cluster of this kernel
int cluster_ctas_for_device() {
	static std::atomic<int> chosen[64];
	int device = 0;
	if (cudaGetDevice(&device) != cudaSuccess) return -1;
	int ctas = chosen[device & 63].load(std::memory_order_acquire);
	if (ctas != 0) return ctas;
	ctas = -1;
	cudaFuncSetAttribute(corner_cluster_factor_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
	for (int candidate : {16, 8, 4}) {
		cudaLaunchConfig_t probe{};
		probe.gridDim = dim3(candidate); probe.blockDim = dim3(kClusterThreads * kGroups);
		cudaLaunchAttribute attribute{};
		attribute.id = cudaLaunchAttributeClusterDimension;
		attribute.val.clusterDim.x = candidate; attribute.val.clusterDim.y = 1; attribute.val.clusterDim.z = 1;
		probe.attrs = &attribute; probe.numAttrs = 1;
		int clusters = 0;
		if (cudaOccupancyMaxActiveClusters(&clusters, corner_cluster_factor_kernel, &probe) == cudaSuccess && clusters > 0) { ctas = candidate; break; }
	}
	cudaGetLastError();
	chosen[device & 63].store(ctas, std::memory_order_release);
	return ctas;
}

}  // namespace

// Factorisation + forward substitution of the dense corner system (lower triangle of schur [n, n], corner_rhs [n]) in one launch. The
// strict upper triangle of every 32 x 32 diagonal block is overwritten (inverse of the block's factor). kUnsupported: no cluster here.
int launch_corner_cluster_factor(float* schur, float* corner_rhs, int n, int* flags, cudaStream_t s) {
	if ((int64_t) n * n >= (1ll << 31)) return kUnsupported;
	int ctas = cluster_ctas_for_device();
	if (ctas < 0) return kUnsupported;
	while (ctas > 4 && n <= kTrail * (ctas / 2 - 1)) ctas /= 2;  // small corners: fewer CTAs, cheaper barriers
	cudaLaunchConfig_t config{};
	config.gridDim = dim3(ctas); config.blockDim = dim3(kClusterThreads * kGroups); config.stream = s;
	cudaLaunchAttribute attribute{};
	attribute.id = cudaLaunchAttributeClusterDimension;
	attribute.val.clusterDim.x = ctas; attribute.val.clusterDim.y = 1; attribute.val.clusterDim.z = 1;
	config.attrs = &attribute; config.numAttrs = 1;
	if (cudaLaunchKernelEx(&config, corner_cluster_factor_kernel, schur, corner_rhs, n, flags) != cudaSuccess) {
		set_last_error("corner_cluster_factor: cluster launch failed (%d CTAs)", ctas);
		return kCudaError;
	}
	return kOk;
}

// Back substitution after launch_corner_cluster_factor (reads the block inverses it left), stem back substitution, node update.
int launch_corner_cluster_finish(const ArrowheadSystem& sys, const ReducedSystem* red, int n, float* x, float* node_rotations, float* node_translations,
                                 cudaStream_t s) {
	const size_t bytes = (size_t) n * sizeof(float);
	if (bytes > 160 * 1024) return kUnsupported;
	static SharedMemoryOptIn opt_in, reduced_opt_in;
	if (red) {
		if (reduced_opt_in.ensure(corner_cluster_finish_kernel<true>, 160 * 1024) != kOk) return kCudaError;
		corner_cluster_finish_kernel<true><<<1, kFinishThreads, bytes, s>>>(sys, *red, sys.schur, sys.corner_rhs, n, x, node_rotations, node_translations);
	} else {
		if (opt_in.ensure(corner_cluster_finish_kernel<false>, 160 * 1024) != kOk) return kCudaError;
		corner_cluster_finish_kernel<false><<<1, kFinishThreads, bytes, s>>>(sys, ReducedSystem{}, sys.schur, sys.corner_rhs, n, x, node_rotations, node_translations);
	}
	NNRT_LAUNCH_CHECK();
	return kOk;
}

}  // namespace nnrt_b200
