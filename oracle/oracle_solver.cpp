// ORACLE (test infrastructure only) — ARAP term, block-sparse arrowhead / block-diagonal
// Cholesky solves, LM damping, Rodrigues, node update. See oracle_common.h for status.
// LAPACK (potrf/potrs/trsm) and BLAS batched gemm of the reference's CPU path are restated with
// hand-rolled fp32 Cholesky / triangular solves: same algorithm class, agreement to fp32 round-off
// (inside the 1e-4 / 1e-3 bars; SURVEY.md §8(c) "third-party arithmetic").
#include "oracle_common.h"

using namespace oracle;

// cpp/alignment/kernel/DeformableMeshToImageFitterImpl.h:644-784.
// edge_layer_indices != nullptr -> FixedCoverageWeight: weight = radii[edge_layer_indices[node_j]] (sic, :708);
// else VariableCoverageWeight: weight = max(w_i, w_j).
ORACLE_API void oracle_arap_residuals(
		const int32_t* edges, int64_t edge_count, const int8_t* edge_layer_indices, const float* layer_decimation_radii,
		const float* node_coverage_weights, const float* nodes, const float* node_translations, const float* node_rotations,
		float regularization_weight, float* edge_residuals /*[3E]*/) {
	for (int64_t e = 0; e < edge_count; e++) {
		int i = edges[e * 2], j = edges[e * 2 + 1];
		V3 gi = load3(nodes + i * 3), gj = load3(nodes + j * 3);
		V3 ti = load3(node_translations + i * 3), tj = load3(node_translations + j * 3);
		float edge_weight = edge_layer_indices ? layer_decimation_radii[edge_layer_indices[j]]
		                                       : fmaxf(node_coverage_weights[i], node_coverage_weights[j]);
		V3 inner = sub(sub(add(gi, ti), add(gj, tj)), matvec(node_rotations + i * 9, sub(gi, gj)));
		store3(edge_residuals + e * 3, scale(regularization_weight * edge_weight, inner));
	}
}

// cpp/alignment/functional/kernel/ArapJacobianImpl.h:36-193: [E,5] = (-lw R_i (g_i - g_j), +lw, -lw);
// FixedCoverageWeight indexes radii by edge_layer_indices[i_edge] (:89,107).
ORACLE_API void oracle_arap_edge_jacobians(
		const int32_t* edges, int64_t edge_count, const int8_t* edge_layer_indices, const float* layer_decimation_radii,
		const float* node_coverage_weights, const float* nodes, const float* node_rotations,
		float regularization_weight, float* edge_jacobians /*[E,5]*/) {
	for (int64_t e = 0; e < edge_count; e++) {
		int i = edges[e * 2], j = edges[e * 2 + 1];
		float edge_weight = edge_layer_indices ? layer_decimation_radii[edge_layer_indices[e]]
		                                       : fmaxf(node_coverage_weights[i], node_coverage_weights[j]);
		V3 rv = matvec(node_rotations + i * 9, sub(load3(nodes + i * 3), load3(nodes + j * 3)));
		store3(edge_jacobians + e * 5, scale(-regularization_weight * edge_weight, rv));
		edge_jacobians[e * 5 + 3] = regularization_weight * edge_weight;
		edge_jacobians[e * 5 + 4] = -regularization_weight * edge_weight;
	}
}

namespace {
// dE_i = [skew(s) | a I] (3x6), skew = [[0,-sz,sy],[sz,0,-sx],[-sy,sx,0]]
void make_dEi(float dE[3][6], const float* s, float a) {
	const float S[3][3] = {{0, -s[2], s[1]}, {s[2], 0, -s[0]}, {-s[1], s[0], 0}};
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) { dE[r][c] = S[r][c]; dE[r][3 + c] = (r == c) ? a : 0.f; }
}
void make_dEj(float dE[3][6], float b) {
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) { dE[r][c] = 0.f; dE[r][3 + c] = (r == c) ? b : 0.f; }
}
void AtB(float* out36, const float A[3][6], const float B[3][6], bool accumulate) {
	for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) {
		float v = sum3(A[0][r] * B[0][c], A[1][r] * B[1][c], A[2][r] * B[2][c]);
		if (accumulate) out36[r * 6 + c] += v; else out36[r * 6 + c] = v;
	}
}
} // namespace

// cpp/alignment/functional/kernel/ArapHessianImpl.h:44-193 as invoked by
// cpp/alignment/functional/ArapHessian.cpp:30-48 (second_layer_node_count = 0 => every edge is an
// upper-wing block, block id = edge id; requires i < arrow_base for all edges — the only defined regime).
// Diagonal sums are accumulated in ascending (edge, i-then-j) order (the reference uses float atomics).
// Returns -1 if an edge violates i < arrow_base <= j.
ORACLE_API int oracle_arap_hessian(
		const int32_t* edges, int64_t edge_count, const float* edge_jacobians, int arrow_base, int node_count,
		float* diagonal_blocks /*[N,36]*/, float* wing_blocks /*[E,36]*/, int32_t* wing_coordinates /*[E,2]*/) {
	memset(diagonal_blocks, 0, sizeof(float) * node_count * 36);
	for (int64_t e = 0; e < edge_count; e++) {
		int i = edges[e * 2], j = edges[e * 2 + 1];
		if (!(i < arrow_base && j >= arrow_base && j < node_count && i >= 0)) return -1;
		float dEi[3][6], dEj[3][6];
		make_dEi(dEi, edge_jacobians + e * 5, edge_jacobians[e * 5 + 3]);
		make_dEj(dEj, edge_jacobians[e * 5 + 4]);
		AtB(wing_blocks + e * 36, dEi, dEj, false);
		wing_coordinates[e * 2] = i; wing_coordinates[e * 2 + 1] = j;
		AtB(diagonal_blocks + i * 36, dEi, dEi, true);
		AtB(diagonal_blocks + j * 36, dEj, dEj, true);
	}
	return 0;
}

// cpp/alignment/kernel/DeformableMeshToImageFitterImpl.h:464-642 (ALL): g -= dE^T e, ascending edge order.
ORACLE_API void oracle_negative_arap_gradient(
		const float* edge_residuals, const float* edge_jacobians, const int32_t* edges, int64_t edge_count,
		int node_count, float* negative_gradient /*[N*6]*/) {
	memset(negative_gradient, 0, sizeof(float) * node_count * 6);
	for (int64_t e = 0; e < edge_count; e++) {
		int i = edges[e * 2], j = edges[e * 2 + 1];
		const float* s = edge_jacobians + e * 5;
		V3 r = load3(edge_residuals + e * 3);
		// skew(s)^T * r
		const float ST[3][3] = {{0, s[2], -s[1]}, {-s[2], 0, s[0]}, {s[1], -s[0], 0}};
		float a = edge_jacobians[e * 5 + 3], b = edge_jacobians[e * 5 + 4];
		for (int c = 0; c < 3; c++) {
			negative_gradient[i * 6 + c] -= sum3(ST[c][0] * r.x, ST[c][1] * r.y, ST[c][2] * r.z);
		}
		negative_gradient[i * 6 + 3] -= a * r.x; negative_gradient[i * 6 + 4] -= a * r.y; negative_gradient[i * 6 + 5] -= a * r.z;
		negative_gradient[j * 6 + 3] -= b * r.x; negative_gradient[j * 6 + 4] -= b * r.y; negative_gradient[j * 6 + 5] -= b * r.z;
	}
}

// cpp/core/linalg/PreconditionDiagonalBlocksImpl.h:29-51: block[d][d] += lambda
ORACLE_API void oracle_precondition_diagonal_blocks(float* blocks, int block_count, int block_size, float lambda) {
	for (int b = 0; b < block_count; b++)
		for (int d = 0; d < block_size; d++) blocks[(int64_t) b * block_size * block_size + d * block_size + d] += lambda;
}

namespace {
// dense in-place Cholesky A = L L^T (lower, row-major n x n); returns false if not PD
bool cholesky_lower(float* A, int n) {
	for (int j = 0; j < n; j++) {
		float d = A[j * n + j];
		for (int k = 0; k < j; k++) d -= A[j * n + k] * A[j * n + k];
		if (!(d > 0.f)) return false;
		d = sqrtf(d);
		A[j * n + j] = d;
		for (int i = j + 1; i < n; i++) {
			float s = A[i * n + j];
			for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k];
			A[i * n + j] = s / d;
		}
	}
	return true;
}
void cholesky_solve(const float* L, int n, float* x /* in: b, out: x */) {
	for (int i = 0; i < n; i++) {
		float s = x[i];
		for (int k = 0; k < i; k++) s -= L[i * n + k] * x[k];
		x[i] = s / L[i * n + i];
	}
	for (int i = n - 1; i >= 0; i--) {
		float s = x[i];
		for (int k = i + 1; k < n; k++) s -= L[k * n + i] * x[k];
		x[i] = s / L[i * n + i];
	}
}
} // namespace

// cpp/core/linalg/SolveBlockDiagonalCholesky.cpp + SolveBlockDiagonalGeneric.h:27-90: per-block potrf + 2 trsm.
// Returns number of blocks that failed to factorise.
ORACLE_API int oracle_solve_block_diagonal_cholesky(const float* blocks, const float* b, int block_count, int block_size, float* x) {
	int failures = 0;
	std::vector<float> L((size_t) block_size * block_size);
	for (int i = 0; i < block_count; i++) {
		std::copy(blocks + (int64_t) i * block_size * block_size, blocks + (int64_t) (i + 1) * block_size * block_size, L.begin());
		for (int d = 0; d < block_size; d++) x[i * block_size + d] = b[i * block_size + d];
		if (!cholesky_lower(L.data(), block_size)) { failures++; continue; }
		cholesky_solve(L.data(), block_size, x + i * block_size);
	}
	return failures;
}

// cpp/core/linalg/SchurComplement.cpp:44-77 + InvertBlocks.cpp:82-127 + MatmulBlockSparseImpl.h:
// S = C - B^T D^-1 B (dense, (6 Nc)^2), with C = corner diagonal blocks (+ optional corner upper blocks, mirrored).
// Also returns D^-1 [Ns,36] and D^-1 B [W,36] when the pointers are non-null.
ORACLE_API int oracle_arrowhead_schur_complement(
		const float* diagonal_blocks, int node_count, int arrow_base,
		const float* wing_blocks, const int32_t* wing_coordinates, int64_t wing_count,
		const float* corner_upper_blocks, const int32_t* corner_upper_coordinates, int64_t corner_upper_count,
		float* schur /*[(6Nc)^2]*/, float* inverted_stem /*[Ns,36] or null*/, float* inverted_stem_wing /*[W,36] or null*/) {
	const int nc = node_count - arrow_base, n = nc * 6;
	std::vector<float> Dinv((size_t) arrow_base * 36), DinvB((size_t) wing_count * 36);
	int failures = 0;
	for (int i = 0; i < arrow_base; i++) {  // potrf + potrs(I)
		float L[36];
		std::copy(diagonal_blocks + i * 36, diagonal_blocks + (i + 1) * 36, L);
		if (!cholesky_lower(L, 6)) { failures++; continue; }
		for (int c = 0; c < 6; c++) {
			float col[6] = {0, 0, 0, 0, 0, 0};
			col[c] = 1.f;
			cholesky_solve(L, 6, col);
			for (int r = 0; r < 6; r++) Dinv[(size_t) i * 36 + r * 6 + c] = col[r];
		}
	}
	for (int64_t w = 0; w < wing_count; w++) {
		const float* Di = Dinv.data() + (size_t) wing_coordinates[w * 2] * 36;
		const float* B = wing_blocks + w * 36;
		for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) {
			float acc = 0.f;
			for (int k = 0; k < 6; k++) acc += Di[r * 6 + k] * B[k * 6 + c];
			DinvB[(size_t) w * 36 + r * 6 + c] = acc;
		}
	}
	std::fill(schur, schur + (size_t) n * n, 0.f);
	for (int bi = 0; bi < nc; bi++)
		for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++)
			schur[(size_t) (bi * 6 + r) * n + bi * 6 + c] = diagonal_blocks[(size_t) (arrow_base + bi) * 36 + r * 6 + c];
	for (int64_t u = 0; u < corner_upper_count; u++) {
		int bi = corner_upper_coordinates[u * 2] - arrow_base, bj = corner_upper_coordinates[u * 2 + 1] - arrow_base;
		for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) {
			schur[(size_t) (bi * 6 + r) * n + bj * 6 + c] = corner_upper_blocks[u * 36 + r * 6 + c];
			schur[(size_t) (bj * 6 + c) * n + bi * 6 + r] = corner_upper_blocks[u * 36 + r * 6 + c];
		}
	}
	// B^T (D^-1 B): products of wing blocks sharing a stem row
	std::vector<std::vector<int64_t>> row_blocks(arrow_base);
	for (int64_t w = 0; w < wing_count; w++) row_blocks[wing_coordinates[w * 2]].push_back(w);
	for (int i = 0; i < arrow_base; i++) {
		for (int64_t wa : row_blocks[i]) for (int64_t wb : row_blocks[i]) {
			int ca = wing_coordinates[wa * 2 + 1] - arrow_base, cb = wing_coordinates[wb * 2 + 1] - arrow_base;
			const float* Ba = wing_blocks + wa * 36;
			const float* Xb = DinvB.data() + (size_t) wb * 36;
			for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) {
				float acc = 0.f;
				for (int k = 0; k < 6; k++) acc += Ba[k * 6 + r] * Xb[k * 6 + c];
				schur[(size_t) (ca * 6 + r) * n + cb * 6 + c] -= acc;
			}
		}
	}
	if (inverted_stem) std::copy(Dinv.begin(), Dinv.end(), inverted_stem);
	if (inverted_stem_wing) std::copy(DinvB.begin(), DinvB.end(), inverted_stem_wing);
	return failures;
}

// cpp/core/linalg/SolveBlockSparseArrowheadCholesky.cpp:31-94. x = concat(x_stem, x_corner).
ORACLE_API int oracle_solve_arrowhead_cholesky(
		const float* diagonal_blocks, int node_count, int arrow_base,
		const float* wing_blocks, const int32_t* wing_coordinates, int64_t wing_count,
		const float* corner_upper_blocks, const int32_t* corner_upper_coordinates, int64_t corner_upper_count,
		const float* b, float* x) {
	const int nc = node_count - arrow_base, n = nc * 6;
	std::vector<float> S((size_t) n * n), Dinv((size_t) arrow_base * 36), DinvB((size_t) wing_count * 36);
	int failures = oracle_arrowhead_schur_complement(diagonal_blocks, node_count, arrow_base, wing_blocks, wing_coordinates, wing_count,
	                                                 corner_upper_blocks, corner_upper_coordinates, corner_upper_count,
	                                                 S.data(), Dinv.data(), DinvB.data());
	// b_C' = b_C - (D^-1 B)^T b_D
	std::vector<float> update((size_t) n, 0.f);
	for (int64_t w = 0; w < wing_count; w++) {
		int i = wing_coordinates[w * 2], cj = wing_coordinates[w * 2 + 1] - arrow_base;
		for (int c = 0; c < 6; c++) {
			float acc = 0.f;
			for (int k = 0; k < 6; k++) acc += DinvB[(size_t) w * 36 + k * 6 + c] * b[i * 6 + k];
			update[cj * 6 + c] += acc;
		}
	}
	float* x_corner = x + arrow_base * 6;
	for (int k = 0; k < n; k++) x_corner[k] = b[arrow_base * 6 + k] - update[k];
	if (n > 0) {
		if (!cholesky_lower(S.data(), n)) failures++;
		else cholesky_solve(S.data(), n, x_corner);
	}
	// x_D = D^-1 (b_D - B x_C)
	std::vector<float> rhs(b, b + arrow_base * 6);
	std::vector<float> bx((size_t) arrow_base * 6, 0.f);
	for (int64_t w = 0; w < wing_count; w++) {
		int i = wing_coordinates[w * 2], cj = wing_coordinates[w * 2 + 1] - arrow_base;
		for (int r = 0; r < 6; r++) {
			float acc = 0.f;
			for (int k = 0; k < 6; k++) acc += wing_blocks[w * 36 + r * 6 + k] * x_corner[cj * 6 + k];
			bx[i * 6 + r] += acc;
		}
	}
	for (int i = 0; i < arrow_base; i++)
		for (int r = 0; r < 6; r++) {
			float acc = 0.f;
			for (int k = 0; k < 6; k++) acc += Dinv[(size_t) i * 36 + r * 6 + k] * (rhs[i * 6 + k] - bx[i * 6 + k]);
			x[i * 6 + r] = acc;
		}
	return failures;
}

// cpp/core/linalg/RodriguesImpl.h:66-100 (no small-angle guard: NaN at zero, as in the reference)
ORACLE_API void oracle_rodrigues(const float* axis_angles, int count, float* matrices) {
	for (int i = 0; i < count; i++) {
		V3 v = load3(axis_angles + i * 3);
		float angle = sqrtf(dot3(v, v));
		V3 k = {v.x / angle, v.y / angle, v.z / angle};
		const float K[3][3] = {{0, -k.z, k.y}, {k.z, 0, -k.x}, {-k.y, k.x, 0}};
		float s = sinf(angle), c1 = 1 - cosf(angle);
		for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) {
			float kk = sum3(K[r][0] * K[0][c], K[r][1] * K[1][c], K[r][2] * K[2][c]);
			matrices[i * 9 + r * 3 + c] = ((r == c ? 1.f : 0.f) + s * K[r][c]) + c1 * kk;
		}
	}
}

// cpp/geometry/HierarchicalGraphWarpField.cpp:261-284 with identity node permutation:
// t += dt ; R <- R * dR (cpp/core/linalg/Matmul3D_CPU.cpp batched 3x3 gemm)
ORACLE_API void oracle_apply_node_updates(const float* deltas /*[N,6]*/, int node_count, float* node_rotations, float* node_translations) {
	std::vector<float> dR((size_t) node_count * 9), w((size_t) node_count * 3);
	for (int i = 0; i < node_count; i++) for (int c = 0; c < 3; c++) w[i * 3 + c] = deltas[i * 6 + c];
	oracle_rodrigues(w.data(), node_count, dR.data());
	for (int i = 0; i < node_count; i++) {
		for (int c = 0; c < 3; c++) node_translations[i * 3 + c] = node_translations[i * 3 + c] + deltas[i * 6 + 3 + c];
		float R[9];
		std::copy(node_rotations + i * 9, node_rotations + i * 9 + 9, R);
		for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) {
			float acc = 0.f;
			for (int k = 0; k < 3; k++) acc += R[r * 3 + k] * dR[i * 9 + k * 3 + c];
			node_rotations[i * 9 + r * 3 + c] = acc;
		}
	}
}
