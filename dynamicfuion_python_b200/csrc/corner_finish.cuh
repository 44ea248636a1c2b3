// Device helpers shared by the corner kernels (solver_kernels.cu, corner_cluster.cu): the node update applied at the end of a solve and
// the back substitution of the reduced fitter system.
#pragma once
#include "solver_kernels.cuh"

namespace nnrt_b200 {

// core::linalg::AxisAngleVectorsToMatricesRodrigues (cpp/core/linalg/RodriguesImpl.h:66-100), no small-angle guard (as the reference)
__device__ inline void rodrigues(const float* __restrict__ w, float* __restrict__ R) {
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
// `mode` follows IterationMode (DeformableMeshToImageFitter.cpp:258-273): TRANSLATION_ONLY (1) never evaluates Rodrigues (a zero
// rotation update would be 0 / 0 there), ROTATION_ONLY (2) leaves the translations untouched.
__device__ __forceinline__ void apply_node_update(const float* __restrict__ dlt, float* __restrict__ rotation, float* __restrict__ translation,
                                                  int mode) {
	if (mode != 1) {
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
	}
	if (mode != 2) {
#pragma unroll
		for (int c = 0; c < 3; c++) translation[c] = translation[c] + dlt[3 + c];
	}
}

// Reduced fitter system, after the dense solve of the corner translations xt (shared memory, 3 per corner node): corner rotations
// x_r = A^-1 b_r - (A^-1 B) x_t, stem unknowns x_i = y_i - sum_e Y_e x_t[col(e)]; x is [N, 6] = (rotation, translation) per node.
__device__ __forceinline__ void reduced_back_substitution(const ReducedSystem& red, const float* __restrict__ xt, float* __restrict__ x, int tid, int T) {
	const int nc = red.node_count - red.arrow_base;
	for (int idx = tid; idx < nc * 6; idx += T) {
		const int c = idx / 6, r = idx - c * 6;
		float v;
		if (r >= 3) v = xt[c * 3 + r - 3];
		else {
			v = red.corner_zr[c * 3 + r];
#pragma unroll
			for (int k = 0; k < 3; k++) v = fmaf(-red.corner_ainvb[c * 9 + r * 3 + k], xt[c * 3 + k], v);
		}
		x[(red.arrow_base + c) * 6 + r] = v;
	}
	for (int idx = tid; idx < red.arrow_base * 6; idx += T) {
		const int i = idx / 6, r = idx - i * 6;
		float acc = red.stem_y[idx];
		for (int e = red.row_offsets[i]; e < red.row_offsets[i + 1]; e++) {
			const int w = red.row_blocks[e];
			const float* Y = red.wing_y + (int64_t) w * 18 + r * 3;
			const float* xv = xt + (red.wing_columns[w] - red.arrow_base) * 3;
#pragma unroll
			for (int k = 0; k < 3; k++) acc = fmaf(-Y[k], xv[k], acc);
		}
		x[idx] = acc;
	}
}

}  // namespace nnrt_b200
