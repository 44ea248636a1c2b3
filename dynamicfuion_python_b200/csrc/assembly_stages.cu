// Un-fused stage entry points of the assembly chain, for parity tests against the reference's intermediate tensors. The fitter's
// iteration never calls these (assemble_kernel / system_stem_kernel fuse them); they exist so that the per-pixel Jacobians and the
// per-edge ARAP quantities can be compared BEFORE any per-node summation.
//   nnrt_b200_pixel_vertex_anchor_jacobians : alignment::functional::PixelVertexAnchorJacobiansAndNodeAssociations
//       (cpp/alignment/functional/PixelVertexAnchorJacobians.cpp:29-56, kernel/PixelVertexAnchorJacobiansImpl.h:40-434)
//   nnrt_b200_arap_term : ComputeEdgeResiduals (cpp/alignment/DeformableMeshToImageFitter.cpp:392-448,
//       kernel/DeformableMeshToImageFitterImpl.h:644-784), ComputeDenseArapEdgeJacobians (functional/ArapJacobian.cpp:25-72),
//       ComputeArapBlockSparseHessianApproximation (functional/ArapHessian.cpp:30-48), ComputeNegativeArapGradient (...Impl.h:464-642)
#include "host_util.h"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {

// One thread per pixel. `mode`: 0 ALL (6 floats per entry: rotation xyz, translation xyz), 1 TRANSLATION_ONLY, 2 ROTATION_ONLY
// (3 floats per entry), as IterationMode selects the stride in the reference.
__global__ void __launch_bounds__(128) pixel_vertex_anchor_jacobians_kernel(
		const float* __restrict__ rv_jac /*[PX,27]*/, const float* __restrict__ rn_jac /*[PX,30]*/, const float* __restrict__ wv_jac /*[V,A,4]*/,
		const float* __restrict__ wn_jac /*[V,A,3]*/, const float* __restrict__ point_map /*[PX,3]*/, const float* __restrict__ normals /*[PX,3]*/,
		const uint8_t* __restrict__ residual_mask, const int64_t* __restrict__ pixel_faces, int faces_per_pixel,
		const int64_t* __restrict__ triangles, const int32_t* __restrict__ face_entries /*[F,3A,2]*/, const int32_t* __restrict__ face_counts,
		int64_t pixel_count, int A, int node_count, int mode, int use_tukey, float tukey_cutoff,
		float* __restrict__ pixel_jacobians /*[PX,3A,S]*/, int32_t* __restrict__ pixel_counts, int32_t* __restrict__ node_pixel_counts) {
	const int64_t px = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (px >= pixel_count) return;
	const int S = mode == 0 ? 6 : 3;
	float* pj = pixel_jacobians + px * (3 * A * S);
	for (int k = 0; k < 3 * A * S; k++) pj[k] = 0.f;
	pixel_counts[px] = 0;
	if (!residual_mask[px]) return;
	const float3 pm = make_float3(point_map[px * 3], point_map[px * 3 + 1], point_map[px * 3 + 2]);
	const float3 nl = make_float3(normals[px * 3], normals[px * 3 + 1], normals[px * 3 + 2]);
	float3 dr_dnl = pm, dr_dwl = nl;
	if (use_tukey) {  // PixelVertexAnchorJacobiansImpl.h:200-219
		const float r = dot3(nl.x, nl.y, nl.z, pm.x, pm.y, pm.z);
		if (fabsf(r) > tukey_cutoff) return;
		const float q = r / tukey_cutoff;
		float psi = 1 - q * q;
		psi = r * psi * psi;
		dr_dnl = psi * pm;
		dr_dwl = psi * nl;
	}
	const float* dwl_dV = rv_jac + px * 27;
	const float* dnl_dV = rn_jac + px * 30;
	const float* rho = dnl_dV + 27;
	float dr_dV[9], dr_dN[9];
#pragma unroll
	for (int c = 0; c < 9; c++) {
		const float a = sum3(dr_dwl.x * dwl_dV[c], dr_dwl.y * dwl_dV[9 + c], dr_dwl.z * dwl_dV[18 + c]);
		const float b = sum3(dr_dnl.x * dnl_dV[c], dr_dnl.y * dnl_dV[9 + c], dr_dnl.z * dnl_dV[18 + c]);
		dr_dV[c] = a + b;
		dr_dN[c] = (c % 3 == 0 ? dr_dnl.x : c % 3 == 1 ? dr_dnl.y : dr_dnl.z) * rho[c / 3];
	}
	const int64_t f = pixel_faces[px * faces_per_pixel];
	const int count = face_counts[f];
	const int32_t* entries = face_entries + f * (3 * A) * 2;
	pixel_counts[px] = count;
	for (int k = 0; k < count; k++) {
		const int node = entries[k * 2];
		const uint32_t packed = (uint32_t) entries[k * 2 + 1];
		float Jr[3] = {0.f, 0.f, 0.f}, Jt[3] = {0.f, 0.f, 0.f};
		for (int fv = 0; fv < 3; fv++) {
			const uint32_t slot = (packed >> (8 * fv)) & 0xFFu;
			if (slot == 0xFFu) continue;
			const int64_t vi = triangles[f * 3 + fv];
			const float* wj = wv_jac + (vi * A + slot) * 4;
			const float* nj = wn_jac + (vi * A + slot) * 3;
			const float w = wj[3];
			const float3 dv = make_float3(dr_dV[fv * 3], dr_dV[fv * 3 + 1], dr_dV[fv * 3 + 2]);
			const float3 dn = make_float3(dr_dN[fv * 3], dr_dN[fv * 3 + 1], dr_dN[fv * 3 + 2]);
			Jt[0] += dv.x * w; Jt[1] += dv.y * w; Jt[2] += dv.z * w;
			// row * skew(s), skew = [[0,-sz,sy],[sz,0,-sx],[-sy,sx,0]]
			const float pv0 = dv.y * wj[2] + dv.z * -wj[1], pv1 = dv.x * -wj[2] + dv.z * wj[0], pv2 = dv.x * wj[1] + dv.y * -wj[0];
			const float pn0 = dn.y * nj[2] + dn.z * -nj[1], pn1 = dn.x * -nj[2] + dn.z * nj[0], pn2 = dn.x * nj[1] + dn.y * -nj[0];
			Jr[0] += pv0 + pn0; Jr[1] += pv1 + pn1; Jr[2] += pv2 + pn2;
		}
		float* J = pj + k * S;
		if (mode == 0) { J[0] = Jr[0]; J[1] = Jr[1]; J[2] = Jr[2]; J[3] = Jt[0]; J[4] = Jt[1]; J[5] = Jt[2]; }
		else if (mode == 1) { J[0] = Jt[0]; J[1] = Jt[1]; J[2] = Jt[2]; }
		else { J[0] = Jr[0]; J[1] = Jr[1]; J[2] = Jr[2]; }
		if (node >= 0 && node < node_count) atomicAdd(node_pixel_counts + node, 1);
	}
}

struct ArapInputs {
	const int32_t* edges; int64_t edge_count;
	const int8_t* edge_layer_indices; const float* radii; const float* coverage;
	const float* nodes; const float* rotations; const float* translations;
	float weight; int use_huber; float huber_constant;
};

__device__ __forceinline__ void arap_edge_quantities(const ArapInputs& in, int64_t e, float (&jac)[5], float3& residual) {
	const int i = in.edges[e * 2], j = in.edges[e * 2 + 1];
	const float3 gi = make_float3(in.nodes[i * 3], in.nodes[i * 3 + 1], in.nodes[i * 3 + 2]);
	const float3 gj = make_float3(in.nodes[j * 3], in.nodes[j * 3 + 1], in.nodes[j * 3 + 2]);
	const float3 ti = make_float3(in.translations[i * 3], in.translations[i * 3 + 1], in.translations[i * 3 + 2]);
	const float3 tj = make_float3(in.translations[j * 3], in.translations[j * 3 + 1], in.translations[j * 3 + 2]);
	// sic: the residual indexes the layer table by NODE j (kernel/DeformableMeshToImageFitterImpl.h:708), the Jacobian by edge
	const float w_res = in.edge_layer_indices ? in.radii[in.edge_layer_indices[j]] : fmaxf(in.coverage[i], in.coverage[j]);
	const float w_jac = in.edge_layer_indices ? in.radii[in.edge_layer_indices[e]] : fmaxf(in.coverage[i], in.coverage[j]);
	const float3 rd = matvec3(in.rotations + (int64_t) i * 9, gi - gj);
	const float3 s = (-in.weight * w_jac) * rd;
	jac[0] = s.x; jac[1] = s.y; jac[2] = s.z; jac[3] = in.weight * w_jac; jac[4] = -in.weight * w_jac;
	residual = (in.weight * w_res) * (((gi + ti) - (gj + tj)) - rd);
	if (in.use_huber) {  // DeformableMeshToImageFitter.cpp:435-444
		const float dl = in.huber_constant, half = 0.5f * dl * dl;
		residual.x = residual.x >= dl ? fabsf(residual.x) - half : 0.5f * residual.x * residual.x;
		residual.y = residual.y >= dl ? fabsf(residual.y) - half : 0.5f * residual.y * residual.y;
		residual.z = residual.z >= dl ? fabsf(residual.z) - half : 0.5f * residual.z * residual.z;
	}
}

// dE_i = [skew(s) | a I], dE_j = [0 | b I]  (ArapHessianImpl.h:117-189)
__device__ __forceinline__ float dEi_entry(const float (&jac)[5], int r, int c) {
	if (c >= 3) return (r == c - 3) ? jac[3] : 0.f;
	if (r == c) return 0.f;
	const int k = 3 - r - c;  // the remaining axis
	const bool positive = (r == 0 && c == 2) || (r == 1 && c == 0) || (r == 2 && c == 1);
	return positive ? jac[k] : -jac[k];
}

__global__ void __launch_bounds__(128) arap_edge_kernel(ArapInputs in, float* __restrict__ edge_residuals, float* __restrict__ edge_jacobians,
                                                         float* __restrict__ wing_blocks, int32_t* __restrict__ wing_coordinates) {
	const int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (e >= in.edge_count) return;
	float jac[5];
	float3 r;
	arap_edge_quantities(in, e, jac, r);
	if (edge_residuals) { edge_residuals[e * 3] = r.x; edge_residuals[e * 3 + 1] = r.y; edge_residuals[e * 3 + 2] = r.z; }
	if (edge_jacobians)
		for (int k = 0; k < 5; k++) edge_jacobians[e * 5 + k] = jac[k];
	if (wing_blocks) {
		const float b = jac[4];
		for (int rr = 0; rr < 6; rr++)
			for (int c = 0; c < 6; c++) {
				// (dE_i^T dE_j)[rr][c] = sum_k dE_i[k][rr] * dE_j[k][c]; dE_j[k][c] = b if c == 3 + k
				wing_blocks[e * 36 + rr * 6 + c] =
						sum3(dEi_entry(jac, 0, rr) * (c == 3 ? b : 0.f), dEi_entry(jac, 1, rr) * (c == 4 ? b : 0.f), dEi_entry(jac, 2, rr) * (c == 5 ? b : 0.f));
			}
	}
	if (wing_coordinates) { wing_coordinates[e * 2] = in.edges[e * 2]; wing_coordinates[e * 2 + 1] = in.edges[e * 2 + 1]; }
}

// One thread per node: diagonal block and gradient summed over the node's edges in ASCENDING edge order (the reference's float
// atomics make its own order scheduling-dependent; ascending is the single-threaded CPU order).
__global__ void __launch_bounds__(128) arap_node_kernel(ArapInputs in, int node_count, float* __restrict__ diagonal_blocks,
                                                         float* __restrict__ negative_gradient) {
	const int n = blockIdx.x * blockDim.x + threadIdx.x;
	if (n >= node_count) return;
	float H[36], g[6];
	for (int k = 0; k < 36; k++) H[k] = 0.f;
	for (int k = 0; k < 6; k++) g[k] = 0.f;
	for (int64_t e = 0; e < in.edge_count; e++) {
		const int i = in.edges[e * 2], j = in.edges[e * 2 + 1];
		if (i != n && j != n) continue;
		float jac[5];
		float3 r;
		arap_edge_quantities(in, e, jac, r);
		if (i == n) {
			for (int rr = 0; rr < 6; rr++)
				for (int c = 0; c < 6; c++)
					H[rr * 6 + c] += sum3(dEi_entry(jac, 0, rr) * dEi_entry(jac, 0, c), dEi_entry(jac, 1, rr) * dEi_entry(jac, 1, c),
					                      dEi_entry(jac, 2, rr) * dEi_entry(jac, 2, c));
			for (int c = 0; c < 6; c++) g[c] -= sum3(dEi_entry(jac, 0, c) * r.x, dEi_entry(jac, 1, c) * r.y, dEi_entry(jac, 2, c) * r.z);
		}
		if (j == n) {
			const float b = jac[4];
			H[21] += b * b; H[28] += b * b; H[35] += b * b;
			g[3] -= b * r.x; g[4] -= b * r.y; g[5] -= b * r.z;
		}
	}
	if (diagonal_blocks)
		for (int k = 0; k < 36; k++) diagonal_blocks[(int64_t) n * 36 + k] = H[k];
	if (negative_gradient)
		for (int k = 0; k < 6; k++) negative_gradient[n * 6 + k] = g[k];
}

}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_pixel_vertex_anchor_jacobians(
		const float* rasterized_vertex_position_jacobians, const float* rasterized_vertex_normal_jacobians,
		const float* warped_vertex_position_jacobians, const float* warped_vertex_normal_jacobians, const float* point_map_vectors,
		const float* rasterized_normals, const uint8_t* residual_mask, const int64_t* pixel_faces, int32_t faces_per_pixel,
		const int64_t* triangles, const int32_t* face_anchor_entries, const int32_t* face_anchor_counts, int64_t pixel_count,
		int32_t anchor_count, int32_t node_count, int32_t iteration_mode, int32_t use_tukey_penalty, float tukey_penalty_cutoff,
		float* pixel_jacobians, int32_t* pixel_node_jacobian_counts, int32_t* node_pixel_jacobian_counts, void* stream) {
	NNRT_CHECK_ARG(rasterized_vertex_position_jacobians && rasterized_vertex_normal_jacobians && warped_vertex_position_jacobians &&
	               warped_vertex_normal_jacobians && point_map_vectors && rasterized_normals && residual_mask && pixel_faces && triangles &&
	               face_anchor_entries && face_anchor_counts && pixel_jacobians && pixel_node_jacobian_counts && node_pixel_jacobian_counts,
	               "pixel_vertex_anchor_jacobians: null pointer argument");
	// PixelVertexAnchorJacobiansImpl.h:77-80
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "Supplied anchor count per vertex, %d, exceeds the set maximum, %d", anchor_count, kMaxAnchors);
	NNRT_CHECK_ARG(iteration_mode >= 0 && iteration_mode <= 2, "iteration mode %d is not one of ALL (0), TRANSLATION_ONLY (1), ROTATION_ONLY (2)", iteration_mode);
	NNRT_CHECK_ARG(faces_per_pixel >= 1 && node_count >= 1 && pixel_count >= 0, "pixel_vertex_anchor_jacobians: bad counts");
	auto s = (cudaStream_t) stream;
	NNRT_CUDA(cudaMemsetAsync(node_pixel_jacobian_counts, 0, sizeof(int32_t) * node_count, s));
	if (pixel_count == 0) return kOk;
	pixel_vertex_anchor_jacobians_kernel<<<blocks_for(pixel_count, 128), 128, 0, s>>>(
			rasterized_vertex_position_jacobians, rasterized_vertex_normal_jacobians, warped_vertex_position_jacobians, warped_vertex_normal_jacobians,
			point_map_vectors, rasterized_normals, residual_mask, pixel_faces, faces_per_pixel, triangles, face_anchor_entries, face_anchor_counts,
			pixel_count, anchor_count, node_count, iteration_mode, use_tukey_penalty, tukey_penalty_cutoff, pixel_jacobians,
			pixel_node_jacobian_counts, node_pixel_jacobian_counts);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_arap_term(const int32_t* edges, int64_t edge_count, const int8_t* edge_layer_indices, const float* layer_decimation_radii,
                                   const float* node_coverage_weights, const float* nodes, const float* node_rotations,
                                   const float* node_translations, int32_t node_count, float arap_term_weight, int32_t use_huber_penalty,
                                   float huber_penalty_constant, float* edge_residuals, float* edge_jacobians, float* hessian_diagonal_blocks,
                                   float* upper_wing_blocks, int32_t* upper_wing_block_coordinates, float* negative_gradient, void* stream) {
	NNRT_CHECK_ARG(edges && nodes && node_rotations && node_translations, "arap_term: null pointer argument");
	NNRT_CHECK_ARG((edge_layer_indices && layer_decimation_radii) || node_coverage_weights,
	               "arap_term: need edge_layer_indices + layer_decimation_radii (fixed coverage) or node_coverage_weights (variable coverage)");
	NNRT_CHECK_ARG(node_count >= 1 && edge_count >= 0, "arap_term: bad counts");
	auto s = (cudaStream_t) stream;
	const ArapInputs in{edges, edge_count, edge_layer_indices, layer_decimation_radii, node_coverage_weights, nodes, node_rotations, node_translations,
	                    arap_term_weight, use_huber_penalty, huber_penalty_constant};
	if (edge_count > 0 && (edge_residuals || edge_jacobians || upper_wing_blocks || upper_wing_block_coordinates))
		arap_edge_kernel<<<blocks_for(edge_count, 128), 128, 0, s>>>(in, edge_residuals, edge_jacobians, upper_wing_blocks, upper_wing_block_coordinates);
	if (hessian_diagonal_blocks || negative_gradient)
		arap_node_kernel<<<blocks_for(node_count, 128), 128, 0, s>>>(in, node_count, hessian_diagonal_blocks, negative_gradient);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
