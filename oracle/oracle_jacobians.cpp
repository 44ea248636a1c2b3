// ORACLE (test infrastructure only) — rasterised-surface Jacobians, face/anchor association,
// per-pixel node Jacobians, data-term J^T J blocks and -J^T r. See oracle_common.h for status.
#include "oracle_common.h"

using namespace oracle;

namespace {
inline float SignedParallelogramArea(const V2& p, const V2& v0, const V2& v1) {
	return (p.x - v0.x) * (v0.y - v1.y) - (p.y - v0.y) * (v0.x - v1.x);
}

// cpp/alignment/functional/kernel/BarycentricCoordinateJacobians.h:43-160 (ClockWise).
// d_ndc[i_vertex] is 3x2 row-major: d rho_r / d p_{i_vertex}.
void BaryWrtNdc(float d_ndc[3][3][2], const V2& ray, const V2& v0, const V2& v1, const V2& v2,
                float face_area, const float sub_area[3]) {
	const float area_sq = face_area * face_area;
	// Jacobian_SignedFaceParallelogramAreaWrt2dVertices :43-61
	const V2 dA[3] = {{v1.y - v2.y, v2.x - v1.x}, {v2.y - v0.y, v0.x - v2.x}, {v0.y - v1.y, v1.x - v0.x}};
	// Jacobian_SignedSubFaceParallelogramAreaWrtIntersectionPointAndNdcVertices :63-80, called with
	// (ray, v1, v2), (ray, v2, v0), (ray, v0, v1): returns (d/d first vertex, d/d second vertex)
	auto sub_jac = [&](const V2& a, const V2& b, V2& d_first, V2& d_second) {
		d_first = {b.y - ray.y, ray.x - b.x};
		d_second = {ray.y - a.y, a.x - ray.x};
	};
	V2 s0_v1, s0_v2, s1_v2, s1_v0, s2_v0, s2_v1;
	sub_jac(v1, v2, s0_v1, s0_v2);
	sub_jac(v2, v0, s1_v2, s1_v0);
	sub_jac(v0, v1, s2_v0, s2_v1);
	const float den = area_sq + K_EPSILON;
	auto q = [&](float A_f, const V2& dsub, float sub, const V2& dface, float* out) {
		out[0] = (A_f * dsub.x - sub * dface.x) / den;
		out[1] = (A_f * dsub.y - sub * dface.y) / den;
	};
	auto q0 = [&](float sub, const V2& dface, float* out) {
		out[0] = (-sub * dface.x) / den;
		out[1] = (-sub * dface.y) / den;
	};
	// rows: rho0, rho1, rho2 ; [i_vertex][row][xy]
	q0(sub_area[0], dA[0], d_ndc[0][0]);
	q(face_area, s0_v1, sub_area[0], dA[1], d_ndc[1][0]);
	q(face_area, s0_v2, sub_area[0], dA[2], d_ndc[2][0]);
	q(face_area, s1_v0, sub_area[1], dA[0], d_ndc[0][1]);
	q0(sub_area[1], dA[1], d_ndc[1][1]);
	q(face_area, s1_v2, sub_area[1], dA[2], d_ndc[2][1]);
	q(face_area, s2_v0, sub_area[2], dA[0], d_ndc[0][2]);
	q(face_area, s2_v1, sub_area[2], dA[1], d_ndc[1][2]);
	q0(sub_area[2], dA[2], d_ndc[2][2]);
}

// BarycentricCoordinateJacobians.h:289-419 -> d rho / d V, 3x9 row-major
void BaryWrtCameraVertices(float out[3][9], const V2& ray, const V3 v[3], const TransformIndexer& proj,
                           bool perspective_correct, const float* stored_bary) {
	V2 ndc[3];
	for (int k = 0; k < 3; k++) proj.Project(v[k].x, v[k].y, v[k].z, &ndc[k].x, &ndc[k].y);
	float face_area = SignedParallelogramArea(ndc[0], ndc[1], ndc[2]) + K_EPSILON;
	float sub_area[3];
	float distorted[3];
	if (perspective_correct) {  // BarycentricCoordinates_PreserveAreas, BarycentricCoordinates.h:66-85
		sub_area[0] = SignedParallelogramArea(ray, ndc[1], ndc[2]);
		sub_area[1] = SignedParallelogramArea(ray, ndc[2], ndc[0]);
		sub_area[2] = SignedParallelogramArea(ray, ndc[0], ndc[1]);
		for (int k = 0; k < 3; k++) distorted[k] = sub_area[k] / face_area;
	} else {  // sub areas recovered from the stored (distorted) barycentrics :169-176
		for (int k = 0; k < 3; k++) sub_area[k] = stored_bary[k] * face_area;
	}
	float d_ndc[3][3][2];
	BaryWrtNdc(d_ndc, ray, ndc[0], ndc[1], ndc[2], face_area, sub_area);
	const float fxn = proj.fx, fyn = proj.fy;
	for (int iv = 0; iv < 3; iv++) {
		// CameraToNdcSpaceProjectionJacobian, ProjectionJacobians.h:26-39 (2x3)
		const float z = v[iv].z, z_sq = z * z;
		const float P[2][3] = {{fxn / z, 0.f, -fxn * v[iv].x / z_sq}, {0.f, fyn / z, -fyn * v[iv].y / z_sq}};
		for (int r = 0; r < 3; r++)
			for (int c = 0; c < 3; c++)
				out[r][iv * 3 + c] = d_ndc[iv][r][0] * P[0][c] + d_ndc[iv][r][1] * P[1][c];
	}
	if (perspective_correct) {  // :203-282 and :369-389
		const float z0 = v[0].z, z1 = v[1].z, z2 = v[2].z;
		const float v1z_x_v2z = z1 * z2, v0z_x_v2z = z0 * z2, v0z_x_v1z = z0 * z1;
		const float n0 = distorted[0] * v1z_x_v2z, n1 = distorted[1] * v0z_x_v2z, n2 = distorted[2] * v0z_x_v1z;
		const float den = fmaxf(n0 + n1 + n2, K_EPSILON);
		const float den_sq = den * den;
		float Dd[3][3] = {{(den - n0) * v1z_x_v2z, -n0 * v0z_x_v2z, -n0 * v0z_x_v1z},
		                  {-n1 * v1z_x_v2z, (den - n1) * v0z_x_v2z, -n1 * v0z_x_v1z},
		                  {-n2 * v1z_x_v2z, -n2 * v0z_x_v2z, (den - n2) * v0z_x_v1z}};
		for (auto& row : Dd) for (float& e : row) e /= den_sq;
		const float pz0 = distorted[1] * z2 + z1 * distorted[2];
		const float pz1 = distorted[0] * z2 + z0 * distorted[2];
		const float pz2 = distorted[0] * z1 + z0 * distorted[1];
		float Dz[3][3] = {{-n0 * pz0, den * distorted[0] * z2 - n0 * pz1, den * distorted[0] * z1 - n0 * pz2},
		                  {den * distorted[1] * z2 - n1 * pz0, -n1 * pz1, den * distorted[1] * z0 - n1 * pz2},
		                  {den * distorted[2] * z1 - n2 * pz0, den * distorted[2] * z0 - n2 * pz1, -n2 * pz2}};
		for (auto& row : Dz) for (float& e : row) e /= den_sq;
		float tmp[3][9];
		for (int r = 0; r < 3; r++)
			for (int c = 0; c < 9; c++)
				tmp[r][c] = sum3(Dd[r][0] * out[0][c], Dd[r][1] * out[1][c], Dd[r][2] * out[2][c]);
		for (int r = 0; r < 3; r++) {
			for (int c = 0; c < 9; c++) out[r][c] = tmp[r][c];
			for (int iv = 0; iv < 3; iv++) out[r][iv * 3 + 2] += Dz[r][iv];
		}
	}
}
} // namespace

// cpp/alignment/functional/kernel/RasterizedSurfaceJacobiansImpl.h:43-201 (normal Jacobians on).
// Only face slot k=0 of each pixel is used (:125). Outputs [PX,3,9] and [PX,3,10] (zeros if uncovered).
ORACLE_API void oracle_rasterized_surface_jacobians(
		const float* warped_positions, const float* warped_normals, const int64_t* triangles,
		const int64_t* pixel_faces, const float* pixel_bary, int height, int width, int faces_per_pixel,
		const double* ndc_intrinsics33, int perspective_correct, int parallel,
		float* rendered_vertex_jacobians, float* rendered_normal_jacobians) {
	TransformIndexer proj(ndc_intrinsics33, IDENTITY4, 1.0f);
	const int64_t px_count = (int64_t) height * width;
	memset(rendered_vertex_jacobians, 0, sizeof(float) * px_count * 27);
	memset(rendered_normal_jacobians, 0, sizeof(float) * px_count * 30);
#pragma omp parallel for schedule(static) if (parallel)
	for (int64_t i = 0; i < px_count; i++) {
		const int v_image = (int) (i / width), u_image = (int) (i % width);
		V2 ray = {PixelToNdc(u_image, width, height), PixelToNdc(v_image, height, width)};
		int64_t f = pixel_faces[i * faces_per_pixel];
		if (f == -1) continue;
		V3 fv[3], fn[3];
		for (int k = 0; k < 3; k++) {
			fv[k] = load3(warped_positions + triangles[f * 3 + k] * 3);
			fn[k] = load3(warped_normals + triangles[f * 3 + k] * 3);
		}
		const float* bary = pixel_bary + i * faces_per_pixel * 3;
		float dB[3][9];
		BaryWrtCameraVertices(dB, ray, fv, proj, perspective_correct != 0, bary);
		float* out_v = rendered_vertex_jacobians + i * 27;
		float* out_n = rendered_normal_jacobians + i * 30;
		const float Vm[3][3] = {{fv[0].x, fv[1].x, fv[2].x}, {fv[0].y, fv[1].y, fv[2].y}, {fv[0].z, fv[1].z, fv[2].z}};
		const float Nm[3][3] = {{fn[0].x, fn[1].x, fn[2].x}, {fn[0].y, fn[1].y, fn[2].y}, {fn[0].z, fn[1].z, fn[2].z}};
		for (int r = 0; r < 3; r++) {
			for (int c = 0; c < 9; c++) {
				float val = sum3(Vm[r][0] * dB[0][c], Vm[r][1] * dB[1][c], Vm[r][2] * dB[2][c]);
				// + kron(rho, I3): entry (r, 3*k + r) += rho_k   (core/linalg/KroneckerTensorProduct.h)
				if (c % 3 == r) val += bary[c / 3];
				out_v[r * 9 + c] = val;
				out_n[r * 9 + c] = sum3(Nm[r][0] * dB[0][c], Nm[r][1] * dB[1][c], Nm[r][2] * dB[2][c]);
			}
		}
		out_n[27] = bary[0]; out_n[28] = bary[1]; out_n[29] = bary[2];
	}
}

// cpp/alignment/functional/FaceNodeAnchors.h:27-31 restated as a packed POD for ctypes:
// node_index, vertices[3] (int64 in the reference; int32 here), vertex_anchor_indices[3]
struct FaceNodeAnchor {
	int32_t node_index;
	int32_t vertices[3];
	int32_t vertex_anchor_indices[3];
};

// cpp/alignment/functional/kernel/AssociateFacesWithAnchorsImpl.h:33-107.
// out_face_anchors: [F, 3*MAX_ANCHOR_COUNT] FaceNodeAnchor (7 int32 each), out_counts [F].
ORACLE_API void oracle_associate_faces_with_anchors(
		const int64_t* triangles, int64_t face_count, const int32_t* vertex_anchors, int anchor_count,
		int32_t* out_face_anchors, int32_t* out_counts) {
	const int stride = 3 * MAX_ANCHOR_COUNT;
	const int max_face_anchor_count = 3 * anchor_count;
	auto* out = reinterpret_cast<FaceNodeAnchor*>(out_face_anchors);
	for (int64_t f = 0; f < face_count; f++) {
		FaceNodeAnchor cur[3 * MAX_ANCHOR_COUNT];
		for (auto& e : cur) e = {-1, {-1, -1, -1}, {-1, -1, -1}};
		int count = 0;
		for (int fv = 0; fv < 3; fv++) {
			int64_t i_vertex = triangles[f * 3 + fv];
			for (int a = 0; a < anchor_count; a++) {
				int node = vertex_anchors[i_vertex * anchor_count + a];
				if (node == -1) continue;
				int k = 0;
				int inspected = cur[k].node_index;
				while (inspected != node && inspected != -1 && k + 1 < max_face_anchor_count) {
					k++;
					inspected = cur[k].node_index;
				}
				if (inspected != node) { count++; cur[k].node_index = node; }
				cur[k].vertices[fv] = (int32_t) i_vertex;
				cur[k].vertex_anchor_indices[fv] = a;
			}
		}
		out_counts[f] = count;
		memcpy(out + f * stride, cur, sizeof(cur));
	}
}

// cpp/alignment/functional/kernel/PixelVertexAnchorJacobiansImpl.h:40-365, IterationMode::ALL
// (6 floats per (pixel, face-anchor): rotation xyz, translation xyz).
// Node->pixel lists: appended in ascending pixel order (single-threaded); max_pixels_per_node <= 0
// lifts the reference's MAX_PIXELS_PER_NODE=4000 cap (SURVEY.md §8(c) oracle policy).
// node_pixel_lists is a caller-provided [N, list_capacity] int32 array; returns the number of dropped entries.
ORACLE_API int64_t oracle_pixel_vertex_anchor_jacobians(
		const float* rasterized_vertex_position_jacobians /*[PX,27]*/, const float* rasterized_vertex_normal_jacobians /*[PX,30]*/,
		const float* warped_vertex_position_jacobians /*[V,A,4]*/, const float* warped_vertex_normal_jacobians /*[V,A,3]*/,
		const float* point_map_vectors /*[PX,3]*/, const float* rasterized_normals /*[PX,3]*/, const uint8_t* residual_mask,
		const int64_t* pixel_faces, int faces_per_pixel, const int32_t* face_anchors, const int32_t* face_anchor_counts,
		int64_t pixel_count, int anchor_count, int node_count, int use_tukey, float tukey_cutoff,
		int64_t list_capacity, int max_pixels_per_node,
		float* pixel_jacobians /*[PX,3A,6]*/, int32_t* pixel_node_jacobian_counts /*[PX]*/,
		int32_t* node_pixel_lists /*[N,list_capacity]*/, int32_t* node_pixel_counts /*[N]*/) {
	const int S = 6;
	const auto* fa_all = reinterpret_cast<const FaceNodeAnchor*>(face_anchors);
	memset(pixel_jacobians, 0, sizeof(float) * pixel_count * anchor_count * 3 * S);
	memset(pixel_node_jacobian_counts, 0, sizeof(int32_t) * pixel_count);
	for (int64_t i = 0; i < (int64_t) node_count * list_capacity; i++) node_pixel_lists[i] = -1;
	memset(node_pixel_counts, 0, sizeof(int32_t) * node_count);
	int64_t dropped = 0;
	for (int64_t px = 0; px < pixel_count; px++) {
		if (!residual_mask[px]) continue;
		V3 wl_minus_ol = load3(point_map_vectors + px * 3);
		V3 nl = load3(rasterized_normals + px * 3);
		V3 dr_dnl, dr_dwl;
		if (use_tukey) {
			float r = dot3(nl, wl_minus_ol);
			if (fabsf(r) > tukey_cutoff) continue;
			float q = r / tukey_cutoff;
			float psi = 1 - q * q;
			psi = r * psi * psi;
			dr_dnl = scale(psi, wl_minus_ol);
			dr_dwl = scale(psi, nl);
		} else {
			dr_dnl = wl_minus_ol;
			dr_dwl = nl;
		}
		const float* dwl_dV = rasterized_vertex_position_jacobians + px * 27;
		const float* dnl_dV = rasterized_vertex_normal_jacobians + px * 30;
		const float* rho = dnl_dV + 27;
		float dr_dV[9], dr_dN[9];
		for (int c = 0; c < 9; c++) {
			float a = sum3(dr_dwl.x * dwl_dV[c], dr_dwl.y * dwl_dV[9 + c], dr_dwl.z * dwl_dV[18 + c]);
			float b = sum3(dr_dnl.x * dnl_dV[c], dr_dnl.y * dnl_dV[9 + c], dr_dnl.z * dnl_dV[18 + c]);
			dr_dV[c] = a + b;
			// dr_dnl * kron(rho, I3): the two zero products are exact zeros
			const float comp = (c % 3 == 0) ? dr_dnl.x : (c % 3 == 1) ? dr_dnl.y : dr_dnl.z;
			dr_dN[c] = comp * rho[c / 3];
		}
		const int64_t f = pixel_faces[px * faces_per_pixel];
		const int list_address = (int) px * (anchor_count * 3 * S);
		float* pj = pixel_jacobians + list_address;
		const FaceNodeAnchor* fa = fa_all + f * (3 * MAX_ANCHOR_COUNT);
		const int fa_count = face_anchor_counts[f];
		pixel_node_jacobian_counts[px] = fa_count;
		for (int k = 0; k < fa_count; k++) {
			const int node = fa[k].node_index;
			float* J_rot = pj + k * S;
			float* J_tr = pj + k * S + 3;
			for (int fv = 0; fv < 3; fv++) {
				const int i_vertex = fa[k].vertices[fv];
				if (i_vertex == -1) continue;
				const int a = fa[k].vertex_anchor_indices[fv];
				const float* wj = warped_vertex_position_jacobians + ((int64_t) i_vertex * anchor_count + a) * 4;
				const float* nj = warped_vertex_normal_jacobians + ((int64_t) i_vertex * anchor_count + a) * 3;
				const float w = wj[3];
				const V3 dr_dv = {dr_dV[fv * 3], dr_dV[fv * 3 + 1], dr_dV[fv * 3 + 2]};
				const V3 dr_dn = {dr_dN[fv * 3], dr_dN[fv * 3 + 1], dr_dN[fv * 3 + 2]};
				// translation: dr_dv * (I * w): row-vector times scaled identity, off-diagonal products are +-0
				J_tr[0] += sum3(dr_dv.x * w, dr_dv.y * 0.f, dr_dv.z * 0.f);
				J_tr[1] += sum3(dr_dv.x * 0.f, dr_dv.y * w, dr_dv.z * 0.f);
				J_tr[2] += sum3(dr_dv.x * 0.f, dr_dv.y * 0.f, dr_dv.z * w);
				// rotation: row * skew(s) with skew = [[0,-sz,sy],[sz,0,-sx],[-sy,sx,0]]
				auto row_times_skew = [](const V3& r, const float* s, float* o) {
					o[0] = sum3(r.x * 0.f, r.y * s[2], r.z * -s[1]);
					o[1] = sum3(r.x * -s[2], r.y * 0.f, r.z * s[0]);
					o[2] = sum3(r.x * s[1], r.y * -s[0], r.z * 0.f);
				};
				float pv[3], pn[3];
				row_times_skew(dr_dv, wj, pv);
				row_times_skew(dr_dn, nj, pn);
				for (int c = 0; c < 3; c++) J_rot[c] += pv[c] + pn[c];
			}
			int slot = node_pixel_counts[node];
			if (max_pixels_per_node > 0 && slot > max_pixels_per_node) { dropped++; continue; }  // sic: '>' (…Impl.h:350)
			if (slot >= list_capacity) { dropped++; continue; }
			node_pixel_lists[(int64_t) node * list_capacity + slot] = list_address + k * S;
			node_pixel_counts[node] = slot + 1;
		}
	}
	return dropped;
}

// cpp/alignment/kernel/DeformableMeshToImageFitterImpl.h:199-317 (ALL): per node, 21 unique
// column products accumulated in list order, mirrored into a 6x6 block.
ORACLE_API void oracle_depth_hessian_blocks(
		const float* pixel_jacobians, const int32_t* node_pixel_lists, const int32_t* node_pixel_counts,
		int node_count, int64_t list_capacity, int parallel, float* hessian_blocks /*[N,6,6]*/) {
#pragma omp parallel for schedule(dynamic, 1) if (parallel)
	for (int n = 0; n < node_count; n++) {
		const int len = node_pixel_counts[n];
		for (int c0 = 0; c0 < 6; c0++) {
			for (int c1 = c0; c1 < 6; c1++) {
				float acc = 0.0f;
				for (int k = 0; k < len; k++) {
					const float* J = pixel_jacobians + node_pixel_lists[(int64_t) n * list_capacity + k];
					acc += J[c0] * J[c1];
				}
				hessian_blocks[n * 36 + c0 * 6 + c1] = acc;
				hessian_blocks[n * 36 + c1 * 6 + c0] = acc;
			}
		}
	}
}

// cpp/alignment/kernel/DeformableMeshToImageFitterImpl.h:340-456 (ALL): -sum_px J * r
ORACLE_API void oracle_negative_depth_gradient(
		const float* residuals, const uint8_t* residual_mask, const float* pixel_jacobians,
		const int32_t* node_pixel_lists, const int32_t* node_pixel_counts, int node_count, int64_t list_capacity,
		int anchor_count, int parallel, float* negative_gradient /*[N*6]*/) {
#pragma omp parallel for schedule(dynamic, 1) if (parallel)
	for (int n = 0; n < node_count; n++) {
		float g[6] = {0, 0, 0, 0, 0, 0};
		const int len = node_pixel_counts[n];
		for (int k = 0; k < len; k++) {
			const int address = node_pixel_lists[(int64_t) n * list_capacity + k];
			const int px = address / (anchor_count * 3 * 6);
			if (!residual_mask[px]) continue;
			const float* J = pixel_jacobians + address;
			const float r = residuals[px];
			for (int c = 0; c < 6; c++) g[c] += J[c] * r;
		}
		for (int c = 0; c < 6; c++) negative_gradient[n * 6 + c] = 0.f - g[c];
	}
}
