// ORACLE (test infrastructure only) — the GN/LM loop of
// nnrt::alignment::DeformableMeshToImageFitter::FitToImage
// (cpp/alignment/DeformableMeshToImageFitter.cpp:85-276, ComputeDepthResiduals :331-390,
// ComputeEdgeResiduals :392-448), keeping the reference's data flow: dense [PX,27], [PX,30],
// [PX,3A,6] temporaries and node->pixel lists (MAX_PIXELS_PER_NODE cap lifted unless asked for).
// IterationMode::ALL only; node order = virtual order with identity permutation.
#include "oracle_common.h"
#include <cstdio>

using namespace oracle;

extern "C" {
void oracle_warp_points_normals(const float*, const float*, int64_t, const float*, const float*, const float*, const int32_t*,
                                const float*, int, int, const double*, float*, float*);
void oracle_intrinsics_to_ndc(const double*, int, int, double*, float*);
void oracle_face_vertices_ndc(const float*, const int64_t*, int64_t, const double*, int, int, float, float, float*, uint8_t*);
int64_t oracle_rasterize_ndc_triangles(const float*, const uint8_t*, int64_t, int, int, float, int, int, int, int, int, int, int,
                                       int64_t*, float*, float*, float*);
void oracle_unproject_depth(const float*, int, int, const double*, const double*, float, float, float*, uint8_t*);
void oracle_point_to_plane_distances(const float*, const float*, const float*, int64_t, float*);
void oracle_warped_surface_jacobians(const float*, const float*, int64_t, const float*, const float*, const int32_t*, const float*, int,
                                     float*, float*);
void oracle_rasterized_surface_jacobians(const float*, const float*, const int64_t*, const int64_t*, const float*, int, int, int,
                                         const double*, int, int, float*, float*);
void oracle_associate_faces_with_anchors(const int64_t*, int64_t, const int32_t*, int, int32_t*, int32_t*);
int64_t oracle_pixel_vertex_anchor_jacobians(const float*, const float*, const float*, const float*, const float*, const float*,
                                             const uint8_t*, const int64_t*, int, const int32_t*, const int32_t*, int64_t, int, int, int,
                                             float, int64_t, int, float*, int32_t*, int32_t*, int32_t*);
void oracle_depth_hessian_blocks(const float*, const int32_t*, const int32_t*, int, int64_t, int, float*);
void oracle_negative_depth_gradient(const float*, const uint8_t*, const float*, const int32_t*, const int32_t*, int, int64_t, int, int,
                                    float*);
void oracle_arap_residuals(const int32_t*, int64_t, const int8_t*, const float*, const float*, const float*, const float*, const float*,
                           float, float*);
void oracle_arap_edge_jacobians(const int32_t*, int64_t, const int8_t*, const float*, const float*, const float*, const float*, float,
                                float*);
int oracle_arap_hessian(const int32_t*, int64_t, const float*, int, int, float*, float*, int32_t*);
void oracle_negative_arap_gradient(const float*, const float*, const int32_t*, int64_t, int, float*);
void oracle_precondition_diagonal_blocks(float*, int, int, float);
int oracle_solve_block_diagonal_cholesky(const float*, const float*, int, int, float*);
int oracle_solve_arrowhead_cholesky(const float*, int, int, const float*, const int32_t*, int64_t, const float*, const int32_t*, int64_t,
                                    const float*, float*);
void oracle_apply_node_updates(const float*, int, float*, float*);
}

// Plain-C parameter block shared with ctypes (tests/oracle_bindings.py mirrors the field order).
struct OracleFitParams {
	// mesh
	const float* vertex_positions; const float* vertex_normals; const int64_t* triangles; int64_t vertex_count; int64_t face_count;
	// graph (virtual node order)
	const float* nodes; int32_t node_count; const int32_t* edges; int64_t edge_count; int32_t arrow_base;
	const int8_t* edge_layer_indices; const float* layer_decimation_radii; const float* node_coverage_weights;
	// anchors
	const int32_t* anchors; const float* anchor_weights; int32_t anchor_count;
	// observation
	const float* reference_points; const uint8_t* reference_mask; int32_t height; int32_t width;
	const double* intrinsics; const double* extrinsics;
	// settings (DeformableMeshToImageFitter ctor, cpp/alignment/DeformableMeshToImageFitter.h:34-46)
	int32_t iteration_count; int32_t use_perspective_correction; float max_depth; int32_t use_tukey; float tukey_cutoff;
	float levenberg_marquardt_factor; float arap_term_weight; int32_t use_huber; float huber_constant;
	int32_t max_pixels_per_node;  // <= 0: cap lifted
	int32_t parallel;
	// state in/out
	float* node_rotations; float* node_translations;
	// optional traces, each may be null: [iters,N,36], [iters,N,6], [iters,N,6]; last-iteration fragments/residuals
	float* trace_hessian_diag; float* trace_negative_gradient; float* trace_deltas;
	int64_t* last_pixel_faces; float* last_residuals; uint8_t* last_residual_mask;
	double* stage_seconds;  // optional [8]: warp, faces, raster, residual, jacobians, assembly, arap, solve+update
	// IterationMode sequence (cpp/alignment/IterationMode.h: 0 ALL, 1 TRANSLATION_ONLY, 2 ROTATION_ONLY), cycled by
	// iteration % count (DeformableMeshToImageFitter.cpp:112); null / 0 = {ALL}
	const int32_t* iteration_modes; int32_t iteration_mode_count;
};

#include <chrono>
static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

ORACLE_API int oracle_fit_to_image(OracleFitParams* p) {
	const int64_t V = p->vertex_count, F = p->face_count;
	const int N = p->node_count, A = p->anchor_count, H = p->height, W = p->width;
	const int64_t PX = (int64_t) H * W, E = p->edge_count;
	const int par = p->parallel;
	double ndc_K[9]; float ndc_range[4];
	oracle_intrinsics_to_ndc(p->intrinsics, H, W, ndc_K, ndc_range);
	std::vector<int32_t> face_anchors((size_t) F * 3 * MAX_ANCHOR_COUNT * 7), face_anchor_counts((size_t) F);
	oracle_associate_faces_with_anchors(p->triangles, F, p->anchors, A, face_anchors.data(), face_anchor_counts.data());
	const bool use_edges = E > 0;

	std::vector<float> wpos((size_t) V * 3), wnrm((size_t) V * 3), face_ndc((size_t) F * 9);
	std::vector<uint8_t> clip_mask((size_t) F);
	std::vector<int64_t> pix_face((size_t) PX);
	std::vector<float> pix_depth((size_t) PX), pix_bary((size_t) PX * 3), pix_dist((size_t) PX);
	std::vector<float> face_normals((size_t) F * 9), rendered_normals((size_t) PX * 3), raster_points((size_t) PX * 3);
	std::vector<uint8_t> rendered_mask((size_t) PX), residual_mask((size_t) PX);
	std::vector<float> residuals((size_t) PX), point_map((size_t) PX * 3);
	std::vector<float> wv_jac((size_t) V * A * 4), wn_jac((size_t) V * A * 3);
	std::vector<float> rv_jac((size_t) PX * 27), rn_jac((size_t) PX * 30);
	std::vector<float> pixel_jac((size_t) PX * 3 * A * 6);
	std::vector<int32_t> pixel_counts((size_t) PX), node_counts((size_t) N);
	std::vector<int32_t> node_lists;
	std::vector<float> Hd((size_t) N * 36), g((size_t) N * 6), delta((size_t) N * 6);
	std::vector<float> edge_res((size_t) E * 3), edge_jac((size_t) E * 5), Hdiag((size_t) N * 36), wing((size_t) E * 36), g_arap((size_t) N * 6);
	std::vector<int32_t> wing_coord((size_t) E * 2);
	double st[8] = {0, 0, 0, 0, 0, 0, 0, 0};

	for (int iteration = 0; iteration < p->iteration_count; iteration++) {
		const int mode = (p->iteration_modes && p->iteration_mode_count > 0) ? p->iteration_modes[iteration % p->iteration_mode_count] : 0;
		// The reference's ARAP Hessian only exists for 6 x 6 blocks (ArapHessianImpl.h:54-56) while the depth blocks of a
		// TRANSLATION_ONLY / ROTATION_ONLY iteration are 3 x 3 (.cpp:198, kernel/DeformableMeshToImageFitterImpl.h:219-226): their sum
		// at .cpp:235 cannot be formed, so those modes are defined without regularisation edges only.
		if (mode != 0 && use_edges) return -3;
		double t0 = now_s();
		oracle_warp_points_normals(p->vertex_positions, p->vertex_normals, V, p->nodes, p->node_rotations, p->node_translations,
		                           p->anchors, p->anchor_weights, A, 0, p->extrinsics, wpos.data(), wnrm.data());
		double t1 = now_s(); st[0] += t1 - t0;
		oracle_face_vertices_ndc(wpos.data(), p->triangles, F, p->intrinsics, H, W, 0.0f, 10.0f, face_ndc.data(), clip_mask.data());
		double t2 = now_s(); st[1] += t2 - t1;
		oracle_rasterize_ndc_triangles(face_ndc.data(), clip_mask.data(), F, H, W, 0.5f, 1, -1, 0x7fffffff, p->use_perspective_correction,
		                               0, 1, par, pix_face.data(), pix_depth.data(), pix_bary.data(), pix_dist.data());
		double t3 = now_s(); st[2] += t3 - t2;
		// ComputeDepthResiduals :331-390
		for (int64_t f = 0; f < F; f++)
			for (int k = 0; k < 3; k++)
				for (int c = 0; c < 3; c++) face_normals[f * 9 + k * 3 + c] = wnrm[p->triangles[f * 3 + k] * 3 + c];
		for (int64_t i = 0; i < PX; i++) {  // InterpolateVertexAttributes, K=1, C=3
			float* o = rendered_normals.data() + i * 3;
			o[0] = o[1] = o[2] = 0.f;
			int64_t f = pix_face[i];
			if (f < 0) continue;
			for (int c = 0; c < 3; c++) {
				float acc = 0.0f;
				for (int k = 0; k < 3; k++) acc += pix_bary[i * 3 + k] * face_normals[f * 9 + k * 3 + c];
				o[c] = acc;
			}
		}
		oracle_unproject_depth(pix_depth.data(), H, W, p->intrinsics, IDENTITY4, 1.0f, p->max_depth, raster_points.data(), rendered_mask.data());
		oracle_point_to_plane_distances(rendered_normals.data(), raster_points.data(), p->reference_points, PX, residuals.data());
		for (int64_t i = 0; i < PX; i++) {
			residual_mask[i] = p->reference_mask[i] && rendered_mask[i];
			if (!residual_mask[i]) residuals[i] = 0.0f;
		}
		if (p->use_tukey) {  // :377-389, including the inverted overwrite condition (distances <= c)
			const float c = p->tukey_cutoff, c2_6 = c * c / 6.f;
			for (int64_t i = 0; i < PX; i++) {
				float d = residuals[i], q = d / c, l = 1.f - q * q;
				float r = c2_6 * (1.f - l * l * l);
				if (d <= c) r = c2_6;
				residuals[i] = r;
			}
		}
		double t4 = now_s(); st[3] += t4 - t3;
		if (use_edges) {
			oracle_arap_residuals(p->edges, E, p->edge_layer_indices, p->layer_decimation_radii, p->node_coverage_weights, p->nodes,
			                      p->node_translations, p->node_rotations, p->arap_term_weight, edge_res.data());
			if (p->use_huber) {  // :435-444
				const float dlt = p->huber_constant, half = 0.5f * dlt * dlt;
				for (auto& r : edge_res) r = (r >= dlt) ? (fabsf(r) - half) : (0.5f * r * r);
			}
		}
		oracle_warped_surface_jacobians(p->vertex_positions, p->vertex_normals, V, p->nodes, p->node_rotations, p->anchors,
		                                p->anchor_weights, A, wv_jac.data(), wn_jac.data());
		oracle_rasterized_surface_jacobians(wpos.data(), wnrm.data(), p->triangles, pix_face.data(), pix_bary.data(), H, W, 1, ndc_K,
		                                    p->use_perspective_correction, par, rv_jac.data(), rn_jac.data());
		for (int64_t i = 0; i < PX * 3; i++) point_map[i] = raster_points[i] - p->reference_points[i];
		// exact list capacity (cap lifted): per-node count pre-pass
		std::fill(node_counts.begin(), node_counts.end(), 0);
		for (int64_t i = 0; i < PX; i++) {
			if (!residual_mask[i]) continue;
			const int64_t f = pix_face[i];
			for (int k = 0; k < face_anchor_counts[f]; k++) node_counts[face_anchors[((size_t) f * 3 * MAX_ANCHOR_COUNT + k) * 7]]++;
		}
		int64_t cap = 1;
		for (int c : node_counts) cap = std::max<int64_t>(cap, c);
		if (p->max_pixels_per_node > 0) cap = std::max<int64_t>(cap, p->max_pixels_per_node + 1);
		node_lists.resize((size_t) N * cap);
		oracle_pixel_vertex_anchor_jacobians(rv_jac.data(), rn_jac.data(), wv_jac.data(), wn_jac.data(), point_map.data(),
		                                     rendered_normals.data(), residual_mask.data(), pix_face.data(), 1, face_anchors.data(),
		                                     face_anchor_counts.data(), PX, A, N, p->use_tukey, p->tukey_cutoff, cap,
		                                     p->max_pixels_per_node, pixel_jac.data(), pixel_counts.data(), node_lists.data(),
		                                     node_counts.data());
		double t5 = now_s(); st[4] += t5 - t4;
		oracle_depth_hessian_blocks(pixel_jac.data(), node_lists.data(), node_counts.data(), N, cap, par, Hd.data());
		oracle_negative_depth_gradient(residuals.data(), residual_mask.data(), pixel_jac.data(), node_lists.data(), node_counts.data(), N, cap,
		                               A, par, g.data());
		double t6 = now_s(); st[5] += t6 - t5;
		int failures = 0;
		double t7;
		if (use_edges) {
			oracle_arap_edge_jacobians(p->edges, E, p->edge_layer_indices, p->layer_decimation_radii, p->node_coverage_weights, p->nodes,
			                           p->node_rotations, p->arap_term_weight, edge_jac.data());
			if (oracle_arap_hessian(p->edges, E, edge_jac.data(), p->arrow_base, N, Hdiag.data(), wing.data(), wing_coord.data()) != 0) return -2;
			for (size_t i = 0; i < Hdiag.size(); i++) Hdiag[i] += Hd[i];
			if (p->levenberg_marquardt_factor > 0.f) oracle_precondition_diagonal_blocks(Hdiag.data(), N, 6, p->levenberg_marquardt_factor);
			oracle_negative_arap_gradient(edge_res.data(), edge_jac.data(), p->edges, E, N, g_arap.data());
			for (size_t i = 0; i < g.size(); i++) g[i] += g_arap[i];
			t7 = now_s(); st[6] += t7 - t6;
			failures = oracle_solve_arrowhead_cholesky(Hdiag.data(), N, p->arrow_base, wing.data(), wing_coord.data(), E, nullptr, nullptr, 0,
			                                           g.data(), delta.data());
			if (p->trace_hessian_diag) std::copy(Hdiag.begin(), Hdiag.end(), p->trace_hessian_diag + (size_t) iteration * N * 36);
		} else if (mode == 0) {
			if (p->levenberg_marquardt_factor > 0.f) oracle_precondition_diagonal_blocks(Hd.data(), N, 6, p->levenberg_marquardt_factor);
			t7 = now_s();
			failures = oracle_solve_block_diagonal_cholesky(Hd.data(), g.data(), N, 6, delta.data());
			if (p->trace_hessian_diag) std::copy(Hd.begin(), Hd.end(), p->trace_hessian_diag + (size_t) iteration * N * 36);
		} else {
			// 3-wide pixel Jacobians hold exactly the translation (resp. rotation) half of the 6-wide ones
			// (PixelVertexAnchorJacobiansImpl.h:278-330: same addends, same order), and every block entry is its own sum over the node's
			// pixel list (kernel/DeformableMeshToImageFitterImpl.h:264-284), so the 3 x 3 blocks and 3-vectors are sub-blocks of Hd and g.
			const int off = mode == 1 ? 3 : 0;
			std::vector<float> H3((size_t) N * 9), g3((size_t) N * 3), d3((size_t) N * 3);
			for (int n = 0; n < N; n++)
				for (int r = 0; r < 3; r++) {
					g3[n * 3 + r] = g[n * 6 + off + r];
					for (int c = 0; c < 3; c++) H3[n * 9 + r * 3 + c] = Hd[(size_t) n * 36 + (off + r) * 6 + off + c];
				}
			if (p->levenberg_marquardt_factor > 0.f) oracle_precondition_diagonal_blocks(H3.data(), N, 3, p->levenberg_marquardt_factor);
			t7 = now_s();
			failures = oracle_solve_block_diagonal_cholesky(H3.data(), g3.data(), N, 3, d3.data());
			// traces keep the 6-wide layout: the inactive half is reported as zero
			std::fill(delta.begin(), delta.end(), 0.f);
			std::vector<float> Ht((size_t) N * 36, 0.f), gt((size_t) N * 6, 0.f);
			for (int n = 0; n < N; n++)
				for (int r = 0; r < 3; r++) {
					delta[n * 6 + off + r] = d3[n * 3 + r];
					gt[n * 6 + off + r] = g3[n * 3 + r];
					for (int c = 0; c < 3; c++) Ht[(size_t) n * 36 + (off + r) * 6 + off + c] = H3[n * 9 + r * 3 + c];
				}
			g = gt;
			if (p->trace_hessian_diag) std::copy(Ht.begin(), Ht.end(), p->trace_hessian_diag + (size_t) iteration * N * 36);
		}
		(void) failures;
		if (p->trace_negative_gradient) std::copy(g.begin(), g.end(), p->trace_negative_gradient + (size_t) iteration * N * 6);
		if (p->trace_deltas) std::copy(delta.begin(), delta.end(), p->trace_deltas + (size_t) iteration * N * 6);
		if (mode == 0) {
			oracle_apply_node_updates(delta.data(), N, p->node_rotations, p->node_translations);
		} else if (mode == 1) {  // .cpp:266-268: TranslateNodes only
			for (int i = 0; i < N * 3; i++) p->node_translations[i] = p->node_translations[i] + delta[(i / 3) * 6 + 3 + i % 3];
		} else {                 // .cpp:269-272: Rodrigues + RotateNodes only
			std::vector<float> saved(p->node_translations, p->node_translations + (size_t) N * 3);
			oracle_apply_node_updates(delta.data(), N, p->node_rotations, p->node_translations);
			std::copy(saved.begin(), saved.end(), p->node_translations);
		}
		st[7] += now_s() - t7;
	}
	if (p->last_pixel_faces) std::copy(pix_face.begin(), pix_face.end(), p->last_pixel_faces);
	if (p->last_residuals) std::copy(residuals.begin(), residuals.end(), p->last_residuals);
	if (p->last_residual_mask) std::copy(residual_mask.begin(), residual_mask.end(), p->last_residual_mask);
	if (p->stage_seconds) for (int i = 0; i < 8; i++) p->stage_seconds[i] = st[i];
	return 0;
}

// Thread control for the timed CPU baseline: launchers such as torchrun export OMP_NUM_THREADS=1, which would silently turn the
// "all host cores" baseline into a single-thread one.
#ifdef _OPENMP
#include <omp.h>
#endif
extern "C" __attribute__((visibility("default"))) int oracle_set_num_threads(int n) {
#ifdef _OPENMP
	if (n > 0) omp_set_num_threads(n);
	return omp_get_max_threads();
#else
	(void) n;
	return 1;
#endif
}
