// ORACLE VALIDATION SCAFFOLDING ONLY — never on the product path, never shipped as source of truth
// for anything but checking oracle/*.cpp. Includes the reference's Eigen-only device-function
// headers VERBATIM from /root/reference/cpp (nothing is copied into this repo) against the stub
// Open3D shim in ref_check/shim, and exposes them through a small C ABI for tests/test_oracle_vs_ref.py.
#include <cstdint>
#include <cmath>
#include <algorithm>
#include <Eigen/Dense>
#include <open3d/t/geometry/kernel/GeometryIndexer.h>
#include "core/kernel/KnnUtilities.h"
#include "geometry/functional/kernel/WarpUtilities.h"
#include "rendering/kernel/RayFaceIntersection.h"
#include "alignment/functional/kernel/BarycentricCoordinateJacobians.h"

#define REF_API extern "C" __attribute__((visibility("default")))
namespace o3tgk_ = open3d::t::geometry::kernel;
constexpr auto CPU = open3d::core::Device::DeviceType::CPU;

// anchors for a batch of points; variant: 0 fixed, 1 variable, 2 fixed+threshold, 3 variable+threshold
REF_API void ref_anchors(const float* points, int64_t n_points, float* nodes, int n_nodes, const float* node_weights,
                         float node_coverage_squared, int anchor_count, int min_valid, int variant, int32_t* anchors, float* weights) {
	o3tgk_::NDArrayIndexer node_indexer(nodes, 3 * sizeof(float));
	using namespace nnrt::geometry::functional::kernel::warp;
	for (int64_t i = 0; i < n_points; i++) {
		Eigen::Vector3f point(points[i * 3], points[i * 3 + 1], points[i * 3 + 2]);
		int32_t* a = anchors + i * anchor_count; float* w = weights + i * anchor_count;
		switch (variant) {
			case 0: FindAnchorsAndWeightsForPoint_Euclidean_FixedNodeCoverageWeight<CPU>(a, w, anchor_count, n_nodes, point, node_indexer, node_coverage_squared); break;
			case 1: FindAnchorsAndWeightsForPoint_Euclidean_VariableNodeCoverageWeight<CPU>(a, w, anchor_count, n_nodes, point, node_indexer, node_weights); break;
			case 2: FindAnchorsAndWeightsForPoint_Euclidean_Threshold_FixedNodeCoverageWeight<CPU>(a, w, anchor_count, min_valid, n_nodes, point, node_indexer, node_coverage_squared); break;
			case 3: FindAnchorsAndWeightsForPoint_Euclidean_Threshold_VariableNodeCoverageWeight<CPU>(a, w, anchor_count, min_valid, n_nodes, point, node_indexer, node_weights); break;
		}
	}
}

REF_API void ref_blend_warp(const float* points, const float* normals, int64_t n_points, float* nodes, float* rotations, float* translations,
                            const int32_t* anchors, const float* weights, int anchor_count, float* out_points, float* out_normals) {
	const o3tgk_::NDArrayIndexer node_indexer(nodes, 3 * sizeof(float)), rot_indexer(rotations, 9 * sizeof(float)), tr_indexer(translations, 3 * sizeof(float));
	for (int64_t i = 0; i < n_points; i++) {
		Eigen::Vector3f p(points[i * 3], points[i * 3 + 1], points[i * 3 + 2]), n(normals[i * 3], normals[i * 3 + 1], normals[i * 3 + 2]);
		out_points[i * 3] = out_points[i * 3 + 1] = out_points[i * 3 + 2] = 0.f;
		out_normals[i * 3] = out_normals[i * 3 + 1] = out_normals[i * 3 + 2] = 0.f;
		Eigen::Map<Eigen::Vector3f> wp(out_points + i * 3), wn(out_normals + i * 3);
		nnrt::geometry::functional::kernel::warp::BlendWarp(wp, wn, anchors + i * anchor_count, weights + i * anchor_count, anchor_count, p, n,
		                                                    node_indexer, rot_indexer, tr_indexer);
	}
}

// brute-force rasterisation of every pixel against every face through the reference's per-pixel routine
REF_API void ref_rasterize_bruteforce(float* face_vertices, int64_t n_faces, int height, int width, float blur_radius_ndc, int faces_per_pixel,
                                      int perspective_correct, int clip_bary, int cull_back, int64_t* out_faces, float* out_depths,
                                      float* out_bary, float* out_dist) {
	using namespace nnrt::rendering::kernel;
	o3tgk_::TArrayIndexer<t_face_index> indexer(face_vertices, 9 * sizeof(float));
	for (int64_t i = 0; i < (int64_t) height * width; i++) {
		int v = (int) (i / width), u = (int) (i % width);
		const float y = ImageSpacePixelAlongDimensionToNdc(v, height, width);
		const float x = ImageSpacePixelAlongDimensionToNdc(u, width, height);
		Eigen::Vector2f pixel(x, y);
		RayFaceIntersection queue[8];
		int queue_size = 0; float queue_max_depth = -1000.f; int queue_max_depth_at = -1;
		for (int64_t f = 0; f < n_faces; f++)
			UpdateQueueIfPixelInsideFace(indexer, (t_face_index) f, queue, queue_size, queue_max_depth, queue_max_depth_at, blur_radius_ndc, pixel,
			                             faces_per_pixel, perspective_correct != 0, clip_bary != 0, cull_back != 0);
		std::sort(std::begin(queue), std::begin(queue) + queue_size);
		for (int k = 0; k < faces_per_pixel; k++) {
			int64_t o = i * faces_per_pixel + k;
			bool have = k < queue_size;
			out_faces[o] = have ? queue[k].face_index : -1;
			out_depths[o] = have ? queue[k].depth : -1.f;
			out_dist[o] = have ? queue[k].distance : -1.f;
			for (int c = 0; c < 3; c++) out_bary[o * 3 + c] = have ? queue[k].barycentric_coordinates(c) : -1.f;
		}
	}
}

// d(barycentric)/d(camera-space face vertices), 3x9 row-major
REF_API void ref_barycentric_jacobian(const float* ray_xy, const float* v0, const float* v1, const float* v2, const double* ndc_K,
                                      int perspective_correct, const float* stored_bary, float* out27) {
	using namespace nnrt::alignment::functional::kernel;
	static const double I4[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
	o3tgk_::TransformIndexer proj(ndc_K, I4, 1.0f);
	Eigen::Vector2f ray(ray_xy[0], ray_xy[1]);
	Eigen::Map<const Eigen::Vector3f> a(v0), b(v1), c(v2);
	nnrt::core::kernel::Matrix3x9f J;
	if (perspective_correct) {
		J = Jacobian_BarycentricCoordinatesWrtCameraSpaceVertices_WithPerspectiveCorrection(ray, a, b, c, proj);
	} else {
		Eigen::Map<const Eigen::RowVector3f> bary(stored_bary);
		J = Jacobian_BarycentricCoordinatesWrtCameraSpaceVertices_WithoutPerspectiveCorrection(ray, a, b, c, proj, bary);
	}
	for (int r = 0; r < 3; r++) for (int k = 0; k < 9; k++) out27[r * 9 + k] = J(r, k);
}
