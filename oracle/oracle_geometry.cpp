// ORACLE (test infrastructure only) — warp-field geometry: anchors, LBS warp, unprojection,
// point-to-plane, warped-surface Jacobians. See oracle_common.h for status.
#include "oracle_common.h"

using namespace oracle;

// cpp/core/kernel/KnnUtilities.h:37-118 (brute-force KNN, replace-current-max, strict <,
// first-max scan) + cpp/geometry/functional/kernel/WarpUtilities.h:35-247 (weights) +
// cpp/geometry/functional/kernel/WarpAnchorComputationImpl.h:42-139 (plumbing).
// node_weights == nullptr -> FixedNodeWeight with node_coverage; else VariableNodeWeight
// (node_weights[n] is used directly as "node_coverage_squared", WarpUtilities.h:66).
// minimum_valid_anchor_count > 0 selects the threshold variants
// (cpp/geometry/functional/WarpAnchorComputation.cpp:74-148).
ORACLE_API void oracle_compute_anchors_euclidean(
		const float* points, int64_t point_count, const float* nodes, int node_count,
		const float* node_weights, float node_coverage, int anchor_count, int minimum_valid_anchor_count,
		int32_t* anchors, float* weights) {
	const float fixed_c2 = node_coverage * node_coverage;
	for (int64_t i = 0; i < point_count; i++) {
		int32_t* idx = anchors + i * anchor_count;
		float* w = weights + i * anchor_count;  // doubles as squared-distance scratch (WarpUtilities.h:190)
		V3 p = load3(points + i * 3);
		for (int a = 0; a < anchor_count; a++) { idx[a] = -1; w[a] = INFINITY; }
		int max_at = 0;
		float max_d2 = INFINITY;
		for (int n = 0; n < node_count; n++) {
			V3 d = sub(load3(nodes + n * 3), p);
			float d2 = dot3(d, d);
			if (d2 < max_d2) {
				w[max_at] = d2;
				idx[max_at] = n;
				max_at = 0;
				max_d2 = w[0];
				for (int a = 1; a < anchor_count; a++) {
					if (w[a] > max_d2) { max_at = a; max_d2 = w[a]; }
				}
			}
		}
		float weight_sum = 0.0f;
		int valid = anchor_count;
		if (minimum_valid_anchor_count > 0) {
			valid = 0;
			for (int a = 0; a < anchor_count; a++) {
				float d2 = w[a];
				float c2 = node_weights ? (idx[a] >= 0 ? node_weights[idx[a]] : fixed_c2) : fixed_c2;
				if (d2 > 4 * c2) { idx[a] = -1; continue; }  // weight slot keeps d2 (reference leaves it)
				float weight = expf(-d2 / (2 * c2));
				weight_sum += weight;
				w[a] = weight;
				valid++;
			}
			if (valid < minimum_valid_anchor_count) continue;  // point rejected, no normalisation
		} else {
			for (int a = 0; a < anchor_count; a++) {
				float d2 = w[a];
				float c2 = node_weights ? (idx[a] >= 0 ? node_weights[idx[a]] : fixed_c2) : fixed_c2;
				float weight = expf(-d2 / (2 * c2));
				weight_sum += weight;
				w[a] = weight;
			}
		}
		// NormalizeAnchorWeights, WarpUtilities.h:35-46
		if (weight_sum > 0.0f) {
			for (int a = 0; a < anchor_count; a++) w[a] /= weight_sum;
		} else if (valid > 0) {
			for (int a = 0; a < anchor_count; a++) w[a] = 1.0f / (float) valid;
		}
	}
}

// cpp/geometry/functional/kernel/Warp3dPointsAndNormalsImpl.h:334-438 + BlendWarp
// WarpUtilities.h:447-467 (+ :527-541 threshold variant). normals may be nullptr.
// minimum_valid_anchor_count <= 0 -> no thresholding.
ORACLE_API void oracle_warp_points_normals(
		const float* points, const float* normals, int64_t point_count,
		const float* nodes, const float* node_rotations, const float* node_translations,
		const int32_t* anchors, const float* anchor_weights, int anchor_count, int minimum_valid_anchor_count,
		const double* extrinsics44, float* warped_points, float* warped_normals) {
	TransformIndexer transform(IDENTITY3, extrinsics44, 1.0f);
	for (int64_t i = 0; i < point_count; i++) {
		V3 pc, nc = {0, 0, 0};
		transform.RigidTransform(points[i * 3], points[i * 3 + 1], points[i * 3 + 2], &pc.x, &pc.y, &pc.z);
		if (normals) transform.Rotate(normals[i * 3], normals[i * 3 + 1], normals[i * 3 + 2], &nc.x, &nc.y, &nc.z);
		V3 wp = {0, 0, 0}, wn = {0, 0, 0};
		const int32_t* idx = anchors + i * anchor_count;
		const float* w = anchor_weights + i * anchor_count;
		bool do_warp = true;
		if (minimum_valid_anchor_count > 0) {
			int valid = 0;
			for (int a = 0; a < anchor_count; a++) valid += (idx[a] != -1);
			do_warp = valid >= minimum_valid_anchor_count;
		}
		if (do_warp) {
			for (int a = 0; a < anchor_count; a++) {
				int n = idx[a];
				if (n == -1) continue;
				float weight = w[a];
				const float* R = node_rotations + n * 9;
				V3 g = load3(nodes + n * 3);
				V3 t = load3(node_translations + n * 3);
				// warped_point += w * (node + R*(p - node) + t)
				V3 rotated = matvec(R, sub(pc, g));
				wp = add(wp, scale(weight, add(add(g, rotated), t)));
				if (normals) {
					// `anchor_weight * (node_rotation * source_normal)` (WarpUtilities.h:466): Eigen folds the scalar into the lazy
					// 3x3 product, i.e. each coefficient is sum_k (w * R_ik) * n_k — verified bit-exact against the compiled header
					// (tests/test_oracle_vs_ref.py); w * (R n) differs in the last ulp
					const float wR[9] = {weight * R[0], weight * R[1], weight * R[2], weight * R[3], weight * R[4], weight * R[5],
					                     weight * R[6], weight * R[7], weight * R[8]};
					wn = add(wn, matvec(wR, nc));
				}
			}
		}
		store3(warped_points + i * 3, wp);
		if (normals && warped_normals) store3(warped_normals + i * 3, wn);
	}
}

// cpp/geometry/functional/kernel/PerspectiveProjectionImpl.h:62-150.
// depth is float32 [H,W]; pose = inverse(extrinsics) is computed by the caller in double
// (Open3D InverseTransformation) and passed here as pose44.
ORACLE_API void oracle_unproject_depth(
		const float* depth, int rows, int cols, const double* intrinsics33, const double* pose44,
		float depth_scale, float depth_max, float* points, uint8_t* mask) {
	TransformIndexer transform(intrinsics33, pose44, 1.0f);
	for (int64_t i = 0; i < (int64_t) rows * cols; i++) {
		int64_t y = i / cols, x = i % cols;
		float d = depth[i] / depth_scale;
		float* out = points + i * 3;
		out[0] = out[1] = out[2] = 0.f;
		if (mask) mask[i] = 0;
		if (d > 0 && d < depth_max) {
			float xc, yc, zc;
			transform.Unproject((float) x, (float) y, d, &xc, &yc, &zc);
			transform.RigidTransform(xc, yc, zc, out, out + 1, out + 2);
			if (mask) mask[i] = 1;
		}
	}
}

// cpp/geometry/functional/kernel/PointToPlaneDistancesImpl.h:25-52: n1 . (v1 - v2)
ORACLE_API void oracle_point_to_plane_distances(
		const float* normals1, const float* vertices1, const float* vertices2, int64_t count, float* distances) {
	for (int64_t i = 0; i < count; i++) {
		distances[i] = dot3(load3(normals1 + i * 3), sub(load3(vertices1 + i * 3), load3(vertices2 + i * 3)));
	}
}

// cpp/alignment/functional/kernel/WarpedSurfaceJacobiansImpl.h:33-158 (TVertexRotationOnly=false).
// out_position_jacobians [V,A,4] = (-w R (v - g), w); out_normal_jacobians [V,A,3] = -w R n.
ORACLE_API void oracle_warped_surface_jacobians(
		const float* vertex_positions, const float* vertex_normals, int64_t vertex_count,
		const float* nodes, const float* node_rotations, const int32_t* anchors, const float* anchor_weights,
		int anchor_count, float* out_position_jacobians, float* out_normal_jacobians) {
	memset(out_position_jacobians, 0, sizeof(float) * vertex_count * anchor_count * 4);
	memset(out_normal_jacobians, 0, sizeof(float) * vertex_count * anchor_count * 3);
	for (int64_t wi = 0; wi < vertex_count * anchor_count; wi++) {
		int64_t v = wi / anchor_count;
		int n = anchors[wi];
		if (n == -1) continue;
		float w = anchor_weights[wi];
		const float* R = node_rotations + n * 9;
		V3 rv = matvec(R, sub(load3(vertex_positions + v * 3), load3(nodes + n * 3)));
		store3(out_position_jacobians + wi * 4, scale(-w, rv));
		out_position_jacobians[wi * 4 + 3] = w;
		store3(out_normal_jacobians + wi * 3, scale(-w, matvec(R, load3(vertex_normals + v * 3))));
	}
}

// ---- normals: cpp/geometry/functional/NormalsOperations.cpp:36-93, kernel/NormalsOperationsImpl.h:38-214 --------------------------
namespace {
inline V3 cross3(const V3& a, const V3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// Eigen MatrixBase::normalize()/normalized() (Eigen/src/Core/Dot.h:119-148): divide by sqrt(squaredNorm) when squaredNorm > 0
inline V3 normalized3(const V3& v) {
	float z = dot3(v, v);
	if (z > 0.f) { float n = sqrtf(z); return {v.x / n, v.y / n, v.z / n}; }
	return v;
}
inline V3 normalize_or_up(V3 v) {  // NormalizeVectors3d, NormalsOperationsImpl.h:72-93
	v = normalized3(v);
	if (std::isnan(v.x)) v = {0.f, 0.f, 1.f};
	return v;
}
}  // namespace

ORACLE_API void oracle_compute_triangle_normals(const float* positions, const int64_t* triangles, int64_t face_count, int normalized, float* normals) {
	for (int64_t f = 0; f < face_count; f++) {
		V3 v0 = load3(positions + triangles[f * 3] * 3), v1 = load3(positions + triangles[f * 3 + 1] * 3), v2 = load3(positions + triangles[f * 3 + 2] * 3);
		V3 n = cross3(sub(v1, v0), sub(v2, v0));
		store3(normals + f * 3, normalized ? normalize_or_up(n) : n);
	}
}

// triangles visited in index order (ParallelFor on one thread): every vertex receives its addends in ascending face order
ORACLE_API void oracle_compute_vertex_normals(const int64_t* triangles, int64_t face_count, const float* triangle_normals, int64_t vertex_count,
                                              int normalized, float* vertex_normals) {
	for (int64_t i = 0; i < vertex_count * 3; i++) vertex_normals[i] = 0.f;
	for (int64_t f = 0; f < face_count; f++)
		for (int k = 0; k < 3; k++) {
			float* out = vertex_normals + triangles[f * 3 + k] * 3;
			out[0] += triangle_normals[f * 3]; out[1] += triangle_normals[f * 3 + 1]; out[2] += triangle_normals[f * 3 + 2];
		}
	if (normalized)
		for (int64_t v = 0; v < vertex_count; v++) store3(vertex_normals + v * 3, normalize_or_up(load3(vertex_normals + v * 3)));
}

ORACLE_API void oracle_compute_ordered_point_cloud_normals(const float* points, int height, int width, float* normals) {
	for (int64_t i = 0; i < (int64_t) height * width; i++) {
		int64_t y = i / width, x = i % width;
		V3 n = {0.f, 0.f, 0.f};
		if (!(x == 0 || x == width - 1 || y == 0 || y == height - 1)) {
			V3 d_horizontal = sub(load3(points + (i + 1) * 3), load3(points + (i - 1) * 3));
			V3 d_vertical = sub(load3(points + (i - width) * 3), load3(points + (i + width) * 3));
			n = normalized3(cross3(d_horizontal, d_vertical));
			if (n.z > 0) n = {-n.x, -n.y, -n.z};
		}
		store3(normals + i * 3, n);
	}
}

// ---- shortest-path anchors: cpp/core/kernel/KnnUtilities_PriorityQueue.h:102-160, cpp/core/kernel/KnnUtilities.h:119-142,
// cpp/core/heap/CPU/DeviceHeapCPU.h (std::push_heap / std::pop_heap on a 40-entry array, inserts beyond capacity dropped),
// cpp/geometry/functional/kernel/WarpUtilities.h:36-92,344-375, kernel/WarpAnchorComputationImpl.h:146-240 ----------------------
namespace {
struct DistanceIndexPair { float key; int32_t index; };
constexpr int SHORTEST_PATH_QUEUE_CAPACITY = 40;
}  // namespace

// edges [node_count, graph_degree] int32 adjacency (-1 = none). node_weights == nullptr -> fixed coverage.
// Requires node_count >= anchor_count (the reference reads uninitialised distances / node_weights[-1] otherwise).
ORACLE_API void oracle_compute_anchors_shortest_path(const float* points, int64_t point_count, const float* nodes, int node_count,
                                                     const float* node_weights, float node_coverage, const int32_t* edges, int graph_degree,
                                                     int anchor_count, int32_t* anchors, float* weights) {
	auto compare = [](const DistanceIndexPair& first, const DistanceIndexPair& second) { return first.key > second.key; };  // MinHeapKeyCompare
	const float fixed_c2 = node_coverage * node_coverage;
	for (int64_t i = 0; i < point_count; i++) {
		int32_t* idx = anchors + i * anchor_count;
		float* dist = weights + i * anchor_count;
		for (int a = 0; a < anchor_count; a++) idx[a] = -1;
		V3 p = load3(points + i * 3);
		DistanceIndexPair queue[SHORTEST_PATH_QUEUE_CAPACITY];
		int size = 0;
		auto insert = [&](DistanceIndexPair e) {
			if (size >= SHORTEST_PATH_QUEUE_CAPACITY) return;
			queue[size++] = e;
			std::push_heap(queue, queue + size, compare);
		};
		int discovered = 0;
		while (discovered < anchor_count) {
			// nearest node not yet among the anchors (k = 1 brute force, strict <, ascending index)
			int closest = -1;
			float closest_d2 = INFINITY;
			for (int n = 0; n < node_count; n++) {
				bool excluded = false;
				for (int a = 0; a < anchor_count; a++) if (idx[a] == n) excluded = true;
				if (excluded) continue;
				V3 d = sub(load3(nodes + n * 3), p);
				float d2 = dot3(d, d);
				if (d2 < closest_d2) { closest_d2 = d2; closest = n; }
			}
			if (closest == -1) break;
			insert({sqrtf(closest_d2), closest});
			while (size > 0 && discovered < anchor_count) {
				std::pop_heap(queue, queue + size, compare);
				DistanceIndexPair source = queue[--size];
				bool processed = false;
				for (int a = 0; a < discovered && !processed; a++) if (idx[a] == source.index) processed = true;
				if (processed) continue;
				dist[discovered] = source.key;
				idx[discovered] = source.index;
				discovered++;
				if (discovered >= anchor_count) break;
				for (int e = 0; e < graph_degree; e++) {
					int32_t target = edges[(int64_t) source.index * graph_degree + e];
					if (target > -1) {
						V3 d = sub(load3(nodes + target * 3), load3(nodes + source.index * 3));
						insert({source.key + sqrtf(dot3(d, d)), target});
					}
				}
			}
		}
		float weight_sum = 0.0f;
		for (int a = 0; a < anchor_count; a++) {
			float square_distance = dist[a] * dist[a];
			float c2 = node_weights ? node_weights[idx[a]] : fixed_c2;
			float weight = expf(-square_distance / (2 * c2));
			weight_sum += weight;
			dist[a] = weight;
		}
		if (weight_sum > 0.0f) for (int a = 0; a < anchor_count; a++) dist[a] /= weight_sum;
		else for (int a = 0; a < anchor_count; a++) dist[a] = 1.0f / (float) anchor_count;
	}
}
