// ORACLE (test infrastructure only) — non-rigid TSDF integration:
// cpp/geometry/kernel/NonRigidSurfaceVoxelBlockGridImpl.h:50-229 (IntegrateNonRigid_Generic) over caller-owned voxel-block arrays
// (the block hash map itself is Open3D's VoxelBlockGrid; only its key / value tensors appear here). See oracle_common.h for status.
// Parity unpinned by the reference (its tests for this path need downloaded data).
//
// Anchors: WarpField::ComputeAnchorsForPoint<threshold = true> (cpp/geometry/WarpField.h:77-105) searches a kd-tree whose traversal
// order decides the SLOT order of the k nearest nodes, hence the fp32 summation order of the blend. The kd-tree is not restated:
// nodes are visited in ascending index order (same set, barring exact distance ties). Distances are norms (sqrt), squared again for
// the weights (WarpUtilities.h:131-153 with TDistancesAreSquared = false); a dropped anchor keeps its DISTANCE in the weight slot.
#include "oracle_common.h"

using namespace oracle;

// voxel workloads in index order; cos_voxel_ray_to_normal [H,W] is zero-initialised and keeps the LAST voxel's value per pixel.
// depth_dtype: 0 = float32, 1 = uint16. color may be null (no colour integration); color_dtype: 0 = float32 in [0,1], 1 = uint8.
// weight_dtype: 0 = float32, 1 = uint16. The voxel weights are read, never written (as in the reference).
ORACLE_API void oracle_integrate_non_rigid(
		const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf, const void* weight, int weight_dtype,
		float* voxel_color, int block_resolution, float voxel_size, float sdf_truncation_distance,
		const void* depth, int depth_dtype, const void* color, int color_dtype, const float* depth_normals, int height, int width,
		int color_height, int color_width, const double* depth_intrinsics, const double* color_intrinsics, const double* extrinsics,
		const float* nodes, const float* node_rotations, const float* node_translations, int node_count, const float* node_coverage_weights,
		float node_coverage, int anchor_count, int minimum_valid_anchor_count, float depth_scale, float depth_max,
		float* cos_voxel_ray_to_normal) {
	const int64_t r3 = (int64_t) block_resolution * block_resolution * block_resolution;
	TransformIndexer depth_transform(depth_intrinsics, extrinsics, 1.0f);
	TransformIndexer color_transform(color_intrinsics, IDENTITY4, 1.0f);
	const float fixed_c2 = node_coverage * node_coverage;
	for (int64_t i = 0; i < (int64_t) height * width; i++) cos_voxel_ray_to_normal[i] = 0.f;
	auto in_boundary = [](float x, float y, int rows, int cols) {  // Open3D ArrayIndexer::InBoundary
		return y >= 0 && x >= 0 && y <= rows - 1.0f && x <= cols - 1.0f;
	};
	for (int64_t workload = 0; workload < block_count * r3; workload++) {
		const int64_t block_index = block_indices[workload / r3];
		const int64_t voxel_in_block = workload % r3;
		const int32_t* key = block_keys + block_index * 3;
		const int xl = (int) (voxel_in_block % block_resolution), yl = (int) ((voxel_in_block / block_resolution) % block_resolution),
		          zl = (int) (voxel_in_block / ((int64_t) block_resolution * block_resolution));
		const float gx = (float) (key[0] * block_resolution + xl) * voxel_size, gy = (float) (key[1] * block_resolution + yl) * voxel_size,
		            gz = (float) (key[2] * block_resolution + zl) * voxel_size;
		V3 voxel;
		depth_transform.RigidTransform(gx, gy, gz, &voxel.x, &voxel.y, &voxel.z);
		// --- anchors (threshold variant)
		int32_t idx[MAX_ANCHOR_COUNT];
		float w[MAX_ANCHOR_COUNT];
		for (int a = 0; a < anchor_count; a++) { idx[a] = -1; w[a] = INFINITY; }
		int max_at = 0;
		float max_distance = INFINITY;
		for (int n = 0; n < node_count; n++) {
			V3 d = sub(load3(nodes + n * 3), voxel);
			float distance = sqrtf(dot3(d, d));
			if (max_distance > distance) {
				w[max_at] = distance;
				idx[max_at] = n;
				max_at = 0;
				max_distance = w[0];
				for (int a = 1; a < anchor_count; a++)
					if (w[a] > max_distance) { max_at = a; max_distance = w[a]; }
			}
		}
		float weight_sum = 0.0f;
		int valid = 0;
		for (int a = 0; a < anchor_count; a++) {
			float square_distance = w[a] * w[a];
			float c2 = node_coverage_weights ? node_coverage_weights[idx[a] < 0 ? 0 : idx[a]] : fixed_c2;
			if (square_distance > 4 * c2) { idx[a] = -1; continue; }
			float anchor_weight = expf(-square_distance / (2 * c2));
			weight_sum += anchor_weight;
			w[a] = anchor_weight;
			valid++;
		}
		if (valid < minimum_valid_anchor_count) continue;
		if (weight_sum > 0.0f) { for (int a = 0; a < anchor_count; a++) w[a] /= weight_sum; }
		else if (valid > 0) { for (int a = 0; a < anchor_count; a++) w[a] = 1.0f / (float) valid; }
		// --- blend warp (WarpUtilities.h:428-445)
		V3 warped = {0.f, 0.f, 0.f};
		for (int a = 0; a < anchor_count; a++) {
			int n = idx[a];
			if (n == -1) continue;
			V3 g = load3(nodes + n * 3), t = load3(node_translations + n * 3);
			V3 rotated = matvec(node_rotations + n * 9, sub(voxel, g));
			warped = add(warped, scale(w[a], add(add(g, rotated), t)));
		}
		if (warped.z < 0) continue;
		float u_precise, v_precise;
		depth_transform.Project(warped.x, warped.y, warped.z, &u_precise, &v_precise);
		if (!in_boundary(u_precise, v_precise, height, width)) continue;
		int64_t u = (int64_t) roundf(u_precise), v = (int64_t) roundf(v_precise);
		float depth_value = depth_dtype == 0 ? ((const float*) depth)[v * width + u] : (float) ((const uint16_t*) depth)[v * width + u];
		depth_value = depth_value / depth_scale;
		if (depth_value <= 0.0f || depth_value > depth_max) continue;
		float psdf = depth_value - warped.z;
		V3 view = {-warped.x, -warped.y, -warped.z};
		float z = dot3(view, view);
		if (z > 0.f) { float norm = sqrtf(z); view = {view.x / norm, view.y / norm, view.z / norm}; }
		float cosine = dot3(view, load3(depth_normals + (v * width + u) * 3));
		cos_voxel_ray_to_normal[v * width + u] = cosine;
		if (psdf <= -sdf_truncation_distance || cosine > 0.5f) continue;
		const int64_t linear = block_index * r3 + voxel_in_block;
		float tsdf_normalized = (psdf < sdf_truncation_distance ? psdf : sdf_truncation_distance) / sdf_truncation_distance;
		float voxel_weight = weight_dtype == 0 ? ((const float*) weight)[linear] : (float) ((const uint16_t*) weight)[linear];
		float inverted_weight_sum = 1.0f / (voxel_weight + 1);
		tsdf[linear] = (voxel_weight * tsdf[linear] + tsdf_normalized) * inverted_weight_sum;
		if (color && voxel_color) {
			float x, y, zz;
			depth_transform.Unproject((float) u, (float) v, 1.0f, &x, &y, &zz);
			color_transform.Project(x, y, zz, &u_precise, &v_precise);
			if (in_boundary(u_precise, v_precise, color_height, color_width)) {
				int64_t uc = (int64_t) roundf(u_precise), vc = (int64_t) roundf(v_precise);
				const float multiplier = color_dtype == 0 ? 255.0f : 1.0f;
				for (int c = 0; c < 3; c++) {
					float in = color_dtype == 0 ? ((const float*) color)[(vc * color_width + uc) * 3 + c] : (float) ((const uint8_t*) color)[(vc * color_width + uc) * 3 + c];
					voxel_color[linear * 3 + c] = (voxel_weight * voxel_color[linear * 3 + c] + in * multiplier) * inverted_weight_sum;
				}
			}
		}
	}
}
