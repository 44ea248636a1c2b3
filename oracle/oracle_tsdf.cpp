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

#include <cfloat>

// cpp/geometry/kernel/NonRigidSurfaceVoxelBlockGridImpl.h:289-363 (GetBoundingBoxesOfWarpedBlocks), quirks included: corner = block key
// (in blocks) + block side (in metres); fixed coverage always; ComputeAnchorsForPoint's result ignored (a corner with too few valid
// anchors is blended with raw weights); `if (min > x) min = x; else if (max < x) max = x`.
ORACLE_API void oracle_bounding_boxes_of_warped_blocks(const int32_t* block_keys, int64_t block_count, const float* nodes, const float* node_rotations,
                                                       const float* node_translations, int node_count, float node_coverage, int anchor_count,
                                                       int minimum_valid_anchor_count, float voxel_size, int block_resolution, const double* extrinsics,
                                                       float* boxes) {
	TransformIndexer transform(IDENTITY3, extrinsics, 1.0f);
	const float side = (float) block_resolution * voxel_size;
	const float c2 = node_coverage * node_coverage;
	static const int corner_bits[8][3] = {{0, 0, 0}, {0, 0, 1}, {0, 1, 0}, {1, 0, 0}, {0, 1, 1}, {1, 0, 1}, {1, 1, 0}, {1, 1, 1}};
	for (int64_t b = 0; b < block_count; b++) {
		const float x0 = (float) block_keys[b * 3], y0 = (float) block_keys[b * 3 + 1], z0 = (float) block_keys[b * 3 + 2];
		float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
		for (int c = 0; c < 8; c++) {
			V3 corner;
			transform.RigidTransform(corner_bits[c][0] ? x0 + side : x0, corner_bits[c][1] ? y0 + side : y0, corner_bits[c][2] ? z0 + side : z0,
			                         &corner.x, &corner.y, &corner.z);
			int32_t idx[MAX_ANCHOR_COUNT];
			float w[MAX_ANCHOR_COUNT];
			for (int a = 0; a < anchor_count; a++) { idx[a] = -1; w[a] = INFINITY; }
			int max_at = 0;
			float max_distance = INFINITY;
			for (int n = 0; n < node_count; n++) {
				V3 d = sub(load3(nodes + n * 3), corner);
				float distance = sqrtf(dot3(d, d));
				if (max_distance > distance) {
					w[max_at] = distance; idx[max_at] = n;
					max_at = 0; max_distance = w[0];
					for (int a = 1; a < anchor_count; a++) if (w[a] > max_distance) { max_at = a; max_distance = w[a]; }
				}
			}
			float weight_sum = 0.0f;
			int valid = 0;
			for (int a = 0; a < anchor_count; a++) {
				float square_distance = w[a] * w[a];
				if (square_distance > 4 * c2) { idx[a] = -1; continue; }
				float anchor_weight = expf(-square_distance / (2 * c2));
				weight_sum += anchor_weight; w[a] = anchor_weight; valid++;
			}
			if (valid >= minimum_valid_anchor_count) {
				if (weight_sum > 0.0f) { for (int a = 0; a < anchor_count; a++) w[a] /= weight_sum; }
				else if (valid > 0) { for (int a = 0; a < anchor_count; a++) w[a] = 1.0f / (float) valid; }
			}
			V3 warped = {0.f, 0.f, 0.f};
			for (int a = 0; a < anchor_count; a++) {
				int n = idx[a];
				if (n == -1) continue;
				V3 g = load3(nodes + n * 3), t = load3(node_translations + n * 3);
				warped = add(warped, scale(w[a], add(add(g, matvec(node_rotations + n * 9, sub(corner, g))), t)));
			}
			const float v[3] = {warped.x, warped.y, warped.z};
			for (int k = 0; k < 3; k++) {
				if (lo[k] > v[k]) lo[k] = v[k];
				else if (hi[k] < v[k]) hi[k] = v[k];
			}
		}
		for (int k = 0; k < 3; k++) { boxes[b * 6 + k] = lo[k]; boxes[b * 6 + 3 + k] = hi[k]; }
	}
}

// cpp/geometry/kernel/NonRigidSurfaceVoxelBlockGridImpl.h:365-435 + cpp/geometry/kernel/Segment.h:27-73: every (segment, box) pair
ORACLE_API void oracle_boxes_intersecting_surface_mask(const float* boxes, int64_t box_count, const double* intrinsics, const void* depth, int depth_dtype,
                                                       int height, int width, float depth_scale, float depth_max, int stride, float truncation,
                                                       uint8_t* mask) {
	TransformIndexer transform(intrinsics, IDENTITY4, 1.0f);
	struct Segment { V3 origin, inv; int sign[3]; };
	std::vector<Segment> segments;
	const int rows = height / stride, cols = width / stride;
	for (int64_t i = 0; i < (int64_t) rows * cols; i++) {
		int v = (int) (i / cols) * stride, u = (int) (i % cols) * stride;
		float d = (depth_dtype == 0 ? ((const float*) depth)[(int64_t) v * width + u] : (float) ((const uint16_t*) depth)[(int64_t) v * width + u]) / depth_scale;
		if (d > 0 && d < depth_max) {
			V3 start, end;
			transform.Unproject((float) u, (float) v, d - truncation, &start.x, &start.y, &start.z);
			transform.Unproject((float) u, (float) v, d + truncation, &end.x, &end.y, &end.z);
			V3 vec = sub(end, start);
			Segment s{start, {1.f / vec.x, 1.f / vec.y, 1.f / vec.z}, {0, 0, 0}};
			s.sign[0] = s.inv.x < 0; s.sign[1] = s.inv.y < 0; s.sign[2] = s.inv.z < 0;
			segments.push_back(s);
		}
	}
	for (int64_t b = 0; b < box_count; b++) {
		const float* box = boxes + b * 6;
		mask[b] = 0;
		for (const Segment& s : segments) {
			float t_min = (box[s.sign[0] ? 3 : 0] - s.origin.x) * s.inv.x, t_max = (box[s.sign[0] ? 0 : 3] - s.origin.x) * s.inv.x;
			float ty_min = (box[s.sign[1] ? 4 : 1] - s.origin.y) * s.inv.y, ty_max = (box[s.sign[1] ? 1 : 4] - s.origin.y) * s.inv.y;
			if ((t_min > ty_max) || (ty_min > t_max)) continue;
			if (ty_min > t_min) t_min = ty_min;
			if (ty_max < t_max) t_max = ty_max;
			float tz_min = (box[s.sign[2] ? 5 : 2] - s.origin.z) * s.inv.z, tz_max = (box[s.sign[2] ? 2 : 5] - s.origin.z) * s.inv.z;
			if ((t_min > tz_max) || (tz_min > t_max)) continue;
			if (tz_min > t_min) t_min = tz_min;
			if (tz_max < t_max) t_max = tz_max;
			if (!(t_max < 0.0f || t_min > 1.0f)) { mask[b] = 1; break; }
		}
	}
}

// Open3D's VoxelBlockGrid::Integrate (third party v0.17, absent from the reference checkout; re-exported by the reference as
// nnrt.geometry.VoxelBlockGrid.integrate), restated from the published kernel: see csrc/tsdf.cu::integrate_rigid_kernel. Parity unpinned.
ORACLE_API void oracle_integrate_rigid(
		const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf, void* weight, int weight_dtype,
		float* voxel_color, int block_resolution, float voxel_size, float sdf_truncation_distance, const void* depth, int depth_dtype,
		const void* color, int color_dtype, int height, int width, int color_height, int color_width, const double* depth_intrinsics,
		const double* color_intrinsics, const double* extrinsics, float depth_scale, float depth_max) {
	const int64_t r3 = (int64_t) block_resolution * block_resolution * block_resolution;
	TransformIndexer depth_transform(depth_intrinsics, extrinsics, 1.0f);
	TransformIndexer color_transform(color_intrinsics ? color_intrinsics : depth_intrinsics, IDENTITY4, 1.0f);
	auto in_boundary = [](float x, float y, int rows, int cols) { return y >= 0 && x >= 0 && y <= rows - 1.0f && x <= cols - 1.0f; };
	for (int64_t workload = 0; workload < block_count * r3; workload++) {
		const int64_t block_index = block_indices[workload / r3];
		const int64_t voxel_in_block = workload % r3;
		const int32_t* key = block_keys + block_index * 3;
		const int xl = (int) (voxel_in_block % block_resolution), yl = (int) ((voxel_in_block / block_resolution) % block_resolution),
		          zl = (int) (voxel_in_block / ((int64_t) block_resolution * block_resolution));
		const float gx = (float) (key[0] * block_resolution + xl) * voxel_size, gy = (float) (key[1] * block_resolution + yl) * voxel_size,
		            gz = (float) (key[2] * block_resolution + zl) * voxel_size;
		float xc, yc, zc, u, v;
		depth_transform.RigidTransform(gx, gy, gz, &xc, &yc, &zc);
		depth_transform.Project(xc, yc, zc, &u, &v);
		if (!in_boundary(u, v, height, width)) continue;
		const int64_t pixel = (int64_t) roundf(v) * width + (int64_t) roundf(u);
		const float d = (depth_dtype == 0 ? ((const float*) depth)[pixel] : (float) ((const uint16_t*) depth)[pixel]) / depth_scale;
		float sdf = d - zc;
		if (d <= 0.0f || d > depth_max || zc <= 0.0f || sdf < -sdf_truncation_distance) continue;
		sdf = (sdf < sdf_truncation_distance ? sdf : sdf_truncation_distance) / sdf_truncation_distance;
		const int64_t linear = block_index * r3 + voxel_in_block;
		const float w = weight_dtype == 0 ? ((const float*) weight)[linear] : (float) ((const uint16_t*) weight)[linear];
		const float inverted_weight_sum = 1.0f / (w + 1);
		tsdf[linear] = (w * tsdf[linear] + sdf) * inverted_weight_sum;
		if (color && voxel_color) {
			float cu, cv;
			color_transform.Project(xc, yc, zc, &cu, &cv);
			if (in_boundary(cu, cv, color_height, color_width)) {
				const int64_t cpx = (int64_t) roundf(cv) * color_width + (int64_t) roundf(cu);
				const float multiplier = color_dtype == 0 ? 255.0f : 1.0f;
				for (int c = 0; c < 3; c++) {
					const float in = color_dtype == 0 ? ((const float*) color)[cpx * 3 + c] : (float) ((const uint8_t*) color)[cpx * 3 + c];
					voxel_color[linear * 3 + c] = (w * voxel_color[linear * 3 + c] + in * multiplier) * inverted_weight_sum;
				}
			}
		}
		if (weight_dtype == 0) ((float*) weight)[linear] = w + 1;
		else ((uint16_t*) weight)[linear] = (uint16_t) (w + 1);
	}
}
