// ORACLE (test infrastructure only) — NDC face extraction, coarse-to-fine rasteriser,
// attribute interpolation. See oracle_common.h for status.
#include "oracle_common.h"

using namespace oracle;

// cpp/rendering/kernel/CoordinateSystemConversions.h:108-148. out_ndc_intrinsics: 9 doubles,
// out_range: {min_x, max_x, min_y, max_y} floats.
ORACLE_API void oracle_intrinsics_to_ndc(const double* K, int height_px, int width_px, double* out_ndc_intrinsics, float* out_range) {
	double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
	double height = height_px, width = width_px;
	double smaller = std::min(width, height);
	float range_x = GetNdcRange(width_px, height_px);
	float range_y = GetNdcRange(height_px, width_px);
	double fx_ndc = 2.0 * fx / smaller;
	double fy_ndc = -2.0 * fy / smaller;
	double cx_ndc = -(2.0 * cx - width) / smaller;
	double cy_ndc = (2.0 * cy - height) / smaller;
	const double out[9] = {fx_ndc, 0, cx_ndc, 0, fy_ndc, cy_ndc, 0, 0, 1};
	memcpy(out_ndc_intrinsics, out, sizeof(out));
	out_range[0] = (float) (cx_ndc - range_x / 2.f);
	out_range[1] = (float) (cx_ndc + range_x / 2.f);
	out_range[2] = (float) (cy_ndc - range_y / 2.f);
	out_range[3] = (float) (cy_ndc + range_y / 2.f);
}

// cpp/rendering/functional/kernel/ExtractClippedFaceVerticesImpl.h:108-179 (single mesh,
// no normals). Masked faces are written as zeros (reference leaves them uninitialised and its
// test zeroes them before comparing, test_extract_face_vertices.cpp:54-55).
ORACLE_API void oracle_face_vertices_ndc(
		const float* vertex_positions, const int64_t* triangles, int64_t face_count,
		const double* intrinsics33, int height, int width, float near_clip, float far_clip,
		float* face_vertices_ndc, uint8_t* clip_mask) {
	double ndc_K[9];
	float range[4];
	oracle_intrinsics_to_ndc(intrinsics33, height, width, ndc_K, range);
	TransformIndexer projection(ndc_K, IDENTITY4, 1.0f);
	for (int64_t f = 0; f < face_count; f++) {
		float* out = face_vertices_ndc + f * 9;
		for (int k = 0; k < 9; k++) out[k] = 0.f;
		clip_mask[f] = 0;
		V3 v[3];
		for (int k = 0; k < 3; k++) v[k] = load3(vertex_positions + triangles[f * 3 + k] * 3);
		bool in_range = false;
		for (int k = 0; k < 3; k++) {
			in_range |= v[k].z >= near_clip;  // sic: OR of both tests, never clips (…Impl.h:129-137)
			in_range |= v[k].z <= far_clip;
		}
		if (!in_range) continue;
		V2 xy[3];
		bool has_inlier = false;
		for (int k = 0; k < 3; k++) {
			projection.Project(v[k].x, v[k].y, v[k].z, &xy[k].x, &xy[k].y);
			// AxisAligned2dBoundingBox::Contains, cpp/geometry/kernel/AxisAlignedBoundingBox.h:31-33
			has_inlier |= (xy[k].y >= range[2] && xy[k].x >= range[0] && xy[k].y <= range[3] && xy[k].x <= range[1]);
		}
		if (!has_inlier) continue;
		for (int k = 0; k < 3; k++) {
			out[k * 3 + 0] = xy[k].x;
			out[k * 3 + 1] = xy[k].y;
			out[k * 3 + 2] = v[k].z;
		}
		clip_mask[f] = 1;
	}
}

namespace {
// cpp/rendering/functional/kernel/BarycentricCoordinates.h:36-48 (ClockWise)
inline float SignedParallelogramArea(const V2& p, const V2& v0, const V2& v1) {
	return (p.x - v0.x) * (v0.y - v1.y) - (p.y - v0.y) * (v0.x - v1.x);
}
inline float saturatef(float x) { return fminf(fmaxf(x, 0.f), 1.f); }
inline float fmin3f(float a, float b, float c) { return std::min(std::min(a, b), c); }
inline float fmax3f(float a, float b, float c) { return std::max(std::max(a, b), c); }

// cpp/rendering/kernel/RayFaceIntersection.h:121-139
inline float PointSegmentSquareDistance(const V2& p, const V2& s0, const V2& s1) {
	V2 seg = {s1.x - s0.x, s1.y - s0.y};
	float len2 = dot2(seg, seg);
	V2 ps0 = {p.x - s0.x, p.y - s0.y};
	float ratio = dot2(seg, ps0) / len2;
	if (len2 <= K_EPSILON) {
		V2 d = {p.x - s1.x, p.y - s1.y};
		return dot2(d, d);
	}
	ratio = saturatef(ratio);
	V2 closest = {s0.x + ratio * seg.x, s0.y + ratio * seg.y};
	V2 d = {closest.x - p.x, closest.y - p.y};
	return dot2(d, d);
}

struct Hit {
	float depth;
	int32_t face;
	float distance;
	V3 bary;
};
inline bool hit_less(const Hit& a, const Hit& b) { return a.depth < b.depth || (a.depth == b.depth && a.face < b.face); }

// cpp/rendering/kernel/RayFaceIntersection.h:162-255
inline void UpdateQueueIfPixelInsideFace(
		const float* fv, int32_t i_face, Hit* queue, int& queue_size, float& queue_max_depth, int& queue_max_depth_at,
		float blur_radius, const V2& pixel, int faces_per_pixel, bool perspective_correct, bool clip_bary, bool cull_back_faces) {
	V2 v0 = {fv[0], fv[1]}, v1 = {fv[3], fv[4]}, v2 = {fv[6], fv[7]};
	float z0 = fv[2], z1 = fv[5], z2 = fv[8];
	const float face_area = SignedParallelogramArea(v0, v1, v2);
	const bool is_back_face = face_area < 0.f;
	const bool zero_face_area = (face_area <= K_EPSILON && face_area >= -1.f * K_EPSILON);
	// PointOutsideFaceBoundingBox, :83-94 (+ ComputeFace2dBoundingBoxAndCheckZMin :56-78)
	float x_min = fmin3f(v0.x, v1.x, v2.x) - blur_radius, x_max = fmax3f(v0.x, v1.x, v2.x) + blur_radius;
	float y_min = fmin3f(v0.y, v1.y, v2.y) - blur_radius, y_max = fmax3f(v0.y, v1.y, v2.y) + blur_radius;
	bool z_invalid = fmax3f(z0, z1, z2) < K_EPSILON;
	bool outside_bbox = (pixel.x > x_max || pixel.x < x_min || pixel.y > y_max || pixel.y < y_min || z_invalid);
	if (outside_bbox || (cull_back_faces && is_back_face) || zero_face_area) return;

	// BarycentricCoordinates.h:51-64
	const float area_eps = SignedParallelogramArea(v0, v1, v2) + K_EPSILON;
	V3 bary = {SignedParallelogramArea(pixel, v1, v2) / area_eps,
	           SignedParallelogramArea(pixel, v2, v0) / area_eps,
	           SignedParallelogramArea(pixel, v0, v1) / area_eps};
	if (perspective_correct) {  // BarycentricCoordinates.h:88-103
		const float n0 = bary.x * z1 * z2;
		const float n1 = z0 * bary.y * z2;
		const float n2 = z0 * z1 * bary.z;
		const float denominator = fmaxf(n0 + n1 + n2, K_EPSILON);
		bary = {n0 / denominator, n1 / denominator, n2 / denominator};
	}
	V3 bary_clipped = bary;
	if (clip_bary) {  // BarycentricCoordinates.h:106-116 (Eigen normalize(): v / sqrt(squaredNorm))
		V3 c = {fmaxf(bary.x, 0.f), fmaxf(bary.y, 0.f), fmaxf(bary.z, 0.f)};
		float norm2 = dot3(c, c);
		if (norm2 > 0.f) { float nrm = sqrtf(norm2); c = {c.x / nrm, c.y / nrm, c.z / nrm}; }
		bary_clipped = c;
	}
	const float depth = bary_clipped.x * z0 + bary_clipped.y * z1 + bary_clipped.z * z2;
	if (depth < 0.f) return;
	const float dist = fmin3f(PointSegmentSquareDistance(pixel, v0, v1), PointSegmentSquareDistance(pixel, v0, v2),
	                          PointSegmentSquareDistance(pixel, v1, v2));
	const bool inside = bary.x > 0.f && bary.y > 0.f && bary.z > 0.f;
	const float signed_dist = inside ? -dist : dist;
	if (!inside && dist >= blur_radius) return;  // sic: squared distance vs radius
	if (queue_size < faces_per_pixel) {
		queue[queue_size] = {depth, i_face, signed_dist, bary_clipped};
		if (depth > queue_max_depth) { queue_max_depth = depth; queue_max_depth_at = queue_size; }
		queue_size++;
	} else if (depth < queue_max_depth) {
		queue[queue_max_depth_at] = {depth, i_face, signed_dist, bary_clipped};
		queue_max_depth = depth;
		for (int i = 0; i < faces_per_pixel; i++) {
			if (queue[i].depth > queue_max_depth) { queue_max_depth = queue[i].depth; queue_max_depth_at = i; }
		}
	}
}
} // namespace

// cpp/rendering/RasterizeNdcTriangles.cpp:32-128 + kernel/RasterizeNdcTrianglesImpl.h:187-391 +
// kernel/RasterizeNdcTrianglesImplCPU.h:26-93 (bins filled in ascending face order = single-threaded CPU).
// clip_mask may be nullptr. bin_size -1 = auto, 0 = brute force. max_faces_per_bin -1 = auto;
// pass 0x7fffffff-like large values to lift the cap (bins are std::vectors here; the cap only truncates).
// Returns 0, or the number of bin insertions dropped by the cap.
ORACLE_API int64_t oracle_rasterize_ndc_triangles(
		const float* face_vertices_ndc, const uint8_t* clip_mask, int64_t face_count, int height, int width,
		float blur_radius_pixels, int faces_per_pixel, int bin_size, int max_faces_per_bin,
		int perspective_correct, int clip_barycentric, int cull_back_faces, int parallel,
		int64_t* pixel_face_indices, float* pixel_depths, float* pixel_bary, float* pixel_face_distances) {
	const int64_t px_count = (int64_t) height * width;
	for (int64_t i = 0; i < px_count * faces_per_pixel; i++) {
		pixel_face_indices[i] = -1; pixel_depths[i] = -1.f; pixel_face_distances[i] = -1.f;
		pixel_bary[i * 3] = pixel_bary[i * 3 + 1] = pixel_bary[i * 3 + 2] = -1.f;
	}
	if (face_count == 0) return 0;
	int max_dim = std::max(height, width);
	if (bin_size == -1) {
		if (max_dim <= 64) bin_size = 8;
		else bin_size = (int) std::pow(2, std::max((int) std::ceil(std::log2((double) max_dim)) - 4, 4));
	}
	if (max_faces_per_bin == -1) {
		int64_t kept = face_count;
		if (clip_mask) { kept = 0; for (int64_t f = 0; f < face_count; f++) kept += clip_mask[f] != 0; }
		max_faces_per_bin = std::max(10000, (int) (kept / 5));
	}
	// ImageSpaceDistanceToNdc, CoordinateSystemConversions.h:98-102
	const float blur_radius = blur_radius_pixels / ((float) fminf((float) height, (float) width) / 2.0f);

	int bins_y = 1, bins_x = 1;
	std::vector<std::vector<int32_t>> bins;
	int64_t dropped = 0;
	const bool binned = bin_size > 0 && max_faces_per_bin > 0;
	if (binned) {
		bins_y = 1 + (height - 1) / bin_size;
		bins_x = 1 + (width - 1) / bin_size;
		bins.resize((size_t) bins_y * bins_x);
		// sic: x/y swapped in the half-pixel computation (RasterizeNdcTrianglesImpl.h:377-381)
		const float half_range_y = GetNdcRange(width, height) / 2.f;
		const float half_range_x = GetNdcRange(height, width) / 2.f;
		const float half_pixel_y = half_range_y / (float) height;
		const float half_pixel_x = half_range_x / (float) width;
		for (int64_t f = 0; f < face_count; f++) {
			if (clip_mask && !clip_mask[f]) continue;
			const float* fv = face_vertices_ndc + f * 9;
			float x_min = fmin3f(fv[0], fv[3], fv[6]) - blur_radius, x_max = fmax3f(fv[0], fv[3], fv[6]) + blur_radius;
			float y_min = fmin3f(fv[1], fv[4], fv[7]) - blur_radius, y_max = fmax3f(fv[1], fv[4], fv[7]) + blur_radius;
			if (fmax3f(fv[2], fv[5], fv[8]) < K_EPSILON) continue;
			for (int by = 0; by < bins_y; by++) {
				const float bin_y_min = PixelToNdc(by * bin_size, height, width) - half_pixel_y;
				const float bin_y_max = PixelToNdc((by + 1) * bin_size - 1, height, width) + half_pixel_y;
				if (!((y_min <= bin_y_max) && (bin_y_min < y_max))) continue;
				for (int bx = 0; bx < bins_x; bx++) {
					const float bin_x_min = PixelToNdc(bx * bin_size, width, height) - half_pixel_x;
					const float bin_x_max = PixelToNdc((bx + 1) * bin_size - 1, width, height) + half_pixel_x;
					if ((x_min <= bin_x_max) && (bin_x_min < x_max)) {
						auto& bin = bins[(size_t) by * bins_x + bx];
						if ((int64_t) bin.size() < max_faces_per_bin) bin.push_back((int32_t) f); else dropped++;
					}
				}
			}
		}
	}
#pragma omp parallel for schedule(dynamic, 64) if (parallel)
	for (int64_t i = 0; i < px_count; i++) {
		const int v = (int) (i / width), u = (int) (i % width);
		V2 pixel = {PixelToNdc(u, width, height), PixelToNdc(v, height, width)};
		Hit queue[MAX_POINTS_PER_PIXEL];
		int queue_size = 0;
		float queue_max_depth = -1000.f;
		int queue_max_depth_at = -1;
		if (binned) {
			const auto& bin = bins[(size_t) (v / bin_size) * bins_x + (u / bin_size)];
			for (int32_t f : bin) {
				UpdateQueueIfPixelInsideFace(face_vertices_ndc + (int64_t) f * 9, f, queue, queue_size, queue_max_depth,
				                             queue_max_depth_at, blur_radius, pixel, faces_per_pixel, perspective_correct,
				                             clip_barycentric, cull_back_faces);
			}
		} else {  // brute force, RasterizeNdcTrianglesImpl.h:58-185
			for (int64_t f = 0; f < face_count; f++) {
				if (clip_mask && !clip_mask[f]) continue;
				UpdateQueueIfPixelInsideFace(face_vertices_ndc + f * 9, (int32_t) f, queue, queue_size, queue_max_depth,
				                             queue_max_depth_at, blur_radius, pixel, faces_per_pixel, perspective_correct,
				                             clip_barycentric, cull_back_faces);
			}
		}
		std::sort(queue, queue + queue_size, hit_less);
		for (int k = 0; k < queue_size; k++) {
			int64_t o = i * faces_per_pixel + k;
			pixel_face_indices[o] = queue[k].face;
			pixel_depths[o] = queue[k].depth;
			pixel_face_distances[o] = queue[k].distance;
			pixel_bary[o * 3] = queue[k].bary.x; pixel_bary[o * 3 + 1] = queue[k].bary.y; pixel_bary[o * 3 + 2] = queue[k].bary.z;
		}
	}
	return dropped;
}

// cpp/rendering/functional/kernel/InterpolateFaceAttributesImpl.h:31-71 (float attributes)
ORACLE_API void oracle_interpolate_vertex_attributes(
		const int64_t* pixel_face_indices, const float* pixel_bary, int64_t pixel_count, int faces_per_pixel,
		const float* face_attributes, int channels, float* pixel_attributes) {
	memset(pixel_attributes, 0, sizeof(float) * pixel_count * faces_per_pixel * channels);
	for (int64_t w = 0; w < pixel_count * channels; w++) {
		int64_t i_pixel = w / channels;
		int c = (int) (w % channels);
		for (int k = 0; k < faces_per_pixel; k++) {
			int64_t f = pixel_face_indices[i_pixel * faces_per_pixel + k];
			if (f < 0) break;
			float acc = 0.0f;
			for (int tv = 0; tv < 3; tv++) {
				float weight = pixel_bary[(i_pixel * faces_per_pixel + k) * 3 + tv];
				acc += weight * face_attributes[f * 3 * channels + tv * channels + c];
			}
			pixel_attributes[(i_pixel * faces_per_pixel + k) * channels + c] = acc;
		}
	}
}

// ---- shaders: cpp/rendering/kernel/FlatEdgeShaderImpl.h:30-99, cpp/rendering/VertexColorShader.cpp:29-59 ----------------------------
ORACLE_API void oracle_shade_edges_flat(const int64_t* pixel_face_indices, const float* pixel_face_distances, int height, int width, int faces_per_pixel,
                                        float pixel_line_width, const float* color, uint8_t* pixels) {
	// ImageSpaceDistanceToNdc, cpp/rendering/kernel/CoordinateSystemConversions.h:97-101
	const float ndc_line_width = pixel_line_width / (fminf((float) height, (float) width) / 2.0f);
	const float w2 = ndc_line_width * ndc_line_width;
	memset(pixels, 0, (size_t) height * width * 3);
	for (int64_t px = 0; px < (int64_t) height * width; px++) {
		if (pixel_face_indices[px * faces_per_pixel] == -1) continue;
		float distance = fabsf(pixel_face_distances[px * faces_per_pixel]);
		if (distance < w2) {
			float factor = (w2 - distance) / w2;
			for (int c = 0; c < 3; c++) pixels[px * 3 + c] = (uint8_t) (int32_t) (255.f * color[c] * factor);
		}
	}
}

ORACLE_API void oracle_shade_vertex_colors(const int64_t* pixel_face_indices, const float* pixel_bary, int height, int width, int faces_per_pixel,
                                           const float* face_vertex_colors, uint8_t* pixels) {
	const int64_t pixel_count = (int64_t) height * width;
	std::vector<float> interpolated((size_t) pixel_count * faces_per_pixel * 3);
	oracle_interpolate_vertex_attributes(pixel_face_indices, pixel_bary, pixel_count, faces_per_pixel, face_vertex_colors, 3, interpolated.data());
	for (int64_t px = 0; px < pixel_count; px++)  // .Slice(2, 0, 1) * 255.f, To(UInt8)
		for (int c = 0; c < 3; c++) pixels[px * 3 + c] = (uint8_t) (int32_t) (interpolated[px * faces_per_pixel * 3 + c] * 255.f);  // x86 cvttss2si + low byte
}
