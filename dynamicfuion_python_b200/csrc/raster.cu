// Stage-level rendering entry points: NDC face extraction, scatter rasteriser (K = 1..8 by depth peeling),
// attribute interpolation, rasterised-surface Jacobians, face/anchor association.
#include "host_util.h"
#include "raster_core.cuh"
#include "jacobian_math.cuh"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {

__global__ void __launch_bounds__(256) face_vertices_ndc_kernel(const float* __restrict__ positions, const int64_t* __restrict__ triangles,
                                                                 int64_t face_count, NdcCamera cam, float near_clip, float far_clip,
                                                                 float* __restrict__ out, uint8_t* __restrict__ mask) {
	const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= face_count) return;
	float3 v[3];
	float2 xy[3];
	bool in_range = false, has_inlier = false;
#pragma unroll
	for (int k = 0; k < 3; k++) {
		const int64_t vi = triangles[f * 3 + k];
		v[k] = make_float3(positions[vi * 3], positions[vi * 3 + 1], positions[vi * 3 + 2]);
		in_range |= v[k].z >= near_clip;  // sic (reference ORs both tests, ExtractClippedFaceVerticesImpl.h:129-137)
		in_range |= v[k].z <= far_clip;
		xy[k] = project(cam.k, v[k]);
		has_inlier |= (xy[k].y >= cam.min_y && xy[k].x >= cam.min_x && xy[k].y <= cam.max_y && xy[k].x <= cam.max_x);
	}
	const bool keep = in_range && has_inlier;
#pragma unroll
	for (int k = 0; k < 3; k++) {
		out[f * 9 + k * 3 + 0] = keep ? xy[k].x : 0.f;
		out[f * 9 + k * 3 + 1] = keep ? xy[k].y : 0.f;
		out[f * 9 + k * 3 + 2] = keep ? v[k].z : 0.f;
	}
	mask[f] = keep ? 1 : 0;
}

__device__ __forceinline__ void load_face_record(const float* __restrict__ faces, int64_t f, FaceNdc& out) {
	const float* p = faces + f * 9;
	out.v0 = make_float2(p[0], p[1]); out.z0 = p[2];
	out.v1 = make_float2(p[3], p[4]); out.z1 = p[5];
	out.v2 = make_float2(p[6], p[7]); out.z2 = p[8];
}

__global__ void __launch_bounds__(256) raster_faces_kernel(const float* __restrict__ faces, const uint8_t* __restrict__ clip_mask,
                                                            int64_t face_count, RasterParams rp, unsigned long long* __restrict__ keys,
                                                            const unsigned long long* __restrict__ floor_keys, BandItem* __restrict__ queue,
                                                            int* __restrict__ queue_count, int queue_capacity, int* __restrict__ overflow) {
	const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= face_count) return;
	if (clip_mask && !clip_mask[f]) return;
	FaceNdc face;
	load_face_record(faces, f, face);
	PixelBox box;
	if (!face_pixel_box(face, rp, box)) return;
	int rows_per_band;
	const int bands = raster_face_or_count_bands(face, (int) f, rp, box, keys, floor_keys, rows_per_band);
	if (bands > 0) push_bands((int) f, box, rows_per_band, bands, queue, queue_count, queue_capacity, overflow);
}

__global__ void __launch_bounds__(256) raster_bands_kernel(const float* __restrict__ faces, RasterParams rp,
                                                            unsigned long long* __restrict__ keys,
                                                            const unsigned long long* __restrict__ floor_keys,
                                                            const BandItem* __restrict__ queue, const int* __restrict__ queue_count,
                                                            int queue_capacity) {
	drain_band_queue(queue, queue_count, queue_capacity, rp, keys, floor_keys,
	                 [&](int f, FaceNdc& out) { load_face_record(faces, f, out); });
}

// Re-shades the winning face of each pixel and writes the reference's four fragment tensors for layer k.
__global__ void __launch_bounds__(256) resolve_fragments_kernel(const float* __restrict__ faces, RasterParams rp,
                                                                 const unsigned long long* __restrict__ keys, int layer, int faces_per_pixel,
                                                                 int64_t* __restrict__ out_faces, float* __restrict__ out_depths,
                                                                 float* __restrict__ out_bary, float* __restrict__ out_dist) {
	const int64_t p = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= (int64_t) rp.height * rp.width) return;
	const unsigned long long key = keys[p];
	const int64_t o = p * faces_per_pixel + layer;
	if (key == kEmptyKey) {
		out_faces[o] = -1; out_depths[o] = -1.f; out_dist[o] = -1.f;
		out_bary[o * 3] = out_bary[o * 3 + 1] = out_bary[o * 3 + 2] = -1.f;
		return;
	}
	const int f = (int) (key & 0xFFFFFFFFull);
	const int v = (int) (p / rp.width), u = (int) (p % rp.width);
	FaceNdc face;
	load_face_record(faces, f, face);
	Fragment frag;
	shade_pixel(face, make_float2(pixel_to_ndc(u, rp.width, rp.height), pixel_to_ndc(v, rp.height, rp.width)), rp.blur_radius,
	            rp.perspective_correct != 0, rp.clip_bary != 0, rp.cull_back_faces != 0, frag);
	out_faces[o] = f; out_depths[o] = frag.depth; out_dist[o] = frag.distance;
	out_bary[o * 3] = frag.bary.x; out_bary[o * 3 + 1] = frag.bary.y; out_bary[o * 3 + 2] = frag.bary.z;
}

__global__ void __launch_bounds__(256) fill_keys_kernel(unsigned long long* keys, int64_t n) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) keys[i] = kEmptyKey;
}

__global__ void __launch_bounds__(256) interpolate_attributes_kernel(const int64_t* __restrict__ faces, const float* __restrict__ bary,
                                                                      int64_t pixel_count, int faces_per_pixel,
                                                                      const float* __restrict__ attrs, int channels, float* __restrict__ out) {
	const int64_t w = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (w >= pixel_count * channels) return;
	const int64_t px = w / channels;
	const int c = (int) (w % channels);
	bool ended = false;
	for (int k = 0; k < faces_per_pixel; k++) {
		const int64_t f = faces[px * faces_per_pixel + k];
		float acc = 0.0f;
		if (f < 0) ended = true;
		if (!ended) {
			for (int tv = 0; tv < 3; tv++) acc += bary[(px * faces_per_pixel + k) * 3 + tv] * attrs[f * 3 * channels + tv * channels + c];
		}
		out[(px * faces_per_pixel + k) * channels + c] = acc;
	}
}

// rendering::kernel::ShadeEdgesFlat (rendering/kernel/FlatEdgeShaderImpl.h:30-99): the nearest fragment's squared edge distance against
// the squared NDC line width, linear fall-off; pixels without a fragment stay black. One thread per pixel, 3 bytes out.
__global__ void __launch_bounds__(256) shade_edges_flat_kernel(const int64_t* __restrict__ faces, const float* __restrict__ distances, int64_t pixel_count,
                                                               int faces_per_pixel, float ndc_line_width_squared, float r, float g, float b,
                                                               uint8_t* __restrict__ pixels) {
	const int64_t px = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (px >= pixel_count) return;
	uint8_t out[3] = {0, 0, 0};
	if (faces[px * faces_per_pixel] != -1) {
		const float distance = fabsf(distances[px * faces_per_pixel]);
		if (distance < ndc_line_width_squared) {
			const float factor = __fdiv_rn(ndc_line_width_squared - distance, ndc_line_width_squared);
			out[0] = (uint8_t) (int) (255.f * r * factor);
			out[1] = (uint8_t) (int) (255.f * g * factor);
			out[2] = (uint8_t) (int) (255.f * b * factor);
		}
	}
	pixels[px * 3] = out[0]; pixels[px * 3 + 1] = out[1]; pixels[px * 3 + 2] = out[2];
}

// rendering::VertexColorShader::ShadeMeshes (rendering/VertexColorShader.cpp:29-59): barycentric interpolation of the nearest
// fragment's face-vertex colours (InterpolateFaceAttributesImpl.h:31-71, same summation order), times 255, truncated to 8 bits —
// the reference's gather + interpolate over all K fragments + Slice + multiply + To(UInt8) in one pass over the pixels.
__global__ void __launch_bounds__(256) shade_vertex_colors_kernel(const int64_t* __restrict__ faces, const float* __restrict__ bary, int64_t pixel_count,
                                                                  int faces_per_pixel, const float* __restrict__ face_vertex_colors,
                                                                  uint8_t* __restrict__ pixels) {
	const int64_t px = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (px >= pixel_count) return;
	const int64_t f = faces[px * faces_per_pixel];
#pragma unroll
	for (int c = 0; c < 3; c++) {
		float acc = 0.0f;
		if (f >= 0) {
#pragma unroll
			for (int tv = 0; tv < 3; tv++) acc += bary[px * faces_per_pixel * 3 + tv] * face_vertex_colors[f * 9 + tv * 3 + c];
		}
		pixels[px * 3 + c] = (uint8_t) (int) (acc * 255.f);  // low byte of the truncated value: what the x86 reference build's cast yields
		                                                      // for colours pushed outside [0, 255] by barycentrics of blurred (outside-the-face) fragments
	}
}

__global__ void __launch_bounds__(128) rasterized_surface_jacobians_kernel(
		const float* __restrict__ positions, const float* __restrict__ normals, const int64_t* __restrict__ triangles,
		const int64_t* __restrict__ pixel_faces, const float* __restrict__ pixel_bary, int height, int width, int faces_per_pixel,
		Intrinsics ndc, int perspective_correct, float* __restrict__ out_v, float* __restrict__ out_n) {
	const int64_t p = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= (int64_t) height * width) return;
	float* ov = out_v + p * 27;
	float* on = out_n + p * 30;
	const int64_t f = pixel_faces[p * faces_per_pixel];
	if (f == -1) {
		for (int k = 0; k < 27; k++) ov[k] = 0.f;
		for (int k = 0; k < 30; k++) on[k] = 0.f;
		return;
	}
	const int v_image = (int) (p / width), u_image = (int) (p % width);
	const float2 ray = make_float2(pixel_to_ndc(u_image, width, height), pixel_to_ndc(v_image, height, width));
	float3 fv[3], fn[3];
#pragma unroll
	for (int k = 0; k < 3; k++) {
		const int64_t vi = triangles[f * 3 + k];
		fv[k] = make_float3(positions[vi * 3], positions[vi * 3 + 1], positions[vi * 3 + 2]);
		fn[k] = make_float3(normals[vi * 3], normals[vi * 3 + 1], normals[vi * 3 + 2]);
	}
	const float* bary = pixel_bary + p * faces_per_pixel * 3;
	const float rho[3] = {bary[0], bary[1], bary[2]};
	BaryJacobian J;
	barycentric_jacobian(J, ray, fv, ndc, perspective_correct != 0, rho);
#pragma unroll
	for (int r = 0; r < 3; r++)
#pragma unroll
		for (int c = 0; c < 9; c++) {
			float val = col_mix(fv, r, J, c);
			if (c % 3 == r) val += rho[c / 3];
			ov[r * 9 + c] = val;
			on[r * 9 + c] = col_mix(fn, r, J, c);
		}
	on[27] = rho[0]; on[28] = rho[1]; on[29] = rho[2];
}

// Face -> unique anchor nodes, first-appearance order over (face vertex 0..2, anchor slot 0..A-1).
__global__ void __launch_bounds__(256) associate_faces_kernel(const int64_t* __restrict__ triangles, int64_t face_count,
                                                               const int32_t* __restrict__ vertex_anchors, int anchor_count,
                                                               int32_t* __restrict__ entries, int32_t* __restrict__ counts) {
	const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= face_count) return;
	const int max_entries = 3 * anchor_count;
	int32_t* e = entries + f * max_entries * 2;
	int count = 0;
	for (int fv = 0; fv < 3; fv++) {
		const int64_t vi = triangles[f * 3 + fv];
		for (int a = 0; a < anchor_count; a++) {
			const int node = vertex_anchors[vi * anchor_count + a];
			if (node == -1) continue;
			int k = 0;
			while (k < count && e[k * 2] != node) k++;
			if (k == count) { e[k * 2] = node; e[k * 2 + 1] = (int32_t) 0xFFFFFFFF; count++; }
			// byte fv of the packed word = anchor slot (a later duplicate within the same vertex overwrites, as in the reference)
			const uint32_t packed = (uint32_t) e[k * 2 + 1];
			e[k * 2 + 1] = (int32_t) ((packed & ~(0xFFu << (8 * fv))) | ((uint32_t) a << (8 * fv)));
		}
	}
	for (int k = count; k < max_entries; k++) { e[k * 2] = -1; e[k * 2 + 1] = (int32_t) 0xFFFFFFFF; }
	counts[f] = count;
}

}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_face_vertices_ndc(const float* vertex_positions, int64_t vertex_count, const int64_t* triangles, int64_t face_count,
                                           const double* intrinsics_host, int32_t image_height, int32_t image_width, float near_clip,
                                           float far_clip, float* face_vertices_ndc, uint8_t* clip_mask, void* stream) {
	NNRT_CHECK_ARG(vertex_positions && triangles && intrinsics_host && face_vertices_ndc && clip_mask, "face_vertices_ndc: null pointer argument");
	NNRT_CHECK_ARG(near_clip >= 0.0f, "near_clipping_distance cannot be less than 0.0. Got %f.", near_clip);
	NNRT_CHECK_ARG(near_clip <= far_clip, "near_clipping_distance cannot be greater than far_clipping_distance. Got %f and %f, respectively.", near_clip, far_clip);
	(void) vertex_count;
	if (face_count == 0) return kOk;
	face_vertices_ndc_kernel<<<blocks_for(face_count, 256), 256, 0, (cudaStream_t) stream>>>(
			vertex_positions, triangles, face_count, ndc_camera_from_host(intrinsics_host, image_height, image_width), near_clip, far_clip,
			face_vertices_ndc, clip_mask);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_rasterize_ndc_triangles(const float* face_vertices_ndc, const uint8_t* clip_mask, int64_t face_count,
                                                 int32_t image_height, int32_t image_width, float blur_radius_pixels, int32_t faces_per_pixel,
                                                 int32_t bin_size, int32_t max_faces_per_bin, int32_t perspective_correct,
                                                 int32_t clip_barycentric, int32_t cull_back_faces, int64_t* pixel_face_indices,
                                                 float* pixel_depths, float* pixel_barycentric, float* pixel_face_distances, void* stream) {
	NNRT_CHECK_ARG(faces_per_pixel >= 1 && faces_per_pixel <= kMaxFacesPerPixel, "Need faces_per_pixel <= %d. Got: %d", kMaxFacesPerPixel, faces_per_pixel);
	NNRT_CHECK_ARG(image_height > 0 && image_width > 0, "image_size must be positive");
	NNRT_CHECK_ARG(pixel_face_indices && pixel_depths && pixel_barycentric && pixel_face_distances, "rasterize: null output pointer");
	NNRT_CHECK_ARG(face_count == 0 || face_vertices_ndc, "rasterize: null face vertices");
	NNRT_CHECK_ARG(face_count <= 0x7FFFFFFFll, "face_count exceeds int32 face indices");
	if (bin_size > 0) {  // same argument check as cpp/rendering/RasterizeNdcTriangles.cpp:70-76
		const int max_dim = image_height > image_width ? image_height : image_width;
		const int bins = 1 + (max_dim - 1) / bin_size;
		NNRT_CHECK_ARG(bins < 22, "The provided bin_size is too small: computed bin count along maximum dimension is %d, this has to be < 22.", bins);
	}
	(void) max_faces_per_bin;
	auto s = (cudaStream_t) stream;
	const int64_t px = (int64_t) image_height * image_width;
	RasterParams rp;
	rp.height = image_height; rp.width = image_width;
	rp.blur_radius = blur_radius_pixels / (fminf((float) image_height, (float) image_width) / 2.0f);
	rp.perspective_correct = perspective_correct; rp.clip_bary = clip_barycentric; rp.cull_back_faces = cull_back_faces;
	rp.ndc_x = nullptr; rp.ndc_y = nullptr;  // the one-off stage entry computes the pixel centres on the fly
	const int queue_capacity = (int) (face_count > (1 << 20) ? face_count : (1 << 20));
	unsigned long long* keys = nullptr;
	BandItem* queue = nullptr;
	int* counters = nullptr;
	Scratch scratch(s);
	NNRT_CUDA(scratch.alloc(&keys, sizeof(unsigned long long) * px * 2));
	NNRT_CUDA(scratch.alloc(&queue, sizeof(BandItem) * queue_capacity));
	NNRT_CUDA(scratch.alloc(&counters, sizeof(int) * 2));
	unsigned long long* cur = keys;
	unsigned long long* prev = keys + px;
	int overflow_host = 0;
	for (int layer = 0; layer < faces_per_pixel; layer++) {
		fill_keys_kernel<<<blocks_for(px, 256), 256, 0, s>>>(cur, px);
		NNRT_CUDA(cudaMemsetAsync(counters, 0, sizeof(int) * 2, s));
		if (face_count > 0) {
			raster_faces_kernel<<<blocks_for(face_count, 256), 256, 0, s>>>(face_vertices_ndc, clip_mask, face_count, rp, cur,
			                                                                layer == 0 ? nullptr : prev, queue, counters, queue_capacity,
			                                                                counters + 1);
			raster_bands_kernel<<<148 * 4, 256, 0, s>>>(face_vertices_ndc, rp, cur, layer == 0 ? nullptr : prev, queue, counters,
			                                            queue_capacity);
		}
		resolve_fragments_kernel<<<blocks_for(px, 256), 256, 0, s>>>(face_vertices_ndc, rp, cur, layer, faces_per_pixel, pixel_face_indices,
		                                                             pixel_depths, pixel_barycentric, pixel_face_distances);
		NNRT_LAUNCH_CHECK();
		int flag = 0;
		NNRT_CUDA(cudaMemcpyAsync(&flag, counters + 1, sizeof(int), cudaMemcpyDeviceToHost, s));
		NNRT_CUDA(cudaStreamSynchronize(s));
		overflow_host |= flag;
		unsigned long long* t = cur; cur = prev; prev = t;
	}
	if (overflow_host) {
		set_last_error("rasterize: large-face band queue overflow (capacity %d)", queue_capacity);
		return kQueueOverflow;
	}
	return kOk;
}

extern "C" int nnrt_b200_interpolate_vertex_attributes(const int64_t* pixel_face_indices, const float* pixel_barycentric, int64_t pixel_count,
                                                       int32_t faces_per_pixel, const float* face_attributes, int32_t channels,
                                                       float* pixel_attributes, void* stream) {
	NNRT_CHECK_ARG(pixel_face_indices && pixel_barycentric && face_attributes && pixel_attributes, "interpolate_vertex_attributes: null pointer argument");
	if (pixel_count * channels == 0) return kOk;
	interpolate_attributes_kernel<<<blocks_for(pixel_count * channels, 256), 256, 0, (cudaStream_t) stream>>>(
			pixel_face_indices, pixel_barycentric, pixel_count, faces_per_pixel, face_attributes, channels, pixel_attributes);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_rasterized_surface_jacobians(const float* warped_positions, const float* warped_normals, const int64_t* triangles,
                                                      const int64_t* pixel_face_indices, const float* pixel_barycentric, int32_t image_height,
                                                      int32_t image_width, int32_t faces_per_pixel, const double* ndc_intrinsics_host,
                                                      int32_t perspective_correct, float* rendered_vertex_jacobians,
                                                      float* rendered_normal_jacobians, void* stream) {
	NNRT_CHECK_ARG(warped_positions && warped_normals && triangles && pixel_face_indices && pixel_barycentric && ndc_intrinsics_host &&
	               rendered_vertex_jacobians && rendered_normal_jacobians, "rasterized_surface_jacobians: null pointer argument");
	const int64_t px = (int64_t) image_height * image_width;
	if (px == 0) return kOk;
	rasterized_surface_jacobians_kernel<<<blocks_for(px, 128), 128, 0, (cudaStream_t) stream>>>(
			warped_positions, warped_normals, triangles, pixel_face_indices, pixel_barycentric, image_height, image_width, faces_per_pixel,
			intrinsics_from_host(ndc_intrinsics_host), perspective_correct, rendered_vertex_jacobians, rendered_normal_jacobians);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

// device check of common.cuh::SharedDivisor: both quotients of every pair, for a bitwise comparison by the test
__global__ void shared_divisor_quotients_kernel(const float* __restrict__ numerators, const float* __restrict__ divisors, int64_t count,
                                                float* __restrict__ shared_quotients, float* __restrict__ ieee_quotients) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	const float a = numerators[i], d = divisors[i];
	shared_quotients[i] = a / shared_divisor(d);
	ieee_quotients[i] = __fdiv_rn(a, d);
}

extern "C" int nnrt_b200_selfcheck_shared_divisor_quotients(const float* numerators, const float* divisors, int64_t count,
                                                            float* shared_quotients, float* ieee_quotients, void* stream) {
	NNRT_CHECK_ARG(numerators && divisors && shared_quotients && ieee_quotients, "selfcheck_shared_divisor_quotients: null pointer argument");
	if (count == 0) return kOk;
	shared_divisor_quotients_kernel<<<blocks_for(count, 256), 256, 0, (cudaStream_t) stream>>>(numerators, divisors, count, shared_quotients,
	                                                                                           ieee_quotients);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_associate_faces_with_anchors(const int64_t* triangles, int64_t face_count, const int32_t* vertex_anchors,
                                                      int32_t anchor_count, int32_t* face_anchor_entries, int32_t* face_anchor_counts,
                                                      void* stream) {
	NNRT_CHECK_ARG(triangles && vertex_anchors && face_anchor_entries && face_anchor_counts, "associate_faces_with_anchors: null pointer argument");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	if (face_count == 0) return kOk;
	associate_faces_kernel<<<blocks_for(face_count, 256), 256, 0, (cudaStream_t) stream>>>(triangles, face_count, vertex_anchors, anchor_count,
	                                                                                       face_anchor_entries, face_anchor_counts);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_shade_edges_flat(const int64_t* pixel_face_indices, const float* pixel_face_distances, int32_t image_height, int32_t image_width,
                                          int32_t faces_per_pixel, float pixel_line_width, const float* line_color_host, uint8_t* pixels, void* stream) {
	NNRT_CHECK_ARG(pixel_face_indices && pixel_face_distances && line_color_host && pixels, "shade_edges_flat: null pointer argument");
	NNRT_CHECK_ARG(image_height > 0 && image_width > 0 && faces_per_pixel >= 1, "shade_edges_flat: bad image size or faces_per_pixel");
	// kernel::ImageSpaceDistanceToNdc (rendering/kernel/CoordinateSystemConversions.h:97-101)
	const float ndc_line_width = pixel_line_width / ((float) (image_height < image_width ? image_height : image_width) / 2.0f);
	NNRT_CHECK_ARG(ndc_line_width > 0.f, "Expecting a non-zero, positive ndc_line_width. Got: %f.", ndc_line_width);
	const int64_t px = (int64_t) image_height * image_width;
	shade_edges_flat_kernel<<<blocks_for(px, 256), 256, 0, (cudaStream_t) stream>>>(pixel_face_indices, pixel_face_distances, px, faces_per_pixel,
	                                                                              ndc_line_width * ndc_line_width, line_color_host[0], line_color_host[1],
	                                                                              line_color_host[2], pixels);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_shade_vertex_colors(const int64_t* pixel_face_indices, const float* pixel_barycentric, int32_t image_height, int32_t image_width,
                                             int32_t faces_per_pixel, const float* face_vertex_colors, uint8_t* pixels, void* stream) {
	NNRT_CHECK_ARG(pixel_face_indices && pixel_barycentric && face_vertex_colors && pixels, "shade_vertex_colors: null pointer argument");
	NNRT_CHECK_ARG(image_height > 0 && image_width > 0 && faces_per_pixel >= 1, "shade_vertex_colors: bad image size or faces_per_pixel");
	const int64_t px = (int64_t) image_height * image_width;
	shade_vertex_colors_kernel<<<blocks_for(px, 256), 256, 0, (cudaStream_t) stream>>>(pixel_face_indices, pixel_barycentric, px, faces_per_pixel,
	                                                                                 face_vertex_colors, pixels);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
