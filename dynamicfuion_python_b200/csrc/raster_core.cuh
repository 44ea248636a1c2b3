// Scatter rasteriser core shared by the stage-level API (raster.cu) and the fused fitter path (fitter_kernels.cu).
//
// B200 design: no bins. Each face is projected once, its pixel bounding box is walked, every pixel centre is tested
// with the reference's exact per-pixel routine (common.cuh::shade_pixel) and the winner is resolved with a 64-bit
// atomicMin on the key (depth bits << 32 | face index), which equals the reference's (depth, face index) ordering
// (cpp/rendering/kernel/RayFaceIntersection.h:41-45) and is independent of thread scheduling. Faces whose box is at
// most kSmallBoxPixels are rasterised inline by the thread that set them up; larger faces are split into row bands
// of <= ~1024 pixels and pushed to a work queue drained by one warp per band (second kernel).
// K > 1 faces per pixel: depth peeling, pass k only accepts keys greater than the pixel's pass k-1 winner.
#pragma once
#include "common.cuh"

namespace nnrt_b200 {

constexpr int kSmallBoxPixels = 16;
constexpr int kBandPixels = 1024;

struct RasterParams {
	int height, width;
	float blur_radius;      // NDC units
	int perspective_correct, clip_bary, cull_back_faces;
	// optional tables of the pixel centres in NDC, ndc_x[u] = pixel_to_ndc(u, width, height), ndc_y[v] = pixel_to_ndc(v, height, width),
	// filled by that very function (bit-identical values): a load instead of an IEEE division per use. nullptr: computed on the fly.
	const float* ndc_x; const float* ndc_y;
};

__device__ __forceinline__ float pixel_centre(const float* __restrict__ table, int i, int dim1, int dim2) {
	return table ? __ldg(table + i) : pixel_to_ndc(i, dim1, dim2);
}

struct BandItem { int face, y0, y1, pad; };

struct PixelBox { int u0, u1, v0, v1; };  // inclusive, empty if u0 > u1 or v0 > v1

// Exact integer range of pixels whose centres pass the reference's bounding-box test lo <= centre(i) <= hi (the same float
// expression shade_pixel evaluates; it is monotone in i, so a float estimate corrected by stepping is exact and tight).
__device__ __forceinline__ void pixel_range(float lo, float hi, int dim1, int dim2, int& i0, int& i1, const float* __restrict__ table = nullptr) {
	if (!(lo <= hi)) { i0 = 0; i1 = dim1 - 1; return; }  // NaN bounds: shade_pixel's comparisons reject nothing, walk the whole axis
	const float range = ndc_range(dim1, dim2);
	const float offset = range / 2.0f;
	// centre(i) = -offset + (range*i + offset)/dim1  =>  i = ((x + offset)*dim1 - offset)/range
	// (the estimate only has to land near the answer -- the loops below step to the exact range -- so it may use the fast division)
	const float f0 = __fdividef((lo + offset) * (float) dim1 - offset, range);
	const float f1 = __fdividef((hi + offset) * (float) dim1 - offset, range);
	i0 = (int) ceilf(fminf(fmaxf(f0, 0.f), (float) dim1));
	i1 = (int) floorf(fminf(fmaxf(f1, -1.f), (float) (dim1 - 1)));
	while (i0 > 0 && pixel_centre(table, i0 - 1, dim1, dim2) >= lo) i0--;
	while (i0 < dim1 && pixel_centre(table, i0, dim1, dim2) < lo) i0++;
	while (i1 < dim1 - 1 && pixel_centre(table, i1 + 1, dim1, dim2) <= hi) i1++;
	while (i1 >= 0 && pixel_centre(table, i1, dim1, dim2) > hi) i1--;
}

__device__ __forceinline__ bool face_pixel_box(const FaceNdc& f, const RasterParams& rp, PixelBox& box) {
	const float face_area = edge_fn(f.v0, f.v1, f.v2);
	if (rp.cull_back_faces && face_area < 0.f) return false;
	if (face_area <= kEps && face_area >= -1.f * kEps) return false;
	if (fmax3(f.z0, f.z1, f.z2) < kEps) return false;
	const float x_min = fmin3(f.v0.x, f.v1.x, f.v2.x) - rp.blur_radius, x_max = fmax3(f.v0.x, f.v1.x, f.v2.x) + rp.blur_radius;
	const float y_min = fmin3(f.v0.y, f.v1.y, f.v2.y) - rp.blur_radius, y_max = fmax3(f.v0.y, f.v1.y, f.v2.y) + rp.blur_radius;
	pixel_range(x_min, x_max, rp.width, rp.height, box.u0, box.u1, rp.ndc_x);
	pixel_range(y_min, y_max, rp.height, rp.width, box.v0, box.v1, rp.ndc_y);
	return box.u0 <= box.u1 && box.v0 <= box.v1;
}

__device__ __forceinline__ void raster_pixel(const FaceNdc& f, int face, int u, int v, const RasterParams& rp,
                                             unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ floor_keys,
                                             bool within_blur) {
	const float2 pixel = make_float2(pixel_centre(rp.ndc_x, u, rp.width, rp.height), pixel_centre(rp.ndc_y, v, rp.height, rp.width));
	Fragment frag;
	if (!shade_pixel(f, pixel, rp.blur_radius, rp.perspective_correct != 0, rp.clip_bary != 0, rp.cull_back_faces != 0, frag, within_blur)) return;
	const unsigned long long key = pack_key(frag.depth, face);
	const int64_t p = (int64_t) v * rp.width + u;
	if (floor_keys) {  // depth peeling: strictly after the previous layer's winner
		const unsigned long long prev = floor_keys[p];
		if (prev == kEmptyKey || key <= prev) return;
	}
	if (key < keys[p]) atomicMin(keys + p, key);
}

// Inline path for small boxes; returns the number of band items the face needs otherwise (0 = done).
__device__ __forceinline__ int raster_face_or_count_bands(const FaceNdc& f, int face, const RasterParams& rp, const PixelBox& box,
                                                          unsigned long long* __restrict__ keys,
                                                          const unsigned long long* __restrict__ floor_keys, int& rows_per_band) {
	const int bw = box.u1 - box.u0 + 1, bh = box.v1 - box.v0 + 1;
	if (bw * bh <= kSmallBoxPixels) {
		const bool within_blur = face_box_within_blur(f, rp.blur_radius);
		for (int v = box.v0; v <= box.v1; v++)
			for (int u = box.u0; u <= box.u1; u++) raster_pixel(f, face, u, v, rp, keys, floor_keys, within_blur);
		return 0;
	}
	rows_per_band = max(1, kBandPixels / bw);
	return (bh + rows_per_band - 1) / rows_per_band;
}

__device__ __forceinline__ void push_bands(int face, const PixelBox& box, int rows_per_band, int band_count, BandItem* __restrict__ queue,
                                           int* __restrict__ queue_count, int queue_capacity, int* __restrict__ overflow_flag) {
	const int slot = atomicAdd(queue_count, band_count);
	if (slot + band_count > queue_capacity) { atomicExch(overflow_flag, 1); return; }
	for (int b = 0; b < band_count; b++) {
		const int y0 = box.v0 + b * rows_per_band;
		queue[slot + b] = BandItem{face, y0, min(box.v1, y0 + rows_per_band - 1), 0};
	}
}

// One warp per band item; `load_face(face, FaceNdc&)` re-derives the face's NDC vertices.
template<typename TLoadFace>
__device__ __forceinline__ void drain_band_queue(const BandItem* __restrict__ queue, const int* __restrict__ queue_count, int queue_capacity,
                                                 const RasterParams& rp, unsigned long long* __restrict__ keys,
                                                 const unsigned long long* __restrict__ floor_keys, TLoadFace&& load_face) {
	const int lane = threadIdx.x & 31;
	const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int warp_count = (gridDim.x * blockDim.x) >> 5;
	const int count = min(*queue_count, queue_capacity);
	for (int item = warp; item < count; item += warp_count) {
		const BandItem it = queue[item];
		FaceNdc f;
		load_face(it.face, f);
		PixelBox box;
		if (!face_pixel_box(f, rp, box)) continue;
		const int bw = box.u1 - box.u0 + 1;
		const int total = bw * (it.y1 - it.y0 + 1);
		const bool within_blur = face_box_within_blur(f, rp.blur_radius);
		for (int k = lane; k < total; k += 32) raster_pixel(f, it.face, box.u0 + k % bw, it.y0 + k / bw, rp, keys, floor_keys, within_blur);
	}
}

}  // namespace nnrt_b200
