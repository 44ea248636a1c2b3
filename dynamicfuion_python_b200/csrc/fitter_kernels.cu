// Fused per-iteration kernels of the dense-depth fitter: K2 warp + vertex-anchor Jacobians, K3 face setup + inline
// scatter raster, K5 band raster for large faces (K6, residual + Jacobian + J^T J / J^T r assembly, lives in assemble_kernel.cu).
#include "fitter_kernels.cuh"

namespace nnrt_b200 {

// K2. One thread per vertex: linear-blend warp (cpp/geometry/functional/kernel/WarpUtilities.h:447-467) fused with
// WarpedSurfaceJacobians (cpp/alignment/functional/kernel/WarpedSurfaceJacobiansImpl.h:128-156), which the reference
// runs as a separate V*A-thread pass over the same node data.
template<int A>
__global__ void __launch_bounds__(256) warp_vertices_kernel(FitterDeviceData d, const float* __restrict__ rotations,
                                                             const float* __restrict__ translations, Rigid extrinsics) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= d.vertex_count) return;
	const float4 p4 = d.positions[i], n4 = d.normals[i];
	const float3 p = make_float3(p4.x, p4.y, p4.z), n = make_float3(n4.x, n4.y, n4.z);
	const float3 pc = rigid_transform(extrinsics, p);
	const float3 nc = rigid_rotate(extrinsics, n);
	float3 wp = make_float3(0.f, 0.f, 0.f), wn = make_float3(0.f, 0.f, 0.f);
	int idx[A];
	float w[A];
	if (A == 4) {
		const int4 a4 = reinterpret_cast<const int4*>(d.anchors)[i];
		const float4 w4 = reinterpret_cast<const float4*>(d.anchor_weights)[i];
		idx[0] = a4.x; idx[1] = a4.y; idx[2] = a4.z; idx[3] = a4.w;
		w[0] = w4.x; w[1] = w4.y; w[2] = w4.z; w[3] = w4.w;
	} else {
#pragma unroll
		for (int a = 0; a < A; a++) { idx[a] = d.anchors[i * A + a]; w[a] = d.anchor_weights[i * A + a]; }
	}
#pragma unroll
	for (int a = 0; a < A; a++) {
		const int node = idx[a];
		float4 rec0 = make_float4(0.f, 0.f, 0.f, 0.f), rec1 = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
		if (node != -1) {
			float R[9];
#pragma unroll
			for (int k = 0; k < 9; k++) R[k] = __ldg(rotations + (int64_t) node * 9 + k);
			const float3 g = make_float3(__ldg(d.nodes + node * 3), __ldg(d.nodes + node * 3 + 1), __ldg(d.nodes + node * 3 + 2));
			const float3 t = make_float3(__ldg(translations + node * 3), __ldg(translations + node * 3 + 1), __ldg(translations + node * 3 + 2));
			wp = wp + w[a] * ((g + matvec3(R, pc - g)) + t);
			{  // scalar folded into the matrix, as Eigen evaluates w * (R * n) (WarpUtilities.h:466)
				float wR[9];
#pragma unroll
				for (int k = 0; k < 9; k++) wR[k] = w[a] * R[k];
				wn = wn + matvec3(wR, nc);
			}
			// Jacobian records use the canonical (un-extrinsic'd) vertex and normal, as the reference does
			const float3 rv = matvec3(R, p - g);
			const float3 rn = matvec3(R, n);
			rec0 = make_float4(-w[a] * rv.x, -w[a] * rv.y, -w[a] * rv.z, w[a]);
			rec1 = make_float4(-w[a] * rn.x, -w[a] * rn.y, -w[a] * rn.z, __int_as_float(node));
		}
		d.vertex_jacobians[(i * A + a) * 2] = rec0;
		d.vertex_jacobians[(i * A + a) * 2 + 1] = rec1;
	}
	d.warped_positions[i] = make_float4(wp.x, wp.y, wp.z, __fdiv_rn(1.0f, wp.z));  // w = 1 / z for the projections downstream
	d.warped_normals[i] = make_float4(wn.x, wn.y, wn.z, 0.f);
}

__device__ __forceinline__ bool load_fitter_face(const FitterDeviceData& d, const NdcCamera& cam, int f, FaceNdc& out) {
	const int4 tri = d.triangles[f];
	const float4 a = d.warped_positions[tri.x], b = d.warped_positions[tri.y], c = d.warped_positions[tri.z];
	out.v0 = project_with_inverse_depth(cam.k, a.x, a.y, a.w); out.z0 = a.z;
	out.v1 = project_with_inverse_depth(cam.k, b.x, b.y, b.w); out.z1 = b.z;
	out.v2 = project_with_inverse_depth(cam.k, c.x, c.y, c.w); out.z2 = c.z;
	// clip mask of GetMeshNdcFaceVerticesAndClipMask(near 0, far 10): the depth test never rejects (OR of both conditions)
	const bool in0 = out.v0.y >= cam.min_y && out.v0.x >= cam.min_x && out.v0.y <= cam.max_y && out.v0.x <= cam.max_x;
	const bool in1 = out.v1.y >= cam.min_y && out.v1.x >= cam.min_x && out.v1.y <= cam.max_y && out.v1.x <= cam.max_x;
	const bool in2 = out.v2.y >= cam.min_y && out.v2.x >= cam.min_x && out.v2.y <= cam.max_y && out.v2.x <= cam.max_x;
	const bool depth_ok = (a.z >= 0.0f) || (a.z <= 10.0f) || (b.z >= 0.0f) || (b.z <= 10.0f) || (c.z >= 0.0f) || (c.z <= 10.0f);
	return depth_ok && (in0 || in1 || in2);
}

__device__ __forceinline__ RasterParams raster_params(const DataTermParams& p) {
	RasterParams rp;
	rp.height = p.height; rp.width = p.width; rp.blur_radius = p.blur_radius;
	rp.perspective_correct = p.perspective_correct; rp.clip_bary = 0; rp.cull_back_faces = 1;
	rp.ndc_x = p.ndc_x; rp.ndc_y = p.ndc_y;
	return rp;
}

__device__ __forceinline__ FaceBandItem make_band_item(const FaceNdc& f, int face, const PixelBox& box, int y0, int y1) {
	FaceBandItem it;
	it.face = face; it.y0 = y0; it.y1 = y1; it.u0 = box.u0; it.u1 = box.u1;
	it.z0 = f.z0; it.z1 = f.z1; it.z2 = f.z2; it.v0 = f.v0; it.v1 = f.v1; it.v2 = f.v2; it.pad0 = 0.f; it.pad1 = 0.f;
	return it;
}

// K3. One thread per face: gather + project (replaces ExtractFaceVertices + the bbox pass + binning), rasterise inline
// when the pixel box is small, otherwise enqueue row bands.
__global__ void __launch_bounds__(256) face_raster_kernel(FitterDeviceData d, DataTermParams p, NdcCamera cam) {
	const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= d.face_count) return;
	FaceNdc face;
	if (!load_fitter_face(d, cam, (int) f, face)) return;
	const RasterParams rp = raster_params(p);
	PixelBox box;
	if (!face_pixel_box(face, rp, box)) return;
	box.v0 = max(box.v0, p.row_begin);  // pixel-sharded fits: this GPU's band of rows only
	box.v1 = min(box.v1, p.row_end - 1);
	if (box.v0 > box.v1) return;
	int rows_per_band;
	const int bands = raster_face_or_count_bands(face, (int) f, rp, box, d.keys, nullptr, rows_per_band);
	if (bands > 0) {
		const int slot = atomicAdd(d.counters, bands);
		if (slot + bands > d.band_queue_capacity) { atomicExch(d.counters + 1, 1); return; }
		for (int b = 0; b < bands; b++) {
			const int y0 = box.v0 + b * rows_per_band;
			d.band_queue[slot + b] = make_band_item(face, (int) f, box, y0, min(box.v1, y0 + rows_per_band - 1));
		}
	}
}

// K3 for frames whose faces mostly take the queue (large images) and for striped pixel shards: the same per-face work, but only rows
// this GPU owns are walked (fitter_kernels.cuh::row_owned), a larger box queues its band items stripe by stripe so that no warp of the
// drain kernel walks rows of another GPU, and queue slots are reserved with ONE atomic per warp (at 3840 x 2160 nearly every face takes
// the queue path: 81 608 atomics on one counter were a fixed 6 us that no amount of sharding removed).
__global__ void __launch_bounds__(256) face_raster_queue_kernel(FitterDeviceData d, DataTermParams p, NdcCamera cam) {
	const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	const int lane = threadIdx.x & 31;
	const RasterParams rp = raster_params(p);
	FaceNdc face;
	PixelBox box;
	bool valid = f < d.face_count && load_fitter_face(d, cam, (int) f, face) && face_pixel_box(face, rp, box);
	if (valid) {
		box.v0 = max(box.v0, p.row_begin);
		box.v1 = min(box.v1, p.row_end - 1);
		valid = box.v0 <= box.v1;
	}
	const bool striped = p.stripe_count > 1;
	int items = 0, rows_per_band = 1, first = 0, last = 0;
	// rows [y0, y1] of the box inside stripe `stripe` (unstriped: the whole box is the one "stripe" 0)
	auto stripe_rows = [&](int stripe, int& y0, int& y1) {
		y0 = striped ? max(box.v0, p.row_begin + stripe * p.stripe_rows) : box.v0;
		y1 = striped ? min(box.v1, p.row_begin + (stripe + 1) * p.stripe_rows - 1) : box.v1;
	};
	const int step = striped ? p.stripe_count : 1;
	if (valid) {
		const int bw = box.u1 - box.u0 + 1, bh = box.v1 - box.v0 + 1;
		if (bw * bh <= kSmallBoxPixels) {
			const bool within_blur = face_box_within_blur(face, rp.blur_radius);
			for (int v = box.v0; v <= box.v1; v++) {
				if (!row_owned(p, v)) continue;
				for (int u = box.u0; u <= box.u1; u++) raster_pixel(face, (int) f, u, v, rp, d.keys, nullptr, within_blur);
			}
		} else {
			rows_per_band = max(1, kBandPixels / bw);
			if (striped) {
				const int s0 = (box.v0 - p.row_begin) / p.stripe_rows;
				last = (box.v1 - p.row_begin) / p.stripe_rows;
				first = s0 + ((p.stripe_index - s0) % p.stripe_count + p.stripe_count) % p.stripe_count;  // first owned stripe >= s0
			}
			for (int stripe = first; stripe <= last; stripe += step) {
				int y0, y1;
				stripe_rows(stripe, y0, y1);
				items += (y1 - y0 + rows_per_band) / rows_per_band;
			}
		}
	}
	// warp-aggregated reservation: inclusive scan of the item counts, one atomicAdd by the last lane
	int inclusive = items;
#pragma unroll
	for (int offset = 1; offset < 32; offset <<= 1) {
		const int up = __shfl_up_sync(0xffffffffu, inclusive, offset);
		if (lane >= offset) inclusive += up;
	}
	const int warp_items = __shfl_sync(0xffffffffu, inclusive, 31);
	if (warp_items == 0) return;
	int base = 0;
	if (lane == 31) base = atomicAdd(d.counters, warp_items);
	base = __shfl_sync(0xffffffffu, base, 31);
	if (items == 0) return;
	int slot = base + inclusive - items;
	if (slot + items > d.band_queue_capacity) { atomicExch(d.counters + 1, 1); return; }
	for (int stripe = first; stripe <= last; stripe += step) {
		int y0, y1;
		stripe_rows(stripe, y0, y1);
		for (int y = y0; y <= y1; y += rows_per_band) d.band_queue[slot++] = make_band_item(face, (int) f, box, y, min(y1, y + rows_per_band - 1));
	}
}

// K5. One warp per queued row band; the next item's four words are requested before the current item's pixels are walked.
__global__ void __launch_bounds__(256) band_raster_kernel(FitterDeviceData d, DataTermParams p, NdcCamera cam) {
	const RasterParams rp = raster_params(p);
	const int lane = threadIdx.x & 31;
	const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int warp_count = (gridDim.x * blockDim.x) >> 5;
	const int count = min(d.counters[0], d.band_queue_capacity);
	const int4* queue = reinterpret_cast<const int4*>(d.band_queue);
	int item = warp;
	if (item >= count) return;
	int4 w0 = queue[item * 4], w1 = queue[item * 4 + 1], w2 = queue[item * 4 + 2], w3 = queue[item * 4 + 3];
	while (true) {
		const int next = item + warp_count;
		int4 n0 = w0, n1 = w1, n2 = w2, n3 = w3;
		if (next < count) { n0 = queue[next * 4]; n1 = queue[next * 4 + 1]; n2 = queue[next * 4 + 2]; n3 = queue[next * 4 + 3]; }
		FaceNdc f;
		f.z0 = __int_as_float(w1.y); f.z1 = __int_as_float(w1.z); f.z2 = __int_as_float(w1.w);
		f.v0 = make_float2(__int_as_float(w2.x), __int_as_float(w2.y)); f.v1 = make_float2(__int_as_float(w2.z), __int_as_float(w2.w));
		f.v2 = make_float2(__int_as_float(w3.x), __int_as_float(w3.y));
		const int face = w0.x, y0 = w0.y, y1 = w0.z, u0 = w0.w, bw = w1.x - u0 + 1;
		const int total = bw * (y1 - y0 + 1);
		const bool within_blur = face_box_within_blur(f, rp.blur_radius);
		for (int k = lane; k < total; k += 32) raster_pixel(f, face, u0 + k % bw, y0 + k / bw, rp, d.keys, nullptr, within_blur);
		if (next >= count) break;
		item = next; w0 = n0; w1 = n1; w2 = n2; w3 = n3;
	}
	(void) cam;
}

__global__ void __launch_bounds__(256) fill_pixel_centres_kernel(float* __restrict__ ndc_x, float* __restrict__ ndc_y, int height, int width) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < width) ndc_x[i] = pixel_to_ndc(i, width, height);
	if (i < height) ndc_y[i] = pixel_to_ndc(i, height, width);
}

__global__ void __launch_bounds__(256) fill_keys_kernel2(unsigned long long* keys, int64_t n) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) keys[i] = kEmptyKey;
}

}  // namespace nnrt_b200

namespace nnrt_b200 {

void launch_warp_vertices(const FitterDeviceData& d, const float* rotations, const float* translations, const Rigid& extrinsics, cudaStream_t s) {
	const unsigned grid = blocks_for(d.vertex_count, 64);  // small CTAs: the mesh fills the GPU barely once, latency hiding needs every SM
	if (grid == 0) return;
	switch (d.anchor_count) {
#define NNRT_CASE(A_) case A_: warp_vertices_kernel<A_><<<grid, 64, 0, s>>>(d, rotations, translations, extrinsics); break;
		NNRT_CASE(1) NNRT_CASE(2) NNRT_CASE(3) NNRT_CASE(4) NNRT_CASE(5) NNRT_CASE(6) NNRT_CASE(7) NNRT_CASE(8)
#undef NNRT_CASE
	}
}

void launch_face_raster(const FitterDeviceData& d, const DataTermParams& p, const NdcCamera& cam, cudaStream_t s) {
	if (d.face_count == 0) return;
	// the queue-oriented form pays a warp scan per 32 faces: worth it when faces are large (more pixels than faces by 4x: most boxes
	// exceed the inline limit) or the rows are striped
	const bool queue_form = p.stripe_count > 1 || (int64_t) p.height * p.width > 4 * d.face_count * (int64_t) kSmallBoxPixels;
	if (queue_form) face_raster_queue_kernel<<<blocks_for(d.face_count, 64), 64, 0, s>>>(d, p, cam);
	else face_raster_kernel<<<blocks_for(d.face_count, 64), 64, 0, s>>>(d, p, cam);
}

void launch_band_raster(const FitterDeviceData& d, const DataTermParams& p, const NdcCamera& cam, cudaStream_t s) {
	band_raster_kernel<<<148 * 3, 256, 0, s>>>(d, p, cam);  // 77 registers: three CTAs per SM
}

void launch_fill_pixel_centres(float* ndc_x, float* ndc_y, int height, int width, cudaStream_t s) {
	fill_pixel_centres_kernel<<<blocks_for(max(height, width), 256), 256, 0, s>>>(ndc_x, ndc_y, height, width);
}

void launch_fill_keys(unsigned long long* keys, int64_t n, cudaStream_t s) {
	if (n > 0) fill_keys_kernel2<<<blocks_for(n, 256), 256, 0, s>>>(keys, n);
}

}  // namespace nnrt_b200
