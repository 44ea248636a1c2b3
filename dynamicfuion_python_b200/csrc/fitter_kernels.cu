// Fused per-iteration kernels of the dense-depth fitter: K2 warp + vertex-anchor Jacobians, K3 face setup + inline
// scatter raster, K5 band raster for large faces, K6 residual + Jacobian + J^T J / J^T r assembly.
#include "fitter_kernels.cuh"
#include "jacobian_math.cuh"

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
	d.warped_positions[i] = make_float4(wp.x, wp.y, wp.z, 0.f);
	d.warped_normals[i] = make_float4(wn.x, wn.y, wn.z, 0.f);
}

__device__ __forceinline__ bool load_fitter_face(const FitterDeviceData& d, const NdcCamera& cam, int f, FaceNdc& out) {
	const int4 tri = d.triangles[f];
	const float4 a = d.warped_positions[tri.x], b = d.warped_positions[tri.y], c = d.warped_positions[tri.z];
	out.v0 = project(cam.k, make_float3(a.x, a.y, a.z)); out.z0 = a.z;
	out.v1 = project(cam.k, make_float3(b.x, b.y, b.z)); out.z1 = b.z;
	out.v2 = project(cam.k, make_float3(c.x, c.y, c.z)); out.z2 = c.z;
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
	return rp;
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
	if (bands > 0) push_bands((int) f, box, rows_per_band, bands, d.band_queue, d.counters, d.band_queue_capacity, d.counters + 1);
}

// K5. One warp per queued row band.
__global__ void __launch_bounds__(256) band_raster_kernel(FitterDeviceData d, DataTermParams p, NdcCamera cam) {
	const RasterParams rp = raster_params(p);
	drain_band_queue(d.band_queue, d.counters, d.band_queue_capacity, rp, d.keys, nullptr,
	                 [&](int f, FaceNdc& out) { load_fitter_face(d, cam, f, out); });
}

__global__ void __launch_bounds__(256) fill_keys_kernel2(unsigned long long* keys, int64_t n) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) keys[i] = kEmptyKey;
}

}  // namespace nnrt_b200

namespace nnrt_b200 {

// Transposed warp reduction: every lane passes 32 values, lane l receives the warp-wide sum of value l (31 shuffles).
__device__ __forceinline__ float warp_transpose_reduce(float (&vals)[32], int lane) {
#pragma unroll
	for (int half = 16; half >= 1; half >>= 1) {
		const bool upper = (lane & half) != 0;
#pragma unroll
		for (int i = 0; i < half; i++) {
			const float send = upper ? vals[i] : vals[i + half];
			const float keep = upper ? vals[i + half] : vals[i];
			vals[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, half);
		}
	}
	return vals[0];
}

// row-vector times skew(s):  r * [[0,-sz,sy],[sz,0,-sx],[-sy,sx,0]]
__device__ __forceinline__ float3 row_times_skew(float3 r, float3 s) {
	return make_float3(sum3(r.x * 0.f, r.y * s.z, r.z * -s.y), sum3(r.x * -s.z, r.y * 0.f, r.z * s.x), sum3(r.x * s.y, r.y * -s.x, r.z * 0.f));
}

__device__ __forceinline__ void cp_async_16(void* shared_dst, const void* global_src) {
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned) __cvta_generic_to_shared(shared_dst)), "l"(global_src) : "memory");
}
__device__ __forceinline__ void cp_async_8(void* shared_dst, const void* global_src) {
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned) __cvta_generic_to_shared(shared_dst)), "l"(global_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

constexpr int kAssembleThreads = 128;
__host__ __device__ constexpr size_t assemble_shared_bytes(int anchor_count) {
	return (size_t) kAssembleThreads * 3 * anchor_count * 2 * sizeof(float4);
}

// K6. One thread per pixel (16x8 tile per CTA, 8x4 pixels per warp). Replaces InterpolateVertexAttributes, the
// per-iteration unprojection, ComputePointToPlaneDistances, masking, RasterizedSurfaceJacobians,
// PixelVertexAnchorJacobiansAndNodeAssociations and both *_UnorderedNodePixels reductions
// (cpp/alignment/DeformableMeshToImageFitter.cpp:133-220): everything per pixel stays on chip, per-node sums are
// formed by a warp-level reduction keyed on node id and flushed with one float atomic per (warp, node, entry).
// The kernel is bound by the latency of its dependent gathers (key -> face -> vertices -> anchor records), so every load
// that hangs off the face index is issued in one burst: the 3A vertex-anchor records and the face's node table travel
// global -> shared with cp.async while the residual and the barycentric chain are computed from registers; the node-voting
// loop then touches shared memory only.
template<int A>
__global__ void __launch_bounds__(kAssembleThreads) assemble_kernel(FitterDeviceData d, DataTermParams p) {
	constexpr int kMaxEntries = 3 * A, T = kAssembleThreads;
	extern __shared__ __align__(16) unsigned char assemble_smem[];
	float4* s_rec = reinterpret_cast<float4*>(assemble_smem);  // [3A * 2][T]: anchor records, then J contributions (+ node id)
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	// tiles are visited in a scrambled order (multiplication by a prime modulo the tile count): the surface covers a compact
	// part of the image, and in raster order its expensive tiles would reach the SMs as one late clump
	const unsigned tile_count = gridDim.x * gridDim.y, linear = blockIdx.y * gridDim.x + blockIdx.x;
	const unsigned tile = (tile_count % 1009u) ? (unsigned) (((unsigned long long) linear * 1009ull) % tile_count) : linear;
	const int tile_x = (int) (tile % gridDim.x) * 16 + (warp & 1) * 8, tile_y = p.row_begin + (int) (tile / gridDim.x) * 8 + (warp >> 1) * 4;
	const int u = tile_x + (lane & 7), v = tile_y + (lane >> 3);
	const bool in_image = u < p.width && v < p.row_end;
	const int64_t px = (int64_t) v * p.width + u;
	unsigned long long key = kEmptyKey;
	if (in_image) key = d.keys[px];
	const bool covered = key != kEmptyKey;
	int face = -1;
	float residual = 0.f;
	bool residual_mask = false;
	unsigned pending = 0u;     // bit q = fvi * A + slot: that (face vertex, anchor slot) pair still has to be accumulated
	int nodes[kMaxEntries];
#pragma unroll
	for (int q = 0; q < kMaxEntries; q++) nodes[q] = -1;
	if (covered) {
		d.keys[px] = kEmptyKey;  // leave the key buffer clean for the next iteration's scatter
		face = (int) (key & 0xFFFFFFFFull);
		const int4 tri = d.triangles[face];
		{
#pragma unroll
			for (int fvi = 0; fvi < 3; fvi++) {
				const int vi = fvi == 0 ? tri.x : fvi == 1 ? tri.y : tri.z;
				const float4* rec = d.vertex_jacobians + (int64_t) vi * A * 2;
#pragma unroll
				for (int q = 0; q < 2 * A; q++) cp_async_16(s_rec + (fvi * 2 * A + q) * T + tid, rec + q);
			}
		}
		float3 fv[3], fn[3];
		{
			const float4 a = d.warped_positions[tri.x], b = d.warped_positions[tri.y], c = d.warped_positions[tri.z];
			fv[0] = make_float3(a.x, a.y, a.z); fv[1] = make_float3(b.x, b.y, b.z); fv[2] = make_float3(c.x, c.y, c.z);
			const float4 na = d.warped_normals[tri.x], nb = d.warped_normals[tri.y], nc = d.warped_normals[tri.z];
			fn[0] = make_float3(na.x, na.y, na.z); fn[1] = make_float3(nb.x, nb.y, nb.z); fn[2] = make_float3(nc.x, nc.y, nc.z);
		}
		const float4 ref = d.reference_points[px];
		const float2 ray = make_float2(pixel_to_ndc(u, p.width, p.height), pixel_to_ndc(v, p.height, p.width));
		FaceNdc fndc;
		fndc.v0 = project(p.ndc, fv[0]); fndc.v1 = project(p.ndc, fv[1]); fndc.v2 = project(p.ndc, fv[2]);
		fndc.z0 = fv[0].z; fndc.z1 = fv[1].z; fndc.z2 = fv[2].z;
		Fragment frag;
		shade_pixel(fndc, ray, p.blur_radius, p.perspective_correct != 0, false, true, frag, true);  // the raster pass accepted this pixel
		const float rho[3] = {frag.bary.x, frag.bary.y, frag.bary.z};
		const float depth = frag.depth;
		// rendered normal (InterpolateFaceAttributesImpl.h:56-62: acc = 0; acc += w*attr)
		float3 nl;
		nl.x = ((0.0f + rho[0] * fn[0].x) + rho[1] * fn[1].x) + rho[2] * fn[2].x;
		nl.y = ((0.0f + rho[0] * fn[0].y) + rho[1] * fn[1].y) + rho[2] * fn[2].y;
		nl.z = ((0.0f + rho[0] * fn[0].z) + rho[1] * fn[1].z) + rho[2] * fn[2].z;
		// unprojection of the rendered depth at the integer pixel with the pinhole intrinsics (PerspectiveProjectionImpl.h:119-131)
		const bool rendered_valid = depth > 0 && depth < p.max_depth;
		float3 wl = make_float3(0.f, 0.f, 0.f);
		if (rendered_valid) wl = make_float3(((float) u - p.pinhole.cx) * depth / p.pinhole.fx, ((float) v - p.pinhole.cy) * depth / p.pinhole.fy, depth);
		const bool mask = rendered_valid && ref.w != 0.f;
		const float3 pm = make_float3(wl.x - ref.x, wl.y - ref.y, wl.z - ref.z);
		const float r_raw = dot3(nl.x, nl.y, nl.z, pm.x, pm.y, pm.z);
		residual = mask ? r_raw : 0.f;
		bool contributes = mask;
		residual_mask = mask;
		float3 dr_dnl = pm, dr_dwl = nl;
		if (p.use_tukey) {
			// residual side (DeformableMeshToImageFitter.cpp:377-389, inverted overwrite kept); Jacobian side (PixelVertexAnchorJacobiansImpl.h:200-219)
			const float c = p.tukey_cutoff, c2_6 = c * c / 6.f;
			const float q = residual / c, l = 1.f - q * q;
			float rt = c2_6 * (1.f - l * l * l);
			if (residual <= c) rt = c2_6;
			if (fabsf(r_raw) > c) contributes = false;
			const float qq = r_raw / c;
			float psi = 1 - qq * qq;
			psi = r_raw * psi * psi;
			dr_dnl = psi * pm;
			dr_dwl = psi * nl;
			residual = rt;
		}
		float dr_dV[9], dr_dN[9];
		if (contributes) {
			BaryJacobian J;
			barycentric_jacobian(J, ray, fv, p.ndc, p.perspective_correct != 0, rho);
#pragma unroll
			for (int c = 0; c < 9; c++) {
				const float w0 = col_mix(fv, 0, J, c) + (c % 3 == 0 ? rho[c / 3] : 0.f);
				const float w1 = col_mix(fv, 1, J, c) + (c % 3 == 1 ? rho[c / 3] : 0.f);
				const float w2 = col_mix(fv, 2, J, c) + (c % 3 == 2 ? rho[c / 3] : 0.f);
				const float a = sum3(dr_dwl.x * w0, dr_dwl.y * w1, dr_dwl.z * w2);
				const float b = sum3(dr_dnl.x * col_mix(fn, 0, J, c), dr_dnl.y * col_mix(fn, 1, J, c), dr_dnl.z * col_mix(fn, 2, J, c));
				dr_dV[c] = a + b;
				dr_dN[c] = (c % 3 == 0 ? dr_dnl.x : c % 3 == 1 ? dr_dnl.y : dr_dnl.z) * rho[c / 3];
			}
		}
		cp_async_wait_all();
		if (contributes) {
			// turn every (face vertex, anchor slot) record into its 1 x 6 Jacobian contribution, in place
#pragma unroll
			for (int fvi = 0; fvi < 3; fvi++) {
				const float3 dv = make_float3(dr_dV[fvi * 3], dr_dV[fvi * 3 + 1], dr_dV[fvi * 3 + 2]);
				const float3 dn = make_float3(dr_dN[fvi * 3], dr_dN[fvi * 3 + 1], dr_dN[fvi * 3 + 2]);
#pragma unroll
				for (int slot = 0; slot < A; slot++) {
					float4* r = s_rec + ((fvi * A + slot) * 2) * T + tid;
					const float4 rec0 = r[0], rec1 = r[T];
					const float3 a = row_times_skew(dv, make_float3(rec0.x, rec0.y, rec0.z));
					const float3 b = row_times_skew(dn, make_float3(rec1.x, rec1.y, rec1.z));
					r[0] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, dv.x * rec0.w);
					r[T] = make_float4(dv.y * rec0.w, dv.z * rec0.w, 0.f, rec1.w);
					const int node = __float_as_int(rec1.w);
					nodes[fvi * A + slot] = node;
					if (node >= 0) pending |= 1u << (fvi * A + slot);
				}
			}
		}
	}
	if (in_image) {
		d.diag_pixel_faces[px] = face;
		d.diag_residuals[px] = residual;
		d.diag_mask[px] = residual_mask ? 1 : 0;
	}
	// node voting: the leader's first pending node is taken up by every lane that has it among its pairs; a node appears at most
	// once per face vertex, and the pairs are summed in face-vertex order like the reference's per-pixel accumulation
	while (true) {
		const unsigned active = __ballot_sync(0xFFFFFFFFu, pending != 0u);
		if (active == 0u) break;
		const int leader = __ffs(active) - 1;
		int my_first_node = -1;
		if (pending) my_first_node = __float_as_int(s_rec[((__ffs(pending) - 1) * 2 + 1) * T + tid].w);
		const int node = __shfl_sync(0xFFFFFFFFu, my_first_node, leader);
		unsigned match = 0u;
#pragma unroll
		for (int q = 0; q < kMaxEntries; q++) match |= (nodes[q] == node ? 1u : 0u) << q;
		match &= pending;
		pending &= ~match;
		float vals[32];
#pragma unroll
		for (int i = 0; i < 32; i++) vals[i] = 0.f;
		if (match) {
			float3 Jr = make_float3(0.f, 0.f, 0.f), Jt = make_float3(0.f, 0.f, 0.f);
			for (unsigned m = match; m; m &= m - 1u) {
				const float4* r = s_rec + ((__ffs(m) - 1) * 2) * T + tid;
				const float4 c0 = r[0], c1 = r[T];
				Jr.x += c0.x; Jr.y += c0.y; Jr.z += c0.z;
				Jt.x += c0.w; Jt.y += c1.x; Jt.z += c1.y;
			}
			const float J6[6] = {Jr.x, Jr.y, Jr.z, Jt.x, Jt.y, Jt.z};
			int o = 0;
#pragma unroll
			for (int c0 = 0; c0 < 6; c0++)
#pragma unroll
				for (int c1 = c0; c1 < 6; c1++) vals[o++] = J6[c0] * J6[c1];
#pragma unroll
			for (int c = 0; c < 6; c++) vals[21 + c] = J6[c] * residual;
		}
		const float total = warp_transpose_reduce(vals, lane);
		if (lane < 27) atomicAdd(d.accumulators + (int64_t) node * kAccumStride + lane, total);
	}
	// the band queue was fully drained before this kernel started: reset its counter for the next iteration
	if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) d.counters[0] = 0;
}

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
	face_raster_kernel<<<blocks_for(d.face_count, 64), 64, 0, s>>>(d, p, cam);
}

void launch_band_raster(const FitterDeviceData& d, const DataTermParams& p, const NdcCamera& cam, cudaStream_t s) {
	band_raster_kernel<<<148 * 2, 256, 0, s>>>(d, p, cam);
}

template<int A>
static int launch_assemble_for(const FitterDeviceData& d, const DataTermParams& p, cudaStream_t s) {
	constexpr size_t bytes = assemble_shared_bytes(A);
	static bool attribute_set = false;
	if (bytes > 48 * 1024 && !attribute_set) {
		if (cudaFuncSetAttribute(assemble_kernel<A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes) != cudaSuccess) return kCudaError;
		attribute_set = true;
	}
	const dim3 grid((p.width + 15) / 16, (p.row_end - p.row_begin + 7) / 8);
	assemble_kernel<A><<<grid, kAssembleThreads, bytes, s>>>(d, p);
	return kOk;
}

int launch_assemble(const FitterDeviceData& d, const DataTermParams& p, cudaStream_t s) {
	switch (d.anchor_count) {
#define NNRT_CASE(A_) case A_: return launch_assemble_for<A_>(d, p, s);
		NNRT_CASE(1) NNRT_CASE(2) NNRT_CASE(3) NNRT_CASE(4) NNRT_CASE(5) NNRT_CASE(6) NNRT_CASE(7) NNRT_CASE(8)
#undef NNRT_CASE
		default: return kInvalidArgument;
	}
	return kOk;
}

void launch_fill_keys(unsigned long long* keys, int64_t n, cudaStream_t s) {
	if (n > 0) fill_keys_kernel2<<<blocks_for(n, 256), 256, 0, s>>>(keys, n);
}

}  // namespace nnrt_b200
