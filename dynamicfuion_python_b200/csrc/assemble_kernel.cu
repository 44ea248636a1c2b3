// K6 of the dense-depth fitter: residual + Jacobian chain + per-node J^T J / J^T r assembly.
// Nothing in this file decides an index (the pixel -> face assignment comes from the raster kernels), but its sums are compared
// with the oracle at 1e-4 relative and contain heavy cancellation, so it is compiled like the rest with -fmad=false and IEEE
// division (see build.py for the measured alternative); the residual chain is written with rounding intrinsics regardless.
#include "fitter_kernels.cuh"
#include "fixed_accum.cuh"
#include "jacobian_math.cuh"

namespace nnrt_b200 {

// Transposed warp reduction: every lane passes 32 values, lane l receives the warp-wide sum of value l (31 shuffles).
// (Measured alternative, round 2: the same per-node Gram matrix M^T M as four k-steps of mma.sync m16n8k8 with a TF32 head / tail split
// of every operand, fragments exchanged through 1 KB of shared memory per warp. It executes a third of the instructions of this
// tree but runs 5.6 us SLOWER per launch (44.9 vs 39.3 us): on sm_100a the legacy mma.sync path has a long issue-to-result latency
// and the kernel is bound by the latency of each warp's dependent chain, not by instruction issue.)
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

// row-vector times skew(s):  r * [[0,-sz,sy],[sz,0,-sx],[-sy,sx,0]]. The products with the matrix's zeros are left out: they add
// +-0 to a sum, which changes nothing but the sign of an exact zero.
__device__ __forceinline__ float3 row_times_skew(float3 r, float3 s) {
	return make_float3(r.y * s.z + r.z * -s.y, r.x * -s.z + r.z * s.x, r.x * s.y + r.y * -s.x);
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
// kSharded = false is the single-GPU kernel (no stripe arithmetic, one accumulator buffer); true adds the striped tile-row mapping and
// the parity-selected accumulator buffer of a pixel-sharded fit.
template<int A, bool kSharded>
__global__ void __launch_bounds__(kAssembleThreads) assemble_kernel(FitterDeviceData d, DataTermParams p) {
	constexpr int kMaxEntries = 3 * A, T = kAssembleThreads;
	extern __shared__ __align__(16) unsigned char assemble_smem[];
	float4* s_rec = reinterpret_cast<float4*>(assemble_smem);  // [3A * 2][T]: anchor records, then J contributions (+ node id)
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	// tiles are visited in a scrambled order (multiplication by a prime modulo the tile count): the surface covers a compact
	// part of the image, and in raster order its expensive tiles would reach the SMs as one late clump
	const unsigned tile_count = gridDim.x * gridDim.y, linear = blockIdx.y * gridDim.x + blockIdx.x;
	const unsigned tile = (tile_count % 1009u) && tile_count < (1u << 22) ? (linear * 1009u) % tile_count : linear;  // 32-bit: no overflow below 2^22 tiles
	// tile row -> image row: tile rows are numbered over the stripes this GPU owns (fitter_kernels.cuh; 8-row stripes dealt to one GPU = every row)
	const int tile_row = (int) (tile / gridDim.x);
	const int tile_x = (int) (tile % gridDim.x) * 16 + (warp & 1) * 8;
	int tile_y = p.row_begin + tile_row * 8 + (warp >> 1) * 4;
	if (kSharded && p.stripe_count > 1) {  // (kept off the unsharded path: three of four tiles of a frame are empty and cost little more than this prologue)
		const int tiles_per_stripe = p.stripe_rows >> 3;
		const int stripe = p.stripe_index + p.stripe_count * (tile_row / tiles_per_stripe);
		tile_y = p.row_begin + stripe * p.stripe_rows + (tile_row % tiles_per_stripe) * 8 + (warp >> 1) * 4;
	}
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
		float inverse_depth[3];
		{
			const float4 a = d.warped_positions[tri.x], b = d.warped_positions[tri.y], c = d.warped_positions[tri.z];
			inverse_depth[0] = a.w; inverse_depth[1] = b.w; inverse_depth[2] = c.w;
			fv[0] = make_float3(a.x, a.y, a.z); fv[1] = make_float3(b.x, b.y, b.z); fv[2] = make_float3(c.x, c.y, c.z);
			const float4 na = d.warped_normals[tri.x], nb = d.warped_normals[tri.y], nc = d.warped_normals[tri.z];
			fn[0] = make_float3(na.x, na.y, na.z); fn[1] = make_float3(nb.x, nb.y, nb.z); fn[2] = make_float3(nc.x, nc.y, nc.z);
		}
		const float4 ref = d.reference_points[px];
		const float2 ray = make_float2(pixel_centre(p.ndc_x, u, p.width, p.height), pixel_centre(p.ndc_y, v, p.height, p.width));
		FaceNdc fndc;
		fndc.v0 = project_with_inverse_depth(p.ndc, fv[0].x, fv[0].y, inverse_depth[0]);
		fndc.v1 = project_with_inverse_depth(p.ndc, fv[1].x, fv[1].y, inverse_depth[1]);
		fndc.v2 = project_with_inverse_depth(p.ndc, fv[2].x, fv[2].y, inverse_depth[2]);
		fndc.z0 = fv[0].z; fndc.z1 = fv[1].z; fndc.z2 = fv[2].z;
		Fragment frag;
		shade_pixel(fndc, ray, p.blur_radius, p.perspective_correct != 0, false, true, frag, true);  // the raster pass accepted this pixel
		const float rho[3] = {frag.bary.x, frag.bary.y, frag.bary.z};
		const float depth = frag.depth;
		// rendered normal (InterpolateFaceAttributesImpl.h:56-62: acc = 0; acc += w*attr)
		// (the residual is a small difference of large numbers: its chain keeps the reference's unfused, correctly rounded operations
		// through intrinsics, whatever the compile flags)
		auto blend = [&](float a, float b, float c) {
			return __fadd_rn(__fadd_rn(__fadd_rn(0.0f, __fmul_rn(rho[0], a)), __fmul_rn(rho[1], b)), __fmul_rn(rho[2], c));
		};
		const float3 nl = make_float3(blend(fn[0].x, fn[1].x, fn[2].x), blend(fn[0].y, fn[1].y, fn[2].y), blend(fn[0].z, fn[1].z, fn[2].z));
		// unprojection of the rendered depth at the integer pixel with the pinhole intrinsics (PerspectiveProjectionImpl.h:119-131)
		const bool rendered_valid = depth > 0 && depth < p.max_depth;
		float3 wl = make_float3(0.f, 0.f, 0.f);
		if (rendered_valid)
			wl = make_float3(__fmul_rn(__fsub_rn((float) u, p.pinhole.cx), depth) / SharedDivisor{p.pinhole.fx, p.pinhole_inverse_fx},
			                 __fmul_rn(__fsub_rn((float) v, p.pinhole.cy), depth) / SharedDivisor{p.pinhole.fy, p.pinhole_inverse_fy}, depth);
		const bool mask = rendered_valid && ref.w != 0.f;
		const float3 pm = make_float3(__fsub_rn(wl.x, ref.x), __fsub_rn(wl.y, ref.y), __fsub_rn(wl.z, ref.z));
		const float r_raw = __fadd_rn(__fmul_rn(nl.x, pm.x), __fadd_rn(__fmul_rn(nl.y, pm.y), __fmul_rn(nl.z, pm.z)));  // dot3 = x + (y + z)
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
			const float2 projected[3] = {fndc.v0, fndc.v1, fndc.v2};
			barycentric_jacobian(J, ray, fv, projected, inverse_depth, p.ndc, p.perspective_correct != 0, rho);
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
	if (in_image && p.write_diagnostics) {
		d.diag_pixel_faces[px] = face;
		d.diag_residuals[px] = residual;
		d.diag_mask[px] = residual_mask ? 1 : 0;
	}
	// node voting: the leader's first pending node is taken up by every lane that has it among its pairs; a node appears at most
	// once per face vertex, and the pairs are summed in face-vertex order like the reference's per-pixel accumulation
	// peer-reduced pixel shard (fitter_kernels.cuh::PeerReduce): two buffers alternate by iteration parity
	unsigned long long* accumulators = d.accumulators;
	if (kSharded && d.accum_epoch) accumulators += (long long) (*d.accum_epoch & 1ull) * d.accum_parity_words;
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
		// integer (fixed-point) atomics: the per-node sums do not depend on the order in which warps arrive (fixed_accum.cuh)
		if (lane < 27) fixed_accum_add(accumulators + ((int64_t) node * kAccumStride + lane) * 2, total, d.counters + 3);
	}
	// the band queue was fully drained before this kernel started: reset its counter for the next iteration
	if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) d.counters[0] = 0;
}

template<int A>
static int launch_assemble_for(const FitterDeviceData& d, const DataTermParams& p, cudaStream_t s) {
	constexpr size_t bytes = assemble_shared_bytes(A);
	static SharedMemoryOptIn opt_in;
	static SharedMemoryOptIn opt_in_sharded;
	const bool sharded = p.stripe_count > 1 || d.accum_epoch != nullptr;
	if (bytes > 48 * 1024 && (sharded ? opt_in_sharded.ensure(assemble_kernel<A, true>, bytes) : opt_in.ensure(assemble_kernel<A, false>, bytes)) != kOk) return kCudaError;
	const int stripes = (p.row_end - p.row_begin + p.stripe_rows - 1) / p.stripe_rows;
	const int owned = p.stripe_index < stripes ? (stripes - p.stripe_index + p.stripe_count - 1) / p.stripe_count : 0;
	const dim3 grid((p.width + 15) / 16, max(1, owned * (p.stripe_rows / 8)));  // a GPU that owns nothing still resets the band queue counter
	if (sharded) assemble_kernel<A, true><<<grid, kAssembleThreads, bytes, s>>>(d, p);
	else assemble_kernel<A, false><<<grid, kAssembleThreads, bytes, s>>>(d, p);
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


}  // namespace nnrt_b200
