// Non-rigid TSDF integration: NonRigidSurfaceVoxelBlockGrid::IntegrateNonRigid's voxel kernel
// (geometry/kernel/NonRigidSurfaceVoxelBlockGridImpl.h:50-229) over caller-owned voxel-block arrays — the block hash map is Open3D's
// VoxelBlockGrid, only its key and value tensors cross the boundary.
//
// One thread per voxel of the listed blocks: voxel centre -> camera space, K nearest warp nodes (brute force over shared-memory
// tiles; the reference walks a kd-tree, whose traversal order decides only the slot order of the same set), distance-thresholded
// weights, linear-blend warp, projection, depth / normal lookup, truncated signed distance and running-average update of tsdf and
// colour. Quirks kept: a dropped anchor leaves its distance in the weight slot, the voxel weight is read but never advanced.
// `cos_voxel_ray_to_normal` is written by several voxels per pixel in the reference (last writer wins, scheduling-dependent); here
// the winner is the voxel with the highest workload index (64-bit atomicMax of index << 32 | value bits), i.e. the reference's
// single-thread result.
#include "host_util.h"
#include "../../include/nnrt_b200.h"
#include <cfloat>

namespace nnrt_b200 {
namespace {

constexpr int kThreads = 128;
constexpr int kNodeTile = 1024;

// the warp field as the kernels see it
struct NodeField {
	const float *nodes, *rotations, *translations; int node_count; const float* node_weights; float fixed_c2; int min_valid;
};

struct IntegrateParams {
	const int64_t* block_indices; const int32_t* block_keys; float* tsdf; const void* weight; int weight_dtype; float* voxel_color;
	int block_resolution; float voxel_size, truncation;
	const void* depth; int depth_dtype; const void* color; int color_dtype; const float* normals;
	int height, width, color_height, color_width;
	Intrinsics depth_k, color_k; Rigid extrinsics;
	NodeField field; float depth_scale, depth_max;
	unsigned long long* cos_keys; int64_t voxel_count;
};

__device__ __forceinline__ bool in_boundary(float x, float y, int rows, int cols) {  // Open3D ArrayIndexer::InBoundary
	return y >= 0 && x >= 0 && y <= rows - 1.0f && x <= cols - 1.0f;
}

// K nearest nodes by DISTANCE (norm), replace-current-maximum in the order the candidates are offered (core/kernel/KnnUtilities.h:144-218
// without the kd-tree). Candidates are screened on the squared distance: m * m * (1 + 1e-6) exceeds the exact m^2 whatever the
// roundings do; above it the correctly rounded root cannot be below m, so only the few candidates near the current k-th distance pay
// for the IEEE square root the comparison is defined on.
template<int A>
struct NearestNodes {
	int idx[A];
	float w[A], c2s[A];
	int max_at;
	float max_distance, skip_above;
	__device__ __forceinline__ void init(float default_c2) {
#pragma unroll
		for (int a = 0; a < A; a++) { idx[a] = -1; w[a] = INFINITY; c2s[a] = default_c2; }
		max_at = 0;
		max_distance = INFINITY;
		skip_above = INFINITY;
	}
	__device__ __forceinline__ void offer(float4 g, int index, float3 point) {
		const float dx = g.x - point.x, dy = g.y - point.y, dz = g.z - point.z;
		const float d2 = sum3(dx * dx, dy * dy, dz * dz);
		if (d2 > skip_above) return;
		const float distance = sqrtf(d2);
		if (max_distance > distance) {
#pragma unroll
			for (int a = 0; a < A; a++)
				if (a == max_at) { w[a] = distance; idx[a] = index; c2s[a] = g.w; }
			max_at = 0;
			max_distance = w[0];
#pragma unroll
			for (int a = 1; a < A; a++)
				if (w[a] > max_distance) { max_at = a; max_distance = w[a]; }
			skip_above = max_distance * max_distance * 1.000001f;
		}
	}
	// thresholded weights (geometry/functional/kernel/WarpUtilities.h:95-153, distances not squared), normalisation (:36-46) and blend
	// warp (:428-445). Returns false when fewer than min_valid anchors survive; with `warp_anyway` the point is then still blended with
	// the raw, un-normalised weights of the surviving anchors (what a caller that ignores ComputeAnchorsForPoint's result gets).
	__device__ __forceinline__ bool finish(const NodeField& p, float3 point, bool warp_anyway, float3& warped_out) {
		float weight_sum = 0.0f;
		int valid = 0;
#pragma unroll
		for (int a = 0; a < A; a++) {
			const float square_distance = w[a] * w[a];
			if (square_distance > 4 * c2s[a]) { idx[a] = -1; continue; }
			const float anchor_weight = expf(-square_distance / (2 * c2s[a]));
			weight_sum += anchor_weight;
			w[a] = anchor_weight;
			valid++;
		}
		const bool enough = valid >= p.min_valid;
		warped_out = make_float3(0.f, 0.f, 0.f);
		if (!enough && !warp_anyway) return false;
		if (!enough) {
			// weights stay raw
		} else if (weight_sum > 0.0f) {
#pragma unroll
			for (int a = 0; a < A; a++) w[a] = __fdiv_rn(w[a], weight_sum);
		} else if (valid > 0) {
#pragma unroll
			for (int a = 0; a < A; a++) w[a] = 1.0f / (float) valid;
		}
		float3 warped = make_float3(0.f, 0.f, 0.f);
#pragma unroll
		for (int a = 0; a < A; a++) {
			const int n = idx[a];
			if (n == -1) continue;
			const float3 g = make_float3(p.nodes[n * 3], p.nodes[n * 3 + 1], p.nodes[n * 3 + 2]);
			const float3 t = make_float3(p.translations[n * 3], p.translations[n * 3 + 1], p.translations[n * 3 + 2]);
			warped = warped + w[a] * ((g + matvec3(p.rotations + (int64_t) n * 9, point - g)) + t);
		}
		warped_out = warped;
		return enough;
	}
};

// WarpField::ComputeAnchorsForPoint<threshold = true> + WarpPoint (geometry/WarpField.h:77-122) for one point per thread over ALL nodes,
// staged through shared memory in tiles; every thread of the CTA must call it.
template<int A>
__device__ __forceinline__ bool warp_thresholded(const NodeField& p, float3 point, bool active, float4* tile, bool warp_anyway, float3& warped_out) {
	warped_out = make_float3(0.f, 0.f, 0.f);
	NearestNodes<A> nearest;
	nearest.init(p.node_weights ? p.node_weights[0] : p.fixed_c2);
	for (int base = 0; base < p.node_count; base += kNodeTile) {
		const int count = min(kNodeTile, p.node_count - base);
		__syncthreads();
		for (int k = threadIdx.x; k < count; k += blockDim.x) {
			const int n = base + k;
			tile[k] = make_float4(p.nodes[n * 3], p.nodes[n * 3 + 1], p.nodes[n * 3 + 2], p.node_weights ? p.node_weights[n] : p.fixed_c2);
		}
		__syncthreads();
		if (!active) continue;
		for (int k = 0; k < count; k++) nearest.offer(tile[k], base + k, point);
	}
	if (!active) return false;
	return nearest.finish(p, point, warp_anyway, warped_out);
}

// everything after the warp: projection, depth / normal lookup, gates, running averages (NonRigidSurfaceVoxelBlockGridImpl.h:160-226)
__device__ __forceinline__ void integrate_warped_voxel(const IntegrateParams& p, int64_t workload, int64_t block_index, int64_t voxel_in_block, int64_t r3,
                                                       float3 warped) {
	if (warped.z < 0) return;
	const float2 uv = project(p.depth_k, warped);
	if (!in_boundary(uv.x, uv.y, p.height, p.width)) return;
	const int64_t u = (int64_t) roundf(uv.x), v = (int64_t) roundf(uv.y);
	const int64_t pixel = v * p.width + u;
	float depth = p.depth_dtype == 0 ? ((const float*) p.depth)[pixel] : (float) ((const uint16_t*) p.depth)[pixel];
	depth = __fdiv_rn(depth, p.depth_scale);
	if (depth <= 0.0f || depth > p.depth_max) return;
	const float psdf = depth - warped.z;
	float3 view = make_float3(-warped.x, -warped.y, -warped.z);
	const float z = sum3(view.x * view.x, view.y * view.y, view.z * view.z);
	if (z > 0.f) {
		const float norm = sqrtf(z);
		view = make_float3(__fdiv_rn(view.x, norm), __fdiv_rn(view.y, norm), __fdiv_rn(view.z, norm));
	}
	const float cosine = sum3(view.x * p.normals[pixel * 3], view.y * p.normals[pixel * 3 + 1], view.z * p.normals[pixel * 3 + 2]);
	atomicMax(p.cos_keys + pixel, ((unsigned long long) (workload + 1) << 32) | __float_as_uint(cosine));
	if (psdf <= -p.truncation || cosine > 0.5f) return;
	const int64_t linear = block_index * r3 + voxel_in_block;
	const float tsdf_normalized = __fdiv_rn(psdf < p.truncation ? psdf : p.truncation, p.truncation);
	const float voxel_weight = p.weight_dtype == 0 ? ((const float*) p.weight)[linear] : (float) ((const uint16_t*) p.weight)[linear];
	const float inverted_weight_sum = __fdiv_rn(1.0f, voxel_weight + 1);
	p.tsdf[linear] = (voxel_weight * p.tsdf[linear] + tsdf_normalized) * inverted_weight_sum;
	if (p.color && p.voxel_color) {
		// Unproject(u, v, 1) with the depth intrinsics, Project with the colour intrinsics (identity pose)
		const float3 ray = make_float3(__fdiv_rn(((float) u - p.depth_k.cx) * 1.0f, p.depth_k.fx), __fdiv_rn(((float) v - p.depth_k.cy) * 1.0f, p.depth_k.fy), 1.0f);
		const float2 cuv = project(p.color_k, ray);
		if (in_boundary(cuv.x, cuv.y, p.color_height, p.color_width)) {
			const int64_t cpx = (int64_t) roundf(cuv.y) * p.color_width + (int64_t) roundf(cuv.x);
			const float multiplier = p.color_dtype == 0 ? 255.0f : 1.0f;
#pragma unroll
			for (int c = 0; c < 3; c++) {
				const float in = p.color_dtype == 0 ? ((const float*) p.color)[cpx * 3 + c] : (float) ((const uint8_t*) p.color)[cpx * 3 + c];
				p.voxel_color[linear * 3 + c] = (voxel_weight * p.voxel_color[linear * 3 + c] + in * multiplier) * inverted_weight_sum;
			}
		}
	}
}

__device__ __forceinline__ float3 voxel_camera_position(const IntegrateParams& p, int64_t block_index, int64_t voxel_in_block) {
	const int r = p.block_resolution;
	const int32_t* key = p.block_keys + block_index * 3;
	const int xl = (int) (voxel_in_block % r), yl = (int) ((voxel_in_block / r) % r), zl = (int) (voxel_in_block / ((int64_t) r * r));
	return rigid_transform(p.extrinsics, make_float3((float) (key[0] * r + xl) * p.voxel_size, (float) (key[1] * r + yl) * p.voxel_size,
	                                                 (float) (key[2] * r + zl) * p.voxel_size));
}

// Rigid integration: Open3D's t::geometry::VoxelBlockGrid::Integrate (third party, Open3D v0.17, not in the reference checkout; the
// reference re-exports it as nnrt.geometry.VoxelBlockGrid.integrate, pybind/geometry/geometry.cpp:170-203, and its fusion pipeline
// integrates the first frame with it). Restated from the published kernel (VoxelBlockGridImpl.h, IntegrateCPU/CUDA): voxel -> camera
// -> pixel (rounded); skip if depth is invalid, the voxel is behind the camera or more than the truncation behind the surface;
// running average of the clamped, normalised projective distance (and of the colour sampled through the colour camera) with the
// voxel weight, which is then raised by one. Parity unpinned by the reference. One thread per voxel of the listed blocks.
__global__ void __launch_bounds__(kThreads) integrate_rigid_kernel(IntegrateParams p) {
	const int64_t workload = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (workload >= p.voxel_count) return;
	const int64_t r3 = (int64_t) p.block_resolution * p.block_resolution * p.block_resolution;
	const int64_t block_index = p.block_indices[workload / r3], voxel_in_block = workload % r3;
	const float3 voxel = voxel_camera_position(p, block_index, voxel_in_block);
	const float2 uv = project(p.depth_k, voxel);
	if (!in_boundary(uv.x, uv.y, p.height, p.width)) return;
	const int64_t pixel = (int64_t) roundf(uv.y) * p.width + (int64_t) roundf(uv.x);
	float depth = p.depth_dtype == 0 ? ((const float*) p.depth)[pixel] : (float) ((const uint16_t*) p.depth)[pixel];
	depth = __fdiv_rn(depth, p.depth_scale);
	float sdf = depth - voxel.z;
	if (depth <= 0.0f || depth > p.depth_max || voxel.z <= 0.0f || sdf < -p.truncation) return;
	sdf = __fdiv_rn(sdf < p.truncation ? sdf : p.truncation, p.truncation);
	const int64_t linear = block_index * r3 + voxel_in_block;
	const float voxel_weight = p.weight_dtype == 0 ? ((const float*) p.weight)[linear] : (float) ((const uint16_t*) p.weight)[linear];
	const float inverted_weight_sum = __fdiv_rn(1.0f, voxel_weight + 1);
	p.tsdf[linear] = (voxel_weight * p.tsdf[linear] + sdf) * inverted_weight_sum;
	if (p.color && p.voxel_color) {
		const float2 cuv = project(p.color_k, voxel);
		if (in_boundary(cuv.x, cuv.y, p.color_height, p.color_width)) {
			const int64_t cpx = (int64_t) roundf(cuv.y) * p.color_width + (int64_t) roundf(cuv.x);
			const float multiplier = p.color_dtype == 0 ? 255.0f : 1.0f;
#pragma unroll
			for (int c = 0; c < 3; c++) {
				const float in = p.color_dtype == 0 ? ((const float*) p.color)[cpx * 3 + c] : (float) ((const uint8_t*) p.color)[cpx * 3 + c];
				p.voxel_color[linear * 3 + c] = (voxel_weight * p.voxel_color[linear * 3 + c] + in * multiplier) * inverted_weight_sum;
			}
		}
	}
	if (p.weight_dtype == 0) ((float*) const_cast<void*>(p.weight))[linear] = voxel_weight + 1;
	else ((uint16_t*) const_cast<void*>(p.weight))[linear] = (uint16_t) (voxel_weight + 1);
}

// Full-scan form: one thread per voxel, every node offered to every voxel. Kept as the yardstick the shortlist kernel below must equal
// bit for bit (nnrt_b200_integrate_non_rigid_full_scan).
template<int A>
__global__ void __launch_bounds__(kThreads) integrate_non_rigid_kernel(IntegrateParams p) {
	__shared__ float4 tile[kNodeTile];
	const int64_t workload = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	const bool active = workload < p.voxel_count;
	const int64_t r3 = (int64_t) p.block_resolution * p.block_resolution * p.block_resolution;
	int64_t block_index = 0, voxel_in_block = 0;
	float3 voxel = make_float3(0.f, 0.f, 0.f);
	if (active) {
		block_index = p.block_indices[workload / r3];
		voxel_in_block = workload % r3;
		voxel = voxel_camera_position(p, block_index, voxel_in_block);
	}
	float3 warped;
	const bool warp_ok = warp_thresholded<A>(p.field, voxel, active, tile, false, warped);
	if (!active || !warp_ok) return;
	integrate_warped_voxel(p, workload, block_index, voxel_in_block, r3, warped);
}

// Shortlist form: one CTA per voxel block. A node can enter a voxel's anchor slots only while it is closer to that voxel than the
// voxel's current A-th distance. With c the block centre and r the half diagonal, every voxel v of the block has
// |n - v| >= |n - c| - r and (A-th distance of v among the nodes before n) <= U + r, where U is the A-th smallest distance to c among
// nodes before n. So a node with |n - c| >= U + 2r is never inserted for any voxel of the block and may be left out of the scan without
// changing a single slot — provided U is the RUNNING bound over the node order (a node that no voxel keeps in the end can still sit in a
// slot for a while and decide where later nodes land). Warp 0 walks the nodes in chunks of 32 in index order, tests each chunk against
// the bound of the chunks before it (a superset of the nodes before n gives a smaller U, so the chunk-level bound is only more
// conservative), appends the survivors to the shortlist in order and folds the chunk into the running A smallest distances. The voxels
// then scan the shortlist instead of all N nodes: same slots, same weights, same sums (checked bit for bit against the full scan).
constexpr int kShortlist = kNodeTile;  // entries; a block whose shortlist overflows falls back to the full scan
constexpr int kBlockThreads = 256;

template<int A>
__global__ void __launch_bounds__(kBlockThreads) integrate_blocks_kernel(IntegrateParams p) {
	__shared__ float4 s_node[kShortlist];  // doubles as the tile buffer of the fallback
	__shared__ int s_index[kShortlist];
	__shared__ int s_count;
	const int r = p.block_resolution;
	const int64_t r3 = (int64_t) r * r * r;
	const int64_t block_index = p.block_indices[blockIdx.x];
	const NodeField& f = p.field;
	if (threadIdx.x < 32) {
		const int lane = threadIdx.x;
		const int32_t* key = p.block_keys + block_index * 3;
		const float half = 0.5f * (float) (r - 1);
		const float3 centre = rigid_transform(p.extrinsics, make_float3(((float) (key[0] * r) + half) * p.voxel_size, ((float) (key[1] * r) + half) * p.voxel_size,
		                                                              ((float) (key[2] * r) + half) * p.voxel_size));
		// half diagonal of the cube of voxel centres, padded for the fp32 rounding of voxel and centre positions
		const float reach = 2.f * (half * p.voxel_size * 1.7320509f * 1.0001f + 1e-6f);
		float smallest[A];  // running A smallest distances to the centre, ascending, identical in every lane
#pragma unroll
		for (int a = 0; a < A; a++) smallest[a] = INFINITY;
		int count = 0;
		for (int base = 0; base < f.node_count; base += 32) {
			const int n = base + lane;
			float distance = INFINITY;
			float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
			if (n < f.node_count) {
				g = make_float4(f.nodes[n * 3], f.nodes[n * 3 + 1], f.nodes[n * 3 + 2], f.node_weights ? f.node_weights[n] : f.fixed_c2);
				const float dx = g.x - centre.x, dy = g.y - centre.y, dz = g.z - centre.z;
				distance = sqrtf(sum3(dx * dx, dy * dy, dz * dz));
			}
			const float bound = (smallest[A - 1] + reach) * 1.00001f;  // inf while fewer than A nodes have been seen
			const bool keep = n < f.node_count && distance < bound;
			const unsigned kept = __ballot_sync(0xFFFFFFFFu, keep);
			const int slot = count + __popc(kept & ((1u << lane) - 1u));
			if (keep && slot < kShortlist) { s_node[slot] = g; s_index[slot] = n; }
			count += __popc(kept);
			if (count > kShortlist) break;  // uniform: falls back
			// fold the chunk into the running A smallest: repeated warp minimum over the chunk, merged into the sorted list
			float mine = distance;
			while (__any_sync(0xFFFFFFFFu, mine < smallest[A - 1])) {
				float best = mine;
#pragma unroll
				for (int offset = 16; offset > 0; offset >>= 1) best = fminf(best, __shfl_xor_sync(0xFFFFFFFFu, best, offset));
				const unsigned owners = __ballot_sync(0xFFFFFFFFu, mine == best);
				if (lane == __ffs(owners) - 1) mine = INFINITY;  // one lane gives up its value per round (duplicates stay for the next)
				float carry = best;  // insertion into the ascending list
#pragma unroll
				for (int a = 0; a < A; a++) {
					const float lower = fminf(smallest[a], carry);
					carry = fmaxf(smallest[a], carry);
					smallest[a] = lower;
				}
			}
		}
		if (lane == 0) s_count = count;
	}
	__syncthreads();
	const int count = s_count;
	const bool fallback = count > kShortlist;
	for (int64_t base = 0; base < r3; base += kBlockThreads) {
		const int64_t voxel_in_block = base + threadIdx.x;
		const bool active = voxel_in_block < r3;
		const float3 voxel = active ? voxel_camera_position(p, block_index, voxel_in_block) : make_float3(0.f, 0.f, 0.f);
		float3 warped;
		bool warp_ok;
		if (fallback) {
			warp_ok = warp_thresholded<A>(f, voxel, active, s_node, false, warped);
		} else {
			NearestNodes<A> nearest;
			nearest.init(f.node_weights ? f.node_weights[0] : f.fixed_c2);
			warp_ok = false;
			if (active) {
				for (int k = 0; k < count; k++) nearest.offer(s_node[k], s_index[k], voxel);
				warp_ok = nearest.finish(f, voxel, false, warped);
			}
		}
		if (active && warp_ok) integrate_warped_voxel(p, (int64_t) blockIdx.x * r3 + voxel_in_block, block_index, voxel_in_block, r3, warped);
	}
}

// GetBoundingBoxesOfWarpedBlocks (geometry/kernel/NonRigidSurfaceVoxelBlockGridImpl.h:289-363): axis-aligned box of the 8 warped
// corners of every block, 8 lanes per block (one corner each; the corner warps are the expensive part). Kept as the reference has them:
// the corner is `block key + block side length` (key in blocks, side in metres), anchors always use the FIXED coverage, a corner with
// too few valid anchors is blended with its raw weights (the reference ignores ComputeAnchorsForPoint's result), and the bounds are
// updated corner by corner with `if (min > x) min = x; else if (max < x) max = x;` (so max can stay -FLT_MAX).
template<int A>
__global__ void __launch_bounds__(kThreads) warped_block_boxes_kernel(const int32_t* __restrict__ block_keys, int64_t block_count, NodeField field,
                                                                      float block_side_length, Rigid extrinsics, float* __restrict__ boxes) {
	__shared__ float4 tile[kNodeTile];
	const int64_t block = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 3;
	const int corner = threadIdx.x & 7;
	const bool active = block < block_count;
	float3 point = make_float3(0.f, 0.f, 0.f);
	if (active) {
		// corner order of the reference: 000, 001, 010, 100, 011, 101, 110, 111 as (x, y, z)
		const int bx = (0xE8 >> corner) & 1, by = (0xD4 >> corner) & 1, bz = (0xB2 >> corner) & 1;
		const float x0 = (float) block_keys[block * 3], y0 = (float) block_keys[block * 3 + 1], z0 = (float) block_keys[block * 3 + 2];
		point = rigid_transform(extrinsics, make_float3(bx ? x0 + block_side_length : x0, by ? y0 + block_side_length : y0, bz ? z0 + block_side_length : z0));
	}
	float3 warped;
	warp_thresholded<A>(field, point, active, tile, true, warped);
	// the 8 lanes of a block stay together (active is uniform over them): lane 0 of the group replays the sequential update
	float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
	const unsigned group_mask = 0xFFu << (threadIdx.x & 24);
#pragma unroll
	for (int c = 0; c < 8; c++) {
		const int src = (threadIdx.x & 24) + c;
		const float v[3] = {__shfl_sync(group_mask, warped.x, src), __shfl_sync(group_mask, warped.y, src), __shfl_sync(group_mask, warped.z, src)};
#pragma unroll
		for (int k = 0; k < 3; k++) {
			if (lo[k] > v[k]) lo[k] = v[k];
			else if (hi[k] < v[k]) hi[k] = v[k];
		}
	}
	if (active && corner == 0) {
#pragma unroll
		for (int k = 0; k < 3; k++) { boxes[block * 6 + k] = lo[k]; boxes[block * 6 + 3 + k] = hi[k]; }
	}
}

// GetAxisAlignedBoxesInterceptingSurfaceMask (NonRigidSurfaceVoxelBlockGridImpl.h:365-435; geometry/kernel/Segment.h:27-73): mask[b] =
// some strided pixel's truncation segment (its ray between depth -+ truncation) intersects box b (slab test, as written there).
// The reference materialises all segments and tests every (segment, box) pair. Here a warp takes a box and visits only the strided
// pixels inside the box's projected rectangle (+-1 pixel): a segment lies on its pixel's ray, and a ray outside that rectangle cannot
// meet a box in front of the camera. Boxes that reach behind the camera or are inverted (the quirk above) fall back to every pixel.
// Stops at the first hit. Identical masks except where the slab test meets 0 * inf (a ray exactly on a box face through the optical axis).
__device__ __forceinline__ bool segment_intersects_box(float3 origin, float3 end, const float* __restrict__ box) {
	const float3 v = end - origin;
	const float3 inv = make_float3(__fdiv_rn(1.f, v.x), __fdiv_rn(1.f, v.y), __fdiv_rn(1.f, v.z));
	const int sx = inv.x < 0, sy = inv.y < 0, sz = inv.z < 0;
	float t_min = (box[sx ? 3 : 0] - origin.x) * inv.x, t_max = (box[sx ? 0 : 3] - origin.x) * inv.x;
	const float ty_min = (box[sy ? 4 : 1] - origin.y) * inv.y, ty_max = (box[sy ? 1 : 4] - origin.y) * inv.y;
	if ((t_min > ty_max) || (ty_min > t_max)) return false;
	if (ty_min > t_min) t_min = ty_min;
	if (ty_max < t_max) t_max = ty_max;
	const float tz_min = (box[sz ? 5 : 2] - origin.z) * inv.z, tz_max = (box[sz ? 2 : 5] - origin.z) * inv.z;
	if ((t_min > tz_max) || (tz_min > t_max)) return false;
	if (tz_min > t_min) t_min = tz_min;
	if (tz_max < t_max) t_max = tz_max;
	return !(t_max < 0.0f || t_min > 1.0f);
}

__global__ void __launch_bounds__(kThreads) boxes_intersecting_surface_kernel(const float* __restrict__ boxes, int64_t box_count, Intrinsics k, const void* __restrict__ depth,
                                                                              int depth_dtype, int height, int width, float depth_scale, float depth_max, int stride,
                                                                              float truncation, uint8_t* __restrict__ mask) {
	const int64_t b = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (b >= box_count) return;  // whole warps
	float box[6];
#pragma unroll
	for (int c = 0; c < 6; c++) box[c] = boxes[b * 6 + c];
	const int rows = height / stride, cols = width / stride;
	int row0 = 0, row1 = rows - 1, col0 = 0, col1 = cols - 1;
	if (box[2] > 0.f && box[0] <= box[3] && box[1] <= box[4] && box[2] <= box[5]) {
		// x / z and y / z over the box are extremal at its corners
		float x_lo = INFINITY, x_hi = -INFINITY, y_lo = INFINITY, y_hi = -INFINITY;
#pragma unroll
		for (int c = 0; c < 4; c++) {
			const float z = (c & 1) ? box[5] : box[2];
			const float x = (c & 2) ? box[3] : box[0], y = (c & 2) ? box[4] : box[1];
			x_lo = fminf(x_lo, x / z); x_hi = fmaxf(x_hi, x / z); y_lo = fminf(y_lo, y / z); y_hi = fmaxf(y_hi, y / z);
		}
		const float u_lo = fminf(k.fx * x_lo, k.fx * x_hi) + k.cx - 1.f, u_hi = fmaxf(k.fx * x_lo, k.fx * x_hi) + k.cx + 1.f;
		const float v_lo = fminf(k.fy * y_lo, k.fy * y_hi) + k.cy - 1.f, v_hi = fmaxf(k.fy * y_lo, k.fy * y_hi) + k.cy + 1.f;
		// strided pixel index i covers pixel i * stride
		col0 = max(0, (int) floorf(fmaxf(u_lo, 0.f) / (float) stride)); col1 = min(cols - 1, (int) fminf(ceilf(u_hi / (float) stride), (float) cols));
		row0 = max(0, (int) floorf(fmaxf(v_lo, 0.f) / (float) stride)); row1 = min(rows - 1, (int) fminf(ceilf(v_hi / (float) stride), (float) rows));
		if (u_hi < 0.f || v_hi < 0.f) { col1 = -1; row1 = -1; }
	}
	bool hit = false;
	const int span = col1 - col0 + 1;
	const int64_t candidates = (row1 >= row0 && span > 0) ? (int64_t) (row1 - row0 + 1) * span : 0;
	for (int64_t base = 0; base < candidates && !hit; base += 32) {
		const int64_t i = base + lane;
		bool mine = false;
		if (i < candidates) {
			const int v = (row0 + (int) (i / span)) * stride, u = (col0 + (int) (i % span)) * stride;
			const int64_t px = (int64_t) v * width + u;
			const float d = __fdiv_rn(depth_dtype == 0 ? ((const float*) depth)[px] : (float) ((const uint16_t*) depth)[px], depth_scale);
			if (d > 0 && d < depth_max) {
				const float d0 = d - truncation, d1 = d + truncation;
				// TransformIndexer::Unproject: x = (u - cx) * d / fx
				const float3 start = make_float3(__fdiv_rn(((float) u - k.cx) * d0, k.fx), __fdiv_rn(((float) v - k.cy) * d0, k.fy), d0);
				const float3 end = make_float3(__fdiv_rn(((float) u - k.cx) * d1, k.fx), __fdiv_rn(((float) v - k.cy) * d1, k.fy), d1);
				mine = segment_intersects_box(start, end, box);
			}
		}
		hit = __any_sync(0xFFFFFFFFu, mine);
	}
	if (lane == 0) mask[b] = hit ? 1 : 0;
}

__global__ void __launch_bounds__(256) resolve_cosines_kernel(const unsigned long long* __restrict__ keys, int64_t pixel_count, float* __restrict__ cosines) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < pixel_count) cosines[i] = keys[i] == 0ull ? 0.f : __uint_as_float((unsigned) (keys[i] & 0xffffffffull));
}

}  // namespace
}  // namespace nnrt_b200

using namespace nnrt_b200;

static int integrate_non_rigid_impl(bool full_scan,
		
		const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf, const void* weight, int32_t weight_dtype,
		float* voxel_color, int32_t block_resolution, float voxel_size, float sdf_truncation_distance, const void* depth, int32_t depth_dtype,
		const void* color, int32_t color_dtype, const float* depth_normals, int32_t image_height, int32_t image_width, int32_t color_height,
		int32_t color_width, const double* depth_intrinsics_host, const double* color_intrinsics_host, const double* extrinsics_host,
		const float* nodes, const float* node_rotations, const float* node_translations, int32_t node_count, const float* node_coverage_weights,
		float node_coverage, int32_t anchor_count, int32_t minimum_valid_anchor_count, float depth_scale, float depth_max,
		float* cos_voxel_ray_to_normal, void* stream) {
	NNRT_CHECK_ARG(block_keys && tsdf && weight && depth && depth_normals && depth_intrinsics_host && nodes && node_rotations && node_translations &&
	               cos_voxel_ray_to_normal, "integrate_non_rigid: null pointer argument");
	NNRT_CHECK_ARG(block_count == 0 || block_indices, "integrate_non_rigid: block_indices is NULL");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	NNRT_CHECK_ARG(block_resolution >= 1 && image_height > 0 && image_width > 0 && node_count > 0, "integrate_non_rigid: bad sizes");
	NNRT_CHECK_ARG((weight_dtype == 0 || weight_dtype == 1) && (depth_dtype == 0 || depth_dtype == 1) && (color_dtype == 0 || color_dtype == 1),
	               "dtype codes must be 0 or 1");
	NNRT_CHECK_ARG(!color || (color_intrinsics_host && color_height > 0 && color_width > 0), "colour image needs its intrinsics and size");
	const int64_t voxel_count = block_count * block_resolution * block_resolution * block_resolution;
	NNRT_CHECK_ARG(voxel_count < (1ll << 32) - 1, "integrate_non_rigid: more than 2^32 voxels in one call");
	cudaStream_t s = (cudaStream_t) stream;
	const int64_t pixel_count = (int64_t) image_height * image_width;
	unsigned long long* keys = nullptr;
	Scratch scratch(s);
	NNRT_CUDA(scratch.alloc(&keys, sizeof(unsigned long long) * pixel_count));
	NNRT_CUDA(cudaMemsetAsync(keys, 0, sizeof(unsigned long long) * pixel_count, s));
	IntegrateParams p{};
	p.block_indices = block_indices; p.block_keys = block_keys; p.tsdf = tsdf; p.weight = weight; p.weight_dtype = weight_dtype; p.voxel_color = voxel_color;
	p.block_resolution = block_resolution; p.voxel_size = voxel_size; p.truncation = sdf_truncation_distance;
	p.depth = depth; p.depth_dtype = depth_dtype; p.color = color; p.color_dtype = color_dtype; p.normals = depth_normals;
	p.height = image_height; p.width = image_width; p.color_height = color_height; p.color_width = color_width;
	p.depth_k = intrinsics_from_host(depth_intrinsics_host);
	p.color_k = color ? intrinsics_from_host(color_intrinsics_host) : p.depth_k;
	p.extrinsics = rigid_from_host(extrinsics_host);
	p.field = NodeField{nodes, node_rotations, node_translations, node_count, node_coverage_weights, node_coverage * node_coverage, minimum_valid_anchor_count};
	p.depth_scale = depth_scale; p.depth_max = depth_max;
	p.cos_keys = keys; p.voxel_count = voxel_count;
	if (voxel_count > 0) {
		const unsigned scan_grid = blocks_for(voxel_count, kThreads);
#define NNRT_TSDF_CASE(A_)                                                                          \
	case A_:                                                                                        \
		if (full_scan) integrate_non_rigid_kernel<A_><<<scan_grid, kThreads, 0, s>>>(p);            \
		else integrate_blocks_kernel<A_><<<(unsigned) block_count, kBlockThreads, 0, s>>>(p);      \
		break;
		switch (anchor_count) {
			NNRT_TSDF_CASE(1) NNRT_TSDF_CASE(2) NNRT_TSDF_CASE(3) NNRT_TSDF_CASE(4) NNRT_TSDF_CASE(5) NNRT_TSDF_CASE(6) NNRT_TSDF_CASE(7) NNRT_TSDF_CASE(8)
		}
#undef NNRT_TSDF_CASE
	}
	resolve_cosines_kernel<<<blocks_for(pixel_count, 256), 256, 0, s>>>(keys, pixel_count, cos_voxel_ray_to_normal);
	const cudaError_t err = cudaGetLastError();
	if (err != cudaSuccess) { set_last_error("integrate_non_rigid: launch failed: %s", cudaGetErrorString(err)); return kCudaError; }
	return kOk;
}

extern "C" int nnrt_b200_integrate_rigid(const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf, void* weight,
                                         int32_t weight_dtype, float* voxel_color, int32_t block_resolution, float voxel_size,
                                         float sdf_truncation_distance, const void* depth, int32_t depth_dtype, const void* color, int32_t color_dtype,
                                         int32_t image_height, int32_t image_width, int32_t color_height, int32_t color_width,
                                         const double* depth_intrinsics_host, const double* color_intrinsics_host, const double* extrinsics_host,
                                         float depth_scale, float depth_max, void* stream) {
	NNRT_CHECK_ARG(block_keys && tsdf && weight && depth && depth_intrinsics_host && extrinsics_host, "integrate_rigid: null pointer argument");
	NNRT_CHECK_ARG(block_count == 0 || block_indices, "integrate_rigid: block_indices is NULL");
	NNRT_CHECK_ARG(block_resolution >= 1 && image_height > 0 && image_width > 0, "integrate_rigid: bad sizes");
	NNRT_CHECK_ARG((weight_dtype == 0 || weight_dtype == 1) && (depth_dtype == 0 || depth_dtype == 1) && (color_dtype == 0 || color_dtype == 1),
	               "dtype codes must be 0 or 1");
	NNRT_CHECK_ARG(!color || (color_intrinsics_host && color_height > 0 && color_width > 0), "colour image needs its intrinsics and size");
	IntegrateParams p{};
	p.block_indices = block_indices; p.block_keys = block_keys; p.tsdf = tsdf; p.weight = weight; p.weight_dtype = weight_dtype; p.voxel_color = voxel_color;
	p.block_resolution = block_resolution; p.voxel_size = voxel_size; p.truncation = sdf_truncation_distance;
	p.depth = depth; p.depth_dtype = depth_dtype; p.color = color; p.color_dtype = color_dtype;
	p.height = image_height; p.width = image_width; p.color_height = color_height; p.color_width = color_width;
	p.depth_k = intrinsics_from_host(depth_intrinsics_host);
	p.color_k = color ? intrinsics_from_host(color_intrinsics_host) : p.depth_k;
	p.extrinsics = rigid_from_host(extrinsics_host);
	p.depth_scale = depth_scale; p.depth_max = depth_max;
	p.voxel_count = block_count * block_resolution * block_resolution * block_resolution;
	if (p.voxel_count > 0) integrate_rigid_kernel<<<blocks_for(p.voxel_count, kThreads), kThreads, 0, (cudaStream_t) stream>>>(p);
	const cudaError_t err = cudaGetLastError();
	if (err != cudaSuccess) { set_last_error("integrate_rigid: launch failed: %s", cudaGetErrorString(err)); return kCudaError; }
	return kOk;
}

extern "C" int nnrt_b200_integrate_non_rigid(
		const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf, const void* weight, int32_t weight_dtype,
		float* voxel_color, int32_t block_resolution, float voxel_size, float sdf_truncation_distance, const void* depth, int32_t depth_dtype,
		const void* color, int32_t color_dtype, const float* depth_normals, int32_t image_height, int32_t image_width, int32_t color_height,
		int32_t color_width, const double* depth_intrinsics_host, const double* color_intrinsics_host, const double* extrinsics_host,
		const float* nodes, const float* node_rotations, const float* node_translations, int32_t node_count, const float* node_coverage_weights,
		float node_coverage, int32_t anchor_count, int32_t minimum_valid_anchor_count, float depth_scale, float depth_max,
		float* cos_voxel_ray_to_normal, void* stream) {
	return integrate_non_rigid_impl(false, block_indices, block_count, block_keys, tsdf, weight, weight_dtype, voxel_color, block_resolution, voxel_size, sdf_truncation_distance, depth, depth_dtype,
	                                 color, color_dtype, depth_normals, image_height, image_width, color_height, color_width, depth_intrinsics_host, color_intrinsics_host,
	                                 extrinsics_host, nodes, node_rotations, node_translations, node_count, node_coverage_weights, node_coverage, anchor_count,
	                                 minimum_valid_anchor_count, depth_scale, depth_max, cos_voxel_ray_to_normal, stream);
}

extern "C" int nnrt_b200_integrate_non_rigid_full_scan(
		const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf, const void* weight, int32_t weight_dtype,
		float* voxel_color, int32_t block_resolution, float voxel_size, float sdf_truncation_distance, const void* depth, int32_t depth_dtype,
		const void* color, int32_t color_dtype, const float* depth_normals, int32_t image_height, int32_t image_width, int32_t color_height,
		int32_t color_width, const double* depth_intrinsics_host, const double* color_intrinsics_host, const double* extrinsics_host,
		const float* nodes, const float* node_rotations, const float* node_translations, int32_t node_count, const float* node_coverage_weights,
		float node_coverage, int32_t anchor_count, int32_t minimum_valid_anchor_count, float depth_scale, float depth_max,
		float* cos_voxel_ray_to_normal, void* stream) {
	return integrate_non_rigid_impl(true, block_indices, block_count, block_keys, tsdf, weight, weight_dtype, voxel_color, block_resolution, voxel_size, sdf_truncation_distance, depth, depth_dtype,
	                                 color, color_dtype, depth_normals, image_height, image_width, color_height, color_width, depth_intrinsics_host, color_intrinsics_host,
	                                 extrinsics_host, nodes, node_rotations, node_translations, node_count, node_coverage_weights, node_coverage, anchor_count,
	                                 minimum_valid_anchor_count, depth_scale, depth_max, cos_voxel_ray_to_normal, stream);
}

extern "C" int nnrt_b200_bounding_boxes_of_warped_blocks(const int32_t* block_keys, int64_t block_count, const float* nodes, const float* node_rotations,
                                                         const float* node_translations, int32_t node_count, float node_coverage, int32_t anchor_count,
                                                         int32_t minimum_valid_anchor_count, float voxel_size, int32_t block_resolution,
                                                         const double* extrinsics_host, float* bounding_boxes, void* stream) {
	if (block_count == 0) return kOk;
	NNRT_CHECK_ARG(block_keys && nodes && node_rotations && node_translations && bounding_boxes, "bounding_boxes_of_warped_blocks: null pointer argument");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	NNRT_CHECK_ARG(block_count > 0 && node_count > 0 && block_resolution >= 1, "bounding_boxes_of_warped_blocks: bad sizes");
	cudaStream_t s = (cudaStream_t) stream;
	const NodeField field{nodes, node_rotations, node_translations, node_count, nullptr, node_coverage * node_coverage, minimum_valid_anchor_count};
	const float side = (float) block_resolution * voxel_size;
	const Rigid extrinsics = rigid_from_host(extrinsics_host);
	const unsigned grid = blocks_for(block_count * 8, kThreads);
#define NNRT_BOX_CASE(A_) case A_: warped_block_boxes_kernel<A_><<<grid, kThreads, 0, s>>>(block_keys, block_count, field, side, extrinsics, bounding_boxes); break;
	switch (anchor_count) {
		NNRT_BOX_CASE(1) NNRT_BOX_CASE(2) NNRT_BOX_CASE(3) NNRT_BOX_CASE(4) NNRT_BOX_CASE(5) NNRT_BOX_CASE(6) NNRT_BOX_CASE(7) NNRT_BOX_CASE(8)
	}
#undef NNRT_BOX_CASE
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_boxes_intersecting_surface_mask(const float* boxes, int64_t box_count, const double* intrinsics_host, const void* depth,
                                                         int32_t depth_dtype, int32_t image_height, int32_t image_width, float depth_scale, float depth_max,
                                                         int32_t stride, float truncation_distance, uint8_t* mask, void* stream) {
	if (box_count == 0) return kOk;
	NNRT_CHECK_ARG(boxes && intrinsics_host && depth && mask, "boxes_intersecting_surface_mask: null pointer argument");
	NNRT_CHECK_ARG(box_count > 0 && image_height > 0 && image_width > 0 && stride >= 1, "boxes_intersecting_surface_mask: bad sizes");
	NNRT_CHECK_ARG(depth_dtype == 0 || depth_dtype == 1, "depth dtype must be float32 (0) or uint16 (1)");
	boxes_intersecting_surface_kernel<<<blocks_for(box_count * 32, kThreads), kThreads, 0, (cudaStream_t) stream>>>(
			boxes, box_count, intrinsics_from_host(intrinsics_host), depth, depth_dtype, image_height, image_width, depth_scale, depth_max, stride,
			truncation_distance, mask);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
