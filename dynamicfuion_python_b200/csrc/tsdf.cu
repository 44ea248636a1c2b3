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

namespace nnrt_b200 {
namespace {

constexpr int kThreads = 128;
constexpr int kNodeTile = 1024;

struct IntegrateParams {
	const int64_t* block_indices; const int32_t* block_keys; float* tsdf; const void* weight; int weight_dtype; float* voxel_color;
	int block_resolution; float voxel_size, truncation;
	const void* depth; int depth_dtype; const void* color; int color_dtype; const float* normals;
	int height, width, color_height, color_width;
	Intrinsics depth_k, color_k; Rigid extrinsics;
	const float *nodes, *rotations, *translations; int node_count; const float* node_weights; float fixed_c2;
	int min_valid; float depth_scale, depth_max;
	unsigned long long* cos_keys; int64_t voxel_count;
};

__device__ __forceinline__ bool in_boundary(float x, float y, int rows, int cols) {  // Open3D ArrayIndexer::InBoundary
	return y >= 0 && x >= 0 && y <= rows - 1.0f && x <= cols - 1.0f;
}

template<int A>
__global__ void __launch_bounds__(kThreads) integrate_non_rigid_kernel(IntegrateParams p) {
	__shared__ float4 tile[kNodeTile];
	const int64_t workload = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	const bool active = workload < p.voxel_count;
	const int r = p.block_resolution;
	const int64_t r3 = (int64_t) r * r * r;
	int64_t block_index = 0, voxel_in_block = 0;
	float3 voxel = make_float3(0.f, 0.f, 0.f);
	if (active) {
		block_index = p.block_indices[workload / r3];
		voxel_in_block = workload % r3;
		const int32_t* key = p.block_keys + block_index * 3;
		const int xl = (int) (voxel_in_block % r), yl = (int) ((voxel_in_block / r) % r), zl = (int) (voxel_in_block / ((int64_t) r * r));
		const float3 metric = make_float3((float) (key[0] * r + xl) * p.voxel_size, (float) (key[1] * r + yl) * p.voxel_size,
		                                  (float) (key[2] * r + zl) * p.voxel_size);
		voxel = rigid_transform(p.extrinsics, metric);
	}
	// --- K nearest nodes by distance (norm), replace-current-maximum in ascending node order
	int idx[A];
	float w[A], c2s[A];
#pragma unroll
	for (int a = 0; a < A; a++) { idx[a] = -1; w[a] = INFINITY; c2s[a] = p.node_weights ? p.node_weights[0] : p.fixed_c2; }
	int max_at = 0;
	float max_distance = INFINITY, skip_above = INFINITY;
	for (int base = 0; base < p.node_count; base += kNodeTile) {
		const int count = min(kNodeTile, p.node_count - base);
		__syncthreads();
		for (int k = threadIdx.x; k < count; k += blockDim.x) {
			const int n = base + k;
			tile[k] = make_float4(p.nodes[n * 3], p.nodes[n * 3 + 1], p.nodes[n * 3 + 2], p.node_weights ? p.node_weights[n] : p.fixed_c2);
		}
		__syncthreads();
		if (!active) continue;
		for (int k = 0; k < count; k++) {
			const float4 g = tile[k];
			const float dx = g.x - voxel.x, dy = g.y - voxel.y, dz = g.z - voxel.z;
			const float d2 = sum3(dx * dx, dy * dy, dz * dz);
			if (d2 > skip_above) continue;  // screened on the squared distance: sqrt(d2) >= max_distance for certain (see below)
			const float distance = sqrtf(d2);
			if (max_distance > distance) {
#pragma unroll
				for (int a = 0; a < A; a++)
					if (a == max_at) { w[a] = distance; idx[a] = base + k; c2s[a] = g.w; }
				max_at = 0;
				max_distance = w[0];
#pragma unroll
				for (int a = 1; a < A; a++)
					if (w[a] > max_distance) { max_at = a; max_distance = w[a]; }
				// m * m * (1 + 1e-6) exceeds the exact m^2 whatever the roundings do; above it the correctly rounded root cannot be
				// below m, so only the few candidates near the current k-th distance pay for the IEEE square root
				skip_above = max_distance * max_distance * 1.000001f;
			}
		}
	}
	if (!active) return;
	// --- thresholded weights (geometry/functional/kernel/WarpUtilities.h:95-153, distances not squared) and normalisation (:36-46)
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
	if (valid < p.min_valid) return;
	if (weight_sum > 0.0f) {
#pragma unroll
		for (int a = 0; a < A; a++) w[a] = __fdiv_rn(w[a], weight_sum);
	} else if (valid > 0) {
#pragma unroll
		for (int a = 0; a < A; a++) w[a] = 1.0f / (float) valid;
	}
	// --- blend warp (WarpUtilities.h:428-445)
	float3 warped = make_float3(0.f, 0.f, 0.f);
#pragma unroll
	for (int a = 0; a < A; a++) {
		const int n = idx[a];
		if (n == -1) continue;
		const float3 g = make_float3(p.nodes[n * 3], p.nodes[n * 3 + 1], p.nodes[n * 3 + 2]);
		const float3 t = make_float3(p.translations[n * 3], p.translations[n * 3 + 1], p.translations[n * 3 + 2]);
		warped = warped + w[a] * ((g + matvec3(p.rotations + (int64_t) n * 9, voxel - g)) + t);
	}
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

__global__ void __launch_bounds__(256) resolve_cosines_kernel(const unsigned long long* __restrict__ keys, int64_t pixel_count, float* __restrict__ cosines) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < pixel_count) cosines[i] = keys[i] == 0ull ? 0.f : __uint_as_float((unsigned) (keys[i] & 0xffffffffull));
}

}  // namespace
}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_integrate_non_rigid(
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
	NNRT_CUDA(scratch_malloc_async(&keys, sizeof(unsigned long long) * pixel_count, s));
	NNRT_CUDA(cudaMemsetAsync(keys, 0, sizeof(unsigned long long) * pixel_count, s));
	IntegrateParams p{};
	p.block_indices = block_indices; p.block_keys = block_keys; p.tsdf = tsdf; p.weight = weight; p.weight_dtype = weight_dtype; p.voxel_color = voxel_color;
	p.block_resolution = block_resolution; p.voxel_size = voxel_size; p.truncation = sdf_truncation_distance;
	p.depth = depth; p.depth_dtype = depth_dtype; p.color = color; p.color_dtype = color_dtype; p.normals = depth_normals;
	p.height = image_height; p.width = image_width; p.color_height = color_height; p.color_width = color_width;
	p.depth_k = intrinsics_from_host(depth_intrinsics_host);
	p.color_k = color ? intrinsics_from_host(color_intrinsics_host) : p.depth_k;
	p.extrinsics = rigid_from_host(extrinsics_host);
	p.nodes = nodes; p.rotations = node_rotations; p.translations = node_translations; p.node_count = node_count;
	p.node_weights = node_coverage_weights; p.fixed_c2 = node_coverage * node_coverage;
	p.min_valid = minimum_valid_anchor_count; p.depth_scale = depth_scale; p.depth_max = depth_max;
	p.cos_keys = keys; p.voxel_count = voxel_count;
	if (voxel_count > 0) {
		const unsigned grid = blocks_for(voxel_count, kThreads);
#define NNRT_TSDF_CASE(A_) case A_: integrate_non_rigid_kernel<A_><<<grid, kThreads, 0, s>>>(p); break;
		switch (anchor_count) {
			NNRT_TSDF_CASE(1) NNRT_TSDF_CASE(2) NNRT_TSDF_CASE(3) NNRT_TSDF_CASE(4) NNRT_TSDF_CASE(5) NNRT_TSDF_CASE(6) NNRT_TSDF_CASE(7) NNRT_TSDF_CASE(8)
		}
#undef NNRT_TSDF_CASE
	}
	resolve_cosines_kernel<<<blocks_for(pixel_count, 256), 256, 0, s>>>(keys, pixel_count, cos_voxel_ray_to_normal);
	const cudaError_t err = cudaGetLastError();
	cudaFreeAsync(keys, s);
	if (err != cudaSuccess) { set_last_error("integrate_non_rigid: launch failed: %s", cudaGetErrorString(err)); return kCudaError; }
	return kOk;
}
