// Warp-field geometry kernels: Euclidean anchors/weights, linear-blend warp, depth unprojection,
// point-to-plane distances, warped-surface Jacobians (stage-level C ABI entry points).
#include "host_util.h"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {

// K1. One thread per point; nodes staged through shared memory as float4 (x, y, z, coverage^2) in tiles so every
// thread streams the same broadcast node record. Top-A selection follows the reference's replace-current-max
// algorithm (cpp/core/kernel/KnnUtilities.h:37-118): slot order is algorithm-defined and reproduced exactly.
template<int A>
__global__ void __launch_bounds__(256) anchors_knn_kernel(const float* __restrict__ points, int64_t point_count,
                                                           const float* __restrict__ nodes, int node_count,
                                                           const float* __restrict__ node_weights, float fixed_c2, int min_valid,
                                                           int32_t* __restrict__ anchors, float* __restrict__ weights) {
	constexpr int kTile = 1024;
	__shared__ float4 tile[kTile];
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	const bool active = i < point_count;
	float3 p = make_float3(0.f, 0.f, 0.f);
	if (active) p = make_float3(points[i * 3], points[i * 3 + 1], points[i * 3 + 2]);
	int idx[A];
	float d2s[A], c2s[A];
#pragma unroll
	for (int a = 0; a < A; a++) { idx[a] = -1; d2s[a] = INFINITY; c2s[a] = fixed_c2; }
	int max_at = 0;
	float max_d2 = INFINITY;
	for (int base = 0; base < node_count; base += kTile) {
		const int count = min(kTile, node_count - base);
		__syncthreads();
		for (int k = threadIdx.x; k < count; k += blockDim.x) {
			const int n = base + k;
			tile[k] = make_float4(nodes[n * 3], nodes[n * 3 + 1], nodes[n * 3 + 2], node_weights ? node_weights[n] : fixed_c2);
		}
		__syncthreads();
		if (!active) continue;
		for (int k = 0; k < count; k++) {
			const float4 g = tile[k];
			const float dx = g.x - p.x, dy = g.y - p.y, dz = g.z - p.z;
			const float d2 = sum3(dx * dx, dy * dy, dz * dz);
			if (d2 < max_d2) {
#pragma unroll
				for (int a = 0; a < A; a++)
					if (a == max_at) { d2s[a] = d2; idx[a] = base + k; c2s[a] = g.w; }
				max_at = 0;
				max_d2 = d2s[0];
#pragma unroll
				for (int a = 1; a < A; a++)
					if (d2s[a] > max_d2) { max_at = a; max_d2 = d2s[a]; }
			}
		}
	}
	if (!active) return;
	// weights: cpp/geometry/functional/kernel/WarpUtilities.h:35-152
	float w[A];
	float weight_sum = 0.0f;
	int valid = A;
	if (min_valid > 0) {
		valid = 0;
#pragma unroll
		for (int a = 0; a < A; a++) {
			w[a] = d2s[a];  // the reference leaves the squared distance in the weight slot of a dropped anchor
			if (d2s[a] > 4 * c2s[a]) { idx[a] = -1; continue; }
			const float weight = expf(-d2s[a] / (2 * c2s[a]));
			weight_sum += weight;
			w[a] = weight;
			valid++;
		}
	} else {
#pragma unroll
		for (int a = 0; a < A; a++) {
			const float weight = expf(-d2s[a] / (2 * c2s[a]));
			weight_sum += weight;
			w[a] = weight;
		}
	}
	if (!(min_valid > 0 && valid < min_valid)) {
		if (weight_sum > 0.0f) {
#pragma unroll
			for (int a = 0; a < A; a++) w[a] /= weight_sum;
		} else if (valid > 0) {
#pragma unroll
			for (int a = 0; a < A; a++) w[a] = 1.0f / (float) valid;
		}
	}
#pragma unroll
	for (int a = 0; a < A; a++) { anchors[i * A + a] = idx[a]; weights[i * A + a] = w[a]; }
}

// Stage-level linear-blend warp on [V,3] tensors (the fitter uses the fused float4 variant in fitter_kernels.cu).
__global__ void __launch_bounds__(256) warp_points_kernel(const float* __restrict__ points, const float* __restrict__ normals,
                                                           int64_t point_count, const float* __restrict__ nodes,
                                                           const float* __restrict__ rotations, const float* __restrict__ translations,
                                                           const int32_t* __restrict__ anchors, const float* __restrict__ anchor_weights,
                                                           int anchor_count, int min_valid, Rigid extrinsics,
                                                           float* __restrict__ warped_points, float* __restrict__ warped_normals) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= point_count) return;
	const float3 pc = rigid_transform(extrinsics, make_float3(points[i * 3], points[i * 3 + 1], points[i * 3 + 2]));
	float3 nc = make_float3(0.f, 0.f, 0.f);
	if (normals) nc = rigid_rotate(extrinsics, make_float3(normals[i * 3], normals[i * 3 + 1], normals[i * 3 + 2]));
	float3 wp = make_float3(0.f, 0.f, 0.f), wn = make_float3(0.f, 0.f, 0.f);
	bool do_warp = true;
	if (min_valid > 0) {
		int valid = 0;
		for (int a = 0; a < anchor_count; a++) valid += anchors[i * anchor_count + a] != -1;
		do_warp = valid >= min_valid;
	}
	if (do_warp) {
		for (int a = 0; a < anchor_count; a++) {
			const int n = anchors[i * anchor_count + a];
			if (n == -1) continue;
			const float weight = anchor_weights[i * anchor_count + a];
			const float* R = rotations + (int64_t) n * 9;
			const float3 g = make_float3(nodes[n * 3], nodes[n * 3 + 1], nodes[n * 3 + 2]);
			const float3 t = make_float3(translations[n * 3], translations[n * 3 + 1], translations[n * 3 + 2]);
			wp = wp + weight * ((g + matvec3(R, pc - g)) + t);
			if (normals) {  // scalar folded into the matrix, as Eigen evaluates w * (R * n) (WarpUtilities.h:466)
				float wR[9];
#pragma unroll
				for (int k = 0; k < 9; k++) wR[k] = weight * R[k];
				wn = wn + matvec3(wR, nc);
			}
		}
	}
	warped_points[i * 3] = wp.x; warped_points[i * 3 + 1] = wp.y; warped_points[i * 3 + 2] = wp.z;
	if (normals && warped_normals) { warped_normals[i * 3] = wn.x; warped_normals[i * 3 + 1] = wn.y; warped_normals[i * 3 + 2] = wn.z; }
}

// Dual-quaternion blending (BASELINE.json north_star names it next to linear blending; the reference has no implementation of it —
// cpp/geometry/TransformationMode.h only distinguishes how rotations are STORED — so this follows the published algorithm, Kavan et al.,
// "Geometric Skinning with Approximate Dual Quaternion Blending", 2008): each anchor's rigid motion x -> R (x - g) + g + t becomes a
// unit dual quaternion, the weighted sum (signs aligned with the first valid anchor) is normalised, and the blended rigid motion
// is applied to the point and its rotation to the normal. Same inputs, validity rules and extrinsics handling as warp_points_kernel.
struct Quat { float w, x, y, z; };
__device__ __forceinline__ Quat quat_mul(Quat a, Quat b) {
	return Quat{a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
	            a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}
// rotation matrix (row-major) -> unit quaternion, Shepperd's method (largest of the four squared components first)
__device__ __forceinline__ Quat quat_from_matrix(const float* __restrict__ R) {
	const float trace = R[0] + R[4] + R[8];
	Quat q;
	if (trace > 0.f) {
		const float s = sqrtf(trace + 1.f) * 2.f;
		q = Quat{0.25f * s, (R[7] - R[5]) / s, (R[2] - R[6]) / s, (R[3] - R[1]) / s};
	} else if (R[0] > R[4] && R[0] > R[8]) {
		const float s = sqrtf(1.f + R[0] - R[4] - R[8]) * 2.f;
		q = Quat{(R[7] - R[5]) / s, 0.25f * s, (R[1] + R[3]) / s, (R[2] + R[6]) / s};
	} else if (R[4] > R[8]) {
		const float s = sqrtf(1.f + R[4] - R[0] - R[8]) * 2.f;
		q = Quat{(R[2] - R[6]) / s, (R[1] + R[3]) / s, 0.25f * s, (R[5] + R[7]) / s};
	} else {
		const float s = sqrtf(1.f + R[8] - R[0] - R[4]) * 2.f;
		q = Quat{(R[3] - R[1]) / s, (R[2] + R[6]) / s, (R[5] + R[7]) / s, 0.25f * s};
	}
	const float n = sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
	return Quat{q.w / n, q.x / n, q.y / n, q.z / n};
}

__global__ void __launch_bounds__(256) warp_points_dqb_kernel(const float* __restrict__ points, const float* __restrict__ normals,
                                                               int64_t point_count, const float* __restrict__ nodes,
                                                               const float* __restrict__ rotations, const float* __restrict__ translations,
                                                               const int32_t* __restrict__ anchors, const float* __restrict__ anchor_weights,
                                                               int anchor_count, int min_valid, Rigid extrinsics,
                                                               float* __restrict__ warped_points, float* __restrict__ warped_normals) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= point_count) return;
	const float3 pc = rigid_transform(extrinsics, make_float3(points[i * 3], points[i * 3 + 1], points[i * 3 + 2]));
	float3 nc = make_float3(0.f, 0.f, 0.f);
	if (normals) nc = rigid_rotate(extrinsics, make_float3(normals[i * 3], normals[i * 3 + 1], normals[i * 3 + 2]));
	float3 wp = make_float3(0.f, 0.f, 0.f), wn = make_float3(0.f, 0.f, 0.f);
	int valid = 0;
	for (int a = 0; a < anchor_count; a++) valid += anchors[i * anchor_count + a] != -1;
	if (valid > 0 && !(min_valid > 0 && valid < min_valid)) {
		Quat real{0.f, 0.f, 0.f, 0.f}, dual{0.f, 0.f, 0.f, 0.f}, pivot{1.f, 0.f, 0.f, 0.f};
		bool first = true;
		for (int a = 0; a < anchor_count; a++) {
			const int n = anchors[i * anchor_count + a];
			if (n == -1) continue;
			float weight = anchor_weights[i * anchor_count + a];
			const float* R = rotations + (int64_t) n * 9;
			const float3 g = make_float3(nodes[n * 3], nodes[n * 3 + 1], nodes[n * 3 + 2]);
			const float3 t = make_float3(translations[n * 3], translations[n * 3 + 1], translations[n * 3 + 2]);
			const float3 tau = (g + t) - matvec3(R, g);  // x -> R x + tau
			const Quat q = quat_from_matrix(R);
			const Quat d = quat_mul(Quat{0.f, 0.5f * tau.x, 0.5f * tau.y, 0.5f * tau.z}, q);
			if (first) { pivot = q; first = false; }
			if (q.w * pivot.w + q.x * pivot.x + q.y * pivot.y + q.z * pivot.z < 0.f) weight = -weight;  // q and -q are the same rotation
			real = Quat{real.w + weight * q.w, real.x + weight * q.x, real.y + weight * q.y, real.z + weight * q.z};
			dual = Quat{dual.w + weight * d.w, dual.x + weight * d.x, dual.y + weight * d.y, dual.z + weight * d.z};
		}
		const float norm = sqrtf(real.w * real.w + real.x * real.x + real.y * real.y + real.z * real.z);
		if (norm > 0.f) {
			const Quat q{real.w / norm, real.x / norm, real.y / norm, real.z / norm};
			const Quat d{dual.w / norm, dual.x / norm, dual.y / norm, dual.z / norm};
			const Quat tq = quat_mul(d, Quat{q.w, -q.x, -q.y, -q.z});  // translation = 2 * vector part of d * conj(q)
			const float Rb[9] = {1.f - 2.f * (q.y * q.y + q.z * q.z), 2.f * (q.x * q.y - q.w * q.z), 2.f * (q.x * q.z + q.w * q.y),
			                     2.f * (q.x * q.y + q.w * q.z), 1.f - 2.f * (q.x * q.x + q.z * q.z), 2.f * (q.y * q.z - q.w * q.x),
			                     2.f * (q.x * q.z - q.w * q.y), 2.f * (q.y * q.z + q.w * q.x), 1.f - 2.f * (q.x * q.x + q.y * q.y)};
			wp = matvec3(Rb, pc) + make_float3(2.f * tq.x, 2.f * tq.y, 2.f * tq.z);
			wn = matvec3(Rb, nc);
		}
	}
	warped_points[i * 3] = wp.x; warped_points[i * 3 + 1] = wp.y; warped_points[i * 3 + 2] = wp.z;
	if (normals && warped_normals) { warped_normals[i * 3] = wn.x; warped_normals[i * 3 + 1] = wn.y; warped_normals[i * 3 + 2] = wn.z; }
}

template<typename TDepth>
__global__ void __launch_bounds__(256) unproject_kernel(const TDepth* __restrict__ depth, int rows, int cols, Intrinsics k, Rigid pose,
                                                         float depth_scale, float depth_max, float* __restrict__ points,
                                                         uint8_t* __restrict__ mask) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (int64_t) rows * cols) return;
	const int y = (int) (i / cols), x = (int) (i % cols);
	const float d = (float) depth[i] / depth_scale;
	float3 out = make_float3(0.f, 0.f, 0.f);
	uint8_t m = 0;
	if (d > 0 && d < depth_max) {
		const float3 c = make_float3(((float) x - k.cx) * d / k.fx, ((float) y - k.cy) * d / k.fy, d);
		out = rigid_transform(pose, c);
		m = 1;
	}
	points[i * 3] = out.x; points[i * 3 + 1] = out.y; points[i * 3 + 2] = out.z;
	if (mask) mask[i] = m;
}

__global__ void __launch_bounds__(256) point_to_plane_kernel(const float* __restrict__ n1, const float* __restrict__ v1,
                                                              const float* __restrict__ v2, int64_t count, float* __restrict__ out) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	out[i] = dot3(n1[i * 3], n1[i * 3 + 1], n1[i * 3 + 2], v1[i * 3] - v2[i * 3], v1[i * 3 + 1] - v2[i * 3 + 1],
	              v1[i * 3 + 2] - v2[i * 3 + 2]);
}

__global__ void __launch_bounds__(256) warped_surface_jacobians_kernel(
		const float* __restrict__ positions, const float* __restrict__ normals, int64_t vertex_count, const float* __restrict__ nodes,
		const float* __restrict__ rotations, const int32_t* __restrict__ anchors, const float* __restrict__ weights, int anchor_count,
		float* __restrict__ jv, float* __restrict__ jn) {
	const int64_t wi = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (wi >= vertex_count * anchor_count) return;
	const int64_t v = wi / anchor_count;
	const int n = anchors[wi];
	float4 outv = make_float4(0.f, 0.f, 0.f, 0.f);
	float3 outn = make_float3(0.f, 0.f, 0.f);
	if (n != -1) {
		const float w = weights[wi];
		const float* R = rotations + (int64_t) n * 9;
		const float3 g = make_float3(nodes[n * 3], nodes[n * 3 + 1], nodes[n * 3 + 2]);
		const float3 rv = matvec3(R, make_float3(positions[v * 3], positions[v * 3 + 1], positions[v * 3 + 2]) - g);
		const float3 rn = matvec3(R, make_float3(normals[v * 3], normals[v * 3 + 1], normals[v * 3 + 2]));
		outv = make_float4(-w * rv.x, -w * rv.y, -w * rv.z, w);
		outn = make_float3(-w * rn.x, -w * rn.y, -w * rn.z);
	}
	reinterpret_cast<float4*>(jv)[wi] = outv;
	jn[wi * 3] = outn.x; jn[wi * 3 + 1] = outn.y; jn[wi * 3 + 2] = outn.z;
}


// Shortest-path anchors (core/kernel/KnnUtilities_PriorityQueue.h:102-160, geometry/functional/kernel/WarpUtilities.h:344-375): per
// point, Dijkstra over the node graph seeded with the nearest node that is not an anchor yet (re-seeded when a component is used
// up), first `anchor_count` nodes settled. One thread per point. The 40-entry priority queue lives in shared memory, slot-major
// ([slot][thread]: conflict-free), and follows libstdc++'s std::push_heap / std::pop_heap (the reference's CPU heap,
// core/heap/CPU/DeviceHeapCPU.h) move for move, so equal path lengths settle in the same order; inserts into a full queue are
// dropped, as there. anchors must be pre-filled with -1 (the exclusion scan reads all `anchor_count` slots, as the reference's does).
constexpr int kPathQueueCapacity = 40;
constexpr int kPathThreads = 64;
constexpr int kPathNodeCache = 1536;  // 18 KB next to the 20 KB of queues

struct PathQueue {
	float* keys;  // [kPathQueueCapacity][kPathThreads]
	int* nodes;
	int size;
	__device__ __forceinline__ float& key(int i) { return keys[i * kPathThreads]; }
	__device__ __forceinline__ int& node(int i) { return nodes[i * kPathThreads]; }
	// std::__push_heap with comp(a, b) = a.key > b.key
	__device__ __forceinline__ void sift_up(int hole, int top, float value_key, int value_node) {
		int parent = (hole - 1) / 2;
		while (hole > top && key(parent) > value_key) {
			key(hole) = key(parent); node(hole) = node(parent);
			hole = parent;
			parent = (hole - 1) / 2;
		}
		key(hole) = value_key; node(hole) = value_node;
	}
	__device__ __forceinline__ void insert(float k, int n) {
		if (size >= kPathQueueCapacity) return;
		size++;
		sift_up(size - 1, 0, k, n);
	}
	// std::pop_heap followed by taking the last element: returns the minimum
	__device__ __forceinline__ void pop(float& out_key, int& out_node) {
		out_key = key(0); out_node = node(0);
		size--;
		if (size == 0) return;
		const float value_key = key(size);
		const int value_node = node(size);
		const int len = size;  // std::__adjust_heap(first, 0, len, value)
		int hole = 0, child = 0;
		while (child < (len - 1) / 2) {
			child = 2 * (child + 1);
			if (key(child) > key(child - 1)) child--;
			key(hole) = key(child); node(hole) = node(child);
			hole = child;
		}
		if ((len & 1) == 0 && child == (len - 2) / 2) {
			child = 2 * (child + 1);
			key(hole) = key(child - 1); node(hole) = node(child - 1);
			hole = child - 1;
		}
		sift_up(hole, 0, value_key, value_node);
	}
};

template<int A>
__global__ void __launch_bounds__(kPathThreads) anchors_shortest_path_kernel(const float* __restrict__ points, int64_t point_count, const float* __restrict__ nodes,
                                                                             int node_count, const float* __restrict__ node_weights, float fixed_c2,
                                                                             const int32_t* __restrict__ edges, int graph_degree,
                                                                             int32_t* __restrict__ anchors, float* __restrict__ weights) {
	__shared__ float queue_keys[kPathQueueCapacity * kPathThreads];
	__shared__ int queue_nodes[kPathQueueCapacity * kPathThreads];
	// node positions are read N times per point by the seeding scan and again along every edge: staged in shared memory when they fit
	__shared__ float s_nodes[kPathNodeCache * 3];
	const bool cached = node_count <= kPathNodeCache;
	if (cached)
		for (int k = threadIdx.x; k < node_count * 3; k += blockDim.x) s_nodes[k] = nodes[k];
	__syncthreads();
	const float* __restrict__ node_xyz = cached ? s_nodes : nodes;
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	const bool active = i < point_count;
	const float px = active ? points[i * 3] : 0.f, py = active ? points[i * 3 + 1] : 0.f, pz = active ? points[i * 3 + 2] : 0.f;
	// With more nodes than the cache holds, the FIRST seeding scan (no anchor excluded yet, and the bulk of the work) still goes
	// through shared memory: all threads of the CTA run it together, tile by tile. Later re-seeds (a graph component used up) are rare
	// and read the nodes from global memory.
	int first_seed = -1;
	float first_seed_d2 = INFINITY;
	if (!cached) {
		constexpr int kTile = kPathNodeCache;
		for (int base = 0; base < node_count; base += kTile) {
			const int count = min(kTile, node_count - base);
			__syncthreads();
			for (int k = threadIdx.x; k < count * 3; k += blockDim.x) s_nodes[k] = nodes[(int64_t) base * 3 + k];
			__syncthreads();
			for (int k = 0; k < count; k++) {
				const float dx = s_nodes[k * 3] - px, dy = s_nodes[k * 3 + 1] - py, dz = s_nodes[k * 3 + 2] - pz;
				const float d2 = sum3(dx * dx, dy * dy, dz * dz);
				if (d2 < first_seed_d2) { first_seed_d2 = d2; first_seed = base + k; }
			}
		}
	}
	if (!active) return;
	PathQueue queue{queue_keys + threadIdx.x, queue_nodes + threadIdx.x, 0};
	int idx[A];
	float dist[A];
#pragma unroll
	for (int a = 0; a < A; a++) { idx[a] = -1; dist[a] = 0.f; }
	int discovered = 0;
	while (discovered < A) {
		int closest = -1;
		float closest_d2 = INFINITY;
		if (first_seed >= 0) {
			closest = first_seed; closest_d2 = first_seed_d2;
			first_seed = -1;
		} else {
			for (int n = 0; n < node_count; n++) {
				bool excluded = false;
#pragma unroll
				for (int a = 0; a < A; a++) excluded |= idx[a] == n;
				if (excluded) continue;
				const float dx = node_xyz[n * 3] - px, dy = node_xyz[n * 3 + 1] - py, dz = node_xyz[n * 3 + 2] - pz;
				const float d2 = sum3(dx * dx, dy * dy, dz * dz);
				if (d2 < closest_d2) { closest_d2 = d2; closest = n; }
			}
		}
		if (closest == -1) break;
		queue.insert(sqrtf(closest_d2), closest);
		while (queue.size > 0 && discovered < A) {
			float source_key;
			int source;
			queue.pop(source_key, source);
			bool processed = false;
#pragma unroll
			for (int a = 0; a < A; a++) processed |= a < discovered && idx[a] == source;
			if (processed) continue;
#pragma unroll
			for (int a = 0; a < A; a++)
				if (a == discovered) { dist[a] = source_key; idx[a] = source; }
			discovered++;
			if (discovered >= A) break;
			const float sx = node_xyz[source * 3], sy = node_xyz[source * 3 + 1], sz = node_xyz[source * 3 + 2];
			for (int e = 0; e < graph_degree; e++) {
				const int target = __ldg(edges + (int64_t) source * graph_degree + e);
				if (target > -1) {
					const float dx = node_xyz[target * 3] - sx, dy = node_xyz[target * 3 + 1] - sy, dz = node_xyz[target * 3 + 2] - sz;
					queue.insert(source_key + sqrtf(sum3(dx * dx, dy * dy, dz * dz)), target);
				}
			}
		}
	}
	// weights from PATH lengths (not squared: WarpUtilities.h:61-64,84-87), normalised over all anchors
	float w[A];
	float weight_sum = 0.0f;
#pragma unroll
	for (int a = 0; a < A; a++) {
		const float square_distance = dist[a] * dist[a];
		const float c2 = node_weights ? node_weights[max(idx[a], 0)] : fixed_c2;
		w[a] = expf(-square_distance / (2 * c2));
		weight_sum += w[a];
	}
#pragma unroll
	for (int a = 0; a < A; a++) {
		anchors[i * A + a] = idx[a];
		weights[i * A + a] = weight_sum > 0.0f ? w[a] / weight_sum : 1.0f / (float) A;
	}
}

}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_compute_anchors_shortest_path(const float* points, int64_t point_count, const float* nodes, int32_t node_count,
                                                       const float* node_weights, float node_coverage, const int32_t* edges, int32_t graph_degree,
                                                       int32_t anchor_count, int32_t* anchors, float* weights, void* stream) {
	NNRT_CHECK_ARG(points && nodes && edges && anchors && weights, "compute_anchors_shortest_path: null pointer argument");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	NNRT_CHECK_ARG(graph_degree >= 1, "edges must have shape [node_count, degree] with degree >= 1, got degree %d", graph_degree);
	// with fewer nodes than anchors the reference reads never-written distances and node_weights[-1]
	NNRT_CHECK_ARG(node_count >= anchor_count, "node count (%d) must be at least anchor_count (%d)", node_count, anchor_count);
	NNRT_CHECK_ARG(point_count >= 0, "negative point count");
	if (point_count == 0) return kOk;
	auto s = (cudaStream_t) stream;
	const float c2 = node_coverage * node_coverage;
	const unsigned grid = blocks_for(point_count, kPathThreads);
#define NNRT_PATH_CASE(A_) \
	case A_: anchors_shortest_path_kernel<A_><<<grid, kPathThreads, 0, s>>>(points, point_count, nodes, node_count, node_weights, c2, edges, graph_degree, anchors, weights); break;
	switch (anchor_count) {
		NNRT_PATH_CASE(1) NNRT_PATH_CASE(2) NNRT_PATH_CASE(3) NNRT_PATH_CASE(4)
		NNRT_PATH_CASE(5) NNRT_PATH_CASE(6) NNRT_PATH_CASE(7) NNRT_PATH_CASE(8)
	}
#undef NNRT_PATH_CASE
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_compute_anchors_euclidean(const float* points, int64_t point_count, const float* nodes, int32_t node_count,
                                                   const float* node_weights, float node_coverage, int32_t anchor_count,
                                                   int32_t minimum_valid_anchor_count, int32_t* anchors, float* weights, void* stream) {
	NNRT_CHECK_ARG(points && nodes && anchors && weights, "compute_anchors: null pointer argument");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	NNRT_CHECK_ARG(node_count >= 0 && point_count >= 0, "negative count");
	if (point_count == 0) return kOk;
	auto s = (cudaStream_t) stream;
	const float c2 = node_coverage * node_coverage;
	const unsigned grid = blocks_for(point_count, 256);
#define NNRT_ANCHOR_CASE(A_) \
	case A_: anchors_knn_kernel<A_><<<grid, 256, 0, s>>>(points, point_count, nodes, node_count, node_weights, c2, minimum_valid_anchor_count, anchors, weights); break;
	switch (anchor_count) {
		NNRT_ANCHOR_CASE(1) NNRT_ANCHOR_CASE(2) NNRT_ANCHOR_CASE(3) NNRT_ANCHOR_CASE(4)
		NNRT_ANCHOR_CASE(5) NNRT_ANCHOR_CASE(6) NNRT_ANCHOR_CASE(7) NNRT_ANCHOR_CASE(8)
	}
#undef NNRT_ANCHOR_CASE
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_warp_points_normals(const float* points, const float* normals, int64_t point_count, const float* nodes,
                                             const float* node_rotations, const float* node_translations, int32_t node_count,
                                             const int32_t* anchors, const float* anchor_weights, int32_t anchor_count,
                                             int32_t minimum_valid_anchor_count, const double* extrinsics_host, float* warped_points,
                                             float* warped_normals, void* stream) {
	NNRT_CHECK_ARG(points && nodes && node_rotations && node_translations && anchors && anchor_weights && warped_points,
	               "warp_points_normals: null pointer argument");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	(void) node_count;
	if (point_count == 0) return kOk;
	warp_points_kernel<<<blocks_for(point_count, 256), 256, 0, (cudaStream_t) stream>>>(
			points, normals, point_count, nodes, node_rotations, node_translations, anchors, anchor_weights, anchor_count,
			minimum_valid_anchor_count, rigid_from_host(extrinsics_host), warped_points, warped_normals);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_warp_points_normals_dual_quaternion(const float* points, const float* normals, int64_t point_count, const float* nodes,
                                                             const float* node_rotations, const float* node_translations, int32_t node_count,
                                                             const int32_t* anchors, const float* anchor_weights, int32_t anchor_count,
                                                             int32_t minimum_valid_anchor_count, const double* extrinsics_host, float* warped_points,
                                                             float* warped_normals, void* stream) {
	NNRT_CHECK_ARG(points && nodes && node_rotations && node_translations && anchors && anchor_weights && warped_points,
	               "warp_points_normals_dual_quaternion: null pointer argument");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	(void) node_count;
	if (point_count == 0) return kOk;
	warp_points_dqb_kernel<<<blocks_for(point_count, 256), 256, 0, (cudaStream_t) stream>>>(
			points, normals, point_count, nodes, node_rotations, node_translations, anchors, anchor_weights, anchor_count,
			minimum_valid_anchor_count, rigid_from_host(extrinsics_host), warped_points, warped_normals);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

static void invert_rigid_host(const double* T, double* out) {
	// Open3D t::geometry::InverseTransformation: [R^T | -R^T t]
	for (int i = 0; i < 3; i++) {
		for (int j = 0; j < 3; j++) out[i * 4 + j] = T[j * 4 + i];
		out[i * 4 + 3] = -(T[0 * 4 + i] * T[3] + T[1 * 4 + i] * T[7] + T[2 * 4 + i] * T[11]);
	}
	out[12] = out[13] = out[14] = 0; out[15] = 1;
}

extern "C" int nnrt_b200_unproject_depth(const void* depth, int32_t depth_dtype, int32_t rows, int32_t cols, const double* intrinsics_host,
                                         const double* extrinsics_host, float depth_scale, float depth_max, float* points, uint8_t* mask,
                                         void* stream) {
	NNRT_CHECK_ARG(depth && intrinsics_host && points, "unproject_depth: null pointer argument");
	NNRT_CHECK_ARG(depth_dtype == 0 || depth_dtype == 1, "depth dtype must be float32 (0) or uint16 (1)");
	double pose[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
	if (extrinsics_host) invert_rigid_host(extrinsics_host, pose);
	const int64_t n = (int64_t) rows * cols;
	if (n == 0) return kOk;
	const Intrinsics k = intrinsics_from_host(intrinsics_host);
	const Rigid p = rigid_from_host(pose);
	if (depth_dtype == 0)
		unproject_kernel<float><<<blocks_for(n, 256), 256, 0, (cudaStream_t) stream>>>((const float*) depth, rows, cols, k, p, depth_scale, depth_max, points, mask);
	else
		unproject_kernel<uint16_t><<<blocks_for(n, 256), 256, 0, (cudaStream_t) stream>>>((const uint16_t*) depth, rows, cols, k, p, depth_scale, depth_max, points, mask);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_point_to_plane_distances(const float* normals1, const float* vertices1, const float* vertices2, int64_t count,
                                                  float* distances, void* stream) {
	NNRT_CHECK_ARG(normals1 && vertices1 && vertices2 && distances, "point_to_plane_distances: null pointer argument");
	if (count == 0) return kOk;
	point_to_plane_kernel<<<blocks_for(count, 256), 256, 0, (cudaStream_t) stream>>>(normals1, vertices1, vertices2, count, distances);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_warped_surface_jacobians(const float* vertex_positions, const float* vertex_normals, int64_t vertex_count,
                                                  const float* nodes, const float* node_rotations, const int32_t* anchors,
                                                  const float* anchor_weights, int32_t anchor_count, float* position_jacobians,
                                                  float* normal_jacobians, void* stream) {
	NNRT_CHECK_ARG(vertex_positions && vertex_normals && nodes && node_rotations && anchors && anchor_weights && position_jacobians && normal_jacobians,
	               "warped_surface_jacobians: null pointer argument");
	if (vertex_count == 0) return kOk;
	warped_surface_jacobians_kernel<<<blocks_for(vertex_count * anchor_count, 256), 256, 0, (cudaStream_t) stream>>>(
			vertex_positions, vertex_normals, vertex_count, nodes, node_rotations, anchors, anchor_weights, anchor_count,
			position_jacobians, normal_jacobians);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
