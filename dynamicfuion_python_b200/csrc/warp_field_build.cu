// Warp-field construction on the device: median-grid subsampling of 3D points, brute-force K nearest
// points, node coverage weights, and the regularisation layers + edges of HierarchicalGraphWarpField.
//
// The reference builds these from an Open3D hash map (bin index = buffer index of the cell), atomic slot counters inside a bin
// (bin-sorted order = arrival order) and a kd-tree; the order of bins and of points inside a bin is therefore run-dependent in
// the reference and its tests compare SETS. Here every "first come" rule is resolved by ascending point index, which is the
// reference's single-thread CPU result and what the oracle restates, so all outputs are reproducible and bit-exact:
//   * bins are numbered by their smallest point index (an exclusive scan over "is the first point of its cell" flags);
//   * a bin's points are ranked by index before their distance sums are taken left to right in fp32;
//   * of equal sums the smallest index wins (64-bit atomicMin of sum bits << 32 | index = the reference's strict `<` scan);
//   * K nearest: replace-current-maximum over the indexed points in ascending order (the kd-tree only prunes).
// Distances are Eigen's (a - b).norm(): sqrt(x + (y + z)) with IEEE sqrt/div (-fmad=false, default -prec-sqrt/-prec-div).
#include "host_util.h"
#include "device_scan.cuh"
#include "../../include/nnrt_b200.h"
#include <vector>
#include <cfloat>
#include <type_traits>

namespace nnrt_b200 {
namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float distance3(const float* __restrict__ a, const float* __restrict__ b) {
	const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
	return sqrtf(sum3(dx * dx, dy * dy, dz * dz));
}

// The reference compares DISTANCES (sqrt of the squared norm). The IEEE square root costs several instructions and almost every
// candidate is far outside the current k-th distance, so candidates are first screened on the squared distance: above
// m * m * (1 + 1e-6) (which exceeds the exact m^2 whatever the two roundings do) the correctly rounded root cannot be below m, and
// the exact comparison `m > sqrtf(d2)` is only evaluated for the rest. Decisions are unchanged.
__device__ __forceinline__ float squared_upper_bound(float m) { return m * m * 1.000001f; }

// ---------------------------------------------------------------------------------------------------------------------------
// Median-grid subsampling (geometry/functional/kernel/GeometrySamplingGridBinning.h:30-52, GeometrySamplingMedian.h:36-285)

struct Cell { int x, y, z; };
__device__ __forceinline__ Cell cell_of(const float* __restrict__ p, float cell_size, float offset) {
	// Open3D: points / grid_cell_size + grid_offset (the same on all three axes: 0, or 0.5 for the second pass of the radius variants),
	// Floor(), To(Int32)
	return Cell{(int) floorf(__fdiv_rn(p[0], cell_size) + offset), (int) floorf(__fdiv_rn(p[1], cell_size) + offset),
	            (int) floorf(__fdiv_rn(p[2], cell_size) + offset)};
}
__device__ __forceinline__ unsigned hash_cell(Cell c) {
	unsigned h = (unsigned) c.x * 73856093u ^ (unsigned) c.y * 19349663u ^ (unsigned) c.z * 83492791u;
	h ^= h >> 15; h *= 0x2c1b3c6du; h ^= h >> 12;
	return h;
}

// Open-addressing table over cells. A slot is claimed by the first point that reaches it (owner = its index); later points compare
// their cell with the owner's, recomputed from the immutable point array, so no key is stored. slot_min = smallest index in the cell.
__global__ void __launch_bounds__(kThreads) bin_insert_kernel(const float* __restrict__ points, int64_t n, float cell_size, float offset, int* __restrict__ owner,
                                                              int* __restrict__ slot_min, unsigned mask, int* __restrict__ slot_of_point) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const Cell c = cell_of(points + i * 3, cell_size, offset);
	unsigned slot = hash_cell(c) & mask;
	while (true) {
		const int previous = atomicCAS(owner + slot, -1, (int) i);
		if (previous == -1) break;
		const Cell other = cell_of(points + (int64_t) previous * 3, cell_size, offset);
		if (other.x == c.x && other.y == c.y && other.z == c.z) break;
		slot = (slot + 1) & mask;
	}
	atomicMin(slot_min + slot, (int) i);
	slot_of_point[i] = (int) slot;
}

__global__ void __launch_bounds__(kThreads) flag_first_points_kernel(const int* __restrict__ slot_of_point, const int* __restrict__ slot_min, int64_t n,
                                                                     int* __restrict__ flags) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) flags[i] = slot_min[slot_of_point[i]] == (int) i;
}

// bin of a point = number of first points before its cell's first point
__global__ void __launch_bounds__(kThreads) assign_bins_kernel(const int* __restrict__ slot_of_point, const int* __restrict__ slot_min,
                                                               const int* __restrict__ first_point_rank, int64_t n, int* __restrict__ bin_of_point,
                                                               int* __restrict__ bin_counts) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int bin = first_point_rank[slot_min[slot_of_point[i]]];
	bin_of_point[i] = bin;
	atomicAdd(bin_counts + bin, 1);
}

__global__ void __launch_bounds__(kThreads) scatter_members_kernel(const int* __restrict__ bin_of_point, const int* __restrict__ bin_starts, int64_t n,
                                                                   int* __restrict__ bin_fill, int* __restrict__ members) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int bin = bin_of_point[i];
	members[bin_starts[bin] + atomicAdd(bin_fill + bin, 1)] = (int) i;
}

// arrival order -> ascending index order inside every bin. Light bins (<= kHeavyBin members): every point ranks itself by counting
// the smaller indices of its bin. Heavy bins: their first arrival enlists the bin, and one CTA per enlisted bin sorts its member list
// in shared memory (bitonic, up to kSortCapacity members; larger bins are ranked by counting with the CTA's threads). For the median
// the same CTA then stages the bin's positions in shared memory and forms every member's distance sum from there (ascending order,
// broadcast reads) instead of one thread per point gathering its whole bin from global memory: a 1200-point bin is 1.4 M distances.
constexpr int kHeavyBin = 64, kSortCapacity = 2048;

__global__ void __launch_bounds__(kThreads) rank_members_kernel(const int* __restrict__ bin_of_point, const int* __restrict__ bin_starts,
                                                                const int* __restrict__ members, int64_t n, int* __restrict__ sorted_members,
                                                                int* __restrict__ heavy_bins, int* __restrict__ heavy_count) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int bin = bin_of_point[i];
	const int begin = bin_starts[bin], end = bin_starts[bin + 1];
	if (end - begin > kHeavyBin) {
		if (members[begin] == (int) i) heavy_bins[atomicAdd(heavy_count, 1)] = bin;
		return;
	}
	int rank = 0;
	for (int k = begin; k < end; k++) rank += members[k] < (int) i;
	sorted_members[begin + rank] = (int) i;
}

__global__ void __launch_bounds__(kThreads) sort_heavy_bins_kernel(const int* __restrict__ heavy_bins, const int* __restrict__ heavy_count,
                                                                   const int* __restrict__ bin_starts, const int* __restrict__ members,
                                                                   int* __restrict__ sorted_members, const float* __restrict__ median_points,
                                                                   unsigned long long* __restrict__ best) {
	__shared__ int s_values[kSortCapacity];
	__shared__ float s_x[kSortCapacity], s_y[kSortCapacity], s_z[kSortCapacity];
	if ((int) blockIdx.x >= *heavy_count) return;
	const int bin = heavy_bins[blockIdx.x];
	const int begin = bin_starts[bin], m = bin_starts[bin + 1] - begin;
	if (m > kSortCapacity) {
		for (int k = threadIdx.x; k < m; k += blockDim.x) {
			const int value = members[begin + k];
			int rank = 0;
			for (int j = 0; j < m; j++) rank += members[begin + j] < value;
			sorted_members[begin + rank] = value;
		}
		return;  // its distance sums are left to median_sums_kernel
	}
	int padded = 1;
	while (padded < m) padded <<= 1;
	for (int k = threadIdx.x; k < padded; k += blockDim.x) s_values[k] = k < m ? members[begin + k] : 0x7fffffff;
	__syncthreads();
	for (int span = 2; span <= padded; span <<= 1) {
		for (int stride = span >> 1; stride > 0; stride >>= 1) {
			for (int t = threadIdx.x; t < padded; t += blockDim.x) {
				const int partner = t ^ stride;
				if (partner > t) {
					const int a = s_values[t], b = s_values[partner];
					if ((a > b) == ((t & span) == 0)) { s_values[t] = b; s_values[partner] = a; }
				}
			}
			__syncthreads();
		}
	}
	for (int k = threadIdx.x; k < m; k += blockDim.x) {
		const int i = s_values[k];
		sorted_members[begin + k] = i;
		if (median_points) { s_x[k] = median_points[(int64_t) i * 3]; s_y[k] = median_points[(int64_t) i * 3 + 1]; s_z[k] = median_points[(int64_t) i * 3 + 2]; }
	}
	if (!median_points) return;
	__syncthreads();
	unsigned long long mine = kEmptyKey;
	for (int k = threadIdx.x; k < m; k += blockDim.x) {
		const float sx = s_x[k], sy = s_y[k], sz = s_z[k];
		float sum = 0.f;
		for (int j = 0; j < m; j++) {
			const float dx = s_x[j] - sx, dy = s_y[j] - sy, dz = s_z[j] - sz;
			sum += sqrtf(sum3(dx * dx, dy * dy, dz * dz));
		}
		if (sum < FLT_MAX) mine = min(mine, ((unsigned long long) __float_as_uint(sum) << 32) | (unsigned) s_values[k]);
	}
#pragma unroll
	for (int offset = 16; offset > 0; offset >>= 1) mine = min(mine, __shfl_xor_sync(0xFFFFFFFFu, mine, offset));
	if ((threadIdx.x & 31) == 0 && mine != kEmptyKey) atomicMin(best + bin, mine);
}

__global__ void __launch_bounds__(kThreads) median_sums_kernel(const float* __restrict__ points, const int* __restrict__ bin_of_point,
                                                               const int* __restrict__ bin_starts, const int* __restrict__ sorted_members, int64_t n,
                                                               unsigned long long* __restrict__ best) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int bin = bin_of_point[i];
	const int begin = bin_starts[bin], end = bin_starts[bin + 1];
	if (end - begin > kHeavyBin && end - begin <= kSortCapacity) return;  // summed by the bin's own CTA (sort_heavy_bins_kernel)
	const float* source = points + i * 3;
	float sum = 0.f;
	for (int k = begin; k < end; k++) sum += distance3(points + (int64_t) sorted_members[k] * 3, source);
	if (sum < FLT_MAX)  // non-negative floats order like their bit patterns
		atomicMin(best + bin, ((unsigned long long) __float_as_uint(sum) << 32) | (unsigned) i);
}

__global__ void __launch_bounds__(kThreads) emit_sample_kernel(const unsigned long long* __restrict__ best, int bins, int64_t* __restrict__ sample) {
	const int b = blockIdx.x * blockDim.x + threadIdx.x;
	if (b < bins) sample[b] = best[b] == kEmptyKey ? -1 : (int64_t) (best[b] & 0xffffffffull);
}

struct SubsampleWorkspace {
	int *owner, *slot_min, *slot_of_point, *rank, *bin_of_point, *bin_counts, *bin_fill, *members, *sorted_members, *scan_scratch;
	unsigned long long* best;
	unsigned table_size;
	static size_t align(size_t bytes) { return (bytes + 255) & ~(size_t) 255; }
	static unsigned table_for(int64_t n) {
		unsigned t = 64;
		while ((int64_t) t < 2 * n) t <<= 1;
		return t;
	}
	static size_t bytes(int64_t n) {
		const size_t ints = (size_t) n + 2;
		return 2 * align(sizeof(int) * table_for(n)) + 7 * align(sizeof(int) * ints) + align(sizeof(int) * (size_t) scan_scratch_ints(n + 1)) +
		       align(sizeof(unsigned long long) * ints);
	}
	void carve(char* p, int64_t n) {
		const size_t ints = (size_t) n + 2;
		table_size = table_for(n);
		auto take = [&p](size_t b) { char* r = p; p += align(b); return r; };
		// [owner | best] are filled with 0xFF and [bin_counts | bin_fill] with zeros by one memset each (sizes are multiples of 256 bytes)
		owner = (int*) take(sizeof(int) * table_size);
		best = (unsigned long long*) take(sizeof(unsigned long long) * ints);
		slot_min = (int*) take(sizeof(int) * table_size);
		bin_counts = (int*) take(sizeof(int) * ints);
		bin_fill = (int*) take(sizeof(int) * ints);
		slot_of_point = (int*) take(sizeof(int) * ints);
		rank = (int*) take(sizeof(int) * ints);
		bin_of_point = (int*) take(sizeof(int) * ints);
		members = (int*) take(sizeof(int) * ints);
		sorted_members = (int*) take(sizeof(int) * ints);
		scan_scratch = (int*) take(sizeof(int) * (size_t) scan_scratch_ints(n + 1));
	}
};

// Bins the points into grid cells and leaves every bin's members in ascending point order: bin_of_point, bin_counts (= CSR starts,
// n + 2 entries) and sorted_members are valid on the stream afterwards; *bin_count_host after its ONE host synchronisation.
int bin_points_sorted(const float* points, int64_t n, float cell_size, float offset, SubsampleWorkspace& w, int* bin_count_host, bool median_sums,
                      cudaStream_t s) {
	const unsigned blocks = blocks_for(n, kThreads);
	const size_t ints = (size_t) n + 2;
	NNRT_CUDA(cudaMemsetAsync(w.owner, 0xFF, SubsampleWorkspace::align(sizeof(int) * w.table_size) + sizeof(unsigned long long) * ints, s));  // owner, best
	NNRT_CUDA(cudaMemsetAsync(w.slot_min, 0x7F, sizeof(int) * w.table_size, s));
	NNRT_CUDA(cudaMemsetAsync(w.bin_counts, 0, SubsampleWorkspace::align(sizeof(int) * ints) + sizeof(int) * ints, s));  // bin_counts, bin_fill
	bin_insert_kernel<<<blocks, kThreads, 0, s>>>(points, n, cell_size, offset, w.owner, w.slot_min, w.table_size - 1, w.slot_of_point);
	flag_first_points_kernel<<<blocks, kThreads, 0, s>>>(w.slot_of_point, w.slot_min, n, w.rank);
	NNRT_LAUNCH_CHECK();
	int rc = exclusive_scan(w.rank, w.rank, n, w.scan_scratch, s);  // rank[n] = bin count
	if (rc != kOk) return rc;
	NNRT_CUDA(cudaMemcpyAsync(bin_count_host, w.rank + n, sizeof(int), cudaMemcpyDeviceToHost, s));
	assign_bins_kernel<<<blocks, kThreads, 0, s>>>(w.slot_of_point, w.slot_min, w.rank, n, w.bin_of_point, w.bin_counts);
	NNRT_LAUNCH_CHECK();
	// starts over n + 1 counts (bins beyond the bin count are empty), so bin_starts[bin + 1] is always defined
	rc = exclusive_scan(w.bin_counts, w.bin_counts, n + 1, w.scan_scratch, s);
	if (rc != kOk) return rc;
	scatter_members_kernel<<<blocks, kThreads, 0, s>>>(w.bin_of_point, w.bin_counts, n, w.bin_fill, w.members);
	// the first-point ranks are consumed (assign_bins) and so is the fill counter array: reused as the heavy-bin list and its counter
	// (bin_fill[n + 1] was zeroed with the rest and no bin id reaches it)
	int* heavy_bins = w.rank;
	int* heavy_count = w.bin_fill + n + 1;
	rank_members_kernel<<<blocks, kThreads, 0, s>>>(w.bin_of_point, w.bin_counts, w.members, n, w.sorted_members, heavy_bins, heavy_count);
	sort_heavy_bins_kernel<<<(unsigned) (n / (kHeavyBin + 1) + 1), kThreads, 0, s>>>(heavy_bins, heavy_count, w.bin_counts, w.members, w.sorted_members,
	                                                                               median_sums ? points : nullptr, w.best);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

// Enqueues the whole median subsample on `s`; sample[0 .. *bin_count) is valid once the stream has passed.
int median_subsample(const float* points, int64_t n, float cell_size, float offset, SubsampleWorkspace& w, int64_t* sample, int* bin_count_host,
                     cudaStream_t s) {
	const int rc = bin_points_sorted(points, n, cell_size, offset, w, bin_count_host, true, s);
	if (rc != kOk) return rc;
	median_sums_kernel<<<blocks_for(n, kThreads), kThreads, 0, s>>>(points, w.bin_of_point, w.bin_counts, w.sorted_members, n, w.best);
	NNRT_LAUNCH_CHECK();
	NNRT_CUDA(cudaStreamSynchronize(s));
	emit_sample_kernel<<<blocks_for(*bin_count_host, kThreads), kThreads, 0, s>>>(w.best, *bin_count_host, sample);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

// Mean of every bin's points (geometry/functional/kernel/GeometrySamplingMean.h:37-92): the reference adds with float atomics; here
// the members are added in ascending point order (its single-thread result), then divided by the count. One thread per bin.
__global__ void __launch_bounds__(kThreads) bin_means_kernel(const float* __restrict__ points, const int* __restrict__ bin_starts,
                                                             const int* __restrict__ sorted_members, int bins, float* __restrict__ means) {
	const int b = blockIdx.x * blockDim.x + threadIdx.x;
	if (b >= bins) return;
	const int begin = bin_starts[b], end = bin_starts[b + 1];
	float x = 0.f, y = 0.f, z = 0.f;
	for (int k = begin; k < end; k++) {
		const int64_t i = sorted_members[k];
		x += points[i * 3]; y += points[i * 3 + 1]; z += points[i * 3 + 2];
	}
	const float count = (float) (end - begin);
	means[b * 3] = __fdiv_rn(x, count); means[b * 3 + 1] = __fdiv_rn(y, count); means[b * 3 + 2] = __fdiv_rn(z, count);
}

// FindClosestInBins (GeometrySamplingMean.h:189-212): per bin the point with the smallest squared distance to the bin mean, the
// smallest index among equals (strict `>` in the reference's compare-and-swap loop)
__global__ void __launch_bounds__(kThreads) closest_to_mean_kernel(const float* __restrict__ points, const int* __restrict__ bin_of_point,
                                                                   const float* __restrict__ means, int64_t n, unsigned long long* __restrict__ best) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int b = bin_of_point[i];
	const float dx = means[b * 3] - points[i * 3], dy = means[b * 3 + 1] - points[i * 3 + 1], dz = means[b * 3 + 2] - points[i * 3 + 2];
	const float d2 = sum3(dx * dx, dy * dy, dz * dz);
	atomicMin(best + b, ((unsigned long long) __float_as_uint(d2) << 32) | (unsigned) i);
}

// ---------------------------------------------------------------------------------------------------------------------------
// K nearest indexed points per query (core/kernel/SearchKdTreeImpl.h:118-146, PLAIN strategy) — thread per query, indexed points
// broadcast from shared memory in tiles. Slot order is the replace-current-maximum order; `descending` additionally sorts the slots
// by index, largest first, -1 last (core::functional::SortTensorAlongLastDimension(.., true, DESC)).
template<int K>
__device__ __forceinline__ void k_nearest(const float* __restrict__ indexed, int indexed_count, const float* query, bool active, float4* tile,
                                          int tile_capacity, int (&idx)[K], float (&dist)[K]) {
#pragma unroll
	for (int a = 0; a < K; a++) { idx[a] = -1; dist[a] = INFINITY; }
	int max_at = 0;
	float max_distance = INFINITY, skip_above = INFINITY;
	for (int base = 0; base < indexed_count; base += tile_capacity) {
		const int count = min(tile_capacity, indexed_count - base);
		__syncthreads();
		for (int k = threadIdx.x; k < count; k += blockDim.x)
			tile[k] = make_float4(indexed[(base + k) * 3], indexed[(base + k) * 3 + 1], indexed[(base + k) * 3 + 2], 0.f);
		__syncthreads();
		if (!active) continue;
		for (int k = 0; k < count; k++) {
			const float4 g = tile[k];
			const float dx = g.x - query[0], dy = g.y - query[1], dz = g.z - query[2];
			const float d2 = sum3(dx * dx, dy * dy, dz * dz);
			if (d2 > skip_above) continue;  // sqrt(d2) >= max_distance for certain: the exact test below would fail
			const float d = sqrtf(d2);
			if (max_distance > d) {
#pragma unroll
				for (int a = 0; a < K; a++)
					if (a == max_at) { dist[a] = d; idx[a] = base + k; }
				max_at = 0;
				max_distance = dist[0];
#pragma unroll
				for (int a = 1; a < K; a++)
					if (dist[a] > max_distance) { max_at = a; max_distance = dist[a]; }
				skip_above = squared_upper_bound(max_distance);
			}
		}
	}
}

constexpr int kKnnTile = 1024;

template<int K>
__global__ void __launch_bounds__(kThreads) find_k_nearest_kernel(const float* __restrict__ indexed, int indexed_count, const float* __restrict__ queries,
                                                                  int64_t query_count, bool sort_by_distance, int32_t* __restrict__ out_indices,
                                                                  float* __restrict__ out_distances) {
	__shared__ float4 tile[kKnnTile];
	const int64_t q = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	const bool active = q < query_count;
	float query[3] = {0.f, 0.f, 0.f};
	if (active) { query[0] = queries[q * 3]; query[1] = queries[q * 3 + 1]; query[2] = queries[q * 3 + 2]; }
	int idx[K];
	float dist[K];
	k_nearest<K>(indexed, indexed_count, query, active, tile, kKnnTile, idx, dist);
	if (!active) return;
	if (sort_by_distance) {  // SortTensorAlongLastDimensionByKey(indices, distances, false, ASC), core/KdTree.cpp:71-73 (stable insertion sort)
#pragma unroll
		for (int a = 1; a < K; a++)
#pragma unroll
			for (int b = a; b > 0; b--)
				if (dist[b] < dist[b - 1]) {
					const float d = dist[b]; dist[b] = dist[b - 1]; dist[b - 1] = d;
					const int t = idx[b]; idx[b] = idx[b - 1]; idx[b - 1] = t;
				}
	}
#pragma unroll
	for (int a = 0; a < K; a++) {
		out_indices[q * K + a] = idx[a];
		if (out_distances) out_distances[q * K + a] = dist[a];
	}
}

// One finer-layer node per WARP: every lane keeps the K nearest of its stride of the coarser nodes (same replace-current-maximum rule,
// indices ascending within a lane), then K rounds of a warp-wide minimum over (distance, index) pick the K nearest overall. That set
// equals the sequential rule's: a later candidate replaces only when strictly closer, so ties go to the smaller index. The set is then
// sorted by index, largest first, and written straight into the edge list (geometry/kernel/HierarchicalGraphWarpFieldImpl.h:218-296
// with flip_source_order: row S-1-i; every row holds the same number min(K, coarser count) of valid targets, so dropping the -1 slots
// leaves a dense [S, valid] table and needs no compaction pass). A thread per node left 126 of 148 SMs idle at 5000 nodes.
constexpr int kEdgeWarps = kThreads / 32;

template<int K>
__global__ void __launch_bounds__(kThreads) layer_edges_kernel(const float* __restrict__ coarser, int coarser_count, const float* __restrict__ finer,
                                                               int finer_count, int first_virtual_finer, int first_virtual_coarser, int layer,
                                                               int32_t* __restrict__ edges, int8_t* __restrict__ edge_layers) {
	__shared__ float4 tile[kKnnTile];
	const int lane = threadIdx.x & 31;
	const int i = blockIdx.x * kEdgeWarps + (threadIdx.x >> 5);
	const bool active = i < finer_count;
	float query[3] = {0.f, 0.f, 0.f};
	if (active) { query[0] = finer[i * 3]; query[1] = finer[i * 3 + 1]; query[2] = finer[i * 3 + 2]; }
	int idx[K];
	float dist[K];
#pragma unroll
	for (int a = 0; a < K; a++) { idx[a] = -1; dist[a] = INFINITY; }
	int max_at = 0;
	float max_distance = INFINITY, skip_above = INFINITY;
	for (int base = 0; base < coarser_count; base += kKnnTile) {
		const int count = min(kKnnTile, coarser_count - base);
		__syncthreads();
		for (int k = threadIdx.x; k < count; k += blockDim.x)
			tile[k] = make_float4(coarser[(base + k) * 3], coarser[(base + k) * 3 + 1], coarser[(base + k) * 3 + 2], 0.f);
		__syncthreads();
		if (!active) continue;
		for (int k = lane; k < count; k += 32) {
			const float4 g = tile[k];
			const float dx = g.x - query[0], dy = g.y - query[1], dz = g.z - query[2];
			const float d2 = sum3(dx * dx, dy * dy, dz * dz);
			if (d2 > skip_above) continue;  // sqrt(d2) >= max_distance for certain: the exact test below would fail
			const float d = sqrtf(d2);
			if (max_distance > d) {
#pragma unroll
				for (int a = 0; a < K; a++)
					if (a == max_at) { dist[a] = d; idx[a] = base + k; }
				max_at = 0;
				max_distance = dist[0];
#pragma unroll
				for (int a = 1; a < K; a++)
					if (dist[a] > max_distance) { max_at = a; max_distance = dist[a]; }
				skip_above = squared_upper_bound(max_distance);
			}
		}
	}
	if (!active) return;  // whole warps leave together
	// (distance bits, index) keys order like (distance, index): distances are non-negative, empty slots (inf, -1) sort last
	unsigned long long keys[K];
#pragma unroll
	for (int a = 0; a < K; a++) keys[a] = ((unsigned long long) __float_as_uint(dist[a]) << 32) | (unsigned) idx[a];
	int chosen[K];
#pragma unroll
	for (int round = 0; round < K; round++) {
		unsigned long long mine = keys[0];
#pragma unroll
		for (int a = 1; a < K; a++) mine = min(mine, keys[a]);
		unsigned long long best = mine;
#pragma unroll
		for (int offset = 16; offset > 0; offset >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, offset));
		chosen[round] = (int) (unsigned) (best & 0xffffffffull);  // -1 when fewer than K coarser nodes exist
		if (mine == best && (unsigned) (best & 0xffffffffull) != 0xffffffffu) {  // indices are unique: exactly one lane owns the key
#pragma unroll
			for (int a = 0; a < K; a++)
				if (keys[a] == best) keys[a] = ~0ull;
		}
	}
	if (lane != 0) return;
#pragma unroll
	for (int a = 1; a < K; a++)  // insertion sort, descending (-1 sinks to the end)
#pragma unroll
		for (int b = a; b > 0; b--)
			if (chosen[b] > chosen[b - 1]) { const int t = chosen[b]; chosen[b] = chosen[b - 1]; chosen[b - 1] = t; }
	const int valid = min(K, coarser_count);
	const int64_t row = (int64_t) (finer_count - 1 - i) * valid;
#pragma unroll
	for (int a = 0; a < K; a++) {
		if (a >= valid) break;
		edges[(row + a) * 2] = first_virtual_finer + i;
		edges[(row + a) * 2 + 1] = first_virtual_coarser + chosen[a];
		edge_layers[row + a] = (int8_t) layer;
	}
}

// geometry/WarpField.cpp:249-262: squared distance to the nearest other node
__global__ void __launch_bounds__(kThreads) node_coverage_weights_kernel(const float* __restrict__ nodes, int node_count, float* __restrict__ weights) {
	__shared__ float4 tile[kKnnTile];
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	const bool active = i < node_count;
	float p[3] = {0.f, 0.f, 0.f};
	if (active) { p[0] = nodes[i * 3]; p[1] = nodes[i * 3 + 1]; p[2] = nodes[i * 3 + 2]; }
	float nearest_d2 = INFINITY;
	for (int base = 0; base < node_count; base += kKnnTile) {
		const int count = min(kKnnTile, node_count - base);
		__syncthreads();
		for (int k = threadIdx.x; k < count; k += blockDim.x)
			tile[k] = make_float4(nodes[(base + k) * 3], nodes[(base + k) * 3 + 1], nodes[(base + k) * 3 + 2], 0.f);
		__syncthreads();
		if (!active) continue;
		for (int k = 0; k < count; k++) {
			if (base + k == i) continue;
			const float4 g = tile[k];
			const float dx = g.x - p[0], dy = g.y - p[1], dz = g.z - p[2];
			nearest_d2 = fminf(nearest_d2, sum3(dx * dx, dy * dy, dz * dz));  // sqrt is monotone: the nearest node has the smallest d2
		}
	}
	if (active) { const float nearest = sqrtf(nearest_d2); weights[i] = nearest * nearest; }
}

// --- layer bookkeeping -----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) iota_kernel(int64_t* __restrict__ out, int n) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = i;
}
// coarser layer = the sampled entries of the finer layer; marks them for removal from the finer layer
__global__ void __launch_bounds__(kThreads) promote_sample_kernel(const int64_t* __restrict__ sample, int sample_count, const int64_t* __restrict__ finer_indices,
                                                                  const float* __restrict__ finer_positions, int64_t* __restrict__ coarser_indices,
                                                                  float* __restrict__ coarser_positions, int* __restrict__ removed) {
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= sample_count) return;
	const int64_t j = sample[s];
	coarser_indices[s] = finer_indices[j];
	coarser_positions[s * 3] = finer_positions[j * 3];
	coarser_positions[s * 3 + 1] = finer_positions[j * 3 + 1];
	coarser_positions[s * 3 + 2] = finer_positions[j * 3 + 2];
	removed[j] = 1;
}
// stable removal: a kept entry moves left by the number of removed entries before it
__global__ void __launch_bounds__(kThreads) filter_layer_kernel(const int* __restrict__ removed_before, const int64_t* __restrict__ indices,
                                                                const float* __restrict__ positions, int n, int64_t* __restrict__ out_indices,
                                                                float* __restrict__ out_positions) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (removed_before[i + 1] != removed_before[i]) return;
	const int dst = i - removed_before[i];
	out_indices[dst] = indices[i];
	out_positions[dst * 3] = positions[i * 3];
	out_positions[dst * 3 + 1] = positions[i * 3 + 1];
	out_positions[dst * 3 + 2] = positions[i * 3 + 2];
}

template<typename Launch>
int dispatch_k(int k, Launch&& launch) {
	switch (k) {
		case 1: launch(std::integral_constant<int, 1>()); break;
		case 2: launch(std::integral_constant<int, 2>()); break;
		case 3: launch(std::integral_constant<int, 3>()); break;
		case 4: launch(std::integral_constant<int, 4>()); break;
		case 5: launch(std::integral_constant<int, 5>()); break;
		case 6: launch(std::integral_constant<int, 6>()); break;
		case 7: launch(std::integral_constant<int, 7>()); break;
		case 8: launch(std::integral_constant<int, 8>()); break;
		default: set_last_error("neighbour count must be in [1, 8], got %d", k); return kInvalidArgument;
	}
	return kOk;
}

}  // namespace
}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_median_grid_subsample_3d_points(const float* points, int64_t point_count, float grid_cell_size, float grid_offset,
                                                         int64_t* sample, int64_t* sample_count_host, void* stream) {
	NNRT_CHECK_ARG(sample_count_host, "sample_count_host must not be NULL");
	*sample_count_host = 0;
	if (point_count == 0) return kOk;
	NNRT_CHECK_ARG(points && sample, "points and sample must not be NULL");
	NNRT_CHECK_ARG(point_count > 0 && point_count < (1ll << 30), "point_count must be in [0, 2^30), got %lld", (long long) point_count);
	NNRT_CHECK_ARG(grid_cell_size > 0.f, "grid_cell_size must be positive, got %f", grid_cell_size);
	cudaStream_t s = (cudaStream_t) stream;
	char* buffer = nullptr;
	Scratch scratch(s);
	NNRT_CUDA(scratch.alloc(&buffer, SubsampleWorkspace::bytes(point_count)));
	SubsampleWorkspace w;
	w.carve(buffer, point_count);
	int bins = 0;
	const int rc = median_subsample(points, point_count, grid_cell_size, grid_offset, w, sample, &bins, s);
	if (rc != kOk) return rc;
	*sample_count_host = bins;
	return kOk;
}

// mode 0: means [bins,3] f32 into `out`; mode 1: indices of the points closest to their bin's mean, int64, into `out`
static int grid_mean_impl(int mode, const float* points, int64_t point_count, float grid_cell_size, float grid_offset, void* out, int64_t* count_host,
                          void* stream) {
	NNRT_CHECK_ARG(count_host, "count_host must not be NULL");
	*count_host = 0;
	if (point_count == 0) return kOk;
	NNRT_CHECK_ARG(points && out, "points and the output must not be NULL");
	NNRT_CHECK_ARG(point_count > 0 && point_count < (1ll << 30), "point_count must be in [0, 2^30), got %lld", (long long) point_count);
	NNRT_CHECK_ARG(grid_cell_size > 0.f, "grid_cell_size must be positive, got %f", grid_cell_size);
	cudaStream_t s = (cudaStream_t) stream;
	char* buffer = nullptr;
	const size_t mean_bytes = mode == 1 ? SubsampleWorkspace::align(sizeof(float) * 3 * (size_t) point_count) : 0;
	Scratch scratch(s);
	NNRT_CUDA(scratch.alloc(&buffer, SubsampleWorkspace::bytes(point_count) + mean_bytes));
	SubsampleWorkspace w;
	w.carve(buffer, point_count);
	float* means = mode == 1 ? (float*) (buffer + SubsampleWorkspace::bytes(point_count)) : (float*) out;
	int bins = 0;
	int rc = bin_points_sorted(points, point_count, grid_cell_size, grid_offset, w, &bins, false, s);
	if (rc == kOk && cudaStreamSynchronize(s) != cudaSuccess) rc = kCudaError;
	if (rc == kOk) {
		bin_means_kernel<<<blocks_for(bins, kThreads), kThreads, 0, s>>>(points, w.bin_counts, w.sorted_members, bins, means);
		if (mode == 1) {
			closest_to_mean_kernel<<<blocks_for(point_count, kThreads), kThreads, 0, s>>>(points, w.bin_of_point, means, point_count, w.best);
			emit_sample_kernel<<<blocks_for(bins, kThreads), kThreads, 0, s>>>(w.best, bins, (int64_t*) out);
		}
		if (cudaGetLastError() != cudaSuccess) { set_last_error("grid mean: kernel launch failed"); rc = kCudaError; }
	}
	if (rc != kOk) return rc;
	*count_host = bins;
	return kOk;
}

extern "C" int nnrt_b200_mean_grid_downsample_3d_points(const float* points, int64_t point_count, float grid_cell_size, float grid_offset,
                                                        float* downsampled_points, int64_t* downsampled_count_host, void* stream) {
	return grid_mean_impl(0, points, point_count, grid_cell_size, grid_offset, downsampled_points, downsampled_count_host, stream);
}

extern "C" int nnrt_b200_closest_to_mean_grid_subsample_3d_points(const float* points, int64_t point_count, float grid_cell_size, float grid_offset,
                                                                  int64_t* sample, int64_t* sample_count_host, void* stream) {
	return grid_mean_impl(1, points, point_count, grid_cell_size, grid_offset, sample, sample_count_host, stream);
}

extern "C" int nnrt_b200_find_k_nearest_to_points(const float* indexed_points, int32_t indexed_count, const float* query_points, int64_t query_count,
                                                  int32_t k, int32_t sort_output, int32_t* nearest_indices, float* nearest_distances, void* stream) {
	if (query_count == 0) return kOk;
	NNRT_CHECK_ARG(indexed_points && query_points && nearest_indices, "indexed_points, query_points and nearest_indices must not be NULL");
	NNRT_CHECK_ARG(indexed_count > 0 && query_count > 0, "indexed_count and query_count must be positive");
	cudaStream_t s = (cudaStream_t) stream;
	const int rc = dispatch_k(k, [&](auto K) {
		find_k_nearest_kernel<decltype(K)::value><<<blocks_for(query_count, kThreads), kThreads, 0, s>>>(indexed_points, indexed_count, query_points, query_count,
		                                                                                                 sort_output != 0, nearest_indices, nearest_distances);
	});
	if (rc != kOk) return rc;
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_node_coverage_weights(const float* nodes, int32_t node_count, float node_coverage, float* node_coverage_weights, void* stream) {
	NNRT_CHECK_ARG(nodes && node_coverage_weights && node_count > 0, "nodes / node_coverage_weights must not be NULL and node_count positive");
	cudaStream_t s = (cudaStream_t) stream;
	if (node_count == 1) {
		NNRT_CUDA(cudaMemcpyAsync(node_coverage_weights, &node_coverage, sizeof(float), cudaMemcpyHostToDevice, s));
		NNRT_CUDA(cudaStreamSynchronize(s));  // the source is a stack variable
		return kOk;
	}
	node_coverage_weights_kernel<<<blocks_for(node_count, kThreads), kThreads, 0, s>>>(nodes, node_count, node_coverage_weights);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_build_regularization_layers(const float* nodes, int32_t node_count, const float* layer_decimation_radii_host, int32_t layer_count,
                                                     int32_t max_vertex_degree, int64_t* node_indices, int32_t* layer_node_counts_host, int32_t* edges,
                                                     int8_t* edge_layer_indices, int64_t* edge_count_host, void* stream) {
	NNRT_CHECK_ARG(nodes && node_indices && layer_node_counts_host && edge_count_host && layer_decimation_radii_host, "NULL argument");
	NNRT_CHECK_ARG(node_count > 0, "node_count must be positive, got %d", node_count);
	NNRT_CHECK_ARG(layer_count >= 1, "layer_count must be a positive integer, got %d.", layer_count);
	NNRT_CHECK_ARG(layer_count == 1 || (edges && edge_layer_indices), "edges / edge_layer_indices must not be NULL when layer_count > 1");
	NNRT_CHECK_ARG(layer_count == 1 || (max_vertex_degree >= 1 && max_vertex_degree <= 8), "max_vertex_degree must be in [1, 8], got %d", max_vertex_degree);
	cudaStream_t s = (cudaStream_t) stream;
	const int n = node_count, L = layer_count;
	// per layer: indices [n] i64 + positions [n,3] f32, twice (slot l = the layer as sampled, slot L + l = after its coarser layer's nodes were
	// filtered out); removed flags + scan; sample; subsample workspace
	const size_t idx_bytes = SubsampleWorkspace::align(sizeof(int64_t) * (size_t) n), pos_bytes = SubsampleWorkspace::align(sizeof(float) * 3 * (size_t) n);
	const size_t flag_bytes = SubsampleWorkspace::align(sizeof(int) * ((size_t) n + 2));
	const size_t total = (size_t) (2 * L) * (idx_bytes + pos_bytes) + flag_bytes + idx_bytes + SubsampleWorkspace::bytes(n);
	char* buffer = nullptr;
	Scratch scratch(s);
	NNRT_CUDA(scratch.alloc(&buffer, total));
	char* p = buffer;
	std::vector<int64_t*> layer_indices(2 * L);
	std::vector<float*> layer_positions(2 * L);
	for (int l = 0; l < 2 * L; l++) {
		layer_indices[l] = (int64_t*) p; p += idx_bytes;
		layer_positions[l] = (float*) p; p += pos_bytes;
	}
	int* removed = (int*) p; p += flag_bytes;
	int64_t* sample = (int64_t*) p; p += idx_bytes;
	SubsampleWorkspace w;
	std::vector<int> counts(L, 0);
	int rc = kOk;
	auto fail = [&](int code) { return code; };  // `scratch` returns the buffer
#define NNRT_LAYERS_CUDA(call)                                                                                                      \
	do {                                                                                                                            \
		cudaError_t err__ = (call);                                                                                                 \
		if (err__ != cudaSuccess) { set_last_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(err__), __FILE__, __LINE__); return fail(kCudaError); } \
	} while (0)

	// finest layer: all nodes (geometry/HierarchicalGraphWarpField.cpp:77-81)
	iota_kernel<<<blocks_for(n, kThreads), kThreads, 0, s>>>(layer_indices[0], n);
	NNRT_LAYERS_CUDA(cudaMemcpyAsync(layer_positions[0], nodes, sizeof(float) * 3 * (size_t) n, cudaMemcpyDeviceToDevice, s));
	counts[0] = n;
	std::vector<int> slot_of_layer(L);
	slot_of_layer[0] = 0;
	for (int l = 1; l < L; l++) {
		const int finer_count = counts[l - 1];
		const int finer = l - 1, filtered = L + l - 1, coarser = l;
		// median-grid-subsample the finer layer with cell = 2 * radius (:94)
		w.carve(p, finer_count);
		int sample_count = 0;
		const float cell = layer_decimation_radii_host[l] * 2;
		if (!(cell > 0.f)) { set_last_error("layer %d: decimation radius must be positive, got %f", l, layer_decimation_radii_host[l]); return fail(kInvalidArgument); }
		rc = median_subsample(layer_positions[finer], finer_count, cell, 0.f, w, sample, &sample_count, s);
		if (rc != kOk) return fail(rc);
		if (sample_count == finer_count) {
			set_last_error("Attempting to generate a coarser layer of size %d out of a finer layer of the same size, which is not possible."
			               "Please either reduce the layer count, or, if applicable, increase the node_coverage parameter or "
			               "change the compute_layer_decimation_radius passed-in function to decrease the node density in the finer layer.", sample_count);
			return fail(kInvalidArgument);
		}
		NNRT_LAYERS_CUDA(cudaMemsetAsync(removed, 0, sizeof(int) * ((size_t) finer_count + 1), s));
		promote_sample_kernel<<<blocks_for(sample_count, kThreads), kThreads, 0, s>>>(sample, sample_count, layer_indices[finer], layer_positions[finer],
		                                                                             layer_indices[coarser], layer_positions[coarser], removed);
		NNRT_LAYERS_CUDA(cudaGetLastError());
		rc = exclusive_scan(removed, removed, finer_count, w.scan_scratch, s);
		if (rc != kOk) return fail(rc);
		filter_layer_kernel<<<blocks_for(finer_count, kThreads), kThreads, 0, s>>>(removed, layer_indices[finer], layer_positions[finer], finer_count,
		                                                                         layer_indices[filtered], layer_positions[filtered]);
		NNRT_LAYERS_CUDA(cudaGetLastError());
		slot_of_layer[l - 1] = filtered;
		slot_of_layer[l] = coarser;
		counts[l - 1] = finer_count - sample_count;
		counts[l] = sample_count;
	}
	// virtual order: layers fine to coarse (:112-124, 179-181)
	std::vector<int> first_virtual(L);
	int virtual_count = 0;
	for (int l = 0; l < L; l++) {
		first_virtual[l] = virtual_count;
		layer_node_counts_host[l] = counts[l];
		NNRT_LAYERS_CUDA(cudaMemcpyAsync(node_indices + virtual_count, layer_indices[slot_of_layer[l]], sizeof(int64_t) * (size_t) counts[l], cudaMemcpyDeviceToDevice, s));
		virtual_count += counts[l];
	}
	// edges, coarsest pair first (:130-190)
	int64_t edge_count = 0;
	for (int l = L - 1; l >= 1; l--) {
		const int finer_count = counts[l - 1], coarser_count = counts[l];
		const int32_t first_finer = first_virtual[l - 1], first_coarser = first_virtual[l];
		const float* coarser_positions = layer_positions[slot_of_layer[l]];
		const float* finer_positions = layer_positions[slot_of_layer[l - 1]];
		int32_t* edges_out = edges + edge_count * 2;
		int8_t* layers_out = edge_layer_indices + edge_count;
		dispatch_k(max_vertex_degree, [&](auto K) {
			layer_edges_kernel<decltype(K)::value><<<blocks_for(finer_count, kEdgeWarps), kThreads, 0, s>>>(coarser_positions, coarser_count, finer_positions, finer_count,
			                                                                                            first_finer, first_coarser, l, edges_out, layers_out);
		});
		NNRT_LAYERS_CUDA(cudaGetLastError());
		edge_count += (int64_t) finer_count * (max_vertex_degree < coarser_count ? max_vertex_degree : coarser_count);
	}
	NNRT_LAYERS_CUDA(cudaGetLastError());
#undef NNRT_LAYERS_CUDA
	*edge_count_host = edge_count;
	return kOk;
}
