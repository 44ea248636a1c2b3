// Triangle-mesh extraction from TSDF voxel blocks: the canonical mesh the fitter consumes is produced by
// nnrt::geometry::VoxelBlockGrid::ExtractTriangleMesh (cpp/geometry/VoxelBlockGrid.cpp:461-484), which hands the work to Open3D's
// t::geometry::kernel::voxel_grid::ExtractTriangleMesh (Open3D v0.17.0-1fix6008, third party: not in the reference checkout) over
// Open3D's device hash map. Both are restated here from the published algorithm so integrate -> extract -> fit stays on the device:
//   * block hash: an open-addressing table over the packed block keys, built per call, answers the 27-neighbourhood of every block
//     (what VoxelBlockGrid::BufferRadiusNeighbors provides: nnrt_b200_voxel_block_neighbors);
//   * marching cubes in Open3D's four passes (cube configuration per voxel, one vertex per sign-change edge owned by the voxel at
//     the edge's origin, interpolated position / TSDF-gradient normal / colour, triangles through the per-edge vertex ids).
// Open3D numbers vertices and triangles with atomic counters (run-dependent order) and flags edges by scattering into neighbour voxels;
// here every voxel decides its own three edges (an edge carries a vertex iff its end points differ in sign and one of the up to four
// cubes around it is complete) and the numbering comes from exclusive scans in (block, voxel, edge) order: reproducible, same sets.
// The lookup tables are derived, not copied (tools/generate_marching_cubes_tables.py): the triangulation of a configuration may differ
// from Lorensen's, the vertex set cannot. PARITY UNPINNED by the reference: its extraction tests need downloaded volumes and Open3D.
#include "host_util.h"
#include "device_scan.cuh"
#include "marching_cubes_tables.h"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {
namespace {

constexpr int kThreads = 256;
constexpr unsigned long long kEmptyKey = ~0ull;

__device__ __constant__ int8_t c_triangle_table[256][16];
__device__ __constant__ int8_t c_triangle_count[256];
__device__ __constant__ int c_vertex_shifts[8][3];
__device__ __constant__ int c_edge_shifts[12][4];

__host__ __device__ __forceinline__ unsigned long long pack_key(int x, int y, int z) {
	return ((unsigned long long) ((x + (1 << 20)) & 0x1FFFFF) << 42) | ((unsigned long long) ((y + (1 << 20)) & 0x1FFFFF) << 21) |
	       (unsigned long long) ((z + (1 << 20)) & 0x1FFFFF);
}
__device__ __forceinline__ unsigned hash_slot(unsigned long long key, unsigned mask) {
	key ^= key >> 33; key *= 0xff51afd7ed558ccdull; key ^= key >> 33; key *= 0xc4ceb9fe1a85ec53ull; key ^= key >> 33;
	return (unsigned) key & mask;
}

__global__ void __launch_bounds__(kThreads) hash_insert_kernel(const int32_t* __restrict__ keys, int64_t count, unsigned long long* __restrict__ table_keys,
                                                               int* __restrict__ table_values, unsigned mask) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	const unsigned long long key = pack_key(keys[i * 3], keys[i * 3 + 1], keys[i * 3 + 2]);
	for (unsigned slot = hash_slot(key, mask);; slot = (slot + 1) & mask) {
		const unsigned long long old = atomicCAS(table_keys + slot, kEmptyKey, key);
		if (old == kEmptyKey || old == key) {
			atomicMin(table_values + slot, (int) i);  // a key listed twice resolves to its first block
			return;
		}
	}
}
__device__ __forceinline__ int hash_find(const unsigned long long* __restrict__ table_keys, const int* __restrict__ table_values, unsigned mask, unsigned long long key) {
	for (unsigned slot = hash_slot(key, mask);; slot = (slot + 1) & mask) {
		const unsigned long long k = table_keys[slot];
		if (k == key) return table_values[slot];
		if (k == kEmptyKey) return -1;
	}
}
__global__ void __launch_bounds__(kThreads) neighbors_kernel(const int32_t* __restrict__ keys, int64_t count, const unsigned long long* __restrict__ table_keys,
                                                             const int* __restrict__ table_values, unsigned mask, int32_t* __restrict__ neighbors) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count * 27) return;
	const int64_t b = i / 27;
	const int nb = (int) (i - b * 27), dx = nb % 3 - 1, dy = (nb / 3) % 3 - 1, dz = nb / 9 - 1;
	neighbors[i] = hash_find(table_keys, table_values, mask, pack_key(keys[b * 3] + dx, keys[b * 3 + 1] + dy, keys[b * 3 + 2] + dz));
}

// voxel (x, y, z) relative to block b, coordinates in [-1, R]: linear voxel index across the 27-neighbourhood, -1 where the block is missing
struct Grid {
	const int32_t* neighbors;
	int R;
	__device__ __forceinline__ int64_t voxel(int64_t b, int x, int y, int z) const {
		const int dx = x < 0 ? -1 : (x >= R ? 1 : 0), dy = y < 0 ? -1 : (y >= R ? 1 : 0), dz = z < 0 ? -1 : (z >= R ? 1 : 0);
		const int block = neighbors[b * 27 + (dz + 1) * 9 + (dy + 1) * 3 + (dx + 1)];
		if (block < 0) return -1;
		return ((int64_t) block * R + (z - dz * R)) * R * R + (int64_t) (y - dy * R) * R + (x - dx * R);
	}
};
__device__ __forceinline__ float weight_at(const void* weight, int dtype, int64_t i) {
	return dtype == 1 ? (float) static_cast<const uint16_t*>(weight)[i] : static_cast<const float*>(weight)[i];
}

// pass 0: configuration of the cube whose origin is this voxel; 0 when a corner is missing or too light (Open3D: weight <= threshold)
__global__ void __launch_bounds__(kThreads) cube_configuration_kernel(Grid grid, int64_t voxel_count, const float* __restrict__ tsdf, const void* __restrict__ weight,
                                                                      int weight_dtype, float weight_threshold, int32_t* __restrict__ structure) {
	const int64_t v = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= voxel_count) return;
	const int R = grid.R, x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
	const int64_t b = v / ((int64_t) R * R * R);
	int configuration = 0;
	bool complete = true;
#pragma unroll
	for (int i = 0; i < 8; i++) {
		const int64_t c = grid.voxel(b, x + c_vertex_shifts[i][0], y + c_vertex_shifts[i][1], z + c_vertex_shifts[i][2]);
		if (c < 0 || weight_at(weight, weight_dtype, c) <= weight_threshold) { complete = false; break; }
		configuration |= tsdf[c] < 0.f ? 1 << i : 0;
	}
	structure[v * 4 + 3] = complete ? configuration : 0;
}

// pass 1: the voxel's own x / y / z edges: flagged when the end points differ in sign and one of the four cubes around the edge is complete
__global__ void __launch_bounds__(kThreads) edge_flags_kernel(Grid grid, int64_t voxel_count, const float* __restrict__ tsdf, int32_t* __restrict__ structure,
                                                              int* __restrict__ vertex_counts, int* __restrict__ triangle_counts) {
	const int64_t v = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= voxel_count) return;
	const int R = grid.R, x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
	const int64_t b = v / ((int64_t) R * R * R);
	const bool negative = tsdf[v] < 0.f;
	int flagged = 0;
#pragma unroll
	for (int axis = 0; axis < 3; axis++) {
		const int64_t other = grid.voxel(b, x + (axis == 0), y + (axis == 1), z + (axis == 2));
		bool needs = false;
		if (other >= 0 && (tsdf[other] < 0.f) != negative) {
			// cubes around the edge: origins offset by {0, -1} along the two other axes
			const int a1 = (axis + 1) % 3, a2 = (axis + 2) % 3;
			for (int k = 0; k < 4 && !needs; k++) {
				int o[3] = {x, y, z};
				o[a1] -= k & 1;
				o[a2] -= k >> 1;
				const int64_t cube = grid.voxel(b, o[0], o[1], o[2]);
				if (cube >= 0) {
					const int configuration = structure[cube * 4 + 3];
					needs = configuration != 0 && configuration != 255;
				}
			}
		}
		structure[v * 4 + axis] = needs ? -1 : -2;  // -1: vertex wanted (Open3D's flag), -2: none
		flagged += needs ? 1 : 0;
	}
	vertex_counts[v] = flagged;
	triangle_counts[v] = c_triangle_count[structure[v * 4 + 3]];
}

__global__ void __launch_bounds__(kThreads) number_vertices_kernel(int64_t voxel_count, const int* __restrict__ vertex_offsets, int32_t* __restrict__ structure) {
	const int64_t v = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= voxel_count) return;
	int next = vertex_offsets[v];
#pragma unroll
	for (int axis = 0; axis < 3; axis++)
		if (structure[v * 4 + axis] == -1) structure[v * 4 + axis] = next++;
}

// TSDF gradient by central differences; a component stays 0 when either neighbour's block is missing (Open3D DeviceGetNormal)
__device__ __forceinline__ float3 gradient_at(const Grid& grid, const float* __restrict__ tsdf, int64_t b, int x, int y, int z) {
	float3 n = make_float3(0.f, 0.f, 0.f);
	const int64_t xp = grid.voxel(b, x + 1, y, z), xn = grid.voxel(b, x - 1, y, z), yp = grid.voxel(b, x, y + 1, z), yn = grid.voxel(b, x, y - 1, z),
	              zp = grid.voxel(b, x, y, z + 1), zn = grid.voxel(b, x, y, z - 1);
	if (xp >= 0 && xn >= 0) n.x = tsdf[xp] - tsdf[xn];
	if (yp >= 0 && yn >= 0) n.y = tsdf[yp] - tsdf[yn];
	if (zp >= 0 && zn >= 0) n.z = tsdf[zp] - tsdf[zn];
	return n;
}

// pass 2: vertices of the flagged edges
__global__ void __launch_bounds__(kThreads) emit_vertices_kernel(Grid grid, int64_t voxel_count, const int32_t* __restrict__ block_keys, const float* __restrict__ tsdf,
                                                                 const float* __restrict__ color, float voxel_size, const int32_t* __restrict__ structure,
                                                                 float* __restrict__ vertices, float* __restrict__ normals, float* __restrict__ colors) {
	const int64_t v = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= voxel_count) return;
	const int4 s = reinterpret_cast<const int4*>(structure)[v];
	if (s.x < 0 && s.y < 0 && s.z < 0) return;
	const int R = grid.R, x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
	const int64_t b = v / ((int64_t) R * R * R);
	const float tsdf_o = tsdf[v];
	const float3 no = gradient_at(grid, tsdf, b, x, y, z);
	const float gx = (float) (block_keys[b * 3] * R + x), gy = (float) (block_keys[b * 3 + 1] * R + y), gz = (float) (block_keys[b * 3 + 2] * R + z);
	const int ids[3] = {s.x, s.y, s.z};
#pragma unroll
	for (int axis = 0; axis < 3; axis++) {
		const int id = ids[axis];
		if (id < 0) continue;
		const int ex = x + (axis == 0), ey = y + (axis == 1), ez = z + (axis == 2);
		const int64_t e = grid.voxel(b, ex, ey, ez);
		const float tsdf_e = tsdf[e];
		const float ratio = (0.f - tsdf_o) / (tsdf_e - tsdf_o);
		vertices[(int64_t) id * 3 + 0] = voxel_size * (gx + (axis == 0 ? ratio : 0.f));
		vertices[(int64_t) id * 3 + 1] = voxel_size * (gy + (axis == 1 ? ratio : 0.f));
		vertices[(int64_t) id * 3 + 2] = voxel_size * (gz + (axis == 2 ? ratio : 0.f));
		// the neighbour's gradient is taken in ITS block's frame
		const int R2 = R;
		const int64_t eb = e / ((int64_t) R2 * R2 * R2);
		const float3 ne = gradient_at(grid, tsdf, eb, (int) (e % R2), (int) ((e / R2) % R2), (int) ((e / ((int64_t) R2 * R2)) % R2));
		const float nx = (1.f - ratio) * no.x + ratio * ne.x, ny = (1.f - ratio) * no.y + ratio * ne.y, nz = (1.f - ratio) * no.z + ratio * ne.z;
		const float norm = sqrtf(nx * nx + ny * ny + nz * nz) + 1e-5f;
		normals[(int64_t) id * 3 + 0] = nx / norm;
		normals[(int64_t) id * 3 + 1] = ny / norm;
		normals[(int64_t) id * 3 + 2] = nz / norm;
		if (color != nullptr) {
#pragma unroll
			for (int c = 0; c < 3; c++) colors[(int64_t) id * 3 + c] = ((1.f - ratio) * color[v * 3 + c] + ratio * color[e * 3 + c]) / 255.f;
		}
	}
}

// pass 3: triangles through the vertex ids of the cube's edges
__global__ void __launch_bounds__(kThreads) emit_triangles_kernel(Grid grid, int64_t voxel_count, const int32_t* __restrict__ structure,
                                                                  const int* __restrict__ triangle_offsets, int64_t* __restrict__ triangles) {
	const int64_t v = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= voxel_count) return;
	const int configuration = structure[v * 4 + 3];
	const int count = c_triangle_count[configuration];
	if (count == 0) return;
	const int R = grid.R, x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
	const int64_t b = v / ((int64_t) R * R * R);
	const int64_t first = triangle_offsets[v];
	for (int t = 0; t < count; t++)
#pragma unroll
		for (int k = 0; k < 3; k++) {
			const int edge = c_triangle_table[configuration][t * 3 + k];
			const int64_t owner = grid.voxel(b, x + c_edge_shifts[edge][0], y + c_edge_shifts[edge][1], z + c_edge_shifts[edge][2]);
			triangles[(first + t) * 3 + k] = structure[owner * 4 + c_edge_shifts[edge][3]];
		}
}

int upload_tables() {
	static std::atomic<unsigned long long> done{0};
	int device = 0;
	NNRT_CUDA(cudaGetDevice(&device));
	const unsigned long long bit = 1ull << (device & 63);
	if (done.load(std::memory_order_acquire) & bit) return kOk;
	NNRT_CUDA(cudaMemcpyToSymbol(c_triangle_table, kMcTriangleTable, sizeof(kMcTriangleTable)));
	NNRT_CUDA(cudaMemcpyToSymbol(c_triangle_count, kMcTriangleCount, sizeof(kMcTriangleCount)));
	NNRT_CUDA(cudaMemcpyToSymbol(c_vertex_shifts, kMcVertexShifts, sizeof(kMcVertexShifts)));
	NNRT_CUDA(cudaMemcpyToSymbol(c_edge_shifts, kMcEdgeShifts, sizeof(kMcEdgeShifts)));
	done.fetch_or(bit, std::memory_order_release);
	return kOk;
}

}  // namespace
}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_voxel_block_neighbors(const int32_t* block_keys, int64_t block_count, int32_t* neighbor_indices, void* stream) {
	if (block_count == 0) return kOk;
	NNRT_CHECK_ARG(block_keys && neighbor_indices && block_count > 0 && block_count < (1 << 28), "voxel_block_neighbors: null pointer or block count out of range");
	cudaStream_t s = (cudaStream_t) stream;
	unsigned capacity = 64;
	while (capacity < (unsigned) block_count * 2u) capacity <<= 1;
	unsigned long long* table_keys = nullptr;
	int* table_values = nullptr;
	Scratch scratch(s);
	NNRT_CUDA(scratch.alloc(&table_keys, sizeof(unsigned long long) * capacity));
	NNRT_CUDA(scratch.alloc(&table_values, sizeof(int) * capacity));
	NNRT_CUDA(cudaMemsetAsync(table_keys, 0xFF, sizeof(unsigned long long) * capacity, s));
	NNRT_CUDA(cudaMemsetAsync(table_values, 0x7F, sizeof(int) * capacity, s));
	hash_insert_kernel<<<blocks_for(block_count, kThreads), kThreads, 0, s>>>(block_keys, block_count, table_keys, table_values, capacity - 1);
	neighbors_kernel<<<blocks_for(block_count * 27, kThreads), kThreads, 0, s>>>(block_keys, block_count, table_keys, table_values, capacity - 1, neighbor_indices);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_extract_triangle_mesh_prepare(int64_t block_count, const int32_t* neighbor_indices, const float* tsdf, const void* weight,
                                                       int32_t weight_dtype, int32_t block_resolution, float weight_threshold, int32_t* mesh_structure,
                                                       int32_t* triangle_offsets, int64_t* counts_host, void* stream) {
	NNRT_CHECK_ARG(counts_host, "extract_triangle_mesh_prepare: counts_host is null");
	counts_host[0] = counts_host[1] = 0;
	if (block_count == 0) return kOk;
	NNRT_CHECK_ARG(neighbor_indices && tsdf && weight && mesh_structure && triangle_offsets, "extract_triangle_mesh_prepare: null pointer argument");
	NNRT_CHECK_ARG(block_resolution >= 2 && block_resolution <= 32, "block_resolution must be in [2, 32], got %d", block_resolution);
	NNRT_CHECK_ARG(weight_dtype == 0 || weight_dtype == 1, "weight dtype must be float32 (0) or uint16 (1)");
	const int64_t voxel_count = block_count * block_resolution * block_resolution * block_resolution;
	NNRT_CHECK_ARG(block_count > 0 && voxel_count < (1ll << 31) - 8, "extract_triangle_mesh_prepare: too many voxels (%lld)", (long long) voxel_count);
	if (upload_tables() != kOk) return kCudaError;
	cudaStream_t s = (cudaStream_t) stream;
	const Grid grid{neighbor_indices, block_resolution};
	int *vertex_counts = nullptr, *scan_tiles = nullptr;
	Scratch scratch(s);
	NNRT_CUDA(scratch.alloc(&vertex_counts, sizeof(int) * (size_t) (voxel_count + 1)));
	NNRT_CUDA(scratch.alloc(&scan_tiles, sizeof(int) * (size_t) scan_scratch_ints(voxel_count + 1)));
	const unsigned blocks = blocks_for(voxel_count, kThreads);
	cube_configuration_kernel<<<blocks, kThreads, 0, s>>>(grid, voxel_count, tsdf, weight, weight_dtype, weight_threshold, mesh_structure);
	edge_flags_kernel<<<blocks, kThreads, 0, s>>>(grid, voxel_count, tsdf, mesh_structure, vertex_counts, triangle_offsets);
	int rc = exclusive_scan(vertex_counts, vertex_counts, voxel_count, scan_tiles, s);
	if (rc != kOk) return rc;
	rc = exclusive_scan(triangle_offsets, triangle_offsets, voxel_count, scan_tiles, s);
	if (rc != kOk) return rc;
	number_vertices_kernel<<<blocks, kThreads, 0, s>>>(voxel_count, vertex_counts, mesh_structure);
	NNRT_LAUNCH_CHECK();
	int totals[2] = {0, 0};
	NNRT_CUDA(cudaMemcpyAsync(&totals[0], vertex_counts + voxel_count, sizeof(int), cudaMemcpyDeviceToHost, s));
	NNRT_CUDA(cudaMemcpyAsync(&totals[1], triangle_offsets + voxel_count, sizeof(int), cudaMemcpyDeviceToHost, s));
	NNRT_CUDA(cudaStreamSynchronize(s));  // the counts size the caller's outputs (as Open3D's first counting pass does)
	counts_host[0] = totals[0];
	counts_host[1] = totals[1];
	return kOk;
}

extern "C" int nnrt_b200_extract_triangle_mesh_emit(const int32_t* block_keys, int64_t block_count, const int32_t* neighbor_indices, const float* tsdf,
                                                    const float* color, int32_t block_resolution, float voxel_size, const int32_t* mesh_structure,
                                                    const int32_t* triangle_offsets, float* vertices, float* normals, float* colors, int64_t* triangles,
                                                    void* stream) {
	if (block_count == 0) return kOk;
	NNRT_CHECK_ARG(block_keys && neighbor_indices && tsdf && mesh_structure && triangle_offsets, "extract_triangle_mesh_emit: null pointer argument");
	NNRT_CHECK_ARG((color == nullptr) == (colors == nullptr), "extract_triangle_mesh_emit: color and colors must be given together");
	NNRT_CHECK_ARG(block_resolution >= 2 && block_resolution <= 32, "block_resolution must be in [2, 32], got %d", block_resolution);
	const int64_t voxel_count = block_count * block_resolution * block_resolution * block_resolution;
	if (upload_tables() != kOk) return kCudaError;
	cudaStream_t s = (cudaStream_t) stream;
	const Grid grid{neighbor_indices, block_resolution};
	const unsigned blocks = blocks_for(voxel_count, kThreads);
	if (vertices != nullptr) {
		NNRT_CHECK_ARG(normals != nullptr, "extract_triangle_mesh_emit: normals is null");
		emit_vertices_kernel<<<blocks, kThreads, 0, s>>>(grid, voxel_count, block_keys, tsdf, color, voxel_size, mesh_structure, vertices, normals, colors);
	}
	if (triangles != nullptr) emit_triangles_kernel<<<blocks, kThreads, 0, s>>>(grid, voxel_count, mesh_structure, triangle_offsets, triangles);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
