// ORACLE — TEST INFRASTRUCTURE ONLY. Triangle-mesh extraction from TSDF voxel blocks: VoxelBlockGrid::ExtractTriangleMesh
// (cpp/geometry/VoxelBlockGrid.cpp:461-484) -> Open3D t::geometry::kernel::voxel_grid::ExtractTriangleMesh (Open3D v0.17.0-1fix6008, third
// party, absent from the reference checkout). PARITY UNPINNED: restated from Open3D's published algorithm (VoxelBlockGridImpl.h) with its
// structure — a block map, pass 0 scatters "-1" flags into the voxels that own the cube's sign-change edges, later passes number and
// emit — run sequentially in (block, voxel, edge) order, which is one of the orders Open3D's atomic counters can produce. The lookup
// tables are the derived ones of tools/generate_marching_cubes_tables.py (shared with the CUDA path; tests check their topology).
#include "oracle_common.h"
#include "../dynamicfuion_python_b200/csrc/marching_cubes_tables.h"

#include <unordered_map>

using namespace oracle;

namespace {
struct Blocks {
	const int32_t* keys;
	int64_t count;
	int R;
	std::unordered_map<uint64_t, int64_t> map;
	static uint64_t pack(int x, int y, int z) { return ((uint64_t) (uint32_t) (x + (1 << 20)) << 42) ^ ((uint64_t) (uint32_t) (y + (1 << 20)) << 21) ^ (uint64_t) (uint32_t) (z + (1 << 20)); }
	Blocks(const int32_t* k, int64_t n, int r) : keys(k), count(n), R(r) {
		for (int64_t b = 0; b < n; b++) map.emplace(pack(k[b * 3], k[b * 3 + 1], k[b * 3 + 2]), b);  // first block of a repeated key wins
	}
	// voxel (x, y, z) relative to block b, coordinates may leave the block by one: linear index or -1
	int64_t voxel(int64_t b, int x, int y, int z) const {
		const int dx = x < 0 ? -1 : (x >= R ? 1 : 0), dy = y < 0 ? -1 : (y >= R ? 1 : 0), dz = z < 0 ? -1 : (z >= R ? 1 : 0);
		int64_t block = b;
		if (dx || dy || dz) {
			auto it = map.find(pack(keys[b * 3] + dx, keys[b * 3 + 1] + dy, keys[b * 3 + 2] + dz));
			if (it == map.end()) return -1;
			block = it->second;
		}
		return ((block * R + (z - dz * R)) * R + (y - dy * R)) * R + (x - dx * R);
	}
	void normal(const float* tsdf, int64_t v, float* n) const {  // DeviceGetNormal: central differences, 0 where a neighbour block is missing
		const int x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
		const int64_t b = v / ((int64_t) R * R * R);
		n[0] = n[1] = n[2] = 0.f;
		const int64_t xp = voxel(b, x + 1, y, z), xn = voxel(b, x - 1, y, z), yp = voxel(b, x, y + 1, z), yn = voxel(b, x, y - 1, z), zp = voxel(b, x, y, z + 1),
		              zn = voxel(b, x, y, z - 1);
		if (xp >= 0 && xn >= 0) n[0] = tsdf[xp] - tsdf[xn];
		if (yp >= 0 && yn >= 0) n[1] = tsdf[yp] - tsdf[yn];
		if (zp >= 0 && zn >= 0) n[2] = tsdf[zp] - tsdf[zn];
	}
};
}  // namespace

// BufferRadiusNeighbors: [B, 27] block indices at key + (dx, dy, dz), slot (dz + 1) * 9 + (dy + 1) * 3 + (dx + 1)
ORACLE_API void oracle_voxel_block_neighbors(const int32_t* keys, int64_t count, int32_t* neighbors) {
	Blocks blocks(keys, count, 1);
	for (int64_t b = 0; b < count; b++)
		for (int nb = 0; nb < 27; nb++) {
			const int dx = nb % 3 - 1, dy = (nb / 3) % 3 - 1, dz = nb / 9 - 1;
			auto it = blocks.map.find(Blocks::pack(keys[b * 3] + dx, keys[b * 3 + 1] + dy, keys[b * 3 + 2] + dz));
			neighbors[b * 27 + nb] = it == blocks.map.end() ? -1 : (int32_t) it->second;
		}
}

// counts[0..1] = vertices, triangles; outputs may be null (counting call) and are filled up to the given capacities otherwise
ORACLE_API void oracle_extract_triangle_mesh(const int32_t* keys, int64_t block_count, const float* tsdf, const float* weight, const float* color, int R,
                                             float voxel_size, float weight_threshold, int64_t* counts, float* vertices, float* normals, float* colors,
                                             int64_t* triangles) {
	using namespace nnrt_b200;
	Blocks blocks(keys, block_count, R);
	const int64_t voxels = block_count * R * R * R;
	std::vector<int32_t> structure((size_t) voxels * 4, 0);
	// pass 0: cube configurations; flag (-1) the edges that carry a vertex in the voxel that owns them
	for (int64_t v = 0; v < voxels; v++) {
		const int x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
		const int64_t b = v / ((int64_t) R * R * R);
		int configuration = 0;
		bool complete = true;
		for (int i = 0; i < 8 && complete; i++) {
			const int64_t c = blocks.voxel(b, x + kMcVertexShifts[i][0], y + kMcVertexShifts[i][1], z + kMcVertexShifts[i][2]);
			if (c < 0 || weight[c] <= weight_threshold) complete = false;
			else configuration |= tsdf[c] < 0.f ? 1 << i : 0;
		}
		if (!complete) continue;
		structure[v * 4 + 3] = configuration;
		if (configuration == 0 || configuration == 255) continue;
		for (int e = 0; e < 12; e++)
			if (kMcEdgeTable[configuration] & (1 << e)) {
				const int64_t owner = blocks.voxel(b, x + kMcEdgeShifts[e][0], y + kMcEdgeShifts[e][1], z + kMcEdgeShifts[e][2]);
				structure[owner * 4 + kMcEdgeShifts[e][3]] = -1;
			}
	}
	// passes 1 + 2: number and emit the vertices
	int64_t vertex_count = 0;
	for (int64_t v = 0; v < voxels; v++) {
		const int x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
		const int64_t b = v / ((int64_t) R * R * R);
		for (int e = 0; e < 3; e++) {
			if (structure[v * 4 + e] != -1) continue;
			const int64_t id = vertex_count++;
			structure[v * 4 + e] = (int32_t) id;
			if (!vertices) continue;
			const int64_t other = blocks.voxel(b, x + (e == 0), y + (e == 1), z + (e == 2));
			const float tsdf_o = tsdf[v], tsdf_e = tsdf[other];
			const float ratio = (0.f - tsdf_o) / (tsdf_e - tsdf_o);
			const float g[3] = {(float) (keys[b * 3] * R + x), (float) (keys[b * 3 + 1] * R + y), (float) (keys[b * 3 + 2] * R + z)};
			for (int c = 0; c < 3; c++) vertices[id * 3 + c] = voxel_size * (g[c] + (e == c ? ratio : 0.f));
			float no[3], ne[3], n[3];
			blocks.normal(tsdf, v, no);
			blocks.normal(tsdf, other, ne);
			for (int c = 0; c < 3; c++) n[c] = (1.f - ratio) * no[c] + ratio * ne[c];
			const float norm = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]) + 1e-5f;
			for (int c = 0; c < 3; c++) normals[id * 3 + c] = n[c] / norm;
			if (color && colors)
				for (int c = 0; c < 3; c++) colors[id * 3 + c] = ((1.f - ratio) * color[v * 3 + c] + ratio * color[other * 3 + c]) / 255.f;
		}
	}
	// pass 3: triangles
	int64_t triangle_count = 0;
	for (int64_t v = 0; v < voxels; v++) {
		const int configuration = structure[v * 4 + 3];
		const int x = (int) (v % R), y = (int) ((v / R) % R), z = (int) ((v / ((int64_t) R * R)) % R);
		const int64_t b = v / ((int64_t) R * R * R);
		for (int t = 0; t < 16 && kMcTriangleTable[configuration][t] >= 0; t += 3) {
			const int64_t id = triangle_count++;
			if (!triangles) continue;
			for (int k = 0; k < 3; k++) {
				const int e = kMcTriangleTable[configuration][t + k];
				const int64_t owner = blocks.voxel(b, x + kMcEdgeShifts[e][0], y + kMcEdgeShifts[e][1], z + kMcEdgeShifts[e][2]);
				triangles[id * 3 + k] = structure[owner * 4 + kMcEdgeShifts[e][3]];
			}
		}
	}
	counts[0] = vertex_count;
	counts[1] = triangle_count;
}
