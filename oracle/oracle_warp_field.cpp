// ORACLE (test infrastructure only) — warp-field construction: median-grid subsampling,
// K nearest coarser nodes, adjacency -> edge list, regularisation layers, node coverage weights.
// See oracle_common.h for status. Sequential restatement: every "first come" rule of the reference (hash-map buffer
// index, atomic slot in a bin) is resolved in ascending point-index order, which is what the reference's CPU path does
// when run on one thread; the reference's own tests compare these outputs as SETS for that reason
// (cpp/tests/test_sampling.cpp:150-177, cpp/tests/test_graph_warp_field.cpp:141-339).
#include "oracle_common.h"
#include <cfloat>
#include <map>
#include <array>

using namespace oracle;

namespace {

// Eigen (a - b).norm() of 3-vectors: sqrt of the x + (y + z) reduction
inline float distance3(const float* a, const float* b) {
	V3 d = sub(load3(a), load3(b));
	return sqrtf(dot3(d, d));
}

// cpp/geometry/functional/kernel/GeometrySamplingGridBinning.h:30-52 (points / cell + offset, Floor, To(Int32); bin index =
// hash-map buffer index, here: order of first occurrence) + GeometrySamplingMedian.h:36-190,254-285.
int64_t median_grid_subsample(const float* points, int64_t point_count, float grid_cell_size, int64_t* sample, float grid_offset = 0.0f) {
	std::map<std::array<int32_t, 3>, int32_t> bin_of_cell;
	std::vector<int32_t> point_bin(point_count);
	std::vector<std::vector<int32_t>> bin_points;  // ascending point index = bin-sorted order of a sequential run
	for (int64_t i = 0; i < point_count; i++) {
		std::array<int32_t, 3> cell;
		for (int c = 0; c < 3; c++) cell[c] = (int32_t) floorf(points[i * 3 + c] / grid_cell_size + grid_offset);
		auto found = bin_of_cell.find(cell);
		int32_t bin;
		if (found == bin_of_cell.end()) {
			bin = (int32_t) bin_points.size();
			bin_of_cell.emplace(cell, bin);
			bin_points.emplace_back();
		} else {
			bin = found->second;
		}
		point_bin[i] = bin;
		bin_points[bin].push_back((int32_t) i);
	}
	for (size_t bin = 0; bin < bin_points.size(); bin++) {
		const auto& members = bin_points[bin];
		float min_distance_sum = FLT_MAX;
		int64_t median = -1;  // the reference leaves the entry uninitialised if no sum is below FLT_MAX
		for (int32_t source : members) {
			float sum = 0.f;  // FindMedianPointIndices: row sum of the bin's distance matrix, left to right
			for (int32_t target : members) sum += distance3(points + (int64_t) target * 3, points + (int64_t) source * 3);
			if (sum < min_distance_sum) { min_distance_sum = sum; median = source; }
		}
		sample[bin] = median;
	}
	return (int64_t) bin_points.size();
}

// cpp/core/kernel/SearchKdTreeImpl.h:118-146 (PLAIN strategy: replace the current maximum when strictly closer; distance =
// (a - b).norm(), :64). The kd-tree only prunes; its visiting order matters for exact distance ties alone. Here the indexed
// points are visited in ascending index order. Missing neighbours stay -1 (fewer indexed points than k).
void find_k_nearest(const float* indexed_points, int indexed_count, const float* query_points, int64_t query_count, int k,
                    int32_t* neighbor_indices, float* neighbor_distances) {
	std::vector<float> scratch(k);
	for (int64_t q = 0; q < query_count; q++) {
		int32_t* idx = neighbor_indices + q * k;
		float* dist = neighbor_distances ? neighbor_distances + q * k : scratch.data();
		for (int a = 0; a < k; a++) { idx[a] = -1; dist[a] = INFINITY; }
		int max_at = 0;
		float max_distance = INFINITY;
		for (int n = 0; n < indexed_count; n++) {
			float d = distance3(indexed_points + n * 3, query_points + q * 3);
			if (max_distance > d) {
				dist[max_at] = d;
				idx[max_at] = n;
				max_at = 0;
				max_distance = dist[0];
				for (int a = 1; a < k; a++)
					if (dist[a] > max_distance) { max_at = a; max_distance = dist[a]; }
			}
		}
	}
}

}  // namespace

ORACLE_API int64_t oracle_median_grid_subsample_3d_points(const float* points, int64_t point_count, float grid_cell_size, float grid_offset, int64_t* sample) {
	return median_grid_subsample(points, point_count, grid_cell_size, sample, grid_offset);
}

ORACLE_API void oracle_find_k_nearest(const float* indexed_points, int indexed_count, const float* query_points, int64_t query_count,
                                      int k, int32_t* neighbor_indices, float* neighbor_distances) {
	find_k_nearest(indexed_points, indexed_count, query_points, query_count, k, neighbor_indices, neighbor_distances);
}

// cpp/geometry/WarpField.cpp:249-262: squared distance to the nearest OTHER node (second entry of the sorted 2-NN of the node set
// queried with itself).
ORACLE_API void oracle_node_coverage_weights(const float* nodes, int node_count, float node_coverage, float* weights) {
	if (node_count == 1) { weights[0] = node_coverage; return; }
	for (int i = 0; i < node_count; i++) {
		float nearest = INFINITY;
		for (int j = 0; j < node_count; j++) {
			if (j == i) continue;
			float d = distance3(nodes + j * 3, nodes + i * 3);
			if (d < nearest) nearest = d;
		}
		weights[i] = nearest * nearest;
	}
}

// cpp/geometry/kernel/HierarchicalGraphWarpFieldImpl.h:218-296 (AdjacencyArrayToEdges): slot (source, neighbour) of the dense
// [S, D] table, source order optionally flipped, then every -1 dropped. Returns the edge count.
ORACLE_API int64_t oracle_adjacency_array_to_edges(const int32_t* adjacency, int64_t source_count, int max_degree, const int32_t* source_indices,
                                                   const int32_t* target_indices, int flip_source_order, int32_t* edges) {
	int64_t count = 0;
	for (int64_t row = 0; row < source_count; row++) {
		const int64_t i_source = flip_source_order ? source_count - 1 - row : row;
		for (int a = 0; a < max_degree; a++) {
			const int32_t i_target = adjacency[i_source * max_degree + a];
			if (i_target == -1) continue;
			edges[count * 2] = source_indices[i_source];
			edges[count * 2 + 1] = target_indices[i_target];
			count++;
		}
	}
	return count;
}

// cpp/geometry/HierarchicalGraphWarpField.cpp:74-195 (RebuildRegularizationLayers). layer_radii[i] = the caller's
// compute_layer_decimation_radius(i, node_coverage) for i >= 1 (entry 0 is node_coverage, :79).
// Outputs: node_indices [N] (virtual -> original, layers fine to coarse), layer_node_counts [L], edges [<= N*D, 2] int32 in virtual
// indices (layers coarse to fine, sources descending), edge_layer_indices [edges] int8 (= layer of the targets). Returns the edge
// count, or -(layer) if that layer's sample is as large as the finer layer (the reference's LogError, :96-101).
ORACLE_API int64_t oracle_build_regularization_layers(const float* nodes, int node_count, const float* layer_radii, int layer_count,
                                                      int max_vertex_degree, int64_t* node_indices, int32_t* layer_node_counts,
                                                      int32_t* edges, int8_t* edge_layer_indices) {
	std::vector<std::vector<int64_t>> layer_indices(layer_count);
	std::vector<std::vector<float>> layer_positions(layer_count);
	layer_indices[0].resize(node_count);
	for (int i = 0; i < node_count; i++) layer_indices[0][i] = i;
	layer_positions[0].assign(nodes, nodes + (size_t) node_count * 3);
	for (int i_layer = 1; i_layer < layer_count; i_layer++) {
		auto& finer_idx = layer_indices[i_layer - 1];
		auto& finer_pos = layer_positions[i_layer - 1];
		const int64_t finer_count = (int64_t) finer_idx.size();
		std::vector<int64_t> sample(finer_count);
		const int64_t sample_count = median_grid_subsample(finer_pos.data(), finer_count, layer_radii[i_layer] * 2, sample.data());
		if (sample_count == finer_count) return -(int64_t) i_layer;
		std::vector<char> keep(finer_count, 1);
		for (int64_t s = 0; s < sample_count; s++) {
			keep[sample[s]] = 0;
			layer_indices[i_layer].push_back(finer_idx[sample[s]]);
			for (int c = 0; c < 3; c++) layer_positions[i_layer].push_back(finer_pos[sample[s] * 3 + c]);
		}
		std::vector<int64_t> filtered_idx;
		std::vector<float> filtered_pos;
		for (int64_t i = 0; i < finer_count; i++) {
			if (!keep[i]) continue;
			filtered_idx.push_back(finer_idx[i]);
			for (int c = 0; c < 3; c++) filtered_pos.push_back(finer_pos[i * 3 + c]);
		}
		finer_idx.swap(filtered_idx);
		finer_pos.swap(filtered_pos);
	}
	std::vector<int32_t> first_virtual(layer_count);
	int32_t virtual_count = 0;
	for (int i_layer = 0; i_layer < layer_count; i_layer++) {
		first_virtual[i_layer] = virtual_count;
		layer_node_counts[i_layer] = (int32_t) layer_indices[i_layer].size();
		for (int64_t index : layer_indices[i_layer]) node_indices[virtual_count++] = index;
	}
	// edges: finer layer i_layer - 1 -> layer i_layer, concatenated from the coarsest pair down (:130-190)
	int64_t edge_count = 0;
	for (int i_layer = layer_count - 1; i_layer >= 1; i_layer--) {
		const auto& current_pos = layer_positions[i_layer];
		const auto& finer_pos = layer_positions[i_layer - 1];
		const int current_count = (int) layer_indices[i_layer].size();
		const int64_t finer_count = (int64_t) layer_indices[i_layer - 1].size();
		std::vector<int32_t> adjacency(finer_count * max_vertex_degree);
		find_k_nearest(current_pos.data(), current_count, finer_pos.data(), finer_count, max_vertex_degree, adjacency.data(), nullptr);
		for (int64_t q = 0; q < finer_count; q++)  // SortTensorAlongLastDimension(adjacency, true, DESC)
			std::sort(adjacency.begin() + q * max_vertex_degree, adjacency.begin() + (q + 1) * max_vertex_degree, [](int32_t a, int32_t b) { return a > b; });
		std::vector<int32_t> finer_virtual(finer_count), current_virtual(current_count);
		for (int64_t i = 0; i < finer_count; i++) finer_virtual[i] = first_virtual[i_layer - 1] + (int32_t) i;
		for (int i = 0; i < current_count; i++) current_virtual[i] = first_virtual[i_layer] + i;
		const int64_t added = oracle_adjacency_array_to_edges(adjacency.data(), finer_count, max_vertex_degree, finer_virtual.data(),
		                                                      current_virtual.data(), 1, edges + edge_count * 2);
		// "we record the layer of the targets as the official edge layer" (:176-177): edges of source layer i_layer - 1 get i_layer
		for (int64_t e = 0; e < added; e++) edge_layer_indices[edge_count + e] = (int8_t) i_layer;
		edge_count += added;
	}
	return edge_count;
}

// cpp/geometry/functional/kernel/GeometrySamplingMean.h:37-92,159-236 (grid binning as above): bin means with the members added in
// ascending point order, and per bin the point closest to the mean (squared distance, strict `>` replacement: first among equals).
// mode 0: out_points [bins,3]; mode 1: out_indices [bins]. Returns the bin count.
ORACLE_API int64_t oracle_grid_mean(const float* points, int64_t point_count, float grid_cell_size, float grid_offset, int mode, float* out_points,
                                    int64_t* out_indices) {
	std::map<std::array<int32_t, 3>, int32_t> bin_of_cell;
	std::vector<int32_t> point_bin(point_count);
	std::vector<std::array<float, 3>> sums;
	std::vector<int32_t> counts;
	for (int64_t i = 0; i < point_count; i++) {
		std::array<int32_t, 3> cell;
		for (int c = 0; c < 3; c++) cell[c] = (int32_t) floorf(points[i * 3 + c] / grid_cell_size + grid_offset);
		auto found = bin_of_cell.find(cell);
		int32_t bin;
		if (found == bin_of_cell.end()) {
			bin = (int32_t) sums.size();
			bin_of_cell.emplace(cell, bin);
			sums.push_back({0.f, 0.f, 0.f});
			counts.push_back(0);
		} else {
			bin = found->second;
		}
		point_bin[i] = bin;
		for (int c = 0; c < 3; c++) sums[bin][c] += points[i * 3 + c];
		counts[bin]++;
	}
	const int64_t bins = (int64_t) sums.size();
	for (int64_t b = 0; b < bins; b++)
		for (int c = 0; c < 3; c++) sums[b][c] /= (float) counts[b];
	if (mode == 0) {
		for (int64_t b = 0; b < bins; b++)
			for (int c = 0; c < 3; c++) out_points[b * 3 + c] = sums[b][c];
		return bins;
	}
	std::vector<float> best(bins, INFINITY);
	for (int64_t b = 0; b < bins; b++) out_indices[b] = -1;
	for (int64_t i = 0; i < point_count; i++) {
		const int32_t b = point_bin[i];
		V3 d = sub(V3{sums[b][0], sums[b][1], sums[b][2]}, load3(points + i * 3));
		float d2 = dot3(d, d);
		if (best[b] > d2) { best[b] = d2; out_indices[b] = i; }
	}
	return bins;
}
