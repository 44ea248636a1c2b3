#include "warp_field.h"

#include <algorithm>

namespace nnrt::geometry {

py::object make_mesh(const Dev& positions, const Dev& normals, const Dev& triangles) {
	TriangleMesh mesh;
	mesh.vertex["positions"] = positions.object;
	if (normals.defined()) mesh.vertex["normals"] = normals.object;
	if (triangles.defined()) mesh.triangle["indices"] = triangles.object;
	return py::cast(std::move(mesh));
}

// ComputeAnchorsAndWeights_Euclidean_{Fixed,Variable}NodeWeight (cpp/geometry/functional/WarpAnchorComputation.cpp:74-148)
std::pair<Dev, Dev> compute_anchors_euclidean(const Dev& points, const Dev& nodes, const Dev& node_weights, float node_coverage, int anchor_count,
                                              int minimum_valid_anchor_count) {
	if (points.dim() < 1 || points.size(-1) != 3) log_error("points must have shape [..., 3], got " + points.shape_string());
	if (nodes.dim() != 2 || nodes.size(1) != 3) log_error("nodes must have shape [N, 3], got " + nodes.shape_string());
	if (node_weights.defined() && node_weights.numel() != nodes.size(0)) log_error("node_coverage_weights must have one entry per node");
	const int64_t count = points.numel() / 3;
	std::vector<int64_t> out_shape(points.shape.begin(), points.shape.end() - 1);
	out_shape.push_back(anchor_count);
	Dev anchors = empty(out_shape, Dtype::Int32), weights = empty(out_shape, Dtype::Float32);
	check(nnrt_b200_compute_anchors_euclidean(points.data<float>(), count, nodes.data<float>(), (int32_t) nodes.size(0),
	                                          node_weights.defined() ? node_weights.data<float>() : nullptr, node_coverage, anchor_count,
	                                          minimum_valid_anchor_count, anchors.data<int32_t>(), weights.data<float>(), current_stream()));
	return {anchors, weights};
}

// WarpTriangleMeshUsingSuppliedAnchors / WarpPointCloud (cpp/geometry/functional/Warping.cpp:169-263)
std::pair<Dev, Dev> warp_points_normals(const Dev& points, const Dev& normals, const Dev& nodes, const Dev& rotations, const Dev& translations,
                                        const Dev& anchors, const Dev& anchor_weights, int minimum_valid_anchor_count, const py::handle& extrinsics,
                                        bool dual_quaternion) {
	const int64_t n = nodes.size(0);
	if (rotations.shape != std::vector<int64_t>{n, 3, 3} || translations.shape != std::vector<int64_t>{n, 3})
		log_error("node_rotations must be [N,3,3] and node_translations [N,3]");
	if (anchors.shape != anchor_weights.shape || anchors.dim() != 2 || anchors.size(0) != points.size(0))
		log_error("anchors / anchor_weights must be [point_count, anchor_count]");
	const std::vector<double> ext = host_matrix(extrinsics, 4, 4, true);
	Dev warped_points = empty(points.shape, Dtype::Float32), warped_normals;
	if (normals.defined()) warped_normals = empty(normals.shape, Dtype::Float32);
	auto fn = dual_quaternion ? nnrt_b200_warp_points_normals_dual_quaternion : nnrt_b200_warp_points_normals;
	check(fn(points.data<float>(), normals.defined() ? normals.data<float>() : nullptr, points.size(0), nodes.data<float>(), rotations.data<float>(),
	         translations.data<float>(), (int32_t) n, anchors.data<int32_t>(), anchor_weights.data<float>(), (int32_t) anchors.size(1),
	         minimum_valid_anchor_count, ext.data(), warped_points.data<float>(), normals.defined() ? warped_normals.data<float>() : nullptr,
	         current_stream()));
	return {warped_points, warped_normals};
}

// ---- WarpField (cpp/geometry/WarpField.cpp) -----------------------------------------------------------------------------------
WarpField::WarpField(const py::handle& nodes_in, float node_coverage, bool threshold_nodes_by_distance, int anchor_count, int minimum_valid_anchor_count,
                     WarpNodeCoverageComputationMethod method, int warp_node_coverage_neighbor_count)
		: nodes(as_cuda(nodes_in, Dtype::Float32)), node_coverage(node_coverage), threshold_nodes_by_distance_by_default(threshold_nodes_by_distance),
		  anchor_count(anchor_count), minimum_valid_anchor_count(minimum_valid_anchor_count), warp_node_coverage_computation_method(method),
		  warp_node_coverage_neighbor_count(warp_node_coverage_neighbor_count) {
	if (nodes.dim() != 2 || nodes.size(1) != 3) log_error("nodes must have shape [N, 3], got " + nodes.shape_string());
	const int64_t n = nodes.size(0);
	if (n < anchor_count)  // WarpField.cpp:57-61
		log_error("Anchor count for warp field, " + std::to_string(anchor_count) + ", exceeds node count, which is " + std::to_string(n) +
		          ". Anchors are nodes closest to a specific point using some distance metric, hence there cannot be more anchors than nodes.");
	node_rotations = empty({n, 3, 3}, Dtype::Float32);
	ResetRotations();
	node_translations = zeros({n, 3}, Dtype::Float32);
	if (method == WarpNodeCoverageComputationMethod::MINIMAL_K_NEIGHBOR_NODE_DISTANCE) RecomputeNodeCoverageWeights();
}

// WarpField.cpp:249-262
void WarpField::RecomputeNodeCoverageWeights() {
	node_coverage_weights = empty({node_count()}, Dtype::Float32);
	check(nnrt_b200_node_coverage_weights(nodes.data<float>(), (int32_t) node_count(), node_coverage, node_coverage_weights.data<float>(), current_stream()));
}

// WarpField.cpp:89-91
py::object WarpField::GetWarpedNodes() const {
	Dev out = empty(nodes.shape, Dtype::Float32);
	check(nnrt_b200_add_scaled(nodes.data<float>(), node_translations.data<float>(), 1.f, nodes.numel(), out.data<float>(), current_stream()));
	return out.object;
}

// WarpField.cpp:93-98: [2,3] min / max corner. A few thousand nodes: reduced on the host (one synchronising copy, as Tensor::Min is).
py::object WarpField::GetNodeExtent() const {
	py::object host = nodes.object.attr("cpu")();
	const float* p = reinterpret_cast<const float*>(host.attr("data_ptr")().cast<uintptr_t>());
	float extent[6] = {p[0], p[1], p[2], p[0], p[1], p[2]};
	for (int64_t i = 1; i < node_count(); i++)
		for (int k = 0; k < 3; k++) {
			extent[k] = std::min(extent[k], p[i * 3 + k]);
			extent[3 + k] = std::max(extent[3 + k], p[i * 3 + k]);
		}
	Torch& t = Torch::get();
	std::vector<std::vector<float>> rows = {{extent[0], extent[1], extent[2]}, {extent[3], extent[4], extent[5]}};
	return t.as_tensor(py::cast(rows), py::arg("dtype") = t.dtype(Dtype::Float32), py::arg("device") = t.cuda_device_kw);
}

std::pair<Dev, Dev> WarpField::PrecomputeAnchorsAndWeights(const Mesh& mesh) const {
	if (warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod::FIXED_NODE_COVERAGE)
		return compute_anchors_euclidean(mesh.positions, nodes, Dev{}, node_coverage, anchor_count, minimum_valid_anchor_count);
	return compute_anchors_euclidean(mesh.positions, nodes, node_coverage_weights, 0.f, anchor_count, minimum_valid_anchor_count);
}

// both WarpMesh overloads (WarpField.cpp:101-146): anchors == None computes them
py::object WarpField::WarpMesh(const py::handle& input_mesh, const py::handle& anchors_in, const py::handle& anchor_weights_in,
                               bool disable_neighbor_thresholding, const py::handle& extrinsics) const {
	const Mesh mesh = as_mesh(input_mesh);
	Dev anchors, weights;
	if (anchors_in.is_none()) {
		std::tie(anchors, weights) = PrecomputeAnchorsAndWeights(mesh);
	} else {
		anchors = as_cuda(anchors_in, Dtype::Int32);
		weights = as_cuda(anchor_weights_in, Dtype::Float32);
	}
	const int min_valid = disable_neighbor_thresholding ? 0 : minimum_valid_anchor_count;
	auto [positions, normals] = warp_points_normals(mesh.positions, mesh.normals, nodes, node_rotations, node_translations, anchors, weights, min_valid, extrinsics);
	return make_mesh(positions, normals, mesh.triangles);
}

// WarpField.cpp:151-159
void WarpField::ResetRotations() {
	check(nnrt_b200_fill_identity_rotations(node_rotations.data<float>(), node_count(), current_stream()));
}

// WarpField.cpp:170-183: the field keeps its own copy (the fitter updates rotations / translations in place)
void WarpField::SetNodeRotations(const Dev& rotations) {
	if (rotations.shape != node_rotations.shape) log_error("node_rotations must have shape " + node_rotations.shape_string() + ", got " + rotations.shape_string());
	node_rotations = clone(rotations);
}
void WarpField::SetNodeTranslations(const Dev& translations) {
	if (translations.shape != node_translations.shape)
		log_error("node_translations must have shape " + node_translations.shape_string() + ", got " + translations.shape_string());
	node_translations = clone(translations);
}
// WarpField.cpp:185-187
void WarpField::TranslateNodes(const Dev& deltas) {
	if (deltas.shape != node_translations.shape) log_error("node_translation_deltas must have shape " + node_translations.shape_string());
	check(nnrt_b200_add_scaled(node_translations.data<float>(), deltas.data<float>(), 1.f, deltas.numel(), node_translations.data<float>(), current_stream()));
}
// WarpField.cpp:189-194: R_new = R_old * dR
void WarpField::RotateNodes(const Dev& deltas) {
	if (deltas.shape != node_rotations.shape) log_error("node_rotation_deltas must have shape " + node_rotations.shape_string());
	Dev product = empty(node_rotations.shape, Dtype::Float32);
	check(nnrt_b200_matmul3d_3x3(node_rotations.data<float>(), deltas.data<float>(), (int32_t) node_count(), product.data<float>(), current_stream()));
	node_rotations = product;
}

// ---- HierarchicalGraphWarpField (cpp/geometry/HierarchicalGraphWarpField.cpp) ------------------------------------------------------
static Dev gather_rows(const Dev& source, const Dev& indices, Dtype dtype) {
	std::vector<int64_t> shape = source.shape;
	shape[0] = indices.size(0);
	Dev out = empty(shape, dtype);
	const int64_t row_words = source.size(0) > 0 ? source.numel() / source.size(0) : 0;
	check(nnrt_b200_gather_rows(source.pointer, source.size(0), (int32_t) row_words, indices.data<int64_t>(), indices.size(0), out.pointer, current_stream()));
	return out;
}

HierarchicalGraphWarpField::HierarchicalGraphWarpField(const py::handle& nodes_in, float node_coverage, bool threshold_nodes_by_distance, int anchor_count,
                                                       int minimum_valid_anchor_count, WarpNodeCoverageComputationMethod method, int layer_count,
                                                       int max_vertex_degree, const py::object& compute_layer_decimation_radius)
		: WarpField(nodes_in, node_coverage, threshold_nodes_by_distance, anchor_count, minimum_valid_anchor_count, method, 4) {
	if (layer_count < 1) log_error("layer_count must be a positive integer, got " + std::to_string(layer_count) + ".");  // .cpp:59-61
	std::vector<float> radii(layer_count, node_coverage);
	for (int i = 1; i < layer_count; i++)  // HierarchicalGraphWarpField.h:46-47: default radius(i) = (i + 1) * node_coverage
		radii[i] = compute_layer_decimation_radius.is_none() ? (float) (i + 1) * node_coverage : compute_layer_decimation_radius(i, node_coverage).cast<float>();
	const int64_t n = node_count();
	const int64_t edge_capacity = std::max<int64_t>(n * max_vertex_degree, 1);
	node_indices = empty({n}, Dtype::Int64);
	Dev all_edges = empty({edge_capacity, 2}, Dtype::Int32), all_edge_layers = empty({edge_capacity}, Dtype::Int8);
	std::vector<int32_t> counts(layer_count, 0);
	int64_t edge_count = 0;
	check(nnrt_b200_build_regularization_layers(nodes.data<float>(), (int32_t) n, radii.data(), layer_count, max_vertex_degree, node_indices.data<int64_t>(),
	                                            counts.data(), all_edges.data<int32_t>(), all_edge_layers.data<int8_t>(), &edge_count, current_stream()));
	layer_node_counts.assign(counts.begin(), counts.end());
	identity_order = layer_count == 1;
	if (identity_order) {
		original_to_virtual = node_indices;
	} else {
		original_to_virtual = empty({n}, Dtype::Int64);
		check(nnrt_b200_invert_permutation(node_indices.data<int64_t>(), n, original_to_virtual.data<int64_t>(), current_stream()));
	}
	// storage is the VIRTUAL order (layer 0 first): the order the edge list and the solver work in
	nodes = gather_rows(nodes, node_indices, Dtype::Float32);
	if (node_coverage_weights.defined()) node_coverage_weights = gather_rows(node_coverage_weights, node_indices, Dtype::Float32);
	edges = clone(narrow_rows(all_edges, 0, edge_count));
	// the reference's ARAP weight reads edge_layer_indices at NODE positions: kept at least node_count long
	edge_layer_indices = zeros({std::max(edge_count, n)}, Dtype::Int8);
	if (edge_count > 0) narrow_rows(edge_layer_indices, 0, edge_count).object.attr("copy_")(narrow_rows(all_edge_layers, 0, edge_count).object);
	first_layer_node_count = counts[0];
	Torch& t = Torch::get();
	layer_decimation_radii = wrap(t.as_tensor(py::cast(radii), py::arg("dtype") = t.dtype(Dtype::Float32), py::arg("device") = t.cuda_device_kw));
	int64_t first = 0, edge_first = edge_count;
	for (int i = 0; i < layer_count; i++) {
		RegularizationLayer layer{radii[i], narrow_rows(node_indices, first, counts[i]).object, narrow_rows(nodes, first, counts[i]).object, py::none()};
		if (i < layer_count - 1) {  // edges are stored coarsest pair first
			const int64_t layer_edges = (int64_t) counts[i] * std::min(max_vertex_degree, counts[i + 1]);
			edge_first -= layer_edges;
			layer.edges = narrow_rows(edges, edge_first, layer_edges).object;
		}
		regularization_layers.push_back(std::move(layer));
		first += counts[i];
	}
}

HierarchicalGraphWarpField::HierarchicalGraphWarpField(const py::handle& nodes_in, const py::handle& edges_in, const py::object& first_layer_count,
                                                       float node_coverage, bool threshold_nodes_by_distance, int anchor_count,
                                                       int minimum_valid_anchor_count, WarpNodeCoverageComputationMethod method,
                                                       int warp_node_coverage_neighbor_count, const py::handle& edge_layer_indices_in,
                                                       const py::handle& layer_decimation_radii_in)
		: WarpField(nodes_in, node_coverage, threshold_nodes_by_distance, anchor_count, minimum_valid_anchor_count, method, warp_node_coverage_neighbor_count) {
	const int64_t n = node_count();
	edges = edges_in.is_none() ? zeros({0, 2}, Dtype::Int32) : as_cuda(edges_in, Dtype::Int32);
	if (edges.dim() != 2 || edges.size(1) != 2) log_error("edges must have shape [E, 2], got " + edges.shape_string());
	const int64_t e = edges.size(0);
	first_layer_node_count = first_layer_count.is_none() ? n : first_layer_count.cast<int64_t>();
	edge_layer_indices = edge_layer_indices_in.is_none() ? zeros({std::max(e, n)}, Dtype::Int8) : as_cuda(edge_layer_indices_in, Dtype::Int8);
	Torch& t = Torch::get();
	if (layer_decimation_radii_in.is_none()) {
		std::vector<float> radii = {node_coverage, 2.f * node_coverage};
		layer_decimation_radii = wrap(t.as_tensor(py::cast(radii), py::arg("dtype") = t.dtype(Dtype::Float32), py::arg("device") = t.cuda_device_kw));
	} else {
		layer_decimation_radii = as_cuda(t.as_tensor(layer_decimation_radii_in, py::arg("dtype") = t.dtype(Dtype::Float32)), Dtype::Float32);
	}
	node_indices = wrap(t.module.attr("arange")(n, py::arg("device") = t.cuda_device_kw));
	original_to_virtual = node_indices;
	identity_order = true;
	if (first_layer_node_count < n) layer_node_counts = {(int) first_layer_node_count, (int) (n - first_layer_node_count)};
	else layer_node_counts = {(int) n};
}

Dev HierarchicalGraphWarpField::to_original(const Dev& virtual_tensor, bool use_virtual_ordering) const {
	if (use_virtual_ordering || identity_order) return virtual_tensor;
	return gather_rows(virtual_tensor, original_to_virtual, Dtype::Float32);
}
Dev HierarchicalGraphWarpField::to_virtual(const Dev& tensor, bool use_virtual_ordering) const {
	if (use_virtual_ordering || identity_order) return tensor;
	return gather_rows(tensor, node_indices, Dtype::Float32);
}

// HierarchicalGraphWarpField.cpp:228-256: anchors index the virtual order if asked (the fitter does), else the original order
std::pair<Dev, Dev> HierarchicalGraphWarpField::PrecomputeAnchorsAndWeights(const Mesh& mesh, bool use_virtual_ordering) const {
	if (reference_ordering_quirks && use_virtual_ordering && !identity_order &&
	    warp_node_coverage_computation_method != WarpNodeCoverageComputationMethod::FIXED_NODE_COVERAGE)  // virtual positions, original-order weights (.cpp:249-253)
		return compute_anchors_euclidean(mesh.positions, nodes, to_original(node_coverage_weights, false), 0.f, anchor_count, minimum_valid_anchor_count);
	if (reference_ordering_quirks && !use_virtual_ordering && !identity_order) {
		// the reference searches the original-order node array: same anchor sets as below, in ITS slot order (scan order of the K-nearest search)
		const Dev original_nodes = to_original(nodes, false);
		if (warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod::FIXED_NODE_COVERAGE)
			return compute_anchors_euclidean(mesh.positions, original_nodes, Dev{}, node_coverage, anchor_count, minimum_valid_anchor_count);
		return compute_anchors_euclidean(mesh.positions, original_nodes, to_original(node_coverage_weights, false), 0.f, anchor_count, minimum_valid_anchor_count);
	}
	auto [anchors, weights] = WarpField::PrecomputeAnchorsAndWeights(mesh);
	if (use_virtual_ordering || identity_order) return {anchors, weights};
	Dev mapped = empty(anchors.shape, Dtype::Int32);
	check(nnrt_b200_remap_indices(anchors.data<int32_t>(), anchors.numel(), node_indices.data<int64_t>(), node_count(), mapped.data<int32_t>(), current_stream()));
	return {mapped, weights};
}
void HierarchicalGraphWarpField::RotateNodesOrdered(const Dev& deltas, bool use_virtual_ordering) {
	if (reference_ordering_quirks && use_virtual_ordering && !identity_order) {
		// new[i] = R_original[i] deltas_virtual[i], scattered to original index node_indices[i] = virtual slot i (.cpp:271-277)
		if (deltas.shape != node_rotations.shape) log_error("node_rotation_deltas must have shape " + node_rotations.shape_string());
		const Dev original = to_original(node_rotations, false);
		Dev product = empty(node_rotations.shape, Dtype::Float32);
		check(nnrt_b200_matmul3d_3x3(original.data<float>(), deltas.data<float>(), (int32_t) node_count(), product.data<float>(), current_stream()));
		node_rotations = product;
		return;
	}
	RotateNodes(to_virtual(deltas, use_virtual_ordering));
}

Dev HierarchicalGraphWarpField::anchors_to_virtual(const Dev& anchors) const {
	if (identity_order) return anchors;
	Dev mapped = empty(anchors.shape, Dtype::Int32);
	check(nnrt_b200_remap_indices(anchors.data<int32_t>(), anchors.numel(), original_to_virtual.data<int64_t>(), node_count(), mapped.data<int32_t>(),
	                              current_stream()));
	return mapped;
}

}  // namespace nnrt::geometry
