// nnrt.geometry / nnrt.geometry.functional bindings: names, argument order and defaults of cpp/pybind/geometry/geometry.cpp:278-316 and
// cpp/pybind/geometry/functional/functional.cpp:51-157.
#include "warp_field.h"

using namespace pybind11::literals;

namespace nnrt {

using namespace geometry;

// The reference's compute_*_normals modify the mesh in place (mesh.SetTriangleNormals / SetVertexNormals): a dict mesh gets the key, an
// Open3D-like mesh its .triangle / .vertex map entry; a tuple is immutable and the caller uses the returned tensor.
static void store_mesh_attribute(const py::object& mesh, const char* group, const char* dict_key, const py::object& value) {
	if (py::isinstance<py::dict>(mesh)) mesh[py::str(dict_key)] = value;
	else if (!py::isinstance<py::tuple>(mesh) && !py::isinstance<py::list>(mesh) && py::hasattr(mesh, group)) mesh.attr(group)[py::str("normals")] = value;
}

static py::tuple pair_to_tuple(const std::pair<Dev, Dev>& p) { return py::make_tuple(p.first.object, p.second.object); }

static void bind_functional(py::module_& m) {
	m.def("compute_anchors_and_weights_euclidean_fixed_node_weight",
	      [](const py::object& points, const py::object& nodes, int anchor_count, int minimum_valid_anchor_count, float node_coverage_weight) {
		      return pair_to_tuple(compute_anchors_euclidean(as_cuda(points, Dtype::Float32), as_cuda(nodes, Dtype::Float32), Dev{}, node_coverage_weight,
		                                                     anchor_count, minimum_valid_anchor_count));
	      }, "points"_a, "nodes"_a, "anchor_count"_a, "minimum_valid_anchor_count"_a, "node_coverage_weight"_a);
	m.def("compute_anchors_and_weights_euclidean_variable_node_weight",
	      [](const py::object& points, const py::object& nodes, const py::object& node_coverage_weights, int anchor_count, int minimum_valid_anchor_count) {
		      return pair_to_tuple(compute_anchors_euclidean(as_cuda(points, Dtype::Float32), as_cuda(nodes, Dtype::Float32),
		                                                     as_cuda(node_coverage_weights, Dtype::Float32), 0.f, anchor_count, minimum_valid_anchor_count));
	      }, "points"_a, "nodes"_a, "node_coverage_weights"_a, "anchor_count"_a, "minimum_valid_anchor_count"_a);

	auto shortest_path = [](const py::object& points_in, const py::object& nodes_in, const py::object& weights_in, float node_coverage, const py::object& edges_in,
	                        int anchor_count) {
		Dev points = as_cuda(points_in, Dtype::Float32), nodes = as_cuda(nodes_in, Dtype::Float32);
		if (points.dim() < 2 || points.dim() > 3)  // WarpAnchorComputation.cpp:41-44
			log_error("`points` needs to have 2 or 3 dimensions. Got: " + std::to_string(points.dim()) + " dimensions.");
		if (points.size(-1) != 3) log_error("points must have shape [..., 3], got " + points.shape_string());
		if (anchor_count < 1) log_error("anchor_count needs to be greater than one. Got: " + std::to_string(anchor_count) + ".");
		py::object edges_tensor = to_torch(edges_in);
		if (!has_dtype(edges_tensor, Torch::get().dtype(Dtype::Int32)))  // o3c::AssertTensorDtype(edges, Int32), :169,197
			log_error("edges must be Int32, got " + py::str(edges_tensor.attr("dtype")).cast<std::string>());
		Dev edges = as_cuda(edges_tensor, Dtype::Int32);
		if (edges.dim() != 2 || edges.size(0) != nodes.size(0)) log_error("edges must have shape [node_count, degree], got " + edges.shape_string());
		Dev weights_nodes = as_cuda_or_none(weights_in, Dtype::Float32);
		if (weights_nodes.defined() && weights_nodes.numel() != nodes.size(0)) log_error("node_coverage_weights must have one entry per node");
		std::vector<int64_t> out_shape(points.shape.begin(), points.shape.end() - 1);
		out_shape.push_back(anchor_count);
		Dev anchors = empty(out_shape, Dtype::Int32), weights = empty(out_shape, Dtype::Float32);
		check(nnrt_b200_compute_anchors_shortest_path(points.data<float>(), points.numel() / 3, nodes.data<float>(), (int32_t) nodes.size(0),
		                                              weights_nodes.defined() ? weights_nodes.data<float>() : nullptr, node_coverage, edges.data<int32_t>(),
		                                              (int32_t) edges.size(1), anchor_count, anchors.data<int32_t>(), weights.data<float>(), current_stream()));
		return py::make_tuple(anchors.object, weights.object);
	};
	m.def("compute_anchors_and_weights_shortest_path_fixed_node_weight",
	      [shortest_path](const py::object& points, const py::object& nodes, const py::object& edges, int anchor_count, float node_coverage_weight) {
		      return shortest_path(points, nodes, py::none(), node_coverage_weight, edges, anchor_count);
	      }, "points"_a, "nodes"_a, "edges"_a, "anchor_count"_a, "node_coverage_weight"_a);
	m.def("compute_anchors_and_weights_shortest_path_variable_node_weight",
	      [shortest_path](const py::object& points, const py::object& nodes, const py::object& node_coverage_weights, const py::object& edges, int anchor_count) {
		      return shortest_path(points, nodes, node_coverage_weights, 0.f, edges, anchor_count);
	      }, "points"_a, "nodes"_a, "node_coverage_weights"_a, "edges"_a, "anchor_count"_a);

	// warp_triangle_mesh, both overloads (functional.cpp:77-93): the fifth / sixth arguments are (anchor_count, node_coverage) or (anchors, anchor_weights)
	m.def("warp_triangle_mesh",
	      [](const py::object& input_mesh, const py::object& nodes_in, const py::object& rotations, const py::object& translations, int anchor_count,
	         float node_coverage, bool threshold_nodes_by_distance, int minimum_valid_anchor_count, const py::object& extrinsics) {
		      const Mesh mesh = as_mesh(input_mesh);
		      Dev nodes = as_cuda(nodes_in, Dtype::Float32);
		      const int min_valid = threshold_nodes_by_distance ? minimum_valid_anchor_count : 0;
		      auto [anchors, weights] = compute_anchors_euclidean(mesh.positions, nodes, Dev{}, node_coverage, anchor_count, min_valid);
		      auto [wp, wn] = warp_points_normals(mesh.positions, mesh.normals, nodes, as_cuda(rotations, Dtype::Float32), as_cuda(translations, Dtype::Float32),
		                                          anchors, weights, min_valid, extrinsics);
		      return make_mesh(wp, wn, mesh.triangles);
	      }, "input_mesh"_a, "nodes"_a, "node_rotations"_a, "node_translations"_a, "anchor_count"_a, "node_coverage"_a,
	      "threshold_nodes_by_distance"_a = false, "minimum_valid_anchor_count"_a = 0, "extrinsics"_a = py::none());
	m.def("warp_triangle_mesh",
	      [](const py::object& input_mesh, const py::object& nodes_in, const py::object& rotations, const py::object& translations, const py::object& anchors,
	         const py::object& anchor_weights, bool threshold_nodes_by_distance, int minimum_valid_anchor_count, const py::object& extrinsics) {
		      const Mesh mesh = as_mesh(input_mesh);
		      const int min_valid = threshold_nodes_by_distance ? minimum_valid_anchor_count : 0;
		      auto [wp, wn] = warp_points_normals(mesh.positions, mesh.normals, as_cuda(nodes_in, Dtype::Float32), as_cuda(rotations, Dtype::Float32),
		                                          as_cuda(translations, Dtype::Float32), as_cuda(anchors, Dtype::Int32), as_cuda(anchor_weights, Dtype::Float32),
		                                          min_valid, extrinsics);
		      return make_mesh(wp, wn, mesh.triangles);
	      }, "input_mesh"_a, "nodes"_a, "node_rotations"_a, "node_translations"_a, "anchors"_a, "anchor_weights"_a,
	      "threshold_nodes_by_distance"_a = false, "minimum_valid_anchor_count"_a = 0, "extrinsics"_a = py::none());

	// compute_{triangle,vertex}_normals (functional.cpp:112-113): returns the normals (a tuple / dict mesh cannot be modified in place)
	m.def("compute_triangle_normals", [](const py::object& mesh_in, bool normalized) {
		const Mesh mesh = as_mesh(mesh_in);
		if (!mesh.triangles.defined()) log_error("Argument mesh needs to have both triangle vertex indices and vertex positions.");
		Dev normals = empty({mesh.triangles.size(0), 3}, Dtype::Float32);
		check(nnrt_b200_compute_triangle_normals(mesh.positions.data<float>(), mesh.positions.size(0), mesh.triangles.data<int64_t>(), mesh.triangles.size(0),
		                                         normalized ? 1 : 0, normals.data<float>(), current_stream()));
		store_mesh_attribute(mesh_in, "triangle", "triangle_normals", normals.object);
		return normals.object;
	}, "mesh"_a, "normalized"_a = true);
	m.def("compute_vertex_normals", [](const py::object& mesh_in, bool normalized) {
		const Mesh mesh = as_mesh(mesh_in);
		if (!mesh.triangles.defined()) log_error("Argument mesh needs to have both triangle vertex indices and vertex positions.");
		Dev triangle_normals = mesh.triangle_normals;
		if (!triangle_normals.defined()) {  // NormalsOperations.cpp:52-60: un-normalised triangle normals are computed when the mesh carries none
			triangle_normals = empty({mesh.triangles.size(0), 3}, Dtype::Float32);
			check(nnrt_b200_compute_triangle_normals(mesh.positions.data<float>(), mesh.positions.size(0), mesh.triangles.data<int64_t>(),
			                                         mesh.triangles.size(0), 0, triangle_normals.data<float>(), current_stream()));
		}
		Dev normals = empty(mesh.positions.shape, Dtype::Float32);
		check(nnrt_b200_compute_vertex_normals(mesh.triangles.data<int64_t>(), mesh.triangles.size(0), triangle_normals.data<float>(), mesh.positions.size(0),
		                                       normalized ? 1 : 0, normals.data<float>(), current_stream()));
		store_mesh_attribute(mesh_in, "vertex", "normals", normals.object);
		return normals.object;
	}, "mesh"_a, "normalized"_a = true);

	// mesh extraction from TSDF voxel blocks (cpp/geometry/VoxelBlockGrid.cpp:461-484 -> Open3D's kernel; csrc/mesh_extraction.cu)
	m.def("voxel_block_neighbors", [](const py::object& block_keys) {
		Dev keys = as_cuda(block_keys, Dtype::Int32);
		if (keys.dim() != 2 || keys.size(1) != 3) log_error("block_keys must have shape [B, 3], got " + keys.shape_string());
		Dev out = empty({keys.size(0), 27}, Dtype::Int32);
		check(nnrt_b200_voxel_block_neighbors(keys.data<int32_t>(), keys.size(0), out.data<int32_t>(), current_stream()));
		return out.object;
	}, "block_keys"_a);
	m.def("extract_triangle_mesh", [](const py::object& block_keys, const py::dict& block_values, float voxel_size, float weight_threshold) {
		Dev keys = as_cuda(block_keys, Dtype::Int32), tsdf = as_cuda(block_values["tsdf"], Dtype::Float32);
		if (tsdf.dim() != 4 || tsdf.size(1) != tsdf.size(2) || tsdf.size(1) != tsdf.size(3) || tsdf.size(0) != keys.size(0))
			log_error("tsdf must have shape [B,R,R,R] with one block per key, got " + tsdf.shape_string());
		py::object weight_tensor = to_torch(block_values["weight"]);
		Torch& t = Torch::get();
		int weight_code = 0;
		Dev weight;
		if (has_dtype(weight_tensor, t.module.attr("uint16")) || has_dtype(weight_tensor, t.dtype(Dtype::Int16))) {
			weight = as_cuda(has_dtype(weight_tensor, t.module.attr("uint16")) ? py::object(weight_tensor.attr("view")(t.dtype(Dtype::Int16))) : weight_tensor, Dtype::Int16);
			weight_code = 1;
		} else if (has_dtype(weight_tensor, t.dtype(Dtype::Float32))) {
			weight = as_cuda(weight_tensor, Dtype::Float32);
		} else {
			log_error("weight voxel attribute must be Float32 or UInt16");
		}
		if (weight.shape != tsdf.shape) log_error("weight must have the shape of tsdf");
		Dev color;
		if (block_values.contains("color") && !block_values["color"].is_none()) color = as_cuda(block_values["color"], Dtype::Float32);
		const int64_t b = keys.size(0), r = tsdf.size(1), voxels = b * r * r * r;
		void* stream = current_stream();
		Dev neighbors = empty({b, 27}, Dtype::Int32), structure = empty({voxels, 4}, Dtype::Int32), offsets = empty({voxels + 1}, Dtype::Int32);
		check(nnrt_b200_voxel_block_neighbors(keys.data<int32_t>(), b, neighbors.data<int32_t>(), stream));
		int64_t counts[2] = {0, 0};
		check(nnrt_b200_extract_triangle_mesh_prepare(b, neighbors.data<int32_t>(), tsdf.data<float>(), weight.pointer, weight_code, (int32_t) r, weight_threshold,
		                                              structure.data<int32_t>(), offsets.data<int32_t>(), counts, stream));
		Dev vertices = empty({counts[0], 3}, Dtype::Float32), normals = empty({counts[0], 3}, Dtype::Float32), triangles = empty({counts[1], 3}, Dtype::Int64), colors;
		if (color.defined()) colors = empty({counts[0], 3}, Dtype::Float32);
		check(nnrt_b200_extract_triangle_mesh_emit(keys.data<int32_t>(), b, neighbors.data<int32_t>(), tsdf.data<float>(), color.defined() ? color.data<float>() : nullptr,
		                                           (int32_t) r, voxel_size, structure.data<int32_t>(), offsets.data<int32_t>(), vertices.data<float>(),
		                                           normals.data<float>(), color.defined() ? colors.data<float>() : nullptr, triangles.data<int64_t>(), stream));
		py::object mesh = make_mesh(vertices, normals, triangles);
		if (color.defined()) mesh.attr("vertex")["colors"] = colors.object;
		return mesh;
	}, "block_keys"_a, "block_values"_a, "voxel_size"_a, "weight_threshold"_a = 3.0f);
}

void bind_geometry(py::module_& m) {
	py::enum_<AnchorComputationMethod>(m, "AnchorComputationMethod")
			.value("EUCLIDEAN", AnchorComputationMethod::EUCLIDEAN)
			.value("SHORTEST_PATH", AnchorComputationMethod::SHORTEST_PATH)
			.export_values();
	py::enum_<WarpNodeCoverageComputationMethod>(m, "WarpNodeCoverageComputationMethod")
			.value("FIXED_NODE_COVERAGE", WarpNodeCoverageComputationMethod::FIXED_NODE_COVERAGE)
			.value("MINIMAL_K_NEIGHBOR_NODE_DISTANCE", WarpNodeCoverageComputationMethod::MINIMAL_K_NEIGHBOR_NODE_DISTANCE)
			.export_values();

	py::class_<TriangleMesh>(m, "TriangleMesh", "Result holder standing in for open3d.t.geometry.TriangleMesh: .vertex / .triangle attribute maps.")
			.def_readonly("vertex", &TriangleMesh::vertex)
			.def_readonly("triangle", &TriangleMesh::triangle)
			.def_property_readonly("positions", [](const TriangleMesh& s) { return s.vertex["positions"]; })
			.def_property_readonly("normals", [](const TriangleMesh& s) -> py::object { return s.vertex.contains("normals") ? py::object(s.vertex["normals"]) : py::none(); })
			.def_property_readonly("triangles", [](const TriangleMesh& s) -> py::object { return s.triangle.contains("indices") ? py::object(s.triangle["indices"]) : py::none(); });

	py::class_<WarpField> graph_warp_field(
			m, "GraphWarpField",
			"A warp (motion) field represented by a graph: the motion of a point is a weighted function of the motions of the graph nodes close to it.");
	graph_warp_field.def(py::init([](const py::object& nodes, float node_coverage, bool threshold_nodes_by_distance, int anchor_count, int minimum_valid_anchor_count,
	                                 WarpNodeCoverageComputationMethod method, int neighbor_count) {
		                     return new WarpField(nodes, node_coverage, threshold_nodes_by_distance, anchor_count, minimum_valid_anchor_count, method, neighbor_count);
	                     }),
	                     "nodes"_a, "node_coverage"_a = 0.05, "threshold_nodes_by_distance"_a = false, "anchor_count"_a = 4, "minimum_valid_anchor_count"_a = 0,
	                     "warp_node_coverage_computation_method"_a = WarpNodeCoverageComputationMethod::MINIMAL_K_NEIGHBOR_NODE_DISTANCE,
	                     "warp_node_coverage_neighbor_count"_a = 4);
	graph_warp_field.def("get_warped_nodes", &WarpField::GetWarpedNodes);
	graph_warp_field.def("get_node_extent", &WarpField::GetNodeExtent);
	graph_warp_field.def("warp_mesh",
	                     [](const WarpField& self, const py::object& input_mesh, bool disable_neighbor_thresholding, const py::object& extrinsics) {
		                     return self.WarpMesh(input_mesh, py::none(), py::none(), disable_neighbor_thresholding, extrinsics);
	                     }, "input_mesh"_a, "disable_neighbor_thresholding"_a = true, "extrinsics"_a = py::none());
	graph_warp_field.def("warp_mesh",
	                     [](const WarpField& self, const py::object& input_mesh, const py::object& anchors, const py::object& anchor_weights,
	                        bool disable_neighbor_thresholding, const py::object& extrinsics) {
		                     return self.WarpMesh(input_mesh, anchors, anchor_weights, disable_neighbor_thresholding, extrinsics);
	                     }, "input_mesh"_a, "anchors"_a, "anchor_weights"_a, "disable_neighbor_thresholding"_a = true, "extrinsics"_a = py::none());
	graph_warp_field.def("precompute_anchors_and_weights",
	                     [](const WarpField& self, const py::object& input_mesh) { return pair_to_tuple(self.PrecomputeAnchorsAndWeights(as_mesh(input_mesh))); },
	                     "input_mesh"_a);
	// WarpField::Clone (WarpField.cpp:166-168): node positions are shared (immutable), the motion is copied
	graph_warp_field.def("clone", [](const WarpField& self) {
		auto* other = new WarpField(self);
		other->node_rotations = clone(self.node_rotations);
		other->node_translations = clone(self.node_translations);
		if (self.node_coverage_weights.defined()) other->node_coverage_weights = clone(self.node_coverage_weights);
		return other;
	});
	graph_warp_field.def("reset_rotations", &WarpField::ResetRotations);
	// WarpField::ApplyTransformations (WarpField.cpp:161-164): a new field at the warped node positions, motion reset
	graph_warp_field.def("apply_transformations", [](const WarpField& self) {
		return new WarpField(self.GetWarpedNodes(), self.node_coverage, self.threshold_nodes_by_distance_by_default, self.anchor_count,
		                     self.minimum_valid_anchor_count, self.warp_node_coverage_computation_method, self.warp_node_coverage_neighbor_count);
	});
	graph_warp_field.def("get_node_rotations", [](const WarpField& self) { return self.node_rotations.object; });
	graph_warp_field.def("get_node_translations", [](const WarpField& self) { return self.node_translations.object; });
	graph_warp_field.def("set_node_rotations", [](WarpField& self, const py::object& r) { self.SetNodeRotations(as_cuda(r, Dtype::Float32)); }, "node_rotations"_a);
	graph_warp_field.def("set_node_translations", [](WarpField& self, const py::object& t) { self.SetNodeTranslations(as_cuda(t, Dtype::Float32)); },
	                     "node_translations"_a);
	graph_warp_field.def("translate_nodes", [](WarpField& self, const py::object& d) { self.TranslateNodes(as_cuda(d, Dtype::Float32)); },
	                     "node_translation_deltas"_a);
	graph_warp_field.def("rotate_nodes", [](WarpField& self, const py::object& d) { self.RotateNodes(as_cuda(d, Dtype::Float32)); }, "node_rotation_deltas"_a);
	graph_warp_field.def_property_readonly("nodes", [](const WarpField& self) { return self.nodes.object; });
	graph_warp_field.def_property_readonly("node_rotations", [](const WarpField& self) { return self.node_rotations.object; });
	graph_warp_field.def_property_readonly("node_translations", [](const WarpField& self) { return self.node_translations.object; });
	graph_warp_field.def_property("node_coverage_weights",
	                              [](const WarpField& self) -> py::object { return self.node_coverage_weights.defined() ? self.node_coverage_weights.object : py::none(); },
	                              [](WarpField& self, const py::object& weights) {
		                              Dev w = as_cuda(weights, Dtype::Float32);
		                              if (w.numel() != self.node_count()) log_error("node_coverage_weights must have one entry per node");
		                              self.node_coverage_weights = w;
	                              });
	graph_warp_field.def_readonly("node_coverage", &WarpField::node_coverage);
	graph_warp_field.def_readonly("anchor_count", &WarpField::anchor_count);
	graph_warp_field.def_readonly("minimum_valid_anchor_count", &WarpField::minimum_valid_anchor_count);
	graph_warp_field.def_readonly("warp_node_coverage_computation_method", &WarpField::warp_node_coverage_computation_method);

	using H = HierarchicalGraphWarpField;
	py::class_<H::RegularizationLayer>(m, "RegularizationLayer")
			.def_readonly("decimation_radius", &H::RegularizationLayer::decimation_radius)
			.def_readonly("node_indices", &H::RegularizationLayer::node_indices)
			.def_readonly("node_positions", &H::RegularizationLayer::node_positions)
			.def_readonly("edges", &H::RegularizationLayer::edges);
	py::class_<H, WarpField> hierarchical(m, "HierarchicalGraphWarpField",
	                                      "Warp field with regularisation layers (cpp/geometry/HierarchicalGraphWarpField.h:35-96). Storage is in virtual order "
	                                      "(layer 0 first); accessors take use_virtual_ordering and default to the original node order.");
	hierarchical.def(py::init([](const py::object& nodes, float node_coverage, bool threshold, int anchor_count, int minimum_valid_anchor_count,
	                             WarpNodeCoverageComputationMethod method, int layer_count, int max_vertex_degree, const py::object& radius_function) {
		                 return new H(nodes, node_coverage, threshold, anchor_count, minimum_valid_anchor_count, method, layer_count, max_vertex_degree, radius_function);
	                 }),
	                 "nodes"_a, "node_coverage"_a = 0.05, "threshold_nodes_by_distance_by_default"_a = false, "anchor_count"_a = 4,
	                 "minimum_valid_anchor_count"_a = 0,
	                 "warp_node_coverage_computation_method"_a = WarpNodeCoverageComputationMethod::MINIMAL_K_NEIGHBOR_NODE_DISTANCE, "layer_count"_a = 4,
	                 "max_vertex_degree"_a = 4, "compute_layer_decimation_radius"_a = py::none());
	hierarchical.def_static("from_graph",
	                        [](const py::object& nodes, const py::object& edges, const py::object& first_layer_node_count, float node_coverage, bool threshold,
	                           int anchor_count, int minimum_valid_anchor_count, WarpNodeCoverageComputationMethod method, int neighbor_count,
	                           const py::object& edge_layer_indices, const py::object& layer_decimation_radii) {
		                        return new H(nodes, edges, first_layer_node_count, node_coverage, threshold, anchor_count, minimum_valid_anchor_count, method,
		                                     neighbor_count, edge_layer_indices, layer_decimation_radii);
	                        },
	                        "nodes"_a, "edges"_a = py::none(), "first_layer_node_count"_a = py::none(), "node_coverage"_a = 0.05,
	                        "threshold_nodes_by_distance"_a = false, "anchor_count"_a = 4, "minimum_valid_anchor_count"_a = 0,
	                        "warp_node_coverage_computation_method"_a = WarpNodeCoverageComputationMethod::MINIMAL_K_NEIGHBOR_NODE_DISTANCE,
	                        "warp_node_coverage_neighbor_count"_a = 4, "edge_layer_indices"_a = py::none(), "layer_decimation_radii"_a = py::none());
	hierarchical.def("get_regularization_level", [](const H& self, int i) {
		if (i < 0 || i >= (int) self.regularization_layers.size()) log_error("regularization level " + std::to_string(i) + " does not exist");
		return self.regularization_layers[i];
	}, "i_layer"_a);
	hierarchical.def("get_regularization_level_count", [](const H& self) { return (int) self.layer_node_counts.size(); });
	hierarchical.def("get_node_positions", [](const H& self, bool v) { return self.to_original(self.nodes, v).object; }, "use_virtual_ordering"_a = false);
	hierarchical.def("get_node_rotations", [](const H& self, bool v) { return self.to_original(self.node_rotations, v).object; }, "use_virtual_ordering"_a = false);
	hierarchical.def("get_node_translations", [](const H& self, bool v) { return self.to_original(self.node_translations, v).object; },
	                 "use_virtual_ordering"_a = false);
	hierarchical.def("get_node_coverage_weights", [](const H& self, bool v) -> py::object {
		if (self.reference_ordering_quirks && !v) return self.to_original(self.node_rotations, false).object;  // the reference returns node_rotations here (.cpp:222-224)
		return self.node_coverage_weights.defined() ? self.to_original(self.node_coverage_weights, v).object : py::none();
	}, "use_virtual_ordering"_a = false);
	hierarchical.def("get_warped_nodes", [](const H& self, bool v) { return self.to_original(wrap(self.GetWarpedNodes()), v).object; },
	                 "use_virtual_ordering"_a = false);
	hierarchical.def("get_edges", [](const H& self) { return self.edges.object; });
	hierarchical.def("get_virtual_node_indices", [](const H& self) { return self.node_indices.object; });
	hierarchical.def("get_edge_layer_indices", [](const H& self) { return self.edge_layer_indices.object; });
	hierarchical.def("get_layer_decimation_radii", [](const H& self) { return self.layer_decimation_radii.object; });
	hierarchical.def("set_node_rotations", [](H& self, const py::object& r, bool v) { self.SetNodeRotations(self.to_virtual(as_cuda(r, Dtype::Float32), v)); },
	                 "node_rotations"_a, "use_virtual_ordering"_a = false);
	hierarchical.def("set_node_translations", [](H& self, const py::object& t, bool v) { self.SetNodeTranslations(self.to_virtual(as_cuda(t, Dtype::Float32), v)); },
	                 "node_translations"_a, "use_virtual_ordering"_a = false);
	hierarchical.def("translate_nodes", [](H& self, const py::object& d, bool v) { self.TranslateNodes(self.to_virtual(as_cuda(d, Dtype::Float32), v)); },
	                 "node_translation_deltas"_a, "use_virtual_ordering"_a = false);
	hierarchical.def("rotate_nodes", [](H& self, const py::object& d, bool v) { self.RotateNodesOrdered(as_cuda(d, Dtype::Float32), v); },
	                 "node_rotation_deltas"_a, "use_virtual_ordering"_a = false);
	hierarchical.def("precompute_anchors_and_weights", [](const H& self, const py::object& mesh, bool v) {
		return pair_to_tuple(self.PrecomputeAnchorsAndWeights(as_mesh(mesh), v));
	}, "input_mesh"_a, "use_virtual_ordering"_a = false);
	hierarchical.def("warp_mesh", [](const H& self, const py::object& input_mesh, bool disable_neighbor_thresholding, const py::object& extrinsics) {
		return self.WarpMesh(input_mesh, py::none(), py::none(), disable_neighbor_thresholding, extrinsics);
	}, "input_mesh"_a, "disable_neighbor_thresholding"_a = true, "extrinsics"_a = py::none());
	hierarchical.def("warp_mesh", [](const H& self, const py::object& input_mesh, const py::object& anchors, const py::object& anchor_weights,
	                                 bool disable_neighbor_thresholding, const py::object& extrinsics, bool use_virtual_ordering) {
		Dev a = as_cuda(anchors, Dtype::Int32);
		// supplied anchors index the original order unless told otherwise (with reference_ordering_quirks always: WarpField::WarpMesh has no such flag)
		if (!use_virtual_ordering || self.reference_ordering_quirks) a = self.anchors_to_virtual(a);
		return self.WarpMesh(input_mesh, a.object, anchor_weights, disable_neighbor_thresholding, extrinsics);
	}, "input_mesh"_a, "anchors"_a, "anchor_weights"_a, "disable_neighbor_thresholding"_a = true, "extrinsics"_a = py::none(), "use_virtual_ordering"_a = false);
	hierarchical.def("clone", [](const H& self) {
		auto* other = new H(self);
		other->node_rotations = clone(self.node_rotations);
		other->node_translations = clone(self.node_translations);
		if (self.node_coverage_weights.defined()) other->node_coverage_weights = clone(self.node_coverage_weights);
		return other;
	});
	hierarchical.def_property_readonly("edges", [](const H& self) { return self.edges.object; });
	hierarchical.def_property_readonly("edge_layer_indices", [](const H& self) { return self.edge_layer_indices.object; });
	hierarchical.def_property_readonly("layer_decimation_radii", [](const H& self) { return self.layer_decimation_radii.object; });
	hierarchical.def_property_readonly("node_indices", [](const H& self) { return self.node_indices.object; });
	hierarchical.def_readonly("_identity_order", &H::identity_order);
	hierarchical.def_readwrite("reference_ordering_quirks", &H::reference_ordering_quirks);
	hierarchical.def_readonly("first_layer_node_count", &H::first_layer_node_count);
	hierarchical.def_readonly("layer_node_counts", &H::layer_node_counts);

	py::module_ functional = m.def_submodule("functional", "Stateless functions on geometry (cpp/pybind/geometry/functional/functional.cpp).");
	bind_functional(functional);
	bind_geometry_functional_more(functional);
	bind_tsdf(functional);
}

}  // namespace nnrt
