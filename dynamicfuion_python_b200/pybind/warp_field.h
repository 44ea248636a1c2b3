// C++ host classes of the warp field behind the pybind11 module: nnrt::geometry::WarpField (bound as GraphWarpField,
// cpp/pybind/geometry/geometry.cpp:278-316; cpp/geometry/WarpField.h:29-171) and HierarchicalGraphWarpField
// (cpp/geometry/HierarchicalGraphWarpField.h:35-96). Device work goes through the C ABI; tensors are held as torch CUDA tensors.
#pragma once
#include "bridge.h"

namespace nnrt::geometry {

enum class AnchorComputationMethod { EUCLIDEAN = 0, SHORTEST_PATH = 1 };
enum class WarpNodeCoverageComputationMethod { FIXED_NODE_COVERAGE = 0, MINIMAL_K_NEIGHBOR_NODE_DISTANCE = 1 };

// What warp_mesh returns where the reference returns an open3d TriangleMesh: .vertex["positions"|"normals"], .triangle["indices"].
struct TriangleMesh {
	py::dict vertex, triangle;
};
py::object make_mesh(const Dev& positions, const Dev& normals, const Dev& triangles);

std::pair<Dev, Dev> compute_anchors_euclidean(const Dev& points, const Dev& nodes, const Dev& node_weights, float node_coverage, int anchor_count,
                                              int minimum_valid_anchor_count);
std::pair<Dev, Dev> warp_points_normals(const Dev& points, const Dev& normals, const Dev& nodes, const Dev& rotations, const Dev& translations,
                                        const Dev& anchors, const Dev& anchor_weights, int minimum_valid_anchor_count, const py::handle& extrinsics,
                                        bool dual_quaternion = false);

class WarpField {
public:
	WarpField(const py::handle& nodes, float node_coverage, bool threshold_nodes_by_distance, int anchor_count, int minimum_valid_anchor_count,
	          WarpNodeCoverageComputationMethod method, int warp_node_coverage_neighbor_count);
	virtual ~WarpField() = default;

	int64_t node_count() const { return nodes.size(0); }
	void RecomputeNodeCoverageWeights();
	py::object GetWarpedNodes() const;
	py::object GetNodeExtent() const;
	std::pair<Dev, Dev> PrecomputeAnchorsAndWeights(const Mesh& mesh) const;
	py::object WarpMesh(const py::handle& input_mesh, const py::handle& anchors, const py::handle& anchor_weights, bool disable_neighbor_thresholding,
	                    const py::handle& extrinsics) const;
	void ResetRotations();
	void SetNodeRotations(const Dev& rotations);
	void SetNodeTranslations(const Dev& translations);
	void TranslateNodes(const Dev& deltas);
	void RotateNodes(const Dev& deltas);

	Dev nodes, node_rotations, node_translations, node_coverage_weights;
	float node_coverage;
	bool threshold_nodes_by_distance_by_default;
	int anchor_count, minimum_valid_anchor_count;
	WarpNodeCoverageComputationMethod warp_node_coverage_computation_method;
	int warp_node_coverage_neighbor_count;
};

class HierarchicalGraphWarpField : public WarpField {
public:
	struct RegularizationLayer {
		float decimation_radius;
		py::object node_indices, node_positions, edges;
	};
	// builds the layers on the device (nnrt_b200_build_regularization_layers)
	HierarchicalGraphWarpField(const py::handle& nodes, float node_coverage, bool threshold_nodes_by_distance, int anchor_count, int minimum_valid_anchor_count,
	                           WarpNodeCoverageComputationMethod method, int layer_count, int max_vertex_degree, const py::object& compute_layer_decimation_radius);
	// explicit graph already in virtual order
	HierarchicalGraphWarpField(const py::handle& nodes, const py::handle& edges, const py::object& first_layer_node_count, float node_coverage,
	                           bool threshold_nodes_by_distance, int anchor_count, int minimum_valid_anchor_count, WarpNodeCoverageComputationMethod method,
	                           int warp_node_coverage_neighbor_count, const py::handle& edge_layer_indices, const py::handle& layer_decimation_radii);

	Dev to_original(const Dev& virtual_tensor, bool use_virtual_ordering) const;
	Dev to_virtual(const Dev& tensor, bool use_virtual_ordering) const;
	std::pair<Dev, Dev> PrecomputeAnchorsAndWeights(const Mesh& mesh, bool use_virtual_ordering) const;
	Dev anchors_to_virtual(const Dev& anchors) const;

	Dev edges, edge_layer_indices, layer_decimation_radii, node_indices, original_to_virtual;
	std::vector<int> layer_node_counts;
	std::vector<RegularizationLayer> regularization_layers;
	int64_t first_layer_node_count = 0;
	bool identity_order = true;
	// reproduce the reference's inconsistencies between its two node orders (see the Python mirror's class docstring and
	// HierarchicalGraphWarpField.cpp:222-224, 249-253, 271-277; WarpField.cpp:129-144). Off by default: both orders describe one field.
	bool reference_ordering_quirks = false;
	void RotateNodesOrdered(const Dev& deltas, bool use_virtual_ordering);
};

}  // namespace nnrt::geometry
