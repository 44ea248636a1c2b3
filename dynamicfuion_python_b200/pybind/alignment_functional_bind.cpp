// nnrt.alignment.functional (un-fused stages of the data term: cpp/alignment/functional/*.h, C++-only in the reference) and the two
// fragment shaders of nnrt.rendering (cpp/rendering/FlatEdgeShader.h, VertexColorShader.h), in C++ over the C ABI. The fitter itself runs
// the fused kernels and never materialises the intermediate tensors these return.
#include "warp_field.h"

using namespace pybind11::literals;

namespace nnrt {

using geometry::HierarchicalGraphWarpField;
using geometry::WarpField;
using geometry::WarpNodeCoverageComputationMethod;

void bind_alignment_functional(py::module_& m) {
	// WarpedSurfaceJacobians (cpp/alignment/functional/WarpedSurfaceJacobians.cpp:28-47) -> ([V,A,4], [V,A,3])
	m.def("warped_surface_jacobians", [](const py::object& canonical_mesh, const WarpField& warp_field, const py::object& anchors_in, const py::object& weights_in, bool) {
		const Mesh mesh = as_mesh(canonical_mesh);
		if (!mesh.normals.defined()) log_error("vertex_normals argument needs to be a tensor for a call to WarpedSurfaceJacobians");
		Dev anchors = as_cuda(anchors_in, Dtype::Int32), weights = as_cuda(weights_in, Dtype::Float32);
		const int64_t v = anchors.size(0), k = anchors.size(1);
		Dev jv = empty({v, k, 4}, Dtype::Float32), jn = empty({v, k, 3}, Dtype::Float32);
		check(nnrt_b200_warped_surface_jacobians(mesh.positions.data<float>(), mesh.normals.data<float>(), v, warp_field.nodes.data<float>(),
		                                         warp_field.node_rotations.data<float>(), anchors.data<int32_t>(), weights.data<float>(), (int32_t) k, jv.data<float>(),
		                                         jn.data<float>(), current_stream()));
		return py::make_tuple(jv.object, jn.object);
	}, "canonical_mesh"_a, "warp_field"_a, "warp_anchors"_a, "warp_anchor_weights"_a, "use_virtual_ordering"_a = true);
	// RasterizedSurfaceJacobians (cpp/alignment/functional/RasterizedSurfaceJacobians.cpp:30-52) -> ([H,W,3,9], [H,W,3,10])
	m.def("rasterized_surface_jacobians", [](const py::object& warped_mesh, const py::object& pixel_faces, const py::object& barycentrics, const py::object& ndc_K,
	                                         bool perspective_corrected) {
		const Mesh mesh = as_mesh(warped_mesh);
		if (!mesh.normals.defined() || !mesh.triangles.defined()) log_error("warped mesh needs vertex normals and triangle indices");
		Dev faces = as_cuda(pixel_faces, Dtype::Int64), bary = as_cuda(barycentrics, Dtype::Float32);
		if (faces.dim() != 3) log_error("pixel_faces must have shape [H,W,K], got " + faces.shape_string());
		const std::vector<double> K = host_matrix(ndc_K, 3, 3, false);
		const int64_t h = faces.size(0), w = faces.size(1), k = faces.size(2);
		Dev jv = empty({h, w, 3, 9}, Dtype::Float32), jn = empty({h, w, 3, 10}, Dtype::Float32);
		check(nnrt_b200_rasterized_surface_jacobians(mesh.positions.data<float>(), mesh.normals.data<float>(), mesh.triangles.data<int64_t>(), faces.data<int64_t>(),
		                                             bary.data<float>(), (int32_t) h, (int32_t) w, (int32_t) k, K.data(), perspective_corrected ? 1 : 0, jv.data<float>(),
		                                             jn.data<float>(), current_stream()));
		return py::make_tuple(jv.object, jn.object);
	}, "warped_mesh"_a, "pixel_faces"_a, "barycentric_coordinates"_a, "ndc_intrinsic_matrix"_a, "perspective_corrected_barycentric_coordinates"_a);
	// AssociateFacesWithAnchors (cpp/alignment/functional/AssociateFacesWithAnchors.cpp:29-37): compact table [F,3A,2] + counts [F]
	m.def("associate_faces_with_anchors", [](const py::object& face_vertex_indices, const py::object& warp_vertex_anchors) {
		Dev triangles = as_cuda(face_vertex_indices, Dtype::Int64), anchors = as_cuda(warp_vertex_anchors, Dtype::Int32);
		const int64_t f = triangles.size(0), a = anchors.size(1);
		Dev entries = empty({f, 3 * a, 2}, Dtype::Int32), counts = empty({f}, Dtype::Int32);
		check(nnrt_b200_associate_faces_with_anchors(triangles.data<int64_t>(), f, anchors.data<int32_t>(), (int32_t) a, entries.data<int32_t>(), counts.data<int32_t>(),
		                                             current_stream()));
		return py::make_tuple(entries.object, counts.object);
	}, "face_vertex_indices"_a, "warp_vertex_anchors"_a);
	// PixelVertexAnchorJacobiansAndNodeAssociations (cpp/alignment/functional/PixelVertexAnchorJacobians.cpp:29-56)
	m.def("pixel_vertex_anchor_jacobians_and_node_associations",
	      [](const py::object& rv_in, const py::object& rn_in, const py::object& wv_in, const py::object& wn_in, const py::object& point_map_vectors,
	         const py::object& rasterized_normals, const py::object& residual_mask, const py::object& pixel_faces, const py::object& face_vertex_indices,
	         const py::object& face_node_anchors, const py::object& face_node_anchor_counts, int node_count, float tukey_penalty_cutoff, int iteration_mode,
	         bool use_tukey_penalty) {
		      Dev rv = as_cuda(rv_in, Dtype::Float32), rn = as_cuda(rn_in, Dtype::Float32), wv = as_cuda(wv_in, Dtype::Float32), wn = as_cuda(wn_in, Dtype::Float32);
		      Dev pm = as_cuda(to_torch(point_map_vectors).attr("reshape")(-1, 3), Dtype::Float32), nl = as_cuda(to_torch(rasterized_normals).attr("reshape")(-1, 3), Dtype::Float32);
		      Dev mask = as_cuda(to_torch(residual_mask).attr("reshape")(-1), Dtype::UInt8);
		      const int64_t px = mask.size(0);
		      Dev faces = as_cuda(to_torch(pixel_faces).attr("reshape")(px, -1), Dtype::Int64);
		      Dev triangles = as_cuda(face_vertex_indices, Dtype::Int64), entries = as_cuda(face_node_anchors, Dtype::Int32), counts_in = as_cuda(face_node_anchor_counts, Dtype::Int32);
		      const int64_t a = wv.size(1), s = iteration_mode == 0 ? 6 : 3;
		      Dev jacobians = empty({px, 3 * a, s}, Dtype::Float32), counts = empty({px}, Dtype::Int32), node_counts = empty({node_count}, Dtype::Int32);
		      check(nnrt_b200_pixel_vertex_anchor_jacobians(rv.data<float>(), rn.data<float>(), wv.data<float>(), wn.data<float>(), pm.data<float>(), nl.data<float>(),
		                                                    mask.data<uint8_t>(), faces.data<int64_t>(), (int32_t) faces.size(1), triangles.data<int64_t>(), entries.data<int32_t>(),
		                                                    counts_in.data<int32_t>(), px, (int32_t) a, node_count, iteration_mode, use_tukey_penalty ? 1 : 0,
		                                                    tukey_penalty_cutoff, jacobians.data<float>(), counts.data<int32_t>(), node_counts.data<int32_t>(), current_stream()));
		      return py::make_tuple(jacobians.object, counts.object, node_counts.object);
	      },
	      "rasterized_vertex_position_jacobians"_a, "rasterized_vertex_normal_jacobians"_a, "warped_vertex_position_jacobians"_a, "warped_vertex_normal_jacobians"_a,
	      "point_map_vectors"_a, "rasterized_normals"_a, "residual_mask"_a, "pixel_faces"_a, "face_vertex_indices"_a, "face_node_anchors"_a,
	      "face_node_anchor_counts"_a, "node_count"_a, "tukey_penalty_cutoff"_a, "iteration_mode"_a = 0, "use_tukey_penalty"_a = false);
	// the ARAP term of one iteration as the reference materialises it (ComputeEdgeResiduals, ComputeDenseArapEdgeJacobians,
	// ComputeArapBlockSparseHessianApproximation, ComputeNegativeArapGradient)
	m.def("arap_term", [](const HierarchicalGraphWarpField& field, float arap_term_weight, bool use_huber_penalty, float huber_penalty_constant) {
		const int64_t e = field.edges.size(0), n = field.node_count();
		const bool fixed = field.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod::FIXED_NODE_COVERAGE;
		Dev residuals = empty({3 * e}, Dtype::Float32), jacobians = empty({e, 5}, Dtype::Float32), diagonal = empty({n, 6, 6}, Dtype::Float32),
		    wing = empty({e, 6, 6}, Dtype::Float32), coordinates = empty({e, 2}, Dtype::Int32), gradient = empty({6 * n}, Dtype::Float32);
		check(nnrt_b200_arap_term(field.edges.data<int32_t>(), e, fixed ? field.edge_layer_indices.data<int8_t>() : nullptr,
		                          fixed ? field.layer_decimation_radii.data<float>() : nullptr, fixed ? nullptr : field.node_coverage_weights.data<float>(),
		                          field.nodes.data<float>(), field.node_rotations.data<float>(), field.node_translations.data<float>(), (int32_t) n, arap_term_weight,
		                          use_huber_penalty ? 1 : 0, huber_penalty_constant, residuals.data<float>(), jacobians.data<float>(), diagonal.data<float>(),
		                          wing.data<float>(), coordinates.data<int32_t>(), gradient.data<float>(), current_stream()));
		py::dict out;
		out["edge_residuals"] = residuals.object;
		out["edge_jacobians"] = jacobians.object;
		out["hessian_diagonal_blocks"] = diagonal.object;
		out["upper_wing_blocks"] = wing.object;
		out["upper_wing_block_coordinates"] = coordinates.object;
		out["negative_gradient"] = gradient.object;
		return out;
	}, "warp_field"_a, "arap_term_weight"_a = 200.f, "use_huber_penalty"_a = false, "huber_penalty_constant"_a = 0.0001f);
}

// ---- shaders over rasteriser fragments (headless: [H,W,3] uint8 device tensors) ---------------------------------------------------
namespace {
struct FlatEdgeShader {  // cpp/rendering/FlatEdgeShader.h:27-52
	float pixel_line_width;
	std::vector<float> line_color;
};
struct VertexColorShader {};  // cpp/rendering/VertexColorShader.cpp:29-59

py::tuple colors_and_triangles(const py::handle& mesh) {
	if (py::isinstance<py::dict>(mesh)) {
		py::dict d = py::reinterpret_borrow<py::dict>(mesh);
		return py::make_tuple(d["colors"], d.contains("triangles") ? d["triangles"] : d["indices"]);
	}
	if (py::isinstance<py::tuple>(mesh) || py::isinstance<py::list>(mesh)) return py::make_tuple(mesh[py::int_(1)], mesh[py::int_(2)]);
	return py::make_tuple(mesh.attr("vertex")["colors"], mesh.attr("triangle")["indices"]);
}
}  // namespace

void bind_shaders(py::module_& m) {
	py::class_<FlatEdgeShader>(m, "FlatEdgeShader")
			.def(py::init([](float width, const std::vector<float>& color) {
				if (color.size() != 3) log_error("line_color must have three components");
				return new FlatEdgeShader{width, color};
			}), "pixel_line_width"_a, "line_color"_a)
			.def("set_pixel_line_width", [](FlatEdgeShader& s, float width) { s.pixel_line_width = width; }, "width"_a)
			.def("set_line_color", [](FlatEdgeShader& s, const std::vector<float>& color) {
				if (color.size() != 3) log_error("line_color must have three components");
				s.line_color = color;
			}, "color"_a)
			.def("get_ndc_line_width", [](const FlatEdgeShader& s, int height, int width) { return s.pixel_line_width / ((float) std::min(height, width) / 2.0f); },
			     "image_height"_a, "image_width"_a)
			.def_readonly("pixel_line_width", &FlatEdgeShader::pixel_line_width)
			.def("shade_meshes", [](const FlatEdgeShader& s, const py::object& pixel_face_indices, const py::object&, const py::object&,
			                         const py::object& pixel_face_distances, const py::object&) {
				Dev faces = as_cuda(pixel_face_indices, Dtype::Int64), distances = as_cuda(pixel_face_distances, Dtype::Float32);
				if (faces.dim() != 3 || distances.shape != faces.shape) log_error("pixel_face_indices and pixel_face_distances must have shape [H,W,K]");
				Dev pixels = empty({faces.size(0), faces.size(1), 3}, Dtype::UInt8);
				check(nnrt_b200_shade_edges_flat(faces.data<int64_t>(), distances.data<float>(), (int32_t) faces.size(0), (int32_t) faces.size(1), (int32_t) faces.size(2),
				                                 s.pixel_line_width, s.line_color.data(), pixels.data<uint8_t>(), current_stream()));
				return pixels.object;
			}, "pixel_face_indices"_a, "pixel_depths"_a, "pixel_barycentric_coordinates"_a, "pixel_face_distances"_a, "meshes"_a = py::none());
	py::class_<VertexColorShader>(m, "VertexColorShader")
			.def(py::init<>())
			.def("shade_meshes", [](const VertexColorShader&, const py::object& pixel_face_indices, const py::object&, const py::object& barycentrics,
			                         const py::object&, const py::object& meshes) {
				if (meshes.is_none() || py::len(meshes) == 0)
					log_error("VertexColorShader requires meshes argument to be filled with meshes that were used during rasterization resulting in its other "
					          "arguments, in order to retrieve and interpolate the vertex colors.");
				Torch& t = Torch::get();
				py::list per_mesh;
				for (const py::handle& mesh : meshes) {
					py::tuple ct = colors_and_triangles(mesh);
					per_mesh.append(as_cuda(ct[0], Dtype::Float32).object.attr("__getitem__")(as_cuda(ct[1], Dtype::Int64).object));  // [F,3,3] gather: plumbing
				}
				Dev face_vertex_colors = wrap(py::object(t.cat(per_mesh, 0).attr("contiguous")()));
				Dev faces = as_cuda(pixel_face_indices, Dtype::Int64), bary = as_cuda(barycentrics, Dtype::Float32);
				if (faces.dim() != 3 || bary.dim() != 4) log_error("pixel_face_indices must have shape [H,W,K] and pixel_barycentric_coordinates [H,W,K,3]");
				Dev pixels = empty({faces.size(0), faces.size(1), 3}, Dtype::UInt8);
				check(nnrt_b200_shade_vertex_colors(faces.data<int64_t>(), bary.data<float>(), (int32_t) faces.size(0), (int32_t) faces.size(1), (int32_t) faces.size(2),
				                                    face_vertex_colors.data<float>(), pixels.data<uint8_t>(), current_stream()));
				return pixels.object;
			}, "pixel_face_indices"_a, "pixel_depths"_a, "pixel_barycentric_coordinates"_a, "pixel_face_distances"_a, "meshes"_a = py::none());
}

}  // namespace nnrt
