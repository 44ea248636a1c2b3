// nnrt.geometry.functional, second part: the point-cloud / sampling / comparison names the reference binds
// (cpp/pybind/geometry/functional/functional.cpp:94-157) and the dual-quaternion warps, in C++ over the C ABI.
#include "warp_field.h"

#include <cmath>

using namespace pybind11::literals;

namespace nnrt {

using namespace geometry;

namespace {

Dev point_positions(const py::object& cloud) {  // an Open3D-like point cloud (.point["positions"]) or a tensor
	py::object positions = py::hasattr(cloud, "point") ? py::object(cloud.attr("point")["positions"]) : cloud;
	Dev points = as_cuda(positions, Dtype::Float32);
	if (points.dim() != 2 || points.size(1) != 3) log_error("points must have shape [N, 3], got " + points.shape_string());
	return points;
}

// depth image (tensor or Open3D-like image with as_tensor()) as [H,W] f32 or uint16 (viewed as int16): returns the dtype code of the C ABI
int depth_tensor(const py::object& depth_in, Dev& depth) {
	py::object d = py::hasattr(depth_in, "as_tensor") ? py::object(depth_in.attr("as_tensor")()) : to_torch(depth_in);
	if (d.attr("dim")().cast<int>() == 3) d = d.attr("select")(2, 0);
	if (d.attr("dim")().cast<int>() != 2) log_error("depth must have shape [H, W] or [H, W, 1]");
	Torch& t = Torch::get();
	if (has_dtype(d, t.module.attr("uint16"))) {
		depth = as_cuda(d.attr("contiguous")().attr("view")(t.dtype(Dtype::Int16)), Dtype::Int16);
		return 1;
	}
	if (has_dtype(d, t.dtype(Dtype::Int16))) {
		depth = as_cuda(d, Dtype::Int16);
		return 1;
	}
	if (!has_dtype(d, t.dtype(Dtype::Float32))) log_error("depth must be UInt16 or Float32");
	depth = as_cuda(d, Dtype::Float32);
	return 0;
}

using GridEntry = int (*)(const float*, int64_t, float, float, int64_t*, int64_t*, void*);

// per-cell index selection (median / closest-to-mean): int64 indices into `points`, cells in order of their first point
py::object grid_subsample(GridEntry entry, const Dev& points, float cell, float offset) {
	Dev sample = empty({points.size(0)}, Dtype::Int64);
	int64_t count = 0;
	check(entry(points.data<float>(), points.size(0), cell, offset, sample.data<int64_t>(), &count, current_stream()));
	return narrow_rows(sample, 0, count).object;
}
py::object mean_downsample(const Dev& points, float cell, float offset) {
	Dev means = empty({points.size(0), 3}, Dtype::Float32);
	int64_t count = 0;
	check(nnrt_b200_mean_grid_downsample_3d_points(points.data<float>(), points.size(0), cell, offset, means.data<float>(), &count, current_stream()));
	return narrow_rows(means, 0, count).object;
}
// sqrtf(2 * (radius * radius)) * 2 in fp32 (kernel/GeometrySamplingImpl.h:47-91)
float extended_cell(float radius) { return std::sqrt(2.f * (radius * radius)) * 2.f; }

Dev points_of(const py::object& points) {
	Dev p = as_cuda(points, Dtype::Float32);
	if (p.dim() != 2 || p.size(1) != 3) log_error("points must have shape [N, 3], got " + p.shape_string());
	return p;
}

}  // namespace

void bind_geometry_functional_more(py::module_& m) {
	// warp_point_cloud, both overloads (functional.cpp:95-107)
	m.def("warp_point_cloud",
	      [](const py::object& cloud, const py::object& nodes_in, const py::object& rotations, const py::object& translations, int anchor_count, float node_coverage,
	         int minimum_valid_anchor_count, const py::object& extrinsics) {
		      Dev points = point_positions(cloud), nodes = as_cuda(nodes_in, Dtype::Float32);
		      auto [anchors, weights] = compute_anchors_euclidean(points, nodes, Dev{}, node_coverage, anchor_count, minimum_valid_anchor_count);
		      return warp_points_normals(points, Dev{}, nodes, as_cuda(rotations, Dtype::Float32), as_cuda(translations, Dtype::Float32), anchors, weights,
		                                 minimum_valid_anchor_count, extrinsics).first.object;
	      }, "input_point_cloud"_a, "nodes"_a, "node_rotations"_a, "node_translations"_a, "anchor_count"_a, "node_coverage"_a, "minimum_valid_anchor_count"_a,
	      "extrinsics"_a = py::none());
	m.def("warp_point_cloud",
	      [](const py::object& cloud, const py::object& nodes_in, const py::object& rotations, const py::object& translations, const py::object& anchors,
	         const py::object& anchor_weights, int minimum_valid_anchor_count, const py::object& extrinsics) {
		      return warp_points_normals(point_positions(cloud), Dev{}, as_cuda(nodes_in, Dtype::Float32), as_cuda(rotations, Dtype::Float32),
		                                 as_cuda(translations, Dtype::Float32), as_cuda(anchors, Dtype::Int32), as_cuda(anchor_weights, Dtype::Float32),
		                                 minimum_valid_anchor_count, extrinsics).first.object;
	      }, "input_point_cloud"_a, "nodes"_a, "node_rotations"_a, "node_translations"_a, "anchors"_a, "anchor_weights"_a, "minimum_valid_anchor_count"_a,
	      "extrinsics"_a = py::none());
	// dual-quaternion blending (named by BASELINE.json; no reference implementation): supplied anchors
	m.def("warp_triangle_mesh_dual_quaternion",
	      [](const py::object& input_mesh, const py::object& nodes_in, const py::object& rotations, const py::object& translations, const py::object& anchors,
	         const py::object& anchor_weights, int minimum_valid_anchor_count, const py::object& extrinsics) {
		      const Mesh mesh = as_mesh(input_mesh);
		      auto [wp, wn] = warp_points_normals(mesh.positions, mesh.normals, as_cuda(nodes_in, Dtype::Float32), as_cuda(rotations, Dtype::Float32),
		                                          as_cuda(translations, Dtype::Float32), as_cuda(anchors, Dtype::Int32), as_cuda(anchor_weights, Dtype::Float32),
		                                          minimum_valid_anchor_count, extrinsics, true);
		      return make_mesh(wp, wn, mesh.triangles);
	      }, "input_mesh"_a, "nodes"_a, "node_rotations"_a, "node_translations"_a, "anchors"_a, "anchor_weights"_a, "minimum_valid_anchor_count"_a = 0,
	      "extrinsics"_a = py::none());
	m.def("warp_point_cloud_dual_quaternion",
	      [](const py::object& cloud, const py::object& nodes_in, const py::object& rotations, const py::object& translations, const py::object& anchors,
	         const py::object& anchor_weights, int minimum_valid_anchor_count, const py::object& extrinsics) {
		      return warp_points_normals(point_positions(cloud), Dev{}, as_cuda(nodes_in, Dtype::Float32), as_cuda(rotations, Dtype::Float32),
		                                 as_cuda(translations, Dtype::Float32), as_cuda(anchors, Dtype::Int32), as_cuda(anchor_weights, Dtype::Float32),
		                                 minimum_valid_anchor_count, extrinsics, true).first.object;
	      }, "input_point_cloud"_a, "nodes"_a, "node_rotations"_a, "node_translations"_a, "anchors"_a, "anchor_weights"_a, "minimum_valid_anchor_count"_a = 0,
	      "extrinsics"_a = py::none());

	// compute_ordered_point_cloud_normals (functional.cpp:114; NormalsOperations.cpp:73-93)
	m.def("compute_ordered_point_cloud_normals", [](const py::object& cloud, const std::vector<int64_t>& source_image_size) {
		if (source_image_size.size() != 2) log_error("Source image size must have two dimensions. Got " + std::to_string(source_image_size.size()) + ".");
		py::object positions = py::hasattr(cloud, "point") ? py::object(cloud.attr("point")["positions"]) : cloud;
		Dev points = as_cuda(to_torch(positions).attr("reshape")(-1, 3), Dtype::Float32);
		const int64_t h = source_image_size[0], w = source_image_size[1];
		if (points.size(0) != h * w)
			log_error("Point cloud point count (got " + std::to_string(points.size(0)) + ") must equal the multiple of dimensions (got " + std::to_string(h) +
			          " * " + std::to_string(w) + " = " + std::to_string(h * w) + ").");
		Dev normals = empty({h * w, 3}, Dtype::Float32);
		check(nnrt_b200_compute_ordered_point_cloud_normals(points.data<float>(), (int32_t) h, (int32_t) w, normals.data<float>(), current_stream()));
		return normals.object;
	}, "point_cloud"_a, "source_image_size"_a);

	// compute_point_to_plane_distances (functional.cpp:119-124): n1 . (v1 - v2), mesh against mesh or point cloud
	m.def("compute_point_to_plane_distances", [](const py::object& mesh1, const py::object& other) {
		const Mesh m1 = as_mesh(mesh1);
		if (!m1.normals.defined()) log_error("mesh1 needs to have vertex normals defined.");
		Dev v2;
		if (py::hasattr(other, "point") || py::isinstance(other, Torch::get().tensor_type) || py::isinstance(other, Torch::get().ndarray_type)) v2 = point_positions(other);
		else v2 = as_mesh(other).positions;
		if (v2.shape != m1.positions.shape) log_error("both point sets must have the same number of points");
		Dev out = empty({m1.positions.size(0)}, Dtype::Float32);
		check(nnrt_b200_point_to_plane_distances(m1.normals.data<float>(), m1.positions.data<float>(), v2.data<float>(), v2.size(0), out.data<float>(), current_stream()));
		return out.object;
	}, "mesh1"_a, "mesh2_or_point_cloud"_a);

	// unproject_raster_depth_without_filtering (functional.cpp:128-141) -> (points, mask)
	m.def("unproject_raster_depth_without_filtering",
	      [](const py::object& depth_in, const py::object& intrinsics, const py::object& extrinsics, float depth_scale, float depth_max, bool preserve_pixel_layout) -> py::tuple {
		      Dev depth;
		      const int code = depth_tensor(depth_in, depth);
		      const int64_t rows = depth.size(0), cols = depth.size(1);
		      const std::vector<double> K = host_matrix(intrinsics, 3, 3, false), ext = host_matrix(extrinsics, 4, 4, true);
		      Dev points = empty({rows * cols, 3}, Dtype::Float32), mask = empty({rows * cols}, Dtype::UInt8);
		      check(nnrt_b200_unproject_depth(depth.pointer, code, (int32_t) rows, (int32_t) cols, K.data(), ext.data(), depth_scale, depth_max, points.data<float>(),
		                                      mask.data<uint8_t>(), current_stream()));
		      py::object mask_bool = to_bool(mask);
		      if (preserve_pixel_layout)
			      return py::make_tuple(points.object.attr("reshape")(rows, cols, 3), mask_bool.attr("reshape")(rows, cols));
		      return py::make_tuple(points.object, mask_bool);
	      }, "depth"_a, "intrinsics"_a, "extrinsics"_a = py::none(), "depth_scale"_a = 1000.f, "depth_max"_a = 3.f, "preserve_pixel_layout"_a = false);

	// down / sub-sampling family (functional.cpp:143-157)
	m.def("mean_grid_downsample_3d_points", [](const py::object& points, float cell) { return mean_downsample(points_of(points), cell, 0.f); }, "points"_a,
	      "grid_cell_size"_a);
	m.def("closest_to_mean_grid_subsample_3d_points", [](const py::object& points, float cell) {
		return grid_subsample(nnrt_b200_closest_to_mean_grid_subsample_3d_points, points_of(points), cell, 0.f);
	}, "points"_a, "grid_cell_size"_a);
	m.def("median_grid_subsample_3d_points", [](const py::object& points, float cell) {
		return grid_subsample(nnrt_b200_median_grid_subsample_3d_points, points_of(points), cell, 0.f);
	}, "points"_a, "grid_cell_size"_a);
	// two passes on a grid of 2 sqrt(2) radius, the second shifted by half a cell (kernel/GeometrySamplingImpl.h:47-91)
	m.def("fast_mean_radius_downsample_3d_points", [](const py::object& points, float radius) {
		const float cell = extended_cell(radius);
		Dev first = wrap(py::object(mean_downsample(points_of(points), cell, 0.f).attr("contiguous")()));
		return mean_downsample(first, cell, 0.5f);
	}, "points"_a, "grid_cell_size"_a);
	m.def("fast_median_radius_subsample_3d_points", [](const py::object& points_in, float radius) {
		const float cell = extended_cell(radius);
		Dev points = points_of(points_in);
		Dev stage1 = wrap(py::object(grid_subsample(nnrt_b200_median_grid_subsample_3d_points, points, cell, 0.f).attr("contiguous")()));
		Dev picked = empty({stage1.size(0), 3}, Dtype::Float32);
		check(nnrt_b200_gather_rows(points.pointer, points.size(0), 3, stage1.data<int64_t>(), stage1.size(0), picked.pointer, current_stream()));
		Dev stage2 = wrap(py::object(grid_subsample(nnrt_b200_median_grid_subsample_3d_points, picked, cell, 0.5f).attr("contiguous")()));
		Dev out = empty({stage2.size(0)}, Dtype::Int64);
		check(nnrt_b200_gather_rows(stage1.pointer, stage1.size(0), 2, stage2.data<int64_t>(), stage2.size(0), out.pointer, current_stream()));  // int64 = 2 words
		return out.object;
	}, "points"_a, "grid_cell_size"_a);
}

}  // namespace nnrt
