// nnrt.geometry.functional, third part: the non-rigid TSDF integration stages of geometry::NonRigidSurfaceVoxelBlockGrid
// (cpp/geometry/NonRigidSurfaceVoxelBlockGrid.cpp:32-148; bound at cpp/pybind/geometry/geometry.cpp:340-368) over a voxel block
// grid's own tensors. Open3D's hash map is absent, so the block keys / value tensors are passed explicitly (with Open3D present they
// come from grid.hashmap().key_tensor() and grid.attribute(name) through DLPack, zero copy).
#include "warp_field.h"

using namespace pybind11::literals;

namespace nnrt {

using namespace geometry;

namespace {

// a voxel attribute that the kernel updates in place: must already be a contiguous CUDA tensor (a converted copy would drop the update)
Dev voxel_attribute(const py::dict& values, const char* name, bool required) {
	if (!values.contains(name)) {
		if (required) log_error(std::string("block_values['") + name + "'] is required");
		return Dev();
	}
	py::object tensor = values[name];
	if (tensor.is_none()) return Dev();
	Torch& t = Torch::get();
	if (!py::isinstance(tensor, t.tensor_type) || !tensor.attr("is_cuda").cast<bool>() || !tensor.attr("is_contiguous")().cast<bool>())
		log_error(std::string("block_values['") + name + "'] must be a contiguous CUDA tensor (it is updated in place)");
	return wrap(tensor);
}

// depth image (tensor or Open3D-like image) as [H,W] f32 or 16-bit; returns the C ABI's dtype code
int depth_image(const py::object& depth_in, Dev& depth) {
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

py::object integrate(const py::object& block_indices, const py::object& block_keys, const py::dict& block_values, int block_resolution, float voxel_size,
                     const WarpField& warp_field, const py::object& depth_in, const py::object& color_in, const py::object& depth_normals,
                     const py::object& depth_intrinsics, const py::object& color_intrinsics, const py::object& extrinsics, float depth_scale, float depth_max,
                     float truncation_voxel_multiplier, bool full_scan) {
	Torch& t = Torch::get();
	Dev indices = as_cuda(block_indices, Dtype::Int64), keys = as_cuda(block_keys, Dtype::Int32);
	Dev tsdf = voxel_attribute(block_values, "tsdf", true), weight = voxel_attribute(block_values, "weight", true),
	    voxel_color = voxel_attribute(block_values, "color", false);
	if (!has_dtype(tsdf.object, t.dtype(Dtype::Float32)) || (voxel_color.defined() && !has_dtype(voxel_color.object, t.dtype(Dtype::Float32))))
		log_error("tsdf and color voxel attributes must be Float32");
	int weight_code;
	if (has_dtype(weight.object, t.dtype(Dtype::Float32)))
		weight_code = 0;
	else if (has_dtype(weight.object, t.module.attr("uint16")) || has_dtype(weight.object, t.dtype(Dtype::Int16)))
		weight_code = 1;
	else
		log_error("weight voxel attribute must be Float32 or UInt16");
	Dev depth;
	const int depth_code = depth_image(depth_in, depth);
	const int64_t rows = depth.size(0), cols = depth.size(1);

	Dev color;
	int color_code = 0;
	int64_t color_rows = 0, color_cols = 0;
	if (!color_in.is_none() && voxel_color.defined()) {
		py::object c = py::hasattr(color_in, "as_tensor") ? py::object(color_in.attr("as_tensor")()) : to_torch(color_in);
		if (c.attr("numel")().cast<int64_t>() != 0) {
			color_code = has_dtype(c, t.dtype(Dtype::Float32)) ? 0 : 1;
			color = as_cuda(c, color_code == 0 ? Dtype::Float32 : Dtype::UInt8);
			color_rows = color.size(0), color_cols = color.size(1);
		}
	}
	Dev normals = as_cuda(depth_normals, Dtype::Float32);
	if (normals.numel() != rows * cols * 3) log_error("depth_normals must have shape [H,W,3]");
	const std::vector<double> Kd = host_matrix(depth_intrinsics, 3, 3, false);
	const std::vector<double> Kc = host_matrix(color_intrinsics.is_none() ? depth_intrinsics : color_intrinsics, 3, 3, false);
	const std::vector<double> ext = host_matrix(extrinsics, 4, 4, true);
	const bool fixed = warp_field.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod::FIXED_NODE_COVERAGE;
	const float* coverage_weights = (fixed || !warp_field.node_coverage_weights.defined()) ? nullptr : warp_field.node_coverage_weights.data<float>();
	Dev cosines = empty({rows, cols}, Dtype::Float32);
	auto entry = full_scan ? nnrt_b200_integrate_non_rigid_full_scan : nnrt_b200_integrate_non_rigid;
	check(entry(indices.data<int64_t>(), indices.size(0), keys.data<int32_t>(), tsdf.data<float>(), weight.pointer, weight_code,
	            color.defined() ? voxel_color.data<float>() : nullptr, block_resolution, voxel_size, voxel_size * truncation_voxel_multiplier, depth.pointer, depth_code,
	            color.defined() ? color.pointer : nullptr, color_code, normals.data<float>(), (int32_t) rows, (int32_t) cols, (int32_t) color_rows, (int32_t) color_cols,
	            Kd.data(), Kc.data(), ext.data(), warp_field.nodes.data<float>(), warp_field.node_rotations.data<float>(), warp_field.node_translations.data<float>(),
	            (int32_t) warp_field.node_count(), coverage_weights, warp_field.node_coverage, warp_field.anchor_count, warp_field.minimum_valid_anchor_count,
	            depth_scale, depth_max, cosines.data<float>(), current_stream()));
	return cosines.object;
}

}  // namespace

void bind_tsdf(py::module_& m) {
	// NonRigidSurfaceVoxelBlockGrid.integrate_non_rigid's voxel kernel (NonRigidSurfaceVoxelBlockGrid.cpp:32-74) -> cos_voxel_ray_to_normal [H,W]
	m.def("integrate_non_rigid_voxel_blocks", &integrate, "block_indices"_a, "block_keys"_a, "block_values"_a, "block_resolution"_a, "voxel_size"_a,
	      "warp_field"_a, "depth"_a, "color"_a, "depth_normals"_a, "depth_intrinsics"_a, "color_intrinsics"_a = py::none(), "extrinsics"_a = py::none(),
	      "depth_scale"_a = 1000.f, "depth_max"_a = 3.f, "truncation_voxel_multiplier"_a = 2.f, "_full_scan"_a = false);

	// GetBoundingBoxesOfWarpedBlocks (NonRigidSurfaceVoxelBlockGrid.cpp:120-130) -> [B,6] f32
	m.def("bounding_boxes_of_warped_blocks",
	      [](const py::object& block_keys, const WarpField& warp_field, float voxel_size, int block_resolution, const py::object& extrinsics) {
		      Dev keys = as_cuda(block_keys, Dtype::Int32);
		      if (keys.dim() != 2 || keys.size(1) != 3) log_error("block_keys must have shape [B, 3], got " + keys.shape_string());
		      const std::vector<double> ext = host_matrix(extrinsics, 4, 4, true);
		      Dev boxes = empty({keys.size(0), 6}, Dtype::Float32);
		      check(nnrt_b200_bounding_boxes_of_warped_blocks(keys.data<int32_t>(), keys.size(0), warp_field.nodes.data<float>(), warp_field.node_rotations.data<float>(),
		                                                      warp_field.node_translations.data<float>(), (int32_t) warp_field.node_count(), warp_field.node_coverage,
		                                                      warp_field.anchor_count, warp_field.minimum_valid_anchor_count, voxel_size, block_resolution, ext.data(),
		                                                      boxes.data<float>(), current_stream()));
		      return boxes.object;
	      }, "block_keys"_a, "warp_field"_a, "voxel_size"_a, "block_resolution"_a, "extrinsics"_a = py::none());

	// GetAxisAlignedBoxesIntersectingSurfaceMask (NonRigidSurfaceVoxelBlockGrid.cpp:132-148) -> bool mask [B]
	m.def("axis_aligned_boxes_intersecting_surface_mask",
	      [](const py::object& boxes_in, const py::object& depth_in, const py::object& intrinsics, float depth_scale, float depth_max, int downsampling_factor,
	         float truncation_distance) {
		      Dev boxes = as_cuda(boxes_in, Dtype::Float32);
		      if (boxes.dim() != 2 || boxes.size(1) != 6) log_error("boxes must have shape [B, 6], got " + boxes.shape_string());
		      Dev depth;
		      const int code = depth_image(depth_in, depth);
		      const std::vector<double> K = host_matrix(intrinsics, 3, 3, false);
		      Dev mask = empty({boxes.size(0)}, Dtype::UInt8);
		      check(nnrt_b200_boxes_intersecting_surface_mask(boxes.data<float>(), boxes.size(0), K.data(), depth.pointer, code, (int32_t) depth.size(0),
		                                                      (int32_t) depth.size(1), depth_scale, depth_max, downsampling_factor, truncation_distance,
		                                                      mask.data<uint8_t>(), current_stream()));
		      return to_bool(mask);
	      }, "boxes"_a, "depth"_a, "intrinsics"_a, "depth_scale"_a, "depth_max"_a, "downsampling_factor"_a, "truncation_distance"_a);
}

}  // namespace nnrt
