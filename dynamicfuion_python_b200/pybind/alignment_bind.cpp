// nnrt.alignment: DeformableMeshToImageFitter (cpp/alignment/DeformableMeshToImageFitter.h:30-126; C++-only in the reference, bound
// here under its natural name with the constructor's argument names and defaults) as a C++ host class over the nnrt_b200_fitter_* C ABI.
#include "warp_field.h"

#include <algorithm>
#include <array>
#include <cstring>

using namespace pybind11::literals;

namespace nnrt {
namespace alignment {

enum class IterationMode { ALL = 0, TRANSLATION_ONLY = 1, ROTATION_ONLY = 2 };  // cpp/alignment/IterationMode.h:24-28

using geometry::HierarchicalGraphWarpField;
using geometry::WarpField;
using geometry::WarpNodeCoverageComputationMethod;

class DeformableMeshToImageFitter {
public:
	DeformableMeshToImageFitter(int max_iteration_count, std::vector<IterationMode> iteration_mode_sequence, float minimal_update_threshold,
	                            bool use_perspective_correction, float max_depth, bool use_tukey_penalty_for_depth_term, float tukey_penalty_cutoff_cm,
	                            float preconditioning_dampening_factor, float arap_term_weight, bool use_huber_penalty_for_arap_term,
	                            float huber_penalty_constant)
			: max_iteration_count(max_iteration_count), iteration_mode_sequence(std::move(iteration_mode_sequence)),
			  min_update_threshold(minimal_update_threshold) {
		if (preconditioning_dampening_factor < 0.0f || preconditioning_dampening_factor > 1.0f)  // DeformableMeshToImageFitter.cpp:79-82
			log_error("`preconditioning_dampening_factor` should be a small non-negative value between 0 and 1. Got " +
			          std::to_string(preconditioning_dampening_factor) + ".");
		if (this->iteration_mode_sequence.empty()) log_error("iteration_mode_sequence must not be empty");
		nnrt_b200_fitter_default_options(&options);
		options.max_iteration_count = max_iteration_count;
		options.use_perspective_correction = use_perspective_correction;
		options.max_depth = max_depth;
		options.use_tukey_penalty = use_tukey_penalty_for_depth_term;
		options.tukey_penalty_cutoff = tukey_penalty_cutoff_cm;
		options.preconditioning_dampening_factor = preconditioning_dampening_factor;
		options.arap_term_weight = arap_term_weight;
		options.use_huber_penalty = use_huber_penalty_for_arap_term;
		options.huber_penalty_constant = huber_penalty_constant;
	}
	DeformableMeshToImageFitter(const DeformableMeshToImageFitter&) = delete;
	DeformableMeshToImageFitter& operator=(const DeformableMeshToImageFitter&) = delete;
	~DeformableMeshToImageFitter() { release(); }

	// FitToImage, point-cloud overload (DeformableMeshToImageFitter.cpp:85-276)
	void FitToPoints(WarpField& warp_field, const py::handle& canonical_mesh, const py::handle& reference_points, const py::handle& reference_point_mask,
	                 const py::handle& intrinsic_matrix, const py::handle& extrinsic_matrix, const std::vector<int64_t>& rendering_image_size,
	                 const py::handle& anchors, const py::handle& anchor_weights, int iteration_count) {
		if (rendering_image_size.size() != 2) log_error("rendering_image_size should be a SizeVector of size 2. Got size " + std::to_string(rendering_image_size.size()) + ".");
		py::object points_object = py::hasattr(reference_points, "point") ? py::object(reference_points.attr("point")["positions"])
		                                                                   : py::reinterpret_borrow<py::object>(reference_points);
		Dev points = as_cuda(points_object, Dtype::Float32);
		const int height = (int) rendering_image_size[0], width = (int) rendering_image_size[1];
		if (points.numel() != (int64_t) height * width * 3) log_error("reference point cloud must hold one point per pixel of the rendering image, got " + points.shape_string());
		stream = current_stream();
		set_model(warp_field, canonical_mesh, height, width, anchors, anchor_weights);
		Dev mask = mask_or_none(reference_point_mask, (int64_t) height * width);
		check(nnrt_b200_fitter_set_reference_points(handle, points.data<float>(), mask.defined() ? mask.data<uint8_t>() : nullptr, stream));
		fit(warp_field, intrinsic_matrix, extrinsic_matrix, iteration_count);
		keep_reference = {points.object, mask.object};
	}

	// FitToImage, depth-image overload (.cpp:278-314): unprojection and masking happen on the device inside set_reference_depth
	void FitToDepth(WarpField& warp_field, const py::handle& canonical_mesh, const py::handle& reference_depth_image, const py::handle& reference_image_mask,
	                const py::handle& intrinsic_matrix, const py::handle& extrinsic_matrix, float depth_scale, const py::handle& anchors,
	                const py::handle& anchor_weights, int iteration_count) {
		py::object depth_object = py::hasattr(reference_depth_image, "as_tensor") ? py::object(reference_depth_image.attr("as_tensor")())
		                                                                           : to_torch(reference_depth_image);
		if (depth_object.attr("dim")().cast<int>() == 3) depth_object = depth_object.attr("select")(2, 0);
		if (depth_object.attr("dim")().cast<int>() != 2) log_error("reference depth image must have shape [H, W] or [H, W, 1]");
		Torch& t = Torch::get();
		int depth_code = 0;
		Dev depth;
		if (has_dtype(depth_object, t.module.attr("uint16"))) {
			depth = as_cuda(depth_object.attr("contiguous")().attr("view")(t.dtype(Dtype::Int16)), Dtype::Int16);
			depth_code = 1;
		} else {
			depth = as_cuda(depth_object, Dtype::Float32);
		}
		const int height = (int) depth.size(0), width = (int) depth.size(1);
		stream = current_stream();
		set_model(warp_field, canonical_mesh, height, width, anchors, anchor_weights);
		Dev mask = mask_or_none(reference_image_mask, (int64_t) height * width);
		const std::vector<double> K = host_matrix(intrinsic_matrix, 3, 3, false);
		check(nnrt_b200_fitter_set_reference_depth(handle, depth.pointer, depth_code, mask.defined() ? mask.data<uint8_t>() : nullptr, depth_scale, K.data(),
		                                           stream));
		fit(warp_field, intrinsic_matrix, extrinsic_matrix, iteration_count);
		keep_reference = {depth.object, mask.object};
	}

	void CheckStatus() {
		require_handle();
		check(nnrt_b200_fitter_status(handle, current_stream()));
	}

	py::dict Diagnostics() {
		require_handle();
		const int64_t n = shape[2], h = shape[5], w = shape[6];
		Dev hessian = empty({n, 6, 6}, Dtype::Float32), gradient = empty({n, 6}, Dtype::Float32), deltas = empty({n, 6}, Dtype::Float32);
		Dev faces = empty({h, w}, Dtype::Int64), residuals = empty({h, w}, Dtype::Float32), mask = empty({h, w}, Dtype::UInt8);
		check(nnrt_b200_fitter_get_diagnostics(handle, hessian.data<float>(), gradient.data<float>(), deltas.data<float>(), faces.data<int64_t>(),
		                                       residuals.data<float>(), mask.data<uint8_t>(), current_stream()));
		py::dict out;
		out["hessian_diagonal"] = hessian.object;
		out["negative_gradient"] = gradient.object;
		out["deltas"] = deltas.object;
		out["pixel_face_indices"] = faces.object;
		out["residuals"] = residuals.object;
		out["residual_mask"] = to_bool(mask);
		return out;
	}

	// one iteration outside the graph with CUDA events between the stages: ms for [warp_vertices, face_raster, band_raster, assemble,
	// build_system, stem_eliminate, corner_solve_update]
	std::vector<float> ProfileIteration(WarpField& warp_field, const py::handle& intrinsic_matrix, const py::handle& extrinsic_matrix) {
		require_handle();
		const std::vector<double> K = host_matrix(intrinsic_matrix, 3, 3, false), extrinsics = host_matrix(extrinsic_matrix, 4, 4, true);
		std::vector<float> out(7, 0.f);
		check(nnrt_b200_fitter_profile_iteration(handle, warp_field.node_rotations.data<float>(), warp_field.node_translations.data<float>(), K.data(),
		                                         extrinsics.data(), out.data(), current_stream()));
		return out;
	}

	int LaunchesPerIteration() {
		require_handle();
		return nnrt_b200_fitter_launches_per_iteration(handle);
	}
	int64_t WorkspaceBytes() const { return handle ? nnrt_b200_fitter_workspace_bytes(handle) : 0; }

	// Pixel-sharded single-frame mode (no reference counterpart; include/nnrt_b200.h "pixel-sharded"): this process fits the image
	// rows of its stripes (stripe s -> rank s % world_size), the per-node sums of the ranks meet inside the stem kernel over NVLink
	// peer memory. `group` is a torch.distributed process group (None = the default one): it carries the 64-byte IPC handles once, when
	// the workspace exists, and nothing afterwards. world_size 1 turns the mode off. (The all-reduce form of the exchange lives in the
	// ctypes front end only.)
	void ShardPixels(int rank, int world_size, const py::object& group, int stripe_rows) {
		if (world_size > 1 && (rank < 0 || rank >= world_size)) log_error("rank " + std::to_string(rank) + " outside world of " + std::to_string(world_size));
		if (stripe_rows <= 0 || stripe_rows % 8) log_error("stripe_rows must be a positive multiple of 8 (the assembly tile height), got " + std::to_string(stripe_rows));
		shard_rank = rank; shard_world = world_size > 1 ? world_size : 1; shard_stripe_rows = stripe_rows; shard_group = group;
		peer_ready = false;
		if (handle) {
			check(nnrt_b200_fitter_disable_peer_reduce(handle));
			if (shard_world == 1) check(nnrt_b200_fitter_set_pixel_stripes(handle, 8, 0, 1));
		}
	}

	int max_iteration_count;
	std::vector<IterationMode> iteration_mode_sequence;
	float min_update_threshold;  // the reference never updates `maximum_update` (.cpp:100,111): unused there as well

private:
	void release() {
		if (handle) nnrt_b200_fitter_destroy(handle);
		handle = nullptr;
	}
	void require_handle() const {
		if (!handle) log_error("DeformableMeshToImageFitter: no fit has been run yet");
	}
	static Dev mask_or_none(const py::handle& mask, int64_t pixel_count) {
		if (mask.is_none()) return Dev{};
		py::object tensor = to_torch(mask);
		if (tensor.attr("dim")().cast<int>() != 1) tensor = tensor.attr("reshape")(-1);
		Dev m = as_cuda(tensor, Dtype::UInt8);
		if (m.numel() != pixel_count) log_error("reference mask must have one entry per pixel, got " + m.shape_string());
		return m;
	}
	void ensure(const std::array<int64_t, 7>& wanted) {
		if (handle && wanted == shape) return;
		const bool same_problem = handle && std::equal(wanted.begin() + 2, wanted.end(), shape.begin() + 2);
		if (same_problem && wanted[0] <= capacity[0] && wanted[1] <= capacity[1]) {
			// a canonical mesh re-extracted every frame changes its counts: the workspace is kept, only the launch sizes change
			check(nnrt_b200_fitter_set_mesh_size(handle, wanted[0], wanted[1]));
			shape = wanted;
			mesh_key.clear();
			return;
		}
		// a mesh that outgrew the workspace gets headroom; the first workspace of a problem is exact
		capacity = same_problem ? std::array<int64_t, 2>{wanted[0] + wanted[0] / 2, wanted[1] + wanted[1] / 2} : std::array<int64_t, 2>{wanted[0], wanted[1]};
		release();
		check(nnrt_b200_fitter_create(&options, capacity[0], capacity[1], (int32_t) wanted[2], wanted[3], (int32_t) wanted[4], (int32_t) wanted[5],
		                              (int32_t) wanted[6], &handle));
		if (capacity[0] != wanted[0] || capacity[1] != wanted[1]) check(nnrt_b200_fitter_set_mesh_size(handle, wanted[0], wanted[1]));
		shape = wanted;
		graph_key.clear();
		mesh_key.clear();
		peer_ready = false;  // a new workspace: its exchange buffer has not been mapped by the peers
		std::vector<int32_t> modes;
		for (IterationMode m : iteration_mode_sequence) modes.push_back((int32_t) m);
		check(nnrt_b200_fitter_set_iteration_modes(handle, modes.data(), (int32_t) modes.size()));
	}

	// identity of a graph tensor for the upload cache: pointer, torch version counter (in-place writes) and leading size
	static void append_key(std::vector<int64_t>& key, const Dev& d) {
		if (!d.defined()) { key.insert(key.end(), {0, 0, 0}); return; }
		key.push_back((int64_t) reinterpret_cast<uintptr_t>(d.pointer));
		key.push_back(d.object.attr("_version").cast<int64_t>());
		key.push_back(d.dim() > 0 ? d.size(0) : 0);
	}

	void set_model(WarpField& warp_field, const py::handle& canonical_mesh, int height, int width, const py::handle& anchors_in, const py::handle& weights_in) {
		const Mesh mesh = as_mesh(canonical_mesh);
		if (!mesh.normals.defined()) log_error("Input warped mesh needs to have normals defined; it doesn't.");  // .cpp:346-348
		if (!mesh.triangles.defined()) log_error("Argument mesh needs to have both triangle vertex indices and vertex positions.");
		auto* hierarchical = dynamic_cast<HierarchicalGraphWarpField*>(&warp_field);
		Dev anchors, weights;
		if (anchors_in.is_none()) {  // warp_field.PrecomputeAnchorsAndWeights(canonical_mesh, true), .cpp:96-97
			std::tie(anchors, weights) = hierarchical ? hierarchical->PrecomputeAnchorsAndWeights(mesh, true) : warp_field.PrecomputeAnchorsAndWeights(mesh);
		} else {
			anchors = as_cuda(anchors_in, Dtype::Int32);
			weights = as_cuda(weights_in, Dtype::Float32);
		}
		if (anchors.dim() != 2 || anchors.size(0) != mesh.positions.size(0) || anchors.shape != weights.shape)
			log_error("anchors / anchor_weights must be [vertex_count, anchor_count]");
		const int64_t n = warp_field.node_count();
		const int64_t e = hierarchical ? hierarchical->edges.size(0) : 0;
		ensure({mesh.positions.size(0), mesh.triangles.size(0), n, e, anchors.size(1), height, width});
		const bool fixed = warp_field.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod::FIXED_NODE_COVERAGE;
		Dev edge_layers, radii, coverage_weights, edges;
		int layer_count = 0;
		if (e > 0) {
			edges = hierarchical->edges;
			if (fixed) {
				edge_layers = hierarchical->edge_layer_indices;
				radii = hierarchical->layer_decimation_radii;
				layer_count = (int) radii.size(0);
				if (edge_layers.size(0) < std::max(e, n)) log_error("edge_layer_indices must have at least max(edge_count, node_count) entries");
			} else {
				coverage_weights = warp_field.node_coverage_weights;
			}
		}
		const int64_t base = hierarchical ? hierarchical->first_layer_node_count : n;
		// The graph belongs to the warp field and is the same frame after frame: its upload (host-side validation, one stream
		// synchronisation) is redone only when one of its tensors was replaced or written to.
		std::vector<int64_t> key = {base, layer_count};
		append_key(key, warp_field.nodes);
		append_key(key, edges);
		append_key(key, edge_layers);
		append_key(key, radii);
		append_key(key, coverage_weights);
		if (key != graph_key) {
			check(nnrt_b200_fitter_set_graph(handle, warp_field.nodes.data<float>(), e > 0 ? edges.data<int32_t>() : nullptr, (int32_t) base,
			                                 edge_layers.defined() ? edge_layers.data<int8_t>() : nullptr, radii.defined() ? radii.data<float>() : nullptr,
			                                 layer_count, coverage_weights.defined() ? coverage_weights.data<float>() : nullptr, stream));
			graph_key = key;
			keep_graph = {warp_field.nodes.object, edges.object, edge_layers.object, radii.object, coverage_weights.object};
		}
		// the canonical mesh is re-laid out into the workspace only when one of its tensors was replaced or written to
		std::vector<int64_t> mesh_identity;
		append_key(mesh_identity, mesh.positions);
		append_key(mesh_identity, mesh.normals);
		append_key(mesh_identity, mesh.triangles);
		append_key(mesh_identity, anchors);
		append_key(mesh_identity, weights);
		if (mesh_identity != mesh_key) {
			check(nnrt_b200_fitter_set_mesh(handle, mesh.positions.data<float>(), mesh.normals.data<float>(), mesh.triangles.data<int64_t>(),
			                                anchors.data<int32_t>(), weights.data<float>(), stream));
			mesh_key = mesh_identity;
			keep_mesh = {mesh.positions.object, mesh.normals.object, mesh.triangles.object, anchors.object, weights.object};
		}
	}

	void fit(WarpField& warp_field, const py::handle& intrinsic_matrix, const py::handle& extrinsic_matrix, int iteration_count) {
		const std::vector<double> K = host_matrix(intrinsic_matrix, 3, 3, false), extrinsics = host_matrix(extrinsic_matrix, 4, 4, true);
		const int iterations = iteration_count > 0 ? iteration_count : max_iteration_count;
		if (shard_world > 1) {
			check(nnrt_b200_fitter_set_pixel_stripes(handle, shard_stripe_rows, shard_rank, shard_world));
			if (!peer_ready) enable_peer_reduce();
		}
		check(nnrt_b200_fitter_fit(handle, warp_field.node_rotations.data<float>(), warp_field.node_translations.data<float>(), K.data(), extrinsics.data(),
		                           iterations, stream));
	}

	// Collective over the group: every rank clears its exchange buffer, the CUDA IPC handles are gathered, every rank maps its peers.
	void enable_peer_reduce() {
		py::module_ dist = py::module_::import("torch.distributed");
		Torch::get().module.attr("cuda").attr("synchronize")();
		dist.attr("barrier")("group"_a = shard_group);  // no rank is still reading a buffer that is about to be cleared
		unsigned char mine[64];
		const int status = nnrt_b200_fitter_peer_exchange(handle, nullptr, nullptr, mine);
		py::list gathered;
		for (int r = 0; r < shard_world; r++) gathered.append(py::none());
		py::object own = status == NNRT_B200_OK ? py::object(py::bytes(reinterpret_cast<const char*>(mine), 64)) : py::object(py::none());
		dist.attr("all_gather_object")(gathered, own, "group"_a = shard_group);
		std::string blob;
		bool ok = status == NNRT_B200_OK;
		for (int r = 0; r < shard_world && ok; r++) {
			if (gathered[r].is_none()) { ok = false; break; }
			blob += gathered[r].cast<std::string>();
		}
		if (ok) ok = nnrt_b200_fitter_enable_peer_reduce_ipc(handle, shard_rank, shard_world, blob.data()) == NNRT_B200_OK;
		const std::string failure = ok ? std::string() : std::string(nnrt_b200_last_error());
		py::list verdicts;
		for (int r = 0; r < shard_world; r++) verdicts.append(py::none());
		dist.attr("all_gather_object")(verdicts, py::bool_(ok), "group"_a = shard_group);  // also the barrier before the first fit
		for (int r = 0; r < shard_world; r++)
			if (!verdicts[r].cast<bool>()) {
				nnrt_b200_fitter_disable_peer_reduce(handle);
				log_error("shard_pixels: rank " + std::to_string(r) + " could not map its peers' exchange buffers" + (ok ? std::string() : " (" + failure + ")") +
				          "; use the all-reduce form (dynamicfuion_python_b200.nnrt, collective=\"nccl\")");
			}
		peer_ready = true;
	}

	int shard_rank = 0, shard_world = 1, shard_stripe_rows = 32;
	py::object shard_group = py::none();
	bool peer_ready = false;
	nnrt_b200_fitter_options options;
	nnrt_b200_fitter* handle = nullptr;
	std::array<int64_t, 7> shape{};
	std::array<int64_t, 2> capacity{};  // vertex / face counts the workspace was created for
	std::vector<int64_t> graph_key, mesh_key;
	void* stream = nullptr;  // torch's current stream, looked up once per fit_to_image call
	std::vector<py::object> keep_graph, keep_mesh, keep_reference;  // inputs of work still queued on the stream
};

}  // namespace alignment

void bind_alignment(py::module_& m) {
	using namespace alignment;
	py::enum_<IterationMode>(m, "IterationMode")
			.value("ALL", IterationMode::ALL)
			.value("TRANSLATION_ONLY", IterationMode::TRANSLATION_ONLY)
			.value("ROTATION_ONLY", IterationMode::ROTATION_ONLY)
			.export_values();

	py::class_<DeformableMeshToImageFitter> fitter(m, "DeformableMeshToImageFitter");
	fitter.def(py::init<int, std::vector<IterationMode>, float, bool, float, bool, float, float, float, bool, float>(),
	           "max_iteration_count"_a = 100, "iteration_mode_sequence"_a = std::vector<IterationMode>{IterationMode::ALL}, "minimal_update_threshold"_a = 1e-6f,
	           "use_perspective_correction"_a = true, "max_depth"_a = 10.f, "use_tukey_penalty_for_depth_term"_a = false, "tukey_penalty_cutoff_cm"_a = 0.01f,
	           "preconditioning_dampening_factor"_a = 0.f, "arap_term_weight"_a = 200.f, "use_huber_penalty_for_arap_term"_a = false,
	           "huber_penalty_constant"_a = 0.0001f);
	// RGBD overload (.cpp:316-329): reference_image has .color / .depth (open3d RGBDImage) or is a (color, depth) pair
	fitter.def("fit_to_image",
	           [](DeformableMeshToImageFitter& self, WarpField& warp_field, const py::object& canonical_mesh, const py::object& reference_image,
	              const py::object& reference_image_mask, const py::object& intrinsic_matrix, const py::object& extrinsic_matrix, float depth_scale,
	              const py::object& anchors, const py::object& anchor_weights, int iteration_count) {
		           py::object depth = py::hasattr(reference_image, "depth") ? py::object(reference_image.attr("depth")) : py::object(reference_image[py::int_(1)]);
		           self.FitToDepth(warp_field, canonical_mesh, depth, reference_image_mask, intrinsic_matrix, extrinsic_matrix, depth_scale, anchors, anchor_weights,
		                           iteration_count);
	           },
	           "warp_field"_a, "canonical_mesh"_a, "reference_image"_a, "reference_image_mask"_a, "intrinsic_matrix"_a, "extrinsic_matrix"_a, "depth_scale"_a,
	           py::kw_only(), "anchors"_a = py::none(), "anchor_weights"_a = py::none(), "iteration_count"_a = 0);
	// depth-image overload (.cpp:278-314); the colour image is unused by the reference beyond being passed along and may be None
	fitter.def("fit_to_image",
	           [](DeformableMeshToImageFitter& self, WarpField& warp_field, const py::object& canonical_mesh, const py::object&, const py::object& reference_depth_image,
	              const py::object& reference_image_mask, const py::object& intrinsic_matrix, const py::object& extrinsic_matrix, float depth_scale,
	              const py::object& anchors, const py::object& anchor_weights, int iteration_count) {
		           self.FitToDepth(warp_field, canonical_mesh, reference_depth_image, reference_image_mask, intrinsic_matrix, extrinsic_matrix, depth_scale, anchors,
		                           anchor_weights, iteration_count);
	           },
	           "warp_field"_a, "canonical_mesh"_a, "reference_color_image"_a, "reference_depth_image"_a, "reference_image_mask"_a, "intrinsic_matrix"_a,
	           "extrinsic_matrix"_a, "depth_scale"_a, py::kw_only(), "anchors"_a = py::none(), "anchor_weights"_a = py::none(), "iteration_count"_a = 0);
	// point-cloud overload (.cpp:85-276)
	fitter.def("fit_to_image",
	           [](DeformableMeshToImageFitter& self, WarpField& warp_field, const py::object& canonical_mesh, const py::object&, const py::object& reference_point_cloud,
	              const py::object& reference_point_mask, const py::object& intrinsic_matrix, const py::object& extrinsic_matrix,
	              const std::vector<int64_t>& rendering_image_size, const py::object& anchors, const py::object& anchor_weights, int iteration_count) {
		           self.FitToPoints(warp_field, canonical_mesh, reference_point_cloud, reference_point_mask, intrinsic_matrix, extrinsic_matrix, rendering_image_size,
		                            anchors, anchor_weights, iteration_count);
	           },
	           "warp_field"_a, "canonical_mesh"_a, "reference_color_image"_a, "reference_point_cloud"_a, "reference_point_mask"_a, "intrinsic_matrix"_a,
	           "extrinsic_matrix"_a, "rendering_image_size"_a, py::kw_only(), "anchors"_a = py::none(), "anchor_weights"_a = py::none(), "iteration_count"_a = 0);
	fitter.def("shard_pixels", &DeformableMeshToImageFitter::ShardPixels, "rank"_a, "world_size"_a, "group"_a = py::none(), "stripe_rows"_a = 32);
	fitter.def("check_status", &DeformableMeshToImageFitter::CheckStatus,
	           "Waits for the stream and raises if the last fit overflowed its band queue or met a non-positive-definite block.");
	fitter.def("diagnostics", &DeformableMeshToImageFitter::Diagnostics);
	fitter.def("profile_iteration", &DeformableMeshToImageFitter::ProfileIteration, "warp_field"_a, "intrinsic_matrix"_a, "extrinsic_matrix"_a = py::none());
	fitter.def("launches_per_iteration", &DeformableMeshToImageFitter::LaunchesPerIteration);
	fitter.def("workspace_bytes", &DeformableMeshToImageFitter::WorkspaceBytes);
	fitter.def_readonly("max_iteration_count", &DeformableMeshToImageFitter::max_iteration_count);
	fitter.def_readonly("iteration_mode_sequence", &DeformableMeshToImageFitter::iteration_mode_sequence);
}

}  // namespace nnrt
