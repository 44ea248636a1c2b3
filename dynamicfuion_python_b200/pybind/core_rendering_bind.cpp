// nnrt.core (+ .linalg) and nnrt.rendering (+ .functional) bindings: cpp/pybind/core/core.cpp:25-46, cpp/pybind/core/linalg/linalg.cpp:25-31,
// cpp/pybind/rendering/rendering.cpp:28-44, cpp/pybind/rendering/functional/functional.cpp:35-61. The block-sparse arrowhead matrix and
// its solvers are C++-only in the reference (cpp/core/linalg/BlockSparseArrowheadMatrix.h:46-107) and are bound under their natural names.
#include "bridge.h"

#include <cmath>

using namespace pybind11::literals;

namespace nnrt {

namespace core::linalg {

// cpp/core/linalg/BlockSparseArrowheadMatrix.h:46-107: public data members kept; the int16 breadboard is not needed (wing rows are
// grouped through a CSR built from upper_wing_block_coordinates).
struct BlockSparseArrowheadMatrix {
	py::object upper_wing_blocks = py::none(), upper_wing_block_coordinates = py::none(), upper_wing_breadboard = py::none();
	py::object corner_upper_blocks = py::none(), corner_upper_block_coordinates = py::none();
	int arrow_base_block_index = 0, diagonal_block_count = 0;
	Dev diagonal_blocks;

	void SetDiagonalBlocks(const py::handle& blocks) {
		diagonal_blocks = as_cuda(blocks, Dtype::Float32);
		if (diagonal_blocks.dim() != 3 || diagonal_blocks.size(1) != 6 || diagonal_blocks.size(2) != 6)
			log_error("diagonal blocks must have shape [count, 6, 6], got " + diagonal_blocks.shape_string());
	}
	int GetBlockSize() const { return 6; }

	struct Arguments {
		Dev diagonal, wing, wing_coordinates, corner, corner_coordinates;
	};
	Arguments arguments() const {
		if (!diagonal_blocks.defined() || diagonal_blocks.size(0) != diagonal_block_count)
			log_error("diagonal blocks must be set after diagonal_block_count (SetDiagonalBlocks)");
		Arguments a;
		a.diagonal = diagonal_blocks;
		a.wing = as_cuda_or_none(upper_wing_blocks, Dtype::Float32);
		a.wing_coordinates = as_cuda_or_none(upper_wing_block_coordinates, Dtype::Int32);
		a.corner = as_cuda_or_none(corner_upper_blocks, Dtype::Float32);
		a.corner_coordinates = as_cuda_or_none(corner_upper_block_coordinates, Dtype::Int32);
		if (a.wing.defined() != a.wing_coordinates.defined() || (a.wing.defined() && a.wing.size(0) != a.wing_coordinates.size(0)))
			log_error("upper_wing_blocks and upper_wing_block_coordinates must be given together and have one coordinate pair per block");
		if (a.corner.defined() != a.corner_coordinates.defined() || (a.corner.defined() && a.corner.size(0) != a.corner_coordinates.size(0)))
			log_error("corner_upper_blocks and corner_upper_block_coordinates must be given together and have one coordinate pair per block");
		return a;
	}
};

template<typename T>
static const T* pointer_or_null(const Dev& d) { return d.defined() ? d.data<T>() : nullptr; }

static void bind(py::module_& m) {
	auto rodrigues = [](const py::object& vectors) {
		Dev v = as_cuda(vectors, Dtype::Float32);
		if (v.dim() != 2 || v.size(1) != 3) log_error("vectors must have shape [N, 3], got " + v.shape_string());
		Dev out = empty({v.size(0), 3, 3}, Dtype::Float32);
		check(nnrt_b200_rodrigues(v.data<float>(), (int32_t) v.size(0), out.data<float>(), current_stream()));
		return out.object;
	};
	m.def("AxisAngleVectorsToMatricesRodrigues", rodrigues, "vectors"_a);
	m.def("axis_angle_vectors_to_matrices_rodrigues", rodrigues, "vectors"_a);

	py::class_<BlockSparseArrowheadMatrix>(m, "BlockSparseArrowheadMatrix")
			.def(py::init<>())
			.def_readwrite("upper_wing_blocks", &BlockSparseArrowheadMatrix::upper_wing_blocks)
			.def_readwrite("upper_wing_block_coordinates", &BlockSparseArrowheadMatrix::upper_wing_block_coordinates)
			.def_readwrite("upper_wing_breadboard", &BlockSparseArrowheadMatrix::upper_wing_breadboard)
			.def_readwrite("corner_upper_blocks", &BlockSparseArrowheadMatrix::corner_upper_blocks)
			.def_readwrite("corner_upper_block_coordinates", &BlockSparseArrowheadMatrix::corner_upper_block_coordinates)
			.def_readwrite("arrow_base_block_index", &BlockSparseArrowheadMatrix::arrow_base_block_index)
			.def_readwrite("diagonal_block_count", &BlockSparseArrowheadMatrix::diagonal_block_count)
			.def("set_diagonal_blocks", &BlockSparseArrowheadMatrix::SetDiagonalBlocks, "diagonal_blocks"_a)
			.def("diagonal_blocks", [](const BlockSparseArrowheadMatrix& s) -> py::object { return s.diagonal_blocks.defined() ? s.diagonal_blocks.object : py::none(); })
			.def("stem_diagonal_blocks", [](const BlockSparseArrowheadMatrix& s) { return narrow_rows(s.diagonal_blocks, 0, s.arrow_base_block_index).object; })
			.def("corner_diagonal_blocks", [](const BlockSparseArrowheadMatrix& s) {
				return narrow_rows(s.diagonal_blocks, s.arrow_base_block_index, s.diagonal_block_count - s.arrow_base_block_index).object;
			})
			.def("get_block_size", &BlockSparseArrowheadMatrix::GetBlockSize);

	// SolveBlockSparseArrowheadCholesky(x, A, b) (cpp/core/linalg/SolveBlockSparseArrowheadCholesky.cpp:31-94); returns x
	m.def("solve_block_sparse_arrowhead_cholesky", [](const BlockSparseArrowheadMatrix& a, const py::object& b_in) {
		Dev b = as_cuda(b_in, Dtype::Float32);
		if (b.dim() != 1 || b.size(0) != (int64_t) a.diagonal_block_count * 6)
			log_error("Size of lhs tensor b must be {" + std::to_string(a.diagonal_block_count * 6) +
			          "}, matching the leading dimension of rhs matrix a, but is " + b.shape_string() + " instead.");
		const auto args = a.arguments();
		Dev x = empty(b.shape, Dtype::Float32);
		check(nnrt_b200_solve_block_sparse_arrowhead_cholesky(args.diagonal.data<float>(), a.diagonal_block_count, a.arrow_base_block_index,
		                                                      pointer_or_null<float>(args.wing), pointer_or_null<int32_t>(args.wing_coordinates),
		                                                      args.wing.defined() ? args.wing.size(0) : 0, pointer_or_null<float>(args.corner),
		                                                      pointer_or_null<int32_t>(args.corner_coordinates), args.corner.defined() ? args.corner.size(0) : 0,
		                                                      b.data<float>(), x.data<float>(), current_stream()));
		return x.object;
	}, "a"_a, "b"_a);
	// ComputeSchurComplementOfArrowStem_Dense (cpp/core/linalg/SchurComplement.cpp:32-77)
	m.def("compute_schur_complement_of_arrow_stem_dense", [](const BlockSparseArrowheadMatrix& a) {
		const auto args = a.arguments();
		const int64_t n = (int64_t) (a.diagonal_block_count - a.arrow_base_block_index) * 6;
		Dev out = empty({n, n}, Dtype::Float32);
		check(nnrt_b200_arrowhead_schur_complement(args.diagonal.data<float>(), a.diagonal_block_count, a.arrow_base_block_index, pointer_or_null<float>(args.wing),
		                                           pointer_or_null<int32_t>(args.wing_coordinates), args.wing.defined() ? args.wing.size(0) : 0,
		                                           pointer_or_null<float>(args.corner), pointer_or_null<int32_t>(args.corner_coordinates),
		                                           args.corner.defined() ? args.corner.size(0) : 0, out.data<float>(), current_stream()));
		return out.object;
	}, "a"_a);
	// SolveBlockDiagonalCholesky(X, A_blocks, B) (cpp/core/linalg/SolveBlockDiagonalGeneric.h:27-90); single right-hand side
	m.def("solve_block_diagonal_cholesky", [](const py::object& a_blocks, const py::object& b_in) {
		Dev blocks = as_cuda(a_blocks, Dtype::Float32), b = as_cuda(b_in, Dtype::Float32);
		if (blocks.dim() != 3 || blocks.size(1) != blocks.size(2) || b.numel() != blocks.size(0) * blocks.size(1))
			log_error("A_blocks must be [N,S,S] and B must have N*S entries");
		Dev x = empty(b.shape, Dtype::Float32);
		check(nnrt_b200_solve_block_diagonal_cholesky(blocks.data<float>(), b.data<float>(), (int32_t) blocks.size(0), (int32_t) blocks.size(1), x.data<float>(),
		                                              current_stream()));
		return x.object;
	}, "a_blocks"_a, "b"_a);
	// PreconditionDiagonalBlocks (cpp/core/linalg/PreconditionDiagonalBlocksImpl.h:29-51): in place, returns the tensor
	m.def("precondition_diagonal_blocks", [](const py::object& blocks_in, float dampening_factor) {
		Dev blocks = as_cuda(blocks_in, Dtype::Float32);
		if (!blocks.object.is(blocks_in)) log_error("blocks must be a contiguous float32 CUDA tensor (modified in place)");
		if (blocks.dim() != 3 || blocks.size(1) != blocks.size(2)) log_error("blocks must have shape [count, size, size], got " + blocks.shape_string());
		check(nnrt_b200_precondition_diagonal_blocks(blocks.data<float>(), (int32_t) blocks.size(0), (int32_t) blocks.size(1), dampening_factor, current_stream()));
		return blocks.object;
	}, "blocks"_a, "dampening_factor"_a);
}

}  // namespace core::linalg

// nnrt.core.KDTree (cpp/pybind/core/core.cpp:36-46): the index is the point array itself (brute-force scan from shared memory)
struct KdTree {
	Dev points;
};

void bind_core(py::module_& m) {
	py::module_ linalg = m.def_submodule("linalg", "Contains additional linear algebra functions not defined in open3d.core or its submodules.");
	core::linalg::bind(linalg);
	bind_linalg_blocks(linalg);
	m.def("matmul3d", [](const py::object& a_in, const py::object& b_in) {
		Dev a = as_cuda(a_in, Dtype::Float32), b = as_cuda(b_in, Dtype::Float32);
		if (a.dim() != 3 || b.dim() != 3 || a.size(0) != b.size(0) || a.size(2) != b.size(1))
			log_error("matmul3d: incompatible shapes " + a.shape_string() + " and " + b.shape_string());
		if (a.size(1) != 3 || a.size(2) != 3 || b.size(2) != 3) log_error("matmul3d: only [N,3,3] x [N,3,3] is on the dense-depth path of this build");
		Dev c = empty(a.shape, Dtype::Float32);
		check(nnrt_b200_matmul3d_3x3(a.data<float>(), b.data<float>(), (int32_t) a.size(0), c.data<float>(), current_stream()));
		return c.object;
	}, "array_of_matrices_a"_a, "array_of_matrices_b"_a);
	py::class_<KdTree>(m, "KDTree")
			.def(py::init([](const py::object& points) {
				Dev p = as_cuda(points, Dtype::Float32);
				if (p.dim() != 2 || p.size(1) != 3) log_error("KDTree (B200 build) indexes 3D points, got shape " + p.shape_string());
				return new KdTree{p};
			}), "points"_a)
			.def("find_k_nearest_to_points", [](const KdTree& self, const py::object& query_points, int32_t k, bool sort_output) {
				Dev q = as_cuda(query_points, Dtype::Float32);
				if (q.dim() != 2 || q.size(1) != 3)  // KdTree.cpp:58-63
					log_error("Query point array of shape " + q.shape_string() + " is incompatible to the set of points being indexed by the KD Tree "
					          "(reference points), which has shape " + self.points.shape_string() + ".");
				Dev indices = empty({q.size(0), k}, Dtype::Int32), distances = empty({q.size(0), k}, Dtype::Float32);
				check(nnrt_b200_find_k_nearest_to_points(self.points.data<float>(), (int32_t) self.points.size(0), q.data<float>(), q.size(0), k, sort_output ? 1 : 0,
				                                         indices.data<int32_t>(), distances.data<float>(), current_stream()));
				return py::make_tuple(indices.object, distances.object);
			}, "query_points"_a, "k"_a, "sort_output"_a = false);
}

// ---- rendering ------------------------------------------------------------------------------------------------------------------
static py::tuple face_vertices_of(const py::handle& mesh_in, const std::vector<double>& K, const std::vector<int64_t>& image_size, float near_clip, float far_clip) {
	const Mesh mesh = as_mesh(mesh_in);
	if (!mesh.triangles.defined()) log_error("Argument mesh needs to have both triangle vertex indices and vertex positions.");
	const int64_t f = mesh.triangles.size(0);
	Dev vertices = empty({f, 3, 3}, Dtype::Float32), mask = empty({f}, Dtype::UInt8);
	check(nnrt_b200_face_vertices_ndc(mesh.positions.data<float>(), mesh.positions.size(0), mesh.triangles.data<int64_t>(), f, K.data(), (int32_t) image_size[0],
	                                  (int32_t) image_size[1], near_clip, far_clip, vertices.data<float>(), mask.data<uint8_t>(), current_stream()));
	return py::make_tuple(vertices.object, to_bool(mask));
}

void bind_rendering(py::module_& m) {
	m.def("rasterize_ndc_triangles",
	      [](const py::object& ndc_face_vertices, const py::object& clipped_faces_mask, const std::vector<int64_t>& image_size, float blur_radius_pixels,
	         int faces_per_pixel, int bin_size, int max_faces_per_bin, bool perspective_correct, bool clip_barycentric, bool cull_back_faces) {
		      if (image_size.size() != 2) log_error("image_size should be a SizeVector of size 2. Got size " + std::to_string(image_size.size()) + ".");
		      Dev faces = as_cuda(ndc_face_vertices, Dtype::Float32);
		      if (faces.dim() != 3 || faces.size(1) != 3 || faces.size(2) != 3) log_error("ndc_face_vertices must have shape [F,3,3], got " + faces.shape_string());
		      Dev mask = as_cuda_or_none(clipped_faces_mask, Dtype::UInt8);
		      const int64_t h = image_size[0], w = image_size[1], k = faces_per_pixel;
		      Dev indices = empty({h, w, k}, Dtype::Int64), depths = empty({h, w, k}, Dtype::Float32), barycentrics = empty({h, w, k, 3}, Dtype::Float32),
		          distances = empty({h, w, k}, Dtype::Float32);
		      check(nnrt_b200_rasterize_ndc_triangles(faces.data<float>(), mask.defined() ? mask.data<uint8_t>() : nullptr, faces.size(0), (int32_t) h, (int32_t) w,
		                                              blur_radius_pixels, faces_per_pixel, bin_size, max_faces_per_bin, perspective_correct, clip_barycentric,
		                                              cull_back_faces, indices.data<int64_t>(), depths.data<float>(), barycentrics.data<float>(),
		                                              distances.data<float>(), current_stream()));
		      return py::make_tuple(indices.object, depths.object, barycentrics.object, distances.object);
	      },
	      "ndc_face_vertices"_a, "clipped_faces_mask"_a, "image_size"_a, "blur_radius_pixels"_a = 0.f, "faces_per_pixel"_a = 8, "bin_size"_a = -1,
	      "max_faces_per_bin"_a = -1, "perspective_correct_barycentric_coordinates"_a = false, "clip_barycentric_coordinates"_a = false,
	      "cull_back_faces"_a = true);

	py::module_ functional = m.def_submodule("functional", "Stateless rendering functions (cpp/pybind/rendering/functional/functional.cpp).");
	// several meshes: faces concatenated in order, plus the per-mesh face counts (cpp/rendering/functional/ExtractFaceVertices.cpp:87-124)
	functional.def("get_mesh_ndc_face_vertices_and_clip_mask",
	               [](const py::list& camera_space_meshes, const py::object& intrinsic_matrix, const std::vector<int64_t>& image_size, float near_clip, float far_clip) {
		               if (image_size.size() != 2) log_error("image_size should be a SizeVector of size 2. Got size " + std::to_string(image_size.size()) + ".");
		               const std::vector<double> K = host_matrix(intrinsic_matrix, 3, 3, false);
		               py::list vertices, masks;
		               std::vector<int64_t> counts;
		               for (const py::handle& mesh : camera_space_meshes) {
			               py::tuple one = face_vertices_of(mesh, K, image_size, near_clip, far_clip);
			               vertices.append(one[0]);
			               masks.append(one[1]);
			               counts.push_back(one[1].attr("shape").attr("__getitem__")(0).cast<int64_t>());
		               }
		               Torch& t = Torch::get();
		               return py::make_tuple(t.cat(vertices), t.cat(masks), t.as_tensor(py::cast(counts), py::arg("dtype") = t.dtype(Dtype::Int64)));
	               },
	               "camera_space_meshes"_a, "intrinsic_matrix"_a, "image_size"_a, "near_clipping_distance"_a = 0.f, "far_clipping_distance"_a = INFINITY);
	functional.def("get_mesh_ndc_face_vertices_and_clip_mask",
	               [](const py::object& camera_space_mesh, const py::object& intrinsic_matrix, const std::vector<int64_t>& image_size, float near_clip, float far_clip) {
		               if (image_size.size() != 2) log_error("image_size should be a SizeVector of size 2. Got size " + std::to_string(image_size.size()) + ".");
		               return face_vertices_of(camera_space_mesh, host_matrix(intrinsic_matrix, 3, 3, false), image_size, near_clip, far_clip);
	               },
	               "camera_space_mesh"_a, "intrinsic_matrix"_a, "image_size"_a, "near_clipping_distance"_a = 0.f, "far_clipping_distance"_a = INFINITY);
	// InterpolateVertexAttributes (cpp/rendering/functional/kernel/InterpolateFaceAttributesImpl.h:31-71)
	functional.def("interpolate_vertex_attributes",
	               [](const py::object& pixel_face_indices, const py::object& barycentric_coordinates, const py::object& face_vertex_attributes) {
		               Dev indices = as_cuda(pixel_face_indices, Dtype::Int64), barycentrics = as_cuda(barycentric_coordinates, Dtype::Float32),
		                   attributes = as_cuda(face_vertex_attributes, Dtype::Float32);
		               if (indices.dim() != 3 || barycentrics.dim() != 4 || attributes.dim() != 3 || attributes.size(1) != 3)
			               log_error("expected pixel_face_indices [H,W,K], barycentric_coordinates [H,W,K,3] and face_vertex_attributes [F,3,C]");
		               const int64_t h = indices.size(0), w = indices.size(1), k = indices.size(2), c = attributes.size(2);
		               Dev out = empty({h, w, k, c}, Dtype::Float32);
		               check(nnrt_b200_interpolate_vertex_attributes(indices.data<int64_t>(), barycentrics.data<float>(), h * w, (int32_t) k, attributes.data<float>(),
		                                                             (int32_t) c, out.data<float>(), current_stream()));
		               return out.object;
	               },
	               "pixel_face_indices"_a, "barycentric_coordinates"_a, "face_vertex_attributes"_a);
}

}  // namespace nnrt
