// nnrt.core.linalg, second part: the generic block-sparse routines the arrowhead solve is built from (C++-only in the reference:
// cpp/core/linalg/MatmulBlockSparse.h, InvertBlocks.h, TransposeBlocks.h, DiagonalBlocks.h, SparseBlocks.h), bound under the reference's
// function names in snake case, over csrc/linalg_blocks.cu.
#include "bridge.h"

using namespace pybind11::literals;

namespace nnrt {
namespace {

enum class MatrixPreprocessingOperation { NONE = 0, TRANSPOSE = 1 };  // cpp/core/linalg/MatrixPreprocessingOperation.h
enum class UpLoTriangular { LOWER = 0, UPPER = 1 };                   // cpp/core/linalg/UpLoTriangular.h

Dev blocks_of(const py::handle& value) {
	Dev b = as_cuda(value, Dtype::Float32);
	if (b.dim() != 3 || b.size(1) != b.size(2)) log_error("blocks must have shape [count, size, size], got " + b.shape_string());
	return b;
}
// the caller's own tensor, modified in place
Dev in_place(const py::object& value, int dims, const char* what) {
	Dev d = as_cuda(value, Dtype::Float32);
	if (!d.object.is(value) || d.dim() != dims) log_error(std::string(what) + " must be a contiguous float32 CUDA tensor (modified in place)");
	return d;
}
void matrix_blocks(const Dev& matrix, const Dev& blocks, const Dev& coordinates, int row_offset, int column_offset, bool transpose, int gather) {
	check(nnrt_b200_dense_matrix_blocks(matrix.data<float>(), matrix.size(0), matrix.size(1), blocks.data<float>(),
	                                    coordinates.defined() ? coordinates.data<int32_t>() : nullptr, blocks.size(0), (int32_t) blocks.size(1), row_offset,
	                                    column_offset, transpose ? 1 : 0, gather, current_stream()));
}
py::tuple padded_product(const py::object& blocks_a, const py::object& blocks_b, const py::object& coordinates) {
	Dev a = blocks_of(blocks_a), b = blocks_of(blocks_b), bc = as_cuda(coordinates, Dtype::Int32);
	Dev c = empty(b.shape, Dtype::Float32);
	check(nnrt_b200_matmul_block_sparse_row_wise_padded(a.data<float>(), (int32_t) a.size(0), b.data<float>(), bc.data<int32_t>(), (int32_t) b.size(0),
	                                                    (int32_t) a.size(1), c.data<float>(), current_stream()));
	return py::make_tuple(c.object, bc.object, a.size(0));
}

}  // namespace

void bind_linalg_blocks(py::module_& m) {
	py::enum_<MatrixPreprocessingOperation>(m, "MatrixPreprocessingOperation")
			.value("NONE", MatrixPreprocessingOperation::NONE)
			.value("TRANSPOSE", MatrixPreprocessingOperation::TRANSPOSE);
	py::enum_<UpLoTriangular>(m, "UpLoTriangular").value("LOWER", UpLoTriangular::LOWER).value("UPPER", UpLoTriangular::UPPER);

	// MatmulBlockSparseRowWisePadded: one output block per block of B (zeros where A has no row)
	m.def("matmul_block_sparse_row_wise_padded", [](const py::object& a, const py::object& b, const py::object& coordinates) {
		return py::object(padded_product(a, b, coordinates)[0]);
	}, "blocks_a"_a, "blocks_b"_a, "blocks_b_coordinates"_a);
	// MatmulBlockSparseRowWise: (product blocks, their coordinates) for the rows of B that have an A block
	m.def("matmul_block_sparse_row_wise", [](const py::object& a, const py::object& b, const py::object& coordinates) {
		py::tuple r = padded_product(a, b, coordinates);
		py::object keep = r[1].attr("__getitem__")(py::make_tuple(py::slice(py::none(), py::none(), py::none()), 0)).attr("__lt__")(r[2]);
		return py::make_tuple(r[0].attr("__getitem__")(keep).attr("contiguous")(), r[1].attr("__getitem__")(keep).attr("contiguous")());
	}, "blocks_a"_a, "blocks_b"_a, "blocks_b_coordinates"_a);
	// MatmulBlockSparse: (product blocks, coordinates [K,2] int32) in row-major order of the output positions that received a product
	m.def("matmul_block_sparse", [](const py::object& blocks_a, const py::object& board_a, const py::object& op_a, const py::object& blocks_b,
	                                const py::object& board_b, const py::object& op_b) {
		Dev a = blocks_of(blocks_a), b = blocks_of(blocks_b), ab = as_cuda(board_a, Dtype::Int16), bb = as_cuda(board_b, Dtype::Int16);
		const int ta = py::int_(op_a).cast<int>(), tb = py::int_(op_b).cast<int>();
		const int64_t out_rows = ta ? ab.size(1) : ab.size(0), out_cols = tb ? bb.size(0) : bb.size(1), s = a.size(1);
		Dev blocks = empty({out_rows * out_cols, s, s}, Dtype::Float32), mask = empty({out_rows * out_cols}, Dtype::UInt8);
		check(nnrt_b200_matmul_block_sparse(a.data<float>(), ab.data<int16_t>(), (int32_t) ab.size(0), (int32_t) ab.size(1), ta, b.data<float>(), bb.data<int16_t>(),
		                                    (int32_t) bb.size(0), (int32_t) bb.size(1), tb, (int32_t) s, blocks.data<float>(), mask.data<uint8_t>(), current_stream()));
		Torch& t = Torch::get();
		py::object keep = to_bool(mask);
		py::object index = t.module.attr("arange")(out_rows * out_cols, py::arg("device") = t.cuda_device_kw);
		py::object coordinates = t.module.attr("stack")(py::make_tuple(index.attr("__floordiv__")(out_cols), index.attr("__mod__")(out_cols)), 1)
				.attr("to")(t.dtype(Dtype::Int32));
		return py::make_tuple(blocks.object.attr("__getitem__")(keep).attr("contiguous")(), coordinates.attr("__getitem__")(keep).attr("contiguous")());
	}, "blocks_a"_a, "blocks_a_breadboard"_a, "matrix_a_preprocessing"_a, "blocks_b"_a, "blocks_b_breadboard"_a, "matrix_b_preprocessing"_a);
	// BlockSparseAndVectorProduct: op(A) b with A given as blocks + block coordinates (+ offset)
	m.def("block_sparse_and_vector_product", [](const py::object& blocks_a, int64_t other_dimension, const py::object& coordinates, const py::object& offset,
	                                            const py::object& op_a, const py::object& vector_b) {
		Dev a = blocks_of(blocks_a), co = as_cuda(coordinates, Dtype::Int32), b = as_cuda(vector_b, Dtype::Float32);
		int row_offset = 0, column_offset = 0;
		if (!offset.is_none() && py::len(offset) == 2) {
			row_offset = offset[py::int_(0)].cast<int>();
			column_offset = offset[py::int_(1)].cast<int>();
		}
		std::vector<int64_t> shape = b.shape;
		shape[0] = other_dimension;
		Dev c = empty(shape, Dtype::Float32);
		check(nnrt_b200_block_sparse_and_vector_product(a.data<float>(), co.data<int32_t>(), (int32_t) a.size(0), (int32_t) a.size(1), (int32_t) other_dimension,
		                                                py::int_(op_a).cast<int>(), row_offset, column_offset, b.data<float>(), c.data<float>(), current_stream()));
		return c.object;
	}, "blocks_a"_a, "matrix_a_other_dimension"_a, "blocks_a_coordinates"_a, "block_coordinate_offset"_a, "matrix_a_preprocessing"_a, "vector_b"_a);
	m.def("diagonal_block_sparse_and_vector_product", [](const py::object& blocks_d, const py::object& vector_b) {
		Dev d = blocks_of(blocks_d), b = as_cuda(vector_b, Dtype::Float32);
		Dev c = empty(b.shape, Dtype::Float32);
		check(nnrt_b200_diagonal_block_sparse_and_vector_product(d.data<float>(), (int32_t) d.size(0), (int32_t) d.size(1), b.data<float>(), c.data<float>(),
		                                                         current_stream()));
		return c.object;
	}, "blocks_d"_a, "vector_b"_a);
	// InvertPositiveSemidefiniteBlocks (cpp/core/linalg/InvertBlocks.cpp:82-127); raises if a block is not positive definite
	m.def("invert_positive_semidefinite_blocks", [](const py::object& blocks) {
		Dev b = blocks_of(blocks);
		Dev out = empty(b.shape, Dtype::Float32), failures = zeros({1}, Dtype::Int32);
		check(nnrt_b200_invert_positive_semidefinite_blocks(b.data<float>(), (int32_t) b.size(0), (int32_t) b.size(1), out.data<float>(), failures.data<int32_t>(),
		                                                    current_stream()));
		const int failed = failures.object.attr("item")().cast<int>();
		if (failed != 0) log_error(std::to_string(failed) + " block(s) are not positive definite");
		return out.object;
	}, "blocks"_a);
	m.def("invert_triangular_blocks", [](const py::object& blocks, const py::object& uplo) {
		Dev b = blocks_of(blocks);
		Dev out = empty(b.shape, Dtype::Float32);
		check(nnrt_b200_invert_triangular_blocks(b.data<float>(), (int32_t) b.size(0), (int32_t) b.size(1), py::int_(uplo).cast<int>(), out.data<float>(),
		                                         current_stream()));
		return out.object;
	}, "blocks"_a, "uplo"_a);
	m.def("transpose_blocks_in_place", [](const py::object& blocks) {
		Dev b = in_place(blocks, 3, "blocks");
		check(nnrt_b200_transpose_blocks_in_place(b.data<float>(), b.size(0), (int32_t) b.size(1), current_stream()));
		return b.object;
	}, "blocks"_a);
	// FillInDiagonalBlocks / GetDiagonalBlocks / FillInSparseBlocks / GetSparseBlocks (DiagonalBlocksImpl.h, SparseBlocksImpl.h)
	m.def("fill_in_diagonal_blocks", [](const py::object& matrix, const py::object& blocks) {
		Dev m2 = in_place(matrix, 2, "matrix");
		matrix_blocks(m2, blocks_of(blocks), Dev{}, 0, 0, false, 0);
		return m2.object;
	}, "matrix"_a, "blocks"_a);
	m.def("get_diagonal_blocks", [](const py::object& matrix, int block_size) {
		Dev m2 = in_place(matrix, 2, "matrix");
		Dev out = empty({m2.size(0) / block_size, block_size, block_size}, Dtype::Float32);
		matrix_blocks(m2, out, Dev{}, 0, 0, false, 1);
		return out.object;
	}, "matrix"_a, "block_size"_a);
	m.def("fill_in_sparse_blocks", [](const py::object& matrix, const py::object& blocks, const py::object& coordinates, const py::object& offset, bool transpose) {
		Dev m2 = in_place(matrix, 2, "matrix");
		matrix_blocks(m2, blocks_of(blocks), as_cuda(coordinates, Dtype::Int32), offset[py::int_(0)].cast<int>(), offset[py::int_(1)].cast<int>(), transpose, 0);
		return m2.object;
	}, "matrix"_a, "blocks"_a, "coordinates"_a, "block_coordinate_offset"_a = py::make_tuple(0, 0), "transpose"_a = false);
	m.def("get_sparse_blocks", [](const py::object& matrix, int block_size, const py::object& coordinates) {
		Dev m2 = in_place(matrix, 2, "matrix"), co = as_cuda(coordinates, Dtype::Int32);
		Dev out = empty({co.size(0), block_size, block_size}, Dtype::Float32);
		matrix_blocks(m2, out, co, 0, 0, false, 1);
		return out.object;
	}, "matrix"_a, "block_size"_a, "coordinates"_a);
}

}  // namespace nnrt
