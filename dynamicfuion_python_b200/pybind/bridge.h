// nnrt pybind11 module — tensor interchange between Python callers and the C ABI (include/nnrt_b200.h).
//
// The reference's bindings take open3d.core.Tensor (cpp/pybind/nnrt_pybind.cpp:23-40); Open3D is absent here, so every tensor
// argument may be a torch.Tensor, a numpy array or anything with __dlpack__ (SURVEY.md §8(b) "tensor interchange"), and results are
// torch CUDA tensors. torch is used for device memory, host<->device copies and the current stream only: it is reached through the
// Python C API (no libtorch link), every kernel goes through libnnrt_b200.so. There is no CPU path: without CUDA the first call throws.
#pragma once
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/nnrt_b200.h"

namespace py = pybind11;

namespace nnrt {

enum class Dtype { Float32, Float64, Int32, Int64, Int16, Int8, UInt8, Bool };

// The reference reports errors with open3d::utility::LogError, which throws std::runtime_error -> Python RuntimeError.
[[noreturn]] inline void log_error(const std::string& message) { throw std::runtime_error(message); }
inline void check(int status) {
	if (status != NNRT_B200_OK) log_error(nnrt_b200_last_error());
}

// torch objects looked up once per process
struct Torch {
	py::object module, empty, zeros, as_tensor, from_numpy, from_dlpack, tensor_type, ndarray_type, cat, current_raw_stream, current_device, is_available;
	py::object dtypes[8];
	py::object cuda_device_kw;  // "cuda"
	static Torch& get();
	py::object dtype(Dtype d) const { return dtypes[(int) d]; }
};

void* current_stream();
inline bool has_dtype(const py::handle& tensor, const py::handle& dtype) { return py::object(tensor.attr("dtype")).ptr() == dtype.ptr(); }

// A contiguous CUDA tensor of a known dtype: `object` keeps the torch tensor (and the memory) alive, `pointer` is its data_ptr.
struct Dev {
	py::object object;
	void* pointer = nullptr;
	std::vector<int64_t> shape;
	bool defined() const { return !object.is_none() && object.ptr() != nullptr; }
	int64_t dim() const { return (int64_t) shape.size(); }
	int64_t size(int i) const { return shape[i < 0 ? shape.size() + i : i]; }
	int64_t numel() const {
		int64_t n = 1;
		for (auto s : shape) n *= s;
		return n;
	}
	template<typename T>
	T* data() const { return static_cast<T*>(pointer); }
	std::string shape_string() const;
};

py::object to_torch(const py::handle& value);                      // torch.Tensor from tensor / ndarray / __dlpack__ / sequence
Dev as_cuda(const py::handle& value, Dtype dtype);                 // converts / uploads when needed; zero-copy when already right
Dev as_cuda_or_none(const py::handle& value, Dtype dtype);         // None -> undefined Dev
Dev wrap(py::object tensor);                                       // a tensor already known to be contiguous CUDA
Dev empty(const std::vector<int64_t>& shape, Dtype dtype);
Dev zeros(const std::vector<int64_t>& shape, Dtype dtype);
Dev clone(const Dev& source);
Dev reshaped(const Dev& source, const std::vector<int64_t>& shape);
Dev narrow_rows(const Dev& source, int64_t first, int64_t count);  // view of rows [first, first + count)
py::object to_bool(const Dev& mask);                               // uint8 mask -> torch.bool tensor

// Float64 matrix on the host (intrinsics 3x3 / extrinsics 4x4 live on CPU:0 in the reference API); None -> identity
std::vector<double> host_matrix(const py::handle& value, int rows, int cols, bool identity_if_none);

// Mesh arguments: an Open3D-like object (.vertex["positions"|"normals"], .triangle["indices"]), a dict, or a (positions, normals,
// triangles) tuple; results are returned as nnrt.geometry.TriangleMesh-like dict holders (see geometry.cpp).
struct Mesh {
	Dev positions, normals, triangles, triangle_normals;
};
Mesh as_mesh(const py::handle& value);
py::object mesh_result(const Mesh& mesh);

void bind_core(py::module_& m);
void bind_geometry(py::module_& m);
void bind_rendering(py::module_& m);
void bind_alignment(py::module_& m);
void bind_io(py::module_& m);
void bind_alignment_functional(py::module_& m);      // alignment_functional_bind.cpp
void bind_shaders(py::module_& m);                   // FlatEdgeShader / VertexColorShader into nnrt.rendering
void bind_linalg_blocks(py::module_& m);             // linalg_blocks_bind.cpp, into nnrt.core.linalg
void bind_geometry_functional_more(py::module_& m);  // geometry_functional_bind.cpp, into nnrt.geometry.functional
void bind_tsdf(py::module_& m);                      // tsdf_bind.cpp, into nnrt.geometry.functional

}  // namespace nnrt
