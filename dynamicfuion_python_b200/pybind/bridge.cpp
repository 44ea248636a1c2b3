#include "bridge.h"

#include <sstream>

namespace nnrt {

Torch& Torch::get() {
	static Torch* instance = nullptr;  // leaked on purpose: the interpreter may be gone when static destructors run
	if (instance) return *instance;
	auto* t = new Torch();
	t->module = py::module_::import("torch");
	t->empty = t->module.attr("empty");
	t->zeros = t->module.attr("zeros");
	t->as_tensor = t->module.attr("as_tensor");
	t->from_numpy = t->module.attr("from_numpy");
	t->from_dlpack = t->module.attr("from_dlpack");
	t->tensor_type = t->module.attr("Tensor");
	t->cat = t->module.attr("cat");
	t->ndarray_type = py::module_::import("numpy").attr("ndarray");
	py::object cuda = t->module.attr("cuda");
	t->is_available = cuda.attr("is_available");
	t->current_device = cuda.attr("current_device");
	t->current_raw_stream = t->module.attr("_C").attr("_cuda_getCurrentRawStream");
	const char* names[8] = {"float32", "float64", "int32", "int64", "int16", "int8", "uint8", "bool"};
	for (int i = 0; i < 8; i++) t->dtypes[i] = t->module.attr(names[i]);
	t->cuda_device_kw = py::str("cuda");
	if (!t->is_available().cast<bool>()) {
		delete t;
		log_error("nnrt (B200 build) needs a CUDA device; there is no CPU path");
	}
	instance = t;
	return *instance;
}

void* current_stream() {
	Torch& t = Torch::get();
	return reinterpret_cast<void*>(t.current_raw_stream(t.current_device()).cast<uintptr_t>());
}

std::string Dev::shape_string() const {
	std::ostringstream out;
	out << "(";
	for (size_t i = 0; i < shape.size(); i++) out << (i ? ", " : "") << shape[i];
	out << ")";
	return out.str();
}

py::object to_torch(const py::handle& value) {
	Torch& t = Torch::get();
	if (py::isinstance(value, t.tensor_type)) return py::reinterpret_borrow<py::object>(value);
	if (py::isinstance(value, t.ndarray_type)) return t.from_numpy(py::module_::import("numpy").attr("ascontiguousarray")(value));
	if (py::hasattr(value, "__dlpack__")) return t.from_dlpack(value);
	return t.as_tensor(value);
}

static std::vector<int64_t> shape_of(const py::object& tensor) {
	py::tuple s = py::tuple(tensor.attr("shape"));
	std::vector<int64_t> shape(s.size());
	for (size_t i = 0; i < s.size(); i++) shape[i] = s[i].cast<int64_t>();
	return shape;
}

// An empty tensor has no storage (data_ptr() == 0), which the library would read as a missing argument: it gets the address of a small
// placeholder instead (the entry points return early on a zero count). One placeholder per process is enough: it is never dereferenced.
static void* empty_placeholder() {
	static py::object* keep = nullptr;
	static void* address = nullptr;
	if (!address) {
		Torch& t = Torch::get();
		keep = new py::object(t.zeros(py::make_tuple(64), py::arg("dtype") = t.dtype(Dtype::Float32), py::arg("device") = t.cuda_device_kw));
		address = reinterpret_cast<void*>(keep->attr("data_ptr")().cast<uintptr_t>());
	}
	return address;
}

Dev wrap(py::object tensor) {
	Dev d;
	d.pointer = reinterpret_cast<void*>(tensor.attr("data_ptr")().cast<uintptr_t>());
	d.shape = shape_of(tensor);
	if (d.pointer == nullptr && d.numel() == 0) d.pointer = empty_placeholder();
	d.object = std::move(tensor);
	return d;
}

Dev as_cuda(const py::handle& value, Dtype dtype) {
	Torch& t = Torch::get();
	py::object tensor = to_torch(value);
	const py::object want = t.dtype(dtype);
	const bool right = tensor.attr("is_cuda").cast<bool>() && has_dtype(tensor, want) && tensor.attr("is_contiguous")().cast<bool>();
	if (!right) {
		// a pinned host tensor is uploaded asynchronously on the current stream (stream order protects the consumer kernels)
		const bool pinned = !tensor.attr("is_cuda").cast<bool>() && tensor.attr("is_pinned")().cast<bool>();
		tensor = tensor.attr("to")(py::arg("device") = t.cuda_device_kw, py::arg("dtype") = want, py::arg("non_blocking") = pinned).attr("contiguous")();
	}
	return wrap(std::move(tensor));
}

Dev as_cuda_or_none(const py::handle& value, Dtype dtype) {
	if (value.is_none()) return Dev{py::none(), nullptr, {}};
	return as_cuda(value, dtype);
}

Dev empty(const std::vector<int64_t>& shape, Dtype dtype) {
	Torch& t = Torch::get();
	return wrap(t.empty(py::cast(shape), py::arg("dtype") = t.dtype(dtype), py::arg("device") = t.cuda_device_kw));
}

Dev zeros(const std::vector<int64_t>& shape, Dtype dtype) {
	Torch& t = Torch::get();
	return wrap(t.zeros(py::cast(shape), py::arg("dtype") = t.dtype(dtype), py::arg("device") = t.cuda_device_kw));
}

Dev clone(const Dev& source) { return wrap(source.object.attr("clone")()); }

Dev reshaped(const Dev& source, const std::vector<int64_t>& shape) { return wrap(source.object.attr("reshape")(py::cast(shape))); }

Dev narrow_rows(const Dev& source, int64_t first, int64_t count) { return wrap(source.object.attr("narrow")(0, first, count)); }

py::object to_bool(const Dev& mask) { return mask.object.attr("to")(Torch::get().dtype(Dtype::Bool)); }

std::vector<double> host_matrix(const py::handle& value, int rows, int cols, bool identity_if_none) {
	std::vector<double> m((size_t) rows * cols, 0.0);
	if (value.is_none()) {
		if (!identity_if_none) log_error("a " + std::to_string(rows) + " x " + std::to_string(cols) + " matrix is required, got None");
		for (int i = 0; i < rows && i < cols; i++) m[(size_t) i * cols + i] = 1.0;
		return m;
	}
	Torch& t = Torch::get();
	py::object host = to_torch(value);
	// the common case on the hot path (a float64 CPU tensor, as the reference API prescribes) is read in place
	if (host.attr("is_cuda").cast<bool>() || !has_dtype(host, t.dtype(Dtype::Float64)) || !host.attr("is_contiguous")().cast<bool>() ||
	    host.attr("requires_grad").cast<bool>())
		host = host.attr("detach")().attr("to")(py::arg("device") = "cpu", py::arg("dtype") = t.dtype(Dtype::Float64)).attr("contiguous")();
	const std::vector<int64_t> shape = shape_of(host);
	if (shape.size() != 2 || shape[0] != rows || shape[1] != cols) {
		Dev d;
		d.shape = shape;
		log_error("expected a tensor of shape (" + std::to_string(rows) + ", " + std::to_string(cols) + "), got " + d.shape_string());
	}
	const double* data = reinterpret_cast<const double*>(host.attr("data_ptr")().cast<uintptr_t>());
	for (size_t i = 0; i < m.size(); i++) m[i] = data[i];
	return m;
}

static py::object item_or_none(const py::handle& mapping, const char* key) {
	if (mapping.attr("__contains__")(key).cast<bool>()) return mapping[py::str(key)];
	return py::none();
}

Mesh as_mesh(const py::handle& value) {
	py::object positions = py::none(), normals = py::none(), triangles = py::none(), triangle_normals = py::none();
	if (py::isinstance<py::tuple>(value) || py::isinstance<py::list>(value)) {
		py::sequence s = py::reinterpret_borrow<py::sequence>(value);
		if (s.size() > 0) positions = s[0];
		if (s.size() > 1) normals = s[1];
		if (s.size() > 2) triangles = s[2];
	} else if (py::isinstance<py::dict>(value)) {
		positions = item_or_none(value, "positions");
		normals = item_or_none(value, "normals");
		triangles = item_or_none(value, "triangles");
		if (triangles.is_none()) triangles = item_or_none(value, "indices");
		triangle_normals = item_or_none(value, "triangle_normals");
	} else if (py::hasattr(value, "vertex") && py::hasattr(value, "triangle")) {
		py::object vertex = value.attr("vertex"), triangle = value.attr("triangle");
		positions = item_or_none(vertex, "positions");
		normals = item_or_none(vertex, "normals");
		triangles = item_or_none(triangle, "indices");
		triangle_normals = item_or_none(triangle, "normals");
	} else {
		log_error("mesh must be an Open3D t.geometry.TriangleMesh-like object, a dict or a (positions, normals, triangles) tuple");
	}
	if (positions.is_none()) log_error("Argument mesh needs to have vertex positions defined.");
	Mesh mesh;
	mesh.positions = as_cuda(positions, Dtype::Float32);
	mesh.normals = as_cuda_or_none(normals, Dtype::Float32);
	mesh.triangles = as_cuda_or_none(triangles, Dtype::Int64);
	mesh.triangle_normals = as_cuda_or_none(triangle_normals, Dtype::Float32);
	if (mesh.positions.dim() != 2 || mesh.positions.size(1) != 3) log_error("mesh vertex positions must have shape [V, 3], got " + mesh.positions.shape_string());
	return mesh;
}

}  // namespace nnrt
