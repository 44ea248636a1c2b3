// nnrt.io: read_tensor / write_tensor in the reference's on-disk format (cpp/pybind/io.cpp:36-52, cpp/io/TensorIO.cpp:32-60).
//
// Byte layout (little endian; cpp/io/SizeVectorIO.cpp, DtypeIO.cpp, DeviceIO.cpp, BlobIO.cpp):
//   shape    u32 count, i64 x count
//   strides  u32 count, i64 x count            (in ELEMENTS, of the tensor as it was handed to write_tensor)
//   dtype    i32 DtypeCode, i64 byte size, u8 name length, name          (Open3D v0.17: Undefined 0, Bool 1, Int 2, UInt 3, Float 4, Object 5;
//                                                                          names "Float32", "Int64", "UInt8", "Bool", ...)
//   blob     i32 DeviceType (CPU 0), i32 device id, i64 byte count, bytes (always the CONTIGUOUS copy of the data on CPU:0)
// `compressed` wraps the stream in zstr (3rd-party/zstr: zlib deflate with a gzip header, level Z_DEFAULT_COMPRESSION); reading with
// compressed = true auto-detects and also accepts a plain file, as zstr::ifstream does. The reader rebuilds the tensor from the stored
// shape AND strides over the contiguous blob, exactly like the reference (TensorIO.cpp:55: a tensor that was non-contiguous when written
// comes back scrambled there too). Host-side only: tensors are returned on the CPU like the reference's ReadTensor.
#include "bridge.h"

#include <zlib.h>

#include <cstdio>
#include <cstring>

using namespace pybind11::literals;

namespace nnrt {
namespace {

struct DtypeRecord {
	const char* torch_name;
	int32_t code;
	int64_t byte_size;
	const char* open3d_name;
};
const DtypeRecord kDtypes[] = {{"float32", 4, 4, "Float32"}, {"float64", 4, 8, "Float64"}, {"int8", 2, 1, "Int8"},     {"int16", 2, 2, "Int16"},
                               {"int32", 2, 4, "Int32"},     {"int64", 2, 8, "Int64"},     {"uint8", 3, 1, "UInt8"},   {"uint16", 3, 2, "UInt16"},
                               {"uint32", 3, 4, "UInt32"},   {"uint64", 3, 8, "UInt64"},   {"bool", 1, 1, "Bool"}};

// plain or gzip byte sink / source
class Stream {
public:
	Stream(const std::string& path, bool write, bool compressed) : compressed_(compressed) {
		if (compressed) gz_ = gzopen(path.c_str(), write ? "wb" : "rb");
		else file_ = std::fopen(path.c_str(), write ? "wb" : "rb");
		if (!gz_ && !file_) log_error("Failed to open file at \"" + path + "\" for " + (write ? "writing." : "reading."));
	}
	~Stream() {
		if (gz_) gzclose(gz_);
		if (file_) std::fclose(file_);
	}
	void write(const void* data, size_t bytes) {
		const char* p = static_cast<const char*>(data);
		while (bytes > 0) {
			const unsigned chunk = (unsigned) std::min<size_t>(bytes, 1u << 30);
			const size_t done = compressed_ ? (size_t) std::max(0, gzwrite(gz_, p, chunk)) : std::fwrite(p, 1, chunk, file_);
			if (done != chunk) log_error("Failure writing to ostream.");
			p += chunk; bytes -= chunk;
		}
	}
	void read(void* data, size_t bytes) {
		char* p = static_cast<char*>(data);
		while (bytes > 0) {
			const unsigned chunk = (unsigned) std::min<size_t>(bytes, 1u << 30);
			const size_t done = compressed_ ? (size_t) std::max(0, gzread(gz_, p, chunk)) : std::fread(p, 1, chunk, file_);
			if (done != chunk) log_error("Failure reading from istream.");
			p += chunk; bytes -= chunk;
		}
	}
	template<typename T>
	void put(T value) { write(&value, sizeof(T)); }
	template<typename T>
	T get() {
		T value;
		read(&value, sizeof(T));
		return value;
	}
private:
	bool compressed_;
	gzFile gz_ = nullptr;
	std::FILE* file_ = nullptr;
};

void put_size_vector(Stream& out, const std::vector<int64_t>& v) {
	out.put<uint32_t>((uint32_t) v.size());
	for (int64_t e : v) out.put<int64_t>(e);
}
std::vector<int64_t> get_size_vector(Stream& in) {
	const uint32_t count = in.get<uint32_t>();
	if (count > 64) log_error("Failure reading from istream.");
	std::vector<int64_t> v(count);
	for (auto& e : v) e = in.get<int64_t>();
	return v;
}

// torch.Tensor from a tensor / ndarray / __dlpack__ object on ANY device (no CUDA needed for IO)
py::object to_torch_any_device(const py::object& value) {
	py::module_ torch = py::module_::import("torch");
	if (py::isinstance(value, torch.attr("Tensor"))) return value;
	if (py::isinstance(value, py::module_::import("numpy").attr("ndarray"))) return torch.attr("from_numpy")(value);
	if (py::hasattr(value, "__dlpack__")) return torch.attr("from_dlpack")(value);
	return torch.attr("as_tensor")(value);
}

void write_tensor(const std::string& path, const py::object& tensor_in, bool compressed) {
	py::object tensor = to_torch_any_device(tensor_in);
	const std::string dtype_name = py::str(tensor.attr("dtype")).cast<std::string>();  // "torch.float32"
	const DtypeRecord* record = nullptr;
	for (const DtypeRecord& r : kDtypes)
		if (dtype_name == std::string("torch.") + r.torch_name) record = &r;
	if (!record) log_error("write_tensor: unsupported dtype " + dtype_name);
	const std::vector<int64_t> shape = py::tuple(tensor.attr("shape")).cast<std::vector<int64_t>>();
	const std::vector<int64_t> strides = py::tuple(tensor.attr("stride")()).cast<std::vector<int64_t>>();
	py::object host = tensor.attr("detach")().attr("contiguous")().attr("cpu")();
	const int64_t byte_size = host.attr("numel")().cast<int64_t>() * record->byte_size;
	const void* data = reinterpret_cast<const void*>(host.attr("data_ptr")().cast<uintptr_t>());
	Stream out(path, true, compressed);
	put_size_vector(out, shape);
	put_size_vector(out, strides);
	out.put<int32_t>(record->code);
	out.put<int64_t>(record->byte_size);
	const std::string name = record->open3d_name;
	out.put<uint8_t>((uint8_t) name.size());
	out.write(name.data(), name.size());
	out.put<int32_t>(0);  // DeviceType::CPU
	out.put<int32_t>(0);  // device id
	out.put<int64_t>(byte_size);
	out.write(data, (size_t) byte_size);
}

py::object read_tensor(const std::string& path, bool compressed) {
	Stream in(path, false, compressed);
	const std::vector<int64_t> shape = get_size_vector(in), strides = get_size_vector(in);
	const int32_t code = in.get<int32_t>();
	const int64_t element_size = in.get<int64_t>();
	const uint8_t name_size = in.get<uint8_t>();
	std::string name(name_size, '\0');
	in.read(name.data(), name_size);
	const DtypeRecord* record = nullptr;
	for (const DtypeRecord& r : kDtypes)
		if (name == r.open3d_name && code == r.code && element_size == r.byte_size) record = &r;
	if (!record) log_error("read_tensor: unsupported dtype record (code " + std::to_string(code) + ", " + std::to_string(element_size) + " bytes, \"" + name + "\")");
	const int32_t device_type = in.get<int32_t>(), device_id = in.get<int32_t>();
	if (device_type != 0 || device_id != 0)
		log_error("Blob istream input for blobs with device type " + std::to_string(device_type) + ":" + std::to_string(device_id) +
		          " not supported (only CPU:0 device is supported).");
	const int64_t byte_size = in.get<int64_t>();
	if (byte_size < 0 || byte_size % record->byte_size != 0 || shape.size() != strides.size()) log_error("Failure reading from istream.");
	py::module_ torch = py::module_::import("torch");
	py::object flat = torch.attr("empty")(py::make_tuple(byte_size / record->byte_size), py::arg("dtype") = torch.attr(record->torch_name));
	in.read(reinterpret_cast<void*>(flat.attr("data_ptr")().cast<uintptr_t>()), (size_t) byte_size);
	// shape + stored strides over the contiguous blob (what Tensor(shape, strides, blob pointer, dtype, blob) does in the reference)
	int64_t last = 0, count = 1;
	for (size_t i = 0; i < shape.size(); i++) {
		if (shape[i] < 0 || strides[i] < 0) log_error("Failure reading from istream.");
		count *= shape[i];
		if (shape[i] > 0) last += (shape[i] - 1) * strides[i];
	}
	if (count > 0 && last >= byte_size / record->byte_size) log_error("read_tensor: the stored strides address beyond the stored data");
	return torch.attr("as_strided")(flat, py::cast(shape), py::cast(strides));
}

}  // namespace

void bind_io(py::module_& m) {
	m.def("write_tensor", &write_tensor, "Write tensor in binary form.", "path"_a, "tensor"_a, "compressed"_a = true);
	m.def("read_tensor", &read_tensor, "Read tensor from binary form.", "path"_a, "compressed"_a = true);
}

}  // namespace nnrt
