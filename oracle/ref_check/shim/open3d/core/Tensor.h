// Stub of Open3D core/Tensor.h — ORACLE VALIDATION SCAFFOLDING ONLY (types named, never used).
#pragma once
#include <cstdint>
#include <vector>
#include <string>
#include <tuple>
namespace open3d {
namespace utility { template<typename... Args> inline void LogError(const char*, Args&&...) { throw 1; } }
namespace core {
struct Device { enum class DeviceType { CPU = 0, CUDA = 1 }; Device() {} Device(const std::string&) {} };
struct SizeVector : public std::vector<int64_t> { using std::vector<int64_t>::vector; };
struct Dtype {};
static const Dtype Float32{}, Float64{}, Int32{}, Int64{};
class Tensor {
public:
	Tensor() {}
	template<typename T> Tensor(const std::vector<T>&, const SizeVector&, Dtype, const Device& = Device()) {}
	template<typename T> std::vector<T> ToFlatVector() const { return {}; }
};
}}
