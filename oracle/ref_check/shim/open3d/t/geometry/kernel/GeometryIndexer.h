// Stub of Open3D 0.17 t/geometry/kernel/GeometryIndexer.h — ORACLE VALIDATION SCAFFOLDING ONLY.
// Provides just enough of NDArrayIndexer / TArrayIndexer / TransformIndexer for the reference's
// Eigen-only device-function headers to compile verbatim. Semantics restated
// from the reference's call sites and Open3D's public API.
#pragma once
#include <cstdint>
#include <open3d/core/Tensor.h>
#define OPEN3D_HOST_DEVICE
#define OPEN3D_DEVICE
namespace open3d { namespace t { namespace geometry { namespace kernel {
class NDArrayIndexer {
public:
	NDArrayIndexer() : ptr_(nullptr), row_bytes_(0) {}
	NDArrayIndexer(void* ptr, int64_t row_bytes) : ptr_(ptr), row_bytes_(row_bytes) {}
	template<typename T> T* GetDataPtr(int64_t row) const {
		return reinterpret_cast<T*>(static_cast<char*>(ptr_) + row * row_bytes_);
	}
	void* ptr_; int64_t row_bytes_;
};
template<typename TIndex>
class TArrayIndexer : public NDArrayIndexer {
public:
	using NDArrayIndexer::NDArrayIndexer;
};
class TransformIndexer {
public:
	TransformIndexer(const double* K33, const double* T44, float scale = 1.0f) : scale_(scale) {
		fx_ = (float) K33[0]; fy_ = (float) K33[4]; cx_ = (float) K33[2]; cy_ = (float) K33[5];
		for (int i = 0; i < 3; i++) for (int j = 0; j < 4; j++) extrinsic_[i][j] = (float) T44[i * 4 + j];
	}
	void RigidTransform(float x_in, float y_in, float z_in, float* x_out, float* y_out, float* z_out) const {
		x_in *= scale_; y_in *= scale_; z_in *= scale_;
		*x_out = x_in * extrinsic_[0][0] + y_in * extrinsic_[0][1] + z_in * extrinsic_[0][2] + extrinsic_[0][3];
		*y_out = x_in * extrinsic_[1][0] + y_in * extrinsic_[1][1] + z_in * extrinsic_[1][2] + extrinsic_[1][3];
		*z_out = x_in * extrinsic_[2][0] + y_in * extrinsic_[2][1] + z_in * extrinsic_[2][2] + extrinsic_[2][3];
	}
	void Rotate(float x_in, float y_in, float z_in, float* x_out, float* y_out, float* z_out) const {
		x_in *= scale_; y_in *= scale_; z_in *= scale_;
		*x_out = x_in * extrinsic_[0][0] + y_in * extrinsic_[0][1] + z_in * extrinsic_[0][2];
		*y_out = x_in * extrinsic_[1][0] + y_in * extrinsic_[1][1] + z_in * extrinsic_[1][2];
		*z_out = x_in * extrinsic_[2][0] + y_in * extrinsic_[2][1] + z_in * extrinsic_[2][2];
	}
	void Project(float x_in, float y_in, float z_in, float* u_out, float* v_out) const {
		float inv_z = 1.0f / z_in;
		*u_out = fx_ * x_in * inv_z + cx_;
		*v_out = fy_ * y_in * inv_z + cy_;
	}
	void Unproject(float u_in, float v_in, float d_in, float* x_out, float* y_out, float* z_out) const {
		*x_out = (u_in - cx_) * d_in / fx_;
		*y_out = (v_in - cy_) * d_in / fy_;
		*z_out = d_in;
	}
	void GetFocalLength(float* fx, float* fy) const { *fx = fx_; *fy = fy_; }
private:
	float extrinsic_[3][4]; float fx_, fy_, cx_, cy_, scale_;
};
}}}}
