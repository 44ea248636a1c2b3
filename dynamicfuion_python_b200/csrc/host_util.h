// Host-side helpers of the C ABI layer: thread-local error message, CUDA error checks, launch sizing.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdarg>
#include <cstdint>
#include <atomic>
#include "common.cuh"

namespace nnrt_b200 {

void set_last_error(const char* fmt, ...);

#define NNRT_CHECK_ARG(cond, ...)                                 \
	do {                                                          \
		if (!(cond)) {                                            \
			::nnrt_b200::set_last_error(__VA_ARGS__);             \
			return ::nnrt_b200::kInvalidArgument;                 \
		}                                                         \
	} while (0)

#define NNRT_CUDA(call)                                                                                   \
	do {                                                                                                  \
		cudaError_t err__ = (call);                                                                       \
		if (err__ != cudaSuccess) {                                                                       \
			::nnrt_b200::set_last_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(err__), __FILE__, __LINE__); \
			return ::nnrt_b200::kCudaError;                                                               \
		}                                                                                                 \
	} while (0)

#define NNRT_LAUNCH_CHECK() NNRT_CUDA(cudaGetLastError())

// Opt-in to more than 48 KB of dynamic shared memory. The attribute belongs to the (kernel, device) pair, so one flag per call site is
// kept per device ordinal; fitters of different devices or threads may reach a launcher at the same time.
struct SharedMemoryOptIn {
	std::atomic<unsigned long long> done{0};
	template<typename Kernel>
	int ensure(Kernel kernel, size_t bytes) {
		int device = 0;
		if (cudaGetDevice(&device) != cudaSuccess) return kCudaError;
		const unsigned long long bit = 1ull << (device & 63);
		if (done.load(std::memory_order_acquire) & bit) return kOk;
		if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes) != cudaSuccess) return kCudaError;
		done.fetch_or(bit, std::memory_order_release);
		return kOk;
	}
};

// Stream-ordered scratch for the stage-level entry points. The device's default memory pool hands its memory back to the driver at
// every synchronisation unless a release threshold is set, which makes each call pay for mapping fresh pages (measured: ~0.3 ms per
// call of the warp-field construction); the threshold is raised once per device so freed scratch stays in the pool.
template<typename T>
inline cudaError_t scratch_malloc_async(T** pointer, size_t bytes, cudaStream_t stream) {
	static std::atomic<unsigned long long> retained{0};
	int device = 0;
	cudaError_t err = cudaGetDevice(&device);
	if (err != cudaSuccess) return err;
	const unsigned long long bit = 1ull << (device & 63);
	if (!(retained.load(std::memory_order_acquire) & bit)) {
		cudaMemPool_t pool;
		if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
			unsigned long long threshold = ~0ull;
			cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
		}
		retained.fetch_or(bit, std::memory_order_release);
	}
	return cudaMallocAsync((void**) pointer, bytes, stream);
}

inline unsigned blocks_for(int64_t n, int threads) { return (unsigned) ((n + threads - 1) / threads); }

inline Intrinsics intrinsics_from_host(const double* K) { return Intrinsics{(float) K[0], (float) K[4], (float) K[2], (float) K[5]}; }
inline Rigid rigid_from_host(const double* T44) {
	Rigid r;
	for (int i = 0; i < 3; i++)
		for (int j = 0; j < 4; j++) r.e[i * 4 + j] = T44 ? (float) T44[i * 4 + j] : (i == j ? 1.f : 0.f);
	return r;
}
// rendering::kernel::ImageSpaceIntrinsicsToNdc (cpp/rendering/kernel/CoordinateSystemConversions.h:108-148), double precision on the host
struct NdcCamera {
	Intrinsics k;
	float min_x, max_x, min_y, max_y;
};
inline NdcCamera ndc_camera_from_host(const double* K, int height, int width) {
	const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
	const double smaller = width < height ? (double) width : (double) height;
	float range_x = 2.0f, range_y = 2.0f;
	if (width > height) range_x = ((float) width * 2.0f) / (float) height;
	if (height > width) range_y = ((float) height * 2.0f) / (float) width;
	const double fx_ndc = 2.0 * fx / smaller, fy_ndc = -2.0 * fy / smaller;
	const double cx_ndc = -(2.0 * cx - (double) width) / smaller, cy_ndc = (2.0 * cy - (double) height) / smaller;
	NdcCamera c;
	c.k = Intrinsics{(float) fx_ndc, (float) fy_ndc, (float) cx_ndc, (float) cy_ndc};
	c.min_x = (float) (cx_ndc - range_x / 2.f); c.max_x = (float) (cx_ndc + range_x / 2.f);
	c.min_y = (float) (cy_ndc - range_y / 2.f); c.max_y = (float) (cy_ndc + range_y / 2.f);
	return c;
}

}  // namespace nnrt_b200
