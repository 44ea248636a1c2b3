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

// Stream-ordered scratch for the stage-level entry points, taken from a PRIVATE memory pool per device (the device's default pool is
// left untouched: a host application that shares it keeps its own release policy). The pool keeps freed memory (release threshold
// raised once) because handing it back to the driver at every synchronisation makes each call pay for mapping fresh pages (measured:
// ~0.3 ms per call of the warp-field construction).
inline cudaMemPool_t scratch_pool(int device) {
	static std::atomic<cudaMemPool_t> pools[64];
	cudaMemPool_t pool = pools[device & 63].load(std::memory_order_acquire);
	if (pool) return pool;
	cudaMemPoolProps props{};
	props.allocType = cudaMemAllocationTypePinned;
	props.handleTypes = cudaMemHandleTypeNone;
	props.location.type = cudaMemLocationTypeDevice;
	props.location.id = device;
	cudaMemPool_t created = nullptr;
	if (cudaMemPoolCreate(&created, &props) != cudaSuccess) return nullptr;
	unsigned long long threshold = ~0ull;
	cudaMemPoolSetAttribute(created, cudaMemPoolAttrReleaseThreshold, &threshold);
	cudaMemPool_t expected = nullptr;
	if (!pools[device & 63].compare_exchange_strong(expected, created, std::memory_order_acq_rel)) {
		cudaMemPoolDestroy(created);  // another thread won the race
		return expected;
	}
	return created;
}

// Scratch of one entry-point call: every allocation is returned to the pool (stream-ordered) when the object goes out of scope, on
// the error paths as well as on success.
class Scratch {
public:
	explicit Scratch(cudaStream_t stream) : stream_(stream) {}
	Scratch(const Scratch&) = delete;
	Scratch& operator=(const Scratch&) = delete;
	~Scratch() {
		for (int i = 0; i < count_; i++) cudaFreeAsync(pointers_[i], stream_);
	}
	template<typename T>
	cudaError_t alloc(T** pointer, size_t bytes) {
		*pointer = nullptr;
		if (count_ >= kMax) return cudaErrorMemoryAllocation;
		int device = 0;
		cudaError_t err = cudaGetDevice(&device);
		if (err != cudaSuccess) return err;
		cudaMemPool_t pool = scratch_pool(device);
		if (pool == nullptr) return cudaErrorMemoryAllocation;
		err = cudaMallocFromPoolAsync((void**) pointer, bytes > 0 ? bytes : 1, pool, stream_);
		if (err == cudaSuccess) pointers_[count_++] = *pointer;
		return err;
	}
private:
	static constexpr int kMax = 8;
	cudaStream_t stream_;
	void* pointers_[kMax];
	int count_ = 0;
};

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
