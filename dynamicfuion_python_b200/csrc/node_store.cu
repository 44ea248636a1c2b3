// Small device helpers of the warp field's node store (WarpField / HierarchicalGraphWarpField accessors): everything the C++ host
// classes of the pybind11 module do to node tensors besides the warp itself. All are index / elementwise work on a few thousand
// elements: one launch each, no scratch.
#include "host_util.h"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {
namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) add_scaled_kernel(const float* __restrict__ a, const float* __restrict__ b, float alpha, int64_t count,
                                                              float* __restrict__ out) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < count) out[i] = a[i] + alpha * b[i];
}

__global__ void __launch_bounds__(kThreads) identity_rotations_kernel(float* __restrict__ rotations, int64_t count) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < count * 9) {
		const int k = (int) (i % 9);
		rotations[i] = (k == 0 || k == 4 || k == 8) ? 1.f : 0.f;
	}
}

// rows of 4-byte words: out[i, :] = source[indices[i], :]
__global__ void __launch_bounds__(kThreads) gather_rows_kernel(const uint32_t* __restrict__ source, int64_t source_rows, int32_t row_words,
                                                               const int64_t* __restrict__ indices, int64_t count, uint32_t* __restrict__ out,
                                                               int* __restrict__ out_of_range) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count * row_words) return;
	const int64_t row = i / row_words, word = i - row * row_words;
	const int64_t from = indices[row];
	if (from < 0 || from >= source_rows) {
		if (out_of_range) atomicExch(out_of_range, 1);
		out[i] = 0u;
		return;
	}
	out[i] = source[from * row_words + word];
}

__global__ void __launch_bounds__(kThreads) invert_permutation_kernel(const int64_t* __restrict__ permutation, int64_t count, int64_t* __restrict__ inverse) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	const int64_t p = permutation[i];
	if (p >= 0 && p < count) inverse[p] = i;
}

// anchors hold node indices or -1 (no anchor): out = map[anchor], -1 kept
__global__ void __launch_bounds__(kThreads) remap_indices_kernel(const int32_t* __restrict__ indices, int64_t count, const int64_t* __restrict__ map,
                                                                 int64_t map_size, int32_t* __restrict__ out) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	const int32_t v = indices[i];
	out[i] = (v >= 0 && v < map_size) ? (int32_t) map[v] : v;
}

}  // namespace
}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_add_scaled(const float* a, const float* b, float alpha, int64_t count, float* out, void* stream) {
	if (count == 0) return kOk;
	NNRT_CHECK_ARG(a && b && out && count > 0, "add_scaled: null pointer or negative count");
	add_scaled_kernel<<<blocks_for(count, kThreads), kThreads, 0, (cudaStream_t) stream>>>(a, b, alpha, count, out);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_fill_identity_rotations(float* rotations, int64_t count, void* stream) {
	if (count == 0) return kOk;
	NNRT_CHECK_ARG(rotations && count > 0, "fill_identity_rotations: null pointer or negative count");
	identity_rotations_kernel<<<blocks_for(count * 9, kThreads), kThreads, 0, (cudaStream_t) stream>>>(rotations, count);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_gather_rows(const void* source, int64_t source_rows, int32_t row_words, const int64_t* indices, int64_t count, void* out,
                                     void* stream) {
	if (count == 0 || row_words == 0) return kOk;
	NNRT_CHECK_ARG(source && indices && out, "gather_rows: null pointer argument");
	NNRT_CHECK_ARG(source_rows > 0 && row_words > 0 && count > 0, "gather_rows: sizes must be positive");
	gather_rows_kernel<<<blocks_for(count * row_words, kThreads), kThreads, 0, (cudaStream_t) stream>>>((const uint32_t*) source, source_rows, row_words, indices,
	                                                                                               count, (uint32_t*) out, nullptr);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_invert_permutation(const int64_t* permutation, int64_t count, int64_t* inverse, void* stream) {
	if (count == 0) return kOk;
	NNRT_CHECK_ARG(permutation && inverse && count > 0, "invert_permutation: null pointer or negative count");
	invert_permutation_kernel<<<blocks_for(count, kThreads), kThreads, 0, (cudaStream_t) stream>>>(permutation, count, inverse);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_remap_indices(const int32_t* indices, int64_t count, const int64_t* map, int64_t map_size, int32_t* out, void* stream) {
	if (count == 0) return kOk;
	NNRT_CHECK_ARG(indices && map && out && count > 0 && map_size > 0, "remap_indices: null pointer or non-positive size");
	remap_indices_kernel<<<blocks_for(count, kThreads), kThreads, 0, (cudaStream_t) stream>>>(indices, count, map, map_size, out);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
