// Device-wide exclusive scan of int32 shared by the index-building entry points (warp-field construction, vertex-normal CSR).
#pragma once
#include "host_util.h"

namespace nnrt_b200 {

// ---------------------------------------------------------------------------------------------------------------------------
// Device-wide exclusive scan of int32 (n + 1 outputs: out[n] = total). Three launches: per-tile sums, one CTA over the tile
// sums, per-tile rescan. `out` may alias `in`.
constexpr int kScanThreads = 256, kScanItems = 8, kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int block_exclusive_scan(int value, int* shared_warp_totals, int* block_total) {
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
	int inclusive = value;
#pragma unroll
	for (int offset = 1; offset < 32; offset <<= 1) {
		const int other = __shfl_up_sync(0xffffffffu, inclusive, offset);
		if (lane >= offset) inclusive += other;
	}
	__syncthreads();  // protects shared_warp_totals across consecutive calls
	if (lane == 31) shared_warp_totals[warp] = inclusive;
	__syncthreads();
	int warp_base = 0, total = 0;
	for (int w = 0; w < warps; w++) {
		const int t = shared_warp_totals[w];
		if (w < warp) warp_base += t;
		total += t;
	}
	*block_total = total;
	return warp_base + inclusive - value;
}

static __global__ void __launch_bounds__(kScanThreads) scan_tile_sums_kernel(const int* __restrict__ in, int64_t n, int* __restrict__ tile_sums) {
	__shared__ int warp_totals[kScanThreads / 32];
	const int64_t base = (int64_t) blockIdx.x * kScanTile + (int64_t) threadIdx.x * kScanItems;
	int sum = 0;
#pragma unroll
	for (int k = 0; k < kScanItems; k++)
		if (base + k < n) sum += in[base + k];
	int total;
	block_exclusive_scan(sum, warp_totals, &total);
	if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

static __global__ void __launch_bounds__(1024) scan_tile_offsets_kernel(int* __restrict__ tile_sums, int tiles, int* __restrict__ total_out) {
	__shared__ int warp_totals[32];
	int carry = 0;
	for (int base = 0; base < tiles; base += 1024) {
		const int i = base + threadIdx.x;
		const int value = i < tiles ? tile_sums[i] : 0;
		int total;
		const int exclusive = block_exclusive_scan(value, warp_totals, &total);
		if (i < tiles) tile_sums[i] = carry + exclusive;
		carry += total;
	}
	if (threadIdx.x == 0) *total_out = carry;
}

static __global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const int* in, int* out, int64_t n, const int* __restrict__ tile_offsets) {
	__shared__ int warp_totals[kScanThreads / 32];
	const int64_t base = (int64_t) blockIdx.x * kScanTile + (int64_t) threadIdx.x * kScanItems;
	int items[kScanItems];
	int sum = 0;
#pragma unroll
	for (int k = 0; k < kScanItems; k++) {
		items[k] = base + k < n ? in[base + k] : 0;
		sum += items[k];
	}
	int total;
	int running = tile_offsets[blockIdx.x] + block_exclusive_scan(sum, warp_totals, &total);
#pragma unroll
	for (int k = 0; k < kScanItems; k++) {
		if (base + k < n) out[base + k] = running;
		running += items[k];
	}
}

// up to kSmallScan items (warp-field sizes): one CTA, one launch
constexpr int kSmallScan = 1024 * kScanItems;
static __global__ void __launch_bounds__(1024) scan_small_kernel(const int* in, int* out, int n) {
	__shared__ int warp_totals[32];
	const int base = threadIdx.x * kScanItems;
	int items[kScanItems];
	int sum = 0;
#pragma unroll
	for (int k = 0; k < kScanItems; k++) {
		items[k] = base + k < n ? in[base + k] : 0;
		sum += items[k];
	}
	int total;
	int running = block_exclusive_scan(sum, warp_totals, &total);
#pragma unroll
	for (int k = 0; k < kScanItems; k++) {
		if (base + k < n) out[base + k] = running;
		running += items[k];
	}
	if (threadIdx.x == 0) out[n] = total;
}

// scratch: ceil(n / kScanTile) ints; out needs n + 1 ints
inline int exclusive_scan(const int* in, int* out, int64_t n, int* tile_scratch, cudaStream_t s) {
	const int tiles = (int) ((n + kScanTile - 1) / kScanTile);
	if (tiles == 0) { NNRT_CUDA(cudaMemsetAsync(out, 0, sizeof(int), s)); return kOk; }
	if (n <= kSmallScan) {
		scan_small_kernel<<<1, 1024, 0, s>>>(in, out, (int) n);
		NNRT_LAUNCH_CHECK();
		return kOk;
	}
	scan_tile_sums_kernel<<<tiles, kScanThreads, 0, s>>>(in, n, tile_scratch);
	scan_tile_offsets_kernel<<<1, 1024, 0, s>>>(tile_scratch, tiles, out + n);
	scan_apply_kernel<<<tiles, kScanThreads, 0, s>>>(in, out, n, tile_scratch);
	NNRT_LAUNCH_CHECK();
	return kOk;
}
inline int64_t scan_scratch_ints(int64_t n) { return (n + kScanTile - 1) / kScanTile + 1; }


}  // namespace nnrt_b200
