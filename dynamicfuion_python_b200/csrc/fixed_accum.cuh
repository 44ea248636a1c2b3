// Order-independent accumulation of fp32 values: a 103-bit fixed-point number in two 64-bit words,
//   value = (int64) word[0] * 2^-20 + (uint64) word[1] * 2^-60,
// added to with integer atomics. Integer addition is associative, so the per-node sums of the data term (J^T J, J^T r) do not
// depend on the order in which warps, CTAs or GPUs contribute: a fit is bit-identical run to run, and a pixel-sharded fit
// (integer all-reduce of the same words) is bit-identical to the single-GPU fit. Every fp32 addend of magnitude in [2^-36, 2^42]
// is represented exactly (smaller ones are truncated at 2^-60 absolute), so the sum is also more accurate than any fp32 order.
#pragma once
#include <cuda_runtime.h>

namespace nnrt_b200 {

constexpr float kFixedAccumLimit = 4.0e12f;  // |addend| must stay below 2^42; NaN / infinity / larger set the range flag instead

__device__ __forceinline__ void fixed_accum_add(unsigned long long* __restrict__ slot, float x, int* __restrict__ range_flag) {
	if (x == 0.f) return;
	if (!(fabsf(x) < kFixedAccumLimit)) { atomicExch(range_flag, 1); return; }
	const double t = (double) x * 1048576.0;  // exact: a power-of-two scaling of a 24-bit mantissa
	const double whole = floor(t);
	const long long hi = (long long) whole;
	const unsigned long long lo = (unsigned long long) ((t - whole) * 1099511627776.0);  // [0, 2^40), bits under 2^-60 dropped
	if (hi != 0ll) atomicAdd(slot, (unsigned long long) hi);
	if (lo != 0ull) atomicAdd(slot + 1, lo);  // zero for every |x| >= 16: one atomic per addend there
}
// (An integer-only conversion -- exponent extraction, 64-bit shifts, two's-complement negation -- measured 3 us slower per
// assembly launch than these five FP64 instructions.)

__device__ __forceinline__ float fixed_accum_value(unsigned long long hi, unsigned long long lo) {
	return (float) ((double) (long long) hi * (1.0 / 1048576.0) + (double) lo * (1.0 / 1152921504606846976.0));
}

}  // namespace nnrt_b200
