// Developer microbenchmark: single-warp / multi-warp FFMA vs FFMA2 issue rate on 64 register accumulators.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
	unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b), rc = *reinterpret_cast<unsigned long long*>(&c), rd;
	asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
	return *reinterpret_cast<float2*>(&rd);
}
template<int MODE>
__global__ void k(float* out, long long* cycles, int iters) {
	float a[64];
	for (int i = 0; i < 64; i++) a[i] = threadIdx.x * 0.001f + i;
	float pr[8], pk[8];
	for (int i = 0; i < 8; i++) { pr[i] = 1.0f + i * 1e-3f + threadIdx.x * 1e-6f; pk[i] = 0.5f - i * 1e-3f; }
	__syncthreads();
	long long t0 = clock64();
	for (int it = 0; it < iters; it++) {
		if (MODE == 0) {
#pragma unroll
			for (int r = 0; r < 8; r++)
#pragma unroll
				for (int c = 0; c < 8; c++) a[r * 8 + c] = fmaf(-pr[r], pk[c], a[r * 8 + c]);
		} else {
#pragma unroll
			for (int r = 0; r < 8; r++) {
				const float2 mm = make_float2(-pr[r], -pr[r]);
#pragma unroll
				for (int q = 0; q < 4; q++) {
					float2 acc = ffma2(mm, make_float2(pk[2 * q], pk[2 * q + 1]), make_float2(a[r * 8 + 2 * q], a[r * 8 + 2 * q + 1]));
					a[r * 8 + 2 * q] = acc.x; a[r * 8 + 2 * q + 1] = acc.y;
				}
			}
		}
	}
	long long t1 = clock64();
	float s = 0; for (int i = 0; i < 64; i++) s += a[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0) cycles[0] = t1 - t0;
}
int main() {
	float* out; long long* cyc; cudaMalloc(&out, 4 * 1024); cudaMalloc(&cyc, 8);
	for (int threads : {32, 128, 256, 512}) {
		for (int mode = 0; mode < 2; mode++) {
			const int iters = 64; long long h;
			if (mode == 0) k<0><<<1, threads>>>(out, cyc, iters); else k<1><<<1, threads>>>(out, cyc, iters);
			cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
			printf("threads %4d %s: %6.2f cycles per 64-FMA-per-lane block (%d warps, %d per SMSP)\n", threads, mode ? "FFMA2" : "FFMA ", (double) h / iters, threads / 32, (threads / 32 + 3) / 4);
		}
	}
	return 0;
}
