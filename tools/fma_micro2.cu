// Developer microbenchmark: the register-tile rank-8 update (LDS.128 operands + 512 FFMA) at several warp counts.
#include <cstdio>
#include <cuda_runtime.h>
template<int UNROLL_ALL>
__global__ void k(float* out, long long* cycles, int iters, int prow) {
	extern __shared__ __align__(16) float PT[];
	for (int i = threadIdx.x; i < 8 * prow; i += blockDim.x) PT[i] = 1e-3f * (i % 97);
	float a[8][8];
	for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) a[r][c] = threadIdx.x * 0.001f + r * 8 + c;
	const int row0 = 8 * (threadIdx.x % 31), col0 = 8 * (threadIdx.x / 31 % 31);
	__syncthreads();
	long long t0 = clock64();
	for (int it = 0; it < iters; it++) {
		if (UNROLL_ALL) {
#pragma unroll
			for (int c = 0; c < 8; c++) {
				const float4 r0 = *reinterpret_cast<const float4*>(PT + c * prow + row0), r1 = *reinterpret_cast<const float4*>(PT + c * prow + row0 + 4);
				const float4 k0 = *reinterpret_cast<const float4*>(PT + c * prow + col0), k1 = *reinterpret_cast<const float4*>(PT + c * prow + col0 + 4);
				const float pr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
				const float pk[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
#pragma unroll
				for (int r = 0; r < 8; r++)
#pragma unroll
					for (int cc = 0; cc < 8; cc++) a[r][cc] = fmaf(-pr[r], pk[cc], a[r][cc]);
			}
		} else {
#pragma unroll 1
			for (int c = 0; c < 8; c++) {
				const float4 r0 = *reinterpret_cast<const float4*>(PT + c * prow + row0), r1 = *reinterpret_cast<const float4*>(PT + c * prow + row0 + 4);
				const float4 k0 = *reinterpret_cast<const float4*>(PT + c * prow + col0), k1 = *reinterpret_cast<const float4*>(PT + c * prow + col0 + 4);
				const float pr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
				const float pk[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
#pragma unroll
				for (int r = 0; r < 8; r++)
#pragma unroll
					for (int cc = 0; cc < 8; cc++) a[r][cc] = fmaf(-pr[r], pk[cc], a[r][cc]);
			}
		}
		__syncthreads();
	}
	long long t1 = clock64();
	float s = 0; for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) s += a[r][c];
	out[threadIdx.x] = s;
	if (threadIdx.x == 0) cycles[0] = t1 - t0;
}
int main() {
	float* out; long long* cyc; cudaMalloc(&out, 4 * 1024); cudaMalloc(&cyc, 8);
	const int prow = 248, smem = 8 * prow * 4;
	for (int threads : {32, 128, 256, 512}) {
		for (int mode = 0; mode < 2; mode++) {
			const int iters = 31; long long h;
			if (mode == 0) k<0><<<1, threads, smem>>>(out, cyc, iters, prow); else k<1><<<1, threads, smem>>>(out, cyc, iters, prow);
			cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
			printf("threads %4d %s: %7.1f cycles per panel-update (512 FFMA + 32 LDS.128 per lane, + barrier)\n", threads, mode ? "unrolled" : "rolled  ", (double) h / iters);
		}
	}
	printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
	return 0;
}
