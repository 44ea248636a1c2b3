// Developer harness (not shipped): times corner_cluster_factor_kernel on a random SPD system of n unknowns, checks the factor against a
// host float64 Cholesky and prints the cycles each CTA rank spent per phase (-DNNRT_CLUSTER_TIMING).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -std=c++17 -DNNRT_CLUSTER_TIMING -o tools/corner_cluster_bench tools/corner_cluster_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <random>
__device__ long long g_cluster_clock[16 * 16];
#include "../dynamicfuion_python_b200/csrc/corner_cluster.cu"
namespace nnrt_b200 { void set_last_error(const char*, ...) {} }
using namespace nnrt_b200;

int main(int argc, char** argv) {
	const int n = argc > 1 ? atoi(argv[1]) : 867;
	std::mt19937 rng(3); std::normal_distribution<float> g(0.f, 1.f);
	// A = B B^T / k + n I-ish, well conditioned
	std::vector<double> A((size_t) n * n, 0.0);
	std::vector<float> B((size_t) n * 16);
	for (auto& v : B) v = g(rng);
	for (int i = 0; i < n; i++) for (int j = 0; j <= i; j++) { double a = 0; for (int k = 0; k < 16; k++) a += (double) B[i * 16 + k] * B[j * 16 + k]; A[(size_t) i * n + j] = a + (i == j ? 20.0 : 0.0); }
	std::vector<float> Af((size_t) n * n), rhs(n);
	for (size_t i = 0; i < Af.size(); i++) Af[i] = (float) A[i];
	for (auto& v : rhs) v = g(rng);
	// host Cholesky (lower) + forward substitution
	std::vector<double> L = A, y(rhs.begin(), rhs.end());
	for (int j = 0; j < n; j++) {
		double d = L[(size_t) j * n + j]; for (int k = 0; k < j; k++) d -= L[(size_t) j * n + k] * L[(size_t) j * n + k];
		d = std::sqrt(d); L[(size_t) j * n + j] = d;
		for (int i = j + 1; i < n; i++) { double v = L[(size_t) i * n + j]; for (int k = 0; k < j; k++) v -= L[(size_t) i * n + k] * L[(size_t) j * n + k]; L[(size_t) i * n + j] = v / d; }
	}
	for (int i = 0; i < n; i++) { double v = y[i]; for (int k = 0; k < i; k++) v -= L[(size_t) i * n + k] * y[k]; y[i] = v / L[(size_t) i * n + i]; }
	float *dS, *dS0, *drhs, *drhs0; int* flags;
	cudaMalloc(&dS, sizeof(float) * Af.size()); cudaMalloc(&dS0, sizeof(float) * Af.size()); cudaMalloc(&drhs, sizeof(float) * n); cudaMalloc(&drhs0, sizeof(float) * n);
	cudaMalloc(&flags, 64); cudaMemset(flags, 0, 64);
	cudaMemcpy(dS0, Af.data(), sizeof(float) * Af.size(), cudaMemcpyHostToDevice); cudaMemcpy(drhs0, rhs.data(), sizeof(float) * n, cudaMemcpyHostToDevice);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	float ms = 0; const int reps = 10;
	for (int rep = 0; rep < reps + 3; rep++) {
		cudaMemcpy(dS, dS0, sizeof(float) * Af.size(), cudaMemcpyDeviceToDevice); cudaMemcpy(drhs, drhs0, sizeof(float) * n, cudaMemcpyDeviceToDevice);
		if (rep == 3) { long long zero[256] = {}; cudaMemcpyToSymbol(g_cluster_clock, zero, sizeof(zero)); }
		cudaEventRecord(e0); int rc = launch_corner_cluster_factor(dS, drhs, n, flags, 0); cudaEventRecord(e1); cudaEventSynchronize(e1);
		if (rc) { printf("launch rc %d\n", rc); return 1; }
		float a; cudaEventElapsedTime(&a, e0, e1); if (rep >= 3) ms += a;
	}
	printf("cluster factor n = %d: %.2f us (%s)\n", n, 1e3 * ms / reps, cudaGetErrorString(cudaGetLastError()));
	std::vector<float> got(Af.size()), goty(n);
	cudaMemcpy(got.data(), dS, sizeof(float) * got.size(), cudaMemcpyDeviceToHost); cudaMemcpy(goty.data(), drhs, sizeof(float) * n, cudaMemcpyDeviceToHost);
	double worst = 0, worst_y = 0;
	for (int i = 0; i < n; i++) { for (int j = 0; j <= i; j++) worst = std::fmax(worst, std::fabs(got[(size_t) i * n + j] - L[(size_t) i * n + j])); worst_y = std::fmax(worst_y, std::fabs(goty[i] - y[i])); }
	printf("max |L - L64| = %.3e, max |y - y64| = %.3e\n", worst, worst_y);
	long long ph[256]; cudaMemcpyFromSymbol(ph, g_cluster_clock, sizeof(ph));
	const char* names[] = {"prologue", "substitute", "barrier 1", "owner store", "tiles", "barrier 2", "owner update", "owner factor", "sub:B", "sub:A+wait", "sub:mma", "sub:store"};
	for (int r = 0; r < 16; r++) { printf("rank %2d:", r); for (int k = 0; k < 12; k++) printf(" %s %6.1f", names[k], ph[r * 16 + k] / (double) reps / 1965.0); printf("\n"); }
	return 0;
}
