// Developer harness (not shipped): times the REDUCED corner kernel (3 Nc unknowns, gathered from per-producer slots) on a synthetic
// system with `base` stem nodes of degree 4 and `nc` corner nodes, and prints its phase breakdown (-DNNRT_PHASE_TIMING).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -std=c++17 -DNNRT_PHASE_TIMING -o tools/corner_reduced_bench tools/corner_reduced_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <random>
__device__ long long g_phase_clock[16];
__device__ long long g_warp_clock[32];
__device__ long long g_warp_abs[128];
#include "../dynamicfuion_python_b200/csrc/solver_kernels.cu"
namespace nnrt_b200 { void set_last_error(const char*, ...) {} }
using namespace nnrt_b200;
template<typename T> T* upload(const std::vector<T>& v) { T* d; cudaMalloc(&d, sizeof(T) * (v.size() + 64)); cudaMemcpy(d, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice); return d; }

int main(int argc, char** argv) {
	const int base = argc > 1 ? atoi(argv[1]) : 208, nc = argc > 2 ? atoi(argv[2]) : 41, deg = 4, N = base + nc, E = base * deg;
	std::mt19937 rng(7); std::normal_distribution<float> g(0.f, 1.f);
	std::vector<int> out_off(base + 1), out_e(E), wcol(E), in_off(nc + 1, 0), in_e(E), in_slot(E), pair_off(base + 1), pair_slot((size_t) base * deg * deg, -1);
	for (int i = 0; i < base; i++) { out_off[i] = i * deg; pair_off[i] = i * deg * deg; for (int e = 0; e < deg; e++) { out_e[i * deg + e] = i * deg + e; wcol[i * deg + e] = base + (i * 7 + e * 11) % nc; } }
	out_off[base] = E; pair_off[base] = base * deg * deg;
	for (int e = 0; e < E; e++) in_off[wcol[e] - base + 1]++;
	for (int c = 0; c < nc; c++) in_off[c + 1] += in_off[c];
	{ std::vector<int> cur(in_off.begin(), in_off.end() - 1); for (int e = 0; e < E; e++) { in_slot[e] = cur[wcol[e] - base]; in_e[cur[wcol[e] - base]++] = e; } }
	std::vector<int> pc, list_off, counts; std::vector<long long> keys;
	for (int c = 0; c < nc; c++) { pc.push_back(c); pc.push_back(c); counts.push_back(0); keys.push_back((long long) c * nc + c); }
	auto idx_of = [&](int c1, int c2) { long long k = (long long) c1 * nc + c2; for (size_t q = 0; q < keys.size(); q++) if (keys[q] == k) return (int) q; keys.push_back(k); pc.push_back(c1); pc.push_back(c2); counts.push_back(0); return (int) keys.size() - 1; };
	std::vector<int> item((size_t) base * deg * deg, -1);
	for (int i = 0; i < base; i++) for (int q1 = 0; q1 < deg; q1++) for (int q2 = 0; q2 < deg; q2++) { int c1 = wcol[i * deg + q1] - base, c2 = wcol[i * deg + q2] - base; if (c1 < c2 || (c1 == c2 && q1 != q2)) continue; int k = idx_of(c1, c2); item[(size_t) i * deg * deg + q1 * deg + q2] = k; counts[k]++; }
	list_off.assign(counts.size() + 1, 0); for (size_t k = 0; k < counts.size(); k++) list_off[k + 1] = list_off[k] + counts[k];
	{ std::vector<int> cur(list_off.begin(), list_off.end() - 1); for (size_t k = 0; k < item.size(); k++) if (item[k] >= 0) pair_slot[k] = cur[item[k]]++; }
	const int pairs = (int) counts.size();
	std::vector<float> pb((size_t) list_off[pairs] * 12), cc((size_t) nc * 9, 0.f), small((size_t) N * 18 * 4 + 64);
	for (auto& v : pb) v = 0.01f * g(rng);
	for (auto& v : small) v = 0.1f * g(rng);
	for (int c = 0; c < nc; c++) for (int d = 0; d < 3; d++) cc[c * 9 + d * 4] = 50.f;
	ReducedSystem red{};
	red.node_count = N; red.arrow_base = base; red.row_offsets = upload(out_off); red.row_blocks = upload(out_e); red.wing_columns = upload(wcol);
	red.in_offsets = upload(in_off); red.in_edges = upload(in_e); red.pair_offsets = upload(pair_off); red.pair_slots = upload(pair_slot);
	red.list_offsets = upload(list_off); red.in_slots = upload(in_slot); red.block_pair_count = pairs; red.block_pairs = (const int2*) upload(pc);
	red.packed = upload(std::vector<float>(reduced_packed_floats(nc), 0.f));
	red.stem_y = upload(small); red.wing_y = upload(std::vector<float>((size_t) E * 18, 0.01f)); red.pair_blocks = upload(pb); red.edge_rhs = upload(std::vector<float>((size_t) E * 4, 0.01f));
	red.corner_c = upload(cc); red.corner_r = upload(small); red.corner_ainvb = upload(small); red.corner_zr = upload(small);
	red.flags = upload(std::vector<int>(8, 0));
	float* x = upload(std::vector<float>((size_t) N * 6, 0.f));
	std::vector<float> eye((size_t) N * 9, 0.f); for (int i = 0; i < N; i++) eye[i * 9] = eye[i * 9 + 4] = eye[i * 9 + 8] = 1.f;
	float* R = upload(eye); float* t = upload(std::vector<float>((size_t) N * 3, 0.f));
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	float ms = 0; const int reps = 20;
	for (int rep = 0; rep < reps + 3; rep++) {
		cudaMemcpy(R, eye.data(), sizeof(float) * eye.size(), cudaMemcpyHostToDevice);
		cudaEventRecord(e0); int rc = launch_reduced_corner_solve(red, x, R, t, 0); cudaEventRecord(e1); cudaEventSynchronize(e1);
		if (rc) { printf("launch rc %d\n", rc); return 1; }
		float a; cudaEventElapsedTime(&a, e0, e1); if (rep >= 3) ms += a;
	}
	printf("reduced corner (%d stem, %d corner, n = %d): %.2f us (%s)\n", base, nc, nc * 3, 1e3 * ms / reps, cudaGetErrorString(cudaGetLastError()));
	long long ph[16]; cudaMemcpyFromSymbol(ph, g_phase_clock, sizeof(ph));
	const char* names[] = {"gather+stage", "panels", "trailing", "backsub", "stem backsub", "node update", "block inverses", "gather only"};
	for (int k = 0; k < 8; k++) printf("  %-14s %8lld cycles\n", names[k], ph[k]);
	std::vector<float> hx((size_t) N * 6); cudaMemcpy(hx.data(), x, sizeof(float) * N * 6, cudaMemcpyDeviceToHost);
	double s = 0; for (float v : hx) s += v; printf("  checksum %.6f\n", s);
	return 0;
}
