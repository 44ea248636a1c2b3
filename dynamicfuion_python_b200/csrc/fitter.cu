// Host class behind the C ABI of nnrt::alignment::DeformableMeshToImageFitter
// (cpp/alignment/DeformableMeshToImageFitter.h:30-126, .cpp:56-329): a workspace sized once per problem shape, once-per-fit
// re-layout of the inputs into 128-bit SoA records, and the GN/LM loop issued as one CUDA graph (7 kernels / iteration,
// no allocation, no host sync inside the loop).
#include <vector>
#include <algorithm>
#include <mutex>
#include <cstring>
#include "fitter_kernels.cuh"
#include "fixed_accum.cuh"
#include "solver_kernels.cuh"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {

// ---- once-per-fit layout kernels -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_float4_kernel(const float* __restrict__ src, int64_t n, float4* __restrict__ dst) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] = make_float4(src[i * 3], src[i * 3 + 1], src[i * 3 + 2], 0.f);
}
__global__ void __launch_bounds__(256) pack_triangles_kernel(const int64_t* __restrict__ src, int64_t n, int4* __restrict__ dst) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] = make_int4((int) src[i * 3], (int) src[i * 3 + 1], (int) src[i * 3 + 2], 0);
}
__global__ void __launch_bounds__(256) pack_reference_kernel(const float* __restrict__ points, const uint8_t* __restrict__ mask, int64_t n,
                                                              float4* __restrict__ dst) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] = make_float4(points[i * 3], points[i * 3 + 1], points[i * 3 + 2], (mask == nullptr || mask[i]) ? 1.f : 0.f);
}
template<typename TDepth>
__global__ void __launch_bounds__(256) reference_from_depth_kernel(const TDepth* __restrict__ depth, const uint8_t* __restrict__ user_mask,
                                                                    int rows, int cols, Intrinsics k, float depth_scale, float depth_max,
                                                                    float4* __restrict__ dst) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (int64_t) rows * cols) return;
	const int y = (int) (i / cols), x = (int) (i % cols);
	const float d = (float) depth[i] / depth_scale;
	float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
	if (d > 0 && d < depth_max) {
		out = make_float4(((float) x - k.cx) * d / k.fx, ((float) y - k.cy) * d / k.fy, d, 1.f);
		if (user_mask && !user_mask[i]) out.w = 0.f;
	}
	dst[i] = out;
}
// ---- per-iteration: ARAP term + data term -> block-sparse arrowhead system ---------------------------------------
struct GraphDeviceData {
	const float* nodes; int node_count; int arrow_base;
	const int2* edges; int edge_count;
	const float* edge_weight_residual;  // lambda-free edge weight used by the residual (radii[layer[node_j]] or max coverage)
	const float* edge_weight_jacobian;  // edge weight used by the Jacobian (radii[layer[edge]] or max coverage)
	const int* out_offsets; const int* out_edges;  // CSR: edges by stem node i (== wing rows)
	const int* in_offsets; const int* in_edges;    // CSR: edges by corner node j
	float arap_weight; float lm_factor; int use_huber; float huber_constant;
	int iteration_mode;  // IterationMode of this iteration: 0 ALL, 1 TRANSLATION_ONLY, 2 ROTATION_ONLY (non-ALL: edge-free fits only)
};

__device__ __forceinline__ void arap_edge(const GraphDeviceData& g, const float* __restrict__ rotations, const float* __restrict__ translations,
                                          int e, float3& s, float& a, float3& residual) {
	const int2 ij = g.edges[e];
	const float3 gi = make_float3(g.nodes[ij.x * 3], g.nodes[ij.x * 3 + 1], g.nodes[ij.x * 3 + 2]);
	const float3 gj = make_float3(g.nodes[ij.y * 3], g.nodes[ij.y * 3 + 1], g.nodes[ij.y * 3 + 2]);
	const float3 ti = make_float3(translations[ij.x * 3], translations[ij.x * 3 + 1], translations[ij.x * 3 + 2]);
	const float3 tj = make_float3(translations[ij.y * 3], translations[ij.y * 3 + 1], translations[ij.y * 3 + 2]);
	float R[9];
#pragma unroll
	for (int k = 0; k < 9; k++) R[k] = rotations[ij.x * 9 + k];
	const float3 rd = matvec3(R, gi - gj);
	const float wj = g.edge_weight_jacobian[e], wr = g.edge_weight_residual[e];
	a = g.arap_weight * wj;                       // dE_i/dt_i = +a I, dE_j/dt_j = -a I   (ArapJacobianImpl.h:74-110)
	s = (-g.arap_weight * wj) * rd;               // dense rotation Jacobian, used as a skew generator
	residual = (g.arap_weight * wr) * (((gi + ti) - (gj + tj)) - rd);  // DeformableMeshToImageFitterImpl.h:697-713
	if (g.use_huber) {                            // DeformableMeshToImageFitter.cpp:435-444
		const float dl = g.huber_constant, half = 0.5f * dl * dl;
		residual.x = residual.x >= dl ? fabsf(residual.x) - half : 0.5f * residual.x * residual.x;
		residual.y = residual.y >= dl ? fabsf(residual.y) - half : 0.5f * residual.y * residual.y;
		residual.z = residual.z >= dl ? fabsf(residual.z) - half : 0.5f * residual.z * residual.z;
	}
}

// One thread per node: D_n = J_d^T J_d (from the pixel accumulators) + sum of ARAP blocks + lambda I, b_n = -J^T r, and the
// wing blocks B_e = dE_i^T dE_j of the node's outgoing edges (ArapHessianImpl.h:117-189, ComputeNegativeArapGradient
// DeformableMeshToImageFitterImpl.h:496-540, PreconditionDiagonalBlocksImpl.h). Replaces 8 launches + float atomics.
__global__ void __launch_bounds__(128) build_system_kernel(GraphDeviceData g, const float* __restrict__ rotations,
                                                            const float* __restrict__ translations, unsigned long long* __restrict__ accumulators,
                                                            float* __restrict__ diag, float* __restrict__ rhs, float* __restrict__ wing) {
	const int n = blockIdx.x * blockDim.x + threadIdx.x;
	if (n >= g.node_count) return;
	float H[36], b[6];
	unsigned long long* acc = accumulators + (int64_t) n * kAccumStride * 2;
	{
		int o = 0;
#pragma unroll
		for (int c0 = 0; c0 < 6; c0++)
#pragma unroll
			for (int c1 = c0; c1 < 6; c1++) { const float v = fixed_accum_value(acc[o * 2], acc[o * 2 + 1]); o++; H[c0 * 6 + c1] = v; H[c1 * 6 + c0] = v; }
#pragma unroll
		for (int c = 0; c < 6; c++) b[c] = 0.f - fixed_accum_value(acc[(21 + c) * 2], acc[(21 + c) * 2 + 1]);
#pragma unroll
		for (int k = 0; k < 54; k++) acc[k] = 0ull;  // clean for the next iteration's assembly
	}
	if (g.edge_count > 0) {
		if (n < g.arrow_base) {
			for (int q = g.out_offsets[n]; q < g.out_offsets[n + 1]; q++) {
				const int e = g.out_edges[q];
				float3 s, r; float a;
				arap_edge(g, rotations, translations, e, s, a, r);
				// dE_i = [S | a I] with S = skew(s); dE_j = [0 | -a I]
				const float S[9] = {0.f, -s.z, s.y, s.z, 0.f, -s.x, -s.y, s.x, 0.f};
				float dEi[18];
#pragma unroll
				for (int rr = 0; rr < 3; rr++)
#pragma unroll
					for (int c = 0; c < 3; c++) { dEi[rr * 6 + c] = S[rr * 3 + c]; dEi[rr * 6 + 3 + c] = rr == c ? a : 0.f; }
				float* B = wing + (int64_t) e * 36;
				const float bj = -a;
#pragma unroll
				for (int rr = 0; rr < 6; rr++)
#pragma unroll
					for (int c = 0; c < 6; c++) {
						H[rr * 6 + c] += sum3(dEi[rr] * dEi[c], dEi[6 + rr] * dEi[6 + c], dEi[12 + rr] * dEi[12 + c]);
						// dE_j column c: zero for c < 3, bj * unit(c-3) otherwise
						B[rr * 6 + c] = c < 3 ? 0.f : sum3(dEi[rr] * (c == 3 ? bj : 0.f), dEi[6 + rr] * (c == 4 ? bj : 0.f), dEi[12 + rr] * (c == 5 ? bj : 0.f));
					}
				const float rv[3] = {r.x, r.y, r.z};
#pragma unroll
				for (int c = 0; c < 6; c++) b[c] -= sum3(dEi[c] * rv[0], dEi[6 + c] * rv[1], dEi[12 + c] * rv[2]);
			}
		} else {
			const int cj = n - g.arrow_base;
			for (int q = g.in_offsets[cj]; q < g.in_offsets[cj + 1]; q++) {
				const int e = g.in_edges[q];
				float3 s, r; float a;
				arap_edge(g, rotations, translations, e, s, a, r);
				const float bj = -a;
				H[3 * 6 + 3] += bj * bj; H[4 * 6 + 4] += bj * bj; H[5 * 6 + 5] += bj * bj;
				b[3] -= bj * r.x; b[4] -= bj * r.y; b[5] -= bj * r.z;
			}
		}
	}
	if (g.lm_factor > 0.f) {
#pragma unroll
		for (int c = 0; c < 6; c++) H[c * 6 + c] += g.lm_factor;
	}
#pragma unroll
	for (int k = 0; k < 36; k++) diag[(int64_t) n * 36 + k] = H[k];
#pragma unroll
	for (int c = 0; c < 6; c++) rhs[n * 6 + c] = b[c];
}

// Edges between two coarser-layer nodes (hierarchies of more than two layers; general path only): the edge's ARAP blocks are added
// to both endpoints' diagonal blocks and right-hand sides (already formed by build_system_kernel) and dE_i^T dE_j becomes a
// corner_upper block. The reference writes these out of bounds (ArapHessianImpl.h:144-150); this is the intended arithmetic.
__global__ void __launch_bounds__(64) corner_edges_kernel(GraphDeviceData g, const int* __restrict__ edge_ids, int count, const float* __restrict__ rotations,
                                                           const float* __restrict__ translations, float* __restrict__ diag, float* __restrict__ rhs,
                                                           float* __restrict__ upper_blocks, int* __restrict__ upper_coordinates) {
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= count) return;
	const int e = edge_ids[k];
	const int2 ij = g.edges[e];
	float3 s, r; float a;
	arap_edge(g, rotations, translations, e, s, a, r);
	const float S[9] = {0.f, -s.z, s.y, s.z, 0.f, -s.x, -s.y, s.x, 0.f};
	float dEi[18];
#pragma unroll
	for (int rr = 0; rr < 3; rr++)
#pragma unroll
		for (int c = 0; c < 3; c++) { dEi[rr * 6 + c] = S[rr * 3 + c]; dEi[rr * 6 + 3 + c] = rr == c ? a : 0.f; }
	const float bj = -a;
	const float rv[3] = {r.x, r.y, r.z};
	const bool upper = ij.x < ij.y;  // block (i, j) sits in the upper triangle as is, otherwise its transpose at (j, i)
	upper_coordinates[k * 2] = upper ? ij.x : ij.y;
	upper_coordinates[k * 2 + 1] = upper ? ij.y : ij.x;
	for (int rr = 0; rr < 6; rr++)
		for (int c = 0; c < 6; c++) {
			atomicAdd(diag + (int64_t) ij.x * 36 + rr * 6 + c, sum3(dEi[rr] * dEi[c], dEi[6 + rr] * dEi[6 + c], dEi[12 + rr] * dEi[12 + c]));
			const float b = c < 3 ? 0.f : sum3(dEi[rr] * (c == 3 ? bj : 0.f), dEi[6 + rr] * (c == 4 ? bj : 0.f), dEi[12 + rr] * (c == 5 ? bj : 0.f));
			upper_blocks[(int64_t) k * 36 + (upper ? rr * 6 + c : c * 6 + rr)] = b;
		}
	for (int c = 0; c < 6; c++) atomicAdd(rhs + ij.x * 6 + c, -sum3(dEi[c] * rv[0], dEi[6 + c] * rv[1], dEi[12 + c] * rv[2]));
	for (int c = 0; c < 3; c++) {
		atomicAdd(diag + (int64_t) ij.y * 36 + (3 + c) * 7, bj * bj);
		atomicAdd(rhs + ij.y * 6 + 3 + c, -bj * rv[c]);
	}
}

// The fitter's system kernel: ONE WARP PER NODE, no atomics (ReducedSystem, solver_kernels.cuh).
//   stem node  : H = data-term block + sum of its edges' ARAP blocks + lambda I (same summation order as build_system_kernel),
//                D^-1 by six lanes (one unit column each), y = D^-1 b; per out-edge q: G_q = -a_q dE_i(q) (3 x 6, the non-zero
//                half of the wing block B_q = [0 | G_q^T]), Y_q = D^-1 G_q^T; per ordered edge pair the 3 x 3 Schur contribution
//                G_q1 Y_q2 and per edge the right-hand-side contribution G_q y, each WRITTEN to its own slot (the corner kernel
//                gathers them in a fixed order).
//   corner node: H = data-term block + sum over incoming edges (lanes stride over edges, shuffle reduction) + lambda I; its rotation
//                block couples to nothing but its own translation and is eliminated here: C' = C - B^T A^-1 B, r' = b_t - B^T A^-1 b_r.
// Replaces 8 launches + cuSOLVER / MAGMA batched inverses + 2 gemmBatched + atomic block sums of the reference.
constexpr int kStemWarps = 4;

__device__ __forceinline__ float skew_entry(float3 s, int r, int c) {
	// [[0,-sz,sy],[sz,0,-sx],[-sy,sx,0]]
	const float v = (r + c == 1) ? s.z : (r + c == 2) ? s.y : s.x;
	const bool neg = (r == 0 && c == 1) || (r == 1 && c == 2) || (r == 2 && c == 0);
	return r == c ? 0.f : (neg ? -v : v);
}

// ---- peer reduction of a pixel-sharded fit (fitter_kernels.cuh::PeerReduce), fused into the consumer of the sums -----------------
// Iteration k on every GPU: assemble_kernel adds into the GPU's own buffer k & 1; system_stem_kernel then (1) tells every peer "my sums
// of iteration k are complete" (flag k + 1 into the peer's flag row, a release store over NVLink -- the sums were finished by the
// previous kernel), (2) while that travels, evaluates its ARAP edges, (3) waits for the peers' flags in its OWN memory, (4) reads the
// node's 27 words from every rank's buffer k & 1 (peer loads) and adds them as integers -- the same total on every GPU, bit for bit,
// and equal to the single-GPU sums -- and (5) clears its own buffer (k - 1) & 1, which no peer can still be reading: a peer raises
// flag k + 1 only after its iteration k - 1 kernel has ended. A GPU can run at most one iteration ahead of a peer (it needs the peer's
// flag to finish an iteration), so two buffers suffice. The iteration counter lives in device memory and is advanced by the last warp
// of this kernel, so a captured graph replays correctly. Waits are bounded (kPeerWaitNs): a missing peer raises flag 4, never a hang.
constexpr unsigned long long kPeerWaitNs = 4000000000ull;

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
	asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
	unsigned long long v;
	asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ ulonglong2 ld_relaxed_sys_v2(const unsigned long long* p) {
	ulonglong2 v;
	asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
	unsigned long long t;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
	return t;
}
// every warp reports once it no longer needs this iteration's epoch; the last one advances it
__device__ __forceinline__ void peer_warp_done(const PeerReduce* peer, unsigned long long epoch, int lane) {
	if (lane != 0) return;
	__threadfence();
	if (atomicAdd(peer->done, 1u) == gridDim.x * (blockDim.x >> 5) - 1u) {
		*peer->done = 0u;
		*peer->epoch = epoch + 1ull;
	}
}

// kPeer = false is the single-GPU / all-reduce kernel, byte for byte what it was before the peer reduction existed (the peer table by
// value, or a local copy of it, gave this kernel a 300-byte stack frame and 8 more registers: 0.5 us on every single-GPU iteration);
// kPeer = true reads the table from device memory field by field.
template<bool kPeer>
__global__ void __launch_bounds__(kStemWarps * 32) system_stem_kernel(GraphDeviceData g, ReducedSystem red, const float* __restrict__ rotations,
                                                                      const float* __restrict__ translations, unsigned long long* __restrict__ accumulators,
                                                                      float* __restrict__ diag, float* __restrict__ rhs, const PeerReduce* __restrict__ peer) {
	__shared__ float s_H[kStemWarps][36], s_b[kStemWarps][8], s_dinv[kStemWarps][36], s_dE[kStemWarps][20], s_y[kStemWarps][8];
	__shared__ float s_G[kStemWarps][kMaxStemDegree][18], s_Y[kStemWarps][kMaxStemDegree][18];
	__shared__ int s_edge[kStemWarps][kMaxStemDegree];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const int node = blockIdx.x * kStemWarps + wib;
	unsigned long long epoch = 0ull;
	int world = 1, rank = 0;
	if constexpr (kPeer) {
		world = peer->world; rank = peer->rank;
		epoch = *reinterpret_cast<volatile unsigned long long*>(peer->epoch);
		__syncthreads();  // every warp of the CTA holds the epoch before any of them can report
		if (blockIdx.x == 0 && (int) threadIdx.x < world && (int) threadIdx.x != rank) {
			__threadfence_system();
			st_release_sys(peer->peer_flags[threadIdx.x] + rank, epoch + 1ull);
		}
	}
	if (node >= g.node_count) {
		if constexpr (kPeer) peer_warp_done(peer, epoch, lane);
		return;
	}
	float* H = s_H[wib];
	float* b = s_b[wib];
	// ---- stem node: lane q fetches and evaluates outgoing edge q, so the dependent loads of all edges (edge list -> endpoints ->
	// node state) are in flight together and overlap the accumulator read below; the blocks are then summed in edge order
	int e0 = 0, deg = 0, my_edge = 0;
	float3 my_sk = make_float3(0.f, 0.f, 0.f), my_r = make_float3(0.f, 0.f, 0.f);
	float my_a = 0.f;
	if (node < g.arrow_base) {
		e0 = g.out_offsets[node];
		deg = min(g.out_offsets[node + 1] - e0, kMaxStemDegree);
		if (lane < deg) {
			my_edge = g.out_edges[e0 + lane];
			arap_edge(g, rotations, translations, my_edge, my_sk, my_a, my_r);
		}
	}
	// ---- data term from the pixel accumulators (21 upper-triangular + 6), cleaned for the next iteration
	{
		ulonglong2* acc = reinterpret_cast<ulonglong2*>(accumulators) + (int64_t) node * kAccumStride;
		ulonglong2 words = make_ulonglong2(0ull, 0ull);
		if constexpr (kPeer) {
			if (lane < world && lane != rank) {
				const unsigned long long* flag = peer->flags + lane;
				const unsigned long long start = global_timer_ns();
				while (ld_acquire_sys(flag) < epoch + 1ull)
					if (global_timer_ns() - start > kPeerWaitNs) { atomicExch(red.flags + 4, 1); break; }
			}
			__syncwarp();
			const long long words_per_buffer = peer->parity_words;
			const long long current = (long long) (epoch & 1ull) * words_per_buffer, previous = (long long) ((epoch & 1ull) ^ 1ull) * words_per_buffer;
			const long long entry = ((long long) node * kAccumStride + lane) * 2;
			if (lane < 27) {
				for (int r = 0; r < world; r++) {
					const ulonglong2 w = ld_relaxed_sys_v2(peer->peer_accum[r] + current + entry);
					words.x += w.x; words.y += w.y;  // two's-complement and unsigned words: plain 64-bit adds, any order
				}
				*reinterpret_cast<ulonglong2*>(const_cast<unsigned long long*>(peer->peer_accum[rank]) + previous + entry) = make_ulonglong2(0ull, 0ull);
			}
			peer_warp_done(peer, epoch, lane);
		} else if (lane < 27) {
			words = acc[lane];
			acc[lane] = make_ulonglong2(0ull, 0ull);
		}
		if (lane < 27) {
			const float v = fixed_accum_value(words.x, words.y);
			if (lane < 21) {
				int c0 = 0, o = lane;
				while (o >= 6 - c0) { o -= 6 - c0; c0++; }
				const int c1 = c0 + o;
				H[c0 * 6 + c1] = v; H[c1 * 6 + c0] = v;
			} else {
				b[lane - 21] = 0.f - v;
			}
		}
	}
	__syncwarp();
	if (node < g.arrow_base) {
		if (lane < deg) s_edge[wib][lane] = my_edge;
		for (int q = 0; q < deg; q++) {
			const float3 sk = make_float3(__shfl_sync(0xFFFFFFFFu, my_sk.x, q), __shfl_sync(0xFFFFFFFFu, my_sk.y, q), __shfl_sync(0xFFFFFFFFu, my_sk.z, q));
			const float3 r = make_float3(__shfl_sync(0xFFFFFFFFu, my_r.x, q), __shfl_sync(0xFFFFFFFFu, my_r.y, q), __shfl_sync(0xFFFFFFFFu, my_r.z, q));
			const float a = __shfl_sync(0xFFFFFFFFu, my_a, q);
			// dE_i (3 x 6) = [skew(s) | a I] staged per warp
			if (lane < 18) { const int rr = lane / 6, c = lane % 6; s_dE[wib][lane] = c < 3 ? skew_entry(sk, rr, c) : (rr == c - 3 ? a : 0.f); }
			__syncwarp();
			const float* dE = s_dE[wib];
			const float bj = -a;
			for (int k = lane; k < 36; k += 32) {
				const int rr = k / 6, c = k % 6;
				H[k] += sum3(dE[rr] * dE[c], dE[6 + rr] * dE[6 + c], dE[12 + rr] * dE[12 + c]);
			}
			if (lane < 18) s_G[wib][q][lane] = dE[lane] * bj;  // G_q = dE_j^T-scaled dE_i: wing block B_q = [0 | G_q^T]
			if (lane < 6) b[lane] -= sum3(dE[lane] * r.x, dE[6 + lane] * r.y, dE[12 + lane] * r.z);
			__syncwarp();
		}
		if (lane < 6 && g.lm_factor > 0.f) H[lane * 7] += g.lm_factor;
		__syncwarp();
		if (g.iteration_mode != 0) {
			// TRANSLATION_ONLY / ROTATION_ONLY (DeformableMeshToImageFitter.cpp:198, 262-272): the reference solves the 3 x 3 block of the
			// active half. Decoupling the other half (identity block, zero right-hand side) makes the 6 x 6 factorisation perform the
			// same operations on the active block and adds exact zeros elsewhere.
			const int lo = g.iteration_mode == 1 ? 0 : 3;  // first row of the inactive half
			for (int k = lane; k < 36; k += 32) {
				const int rr = k / 6, c = k % 6;
				if ((rr >= lo && rr < lo + 3) || (c >= lo && c < lo + 3)) H[k] = rr == c ? 1.f : 0.f;
			}
			if (lane >= lo && lane < lo + 3) b[lane] = 0.f;
			__syncwarp();
		}
		for (int k = lane; k < 36; k += 32) diag[(int64_t) node * 36 + k] = H[k];
		if (lane < 6) rhs[node * 6 + lane] = b[lane];
		// ---- D^-1: every one of six lanes factors H (redundantly) and solves for its unit column
		if (lane < 6) {
			float L[36];
#pragma unroll
			for (int k = 0; k < 36; k++) L[k] = H[k];
			bool ok = true;
#pragma unroll
			for (int j = 0; j < 6; j++) {
				float dj = L[j * 6 + j];
#pragma unroll
				for (int k = 0; k < j; k++) dj -= L[j * 6 + k] * L[j * 6 + k];
				if (!(dj > 0.f)) ok = false;
				dj = sqrtf(dj);
				L[j * 6 + j] = dj;
#pragma unroll
				for (int i = j + 1; i < 6; i++) {
					float v = L[i * 6 + j];
#pragma unroll
					for (int k = 0; k < j; k++) v -= L[i * 6 + k] * L[j * 6 + k];
					L[i * 6 + j] = v / dj;
				}
			}
			if (!ok && lane == 0) atomicExch(red.flags + 2, 1);
			float xcol[6];
#pragma unroll
			for (int i = 0; i < 6; i++) {
				float v = (i == lane) ? 1.f : 0.f;
#pragma unroll
				for (int k = 0; k < i; k++) v -= L[i * 6 + k] * xcol[k];
				xcol[i] = v / L[i * 6 + i];
			}
#pragma unroll
			for (int i = 5; i >= 0; i--) {
				float v = xcol[i];
#pragma unroll
				for (int k = i + 1; k < 6; k++) v -= L[k * 6 + i] * xcol[k];
				xcol[i] = v / L[i * 6 + i];
			}
#pragma unroll
			for (int i = 0; i < 6; i++) s_dinv[wib][i * 6 + lane] = xcol[i];
		}
		__syncwarp();
		const float* Dinv = s_dinv[wib];
		if (lane < 6) {
			float acc = 0.f;
#pragma unroll
			for (int k = 0; k < 6; k++) acc += Dinv[lane * 6 + k] * b[k];
			s_y[wib][lane] = acc;
			red.stem_y[node * 6 + lane] = acc;
		}
		// Y_q = D^-1 G_q^T (6 x 3): kept in shared memory for the pair products, written out for the back substitution
		for (int idx = lane; idx < deg * 18; idx += 32) {
			const int q = idx / 18, k = idx % 18, rr = k / 3, c = k % 3;
			float acc = 0.f;
#pragma unroll
			for (int m = 0; m < 6; m++) acc += Dinv[rr * 6 + m] * s_G[wib][q][c * 6 + m];
			s_Y[wib][q][k] = acc;
			red.wing_y[(int64_t) s_edge[wib][q] * 18 + k] = acc;
		}
		__syncwarp();
		// Schur contributions of every ordered edge pair: (G_q1 Y_q2), 3 x 3, one slot each
		const int* slots = red.pair_slots + red.pair_offsets[node];
		for (int idx = lane; idx < deg * deg * 9; idx += 32) {
			const int pair = idx / 9, k = idx % 9, q1 = pair / deg, q2 = pair % deg, rr = k / 3, c = k % 3;
			const int slot = slots[pair];
			if (slot < 0) continue;  // col(q1) < col(q2): the mirror image of a stored pair
			float acc = 0.f;
#pragma unroll
			for (int m = 0; m < 6; m++) acc += s_G[wib][q1][rr * 6 + m] * s_Y[wib][q2][m * 3 + c];
			red.pair_blocks[(int64_t) slot * 12 + k] = acc;
		}
		// right-hand-side contributions G_q y
		for (int idx = lane; idx < deg * 3; idx += 32) {
			const int q = idx / 3, rr = idx % 3;
			float acc = 0.f;
#pragma unroll
			for (int m = 0; m < 6; m++) acc += s_G[wib][q][rr * 6 + m] * s_y[wib][m];
			red.edge_rhs[(int64_t) red.in_slots[s_edge[wib][q]] * 4 + rr] = acc;
		}
	} else {
		const int cj = node - g.arrow_base;
		float h_add = 0.f, b3 = 0.f, b4 = 0.f, b5 = 0.f;
		for (int q = g.in_offsets[cj] + lane; q < g.in_offsets[cj + 1]; q += 32) {
			float3 sk, r; float a;
			arap_edge(g, rotations, translations, g.in_edges[q], sk, a, r);
			const float bj = -a;
			h_add += bj * bj; b3 -= bj * r.x; b4 -= bj * r.y; b5 -= bj * r.z;
		}
#pragma unroll
		for (int off = 16; off >= 1; off >>= 1) {
			h_add += __shfl_xor_sync(0xFFFFFFFFu, h_add, off); b3 += __shfl_xor_sync(0xFFFFFFFFu, b3, off);
			b4 += __shfl_xor_sync(0xFFFFFFFFu, b4, off); b5 += __shfl_xor_sync(0xFFFFFFFFu, b5, off);
		}
		if (lane == 0) {
			H[21] += h_add; H[28] += h_add; H[35] += h_add;
			b[3] += b3; b[4] += b4; b[5] += b5;
		}
		__syncwarp();
		if (lane < 6 && g.lm_factor > 0.f) H[lane * 7] += g.lm_factor;
		__syncwarp();
		for (int k = lane; k < 36; k += 32) diag[(int64_t) node * 36 + k] = H[k];
		if (lane < 6) rhs[node * 6 + lane] = b[lane];
		if (lane == 0) {
			// eliminate the node's rotation: A = H[0:3,0:3] = L L^T, [A^-1 B | A^-1 b_r] by two triangular solves per column
			float L[9];
			bool ok = true;
#pragma unroll
			for (int j = 0; j < 3; j++) {
				float dj = H[j * 6 + j];
#pragma unroll
				for (int k = 0; k < j; k++) dj -= L[j * 3 + k] * L[j * 3 + k];
				if (!(dj > 0.f)) ok = false;
				dj = sqrtf(dj);
				L[j * 3 + j] = dj;
#pragma unroll
				for (int i = j + 1; i < 3; i++) {
					float v = H[i * 6 + j];
#pragma unroll
					for (int k = 0; k < j; k++) v -= L[i * 3 + k] * L[j * 3 + k];
					L[i * 3 + j] = v / dj;
				}
			}
			if (!ok) atomicExch(red.flags + 2, 1);
			float sol[4][3];  // columns 0..2: A^-1 B, column 3: A^-1 b_r
#pragma unroll
			for (int c = 0; c < 4; c++) {
#pragma unroll
				for (int i = 0; i < 3; i++) {
					float v = c < 3 ? H[i * 6 + 3 + c] : b[i];
#pragma unroll
					for (int k = 0; k < i; k++) v -= L[i * 3 + k] * sol[c][k];
					sol[c][i] = v / L[i * 3 + i];
				}
#pragma unroll
				for (int i = 2; i >= 0; i--) {
					float v = sol[c][i];
#pragma unroll
					for (int k = i + 1; k < 3; k++) v -= L[k * 3 + i] * sol[c][k];
					sol[c][i] = v / L[i * 3 + i];
				}
			}
#pragma unroll
			for (int rr = 0; rr < 3; rr++) {
#pragma unroll
				for (int c = 0; c < 3; c++) {
					float v = H[(3 + rr) * 6 + 3 + c];
#pragma unroll
					for (int k = 0; k < 3; k++) v -= H[k * 6 + 3 + rr] * sol[c][k];  // B^T (A^-1 B)
					red.corner_c[cj * 9 + rr * 3 + c] = v;
					red.corner_ainvb[cj * 9 + rr * 3 + c] = sol[c][rr];
				}
				float v = b[3 + rr];
#pragma unroll
				for (int k = 0; k < 3; k++) v -= H[k * 6 + 3 + rr] * sol[3][k];
				red.corner_r[cj * 3 + rr] = v;
				red.corner_zr[cj * 3 + rr] = sol[3][rr];
			}
		}
	}
}

__global__ void __launch_bounds__(256) diagnostics_convert_kernel(const int32_t* __restrict__ faces, const uint8_t* __restrict__ mask_in, int64_t n,
                                                                   int64_t* __restrict__ faces_out, uint8_t* __restrict__ mask_out) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (faces_out) faces_out[i] = faces[i];
	if (mask_out) mask_out[i] = mask_in[i];
}

}  // namespace nnrt_b200

// ---- host class --------------------------------------------------------------------------------------------------
struct nnrt_b200_fitter {
	nnrt_b200_fitter_options opt;
	int64_t V, F, E; int N, A, H, W;
	int64_t V_cap = 0, F_cap = 0;  // the workspace is sized for these; V, F are the counts of the mesh currently set (fitter_set_mesh_size)
	int arrow_base = 0, corner_blocks = 0;
	bool graph_set = false, mesh_set = false, reference_set = false, fused_stem = true;
	int row_begin = 0, row_end = 0;  // pixel band of a pixel-sharded fit; row_end == 0: whole image
	int stripe_rows = 8, stripe_index = 0, stripe_count = 1;  // striped pixel shard (fitter_kernels.cuh); 8, 0, 1: every row of the band
	// peer-reduced pixel shard: accumulators, flags and the iteration counter in a cudaMalloc'ed buffer of its own (it is exported to the
	// peers by IPC handle, which the arena must not be)
	char* exchange = nullptr; size_t exchange_bytes = 0;
	nnrt_b200::PeerReduce peer{};
	void* peer_mapped[nnrt_b200::kMaxPeers]{};  // bases opened with cudaIpcOpenMemHandle
	std::vector<int> modes{0};  // IterationMode sequence, cycled by iteration index (DeformableMeshToImageFitter.cpp:112)
	// arena
	char* arena = nullptr; size_t arena_bytes = 0, arena_used = 0;
	int* host_flags = nullptr;  // pinned
	// device views
	nnrt_b200::FitterDeviceData d{};
	nnrt_b200::GraphDeviceData g{};
	nnrt_b200::ArrowheadSystem sys{};
	float4 *positions, *normals, *wpos, *wnrm, *vjac, *refpts;
	int4* tris;
	int32_t* anchors; float* weights; float* nodes;
	int2* edges; float *ew_res, *ew_jac; int *out_offsets, *out_edges, *in_offsets, *in_edges, *wing_columns;
	unsigned long long* accum;
	float *sys_diag, *sys_rhs, *wing, *dinv, *stem_y, *wing_x, *schur, *corner_rhs, *x;
	// reduced (deterministic) solve: per-producer slots gathered by the corner kernel (solver_kernels.cuh ReducedSystem)
	float *wing_y, *edge_rhs, *corner_c, *corner_r, *corner_ainvb, *corner_zr;
	char* graph_buffer = nullptr;  // sized by the graph: pair_offsets, list_offsets, list_items, pair_blocks, dense corner (if beyond one SM)
	int *pair_offsets = nullptr, *list_offsets = nullptr, *pair_slots = nullptr, *in_slots = nullptr, listed_pairs = 0;
	int2* block_pairs = nullptr;
	int *corner_edge_ids = nullptr, *corner_upper_coords = nullptr, corner_edge_count = 0;
	float* corner_upper = nullptr;
	float *pair_blocks = nullptr, *reduced_packed = nullptr, *reduced_dense = nullptr, *reduced_dense_rhs = nullptr;
	unsigned long long* keys; nnrt_b200::FaceBandItem* queue; int* counters;
	int32_t* diag_faces; float* diag_res; uint8_t* diag_mask;
	float *ndc_x, *ndc_y;  // pixel centres in NDC (raster_core.cuh)
	// cached graph
	cudaGraphExec_t graph_exec = nullptr;
	struct GraphKey { float* R; float* t; double K[9]; double T[16]; int iters; } graph_key{};  // an instantiated graph launches into any stream

	template<typename T> T* take(size_t count) {
		const size_t bytes = (count * sizeof(T) + 255) & ~size_t(255);
		T* p = reinterpret_cast<T*>(arena + arena_used);
		arena_used += bytes;
		return p;
	}
};

namespace nnrt_b200 {

static size_t layout(nnrt_b200_fitter* f, bool assign) {
	// two passes: measure (arena == nullptr), then assign
	f->arena_used = 0;
	const int64_t V = f->V_cap, F = f->F_cap, E = f->E, PX = (int64_t) f->H * f->W;
	const int N = f->N, A = f->A;
	const int64_t nmax = (int64_t) N * 6;  // corner can be at most all nodes
	(void) assign;
	f->positions = f->take<float4>(V); f->normals = f->take<float4>(V);
	f->wpos = f->take<float4>(V); f->wnrm = f->take<float4>(V); f->vjac = f->take<float4>(V * A * 2);
	f->tris = f->take<int4>(F);
	f->anchors = f->take<int32_t>(V * A); f->weights = f->take<float>(V * A); f->nodes = f->take<float>((int64_t) N * 3);
	f->edges = f->take<int2>(E + 1); f->ew_res = f->take<float>(E + 1); f->ew_jac = f->take<float>(E + 1);
	f->out_offsets = f->take<int>(N + 2); f->out_edges = f->take<int>(E + 1);
	f->in_offsets = f->take<int>(N + 2); f->in_edges = f->take<int>(E + 1); f->wing_columns = f->take<int>(E + 1);
	f->refpts = f->take<float4>(PX);
	f->keys = f->take<unsigned long long>(PX);
	f->counters = f->take<int>(8);
	f->accum = f->take<unsigned long long>((int64_t) N * kAccumStride * 2);
	f->sys_diag = f->take<float>((int64_t) N * 36); f->sys_rhs = f->take<float>(nmax); f->wing = f->take<float>((E + 1) * 36);
	f->dinv = f->take<float>((int64_t) N * 36); f->stem_y = f->take<float>(nmax); f->wing_x = f->take<float>((E + 1) * 36);
	// dense corner: sized for the worst case the graph can ask for is N blocks; allocate lazily bounded by what set_model needs
	f->x = f->take<float>(nmax);
	f->wing_y = f->take<float>((E + 1) * 18); f->edge_rhs = f->take<float>((E + 1) * 4);
	f->corner_c = f->take<float>((int64_t) N * 9); f->corner_r = f->take<float>((int64_t) N * 3);
	f->corner_ainvb = f->take<float>((int64_t) N * 9); f->corner_zr = f->take<float>((int64_t) N * 3);
	f->diag_faces = f->take<int32_t>(PX); f->diag_res = f->take<float>(PX); f->diag_mask = f->take<uint8_t>(PX);
	f->ndc_x = f->take<float>(f->W); f->ndc_y = f->take<float>(f->H);
	f->queue = f->take<FaceBandItem>(F + 65536);  // last: the largest buffer of a small scene, kept away from the accumulators
	return f->arena_used;
}

}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" void nnrt_b200_fitter_default_options(nnrt_b200_fitter_options* o) {
	if (!o) return;
	o->max_iteration_count = 100; o->use_perspective_correction = 1; o->max_depth = 10.0f; o->use_tukey_penalty = 0;
	o->tukey_penalty_cutoff = 0.01f; o->preconditioning_dampening_factor = 0.0f; o->arap_term_weight = 200.f;
	o->use_huber_penalty = 0; o->huber_penalty_constant = 0.0001f;
}

extern "C" int nnrt_b200_fitter_create(const nnrt_b200_fitter_options* options, int64_t vertex_count, int64_t face_count, int32_t node_count,
                                       int64_t edge_count, int32_t anchor_count, int32_t image_height, int32_t image_width,
                                       nnrt_b200_fitter** out_fitter) {
	NNRT_CHECK_ARG(out_fitter, "fitter_create: null output");
	NNRT_CHECK_ARG(vertex_count > 0 && face_count > 0 && node_count > 0 && edge_count >= 0, "fitter_create: counts must be positive");
	NNRT_CHECK_ARG(anchor_count >= 1 && anchor_count <= kMaxAnchors, "anchor_count must be in [1, %d], got %d", kMaxAnchors, anchor_count);
	NNRT_CHECK_ARG(image_height > 0 && image_width > 0, "fitter_create: image size must be positive");
	NNRT_CHECK_ARG(node_count <= 32767, "node_count (currently, %d) cannot exceed int16_t max, which is 32767.", node_count);
	nnrt_b200_fitter_options opt;
	nnrt_b200_fitter_default_options(&opt);
	if (options) opt = *options;
	// cpp/alignment/DeformableMeshToImageFitter.cpp:79-82
	NNRT_CHECK_ARG(opt.preconditioning_dampening_factor >= 0.f && opt.preconditioning_dampening_factor <= 1.f,
	               "`preconditioning_dampening_factor` should be a small non-negative value between 0 and 1. Got %f.", opt.preconditioning_dampening_factor);
	auto* f = new nnrt_b200_fitter();
	f->opt = opt; f->V = f->V_cap = vertex_count; f->F = f->F_cap = face_count; f->E = edge_count; f->N = node_count; f->A = anchor_count;
	f->H = image_height; f->W = image_width;
	const size_t bytes = layout(f, false);
	if (cudaMalloc(&f->arena, bytes) != cudaSuccess) { delete f; set_last_error("fitter_create: cudaMalloc of %zu bytes failed", bytes); return kCudaError; }
	f->arena_bytes = bytes;
	layout(f, true);
	if (cudaMallocHost(&f->host_flags, 8 * sizeof(int)) != cudaSuccess) { cudaFree(f->arena); delete f; return kCudaError; }
	memset(f->host_flags, 0, 8 * sizeof(int));
	f->schur = nullptr; f->corner_rhs = nullptr;
	cudaMemset(f->arena, 0, bytes);
	launch_fill_pixel_centres(f->ndc_x, f->ndc_y, f->H, f->W, nullptr);
	launch_fill_keys(f->keys, (int64_t) f->H * f->W, nullptr);
	if (cudaDeviceSynchronize() != cudaSuccess) { nnrt_b200_fitter_destroy(f); set_last_error("fitter_create: workspace initialisation failed"); return kCudaError; }
	*out_fitter = f;
	return kOk;
}

// Exchange buffers of the peer reduction are exported to other processes (CUDA IPC). Freeing exported memory while an importer still
// has it mapped is undefined, and a fitter cannot know when its peers have unmapped it, so these buffers (~130 KB each at 249 nodes)
// are never returned to the driver: a destroyed fitter's buffer goes to a per-process free list and is handed to the next fitter that
// asks for the same size on the same device.
namespace {
struct PooledExchange { void* pointer; size_t bytes; int device; };
std::mutex g_exchange_mutex;
std::vector<PooledExchange> g_exchange_pool;

void release_exchange_buffer(void* pointer, size_t bytes) {
	if (!pointer) return;
	int device = 0;
	cudaPointerAttributes attributes{};
	if (cudaPointerGetAttributes(&attributes, pointer) == cudaSuccess) device = attributes.device; else cudaGetLastError();
	std::lock_guard<std::mutex> lock(g_exchange_mutex);
	g_exchange_pool.push_back({pointer, bytes, device});
}
void* acquire_exchange_buffer(size_t bytes) {
	int device = 0;
	cudaGetDevice(&device);
	{
		std::lock_guard<std::mutex> lock(g_exchange_mutex);
		for (size_t i = 0; i < g_exchange_pool.size(); i++)
			if (g_exchange_pool[i].bytes == bytes && g_exchange_pool[i].device == device) {
				void* pointer = g_exchange_pool[i].pointer;
				g_exchange_pool.erase(g_exchange_pool.begin() + (long) i);
				return pointer;
			}
	}
	void* pointer = nullptr;
	if (cudaMalloc(&pointer, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
	return pointer;
}
}  // namespace

extern "C" void nnrt_b200_fitter_destroy(nnrt_b200_fitter* f) {
	if (!f) return;
	nnrt_b200_fitter_disable_peer_reduce(f);  // waits for this fitter's kernels if they read peers' buffers, unmaps them
	if (f->graph_exec) { cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr; }
	release_exchange_buffer(f->exchange, f->exchange_bytes);
	cudaFree(f->arena);
	cudaFree(f->schur);  // corner buffers are allocated by set_model (their size depends on arrow_base)
	cudaFree(f->graph_buffer);
	cudaFreeHost(f->host_flags);
	delete f;
}

extern "C" int64_t nnrt_b200_fitter_workspace_bytes(const nnrt_b200_fitter* f) { return f ? (int64_t) f->arena_bytes : 0; }
extern "C" int32_t nnrt_b200_fitter_launches_per_iteration(const nnrt_b200_fitter* f) { return 7; (void) f; }

extern "C" int nnrt_b200_fitter_set_graph(nnrt_b200_fitter* f, const float* nodes, const int32_t* edges, int32_t arrow_base,
                                          const int8_t* edge_layer_indices, const float* layer_decimation_radii, int32_t layer_count,
                                          const float* node_coverage_weights, void* stream) {
	NNRT_CHECK_ARG(f && nodes, "fitter_set_graph: null pointer argument");
	auto s = (cudaStream_t) stream;
	const int64_t E = f->E;
	const int N = f->N;
	bool fused_stem = true;
	if (E > 0) {
		NNRT_CHECK_ARG(edges, "fitter_set_model: edge_count > 0 but edges is null");
		NNRT_CHECK_ARG((edge_layer_indices && layer_decimation_radii && layer_count > 0) || node_coverage_weights,
		               "fitter_set_model: edges need edge_layer_indices + layer_decimation_radii, or node_coverage_weights");
		NNRT_CHECK_ARG(arrow_base >= 1 && arrow_base <= N, "first_layer_node_count (currently, %d) must be in [1, node_count = %d].", arrow_base, N);
	} else {
		arrow_base = N;  // no regularisation edges: block-diagonal system (DeformableMeshToImageFitter.cpp:248-254)
	}
	NNRT_CUDA(cudaMemcpyAsync(f->nodes, nodes, sizeof(float) * N * 3, cudaMemcpyDeviceToDevice, s));
	// graph topology is static for the fit: build CSR + per-edge weights on the host once (the reference re-derives them per iteration)
	std::vector<int32_t> h_edges((size_t) E * 2);
	std::vector<int> out_off(N + 2, 0), in_off(N + 2, 0), out_e((size_t) E + 1), in_e((size_t) E + 1), wcol((size_t) E + 1);
	std::vector<float> ew_res((size_t) E + 1, 0.f), ew_jac((size_t) E + 1, 0.f);
	std::vector<int> corner_edges;  // edges between two coarser-layer nodes
	if (E > 0) {
		NNRT_CUDA(cudaMemcpyAsync(h_edges.data(), edges, sizeof(int32_t) * E * 2, cudaMemcpyDeviceToHost, s));
		std::vector<int8_t> h_layers;
		std::vector<float> h_radii, h_cov;
		const int64_t layer_len = E > N ? E : N;
		if (edge_layer_indices && layer_decimation_radii) {
			h_layers.resize((size_t) layer_len); h_radii.resize((size_t) layer_count);
			NNRT_CUDA(cudaMemcpyAsync(h_layers.data(), edge_layer_indices, (size_t) layer_len, cudaMemcpyDeviceToHost, s));
			NNRT_CUDA(cudaMemcpyAsync(h_radii.data(), layer_decimation_radii, sizeof(float) * layer_count, cudaMemcpyDeviceToHost, s));
		} else {
			h_cov.resize((size_t) N);
			NNRT_CUDA(cudaMemcpyAsync(h_cov.data(), node_coverage_weights, sizeof(float) * N, cudaMemcpyDeviceToHost, s));
		}
		NNRT_CUDA(cudaStreamSynchronize(s));
		const int nc = N - arrow_base;
		for (int64_t e = 0; e < E; e++) {
			const int i = h_edges[e * 2], j = h_edges[e * 2 + 1];
			// An edge from a first-layer node to a coarser node is a wing block: the only regime in which the reference's ARAP Hessian is
			// defined. An edge between two coarser nodes (hierarchies of more than two layers) is out of bounds
			// there; here it gets the evidently intended semantics: its blocks go to the corner (diagonal + corner_upper block).
			NNRT_CHECK_ARG(i >= 0 && i < N && j >= arrow_base && j < N && i != j,
			               "edge %lld = (%d, %d) must lead from any node to a different node of a coarser layer (>= %d)", (long long) e, i, j, arrow_base);
			wcol[e] = j;
			if (i >= arrow_base) corner_edges.push_back((int) e);
			else { out_off[i + 1]++; in_off[j - arrow_base + 1]++; }
			if (!h_layers.empty()) {
				const int lr = h_layers[j], lj = h_layers[e];  // sic: residual indexes the layer table by node j (…FitterImpl.h:708), Jacobian by edge
				NNRT_CHECK_ARG(lr >= 0 && lr < layer_count && lj >= 0 && lj < layer_count, "edge layer index out of range");
				ew_res[e] = h_radii[lr]; ew_jac[e] = h_radii[lj];
			} else {
				ew_res[e] = ew_jac[e] = h_cov[i] > h_cov[j] ? h_cov[i] : h_cov[j];
			}
		}
		int max_degree = 0;
		for (int i = 0; i < arrow_base; i++) max_degree = out_off[i + 1] > max_degree ? out_off[i + 1] : max_degree;
		fused_stem = max_degree <= kMaxStemDegree && corner_edges.empty();  // coarse-to-coarse edges take the general (6 Nc, corner_upper) path
		for (int i = 0; i < arrow_base; i++) out_off[i + 1] += out_off[i];
		for (int c = 0; c < nc; c++) in_off[c + 1] += in_off[c];
		std::vector<int> oc(out_off.begin(), out_off.end()), ic(in_off.begin(), in_off.end());
		for (int64_t e = 0; e < E; e++) {
			if (h_edges[e * 2] >= arrow_base) continue;
			out_e[oc[h_edges[e * 2]]++] = (int) e;
			in_e[ic[h_edges[e * 2 + 1] - arrow_base]++] = (int) e;
		}
		for (int i = 0; i < arrow_base; i++)
			for (int q = out_off[i]; q < out_off[i + 1]; q++)
				for (int q2 = q + 1; q2 < out_off[i + 1]; q2++)
					NNRT_CHECK_ARG(wcol[out_e[q]] != wcol[out_e[q2]], "duplicate edge (%d, %d)", i, wcol[out_e[q]]);
	}
	NNRT_CUDA(cudaMemcpyAsync(f->edges, h_edges.data(), sizeof(int32_t) * E * 2, cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->ew_res, ew_res.data(), sizeof(float) * (E + 1), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->ew_jac, ew_jac.data(), sizeof(float) * (E + 1), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->out_offsets, out_off.data(), sizeof(int) * (N + 2), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->in_offsets, in_off.data(), sizeof(int) * (N + 2), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->out_edges, out_e.data(), sizeof(int) * (E + 1), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->in_edges, in_e.data(), sizeof(int) * (E + 1), cudaMemcpyHostToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->wing_columns, wcol.data(), sizeof(int) * (E + 1), cudaMemcpyHostToDevice, s));
	// ---- reduced solve: slot of every ordered edge pair of a stem node, and per lower corner block pair (c1 >= c2) the list of the
	// pair slots that contribute to it, in ascending (stem node, q1, q2) order: the corner kernel sums them in exactly this order
	const int corner_blocks = N - arrow_base;
	std::vector<int> pair_off((size_t) arrow_base + 1, 0);
	for (int i = 0; i < arrow_base; i++) {
		const int deg = out_off[i + 1] - out_off[i];
		pair_off[i + 1] = pair_off[i] + (fused_stem ? deg * deg : 0);
	}
	// listed block pairs: every diagonal pair first (they carry the corner's own block), then the off-diagonal pairs (c1 > c2) that
	// share at least one stem node, in order of first appearance
	std::vector<int> pair_c1c2, list_off, pair_slot((size_t) pair_off[arrow_base] + 1, -1), in_slot((size_t) E + 1, 0);
	if (fused_stem) {
		std::vector<int> counts;
		std::vector<std::pair<long long, int>> seen;  // (c1 * Nc + c2, listed index) kept sorted for lookup
		for (int c = 0; c < corner_blocks; c++) { pair_c1c2.push_back(c); pair_c1c2.push_back(c); counts.push_back(0); }
		auto find_or_add = [&](int c1, int c2) -> int {
			if (c1 == c2) return c1;
			const long long key = (long long) c1 * corner_blocks + c2;
			auto it = std::lower_bound(seen.begin(), seen.end(), std::make_pair(key, -1));
			if (it != seen.end() && it->first == key) return it->second;
			const int idx = (int) counts.size();
			seen.insert(it, std::make_pair(key, idx));
			pair_c1c2.push_back(c1); pair_c1c2.push_back(c2); counts.push_back(0);
			return idx;
		};
		std::vector<int> item_pair((size_t) pair_off[arrow_base], -1);
		for (int i = 0; i < arrow_base; i++) {
			const int deg = out_off[i + 1] - out_off[i];
			for (int q1 = 0; q1 < deg; q1++)
				for (int q2 = 0; q2 < deg; q2++) {
					const int c1 = wcol[out_e[out_off[i] + q1]] - arrow_base, c2 = wcol[out_e[out_off[i] + q2]] - arrow_base;
					if (c1 < c2) continue;
					const int idx = find_or_add(c1, c2);
					item_pair[pair_off[i] + q1 * deg + q2] = idx;
					counts[idx]++;
				}
		}
		list_off.assign(counts.size() + 1, 0);
		for (size_t k = 0; k < counts.size(); k++) list_off[k + 1] = list_off[k] + counts[k];
		std::vector<int> cursor(list_off.begin(), list_off.end() - 1);
		for (size_t k = 0; k < item_pair.size(); k++)
			if (item_pair[k] >= 0) pair_slot[k] = cursor[item_pair[k]]++;
		for (int c = 0; c < corner_blocks; c++)
			for (int q = in_off[c]; q < in_off[c + 1]; q++) in_slot[in_e[q]] = q;
	} else {
		list_off.assign(1, 0);
	}
	const int listed_pairs = (int) list_off.size() - 1;
	{
		auto al = [](size_t v) { return (v + 255) & ~size_t(255); };
		const bool wide = fused_stem && !reduced_corner_fits_one_cta(corner_blocks);
		const size_t n3 = (size_t) corner_blocks * 3;
		const size_t sz_po = al(sizeof(int) * pair_off.size()), sz_lo = al(sizeof(int) * list_off.size()), sz_ps = al(sizeof(int) * pair_slot.size()),
		             sz_is = al(sizeof(int) * in_slot.size()), sz_pc = al(sizeof(int) * (pair_c1c2.size() + 2)),
		             sz_pb = al(sizeof(float) * ((size_t) list_off[listed_pairs] * 12 + 64)),
		             sz_pk = wide ? 0 : al(sizeof(float) * reduced_packed_floats(corner_blocks)),
		             sz_d = wide ? al(sizeof(float) * n3 * n3) : 0, sz_dr = wide ? al(sizeof(float) * (n3 + 16)) : 0;
		const size_t ce = corner_edges.size(), sz_ce = al(sizeof(int) * (ce + 1)), sz_cu = al(sizeof(float) * (ce * 36 + 36)), sz_cc = al(sizeof(int) * (ce * 2 + 2));
		if (f->graph_buffer) { cudaFree(f->graph_buffer); f->graph_buffer = nullptr; }
		const size_t total = sz_po + sz_lo + sz_ps + sz_is + sz_pc + sz_pb + sz_pk + sz_d + sz_dr + sz_ce + sz_cu + sz_cc;
		NNRT_CUDA(cudaMalloc(&f->graph_buffer, total));
		NNRT_CUDA(cudaMemsetAsync(f->graph_buffer, 0, total, s));  // the packed system's never-written entries (padding, strict upper tiles) stay zero
		char* q = f->graph_buffer;
		f->pair_offsets = (int*) q; q += sz_po;
		f->list_offsets = (int*) q; q += sz_lo;
		f->pair_slots = (int*) q; q += sz_ps;
		f->in_slots = (int*) q; q += sz_is;
		f->block_pairs = (int2*) q; q += sz_pc;
		f->pair_blocks = (float*) q; q += sz_pb;
		f->reduced_packed = wide ? nullptr : (float*) q; q += sz_pk;
		f->reduced_dense = wide ? (float*) q : nullptr; q += sz_d;
		f->reduced_dense_rhs = wide ? (float*) q : nullptr; q += sz_dr;
		f->corner_edge_ids = (int*) q; q += sz_ce;
		f->corner_upper = (float*) q; q += sz_cu;
		f->corner_upper_coords = (int*) q;
		f->corner_edge_count = (int) ce;
		if (ce > 0) NNRT_CUDA(cudaMemcpyAsync(f->corner_edge_ids, corner_edges.data(), sizeof(int) * ce, cudaMemcpyHostToDevice, s));
		f->listed_pairs = listed_pairs;
		NNRT_CUDA(cudaMemcpyAsync(f->pair_offsets, pair_off.data(), sizeof(int) * pair_off.size(), cudaMemcpyHostToDevice, s));
		NNRT_CUDA(cudaMemcpyAsync(f->list_offsets, list_off.data(), sizeof(int) * list_off.size(), cudaMemcpyHostToDevice, s));
		NNRT_CUDA(cudaMemcpyAsync(f->pair_slots, pair_slot.data(), sizeof(int) * pair_slot.size(), cudaMemcpyHostToDevice, s));
		NNRT_CUDA(cudaMemcpyAsync(f->in_slots, in_slot.data(), sizeof(int) * in_slot.size(), cudaMemcpyHostToDevice, s));
		if (!pair_c1c2.empty()) NNRT_CUDA(cudaMemcpyAsync(f->block_pairs, pair_c1c2.data(), sizeof(int) * pair_c1c2.size(), cudaMemcpyHostToDevice, s));
	}
	NNRT_CUDA(cudaStreamSynchronize(s));  // host staging vectors go out of scope

	// dense corner buffers of the unfused fallback (size depends on arrow_base)
	bool reallocated = false;
	if (corner_blocks != f->corner_blocks || f->schur == nullptr) {
		reallocated = true;
		if (f->schur) { cudaFree(f->schur); f->schur = nullptr; }
		const size_t n = (size_t) corner_blocks * 6;
		const size_t schur_floats = corner_schur_floats(corner_blocks);
		const size_t total = (schur_floats + n + 16) * sizeof(float);
		NNRT_CUDA(cudaMalloc(&f->schur, total));
		NNRT_CUDA(cudaMemset(f->schur, 0, total));
		f->corner_rhs = f->schur + schur_floats;
	}
	// the captured iteration graph bakes in the arrowhead shape and the corner buffers: keep it unless one of them moved
	(void) reallocated;
	if (f->graph_exec) {  // the reduced solve's lists and slots were just reallocated: the captured pointers are stale
		cudaGraphExecDestroy(f->graph_exec);
		f->graph_exec = nullptr;
	}
	f->arrow_base = arrow_base; f->corner_blocks = corner_blocks; f->fused_stem = fused_stem;
	f->graph_set = true;
	return kOk;
}

extern "C" int nnrt_b200_fitter_set_mesh(nnrt_b200_fitter* f, const float* vertex_positions, const float* vertex_normals,
                                         const int64_t* triangles, const int32_t* anchors, const float* anchor_weights, void* stream) {
	NNRT_CHECK_ARG(f, "fitter_set_mesh: null fitter");
	NNRT_CHECK_ARG(vertex_positions && triangles && anchors && anchor_weights, "fitter_set_mesh: null pointer argument");
	// cpp/alignment/DeformableMeshToImageFitter.cpp:346-348
	NNRT_CHECK_ARG(vertex_normals, "Input warped mesh needs to have normals defined; it doesn't.");
	auto s = (cudaStream_t) stream;
	const int64_t V = f->V, F = f->F;
	const int A = f->A;
	pack_float4_kernel<<<blocks_for(V, 256), 256, 0, s>>>(vertex_positions, V, f->positions);
	pack_float4_kernel<<<blocks_for(V, 256), 256, 0, s>>>(vertex_normals, V, f->normals);
	pack_triangles_kernel<<<blocks_for(F, 256), 256, 0, s>>>(triangles, F, f->tris);
	NNRT_CUDA(cudaMemcpyAsync(f->anchors, anchors, sizeof(int32_t) * V * A, cudaMemcpyDeviceToDevice, s));
	NNRT_CUDA(cudaMemcpyAsync(f->weights, anchor_weights, sizeof(float) * V * A, cudaMemcpyDeviceToDevice, s));
	NNRT_LAUNCH_CHECK();

	f->mesh_set = true;
	return kOk;
}

extern "C" int nnrt_b200_fitter_set_mesh_size(nnrt_b200_fitter* f, int64_t vertex_count, int64_t face_count) {
	NNRT_CHECK_ARG(f, "fitter_set_mesh_size: null fitter");
	NNRT_CHECK_ARG(vertex_count > 0 && face_count > 0, "fitter_set_mesh_size: counts must be positive");
	NNRT_CHECK_ARG(vertex_count <= f->V_cap && face_count <= f->F_cap, "the workspace holds up to %lld vertices and %lld faces, asked for %lld and %lld",
	               (long long) f->V_cap, (long long) f->F_cap, (long long) vertex_count, (long long) face_count);
	if (vertex_count == f->V && face_count == f->F) return kOk;
	if (f->graph_exec) { cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr; }  // launch sizes are part of the captured graph
	f->V = vertex_count; f->F = face_count;
	f->mesh_set = false;
	return kOk;
}

extern "C" int nnrt_b200_fitter_set_model(nnrt_b200_fitter* f, const float* vertex_positions, const float* vertex_normals,
                                          const int64_t* triangles, const float* nodes, const int32_t* edges, int32_t arrow_base,
                                          const int8_t* edge_layer_indices, const float* layer_decimation_radii, int32_t layer_count,
                                          const float* node_coverage_weights, const int32_t* anchors, const float* anchor_weights,
                                          void* stream) {
	NNRT_CHECK_ARG(f, "fitter_set_model: null fitter");
	NNRT_CHECK_ARG(vertex_positions && triangles && nodes && anchors && anchor_weights, "fitter_set_model: null pointer argument");
	const int rc = nnrt_b200_fitter_set_mesh(f, vertex_positions, vertex_normals, triangles, anchors, anchor_weights, stream);
	if (rc != kOk) return rc;
	return nnrt_b200_fitter_set_graph(f, nodes, edges, arrow_base, edge_layer_indices, layer_decimation_radii, layer_count, node_coverage_weights, stream);
}


extern "C" int nnrt_b200_fitter_set_reference_points(nnrt_b200_fitter* f, const float* reference_points, const uint8_t* reference_mask,
                                                     void* stream) {
	NNRT_CHECK_ARG(f && reference_points, "fitter_set_reference_points: null pointer argument");
	const int64_t px = (int64_t) f->H * f->W;
	// a new observation starts a new fit: its status must not inherit the overflow / not-positive-definite flags of an earlier one
	NNRT_CUDA(cudaMemsetAsync(f->counters + 1, 0, 4 * sizeof(int), (cudaStream_t) stream));
	pack_reference_kernel<<<blocks_for(px, 256), 256, 0, (cudaStream_t) stream>>>(reference_points, reference_mask, px, f->refpts);
	NNRT_LAUNCH_CHECK();
	f->reference_set = true;
	return kOk;
}

extern "C" int nnrt_b200_fitter_set_reference_depth(nnrt_b200_fitter* f, const void* depth, int32_t depth_dtype, const uint8_t* user_mask,
                                                    float depth_scale, const double* intrinsics_host, void* stream) {
	NNRT_CHECK_ARG(f && depth && intrinsics_host, "fitter_set_reference_depth: null pointer argument");
	NNRT_CHECK_ARG(depth_dtype == 0 || depth_dtype == 1, "depth dtype must be float32 (0) or uint16 (1)");
	const int64_t px = (int64_t) f->H * f->W;
	const Intrinsics k = intrinsics_from_host(intrinsics_host);
	auto s = (cudaStream_t) stream;
	NNRT_CUDA(cudaMemsetAsync(f->counters + 1, 0, 4 * sizeof(int), s));  // see set_reference_points
	if (depth_dtype == 0)
		reference_from_depth_kernel<float><<<blocks_for(px, 256), 256, 0, s>>>((const float*) depth, user_mask, f->H, f->W, k, depth_scale, f->opt.max_depth, f->refpts);
	else
		reference_from_depth_kernel<uint16_t><<<blocks_for(px, 256), 256, 0, s>>>((const uint16_t*) depth, user_mask, f->H, f->W, k, depth_scale, f->opt.max_depth, f->refpts);
	NNRT_LAUNCH_CHECK();
	f->reference_set = true;
	return kOk;
}

namespace nnrt_b200 {

struct IterationContext {
	FitterDeviceData d; GraphDeviceData g; ArrowheadSystem sys; ReducedSystem red; DataTermParams p; NdcCamera cam; Rigid extrinsics;
	PeerReduce peer; const PeerReduce* peer_device;  // host copy (world > 1: on) and the copy the stem kernel reads
};

// exchange buffer: [2 accumulator buffers][flag row, kMaxPeers words][iteration counter][done counter] ... [PeerReduce table at +512]
static size_t exchange_flags_offset(const nnrt_b200_fitter* f) { return ((size_t) f->N * kAccumStride * 2 * 2 * sizeof(unsigned long long) + 255) & ~size_t(255); }
static size_t exchange_table_offset(const nnrt_b200_fitter* f) { return exchange_flags_offset(f) + 512; }
static_assert(sizeof(PeerReduce) <= 512 && kMaxPeers * sizeof(unsigned long long) + 16 <= 512, "exchange buffer tail layout");

static IterationContext make_context(nnrt_b200_fitter* f, const double* K, const double* T) {
	IterationContext c{};
	c.d.positions = f->positions; c.d.normals = f->normals; c.d.triangles = f->tris; c.d.vertex_count = f->V; c.d.face_count = f->F;
	c.d.anchors = f->anchors; c.d.anchor_weights = f->weights; c.d.anchor_count = f->A;
	c.d.nodes = f->nodes; c.d.node_count = f->N; c.d.arrow_base = f->arrow_base;
	c.d.warped_positions = f->wpos; c.d.warped_normals = f->wnrm; c.d.vertex_jacobians = f->vjac;
	c.d.keys = f->keys; c.d.band_queue = f->queue; c.d.counters = f->counters; c.d.band_queue_capacity = (int) (f->F_cap + 65536);
	c.d.reference_points = f->refpts; c.d.accumulators = f->accum;
	c.peer = f->peer; c.peer_device = f->exchange ? reinterpret_cast<const PeerReduce*>(f->exchange + exchange_table_offset(f)) : nullptr;
	if (f->peer.world > 1) {  // the sums go to the exported buffer of this iteration's parity
		c.d.accumulators = reinterpret_cast<unsigned long long*>(f->exchange);
		c.d.accum_epoch = f->peer.epoch; c.d.accum_parity_words = f->peer.parity_words;
	}
	c.d.diag_pixel_faces = f->diag_faces; c.d.diag_residuals = f->diag_res; c.d.diag_mask = f->diag_mask;
	c.g.nodes = f->nodes; c.g.node_count = f->N; c.g.arrow_base = f->arrow_base; c.g.edges = f->edges; c.g.edge_count = (int) f->E;
	c.g.edge_weight_residual = f->ew_res; c.g.edge_weight_jacobian = f->ew_jac;
	c.g.out_offsets = f->out_offsets; c.g.out_edges = f->out_edges; c.g.in_offsets = f->in_offsets; c.g.in_edges = f->in_edges;
	c.g.arap_weight = f->opt.arap_term_weight; c.g.lm_factor = f->opt.preconditioning_dampening_factor;
	c.g.use_huber = f->opt.use_huber_penalty; c.g.huber_constant = f->opt.huber_penalty_constant;
	c.sys.node_count = f->N; c.sys.arrow_base = f->arrow_base; c.sys.diagonal_blocks = f->sys_diag; c.sys.rhs = f->sys_rhs;
	c.sys.wing_blocks = f->wing; c.sys.wing_columns = f->wing_columns; c.sys.row_offsets = f->out_offsets; c.sys.row_blocks = f->out_edges;
	c.sys.corner_upper_blocks = f->corner_upper; c.sys.corner_upper_coordinates = f->corner_upper_coords; c.sys.corner_upper_count = f->corner_edge_count;
	c.sys.inverted_stem = f->dinv; c.sys.stem_y = f->stem_y; c.sys.wing_x = f->wing_x; c.sys.schur = f->schur; c.sys.corner_rhs = f->corner_rhs;
	c.sys.schur_tile_packed = corner_uses_tile_packed_schur(f->N - f->arrow_base) ? 1 : 0;
	c.sys.flags = f->counters;
	c.red.node_count = f->N; c.red.arrow_base = f->arrow_base; c.red.row_offsets = f->out_offsets; c.red.row_blocks = f->out_edges;
	c.red.wing_columns = f->wing_columns; c.red.in_offsets = f->in_offsets; c.red.in_edges = f->in_edges;
	c.red.pair_offsets = f->pair_offsets; c.red.pair_slots = f->pair_slots; c.red.list_offsets = f->list_offsets; c.red.in_slots = f->in_slots;
	c.red.stem_y = f->stem_y; c.red.wing_y = f->wing_y; c.red.pair_blocks = f->pair_blocks; c.red.edge_rhs = f->edge_rhs;
	c.red.corner_c = f->corner_c; c.red.corner_r = f->corner_r; c.red.corner_ainvb = f->corner_ainvb; c.red.corner_zr = f->corner_zr;
	c.red.block_pair_count = f->listed_pairs; c.red.block_pairs = f->block_pairs; c.red.packed = f->reduced_packed;
	c.red.dense = f->reduced_dense; c.red.dense_rhs = f->reduced_dense_rhs; c.red.flags = f->counters;
	c.cam = ndc_camera_from_host(K, f->H, f->W);
	c.p.height = f->H; c.p.width = f->W; c.p.pinhole = intrinsics_from_host(K); c.p.pinhole_inverse_fx = 1.0f / c.p.pinhole.fx; c.p.pinhole_inverse_fy = 1.0f / c.p.pinhole.fy;
	c.p.ndc = c.cam.k; c.p.max_depth = f->opt.max_depth;
	c.p.perspective_correct = f->opt.use_perspective_correction; c.p.use_tukey = f->opt.use_tukey_penalty;
	c.p.tukey_cutoff = f->opt.tukey_penalty_cutoff;
	// RasterizeNdcTriangles(..., blur 0.5 px, ...) (DeformableMeshToImageFitter.cpp:124-127); ImageSpaceDistanceToNdc
	c.p.blur_radius = 0.5f / (fminf((float) f->H, (float) f->W) / 2.0f);
	c.p.row_begin = f->row_begin; c.p.row_end = f->row_end > 0 ? f->row_end : f->H;
	c.p.stripe_rows = f->stripe_rows; c.p.stripe_index = f->stripe_index; c.p.stripe_count = f->stripe_count;
	c.p.ndc_x = f->ndc_x; c.p.ndc_y = f->ndc_y;
	c.p.write_diagnostics = 1;
	c.extrinsics = rigid_from_host(T);
	return c;
}

// One GN/LM iteration = 7 launches. `ev` (optional, 8 events = 7 stages) brackets the stages for profile_iteration.
static int enqueue_data_term(const IterationContext& c, float* R, float* t, cudaStream_t s, cudaEvent_t* ev);
static int enqueue_solve(nnrt_b200_fitter* f, const IterationContext& base, int iteration_mode, float* R, float* t, cudaStream_t s,
                         cudaEvent_t* ev);

static int enqueue_iteration(nnrt_b200_fitter* f, const IterationContext& base, int iteration_mode, float* R, float* t, cudaStream_t s,
                             cudaEvent_t* ev) {
	const int rc = enqueue_data_term(base, R, t, s, ev);
	return rc != kOk ? rc : enqueue_solve(f, base, iteration_mode, R, t, s, ev);
}

// first half of an iteration: warp, rasterise, residual + Jacobian + per-node J^T J / J^T r sums into the accumulators
static int enqueue_data_term(const IterationContext& c, float* R, float* t, cudaStream_t s, cudaEvent_t* ev) {
	if (ev) cudaEventRecord(ev[0], s);
	launch_warp_vertices(c.d, R, t, c.extrinsics, s);
	if (ev) cudaEventRecord(ev[1], s);
	launch_face_raster(c.d, c.p, c.cam, s);
	if (ev) cudaEventRecord(ev[2], s);
	launch_band_raster(c.d, c.p, c.cam, s);
	if (ev) cudaEventRecord(ev[3], s);
	if (launch_assemble(c.d, c.p, s) != kOk) return kInvalidArgument;
	if (ev) cudaEventRecord(ev[4], s);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

// second half: ARAP term + system, arrowhead solve, node update (consumes and clears the accumulators)
static int enqueue_solve(nnrt_b200_fitter* f, const IterationContext& base, int iteration_mode, float* R, float* t, cudaStream_t s,
                         cudaEvent_t* ev) {
	IterationContext c = base;
	c.g.iteration_mode = iteration_mode;
	c.sys.update_mode = iteration_mode;
	if (f->fused_stem) {
		c.red.update_mode = iteration_mode;
		const unsigned stem_grid = (f->N + kStemWarps - 1) / kStemWarps;
		if (c.peer.world > 1) system_stem_kernel<true><<<stem_grid, kStemWarps * 32, 0, s>>>(c.g, c.red, R, t, f->accum, f->sys_diag, f->sys_rhs, c.peer_device);
		else system_stem_kernel<false><<<stem_grid, kStemWarps * 32, 0, s>>>(c.g, c.red, R, t, f->accum, f->sys_diag, f->sys_rhs, nullptr);
		if (ev) { cudaEventRecord(ev[5], s); cudaEventRecord(ev[6], s); }
		const int rc = launch_reduced_corner_solve(c.red, f->x, R, t, s);
		if (ev) cudaEventRecord(ev[7], s);
		if (rc != kOk) return rc;
		NNRT_LAUNCH_CHECK();
		return kOk;
	} else {  // stem rows with more than kMaxStemDegree edges: unfused pair of kernels
		if (c.peer.world > 1) { set_last_error("fitter: the peer reduction is fused into the stem kernel, which this graph (a stem row with more than %d edges) does not use; use the all-reduce path", kMaxStemDegree); return kInvalidArgument; }
		build_system_kernel<<<(f->N + 127) / 128, 128, 0, s>>>(c.g, R, t, f->accum, f->sys_diag, f->sys_rhs, f->wing);
		if (f->corner_edge_count > 0)
			corner_edges_kernel<<<(f->corner_edge_count + 63) / 64, 64, 0, s>>>(c.g, f->corner_edge_ids, f->corner_edge_count, R, t, f->sys_diag, f->sys_rhs,
			                                                                     f->corner_upper, f->corner_upper_coords);
		if (ev) cudaEventRecord(ev[5], s);
		launch_stem_eliminate(c.sys, s);
		if (ev) cudaEventRecord(ev[6], s);
	}
	const int rc = launch_corner_solve(c.sys, f->x, R, t, true, s);
	if (ev) cudaEventRecord(ev[7], s);
	if (rc != kOk) return rc;
	NNRT_LAUNCH_CHECK();
	return kOk;
}

}  // namespace nnrt_b200

extern "C" int nnrt_b200_fitter_set_iteration_modes(nnrt_b200_fitter* f, const int32_t* modes_host, int32_t mode_count) {
	NNRT_CHECK_ARG(f && modes_host && mode_count > 0, "fitter_set_iteration_modes: need at least one mode");
	std::vector<int> modes(modes_host, modes_host + mode_count);
	for (int m : modes) NNRT_CHECK_ARG(m >= 0 && m <= 2, "iteration mode %d is not one of ALL (0), TRANSLATION_ONLY (1), ROTATION_ONLY (2)", m);
	if (modes != f->modes && f->graph_exec) { cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr; }
	f->modes = modes;
	return kOk;
}

extern "C" int nnrt_b200_fitter_set_pixel_rows(nnrt_b200_fitter* f, int32_t row_begin, int32_t row_end) {
	NNRT_CHECK_ARG(f, "fitter_set_pixel_rows: null fitter");
	NNRT_CHECK_ARG(row_begin >= 0 && row_begin < row_end && row_end <= f->H, "pixel rows [%d, %d) must lie inside the image (height %d)", row_begin, row_end, f->H);
	if ((row_begin != f->row_begin || row_end != f->row_end) && f->graph_exec) { cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr; }
	f->row_begin = row_begin; f->row_end = row_end;
	return kOk;
}

extern "C" int nnrt_b200_fitter_set_pixel_stripes(nnrt_b200_fitter* f, int32_t stripe_rows, int32_t stripe_index, int32_t stripe_count) {
	NNRT_CHECK_ARG(f, "fitter_set_pixel_stripes: null fitter");
	NNRT_CHECK_ARG(stripe_rows > 0 && stripe_rows % 8 == 0, "stripe_rows must be a positive multiple of the 8-row assembly tile, got %d", stripe_rows);
	NNRT_CHECK_ARG(stripe_count >= 1 && stripe_index >= 0 && stripe_index < stripe_count, "stripe index %d outside [0, %d)", stripe_index, stripe_count);
	if ((stripe_rows != f->stripe_rows || stripe_index != f->stripe_index || stripe_count != f->stripe_count) && f->graph_exec) {
		cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr;
	}
	f->stripe_rows = stripe_rows; f->stripe_index = stripe_index; f->stripe_count = stripe_count;
	return kOk;
}

// ---- peer-reduced pixel shard ---------------------------------------------------------------------------------------------
extern "C" int nnrt_b200_fitter_peer_exchange(nnrt_b200_fitter* f, void** device_pointer, int64_t* byte_count, void* ipc_handle_out) {
	NNRT_CHECK_ARG(f, "fitter_peer_exchange: null fitter");
	nnrt_b200_fitter_disable_peer_reduce(f);  // the buffer (and the peer table in it) is cleared below: the mode restarts with enable_peer_reduce
	if (f->exchange == nullptr) {
		f->exchange_bytes = exchange_table_offset(f) + 512;
		f->exchange = static_cast<char*>(acquire_exchange_buffer(f->exchange_bytes));
		if (f->exchange == nullptr) { set_last_error("fitter_peer_exchange: cudaMalloc of %zu bytes failed", f->exchange_bytes); return kCudaError; }
	}
	// a fresh exchange: zero sums, no flags raised, iteration counter 0 (every rank does this before the handles are gathered)
	NNRT_CUDA(cudaDeviceSynchronize());
	NNRT_CUDA(cudaMemset(f->exchange, 0, f->exchange_bytes));
	NNRT_CUDA(cudaDeviceSynchronize());
	if (device_pointer) *device_pointer = f->exchange;
	if (byte_count) *byte_count = (int64_t) f->exchange_bytes;
	if (ipc_handle_out) {
		static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the C ABI states 64 bytes per IPC handle");
		cudaIpcMemHandle_t handle;
		NNRT_CUDA(cudaIpcGetMemHandle(&handle, f->exchange));
		memcpy(ipc_handle_out, &handle, sizeof(handle));
	}
	return kOk;
}

extern "C" int nnrt_b200_fitter_disable_peer_reduce(nnrt_b200_fitter* f) {
	if (!f) return kOk;
	if (f->peer.world > 1) {
		cudaDeviceSynchronize();  // this fitter's kernels read the peers' buffers: they must have ended before the mappings go
		if (f->graph_exec) { cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr; }
	}
	for (int r = 0; r < kMaxPeers; r++)
		if (f->peer_mapped[r]) { cudaIpcCloseMemHandle(f->peer_mapped[r]); f->peer_mapped[r] = nullptr; }
	f->peer = PeerReduce{};
	return kOk;
}

extern "C" int nnrt_b200_fitter_enable_peer_reduce(nnrt_b200_fitter* f, int32_t rank, int32_t world, void* const* exchange_pointers) {
	NNRT_CHECK_ARG(f && exchange_pointers, "fitter_enable_peer_reduce: null pointer argument");
	NNRT_CHECK_ARG(world >= 2 && world <= kMaxPeers && rank >= 0 && rank < world, "peer reduction needs 2..%d ranks and a rank inside them, got rank %d of %d", kMaxPeers, rank, world);
	NNRT_CHECK_ARG(f->exchange, "fitter_enable_peer_reduce: call fitter_peer_exchange first");
	NNRT_CHECK_ARG(!f->graph_set || f->fused_stem, "the peer reduction is fused into the stem kernel, which this graph (a stem row with more than %d edges) does not use", kMaxStemDegree);
	if (f->graph_exec) { cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr; }
	PeerReduce p{};
	p.world = world; p.rank = rank; p.parity_words = (long long) f->N * kAccumStride * 2;
	const size_t flags_at = exchange_flags_offset(f);
	for (int r = 0; r < world; r++) {
		char* base = r == rank ? f->exchange : static_cast<char*>(exchange_pointers[r]);
		NNRT_CHECK_ARG(base, "fitter_enable_peer_reduce: no exchange pointer for rank %d", r);
		p.peer_accum[r] = reinterpret_cast<const unsigned long long*>(base);
		p.peer_flags[r] = reinterpret_cast<unsigned long long*>(base + flags_at);
	}
	p.flags = p.peer_flags[rank];
	p.epoch = p.flags + kMaxPeers;
	p.done = reinterpret_cast<unsigned int*>(p.flags + kMaxPeers + 1);
	NNRT_CUDA(cudaMemcpy(f->exchange + exchange_table_offset(f), &p, sizeof(p), cudaMemcpyHostToDevice));
	f->peer = p;
	return kOk;
}

extern "C" int nnrt_b200_fitter_enable_peer_reduce_ipc(nnrt_b200_fitter* f, int32_t rank, int32_t world, const void* ipc_handles) {
	NNRT_CHECK_ARG(f && ipc_handles, "fitter_enable_peer_reduce_ipc: null pointer argument");
	NNRT_CHECK_ARG(world >= 2 && world <= kMaxPeers && rank >= 0 && rank < world, "peer reduction needs 2..%d ranks and a rank inside them, got rank %d of %d", kMaxPeers, rank, world);
	nnrt_b200_fitter_disable_peer_reduce(f);
	void* pointers[kMaxPeers] = {};
	for (int r = 0; r < world; r++) {
		if (r == rank) continue;
		cudaIpcMemHandle_t handle;
		memcpy(&handle, static_cast<const char*>(ipc_handles) + (size_t) r * sizeof(handle), sizeof(handle));
		const cudaError_t e = cudaIpcOpenMemHandle(&pointers[r], handle, cudaIpcMemLazyEnablePeerAccess);
		if (e != cudaSuccess) {
			cudaGetLastError();
			for (int q = 0; q < r; q++) if (pointers[q]) cudaIpcCloseMemHandle(pointers[q]);
			set_last_error("fitter_enable_peer_reduce_ipc: cannot map rank %d's exchange buffer (%s)", r, cudaGetErrorString(e));
			return kCudaError;
		}
	}
	const int rc = nnrt_b200_fitter_enable_peer_reduce(f, rank, world, pointers);
	if (rc != kOk) { for (int r = 0; r < world; r++) if (pointers[r]) cudaIpcCloseMemHandle(pointers[r]); return rc; }
	for (int r = 0; r < world; r++) f->peer_mapped[r] = pointers[r];
	return kOk;
}

extern "C" int64_t* nnrt_b200_fitter_accumulators(nnrt_b200_fitter* f, int64_t* word_count) {
	if (f == nullptr) return nullptr;
	if (word_count) *word_count = (int64_t) f->N * kAccumStride * 2;
	return reinterpret_cast<int64_t*>(f->accum);
}

extern "C" int nnrt_b200_fitter_iteration_data_term(nnrt_b200_fitter* f, float* node_rotations, float* node_translations,
                                                    const double* intrinsics_host, const double* extrinsics_host, void* stream) {
	NNRT_CHECK_ARG(f && node_rotations && node_translations && intrinsics_host, "fitter_iteration_data_term: null pointer argument");
	NNRT_CHECK_ARG(f->graph_set && f->mesh_set && f->reference_set, "fitter_iteration_data_term: model / reference not set");
	static const double I4[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
	const IterationContext c = make_context(f, intrinsics_host, extrinsics_host ? extrinsics_host : I4);
	return enqueue_data_term(c, node_rotations, node_translations, (cudaStream_t) stream, nullptr);
}

extern "C" int nnrt_b200_fitter_iteration_solve(nnrt_b200_fitter* f, float* node_rotations, float* node_translations,
                                                const double* intrinsics_host, const double* extrinsics_host, int32_t iteration_index,
                                                void* stream) {
	NNRT_CHECK_ARG(f && node_rotations && node_translations && intrinsics_host && iteration_index >= 0, "fitter_iteration_solve: bad argument");
	NNRT_CHECK_ARG(f->graph_set && f->mesh_set && f->reference_set, "fitter_iteration_solve: model / reference not set");
	static const double I4[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
	const IterationContext c = make_context(f, intrinsics_host, extrinsics_host ? extrinsics_host : I4);
	const int mode = f->modes[(size_t) iteration_index % f->modes.size()];
	NNRT_CHECK_ARG(mode == 0 || f->E == 0, "TRANSLATION_ONLY / ROTATION_ONLY iterations need a warp field without regularisation edges");
	const int rc = enqueue_solve(f, c, mode, node_rotations, node_translations, (cudaStream_t) stream, nullptr);
	if (rc == kOk) NNRT_CUDA(cudaMemcpyAsync(f->host_flags, f->counters, 5 * sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t) stream));
	return rc;
}

extern "C" int nnrt_b200_fitter_fit(nnrt_b200_fitter* f, float* node_rotations, float* node_translations, const double* intrinsics_host,
                                    const double* extrinsics_host, int32_t iteration_count, void* stream) {
	NNRT_CHECK_ARG(f && node_rotations && node_translations && intrinsics_host, "fitter_fit: null pointer argument");
	NNRT_CHECK_ARG(f->graph_set && f->mesh_set, "fitter_fit: set_model has not been called");
	NNRT_CHECK_ARG(f->reference_set, "fitter_fit: no reference observation set");
	if (iteration_count <= 0) iteration_count = f->opt.max_iteration_count;
	if (f->E > 0)
		for (int m : f->modes)
			NNRT_CHECK_ARG(m == 0, "TRANSLATION_ONLY / ROTATION_ONLY iterations need a warp field without regularisation edges: the reference's ARAP "
			                       "Hessian has 6 x 6 blocks only (ArapHessianImpl.h:54-56) and cannot be added to the 3 x 3 depth blocks of those modes");
	auto s = (cudaStream_t) stream;
	static const double I4[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
	const double* T = extrinsics_host ? extrinsics_host : I4;
	nnrt_b200_fitter::GraphKey key{};
	key.R = node_rotations; key.t = node_translations; key.iters = iteration_count;
	memcpy(key.K, intrinsics_host, sizeof(key.K)); memcpy(key.T, T, sizeof(key.T));
	if (f->graph_exec == nullptr || memcmp(&key, &f->graph_key, sizeof(key)) != 0) {
		if (f->graph_exec) { cudaGraphExecDestroy(f->graph_exec); f->graph_exec = nullptr; }
		const IterationContext c = make_context(f, intrinsics_host, T);
		cudaStream_t cs;
		NNRT_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
		NNRT_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
		int rc = kOk;
		for (int it = 0; it < iteration_count && rc == kOk; it++) {
			IterationContext ci = c;
			ci.p.write_diagnostics = it == iteration_count - 1 ? 1 : 0;  // get_diagnostics reports the last iteration
			rc = enqueue_iteration(f, ci, f->modes[(size_t) it % f->modes.size()], node_rotations, node_translations, cs, nullptr);
		}
		cudaGraph_t graph = nullptr;
		const cudaError_t end = cudaStreamEndCapture(cs, &graph);
		cudaStreamDestroy(cs);
		if (rc != kOk || end != cudaSuccess) {
			if (graph) cudaGraphDestroy(graph);
			if (rc == kOk) { set_last_error("fitter_fit: graph capture failed: %s", cudaGetErrorString(end)); rc = kCudaError; }
			return rc;
		}
		const cudaError_t inst = cudaGraphInstantiate(&f->graph_exec, graph, 0);
		cudaGraphDestroy(graph);
		if (inst != cudaSuccess) { f->graph_exec = nullptr; set_last_error("fitter_fit: graph instantiation failed: %s", cudaGetErrorString(inst)); return kCudaError; }
		f->graph_key = key;
	}
	NNRT_CUDA(cudaGraphLaunch(f->graph_exec, s));
	NNRT_CUDA(cudaMemcpyAsync(f->host_flags, f->counters, 5 * sizeof(int), cudaMemcpyDeviceToHost, s));
	return kOk;
}

// Waits for `stream` and reports what the last fit recorded on the device: band-queue overflow or a non-positive-definite block.
extern "C" int nnrt_b200_fitter_status(nnrt_b200_fitter* f, void* stream) {
	NNRT_CHECK_ARG(f, "fitter_status: null fitter");
	NNRT_CUDA(cudaStreamSynchronize((cudaStream_t) stream));
	if (f->host_flags[1]) { set_last_error("fitter: large-face band queue overflow"); return kQueueOverflow; }
	if (f->host_flags[2]) { set_last_error("fitter: a Hessian block was not positive definite"); return kNotPositiveDefinite; }
	if (f->host_flags[3]) { set_last_error("fitter: a data-term sum left the accumulator range (non-finite, or magnitude above 4e12)"); return kNotPositiveDefinite; }
	if (f->host_flags[4]) { set_last_error("fitter: peer reduction timed out waiting for another rank's sums (ranks out of step, or a peer stopped)"); return kCudaError; }
	return kOk;
}

extern "C" int nnrt_b200_fitter_get_diagnostics(nnrt_b200_fitter* f, float* hessian_diagonal, float* negative_gradient, float* deltas,
                                                int64_t* pixel_face_indices, float* residuals, uint8_t* residual_mask, void* stream) {
	NNRT_CHECK_ARG(f, "fitter_get_diagnostics: null fitter");
	auto s = (cudaStream_t) stream;
	const int64_t px = (int64_t) f->H * f->W;
	if (hessian_diagonal) NNRT_CUDA(cudaMemcpyAsync(hessian_diagonal, f->sys_diag, sizeof(float) * f->N * 36, cudaMemcpyDeviceToDevice, s));
	if (negative_gradient) NNRT_CUDA(cudaMemcpyAsync(negative_gradient, f->sys_rhs, sizeof(float) * f->N * 6, cudaMemcpyDeviceToDevice, s));
	if (deltas) NNRT_CUDA(cudaMemcpyAsync(deltas, f->x, sizeof(float) * f->N * 6, cudaMemcpyDeviceToDevice, s));
	if (residuals) NNRT_CUDA(cudaMemcpyAsync(residuals, f->diag_res, sizeof(float) * px, cudaMemcpyDeviceToDevice, s));
	if (pixel_face_indices || residual_mask) {
		diagnostics_convert_kernel<<<blocks_for(px, 256), 256, 0, s>>>(f->diag_faces, f->diag_mask, px, pixel_face_indices, residual_mask);
		NNRT_LAUNCH_CHECK();
	}
	return kOk;
}

extern "C" int nnrt_b200_fitter_profile_iteration(nnrt_b200_fitter* f, float* node_rotations, float* node_translations,
                                                  const double* intrinsics_host, const double* extrinsics_host, float* stage_ms7, void* stream) {
	NNRT_CHECK_ARG(f && node_rotations && node_translations && intrinsics_host && stage_ms7, "fitter_profile_iteration: null pointer argument");
	NNRT_CHECK_ARG(f->graph_set && f->mesh_set && f->reference_set, "fitter_profile_iteration: model / reference not set");
	static const double I4[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
	auto s = (cudaStream_t) stream;
	const IterationContext c = make_context(f, intrinsics_host, extrinsics_host ? extrinsics_host : I4);
	cudaEvent_t ev[8];
	for (auto& e : ev) NNRT_CUDA(cudaEventCreate(&e));
	const int rc = enqueue_iteration(f, c, f->modes[0], node_rotations, node_translations, s, ev);
	NNRT_CUDA(cudaStreamSynchronize(s));
	for (int k = 0; k < 7; k++) cudaEventElapsedTime(stage_ms7 + k, ev[k], ev[k + 1]);
	for (auto& e : ev) cudaEventDestroy(e);
	return rc;
}
