// Block-sparse arrowhead solve on device (core::linalg::SolveBlockSparseArrowheadCholesky,
// cpp/core/linalg/SolveBlockSparseArrowheadCholesky.cpp:31-94) in two launches:
//   stem_eliminate: one warp per block row. Stem rows: 6x6 Cholesky inverse D^-1, y = D^-1 b, X_e = D^-1 B_e for the row's wing
//                   blocks, Schur update S -= B_e1^T X_e2 and rhs update b_C -= X_e^T b (float atomics into the dense corner).
//                   Corner rows: S_diag += C_jj, b_C += b_j.
//   corner_solve:   Cholesky of S with the right-hand side carried as an extra row (forward substitution for free), back
//                   substitution, stem back-substitution x_i = y_i - sum_e X_e x_C, optional fused node update (Rodrigues,
//                   R <- R dR, t += dt). Up to 246 unknowns: one CTA, one 8 x 8 tile per thread in registers; up to what one SM's
//                   shared memory holds: one CTA, blocked in shared memory; beyond: per-panel launches over the dense buffer.
// This replaces cuSOLVER potrfBatched + MAGMA potrs_batched + 5 cuBLAS gemmBatched + potrf + 2 trsm + their host syncs.
#pragma once
#include "host_util.h"

namespace nnrt_b200 {

struct ArrowheadSystem {
	int node_count, arrow_base;
	const float* diagonal_blocks;   // [N,36]
	const float* rhs;               // [6N]
	const float* wing_blocks;       // [W,36]
	const int* wing_columns;        // [W] global block column (>= arrow_base)
	const int* row_offsets;         // [arrow_base+1] CSR over wing blocks by stem row
	const int* row_blocks;          // [W] wing block ids grouped by row
	const float* corner_upper_blocks; const int* corner_upper_coordinates; int corner_upper_count;
	// workspace
	float* inverted_stem;           // [arrow_base,36]
	float* stem_y;                  // [arrow_base,6]
	float* wing_x;                  // [W,36]  D^-1 B
	float* schur;                   // dense [(6Nc)^2], or tile-packed lower triangle + rhs row (schur_tile_packed); zero on entry
	float* corner_rhs;              // [6Nc], must be zero on entry (unused when schur_tile_packed)
	int schur_tile_packed;          // 1: the register-tiled corner kernel consumes `schur` as warp-interleaved 8 x 8 tiles (see corner_tile_index)
	int* flags;                     // flags[2] set when a block is not positive definite
	int update_mode;                // node update applied by the corner kernels: 0 rotation + translation, 1 translation only, 2 rotation only
};

// ---- corner layouts ---------------------------------------------------------------------------------------------------------
constexpr int kTile = 8;
constexpr int kCornerThreads = 512;
constexpr int kRegTileMaxRows = 256;

// Tile-packed Schur layout consumed by corner_solve_regtile_kernel: 8 x 8 tiles (bi >= bj) of the augmented lower triangle
// (row n = right-hand side), block columns in DESCENDING order, rows ascending inside a column; tile t is thread t's tile.
__host__ __device__ __forceinline__ int corner_tile_rows(int n) { return (n + 1 + kTile - 1) / kTile; }
__host__ __device__ __forceinline__ int corner_tile_cols(int n) { return (n + kTile - 1) / kTile; }
__host__ __device__ __forceinline__ int corner_tile_count(int n) {
	const int tr = corner_tile_rows(n), tc = corner_tile_cols(n);
	return tc * tr - tc * (tc - 1) / 2;
}
// Block column 0 (never updated, only factored) belongs to warp 0, which doubles as the look-ahead warp; the other block columns
// follow from thread 32 on in DESCENDING order, rows ascending inside a column.
__host__ __device__ __forceinline__ int corner_tile_id(int n, int bi, int bj) {
	if (bj == 0) return bi;
	const int tr = corner_tile_rows(n), tc = corner_tile_cols(n), m = tc - 1 - bj;
	return 32 + m * tr - (bj + tc) * m / 2 + (bi - bj);
}
__host__ __device__ __forceinline__ int corner_tile_slots(int n) { return 32 + corner_tile_count(n) - corner_tile_rows(n); }
// offset of element (row, col), row in [0, n], col in [0, n), row / 8 >= col / 8. Tiles are interleaved in groups of 32 (one
// warp of the corner kernel): 128-bit quad q = 2 * (row % 8) + (col % 8) / 4 of tile t sits at quad ((t / 32) * 16 + q) * 32 + t % 32,
// so a warp's q-th load instruction reads 512 contiguous bytes and every lane receives its own tile directly.
__host__ __device__ __forceinline__ int corner_tile_index(int n, int row, int col) {
	const int t = corner_tile_id(n, row >> 3, col >> 3), q = ((row & 7) << 1) | ((col & 7) >> 2);
	return ((((t >> 5) << 4) + q) * 32 + (t & 31)) * 4 + (col & 3);
}
__host__ __device__ __forceinline__ int corner_tile_floats(int n) { return ((corner_tile_slots(n) + 31) / 32) * 32 * 64; }
// where the stem kernel accumulates S(row, col): -1 = not stored in this layout (strict upper tile triangle)
__device__ __forceinline__ int64_t schur_slot(const ArrowheadSystem& sys, int n, int row, int col) {
	if (!sys.schur_tile_packed) return (int64_t) row * n + col;
	return (col > row) ? -1 : corner_tile_index(n, row, col);
}

// ---- fitter path: deterministic, reduced corner -------------------------------------------------------------------------------
// In the fitter's system the ARAP residual of an edge (i -> j) depends on R_i, t_i and t_j only (DeformableMeshToImageFitterImpl.h:
// 697-713) and the data term couples nothing across nodes, so (two-layer graphs) the ROTATION unknowns of a corner node couple to
// nothing but that node's own translation: they are eliminated block-wise like the stem, which leaves a dense system over the
// corner TRANSLATIONS only (3 Nc unknowns instead of 6 Nc: an eighth of the Cholesky flops, half its dependency chain). Every sum
// that the stem elimination used to scatter with float atomics is written once per producer (pair_blocks, edge_rhs) and gathered by
// the corner kernel in a fixed order through lists built once per graph on the host: bit-identical results run to run.
struct ReducedSystem {
	int node_count, arrow_base;
	const int* row_offsets; const int* row_blocks;   // CSR over out-edges (edge ids) by stem node
	const int* wing_columns;                          // [E] global index of the edge's corner node
	const int* in_offsets; const int* in_edges;       // CSR over in-edges by corner node, ascending edge id
	const int* pair_offsets;                          // [stem + 1] prefix sums of deg_i^2
	const int* pair_slots;                            // [sum deg^2] slot in pair_blocks of the ordered edge pair (q1, q2) of a stem node, -1: not stored
	int block_pair_count;                             // lower corner block pairs (c1 >= c2) that receive a product, plus every diagonal pair
	const int2* block_pairs;                          // [block_pair_count] (c1, c2)
	const int* list_offsets;                          // [block_pair_count + 1] CONTIGUOUS slot range of each listed block pair
	const int* in_slots;                              // [E] position of edge e in the in-edge order of its corner node (slot in edge_rhs)
	float* stem_y;        // [stem, 6]   y_i = D_i^-1 b_i
	float* wing_y;        // [E, 18]     Y_e = D_i^-1 G_e^T (6 x 3), G_e = -a_e dE_i(e) (3 x 6)
	float* pair_blocks;   // [slots, 12]  G_q1 Y_q2 (3 x 3 + pad) of the edge pairs of the stem nodes with col(q1) >= col(q2), grouped by block pair
	float* edge_rhs;      // [E, 4]      G_e y_i (+ pad), grouped by corner node
	float* corner_c;      // [Nc, 9]     C' = C - B^T A^-1 B (translation block after eliminating the node's rotation)
	float* corner_r;      // [Nc, 3]     r' = b_t - B^T A^-1 b_r
	float* corner_ainvb;  // [Nc, 9]     A^-1 B
	float* corner_zr;     // [Nc, 3]     A^-1 b_r
	float* packed;        // tile-packed lower triangle + rhs row of the reduced corner system (corner_tile_index with n = 3 Nc): one SM
	float* dense; float* dense_rhs;  // [(3 Nc)^2], [3 Nc]: corners beyond one SM (multi-kernel path); exactly one of packed / dense is set
	int* flags;           // flags[2] set when a block is not positive definite
	int update_mode;
};
bool reduced_corner_fits_one_cta(int corner_blocks);
size_t reduced_packed_floats(int corner_blocks);  // allocation size of ReducedSystem::packed (zero-initialised once)
int launch_reduced_corner_solve(const ReducedSystem& sys, float* x, float* node_rotations, float* node_translations, cudaStream_t s);

// corner_cluster.cu: in-place Cholesky + forward substitution of a dense corner system (lower triangle of schur [n, n], corner_rhs [n])
// by one thread-block cluster; kUnsupported when the device cannot co-schedule one.
int launch_corner_cluster_factor(float* schur, float* corner_rhs, int n, int* flags, cudaStream_t s);
// back substitution with the block inverses the cluster kernel left behind, stem back substitution and node update (one CTA);
// `red` != nullptr: the fitter's reduced system (sys carries schur / corner_rhs / flags only).
int launch_corner_cluster_finish(const ArrowheadSystem& sys, const ReducedSystem* red, int n, float* x, float* node_rotations, float* node_translations,
                                 cudaStream_t s);

void launch_stem_eliminate(const ArrowheadSystem& sys, cudaStream_t s);
// x [6N]; if node_rotations != nullptr the update is applied in place; clear_corner zeroes schur / corner_rhs after reading.
int launch_corner_solve(const ArrowheadSystem& sys, float* x, float* node_rotations, float* node_translations, bool clear_corner,
                        cudaStream_t s);
bool corner_uses_tile_packed_schur(int corner_blocks);   // true when launch_corner_solve will pick the register-tiled kernel
size_t corner_schur_floats(int corner_blocks);            // allocation size of `schur` valid for either layout

}  // namespace nnrt_b200
