/* nnrt_b200 — C ABI of the sm_100a dense-depth Gauss-Newton path.
 *
 * Drop-in boundary for the hot path of henry123-boy/Dynamicfuion_python (`nnrt` pybind11 module,
 * cpp/pybind/nnrt_pybind.cpp:23-40). Each entry point replaces the kernel(s) behind one reference
 * host function (cited per function, paths relative to the reference's cpp/). All pointers are
 * DEVICE pointers unless the name ends in `_host`; all tensors are contiguous row-major; `stream`
 * is a cudaStream_t passed as void* (NULL = default stream). Every function returns 0 on success,
 * otherwise an nnrt_b200_status; nnrt_b200_last_error() returns a message for the calling thread.
 * No function allocates per call except where a workspace argument is documented as optional.
 */
#ifndef NNRT_B200_H
#define NNRT_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

typedef enum {
	NNRT_B200_OK = 0, NNRT_B200_INVALID_ARGUMENT = 1, NNRT_B200_CUDA_ERROR = 2, NNRT_B200_QUEUE_OVERFLOW = 3,
	NNRT_B200_NOT_POSITIVE_DEFINITE = 4, NNRT_B200_UNSUPPORTED = 5
} nnrt_b200_status;

const char* nnrt_b200_last_error(void);
int nnrt_b200_version(void);

/* geometry::functional::ComputeAnchorsAndWeights_Euclidean_{Fixed,Variable}NodeWeight
 * (geometry/functional/WarpAnchorComputation.cpp:74-148; kernel/WarpAnchorComputationImpl.h:42-139).
 * node_weights == NULL selects the fixed-coverage variant. anchors [V,A] int32, weights [V,A] f32. */
int nnrt_b200_compute_anchors_euclidean(const float* points, int64_t point_count, const float* nodes, int32_t node_count,
                                        const float* node_weights, float node_coverage, int32_t anchor_count,
                                        int32_t minimum_valid_anchor_count, int32_t* anchors, float* weights, void* stream);

/* geometry::functional::ComputeAnchorsAndWeights_ShortestPath_{Fixed,Variable}NodeWeight
 * (geometry/functional/WarpAnchorComputation.cpp:158-214; kernel/WarpAnchorComputationImpl.h:146-240;
 * core/kernel/KnnUtilities_PriorityQueue.h:102-160). edges [N, graph_degree] int32 adjacency (-1 = no neighbour). Anchors are the
 * first anchor_count nodes settled by Dijkstra from the point's nearest node (re-seeded per connected component), weights come
 * from the path lengths. The 40-entry queue and its std::push_heap / std::pop_heap tie order are the reference's (CPU path).
 * node_weights == NULL selects the fixed-coverage variant. Needs node_count >= anchor_count. */
int nnrt_b200_compute_anchors_shortest_path(const float* points, int64_t point_count, const float* nodes, int32_t node_count,
                                            const float* node_weights, float node_coverage, const int32_t* edges, int32_t graph_degree,
                                            int32_t anchor_count, int32_t* anchors, float* weights, void* stream);

/* geometry::functional::WarpTriangleMeshUsingSuppliedAnchors / WarpPointCloud (geometry/functional/Warping.cpp:169-263;
 * kernel/Warp3dPointsAndNormalsImpl.h:334-438). normals / warped_normals may be NULL (points only).
 * extrinsics_host: 4x4 row-major f64 on the host (as in the reference API). */
int nnrt_b200_warp_points_normals(const float* points, const float* normals, int64_t point_count, const float* nodes,
                                  const float* node_rotations, const float* node_translations, int32_t node_count,
                                  const int32_t* anchors, const float* anchor_weights, int32_t anchor_count,
                                  int32_t minimum_valid_anchor_count, const double* extrinsics_host, float* warped_points,
                                  float* warped_normals, void* stream);

/* Dual-quaternion blended warp with the arguments of nnrt_b200_warp_points_normals. BASELINE.json names dual-quaternion warping next
 * to linear blending; the reference has no implementation (geometry/TransformationMode.h only selects how rotations are stored), so
 * this follows Kavan et al. 2008: per anchor the rigid motion x -> R (x - g) + g + t as a unit dual quaternion, weighted sum with
 * signs aligned to the first valid anchor, normalisation, blended rigid motion applied to the point (its rotation to the normal). */
int nnrt_b200_warp_points_normals_dual_quaternion(const float* points, const float* normals, int64_t point_count, const float* nodes,
                                                  const float* node_rotations, const float* node_translations, int32_t node_count,
                                                  const int32_t* anchors, const float* anchor_weights, int32_t anchor_count,
                                                  int32_t minimum_valid_anchor_count, const double* extrinsics_host,
                                                  float* warped_points, float* warped_normals, void* stream);

/* rendering::functional::GetMeshNdcFaceVerticesAndClipMask (rendering/functional/ExtractFaceVertices.cpp:56-85;
 * kernel/ExtractClippedFaceVerticesImpl.h:33-194). triangles int64 [F,3]; out face_vertices_ndc [F,3,3] f32
 * (zeros for masked faces), clip_mask [F] u8 (1 = keep). */
int nnrt_b200_face_vertices_ndc(const float* vertex_positions, int64_t vertex_count, const int64_t* triangles, int64_t face_count,
                                const double* intrinsics_host, int32_t image_height, int32_t image_width, float near_clip,
                                float far_clip, float* face_vertices_ndc, uint8_t* clip_mask, void* stream);

/* rendering::RasterizeNdcTriangles (rendering/RasterizeNdcTriangles.cpp:32-128; kernel/RasterizeNdcTrianglesImpl.h,
 * kernel/RayFaceIntersection.h). clip_mask may be NULL. bin_size / max_faces_per_bin are accepted for API
 * compatibility and validated like the reference (bins-per-side < 22 for an explicit bin_size) but do not affect
 * results: faces are scattered to pixels with 64-bit atomic-min keys, there are no bins to overflow.
 * Outputs: pixel_face_indices [H,W,K] int64, pixel_depths [H,W,K], pixel_barycentric [H,W,K,3], pixel_face_distances [H,W,K]
 * (all -1 where empty). */
int nnrt_b200_rasterize_ndc_triangles(const float* face_vertices_ndc, const uint8_t* clip_mask, int64_t face_count,
                                      int32_t image_height, int32_t image_width, float blur_radius_pixels, int32_t faces_per_pixel,
                                      int32_t bin_size, int32_t max_faces_per_bin, int32_t perspective_correct,
                                      int32_t clip_barycentric, int32_t cull_back_faces, int64_t* pixel_face_indices,
                                      float* pixel_depths, float* pixel_barycentric, float* pixel_face_distances, void* stream);

/* rendering::functional::InterpolateVertexAttributes (rendering/functional/kernel/InterpolateFaceAttributesImpl.h:31-71). */
int nnrt_b200_interpolate_vertex_attributes(const int64_t* pixel_face_indices, const float* pixel_barycentric, int64_t pixel_count,
                                            int32_t faces_per_pixel, const float* face_attributes, int32_t channels,
                                            float* pixel_attributes, void* stream);

/* geometry::functional::UnprojectDepthImageWithoutFiltering (geometry/functional/kernel/PerspectiveProjectionImpl.h:62-150).
 * depth_dtype: 0 = float32, 1 = uint16. points [H*W,3] (zeros where invalid), mask [H*W] u8. */
int nnrt_b200_unproject_depth(const void* depth, int32_t depth_dtype, int32_t rows, int32_t cols, const double* intrinsics_host,
                              const double* extrinsics_host, float depth_scale, float depth_max, float* points, uint8_t* mask,
                              void* stream);

/* geometry::functional::ComputePointToPlaneDistances (geometry/functional/kernel/PointToPlaneDistancesImpl.h:25-52). */
int nnrt_b200_point_to_plane_distances(const float* normals1, const float* vertices1, const float* vertices2, int64_t count,
                                       float* distances, void* stream);

/* alignment::functional::WarpedSurfaceJacobians (alignment/functional/kernel/WarpedSurfaceJacobiansImpl.h:33-158).
 * position_jacobians [V,A,4], normal_jacobians [V,A,3]. */
int nnrt_b200_warped_surface_jacobians(const float* vertex_positions, const float* vertex_normals, int64_t vertex_count,
                                       const float* nodes, const float* node_rotations, const int32_t* anchors,
                                       const float* anchor_weights, int32_t anchor_count, float* position_jacobians,
                                       float* normal_jacobians, void* stream);

/* alignment::functional::RasterizedSurfaceJacobians (alignment/functional/kernel/RasterizedSurfaceJacobiansImpl.h:43-201,
 * BarycentricCoordinateJacobians.h). Outputs [H,W,3,9] and [H,W,3,10] f32 (zeros where uncovered). */
int nnrt_b200_rasterized_surface_jacobians(const float* warped_positions, const float* warped_normals, const int64_t* triangles,
                                           const int64_t* pixel_face_indices, const float* pixel_barycentric, int32_t image_height,
                                           int32_t image_width, int32_t faces_per_pixel, const double* ndc_intrinsics_host,
                                           int32_t perspective_correct, float* rendered_vertex_jacobians,
                                           float* rendered_normal_jacobians, void* stream);

/* Self-check of the division the Jacobian chain uses (csrc/common.cuh::SharedDivisor: one reciprocal per denominator, a corrected
 * multiply per numerator). The reference divides with `/` throughout BarycentricCoordinateJacobians.h:43-419; this entry returns, for
 * `count` (numerator, divisor) pairs on the device, the quotient of the shared-divisor form and of the IEEE division, which the parity
 * test compares bit for bit. */
int nnrt_b200_selfcheck_shared_divisor_quotients(const float* numerators, const float* divisors, int64_t count, float* shared_quotients,
                                                 float* ieee_quotients, void* stream);

/* alignment::functional::AssociateFacesWithAnchors (alignment/functional/kernel/AssociateFacesWithAnchorsImpl.h:33-107).
 * Compact replacement of the 1152 B/face Blob: face_anchor_entries [F, 3*A, 2] int32 = (node index, packed slots:
 * byte k = anchor slot of face vertex k or 0xFF), in the reference's first-appearance order; face_anchor_counts [F] int32. */
int nnrt_b200_associate_faces_with_anchors(const int64_t* triangles, int64_t face_count, const int32_t* vertex_anchors,
                                           int32_t anchor_count, int32_t* face_anchor_entries, int32_t* face_anchor_counts,
                                           void* stream);

/* core::linalg::SolveBlockSparseArrowheadCholesky (core/linalg/SolveBlockSparseArrowheadCholesky.cpp:31-94) on a
 * BlockSparseArrowheadMatrix (core/linalg/BlockSparseArrowheadMatrix.h:46-107): diagonal_blocks [N,6,6],
 * upper_wing_blocks [Wc,6,6] + coordinates [Wc,2] (global block row < arrow_base, col >= arrow_base),
 * corner_upper_blocks [Cu,6,6] + coordinates (may be NULL / 0). b, x: [6N]. */
int nnrt_b200_solve_block_sparse_arrowhead_cholesky(const float* diagonal_blocks, int32_t diagonal_block_count, int32_t arrow_base,
                                                    const float* upper_wing_blocks, const int32_t* upper_wing_coordinates,
                                                    int64_t upper_wing_count, const float* corner_upper_blocks,
                                                    const int32_t* corner_upper_coordinates, int64_t corner_upper_count,
                                                    const float* b, float* x, void* stream);
/* core::linalg::ComputeSchurComplementOfArrowStem_Dense (core/linalg/SchurComplement.cpp:32-77): out [(6Nc),(6Nc)]. */
int nnrt_b200_arrowhead_schur_complement(const float* diagonal_blocks, int32_t diagonal_block_count, int32_t arrow_base,
                                         const float* upper_wing_blocks, const int32_t* upper_wing_coordinates,
                                         int64_t upper_wing_count, const float* corner_upper_blocks,
                                         const int32_t* corner_upper_coordinates, int64_t corner_upper_count, float* schur,
                                         void* stream);
/* core::linalg::SolveBlockDiagonalCholesky (core/linalg/SolveBlockDiagonalGeneric.h:27-90), block_size 3 or 6. */
int nnrt_b200_solve_block_diagonal_cholesky(const float* blocks, const float* b, int32_t block_count, int32_t block_size, float* x,
                                            void* stream);
/* core::linalg::PreconditionDiagonalBlocks (core/linalg/PreconditionDiagonalBlocksImpl.h:29-51), in place. */
int nnrt_b200_precondition_diagonal_blocks(float* blocks, int32_t block_count, int32_t block_size, float lambda, void* stream);
/* core::linalg::AxisAngleVectorsToMatricesRodrigues (core/linalg/RodriguesImpl.h:66-100). */
int nnrt_b200_rodrigues(const float* axis_angles, int32_t count, float* matrices, void* stream);
/* core::linalg::Matmul3D for batched 3x3 (core/linalg/Matmul3D_CUDA.cpp), c = a * b. */
int nnrt_b200_matmul3d_3x3(const float* a, const float* b, int32_t count, float* c, void* stream);

/* ---- nnrt::alignment::DeformableMeshToImageFitter (alignment/DeformableMeshToImageFitter.h:30-126, .cpp:56-329) ---- */
typedef struct nnrt_b200_fitter nnrt_b200_fitter;

typedef struct {
	int32_t max_iteration_count;          /* default 100 */
	int32_t use_perspective_correction;   /* default 1 */
	float max_depth;                      /* default 10 */
	int32_t use_tukey_penalty;            /* default 0 */
	float tukey_penalty_cutoff;           /* default 0.01 */
	float preconditioning_dampening_factor; /* in [0,1], default 0 (LM lambda) */
	float arap_term_weight;               /* default 200 */
	int32_t use_huber_penalty;            /* default 0 */
	float huber_penalty_constant;         /* default 1e-4 */
} nnrt_b200_fitter_options;

void nnrt_b200_fitter_default_options(nnrt_b200_fitter_options* options);
/* Creates a fitter with a workspace sized once for (V, F, N, E, A, H, W); the iteration loop allocates nothing. */
int nnrt_b200_fitter_create(const nnrt_b200_fitter_options* options, int64_t vertex_count, int64_t face_count, int32_t node_count,
                            int64_t edge_count, int32_t anchor_count, int32_t image_height, int32_t image_width,
                            nnrt_b200_fitter** out_fitter);
void nnrt_b200_fitter_destroy(nnrt_b200_fitter* fitter);
int64_t nnrt_b200_fitter_workspace_bytes(const nnrt_b200_fitter* fitter);

/* Once-per-fit inputs (device pointers; copied/re-laid-out into the workspace on `stream`):
 * canonical mesh, graph in virtual node order (edges [E,2] int32 with i < arrow_base <= j; E may be 0),
 * precomputed anchors/weights [V,A] (PrecomputeAnchorsAndWeights, geometry/HierarchicalGraphWarpField.cpp:234-259).
 * edge_layer_indices [>=max(E,N)] int8 + layer_decimation_radii select FIXED_NODE_COVERAGE edge weights; otherwise
 * node_coverage_weights [N] (MINIMAL_K_NEIGHBOR_NODE_DISTANCE). */
int nnrt_b200_fitter_set_model(nnrt_b200_fitter* fitter, const float* vertex_positions, const float* vertex_normals,
                               const int64_t* triangles, const float* nodes, const int32_t* edges, int32_t arrow_base,
                               const int8_t* edge_layer_indices, const float* layer_decimation_radii, int32_t layer_count,
                               const float* node_coverage_weights, const int32_t* anchors, const float* anchor_weights,
                               void* stream);
/* The two halves of set_model, for callers that track a sequence: the warp field's graph (nodes, edges, edge weights) is
 * fixed while the canonical mesh changes with every frame (HierarchicalGraphWarpField owns the edges,
 * geometry/HierarchicalGraphWarpField.h:40-78; FitToImage receives the same field each frame). set_graph validates the
 * topology on the host (one stream synchronisation) and keeps the captured iteration graph when the arrowhead shape is
 * unchanged; set_mesh is fully asynchronous. set_model == set_mesh + set_graph. */
int nnrt_b200_fitter_set_graph(nnrt_b200_fitter* fitter, const float* nodes, const int32_t* edges, int32_t arrow_base,
                               const int8_t* edge_layer_indices, const float* layer_decimation_radii, int32_t layer_count,
                               const float* node_coverage_weights, void* stream);
int nnrt_b200_fitter_set_mesh(nnrt_b200_fitter* fitter, const float* vertex_positions, const float* vertex_normals,
                              const int64_t* triangles, const int32_t* anchors, const float* anchor_weights, void* stream);
/* The workspace is sized by fitter_create's vertex / face counts; a canonical mesh that changes from frame to frame (re-extracted from
 * the TSDF volume) may have any counts up to those: set them before set_mesh. Changing them drops the captured iteration graph (launch
 * sizes are part of it) but keeps the allocation. No reference counterpart (the reference allocates per call). */
int nnrt_b200_fitter_set_mesh_size(nnrt_b200_fitter* fitter, int64_t vertex_count, int64_t face_count);
/* IterationMode sequence (alignment/IterationMode.h:24-28: 0 ALL, 1 TRANSLATION_ONLY, 2 ROTATION_ONLY), host array, cycled by
 * iteration % count (DeformableMeshToImageFitter.cpp:112). Default {ALL}. Non-ALL modes solve the 3 x 3 block of the active half
 * (.cpp:198, 262-272) and are defined for edge-free warp fields only, as in the reference (fit returns an error otherwise). */
int nnrt_b200_fitter_set_iteration_modes(nnrt_b200_fitter* fitter, const int32_t* modes_host, int32_t mode_count);
/* Pixel-sharded single frame (one exchange per iteration): every GPU holds the whole mesh and graph, rasterises and assembles
 * only image rows [row_begin, row_end) (multiples of 8 rows, the assembly tile), and the per-node J^T J / J^T r partial sums are
 * summed across GPUs by the caller between iteration_data_term and iteration_solve; the solve is replicated. The accumulator
 * buffer is node_count x 32 entries (21 upper-triangular + 6 gradient + padding) of two 64-bit words each: a 103-bit FIXED-POINT
 * number (value = int64 word0 * 2^-20 + uint64 word1 * 2^-60) that the assembly kernel adds to with integer atomics. Integer sums
 * are order-independent, so an all-reduce of the words as int64 (ncclAllReduce, ncclInt64, ncclSum) gives every rank bit-identical
 * sums, equal to the single-GPU sums: the replicated solves stay bit-identical without any broadcast. The reference has no
 * multi-GPU path; the sums being exchanged are those of ComputeDepthHessianApproximationBlocks_UnorderedNodePixels /
 * ComputeNegativeDepthGradient_... (alignment/kernel/DeformableMeshToImageFitterImpl.h:199-420). iteration_index selects the
 * IterationMode of the sequence. */
int nnrt_b200_fitter_set_pixel_rows(nnrt_b200_fitter* fitter, int32_t row_begin, int32_t row_end);
/* Balanced form of the same split: the rows (of the band above; the whole image by default) are cut into stripes of stripe_rows rows
 * (a multiple of 8, the assembly tile height) dealt round-robin to stripe_count GPUs; this fitter rasterises and assembles stripes
 * stripe_index, stripe_index + stripe_count, ... A surface that covers the middle of the frame then loads every GPU alike, where
 * contiguous bands leave the outer GPUs idle. (8, 0, 1) = every row. The sums are the same integers either way. */
int nnrt_b200_fitter_set_pixel_stripes(nnrt_b200_fitter* fitter, int32_t stripe_rows, int32_t stripe_index, int32_t stripe_count);
int64_t* nnrt_b200_fitter_accumulators(nnrt_b200_fitter* fitter, int64_t* word_count);
/* The same exchange without a library collective (no reference counterpart; the B200 form of the step): each fitter keeps its sums in
 * a buffer of its own that the peers map over NVLink (CUDA IPC across processes, plain pointers inside one process), and the kernel
 * that consumes the sums gathers and adds the ranks' words itself while its ARAP edge work is in flight. Call order, on every rank:
 *   fitter_peer_exchange    allocates / clears the buffer (synchronises the device); returns its device pointer, size and, if asked,
 *                           the 64-byte cudaIpcMemHandle_t to hand to the other processes;
 *   enable_peer_reduce_ipc  with all ranks' handles (world x 64 bytes, rank-major; the own entry is ignored), or enable_peer_reduce
 *                           with device pointers valid in this process (same-process use; the caller enables peer access);
 *   a host barrier, then fitter_fit (or the split entry points) on every rank, the same number of iterations everywhere.
 * Needs set_pixel_rows / set_pixel_stripes to give each rank its share, and a graph on the fused stem path (no stem row with more than
 * 8 edges, no edges between coarse layers). A rank that waits more than 4 s for a peer raises an error status instead of hanging.
 * Before fitter_peer_exchange (which clears the buffer) and before disable_peer_reduce / fitter_destroy, every rank's stream must be
 * idle and the ranks must have met at a host barrier: a peer may still be reading this rank's sums of the last iteration. Exchange
 * buffers are pooled for the life of the process, never freed, so a late importer never holds a mapping of freed memory. */
int nnrt_b200_fitter_peer_exchange(nnrt_b200_fitter* fitter, void** device_pointer, int64_t* byte_count, void* ipc_handle_out);
int nnrt_b200_fitter_enable_peer_reduce(nnrt_b200_fitter* fitter, int32_t rank, int32_t world_size, void* const* exchange_pointers);
int nnrt_b200_fitter_enable_peer_reduce_ipc(nnrt_b200_fitter* fitter, int32_t rank, int32_t world_size, const void* ipc_handles);
int nnrt_b200_fitter_disable_peer_reduce(nnrt_b200_fitter* fitter);
int nnrt_b200_fitter_iteration_data_term(nnrt_b200_fitter* fitter, float* node_rotations, float* node_translations,
                                         const double* intrinsics_host, const double* extrinsics_host, void* stream);
int nnrt_b200_fitter_iteration_solve(nnrt_b200_fitter* fitter, float* node_rotations, float* node_translations,
                                     const double* intrinsics_host, const double* extrinsics_host, int32_t iteration_index, void* stream);
/* Reference observation: reference_points [H*W,3] f32 + reference_mask [H*W] u8 (FitToImage overload .cpp:85-94),
 * or a depth image (overload .cpp:278-314: depth / depth_scale, unprojected with the camera intrinsics, optional user mask). */
int nnrt_b200_fitter_set_reference_points(nnrt_b200_fitter* fitter, const float* reference_points, const uint8_t* reference_mask,
                                          void* stream);
int nnrt_b200_fitter_set_reference_depth(nnrt_b200_fitter* fitter, const void* depth, int32_t depth_dtype, const uint8_t* user_mask,
                                         float depth_scale, const double* intrinsics_host, void* stream);
/* Runs `iteration_count` GN/LM iterations (<= 0: options.max_iteration_count) as one CUDA graph launch.
 * node_rotations [N,3,3] / node_translations [N,3] are updated in place. */
int nnrt_b200_fitter_fit(nnrt_b200_fitter* fitter, float* node_rotations, float* node_translations, const double* intrinsics_host,
                         const double* extrinsics_host, int32_t iteration_count, void* stream);
/* Diagnostics of the LAST iteration executed (copied out on `stream`): any pointer may be NULL.
 * hessian_diagonal [N,6,6] (data + ARAP + damping), negative_gradient [N,6], deltas [N,6],
 * pixel_face_indices [H*W] int64, residuals [H*W] f32, residual_mask [H*W] u8. */
int nnrt_b200_fitter_get_diagnostics(nnrt_b200_fitter* fitter, float* hessian_diagonal, float* negative_gradient, float* deltas,
                                     int64_t* pixel_face_indices, float* residuals, uint8_t* residual_mask, void* stream);
/* Waits for `stream`, then reports device-side conditions recorded by the last fit (queue overflow, non-PD block). */
int nnrt_b200_fitter_status(nnrt_b200_fitter* fitter, void* stream);
/* Number of kernel launches one iteration issues (for bench.py's gpu_launches). */
int32_t nnrt_b200_fitter_launches_per_iteration(const nnrt_b200_fitter* fitter);
/* ---- Warp-field construction (the once-per-fit setup in front of the loop) ----
 * The reference numbers grid cells by hash-map buffer index and fills a cell's point list in arrival order, so its bin order and
 * (for exact distance-sum ties) its choice are run-dependent; its tests compare sets. These entry points resolve every such rule
 * by ascending point index (the reference's single-thread CPU result) and are reproducible. */

/* geometry::functional::MedianGridSubsample3dPoints (geometry/functional/GeometrySampling.cpp:61-68;
 * kernel/GeometrySamplingGridBinning.h:30-52, kernel/GeometrySamplingMedian.h:36-285): for every occupied cell of a grid with
 * the given cell size, the index of the point with the smallest sum of distances to the cell's points. sample: int64, capacity
 * point_count; cells in order of their first point. grid_offset (in cells, the same on all axes) is 0 except in the second pass of
 * FastMedianRadiusSubsample3dPoints / FastMeanRadiusDownsample3dPoints (kernel/GeometrySamplingImpl.h:47-91), which merge the first pass's
 * result on a grid shifted by 0.5. Synchronises the stream once (the count is a host value, as in the reference). */
int nnrt_b200_median_grid_subsample_3d_points(const float* points, int64_t point_count, float grid_cell_size, float grid_offset,
                                              int64_t* sample, int64_t* sample_count_host, void* stream);

/* geometry::functional::MeanGridDownsample3dPoints (geometry/functional/GeometrySampling.cpp:27-37; kernel/GeometrySamplingMean.h):
 * mean of the points of every occupied grid cell, downsampled_points [<= point_count, 3] f32, cells in order of their first point;
 * members are added in ascending point order (the reference's float atomics are scheduling-dependent). */
int nnrt_b200_mean_grid_downsample_3d_points(const float* points, int64_t point_count, float grid_cell_size, float grid_offset,
                                             float* downsampled_points, int64_t* downsampled_count_host, void* stream);
/* geometry::functional::ClosestToGridMeanSubsample3dPoints (GeometrySampling.cpp:40-47; kernel/GeometrySamplingMean.h:189-236):
 * per occupied cell the index (int64) of the point closest to the cell's mean, the smallest index among equals. */
int nnrt_b200_closest_to_mean_grid_subsample_3d_points(const float* points, int64_t point_count, float grid_cell_size, float grid_offset,
                                                       int64_t* sample, int64_t* sample_count_host, void* stream);

/* core::KdTree::FindKNearestToPoints (core/KdTree.cpp:52-90; kernel/SearchKdTreeImpl.h:118-146, PLAIN
 * strategy), brute force: indexed points are visited in ascending order, which fixes the slot order the kd-tree leaves to its
 * traversal; sort_output != 0 orders each row by ascending distance. nearest_indices [Q,k] int32 (-1 where fewer than k points
 * exist), nearest_distances [Q,k] f32 (may be NULL). k <= 8. */
int nnrt_b200_find_k_nearest_to_points(const float* indexed_points, int32_t indexed_count, const float* query_points,
                                       int64_t query_count, int32_t k, int32_t sort_output, int32_t* nearest_indices,
                                       float* nearest_distances, void* stream);

/* geometry::WarpField::RecomputeNodeCoverageWeights (geometry/WarpField.cpp:249-262): squared distance to the nearest other node
 * (node_coverage itself for a single node). */
int nnrt_b200_node_coverage_weights(const float* nodes, int32_t node_count, float node_coverage, float* node_coverage_weights,
                                    void* stream);

/* geometry::HierarchicalGraphWarpField::RebuildRegularizationLayers (geometry/HierarchicalGraphWarpField.cpp:74-195;
 * kernel/HierarchicalGraphWarpFieldImpl.h:218-296 AdjacencyArrayToEdges). layer_decimation_radii_host [L]: entry i >= 1 is the
 * caller's compute_layer_decimation_radius(i, node_coverage) (entry 0 unused: the finest layer holds every node).
 * Outputs: node_indices [N] int64 (virtual -> original index, layers fine to coarse), layer_node_counts_host [L],
 * edges [<= N * max_vertex_degree, 2] int32 in virtual indices (layer pairs coarse to fine, sources descending, each source's
 * targets descending), edge_layer_indices [same count] int8 (layer of the targets), edge_count_host. max_vertex_degree <= 8.
 * One stream synchronisation per layer beyond the first. Fails like the reference when a layer cannot be decimated. */
int nnrt_b200_build_regularization_layers(const float* nodes, int32_t node_count, const float* layer_decimation_radii_host,
                                          int32_t layer_count, int32_t max_vertex_degree, int64_t* node_indices,
                                          int32_t* layer_node_counts_host, int32_t* edges, int8_t* edge_layer_indices,
                                          int64_t* edge_count_host, void* stream);

/* ---- Normals (nnrt.geometry.functional.compute_*_normals, pybind/geometry/functional/functional.cpp:112-114) ---- */

/* geometry::ComputeTriangleNormals (geometry/functional/NormalsOperations.cpp:36-46; kernel/NormalsOperationsImpl.h:38-93):
 * (v1 - v0) x (v2 - v0) per face, normalised if `normalized` (a NaN result becomes (0, 0, 1), as in the reference).
 * triangles [F,3] int64, triangle_normals [F,3] f32. */
int nnrt_b200_compute_triangle_normals(const float* vertex_positions, int64_t vertex_count, const int64_t* triangles,
                                       int64_t face_count, int32_t normalized, float* triangle_normals, void* stream);
/* geometry::ComputeVertexNormals (NormalsOperations.cpp:52-67; kernel/NormalsOperationsImpl.h:96-165): per vertex, the sum of the
 * given normals of the triangles referencing it, normalised if `normalized`. The reference adds with float atomics (order depends on
 * scheduling); here the addends are taken in ascending face order, which is its single-thread CPU result. */
int nnrt_b200_compute_vertex_normals(const int64_t* triangles, int64_t face_count, const float* triangle_normals, int64_t vertex_count,
                                     int32_t normalized, float* vertex_normals, void* stream);
/* Allocation-free variant of nnrt_b200_compute_vertex_normals. It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid allocation). Same arguments and result (bit-identical) as nnrt_b200_compute_vertex_normals; the CSR scratch is a file-scope
 * __device__ array (about 4.2 MB, held by every GPU that loads the library until exit) instead of a pool allocation, and the call
 * makes no cudaMalloc / cudaFree / allocating library call. Meshes beyond the limits below are rejected with
 * NNRT_B200_INVALID_ARGUMENT and a message in nnrt_b200_last_error(). All calls on a device share that array: calls that overlap on
 * the GPU (different streams without ordering) race; calls ordered on one stream, or by events, are safe. The first call on a device
 * looks up the array's address (cudaGetSymbolAddress); make that call outside stream capture. */
#define NNRT_B200_STATIC_SCRATCH_MAX_VERTICES (1 << 17)
#define NNRT_B200_STATIC_SCRATCH_MAX_FACES (1 << 18)
int nnrt_b200_compute_vertex_normals_static_scratch(const int64_t* triangles, int64_t face_count, const float* triangle_normals,
                                                    int64_t vertex_count, int32_t normalized, float* vertex_normals, void* stream);
/* geometry::NormalizeVectors3d (NormalsOperations.cpp:48-50), in place. */
int nnrt_b200_normalize_vectors_3d(float* vectors, int64_t count, void* stream);
/* geometry::ComputeOrderedPointCloudNormals (NormalsOperations.cpp:73-93; kernel/NormalsOperationsImpl.h:172-214): central
 * differences over the source image grid, zero on the border, flipped towards the camera (z <= 0). */
int nnrt_b200_compute_ordered_point_cloud_normals(const float* point_positions, int32_t image_height, int32_t image_width,
                                                  float* normals, void* stream);

/* ---- Shaders over rasteriser fragments (headless: 8-bit RGB device buffers) ---- */

/* rendering::FlatEdgeShader::ShadeMeshes (rendering/FlatEdgeShader.cpp:32-50; kernel/FlatEdgeShaderImpl.h:30-99): triangle edges
 * drawn from the nearest fragment's face distance, `pixel_line_width` in pixels, line_color_host = 3 floats in [0, 1].
 * pixels [H,W,3] u8. */
int nnrt_b200_shade_edges_flat(const int64_t* pixel_face_indices, const float* pixel_face_distances, int32_t image_height,
                               int32_t image_width, int32_t faces_per_pixel, float pixel_line_width, const float* line_color_host,
                               uint8_t* pixels, void* stream);
/* rendering::VertexColorShader::ShadeMeshes (rendering/VertexColorShader.cpp:29-59): face_vertex_colors [F,3,3] f32 in [0, 1]
 * (vertex colours gathered by triangle, meshes concatenated in rasterisation order); pixels [H,W,3] u8. */
int nnrt_b200_shade_vertex_colors(const int64_t* pixel_face_indices, const float* pixel_barycentric, int32_t image_height,
                                  int32_t image_width, int32_t faces_per_pixel, const float* face_vertex_colors, uint8_t* pixels,
                                  void* stream);

/* ---- Non-rigid TSDF integration (the step on the other side of the tracker) ----
 * geometry::NonRigidSurfaceVoxelBlockGrid::IntegrateNonRigid's voxel kernel (geometry/NonRigidSurfaceVoxelBlockGrid.cpp:32-74;
 * kernel/NonRigidSurfaceVoxelBlockGridImpl.h:50-229) over the caller's voxel-block arrays: block_indices [B] int64 (active buffer
 * indices), block_keys [capacity,3] int32, tsdf [capacity*R^3] f32 and voxel_color [capacity*R^3*3] f32 (may be NULL) updated in
 * place, weight [capacity*R^3] (weight_dtype 0 = f32, 1 = u16) read only, as in the reference. depth [H,W] (depth_dtype 0 = f32,
 * 1 = u16), color [Hc,Wc,3] or NULL (color_dtype 0 = f32 in [0,1], 1 = u8), depth_normals [H,W,3]; the warp field as nodes /
 * rotations / translations with node_coverage_weights (NULL = fixed node_coverage), anchor_count, minimum_valid_anchor_count.
 * Output cos_voxel_ray_to_normal [H,W] f32 (0 where no voxel projected; the highest voxel index wins a pixel). The k nearest nodes
 * are found by brute force in ascending node order (the reference's kd-tree yields the same set in its traversal order), over a
 * per-block shortlist of the nodes that can reach a slot of some voxel of the block (csrc/tsdf.cu). */
int nnrt_b200_integrate_non_rigid(const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf,
                                  const void* weight, int32_t weight_dtype, float* voxel_color, int32_t block_resolution,
                                  float voxel_size, float sdf_truncation_distance, const void* depth, int32_t depth_dtype,
                                  const void* color, int32_t color_dtype, const float* depth_normals, int32_t image_height,
                                  int32_t image_width, int32_t color_height, int32_t color_width,
                                  const double* depth_intrinsics_host, const double* color_intrinsics_host,
                                  const double* extrinsics_host, const float* nodes, const float* node_rotations,
                                  const float* node_translations, int32_t node_count, const float* node_coverage_weights,
                                  float node_coverage, int32_t anchor_count, int32_t minimum_valid_anchor_count, float depth_scale,
                                  float depth_max, float* cos_voxel_ray_to_normal, void* stream);
/* The same with every node offered to every voxel (no per-block shortlist): the yardstick the default entry point equals bit for bit. */
int nnrt_b200_integrate_non_rigid_full_scan(const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf,
                                  const void* weight, int32_t weight_dtype, float* voxel_color, int32_t block_resolution,
                                  float voxel_size, float sdf_truncation_distance, const void* depth, int32_t depth_dtype,
                                  const void* color, int32_t color_dtype, const float* depth_normals, int32_t image_height,
                                  int32_t image_width, int32_t color_height, int32_t color_width,
                                  const double* depth_intrinsics_host, const double* color_intrinsics_host,
                                  const double* extrinsics_host, const float* nodes, const float* node_rotations,
                                  const float* node_translations, int32_t node_count, const float* node_coverage_weights,
                                  float node_coverage, int32_t anchor_count, int32_t minimum_valid_anchor_count, float depth_scale,
                                  float depth_max, float* cos_voxel_ray_to_normal, void* stream);

/* geometry::NonRigidSurfaceVoxelBlockGrid::GetBoundingBoxesOfWarpedBlocks (geometry/NonRigidSurfaceVoxelBlockGrid.cpp:120-130;
 * kernel/NonRigidSurfaceVoxelBlockGridImpl.h:289-363): bounding_boxes [B,6] f32 (min xyz, max xyz) of the warped corners of each block
 * key [B,3] int32. Fixed node coverage, as the reference's kernel. Its quirks are kept (see csrc/tsdf.cu). */
int nnrt_b200_bounding_boxes_of_warped_blocks(const int32_t* block_keys, int64_t block_count, const float* nodes,
                                              const float* node_rotations, const float* node_translations, int32_t node_count,
                                              float node_coverage, int32_t anchor_count, int32_t minimum_valid_anchor_count,
                                              float voxel_size, int32_t block_resolution, const double* extrinsics_host,
                                              float* bounding_boxes, void* stream);
/* Rigid integration of a depth (+ colour) frame into the listed voxel blocks: what nnrt.geometry.VoxelBlockGrid.integrate re-exports
 * from Open3D (pybind/geometry/geometry.cpp:170-203; the fusion pipeline integrates its first frame with it, apps/fusion/pipeline.py).
 * Open3D v0.17 is third party and absent from the reference checkout: restated from its published kernel, parity unpinned. tsdf,
 * weight (float32 or uint16, raised by one per integrated voxel) and voxel_color are updated in place. */
int nnrt_b200_integrate_rigid(const int64_t* block_indices, int64_t block_count, const int32_t* block_keys, float* tsdf, void* weight,
                              int32_t weight_dtype, float* voxel_color, int32_t block_resolution, float voxel_size,
                              float sdf_truncation_distance, const void* depth, int32_t depth_dtype, const void* color, int32_t color_dtype,
                              int32_t image_height, int32_t image_width, int32_t color_height, int32_t color_width,
                              const double* depth_intrinsics_host, const double* color_intrinsics_host, const double* extrinsics_host,
                              float depth_scale, float depth_max, void* stream);
/* geometry::NonRigidSurfaceVoxelBlockGrid::GetAxisAlignedBoxesIntersectingSurfaceMask (NonRigidSurfaceVoxelBlockGrid.cpp:132-148;
 * kernel/NonRigidSurfaceVoxelBlockGridImpl.h:365-435; kernel/Segment.h): mask [B] u8 = the box is met by the truncation segment
 * (depth -+ truncation_distance along the pixel ray) of some pixel sampled with `stride`. Together with the call above this is the
 * device work of FindBlocksIntersectingTruncationRegion (NonRigidSurfaceVoxelBlockGrid.cpp:150-178). */
int nnrt_b200_boxes_intersecting_surface_mask(const float* boxes, int64_t box_count, const double* intrinsics_host, const void* depth,
                                              int32_t depth_dtype, int32_t image_height, int32_t image_width, float depth_scale,
                                              float depth_max, int32_t stride, float truncation_distance, uint8_t* mask, void* stream);

/* ---- Mesh extraction from TSDF voxel blocks (the canonical mesh the fitter consumes) ----
 * geometry::VoxelBlockGrid::ExtractTriangleMesh (geometry/VoxelBlockGrid.cpp:461-484) calls Open3D's
 * t::geometry::kernel::voxel_grid::ExtractTriangleMesh over Open3D's device hash map (Open3D v0.17.0-1fix6008, third party, not in
 * the reference checkout); both are restated from the published algorithm (csrc/mesh_extraction.cu). Parity unpinned by the reference.
 *
 * VoxelBlockGrid::BufferRadiusNeighbors (VoxelBlockGrid.cpp:466-468): the block hash. block_keys [B,3] int32 (block coordinates);
 * neighbor_indices [B,27] int32: index of the block at key + (dx, dy, dz), slot (dz + 1) * 9 + (dy + 1) * 3 + (dx + 1), -1 where absent
 * (an open-addressing table over the packed keys is built per call; a key listed twice resolves to its first block). */
int nnrt_b200_voxel_block_neighbors(const int32_t* block_keys, int64_t block_count, int32_t* neighbor_indices, void* stream);
/* Marching cubes, first half: cube configuration per voxel, one vertex per sign-change edge that lies on a complete cube (all eight
 * corners present with weight > weight_threshold), numbered by exclusive scans in (block, voxel, x / y / z edge) order.
 * tsdf [B,R,R,R] f32 (z, y, x major to minor as Open3D's blocks), weight the same shape (weight_dtype 0 = f32, 1 = uint16).
 * Caller-allocated: mesh_structure [B*R^3, 4] int32 (vertex id of the voxel's x / y / z edge or < 0, cube configuration),
 * triangle_offsets [B*R^3 + 1] int32. counts_host[0..1] = vertex and triangle counts (one stream synchronisation). */
int nnrt_b200_extract_triangle_mesh_prepare(int64_t block_count, const int32_t* neighbor_indices, const float* tsdf, const void* weight,
                                            int32_t weight_dtype, int32_t block_resolution, float weight_threshold,
                                            int32_t* mesh_structure, int32_t* triangle_offsets, int64_t* counts_host, void* stream);
/* Second half: vertices [Nv,3] = voxel_size * (global voxel coordinate + ratio along the edge), ratio = -tsdf_o / (tsdf_e - tsdf_o);
 * normals [Nv,3] = interpolated central-difference TSDF gradients / (norm + 1e-5); colors [Nv,3] = interpolated color / 255 when
 * color [B,R^3,3] f32 is given (else both NULL); triangles [Nt,3] int64. vertices / triangles may be NULL to skip that part. */
int nnrt_b200_extract_triangle_mesh_emit(const int32_t* block_keys, int64_t block_count, const int32_t* neighbor_indices,
                                         const float* tsdf, const float* color, int32_t block_resolution, float voxel_size,
                                         const int32_t* mesh_structure, const int32_t* triangle_offsets, float* vertices, float* normals,
                                         float* colors, int64_t* triangles, void* stream);

/* ---- Node store of the warp field (what the host classes do to node tensors besides the warp) ----
 * WarpField::GetWarpedNodes / TranslateNodes (geometry/WarpField.cpp:89-91,185-187; HierarchicalGraphWarpField.cpp:261-269): out = a + alpha * b. */
int nnrt_b200_add_scaled(const float* a, const float* b, float alpha, int64_t count, float* out, void* stream);
/* WarpField::ResetRotations (geometry/WarpField.cpp:151-159): [count,3,3] identities. */
int nnrt_b200_fill_identity_rotations(float* rotations, int64_t count, void* stream);
/* Virtual <-> original node order (HierarchicalGraphWarpField.cpp:197-226, tensor[indices] with an Int64 index tensor): rows of
 * `row_words` 4-byte words, out[i] = source[indices[i]]. */
int nnrt_b200_gather_rows(const void* source, int64_t source_rows, int32_t row_words, const int64_t* indices, int64_t count, void* out,
                          void* stream);
/* inverse[permutation[i]] = i (original -> virtual index map of HierarchicalGraphWarpField). */
int nnrt_b200_invert_permutation(const int64_t* permutation, int64_t count, int64_t* inverse, void* stream);
/* Anchor indices between the two node orders (HierarchicalGraphWarpField.cpp:228-256): out = map[index], -1 (no anchor) kept. */
int nnrt_b200_remap_indices(const int32_t* indices, int64_t count, const int64_t* map, int64_t map_size, int32_t* out, void* stream);

/* ---- Un-fused stages of the assembly chain (parity-test entry points; the fitter's iteration fuses them) ----
 * alignment::functional::PixelVertexAnchorJacobiansAndNodeAssociations (alignment/functional/PixelVertexAnchorJacobians.cpp:29-56;
 * kernel/PixelVertexAnchorJacobiansImpl.h:40-434). Inputs are the reference's intermediate tensors: rasterized vertex position /
 * normal Jacobians [PX,3,9] / [PX,3,10] (nnrt_b200_rasterized_surface_jacobians), warped vertex position / normal Jacobians
 * [V,A,4] / [V,A,3] (nnrt_b200_warped_surface_jacobians), point_map_vectors [PX,3] (rendered point - reference point),
 * rasterized_normals [PX,3], residual_mask [PX] u8, pixel_faces [PX,K] int64 (k = 0 used), triangles [F,3] int64 and the compact
 * face-anchor table of nnrt_b200_associate_faces_with_anchors. iteration_mode 0 ALL (S = 6 floats per entry: rotation, translation),
 * 1 TRANSLATION_ONLY, 2 ROTATION_ONLY (S = 3). Outputs: pixel_jacobians [PX,3A,S] f32 (zero-filled), pixel_node_jacobian_counts
 * [PX] int32, node_pixel_jacobian_counts [N] int32. The reference's node_pixel_jacobian_indices [N,4000] lists (capped, order
 * scheduling-dependent) have no counterpart: nothing downstream of this library consumes per-node pixel lists. */
int nnrt_b200_pixel_vertex_anchor_jacobians(
		const float* rasterized_vertex_position_jacobians, const float* rasterized_vertex_normal_jacobians,
		const float* warped_vertex_position_jacobians, const float* warped_vertex_normal_jacobians, const float* point_map_vectors,
		const float* rasterized_normals, const uint8_t* residual_mask, const int64_t* pixel_faces, int32_t faces_per_pixel,
		const int64_t* triangles, const int32_t* face_anchor_entries, const int32_t* face_anchor_counts, int64_t pixel_count,
		int32_t anchor_count, int32_t node_count, int32_t iteration_mode, int32_t use_tukey_penalty, float tukey_penalty_cutoff,
		float* pixel_jacobians, int32_t* pixel_node_jacobian_counts, int32_t* node_pixel_jacobian_counts, void* stream);

/* The ARAP term as the reference materialises it: ComputeEdgeResiduals (alignment/DeformableMeshToImageFitter.cpp:392-448,
 * kernel/DeformableMeshToImageFitterImpl.h:644-784; optional Huber), ComputeDenseArapEdgeJacobians (functional/ArapJacobian.cpp:25-72)
 * -> edge_jacobians [E,5], ComputeArapBlockSparseHessianApproximation (functional/ArapHessian.cpp:30-48) -> ARAP part of the diagonal
 * blocks [N,6,6], upper wing blocks [E,6,6] + coordinates [E,2] (block id = edge id), ComputeNegativeArapGradient (...Impl.h:464-642)
 * -> [6N]. edges [E,2] int32 in virtual node order; fixed coverage: edge_layer_indices [max(E,N)] int8 + layer_decimation_radii,
 * variable coverage: node_coverage_weights [N]. Any output pointer may be NULL. Per-node sums run in ascending edge order. */
int nnrt_b200_arap_term(const int32_t* edges, int64_t edge_count, const int8_t* edge_layer_indices, const float* layer_decimation_radii,
                        const float* node_coverage_weights, const float* nodes, const float* node_rotations,
                        const float* node_translations, int32_t node_count, float arap_term_weight, int32_t use_huber_penalty,
                        float huber_penalty_constant, float* edge_residuals, float* edge_jacobians, float* hessian_diagonal_blocks,
                        float* upper_wing_blocks, int32_t* upper_wing_block_coordinates, float* negative_gradient, void* stream);

/* ---- Generic block-sparse routines of core/linalg (block size <= 8), one per reference function; the arrowhead solve above fuses
 * them. Pinned by the reference's inline known-answer tests (cpp/tests/test_linalg_matmul_block_sparse.cpp,
 * cpp/tests/test_linalg_block_routines.cpp). ----
 * core::linalg::MatmulBlockSparseRowWisePadded (core/linalg/MatmulBlockSparseImpl.h:48-150): C_k = A[row(B_k)] B_k, zero where the
 * row has no A block; MatmulBlockSparseRowWise = the same restricted to rows < a_block_count. */
int nnrt_b200_matmul_block_sparse_row_wise_padded(const float* blocks_a, int32_t a_block_count, const float* blocks_b,
                                                  const int32_t* blocks_b_coordinates, int32_t b_block_count, int32_t block_size,
                                                  float* blocks_c, void* stream);
/* core::linalg::MatmulBlockSparse (MatmulBlockSparseImpl.h:171-420): op(A) op(B) of two block lists addressed through int16
 * breadboards [rows, columns] (block index or -1). out_blocks [out_rows*out_columns, s, s] row-major over output positions,
 * out_block_mask [out_rows*out_columns] = position received a product (the reference returns the masked blocks + coordinates). */
int nnrt_b200_matmul_block_sparse(const float* blocks_a, const int16_t* a_breadboard, int32_t a_rows, int32_t a_columns, int32_t transpose_a,
                                  const float* blocks_b, const int16_t* b_breadboard, int32_t b_rows, int32_t b_columns, int32_t transpose_b,
                                  int32_t block_size, float* out_blocks, uint8_t* out_block_mask, void* stream);
/* core::linalg::BlockSparseAndVectorProduct (MatmulBlockSparseImpl.h:441-600): c [output_length] = op(A) b. */
int nnrt_b200_block_sparse_and_vector_product(const float* blocks_a, const int32_t* blocks_a_coordinates, int32_t block_count,
                                              int32_t block_size, int32_t output_length, int32_t transpose_a, int32_t block_row_offset,
                                              int32_t block_column_offset, const float* vector_b, float* vector_c, void* stream);
/* core::linalg::DiagonalBlockSparseAndVectorProduct (MatmulBlockSparseImpl.h:620-700). */
int nnrt_b200_diagonal_block_sparse_and_vector_product(const float* blocks_d, int32_t block_count, int32_t block_size, const float* vector_b,
                                                       float* vector_c, void* stream);
/* core::linalg::InvertPositiveSemidefiniteBlocks (core/linalg/InvertBlocks.cpp:82-127; InvertBlocksCPU.cpp:80-112 potrf + potrs;
 * InvertBlocksCUDA.cu:96-139). failure_count_device [1] int32 = blocks that were not positive definite (left untouched). */
int nnrt_b200_invert_positive_semidefinite_blocks(const float* blocks, int32_t block_count, int32_t block_size, float* inverted_blocks,
                                                  int32_t* failure_count_device, void* stream);
/* core::linalg::InvertTriangularBlocks (core/linalg/InvertBlocks.cpp:33-80); upper = 0 lower, 1 upper triangular. */
int nnrt_b200_invert_triangular_blocks(const float* blocks, int32_t block_count, int32_t block_size, int32_t upper, float* inverted_blocks,
                                       void* stream);
/* core::linalg::TransposeBlocksInPlace (core/linalg/TransposeBlocks.cpp). */
int nnrt_b200_transpose_blocks_in_place(float* blocks, int64_t block_count, int32_t block_size, void* stream);
/* core::linalg::FillInDiagonalBlocks / GetDiagonalBlocks (core/linalg/DiagonalBlocksImpl.h; coordinates == NULL) and FillInSparseBlocks /
 * GetSparseBlocks (core/linalg/SparseBlocksImpl.h:30-140) on a dense row-major matrix; gather = 0 writes blocks into the matrix (with
 * `transpose` block k lands transposed at (column, row)), gather = 1 reads them out. */
int nnrt_b200_dense_matrix_blocks(float* matrix, int64_t matrix_rows, int64_t matrix_columns, float* blocks, const int32_t* coordinates,
                                  int64_t block_count, int32_t block_size, int32_t block_row_offset, int32_t block_column_offset,
                                  int32_t transpose, int32_t gather, void* stream);

/* Per-stage device time of one iteration measured with CUDA events outside graph capture (ms):
 * [warp_vertices, face_raster, band_raster, assemble, build_system, stem_eliminate, corner_solve_update]. */
int nnrt_b200_fitter_profile_iteration(nnrt_b200_fitter* fitter, float* node_rotations, float* node_translations,
                                       const double* intrinsics_host, const double* extrinsics_host, float* stage_ms7, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif
