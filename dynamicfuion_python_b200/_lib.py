"""ctypes binding of libnnrt_b200.so (include/nnrt_b200.h). There is no CPU fallback: if the CUDA library is missing
or fails to load, importing the product path raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libnnrt_b200.so")

P, I32, I64, F32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float

_SIGNATURES = {
    "nnrt_b200_version": (C.c_int, []),
    "nnrt_b200_last_error": (C.c_char_p, []),
    "nnrt_b200_compute_anchors_euclidean": (C.c_int, [P, I64, P, I32, P, F32, I32, I32, P, P, P]),
    "nnrt_b200_compute_anchors_shortest_path": (C.c_int, [P, I64, P, I32, P, F32, P, I32, I32, P, P, P]),
    "nnrt_b200_warp_points_normals": (C.c_int, [P, P, I64, P, P, P, I32, P, P, I32, I32, P, P, P, P]),
    "nnrt_b200_warp_points_normals_dual_quaternion": (C.c_int, [P, P, I64, P, P, P, I32, P, P, I32, I32, P, P, P, P]),
    "nnrt_b200_face_vertices_ndc": (C.c_int, [P, I64, P, I64, P, I32, I32, F32, F32, P, P, P]),
    "nnrt_b200_rasterize_ndc_triangles": (C.c_int, [P, P, I64, I32, I32, F32, I32, I32, I32, I32, I32, I32, P, P, P, P, P]),
    "nnrt_b200_interpolate_vertex_attributes": (C.c_int, [P, P, I64, I32, P, I32, P, P]),
    "nnrt_b200_unproject_depth": (C.c_int, [P, I32, I32, I32, P, P, F32, F32, P, P, P]),
    "nnrt_b200_point_to_plane_distances": (C.c_int, [P, P, P, I64, P, P]),
    "nnrt_b200_warped_surface_jacobians": (C.c_int, [P, P, I64, P, P, P, P, I32, P, P, P]),
    "nnrt_b200_rasterized_surface_jacobians": (C.c_int, [P, P, P, P, P, I32, I32, I32, P, I32, P, P, P]),
    "nnrt_b200_selfcheck_shared_divisor_quotients": (C.c_int, [P, P, I64, P, P, P]),
    "nnrt_b200_associate_faces_with_anchors": (C.c_int, [P, I64, P, I32, P, P, P]),
    "nnrt_b200_solve_block_sparse_arrowhead_cholesky": (C.c_int, [P, I32, I32, P, P, I64, P, P, I64, P, P, P]),
    "nnrt_b200_arrowhead_schur_complement": (C.c_int, [P, I32, I32, P, P, I64, P, P, I64, P, P]),
    "nnrt_b200_solve_block_diagonal_cholesky": (C.c_int, [P, P, I32, I32, P, P]),
    "nnrt_b200_precondition_diagonal_blocks": (C.c_int, [P, I32, I32, F32, P]),
    "nnrt_b200_rodrigues": (C.c_int, [P, I32, P, P]),
    "nnrt_b200_matmul3d_3x3": (C.c_int, [P, P, I32, P, P]),
    "nnrt_b200_median_grid_subsample_3d_points": (C.c_int, [P, I64, F32, F32, P, P, P]),
    "nnrt_b200_mean_grid_downsample_3d_points": (C.c_int, [P, I64, F32, F32, P, P, P]),
    "nnrt_b200_closest_to_mean_grid_subsample_3d_points": (C.c_int, [P, I64, F32, F32, P, P, P]),
    "nnrt_b200_find_k_nearest_to_points": (C.c_int, [P, I32, P, I64, I32, I32, P, P, P]),
    "nnrt_b200_node_coverage_weights": (C.c_int, [P, I32, F32, P, P]),
    "nnrt_b200_build_regularization_layers": (C.c_int, [P, I32, P, I32, I32, P, P, P, P, P, P]),
    "nnrt_b200_compute_triangle_normals": (C.c_int, [P, I64, P, I64, I32, P, P]),
    "nnrt_b200_compute_vertex_normals": (C.c_int, [P, I64, P, I64, I32, P, P]),
    "nnrt_b200_compute_vertex_normals_static_scratch": (C.c_int, [P, I64, P, I64, I32, P, P]),
    "nnrt_b200_normalize_vectors_3d": (C.c_int, [P, I64, P]),
    "nnrt_b200_compute_ordered_point_cloud_normals": (C.c_int, [P, I32, I32, P, P]),
    "nnrt_b200_shade_edges_flat": (C.c_int, [P, P, I32, I32, I32, F32, P, P, P]),
    "nnrt_b200_shade_vertex_colors": (C.c_int, [P, P, I32, I32, I32, P, P, P]),
    "nnrt_b200_integrate_rigid": (C.c_int, [P, I64, P, P, P, I32, P, I32, F32, F32, P, I32, P, I32, I32, I32, I32, I32, P, P, P, F32, F32, P]),
    "nnrt_b200_integrate_non_rigid": (C.c_int, [P, I64, P, P, P, I32, P, I32, F32, F32, P, I32, P, I32, P, I32, I32, I32, I32, P, P, P, P, P, P, I32, P,
                                                F32, I32, I32, F32, F32, P, P]),
    "nnrt_b200_integrate_non_rigid_full_scan": (C.c_int, [P, I64, P, P, P, I32, P, I32, F32, F32, P, I32, P, I32, P, I32, I32, I32, I32, P, P, P, P, P, P, I32, P,
                                                F32, I32, I32, F32, F32, P, P]),
    "nnrt_b200_bounding_boxes_of_warped_blocks": (C.c_int, [P, I64, P, P, P, I32, F32, I32, I32, F32, I32, P, P, P]),
    "nnrt_b200_boxes_intersecting_surface_mask": (C.c_int, [P, I64, P, P, I32, I32, I32, F32, F32, I32, F32, P, P]),
    "nnrt_b200_pixel_vertex_anchor_jacobians": (C.c_int, [P, P, P, P, P, P, P, P, I32, P, P, P, I64, I32, I32, I32, I32, F32, P, P, P, P]),
    "nnrt_b200_arap_term": (C.c_int, [P, I64, P, P, P, P, P, P, I32, F32, I32, F32, P, P, P, P, P, P, P]),
    "nnrt_b200_matmul_block_sparse_row_wise_padded": (C.c_int, [P, I32, P, P, I32, I32, P, P]),
    "nnrt_b200_matmul_block_sparse": (C.c_int, [P, P, I32, I32, I32, P, P, I32, I32, I32, I32, P, P, P]),
    "nnrt_b200_block_sparse_and_vector_product": (C.c_int, [P, P, I32, I32, I32, I32, I32, I32, P, P, P]),
    "nnrt_b200_diagonal_block_sparse_and_vector_product": (C.c_int, [P, I32, I32, P, P, P]),
    "nnrt_b200_invert_positive_semidefinite_blocks": (C.c_int, [P, I32, I32, P, P, P]),
    "nnrt_b200_invert_triangular_blocks": (C.c_int, [P, I32, I32, I32, P, P]),
    "nnrt_b200_transpose_blocks_in_place": (C.c_int, [P, I64, I32, P]),
    "nnrt_b200_dense_matrix_blocks": (C.c_int, [P, I64, I64, P, P, I64, I32, I32, I32, I32, I32, P]),
    "nnrt_b200_voxel_block_neighbors": (C.c_int, [P, I64, P, P]),
    "nnrt_b200_extract_triangle_mesh_prepare": (C.c_int, [I64, P, P, P, I32, I32, F32, P, P, P, P]),
    "nnrt_b200_extract_triangle_mesh_emit": (C.c_int, [P, I64, P, P, P, I32, F32, P, P, P, P, P, P, P]),
    "nnrt_b200_add_scaled": (C.c_int, [P, P, F32, I64, P, P]),
    "nnrt_b200_fill_identity_rotations": (C.c_int, [P, I64, P]),
    "nnrt_b200_gather_rows": (C.c_int, [P, I64, I32, P, I64, P, P]),
    "nnrt_b200_invert_permutation": (C.c_int, [P, I64, P, P]),
    "nnrt_b200_remap_indices": (C.c_int, [P, I64, P, I64, P, P]),
    "nnrt_b200_fitter_default_options": (None, [P]),
    "nnrt_b200_fitter_create": (C.c_int, [P, I64, I64, I32, I64, I32, I32, I32, P]),
    "nnrt_b200_fitter_destroy": (None, [P]),
    "nnrt_b200_fitter_workspace_bytes": (I64, [P]),
    "nnrt_b200_fitter_set_model": (C.c_int, [P, P, P, P, P, P, I32, P, P, I32, P, P, P, P]),
    "nnrt_b200_fitter_set_iteration_modes": (C.c_int, [P, P, I32]),
    "nnrt_b200_fitter_set_mesh_size": (C.c_int, [P, I64, I64]),
    "nnrt_b200_fitter_set_pixel_rows": (C.c_int, [P, I32, I32]),
    "nnrt_b200_fitter_set_pixel_stripes": (C.c_int, [P, I32, I32, I32]),
    "nnrt_b200_fitter_peer_exchange": (C.c_int, [P, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), P]),
    "nnrt_b200_fitter_enable_peer_reduce": (C.c_int, [P, I32, I32, C.POINTER(C.c_void_p)]),
    "nnrt_b200_fitter_enable_peer_reduce_ipc": (C.c_int, [P, I32, I32, P]),
    "nnrt_b200_fitter_disable_peer_reduce": (C.c_int, [P]),
    "nnrt_b200_fitter_accumulators": (P, [P, P]),
    "nnrt_b200_fitter_iteration_data_term": (C.c_int, [P, P, P, P, P, P]),
    "nnrt_b200_fitter_iteration_solve": (C.c_int, [P, P, P, P, P, I32, P]),
    "nnrt_b200_fitter_set_graph": (C.c_int, [P, P, P, I32, P, P, I32, P, P]),
    "nnrt_b200_fitter_set_mesh": (C.c_int, [P, P, P, P, P, P, P]),
    "nnrt_b200_fitter_set_reference_points": (C.c_int, [P, P, P, P]),
    "nnrt_b200_fitter_set_reference_depth": (C.c_int, [P, P, I32, P, F32, P, P]),
    "nnrt_b200_fitter_fit": (C.c_int, [P, P, P, P, P, I32, P]),
    "nnrt_b200_fitter_status": (C.c_int, [P, P]),
    "nnrt_b200_fitter_get_diagnostics": (C.c_int, [P, P, P, P, P, P, P, P]),
    "nnrt_b200_fitter_launches_per_iteration": (I32, [P]),
    "nnrt_b200_fitter_profile_iteration": (C.c_int, [P, P, P, P, P, P, P]),
}


class FitterOptions(C.Structure):
    _fields_ = [("max_iteration_count", I32), ("use_perspective_correction", I32), ("max_depth", F32), ("use_tukey_penalty", I32),
                ("tukey_penalty_cutoff", F32), ("preconditioning_dampening_factor", F32), ("arap_term_weight", F32),
                ("use_huber_penalty", I32), ("huber_penalty_constant", F32)]


_lib = None


def exported_symbols():
    return sorted(_SIGNATURES)


def lib():
    """Loads the CUDA library; raises if it is missing (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not found: build it with `python -m dynamicfuion_python_b200.build` "
                               "(or __graft_entry__.build()); this package has no CPU fallback")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc):
    """LogError semantics of the reference: a failing call raises RuntimeError with the library's message."""
    if rc != 0:
        msg = lib().nnrt_b200_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"nnrt_b200 error {rc}: {msg}")
