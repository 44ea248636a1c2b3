"""nnrt.alignment.functional — un-fused stages of the data term, returning the reference's intermediate tensors (C++-only in the
reference: cpp/alignment/functional/*.h). The fitter itself runs the fused kernel and never materialises these."""
import torch

from .. import _tensor as T


def warped_surface_jacobians(canonical_mesh, warp_field, warp_anchors, warp_anchor_weights, use_virtual_ordering=True):
    """WarpedSurfaceJacobians (cpp/alignment/functional/WarpedSurfaceJacobians.cpp:28-47) -> ([V,A,4], [V,A,3])."""
    mesh = T.MeshView(canonical_mesh)
    if mesh.normals is None:
        raise RuntimeError("vertex_normals argument needs to be a tensor for a call to WarpedSurfaceJacobians")
    a, w = T.cuda(warp_anchors, torch.int32), T.cuda(warp_anchor_weights, torch.float32)
    v, k = a.shape
    jv, jn = T.empty((v, k, 4), torch.float32), T.empty((v, k, 3), torch.float32)
    T.call("nnrt_b200_warped_surface_jacobians", T.ptr(mesh.positions), T.ptr(mesh.normals), v, T.ptr(warp_field.nodes),
           T.ptr(warp_field.node_rotations), T.ptr(a), T.ptr(w), k, T.ptr(jv), T.ptr(jn), T.stream())
    return jv, jn


def rasterized_surface_jacobians(warped_mesh, pixel_faces, barycentric_coordinates, ndc_intrinsic_matrix, perspective_corrected_barycentric_coordinates):
    """RasterizedSurfaceJacobians (cpp/alignment/functional/RasterizedSurfaceJacobians.cpp:30-52) -> ([H,W,3,9], [H,W,3,10])."""
    mesh = T.MeshView(warped_mesh)
    idx, bary = T.cuda(pixel_faces, torch.int64), T.cuda(barycentric_coordinates, torch.float32)
    K = T.host_f64(ndc_intrinsic_matrix, (3, 3))
    h, w, k = idx.shape
    jv, jn = T.empty((h, w, 3, 9), torch.float32), T.empty((h, w, 3, 10), torch.float32)
    T.call("nnrt_b200_rasterized_surface_jacobians", T.ptr(mesh.positions), T.ptr(mesh.normals), T.ptr(mesh.triangles), T.ptr(idx), T.ptr(bary),
           h, w, k, K.data_ptr(), int(bool(perspective_corrected_barycentric_coordinates)), T.ptr(jv), T.ptr(jn), T.stream())
    return jv, jn


def associate_faces_with_anchors(face_vertex_indices, warp_vertex_anchors):
    """AssociateFacesWithAnchors (cpp/alignment/functional/AssociateFacesWithAnchors.cpp:29-37). Compact table instead of the
    reference's 1152 B/face Blob: entries [F, 3A, 2] int32 (node, packed per-vertex anchor slots; 0xFF = none), counts [F]."""
    tri, anchors = T.cuda(face_vertex_indices, torch.int64), T.cuda(warp_vertex_anchors, torch.int32)
    f, a = tri.shape[0], anchors.shape[1]
    entries, counts = T.empty((f, 3 * a, 2), torch.int32), T.empty((f,), torch.int32)
    T.call("nnrt_b200_associate_faces_with_anchors", T.ptr(tri), f, T.ptr(anchors), a, T.ptr(entries), T.ptr(counts), T.stream())
    return entries, counts


def pixel_vertex_anchor_jacobians_and_node_associations(rasterized_vertex_position_jacobians, rasterized_vertex_normal_jacobians,
                                                        warped_vertex_position_jacobians, warped_vertex_normal_jacobians, point_map_vectors,
                                                        rasterized_normals, residual_mask, pixel_faces, face_vertex_indices, face_node_anchors,
                                                        face_node_anchor_counts, node_count, tukey_penalty_cutoff, iteration_mode=0,
                                                        use_tukey_penalty=False):
    """PixelVertexAnchorJacobiansAndNodeAssociations (cpp/alignment/functional/PixelVertexAnchorJacobians.cpp:29-56) ->
    (pixel_jacobians [PX,3A,S], pixel_node_jacobian_counts [PX], node_pixel_jacobian_counts [N]); `face_node_anchors` / counts are
    the compact table of associate_faces_with_anchors (which needs `face_vertex_indices` to address the vertices). The reference's
    capped per-node pixel lists are not produced (the library sums per node without them)."""
    rv, rn = T.cuda(rasterized_vertex_position_jacobians, torch.float32), T.cuda(rasterized_vertex_normal_jacobians, torch.float32)
    wv, wn = T.cuda(warped_vertex_position_jacobians, torch.float32), T.cuda(warped_vertex_normal_jacobians, torch.float32)
    pm, nl = T.cuda(point_map_vectors, torch.float32).reshape(-1, 3), T.cuda(rasterized_normals, torch.float32).reshape(-1, 3)
    mask = T.cuda(T.to_torch(residual_mask).reshape(-1), torch.uint8)
    px = mask.shape[0]
    faces = T.cuda(pixel_faces, torch.int64).reshape(px, -1)
    tri, fa, fc = T.cuda(face_vertex_indices, torch.int64), T.cuda(face_node_anchors, torch.int32), T.cuda(face_node_anchor_counts, torch.int32)
    a, s = wv.shape[1], (6 if int(iteration_mode) == 0 else 3)
    jac, counts, node_counts = T.empty((px, 3 * a, s), torch.float32), T.empty((px,), torch.int32), T.empty((int(node_count),), torch.int32)
    T.call("nnrt_b200_pixel_vertex_anchor_jacobians", T.ptr(rv), T.ptr(rn), T.ptr(wv), T.ptr(wn), T.ptr(pm), T.ptr(nl), T.ptr(mask), T.ptr(faces),
           faces.shape[1], T.ptr(tri), T.ptr(fa), T.ptr(fc), px, a, int(node_count), int(iteration_mode), int(bool(use_tukey_penalty)),
           float(tukey_penalty_cutoff), T.ptr(jac), T.ptr(counts), T.ptr(node_counts), T.stream())
    return jac, counts, node_counts


def arap_term(warp_field, arap_term_weight=200.0, use_huber_penalty=False, huber_penalty_constant=0.0001):
    """The ARAP term of one iteration as the reference materialises it (ComputeEdgeResiduals, ComputeDenseArapEdgeJacobians,
    ComputeArapBlockSparseHessianApproximation, ComputeNegativeArapGradient): dict(edge_residuals [3E], edge_jacobians [E,5],
    hessian_diagonal_blocks [N,6,6], upper_wing_blocks [E,6,6], upper_wing_block_coordinates [E,2], negative_gradient [6N])."""
    from ..geometry import WarpNodeCoverageComputationMethod
    edges = T.cuda(warp_field.edges, torch.int32)
    e, n = edges.shape[0], warp_field.nodes.shape[0]
    fixed = warp_field.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE
    eli = radii = ncw = None
    if fixed:
        eli, radii = warp_field.edge_layer_indices, warp_field.layer_decimation_radii
    else:
        ncw = warp_field.node_coverage_weights
    out = dict(edge_residuals=T.empty((3 * e,), torch.float32), edge_jacobians=T.empty((e, 5), torch.float32),
               hessian_diagonal_blocks=T.empty((n, 6, 6), torch.float32), upper_wing_blocks=T.empty((e, 6, 6), torch.float32),
               upper_wing_block_coordinates=T.empty((e, 2), torch.int32), negative_gradient=T.empty((6 * n,), torch.float32))
    T.call("nnrt_b200_arap_term", T.ptr(edges), e, T.ptr(eli), T.ptr(radii), T.ptr(ncw), T.ptr(warp_field.nodes), T.ptr(warp_field.node_rotations),
           T.ptr(warp_field.node_translations), n, float(arap_term_weight), int(bool(use_huber_penalty)), float(huber_penalty_constant),
           T.ptr(out["edge_residuals"]), T.ptr(out["edge_jacobians"]), T.ptr(out["hessian_diagonal_blocks"]), T.ptr(out["upper_wing_blocks"]),
           T.ptr(out["upper_wing_block_coordinates"]), T.ptr(out["negative_gradient"]), T.stream())
    return out
