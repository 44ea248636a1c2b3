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
