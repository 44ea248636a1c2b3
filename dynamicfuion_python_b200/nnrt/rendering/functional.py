"""nnrt.rendering.functional (cpp/pybind/rendering/functional/functional.cpp:35-61)."""
import torch

from .. import _tensor as T


def get_mesh_ndc_face_vertices_and_clip_mask(camera_space_mesh, intrinsic_matrix, image_size, near_clipping_distance=0.0,
                                             far_clipping_distance=float("inf")):
    """GetMeshNdcFaceVerticesAndClipMask for one mesh or a list of meshes (faces concatenated, cpp/rendering/functional/ExtractFaceVertices.cpp:56-124)."""
    if len(image_size) != 2:
        raise RuntimeError(f"image_size should be a SizeVector of size 2. Got size {len(image_size)}.")
    K = T.host_f64(intrinsic_matrix, (3, 3))
    meshes = camera_space_mesh if isinstance(camera_space_mesh, (list, tuple)) and not _is_mesh_tuple(camera_space_mesh) else [camera_space_mesh]
    outs, masks, counts = [], [], []
    for m in meshes:
        mesh = T.MeshView(m)
        if mesh.triangles is None:
            raise RuntimeError("Argument mesh needs to have both triangle vertex indices and vertex positions.")
        f = mesh.triangles.shape[0]
        out = T.empty((f, 3, 3), torch.float32)
        mask = T.empty((f,), torch.uint8)
        T.call("nnrt_b200_face_vertices_ndc", T.ptr(mesh.positions), mesh.positions.shape[0], T.ptr(mesh.triangles), f, K.data_ptr(),
               int(image_size[0]), int(image_size[1]), float(near_clipping_distance), float(far_clipping_distance), T.ptr(out), T.ptr(mask),
               T.stream())
        outs.append(out); masks.append(mask.to(torch.bool)); counts.append(f)
    if len(meshes) == 1 and meshes[0] is camera_space_mesh:
        return outs[0], masks[0]
    return torch.cat(outs), torch.cat(masks), torch.tensor(counts, dtype=torch.int64)


def _is_mesh_tuple(x):
    return len(x) in (2, 3) and all(hasattr(e, "shape") or e is None for e in x)


def interpolate_vertex_attributes(pixel_face_indices, barycentric_coordinates, face_vertex_attributes):
    """InterpolateVertexAttributes (cpp/rendering/functional/kernel/InterpolateFaceAttributesImpl.h:31-71)."""
    idx = T.cuda(pixel_face_indices, torch.int64)
    bary = T.cuda(barycentric_coordinates, torch.float32)
    attrs = T.cuda(face_vertex_attributes, torch.float32)
    h, w, k = idx.shape
    c = attrs.shape[2]
    out = T.empty((h, w, k, c), torch.float32)
    T.call("nnrt_b200_interpolate_vertex_attributes", T.ptr(idx), T.ptr(bary), h * w, k, T.ptr(attrs), c, T.ptr(out), T.stream())
    return out
