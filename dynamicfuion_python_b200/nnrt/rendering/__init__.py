"""nnrt.rendering — rasterize_ndc_triangles (cpp/pybind/rendering/rendering.cpp:35-42)."""
import torch

from .. import _tensor as T
from . import functional


def rasterize_ndc_triangles(ndc_face_vertices, clipped_faces_mask, image_size, blur_radius_pixels=0.0, faces_per_pixel=8, bin_size=-1,
                            max_faces_per_bin=-1, perspective_correct_barycentric_coordinates=False, clip_barycentric_coordinates=False,
                            cull_back_faces=True):
    """Returns (pixel_face_indices [H,W,K] int64, pixel_depths [H,W,K], pixel_barycentric_coordinates [H,W,K,3], pixel_face_distances [H,W,K])."""
    if len(image_size) != 2:
        raise RuntimeError(f"image_size should be a SizeVector of size 2. Got size {len(image_size)}.")
    faces = T.cuda(ndc_face_vertices, torch.float32)
    if faces.dim() != 3 or tuple(faces.shape[1:]) != (3, 3):
        raise RuntimeError(f"ndc_face_vertices must have shape [F,3,3], got {tuple(faces.shape)}")
    mask = None if clipped_faces_mask is None else T.cuda(clipped_faces_mask, torch.uint8)
    h, w, k = int(image_size[0]), int(image_size[1]), int(faces_per_pixel)
    idx = T.empty((h, w, k), torch.int64)
    depth = T.empty((h, w, k), torch.float32)
    bary = T.empty((h, w, k, 3), torch.float32)
    dist = T.empty((h, w, k), torch.float32)
    T.call("nnrt_b200_rasterize_ndc_triangles", T.ptr(faces), T.ptr(mask), faces.shape[0], h, w, float(blur_radius_pixels), k, int(bin_size),
           int(max_faces_per_bin), int(bool(perspective_correct_barycentric_coordinates)), int(bool(clip_barycentric_coordinates)),
           int(bool(cull_back_faces)), T.ptr(idx), T.ptr(depth), T.ptr(bary), T.ptr(dist), T.stream())
    return idx, depth, bary, dist
