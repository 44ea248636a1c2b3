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


class FlatEdgeShader:
    """rendering::FlatEdgeShader (cpp/rendering/FlatEdgeShader.h:27-52; C++-only in the reference): draws triangle edges from the
    rasteriser's face distances. Headless: shade_meshes returns an [H,W,3] uint8 device tensor instead of an open3d Image."""

    def __init__(self, pixel_line_width, line_color):
        self.set_pixel_line_width(pixel_line_width)
        self.set_line_color(line_color)

    def set_pixel_line_width(self, width):
        self.pixel_line_width = float(width)

    def set_line_color(self, color):
        if len(color) != 3:
            raise RuntimeError("line_color must have three components")
        self.line_color = tuple(float(c) for c in color)

    def get_ndc_line_width(self, image_height, image_width):
        return self.pixel_line_width / (min(image_height, image_width) / 2.0)

    def shade_meshes(self, pixel_face_indices, pixel_depths, pixel_barycentric_coordinates, pixel_face_distances, meshes=None):
        import ctypes as C
        faces = T.cuda(pixel_face_indices, torch.int64)
        dist = T.cuda(pixel_face_distances, torch.float32)
        if faces.dim() != 3 or dist.shape != faces.shape:
            raise RuntimeError("pixel_face_indices and pixel_face_distances must have shape [H,W,K]")
        h, w, k = faces.shape
        pixels = T.empty((h, w, 3), torch.uint8)
        color = (C.c_float * 3)(*self.line_color)
        T.call("nnrt_b200_shade_edges_flat", T.ptr(faces), T.ptr(dist), h, w, k, self.pixel_line_width, color, T.ptr(pixels), T.stream())
        return pixels


class VertexColorShader:
    """rendering::VertexColorShader (cpp/rendering/VertexColorShader.cpp:29-59): barycentric interpolation of vertex colours over the
    nearest fragment. `meshes`: the meshes that were rasterised, in order, each with vertex colours in [0, 1] (`.vertex["colors"]`, a
    dict key "colors", or a (positions, colors, triangles) tuple)."""

    @staticmethod
    def _colors_and_triangles(mesh):
        if isinstance(mesh, dict):
            return mesh["colors"], mesh.get("triangles", mesh.get("indices"))
        if isinstance(mesh, (tuple, list)):
            return mesh[1], mesh[2]
        return mesh.vertex["colors"], mesh.triangle["indices"]

    def shade_meshes(self, pixel_face_indices, pixel_depths, pixel_barycentric_coordinates, pixel_face_distances, meshes=None):
        if not meshes:
            raise RuntimeError("VertexColorShader requires meshes argument to be filled with meshes that were used during rasterization "
                               "resulting in its other arguments, in order to retrieve and interpolate the vertex colors.")
        per_mesh = []
        for mesh in meshes:
            colors, triangles = self._colors_and_triangles(mesh)
            per_mesh.append(T.cuda(colors, torch.float32)[T.cuda(triangles, torch.int64)])  # [F,3,3] gather: plumbing, not arithmetic
        face_vertex_colors = (per_mesh[0] if len(per_mesh) == 1 else torch.cat(per_mesh, 0)).contiguous()
        faces = T.cuda(pixel_face_indices, torch.int64)
        bary = T.cuda(pixel_barycentric_coordinates, torch.float32)
        if faces.dim() != 3 or tuple(bary.shape) != tuple(faces.shape) + (3,):
            raise RuntimeError("pixel_face_indices must have shape [H,W,K] and pixel_barycentric_coordinates [H,W,K,3]")
        h, w, k = faces.shape
        pixels = T.empty((h, w, 3), torch.uint8)
        T.call("nnrt_b200_shade_vertex_colors", T.ptr(faces), T.ptr(bary), h, w, k, T.ptr(face_vertex_colors), T.ptr(pixels), T.stream())
        return pixels
