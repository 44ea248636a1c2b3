"""nnrt.geometry.functional — same names, argument order and defaults as cpp/pybind/geometry/functional/functional.cpp:51-140."""
import torch

from .. import _tensor as T


def _anchors(points, nodes, node_weights, node_coverage, anchor_count, minimum_valid_anchor_count):
    pts = T.cuda(points, torch.float32)
    shape = pts.shape
    if shape[-1] != 3:
        raise RuntimeError(f"points must have shape [..., 3], got {tuple(shape)}")
    flat = pts.reshape(-1, 3)
    nd = T.cuda(nodes, torch.float32)
    nw = None if node_weights is None else T.cuda(node_weights, torch.float32)
    if nw is not None and nw.numel() != nd.shape[0]:
        raise RuntimeError("node_coverage_weights must have one entry per node")
    anchors = T.empty((flat.shape[0], anchor_count), torch.int32)
    weights = T.empty((flat.shape[0], anchor_count), torch.float32)
    T.call("nnrt_b200_compute_anchors_euclidean", T.ptr(flat), flat.shape[0], T.ptr(nd), nd.shape[0], T.ptr(nw), float(node_coverage),
           int(anchor_count), int(minimum_valid_anchor_count), T.ptr(anchors), T.ptr(weights), T.stream())
    out_shape = tuple(shape[:-1]) + (anchor_count,)
    return anchors.reshape(out_shape), weights.reshape(out_shape)


def compute_anchors_and_weights_euclidean_fixed_node_weight(points, nodes, anchor_count, minimum_valid_anchor_count, node_coverage_weight):
    """ComputeAnchorsAndWeights_Euclidean_FixedNodeWeight (cpp/geometry/functional/WarpAnchorComputation.cpp:74-110)."""
    return _anchors(points, nodes, None, node_coverage_weight, anchor_count, minimum_valid_anchor_count)


def compute_anchors_and_weights_euclidean_variable_node_weight(points, nodes, node_coverage_weights, anchor_count, minimum_valid_anchor_count):
    """ComputeAnchorsAndWeights_Euclidean_VariableNodeWeight (cpp/geometry/functional/WarpAnchorComputation.cpp:112-148)."""
    return _anchors(points, nodes, node_coverage_weights, 0.0, anchor_count, minimum_valid_anchor_count)


def _anchors_shortest_path(points, nodes, node_weights, node_coverage, edges, anchor_count):
    pts = T.cuda(points, torch.float32)
    shape = pts.shape
    if pts.dim() < 2 or pts.dim() > 3:  # WarpAnchorComputation.cpp:41-44
        raise RuntimeError(f"`points` needs to have 2 or 3 dimensions. Got: {pts.dim()} dimensions.")
    if shape[-1] != 3:
        raise RuntimeError(f"points must have shape [..., 3], got {tuple(shape)}")
    if anchor_count < 1:
        raise RuntimeError(f"anchor_count needs to be greater than one. Got: {anchor_count}.")
    flat = pts.reshape(-1, 3)
    nd = T.cuda(nodes, torch.float32)
    e = T.to_torch(edges)
    if e.dtype != torch.int32:  # o3c::AssertTensorDtype(edges, Int32), WarpAnchorComputation.cpp:169,197
        raise RuntimeError(f"edges must be Int32, got {e.dtype}")
    e = T.cuda(e, torch.int32)
    if e.dim() != 2 or e.shape[0] != nd.shape[0]:
        raise RuntimeError(f"edges must have shape [node_count, degree], got {tuple(e.shape)}")
    nw = None if node_weights is None else T.cuda(node_weights, torch.float32)
    if nw is not None and nw.numel() != nd.shape[0]:
        raise RuntimeError("node_coverage_weights must have one entry per node")
    anchors = T.empty((flat.shape[0], anchor_count), torch.int32)
    weights = T.empty((flat.shape[0], anchor_count), torch.float32)
    T.call("nnrt_b200_compute_anchors_shortest_path", T.ptr(flat), flat.shape[0], T.ptr(nd), nd.shape[0], T.ptr(nw), float(node_coverage),
           T.ptr(e), e.shape[1], int(anchor_count), T.ptr(anchors), T.ptr(weights), T.stream())
    out_shape = tuple(shape[:-1]) + (anchor_count,)
    return anchors.reshape(out_shape), weights.reshape(out_shape)


def compute_anchors_and_weights_shortest_path_fixed_node_weight(points, nodes, edges, anchor_count, node_coverage_weight):
    """ComputeAnchorsAndWeights_ShortestPath_FixedNodeWeight (cpp/geometry/functional/WarpAnchorComputation.cpp:191-214);
    edges [N, degree] int32 adjacency, -1 = none."""
    return _anchors_shortest_path(points, nodes, None, node_coverage_weight, edges, anchor_count)


def compute_anchors_and_weights_shortest_path_variable_node_weight(points, nodes, node_coverage_weights, edges, anchor_count):
    """ComputeAnchorsAndWeights_ShortestPath_VariableNodeWeight (cpp/geometry/functional/WarpAnchorComputation.cpp:158-189)."""
    return _anchors_shortest_path(points, nodes, node_coverage_weights, 0.0, edges, anchor_count)


def _warp(points, normals, nodes, node_rotations, node_translations, anchors, anchor_weights, minimum_valid_anchor_count, extrinsics,
          dual_quaternion=False):
    nd = T.cuda(nodes, torch.float32)
    R = T.cuda(node_rotations, torch.float32)
    t = T.cuda(node_translations, torch.float32)
    a = T.cuda(anchors, torch.int32)
    w = T.cuda(anchor_weights, torch.float32)
    if tuple(R.shape) != (nd.shape[0], 3, 3) or tuple(t.shape) != (nd.shape[0], 3):
        raise RuntimeError("node_rotations must be [N,3,3] and node_translations [N,3]")
    if a.shape != w.shape or a.shape[0] != points.shape[0]:
        raise RuntimeError("anchors / anchor_weights must be [point_count, anchor_count]")
    ext = T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsics is None else extrinsics, (4, 4))
    wp = torch.empty_like(points)
    wn = None if normals is None else torch.empty_like(normals)
    T.call("nnrt_b200_warp_points_normals_dual_quaternion" if dual_quaternion else "nnrt_b200_warp_points_normals", T.ptr(points), T.ptr(normals), points.shape[0], T.ptr(nd), T.ptr(R), T.ptr(t), nd.shape[0],
           T.ptr(a), T.ptr(w), a.shape[1], int(minimum_valid_anchor_count), ext.data_ptr(), T.ptr(wp), T.ptr(wn), T.stream())
    return wp, wn


def warp_triangle_mesh(input_mesh, nodes, node_rotations, node_translations, *args, threshold_nodes_by_distance=False,
                       minimum_valid_anchor_count=0, extrinsics=None, **kwargs):
    """Both overloads of warp_triangle_mesh (cpp/geometry/functional/Warping.cpp:169-263): (anchor_count, node_coverage, ...) computes
    anchors on the fly; (anchors, anchor_weights, ...) uses supplied anchors."""
    names = ["a0", "a1", "threshold_nodes_by_distance", "minimum_valid_anchor_count", "extrinsics"]
    vals = dict(zip(names, args))
    for key in ("anchor_count", "anchors"):
        if key in kwargs:
            vals["a0"] = kwargs.pop(key)
    for key in ("node_coverage", "anchor_weights"):
        if key in kwargs:
            vals["a1"] = kwargs.pop(key)
    if kwargs:
        raise TypeError(f"unexpected arguments {sorted(kwargs)}")
    threshold = vals.get("threshold_nodes_by_distance", threshold_nodes_by_distance)
    min_valid = vals.get("minimum_valid_anchor_count", minimum_valid_anchor_count)
    ext = vals.get("extrinsics", extrinsics)
    mesh = T.MeshView(input_mesh)
    if isinstance(vals["a0"], int):
        anchors, weights = compute_anchors_and_weights_euclidean_fixed_node_weight(mesh.positions, nodes, vals["a0"],
                                                                                   min_valid if threshold else 0, vals["a1"])
    else:
        anchors, weights = vals["a0"], vals["a1"]
    wp, wn = _warp(mesh.positions, mesh.normals, nodes, node_rotations, node_translations, anchors, weights,
                   min_valid if threshold else 0, ext)
    return mesh.with_vertices(wp, wn)


def warp_point_cloud(input_point_cloud, nodes, node_rotations, node_translations, a0, a1, minimum_valid_anchor_count, extrinsics=None):
    """warp_point_cloud, both overloads (cpp/pybind/geometry/functional/functional.cpp:94-109). Returns warped positions [P,3]."""
    pts = input_point_cloud.point["positions"] if hasattr(input_point_cloud, "point") else input_point_cloud
    pts = T.cuda(pts, torch.float32)
    if isinstance(a0, int):
        a0, a1 = compute_anchors_and_weights_euclidean_fixed_node_weight(pts, nodes, a0, minimum_valid_anchor_count, a1)
    wp, _ = _warp(pts, None, nodes, node_rotations, node_translations, a0, a1, minimum_valid_anchor_count, extrinsics)
    return wp


def compute_point_to_plane_distances(mesh1, mesh2_or_point_cloud):
    """ComputePointToPlaneDistances (cpp/geometry/functional/kernel/PointToPlaneDistancesImpl.h:25-52): n1 . (v1 - v2)."""
    m1 = T.MeshView(mesh1)
    if m1.normals is None:
        raise RuntimeError("mesh1 needs to have vertex normals defined.")
    if hasattr(mesh2_or_point_cloud, "point"):
        v2 = T.cuda(mesh2_or_point_cloud.point["positions"], torch.float32)
    elif isinstance(mesh2_or_point_cloud, (torch.Tensor,)) or type(mesh2_or_point_cloud).__module__ == "numpy":
        v2 = T.cuda(mesh2_or_point_cloud, torch.float32)
    else:
        v2 = T.MeshView(mesh2_or_point_cloud).positions
    if v2.shape != m1.positions.shape:
        raise RuntimeError("both point sets must have the same number of points")
    out = T.empty((m1.positions.shape[0],), torch.float32)
    T.call("nnrt_b200_point_to_plane_distances", T.ptr(m1.normals), T.ptr(m1.positions), T.ptr(v2), v2.shape[0], T.ptr(out), T.stream())
    return out


def unproject_raster_depth_without_filtering(depth, intrinsics, extrinsics=None, depth_scale=1000.0, depth_max=3.0, preserve_pixel_layout=False):
    """UnprojectDepthImageWithoutFiltering (cpp/geometry/functional/kernel/PerspectiveProjectionImpl.h:62-150) -> (points, mask)."""
    d = T.to_torch(depth.as_tensor() if hasattr(depth, "as_tensor") else depth)
    if d.dim() == 3 and d.shape[2] == 1:
        d = d[..., 0]
    if d.dtype in (torch.uint16, torch.int16):
        d = T.cuda(d.view(torch.int16) if d.dtype == torch.uint16 else d, torch.int16)
        dtype_code = 1
    elif d.dtype == torch.float32:
        d = T.cuda(d, torch.float32)
        dtype_code = 0
    else:
        raise RuntimeError("depth must be UInt16 or Float32")
    rows, cols = d.shape
    K = T.host_f64(intrinsics, (3, 3))
    ext = T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsics is None else extrinsics, (4, 4))
    pts = T.empty((rows * cols, 3), torch.float32)
    mask = T.empty((rows * cols,), torch.uint8)
    T.call("nnrt_b200_unproject_depth", T.ptr(d), dtype_code, rows, cols, K.data_ptr(), ext.data_ptr(), float(depth_scale), float(depth_max),
           T.ptr(pts), T.ptr(mask), T.stream())
    mask = mask.to(torch.bool)
    if preserve_pixel_layout:
        return pts.reshape(rows, cols, 3), mask.reshape(rows, cols)
    return pts, mask


def median_grid_subsample_3d_points(points, grid_cell_size, _grid_offset=0.0):
    """MedianGridSubsample3dPoints (cpp/pybind/geometry/functional/functional.cpp:152-154; cpp/geometry/functional/GeometrySampling.cpp:61-68):
    int64 indices of the per-grid-cell median points (smallest sum of distances to the cell's points), cells in order of their first point."""
    import ctypes as C
    pts = T.cuda(points, torch.float32)
    if pts.dim() != 2 or pts.shape[1] != 3:
        raise RuntimeError(f"points must have shape [N, 3], got {tuple(pts.shape)}")
    sample = T.empty((pts.shape[0],), torch.int64)
    count = C.c_int64()
    T.call("nnrt_b200_median_grid_subsample_3d_points", T.ptr(pts), pts.shape[0], float(grid_cell_size), float(_grid_offset), T.ptr(sample), C.byref(count), T.stream())
    return sample[:count.value]


def compute_triangle_normals(mesh, normalized=True):
    """ComputeTriangleNormals (cpp/geometry/functional/NormalsOperations.cpp:36-46): sets the mesh's triangle normals (where the mesh
    argument is mutable) and returns them, [F,3] f32."""
    view = T.MeshView(mesh)
    if view.triangles is None:
        raise RuntimeError("Mesh needs to have both vertex positions and triangle indices to compute triangle normals. "
                           "Vertex positions: present. Triangle indices: absent.")
    normals = T.empty((view.triangles.shape[0], 3), torch.float32)
    T.call("nnrt_b200_compute_triangle_normals", T.ptr(view.positions), view.positions.shape[0], T.ptr(view.triangles), view.triangles.shape[0],
           int(bool(normalized)), T.ptr(normals), T.stream())
    T.store_mesh_attribute(mesh, "triangle", "normals", normals)
    return normals


def compute_vertex_normals(mesh, normalized=True):
    """ComputeVertexNormals (cpp/geometry/functional/NormalsOperations.cpp:52-67): uses the mesh's triangle normals if it has them,
    otherwise computes (and stores) un-normalised ones; sets the mesh's vertex normals and returns them, [V,3] f32."""
    view = T.MeshView(mesh)
    if view.triangles is None:
        raise RuntimeError("Mesh needs to have both vertex positions and triangle indices to compute triangle normals. "
                           "Vertex positions: present. Triangle indices: absent.")
    triangle_normals = view.triangle_normals
    if triangle_normals is None:
        triangle_normals = compute_triangle_normals(mesh, False)
    normals = T.empty((view.positions.shape[0], 3), torch.float32)
    T.call("nnrt_b200_compute_vertex_normals", T.ptr(view.triangles), view.triangles.shape[0], T.ptr(triangle_normals), view.positions.shape[0],
           int(bool(normalized)), T.ptr(normals), T.stream())
    T.store_mesh_attribute(mesh, "vertex", "normals", normals)
    return normals


def compute_ordered_point_cloud_normals(point_cloud, source_image_size):
    """ComputeOrderedPointCloudNormals (cpp/geometry/functional/NormalsOperations.cpp:73-93) -> normals [H*W,3] f32."""
    if len(source_image_size) != 2:
        raise RuntimeError(f"Source image size must have two dimensions. Got {len(source_image_size)}.")
    pts = point_cloud.point["positions"] if hasattr(point_cloud, "point") else point_cloud
    pts = T.cuda(pts, torch.float32).reshape(-1, 3)
    h, w = int(source_image_size[0]), int(source_image_size[1])
    if pts.shape[0] != h * w:
        raise RuntimeError(f"Point cloud point count (got {pts.shape[0]}) must equal the multiple of dimensions (got {h} * {w} = {h * w}).")
    normals = T.empty((h * w, 3), torch.float32)
    T.call("nnrt_b200_compute_ordered_point_cloud_normals", T.ptr(pts), h, w, T.ptr(normals), T.stream())
    return normals


def integrate_non_rigid_voxel_blocks(block_indices, block_keys, block_values, block_resolution, voxel_size, warp_field, depth, color,
                                     depth_normals, depth_intrinsics, color_intrinsics=None, extrinsics=None, depth_scale=1000.0, depth_max=3.0,
                                     truncation_voxel_multiplier=2.0, _full_scan=False):
    """The voxel kernel of NonRigidSurfaceVoxelBlockGrid.integrate_non_rigid (cpp/geometry/NonRigidSurfaceVoxelBlockGrid.cpp:32-74;
    kernel/NonRigidSurfaceVoxelBlockGridImpl.h:50-229) over a voxel block grid's own tensors: `block_indices` [B] the active buffer
    indices, `block_keys` [capacity,3] int32 the hash map's key tensor, `block_values` a dict with "tsdf" [capacity,R,R,R] f32,
    "weight" (f32 or uint16) and optionally "color" [capacity,R,R,R,3] f32 — CUDA tensors updated IN PLACE (with Open3D present:
    `torch.utils.dlpack.from_dlpack(grid.attribute("tsdf").to_dlpack())`). `warp_field` is a GraphWarpField; its
    minimum_valid_anchor_count and coverage mode apply (WarpField::ComputeAnchorsForPoint<threshold = true>).
    Returns cos_voxel_ray_to_normal [H,W] f32."""
    from . import WarpNodeCoverageComputationMethod
    idx = T.cuda(block_indices, torch.int64)
    keys = T.cuda(block_keys, torch.int32)
    tsdf, weight = block_values["tsdf"], block_values["weight"]
    voxel_color = block_values.get("color")
    for name, tensor in (("tsdf", tsdf), ("weight", weight), ("color", voxel_color)):
        if tensor is not None and not (isinstance(tensor, torch.Tensor) and tensor.is_cuda and tensor.is_contiguous()):
            raise RuntimeError(f"block_values['{name}'] must be a contiguous CUDA tensor (it is updated in place)")
    if tsdf.dtype != torch.float32 or (voxel_color is not None and voxel_color.dtype != torch.float32):
        raise RuntimeError("tsdf and color voxel attributes must be Float32")
    if weight.dtype == torch.float32:
        weight_code = 0
    elif weight.dtype in (torch.uint16, torch.int16):
        weight_code = 1
    else:
        raise RuntimeError("weight voxel attribute must be Float32 or UInt16")
    d = T.to_torch(depth.as_tensor() if hasattr(depth, "as_tensor") else depth)
    if d.dim() == 3:
        d = d[..., 0]
    if d.dtype in (torch.uint16, torch.int16):
        d, depth_code = T.cuda(d.view(torch.int16) if d.dtype == torch.uint16 else d, torch.int16), 1
    elif d.dtype == torch.float32:
        d, depth_code = T.cuda(d, torch.float32), 0
    else:
        raise RuntimeError("depth must be UInt16 or Float32")
    h, w = d.shape
    c = None
    color_code, ch, cw = 0, 0, 0
    if color is not None and voxel_color is not None:
        c = T.to_torch(color.as_tensor() if hasattr(color, "as_tensor") else color)
        if c.numel() == 0:
            c = None
        else:
            color_code = 0 if c.dtype == torch.float32 else 1
            c = T.cuda(c, torch.float32 if color_code == 0 else torch.uint8)
            ch, cw = int(c.shape[0]), int(c.shape[1])
    normals = T.cuda(depth_normals, torch.float32)
    if normals.numel() != h * w * 3:
        raise RuntimeError("depth_normals must have shape [H,W,3]")
    Kd = T.host_f64(depth_intrinsics, (3, 3))
    Kc = T.host_f64(depth_intrinsics if color_intrinsics is None else color_intrinsics, (3, 3))
    ext = T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsics is None else extrinsics, (4, 4))
    fixed = warp_field.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE
    ncw = None if fixed else warp_field.node_coverage_weights
    cos = T.empty((h, w), torch.float32)
    T.call("nnrt_b200_integrate_non_rigid_full_scan" if _full_scan else "nnrt_b200_integrate_non_rigid", T.ptr(idx), idx.shape[0], T.ptr(keys), T.ptr(tsdf), T.ptr(weight), weight_code,
           T.ptr(voxel_color) if c is not None else None, int(block_resolution), float(voxel_size), float(voxel_size) * float(truncation_voxel_multiplier),
           T.ptr(d), depth_code, T.ptr(c), color_code, T.ptr(normals), h, w, ch, cw, Kd.data_ptr(), Kc.data_ptr(), ext.data_ptr(),
           T.ptr(warp_field.nodes), T.ptr(warp_field.node_rotations), T.ptr(warp_field.node_translations), warp_field.nodes.shape[0], T.ptr(ncw),
           float(warp_field.node_coverage), warp_field.anchor_count, warp_field.minimum_valid_anchor_count, float(depth_scale), float(depth_max),
           T.ptr(cos), T.stream())
    return cos


def warp_triangle_mesh_dual_quaternion(input_mesh, nodes, node_rotations, node_translations, anchors, anchor_weights,
                                       minimum_valid_anchor_count=0, extrinsics=None):
    """Dual-quaternion blended counterpart of warp_triangle_mesh with supplied anchors (no reference implementation exists; Kavan et
    al. 2008 on the node motions x -> R (x - g) + g + t). Volume-preserving where linear blending of distinct rotations shrinks."""
    mesh = T.MeshView(input_mesh)
    wp, wn = _warp(mesh.positions, mesh.normals, nodes, node_rotations, node_translations, anchors, anchor_weights, minimum_valid_anchor_count,
                   extrinsics, dual_quaternion=True)
    return mesh.with_vertices(wp, wn)


def warp_point_cloud_dual_quaternion(input_point_cloud, nodes, node_rotations, node_translations, anchors, anchor_weights,
                                     minimum_valid_anchor_count=0, extrinsics=None):
    pts = input_point_cloud.point["positions"] if hasattr(input_point_cloud, "point") else input_point_cloud
    pts = T.cuda(pts, torch.float32)
    wp, _ = _warp(pts, None, nodes, node_rotations, node_translations, anchors, anchor_weights, minimum_valid_anchor_count, extrinsics, dual_quaternion=True)
    return wp


def integrate_voxel_blocks(block_indices, block_keys, block_values, block_resolution, voxel_size, depth, color, depth_intrinsics,
                           color_intrinsics=None, extrinsics=None, depth_scale=1000.0, depth_max=3.0, truncation_voxel_multiplier=8.0):
    """The voxel kernel of VoxelBlockGrid.integrate (Open3D's rigid integration, re-exported by the reference at
    cpp/pybind/geometry/geometry.cpp:170-203) over a grid's own tensors, as integrate_non_rigid_voxel_blocks takes them. tsdf, weight
    (raised by one per integrated voxel) and colour are updated IN PLACE."""
    idx = T.cuda(block_indices, torch.int64)
    keys = T.cuda(block_keys, torch.int32)
    tsdf, weight = block_values["tsdf"], block_values["weight"]
    voxel_color = block_values.get("color")
    for name, tensor in (("tsdf", tsdf), ("weight", weight), ("color", voxel_color)):
        if tensor is not None and not (isinstance(tensor, torch.Tensor) and tensor.is_cuda and tensor.is_contiguous()):
            raise RuntimeError(f"block_values['{name}'] must be a contiguous CUDA tensor (it is updated in place)")
    if tsdf.dtype != torch.float32 or (voxel_color is not None and voxel_color.dtype != torch.float32):
        raise RuntimeError("tsdf and color voxel attributes must be Float32")
    if weight.dtype == torch.float32:
        weight_code = 0
    elif weight.dtype in (torch.uint16, torch.int16):
        weight_code = 1
    else:
        raise RuntimeError("weight voxel attribute must be Float32 or UInt16")
    d = T.to_torch(depth.as_tensor() if hasattr(depth, "as_tensor") else depth)
    if d.dim() == 3:
        d = d[..., 0]
    if d.dtype in (torch.uint16, torch.int16):
        d, depth_code = T.cuda(d.view(torch.int16) if d.dtype == torch.uint16 else d, torch.int16), 1
    elif d.dtype == torch.float32:
        d, depth_code = T.cuda(d, torch.float32), 0
    else:
        raise RuntimeError("depth must be UInt16 or Float32")
    h, w = d.shape
    c, color_code, ch, cw = None, 0, 0, 0
    if color is not None and voxel_color is not None:
        c = T.to_torch(color.as_tensor() if hasattr(color, "as_tensor") else color)
        if c.numel() == 0:
            c = None
        else:
            color_code = 0 if c.dtype == torch.float32 else 1
            c = T.cuda(c, torch.float32 if color_code == 0 else torch.uint8)
            ch, cw = int(c.shape[0]), int(c.shape[1])
    Kd = T.host_f64(depth_intrinsics, (3, 3))
    Kc = T.host_f64(depth_intrinsics if color_intrinsics is None else color_intrinsics, (3, 3))
    ext = T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsics is None else extrinsics, (4, 4))
    T.call("nnrt_b200_integrate_rigid", T.ptr(idx), idx.shape[0], T.ptr(keys), T.ptr(tsdf), T.ptr(weight), weight_code,
           T.ptr(voxel_color) if c is not None else None, int(block_resolution), float(voxel_size), float(voxel_size) * float(truncation_voxel_multiplier),
           T.ptr(d), depth_code, T.ptr(c), color_code, h, w, ch, cw, Kd.data_ptr(), Kc.data_ptr(), ext.data_ptr(), float(depth_scale), float(depth_max),
           T.stream())


def bounding_boxes_of_warped_blocks(block_keys, warp_field, voxel_size, block_resolution, extrinsics=None):
    """NonRigidSurfaceVoxelBlockGrid.GetBoundingBoxesOfWarpedBlocks (cpp/geometry/NonRigidSurfaceVoxelBlockGrid.cpp:120-130) -> [B,6] f32."""
    keys = T.cuda(block_keys, torch.int32)
    if keys.dim() != 2 or keys.shape[1] != 3:
        raise RuntimeError(f"block_keys must have shape [B, 3], got {tuple(keys.shape)}")
    ext = T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsics is None else extrinsics, (4, 4))
    boxes = T.empty((keys.shape[0], 6), torch.float32)
    T.call("nnrt_b200_bounding_boxes_of_warped_blocks", T.ptr(keys), keys.shape[0], T.ptr(warp_field.nodes), T.ptr(warp_field.node_rotations),
           T.ptr(warp_field.node_translations), warp_field.nodes.shape[0], float(warp_field.node_coverage), warp_field.anchor_count,
           warp_field.minimum_valid_anchor_count, float(voxel_size), int(block_resolution), ext.data_ptr(), T.ptr(boxes), T.stream())
    return boxes


def axis_aligned_boxes_intersecting_surface_mask(boxes, depth, intrinsics, depth_scale, depth_max, downsampling_factor, truncation_distance):
    """NonRigidSurfaceVoxelBlockGrid.GetAxisAlignedBoxesIntersectingSurfaceMask (cpp/geometry/NonRigidSurfaceVoxelBlockGrid.cpp:132-148);
    `truncation_distance` = voxel_size * truncation_voxel_multiplier. Returns a bool mask [B]."""
    b = T.cuda(boxes, torch.float32)
    if b.dim() != 2 or b.shape[1] != 6:
        raise RuntimeError(f"boxes must have shape [B, 6], got {tuple(b.shape)}")
    d = T.to_torch(depth.as_tensor() if hasattr(depth, "as_tensor") else depth)
    if d.dim() == 3:
        d = d[..., 0]
    if d.dtype in (torch.uint16, torch.int16):
        d, code = T.cuda(d.view(torch.int16) if d.dtype == torch.uint16 else d, torch.int16), 1
    elif d.dtype == torch.float32:
        d, code = T.cuda(d, torch.float32), 0
    else:
        raise RuntimeError("depth must be UInt16 or Float32")
    K = T.host_f64(intrinsics, (3, 3))
    mask = T.empty((b.shape[0],), torch.uint8)
    T.call("nnrt_b200_boxes_intersecting_surface_mask", T.ptr(b), b.shape[0], K.data_ptr(), T.ptr(d), code, int(d.shape[0]), int(d.shape[1]),
           float(depth_scale), float(depth_max), int(downsampling_factor), float(truncation_distance), T.ptr(mask), T.stream())
    return mask.to(torch.bool)


def mean_grid_downsample_3d_points(points, grid_cell_size, _grid_offset=0.0):
    """MeanGridDownsample3dPoints (cpp/pybind/geometry/functional/functional.cpp:143-145): [bins,3] f32 means of the occupied grid cells,
    cells in order of their first point."""
    import ctypes as C
    pts = T.cuda(points, torch.float32)
    if pts.dim() != 2 or pts.shape[1] != 3:
        raise RuntimeError(f"points must have shape [N, 3], got {tuple(pts.shape)}")
    out = T.empty((pts.shape[0], 3), torch.float32)
    count = C.c_int64()
    T.call("nnrt_b200_mean_grid_downsample_3d_points", T.ptr(pts), pts.shape[0], float(grid_cell_size), float(_grid_offset), T.ptr(out), C.byref(count), T.stream())
    return out[:count.value]


def closest_to_mean_grid_subsample_3d_points(points, grid_cell_size):
    """ClosestToGridMeanSubsample3dPoints (cpp/pybind/geometry/functional/functional.cpp:146-148): int64 indices of the points closest to
    the mean of their grid cell."""
    import ctypes as C
    pts = T.cuda(points, torch.float32)
    if pts.dim() != 2 or pts.shape[1] != 3:
        raise RuntimeError(f"points must have shape [N, 3], got {tuple(pts.shape)}")
    out = T.empty((pts.shape[0],), torch.int64)
    count = C.c_int64()
    T.call("nnrt_b200_closest_to_mean_grid_subsample_3d_points", T.ptr(pts), pts.shape[0], float(grid_cell_size), 0.0, T.ptr(out), C.byref(count), T.stream())
    return out[:count.value]


def _extended_cell(radius):
    import numpy as np
    r = np.float32(radius)
    return float(np.sqrt(np.float32(2) * (r * r), dtype=np.float32) * np.float32(2))  # sqrtf(2 * (radius * radius)) * 2


def fast_mean_radius_downsample_3d_points(points, radius):
    """FastMeanRadiusDownsample3dPoints (cpp/geometry/functional/kernel/GeometrySamplingImpl.h:47-59): cell means on a grid of
    2 * sqrt(2) * radius, merged again on the same grid shifted by half a cell."""
    cell = _extended_cell(radius)
    return mean_grid_downsample_3d_points(mean_grid_downsample_3d_points(points, cell), cell, _grid_offset=0.5)


def fast_median_radius_subsample_3d_points(points, radius):
    """FastMedianRadiusSubsample3dPoints (GeometrySamplingImpl.h:78-91): cell medians, then the medians of those on the shifted grid;
    int64 indices into `points`."""
    cell = _extended_cell(radius)
    pts = T.cuda(points, torch.float32)
    stage1 = median_grid_subsample_3d_points(pts, cell)
    stage2 = median_grid_subsample_3d_points(pts[stage1].contiguous(), cell, _grid_offset=0.5)
    return stage1[stage2]


# --- mesh extraction from TSDF voxel blocks (cpp/geometry/VoxelBlockGrid.cpp:461-484 -> Open3D's kernel) ----------
def voxel_block_neighbors(block_keys):
    """The block hash behind VoxelBlockGrid::BufferRadiusNeighbors: for block_keys [B,3] int32 the [B,27] int32 indices of the blocks at
    key + (dx, dy, dz) (slot (dz + 1) * 9 + (dy + 1) * 3 + (dx + 1)), -1 where no such block exists."""
    keys = T.cuda(block_keys, torch.int32)
    if keys.dim() != 2 or keys.shape[1] != 3:
        raise RuntimeError(f"block_keys must have shape [B, 3], got {tuple(keys.shape)}")
    out = T.empty((keys.shape[0], 27), torch.int32)
    T.call("nnrt_b200_voxel_block_neighbors", T.ptr(keys), keys.shape[0], T.ptr(out), T.stream())
    return out


def extract_triangle_mesh(block_keys, block_values, voxel_size, weight_threshold=3.0):
    """VoxelBlockGrid::ExtractTriangleMesh(weight_threshold) on explicit block tensors: block_keys [B,3] int32, block_values a dict with
    "tsdf" [B,R,R,R] f32, "weight" (f32 or uint16) and optionally "color" [B,R,R,R,3] f32 (0..255). Returns a mesh (positions,
    normals = normalised TSDF gradients, triangles int64; `.colors` in [0, 1] when colour is given). Vertices and triangles come in
    (block, voxel, edge) order, reproducibly."""
    import ctypes as C
    keys = T.cuda(block_keys, torch.int32)
    tsdf = T.cuda(block_values["tsdf"], torch.float32)
    if tsdf.dim() != 4 or tsdf.shape[1] != tsdf.shape[2] or tsdf.shape[1] != tsdf.shape[3] or tsdf.shape[0] != keys.shape[0]:
        raise RuntimeError(f"tsdf must have shape [B,R,R,R] with one block per key, got {tuple(tsdf.shape)}")
    weight = T.to_torch(block_values["weight"])
    if weight.dtype in (torch.uint16, torch.int16):
        weight, weight_code = T.cuda(weight.view(torch.int16) if weight.dtype == torch.uint16 else weight, torch.int16), 1
    elif weight.dtype == torch.float32:
        weight, weight_code = T.cuda(weight, torch.float32), 0
    else:
        raise RuntimeError("weight voxel attribute must be Float32 or UInt16")
    if tuple(weight.shape) != tuple(tsdf.shape):
        raise RuntimeError("weight must have the shape of tsdf")
    color = None if block_values.get("color") is None else T.cuda(block_values["color"], torch.float32)
    if color is not None and tuple(color.shape) != tuple(tsdf.shape) + (3,):
        raise RuntimeError("color must have shape [B,R,R,R,3]")
    b, r = keys.shape[0], tsdf.shape[1]
    neighbors = voxel_block_neighbors(keys)
    structure = T.empty((b * r ** 3, 4), torch.int32)
    offsets = T.empty((b * r ** 3 + 1,), torch.int32)
    counts = (C.c_int64 * 2)()
    T.call("nnrt_b200_extract_triangle_mesh_prepare", b, T.ptr(neighbors), T.ptr(tsdf), T.ptr(weight), weight_code, r, float(weight_threshold),
           T.ptr(structure), T.ptr(offsets), counts, T.stream())
    vertices, normals = T.empty((counts[0], 3), torch.float32), T.empty((counts[0], 3), torch.float32)
    colors = None if color is None else T.empty((counts[0], 3), torch.float32)
    triangles = T.empty((counts[1], 3), torch.int64)
    if counts[0] > 0:  # (no surface crossing among sufficiently weighted voxels: an empty mesh, as the reference returns)
        T.call("nnrt_b200_extract_triangle_mesh_emit", T.ptr(keys), b, T.ptr(neighbors), T.ptr(tsdf), T.ptr(color), r, float(voxel_size), T.ptr(structure),
               T.ptr(offsets), T.ptr(vertices), T.ptr(normals), T.ptr(colors), T.ptr(triangles), T.stream())
    mesh = T.MeshView((vertices, normals, triangles))
    mesh.colors = colors
    return mesh
