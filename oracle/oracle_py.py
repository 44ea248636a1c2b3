"""ctypes/numpy bindings for the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module. The product package (dynamicfuion_python_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

MAX_ANCHOR_COUNT = 8


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/cpp"):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.oracle_rasterize_ndc_triangles.restype = C.c_int64
        _LIB.oracle_pixel_vertex_anchor_jacobians.restype = C.c_int64
    return _LIB


def set_num_threads(n):
    """Sets the OpenMP team size of the oracle (returns the size in effect); overrides an inherited OMP_NUM_THREADS."""
    return int(lib().oracle_set_num_threads(C.c_int(int(n))))


def ref_lib():
    """The reference's own device-function headers compiled verbatim (None if not built)."""
    global _REF
    if _REF is None:
        so = os.path.join(_HERE, "_ref", "libnnrt_ref.so")
        if not os.path.exists(so):
            return None
        _REF = C.CDLL(so)
    return _REF


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def compute_anchors(points, nodes, anchor_count=4, min_valid=0, node_coverage=0.05, node_weights=None):
    points, nodes = _f32(points), _f32(nodes)
    nw = None if node_weights is None else _f32(node_weights)
    V = points.shape[0]
    anchors = np.empty((V, anchor_count), np.int32)
    weights = np.empty((V, anchor_count), np.float32)
    lib().oracle_compute_anchors_euclidean(_p(points), C.c_int64(V), _p(nodes), C.c_int(nodes.shape[0]), _p(nw),
                                           C.c_float(node_coverage), C.c_int(anchor_count), C.c_int(min_valid),
                                           _p(anchors), _p(weights))
    return anchors, weights


def warp(points, normals, nodes, rotations, translations, anchors, weights, extrinsics=None, min_valid=0):
    points, normals = _f32(points), _f32(normals)
    nodes, rotations, translations = _f32(nodes), _f32(rotations), _f32(translations)
    anchors, weights = _i32(anchors), _f32(weights)
    ext = _f64(np.eye(4) if extrinsics is None else extrinsics)
    wp, wn = np.empty_like(points), np.empty_like(normals)
    lib().oracle_warp_points_normals(_p(points), _p(normals), C.c_int64(points.shape[0]), _p(nodes), _p(rotations), _p(translations),
                                     _p(anchors), _p(weights), C.c_int(anchors.shape[1]), C.c_int(min_valid), _p(ext), _p(wp), _p(wn))
    return wp, wn


def intrinsics_to_ndc(K, height, width):
    K = _f64(K)
    out = np.empty(9, np.float64)
    rng = np.empty(4, np.float32)
    lib().oracle_intrinsics_to_ndc(_p(K), C.c_int(height), C.c_int(width), _p(out), _p(rng))
    return out.reshape(3, 3), rng


def face_vertices_ndc(positions, triangles, K, height, width, near=0.0, far=float("inf")):
    positions, triangles, K = _f32(positions), _i64(triangles), _f64(K)
    F = triangles.shape[0]
    out = np.empty((F, 3, 3), np.float32)
    mask = np.empty(F, np.uint8)
    lib().oracle_face_vertices_ndc(_p(positions), _p(triangles), C.c_int64(F), _p(K), C.c_int(height), C.c_int(width),
                                   C.c_float(near), C.c_float(far), _p(out), _p(mask))
    return out, mask.astype(bool)


def rasterize(face_ndc, clip_mask, height, width, blur_radius_pixels=0.0, faces_per_pixel=8, bin_size=-1, max_faces_per_bin=-1,
              perspective_correct=False, clip_barycentric=False, cull_back_faces=True, parallel=True):
    face_ndc = _f32(face_ndc)
    mask = None if clip_mask is None else np.ascontiguousarray(clip_mask, dtype=np.uint8)
    K = faces_per_pixel
    idx = np.empty((height, width, K), np.int64)
    depth = np.empty((height, width, K), np.float32)
    bary = np.empty((height, width, K, 3), np.float32)
    dist = np.empty((height, width, K), np.float32)
    dropped = lib().oracle_rasterize_ndc_triangles(
        _p(face_ndc), _p(mask), C.c_int64(face_ndc.shape[0]), C.c_int(height), C.c_int(width), C.c_float(blur_radius_pixels),
        C.c_int(K), C.c_int(bin_size), C.c_int(max_faces_per_bin), C.c_int(perspective_correct), C.c_int(clip_barycentric),
        C.c_int(cull_back_faces), C.c_int(parallel), _p(idx), _p(depth), _p(bary), _p(dist))
    return idx, depth, bary, dist, dropped


def interpolate_vertex_attributes(pixel_faces, bary, face_attributes):
    pixel_faces, bary, face_attributes = _i64(pixel_faces), _f32(bary), _f32(face_attributes)
    H, W, K = pixel_faces.shape
    Cn = face_attributes.shape[2]
    out = np.empty((H, W, K, Cn), np.float32)
    lib().oracle_interpolate_vertex_attributes(_p(pixel_faces), _p(bary), C.c_int64(H * W), C.c_int(K), _p(face_attributes), C.c_int(Cn), _p(out))
    return out


def unproject_depth(depth, K, extrinsics=None, depth_scale=1000.0, depth_max=3.0):
    depth, K = _f32(depth), _f64(K)
    ext = np.eye(4) if extrinsics is None else np.asarray(extrinsics, np.float64)
    pose = _f64(np.linalg.inv(ext))
    H, W = depth.shape[:2]
    pts = np.empty((H * W, 3), np.float32)
    mask = np.empty(H * W, np.uint8)
    lib().oracle_unproject_depth(_p(depth), C.c_int(H), C.c_int(W), _p(K), _p(pose), C.c_float(depth_scale), C.c_float(depth_max), _p(pts), _p(mask))
    return pts, mask.astype(bool)


def point_to_plane(normals1, vertices1, vertices2):
    n1, v1, v2 = _f32(normals1), _f32(vertices1), _f32(vertices2)
    out = np.empty(v1.shape[0], np.float32)
    lib().oracle_point_to_plane_distances(_p(n1), _p(v1), _p(v2), C.c_int64(v1.shape[0]), _p(out))
    return out


def warped_surface_jacobians(positions, normals, nodes, rotations, anchors, weights):
    positions, normals, nodes, rotations = _f32(positions), _f32(normals), _f32(nodes), _f32(rotations)
    anchors, weights = _i32(anchors), _f32(weights)
    V, A = anchors.shape
    jv = np.empty((V, A, 4), np.float32)
    jn = np.empty((V, A, 3), np.float32)
    lib().oracle_warped_surface_jacobians(_p(positions), _p(normals), C.c_int64(V), _p(nodes), _p(rotations), _p(anchors), _p(weights),
                                          C.c_int(A), _p(jv), _p(jn))
    return jv, jn


def rasterized_surface_jacobians(wpos, wnrm, triangles, pixel_faces, bary, ndc_K, perspective_correct=True, parallel=True):
    wpos, wnrm, triangles = _f32(wpos), _f32(wnrm), _i64(triangles)
    pixel_faces, bary, ndc_K = _i64(pixel_faces), _f32(bary), _f64(ndc_K)
    H, W, K = pixel_faces.shape
    jv = np.empty((H, W, 3, 9), np.float32)
    jn = np.empty((H, W, 3, 10), np.float32)
    lib().oracle_rasterized_surface_jacobians(_p(wpos), _p(wnrm), _p(triangles), _p(pixel_faces), _p(bary), C.c_int(H), C.c_int(W), C.c_int(K),
                                              _p(ndc_K), C.c_int(perspective_correct), C.c_int(parallel), _p(jv), _p(jn))
    return jv, jn


def associate_faces_with_anchors(triangles, anchors):
    triangles, anchors = _i64(triangles), _i32(anchors)
    F = triangles.shape[0]
    fa = np.empty((F, 3 * MAX_ANCHOR_COUNT, 7), np.int32)
    counts = np.empty(F, np.int32)
    lib().oracle_associate_faces_with_anchors(_p(triangles), C.c_int64(F), _p(anchors), C.c_int(anchors.shape[1]), _p(fa), _p(counts))
    return fa, counts


def arap_terms(edges, nodes, rotations, translations, arap_weight, arrow_base, edge_layer_indices=None, layer_radii=None,
               node_coverage_weights=None):
    """Returns (edge_residuals [3E], edge_jacobians [E,5], diag [N,6,6], wing [E,6,6], wing_coords [E,2], g_arap [6N])."""
    edges, nodes, rotations, translations = _i32(edges), _f32(nodes), _f32(rotations), _f32(translations)
    E, N = edges.shape[0], nodes.shape[0]
    eli = None if edge_layer_indices is None else np.ascontiguousarray(edge_layer_indices, np.int8)
    radii = None if layer_radii is None else _f32(layer_radii)
    ncw = None if node_coverage_weights is None else _f32(node_coverage_weights)
    res = np.empty(3 * E, np.float32)
    jac = np.empty((E, 5), np.float32)
    lib().oracle_arap_residuals(_p(edges), C.c_int64(E), _p(eli), _p(radii), _p(ncw), _p(nodes), _p(translations), _p(rotations),
                                C.c_float(arap_weight), _p(res))
    lib().oracle_arap_edge_jacobians(_p(edges), C.c_int64(E), _p(eli), _p(radii), _p(ncw), _p(nodes), _p(rotations), C.c_float(arap_weight), _p(jac))
    diag = np.empty((N, 6, 6), np.float32)
    wing = np.empty((E, 6, 6), np.float32)
    coords = np.empty((E, 2), np.int32)
    rc = lib().oracle_arap_hessian(_p(edges), C.c_int64(E), _p(jac), C.c_int(arrow_base), C.c_int(N), _p(diag), _p(wing), _p(coords))
    if rc != 0:
        raise ValueError("edge outside the defined regime i < arrow_base <= j")
    g = np.empty(6 * N, np.float32)
    lib().oracle_negative_arap_gradient(_p(res), _p(jac), _p(edges), C.c_int64(E), C.c_int(N), _p(g))
    return res, jac, diag, wing, coords, g


def solve_block_diagonal(blocks, b):
    blocks, b = _f32(blocks), _f32(b)
    n, s = blocks.shape[0], blocks.shape[1]
    x = np.empty(n * s, np.float32)
    fails = lib().oracle_solve_block_diagonal_cholesky(_p(blocks), _p(b), C.c_int(n), C.c_int(s), _p(x))
    return x, fails


def schur_complement(diag, arrow_base, wing, wing_coords, corner_upper=None, corner_upper_coords=None):
    diag, wing, wing_coords = _f32(diag), _f32(wing), _i32(wing_coords)
    N = diag.shape[0]
    n = (N - arrow_base) * 6
    cu = None if corner_upper is None else _f32(corner_upper)
    cuc = None if corner_upper_coords is None else _i32(corner_upper_coords)
    ncu = 0 if cu is None else cu.shape[0]
    S = np.empty((n, n), np.float32)
    lib().oracle_arrowhead_schur_complement(_p(diag), C.c_int(N), C.c_int(arrow_base), _p(wing), _p(wing_coords), C.c_int64(wing.shape[0]),
                                            _p(cu), _p(cuc), C.c_int64(ncu), _p(S), None, None)
    return S


def solve_arrowhead(diag, arrow_base, wing, wing_coords, b, corner_upper=None, corner_upper_coords=None):
    diag, wing, wing_coords, b = _f32(diag), _f32(wing), _i32(wing_coords), _f32(b)
    N = diag.shape[0]
    cu = None if corner_upper is None else _f32(corner_upper)
    cuc = None if corner_upper_coords is None else _i32(corner_upper_coords)
    ncu = 0 if cu is None else cu.shape[0]
    x = np.empty(6 * N, np.float32)
    fails = lib().oracle_solve_arrowhead_cholesky(_p(diag), C.c_int(N), C.c_int(arrow_base), _p(wing), _p(wing_coords), C.c_int64(wing.shape[0]),
                                                  _p(cu), _p(cuc), C.c_int64(ncu), _p(b), _p(x))
    return x, fails


def rodrigues(vectors):
    v = _f32(vectors)
    out = np.empty((v.shape[0], 3, 3), np.float32)
    lib().oracle_rodrigues(_p(v), C.c_int(v.shape[0]), _p(out))
    return out


class _FitParams(C.Structure):
    _fields_ = [
        ("vertex_positions", C.c_void_p), ("vertex_normals", C.c_void_p), ("triangles", C.c_void_p),
        ("vertex_count", C.c_int64), ("face_count", C.c_int64),
        ("nodes", C.c_void_p), ("node_count", C.c_int32), ("edges", C.c_void_p), ("edge_count", C.c_int64), ("arrow_base", C.c_int32),
        ("edge_layer_indices", C.c_void_p), ("layer_decimation_radii", C.c_void_p), ("node_coverage_weights", C.c_void_p),
        ("anchors", C.c_void_p), ("anchor_weights", C.c_void_p), ("anchor_count", C.c_int32),
        ("reference_points", C.c_void_p), ("reference_mask", C.c_void_p), ("height", C.c_int32), ("width", C.c_int32),
        ("intrinsics", C.c_void_p), ("extrinsics", C.c_void_p),
        ("iteration_count", C.c_int32), ("use_perspective_correction", C.c_int32), ("max_depth", C.c_float),
        ("use_tukey", C.c_int32), ("tukey_cutoff", C.c_float), ("levenberg_marquardt_factor", C.c_float),
        ("arap_term_weight", C.c_float), ("use_huber", C.c_int32), ("huber_constant", C.c_float),
        ("max_pixels_per_node", C.c_int32), ("parallel", C.c_int32),
        ("node_rotations", C.c_void_p), ("node_translations", C.c_void_p),
        ("trace_hessian_diag", C.c_void_p), ("trace_negative_gradient", C.c_void_p), ("trace_deltas", C.c_void_p),
        ("last_pixel_faces", C.c_void_p), ("last_residuals", C.c_void_p), ("last_residual_mask", C.c_void_p),
        ("stage_seconds", C.c_void_p),
        ("iteration_modes", C.c_void_p), ("iteration_mode_count", C.c_int32),
    ]


def fit_to_image(positions, normals, triangles, nodes, edges, arrow_base, anchors, anchor_weights, reference_points, reference_mask,
                 height, width, K, extrinsics=None, rotations=None, translations=None, iteration_count=10,
                 use_perspective_correction=True, max_depth=10.0, use_tukey=False, tukey_cutoff=0.01, lm_factor=0.0,
                 arap_term_weight=200.0, use_huber=False, huber_constant=1e-4, edge_layer_indices=None, layer_radii=None,
                 node_coverage_weights=None, max_pixels_per_node=0, parallel=False, trace=True, iteration_modes=None):
    """The reference GN loop on the CPU. Returns dict with final rotations/translations and per-iteration traces."""
    positions, normals, triangles = _f32(positions), _f32(normals), _i64(triangles)
    nodes = _f32(nodes)
    N = nodes.shape[0]
    edges = np.zeros((0, 2), np.int32) if edges is None else _i32(edges)
    anchors, anchor_weights = _i32(anchors), _f32(anchor_weights)
    ref_pts = _f32(reference_points).reshape(-1, 3)
    ref_mask = np.ascontiguousarray(reference_mask, np.uint8).reshape(-1)
    Kd = _f64(K)
    ext = _f64(np.eye(4) if extrinsics is None else extrinsics)
    R = _f32(np.tile(np.eye(3, dtype=np.float32), (N, 1, 1)) if rotations is None else rotations).copy()
    t = _f32(np.zeros((N, 3), np.float32) if translations is None else translations).copy()
    eli = None if edge_layer_indices is None else np.ascontiguousarray(edge_layer_indices, np.int8)
    radii = None if layer_radii is None else _f32(layer_radii)
    ncw = None if node_coverage_weights is None else _f32(node_coverage_weights)
    if edges.shape[0] > 0 and eli is None and ncw is None:
        raise ValueError("edges need edge_layer_indices+layer_radii or node_coverage_weights")
    PX = height * width
    tr_H = np.zeros((iteration_count, N, 6, 6), np.float32) if trace else None
    tr_g = np.zeros((iteration_count, N, 6), np.float32) if trace else None
    tr_d = np.zeros((iteration_count, N, 6), np.float32) if trace else None
    last_faces = np.empty(PX, np.int64)
    last_res = np.empty(PX, np.float32)
    last_mask = np.empty(PX, np.uint8)
    stage = np.zeros(8, np.float64)
    modes = None if iteration_modes is None else np.ascontiguousarray([int(m) for m in iteration_modes], np.int32)  # 0 ALL, 1 TRANSLATION_ONLY, 2 ROTATION_ONLY
    p = _FitParams(
        _p(positions), _p(normals), _p(triangles), positions.shape[0], triangles.shape[0],
        _p(nodes), N, _p(edges), edges.shape[0], arrow_base, _p(eli), _p(radii), _p(ncw),
        _p(anchors), _p(anchor_weights), anchors.shape[1],
        _p(ref_pts), _p(ref_mask), height, width, _p(Kd), _p(ext),
        iteration_count, int(use_perspective_correction), max_depth, int(use_tukey), tukey_cutoff, lm_factor,
        arap_term_weight, int(use_huber), huber_constant, max_pixels_per_node, int(parallel),
        _p(R), _p(t), _p(tr_H), _p(tr_g), _p(tr_d), _p(last_faces), _p(last_res), _p(last_mask), _p(stage),
        _p(modes), 0 if modes is None else len(modes))
    rc = lib().oracle_fit_to_image(C.byref(p))
    if rc != 0:
        raise RuntimeError(f"oracle_fit_to_image failed: {rc}")
    return dict(rotations=R, translations=t, hessian_diag=tr_H, negative_gradient=tr_g, deltas=tr_d,
                pixel_faces=last_faces.reshape(height, width), residuals=last_res.reshape(height, width),
                residual_mask=last_mask.reshape(height, width).astype(bool), stage_seconds=stage)


# --- warp-field construction -----------------------------------------------------------------------
def median_grid_subsample_3d_points(points, grid_cell_size, grid_offset=0.0):
    """Indices (int64, bins in order of first occurrence) of the per-cell median points."""
    points = _f32(points)
    sample = np.empty(points.shape[0], np.int64)
    lib().oracle_median_grid_subsample_3d_points.restype = C.c_int64
    count = lib().oracle_median_grid_subsample_3d_points(_p(points), C.c_int64(points.shape[0]), C.c_float(grid_cell_size), C.c_float(grid_offset), _p(sample))
    return sample[:count].copy()


def find_k_nearest(indexed_points, query_points, k):
    indexed_points, query_points = _f32(indexed_points), _f32(query_points)
    idx = np.empty((query_points.shape[0], k), np.int32)
    dist = np.empty((query_points.shape[0], k), np.float32)
    lib().oracle_find_k_nearest(_p(indexed_points), C.c_int(indexed_points.shape[0]), _p(query_points), C.c_int64(query_points.shape[0]),
                                C.c_int(k), _p(idx), _p(dist))
    return idx, dist


def node_coverage_weights(nodes, node_coverage):
    nodes = _f32(nodes)
    out = np.empty(nodes.shape[0], np.float32)
    lib().oracle_node_coverage_weights(_p(nodes), C.c_int(nodes.shape[0]), C.c_float(node_coverage), _p(out))
    return out


def adjacency_array_to_edges(adjacency, source_indices, target_indices, flip_source_order):
    adjacency, source_indices, target_indices = _i32(adjacency), _i32(source_indices), _i32(target_indices)
    edges = np.empty((adjacency.size, 2), np.int32)
    lib().oracle_adjacency_array_to_edges.restype = C.c_int64
    count = lib().oracle_adjacency_array_to_edges(_p(adjacency), C.c_int64(adjacency.shape[0]), C.c_int(adjacency.shape[1]), _p(source_indices),
                                                  _p(target_indices), C.c_int(int(flip_source_order)), _p(edges))
    return edges[:count].copy()


def layer_decimation_radii(node_coverage, layer_count, compute_layer_decimation_radius=None):
    """radius(0) = node_coverage; radius(i) = compute_layer_decimation_radius(i, node_coverage), default (i + 1) * node_coverage in fp32
    (cpp/geometry/HierarchicalGraphWarpField.h:46-47)."""
    c = np.float32(node_coverage)
    fn = compute_layer_decimation_radius or (lambda i, cov: np.float32(i + 1) * np.float32(cov))
    return np.array([c] + [np.float32(fn(i, c)) for i in range(1, layer_count)], np.float32)


def build_regularization_layers(nodes, node_coverage, layer_count, max_vertex_degree, compute_layer_decimation_radius=None):
    """dict(node_indices [N] int64, layer_node_counts [L], edges [E,2] int32 (virtual indices), edge_layer_indices [E] int8,
    layer_decimation_radii [L])."""
    nodes = _f32(nodes)
    n = nodes.shape[0]
    radii = layer_decimation_radii(node_coverage, layer_count, compute_layer_decimation_radius)
    node_indices = np.empty(n, np.int64)
    counts = np.empty(layer_count, np.int32)
    edges = np.empty((max(n * max_vertex_degree, 1), 2), np.int32)
    eli = np.empty(max(n * max_vertex_degree, 1), np.int8)
    lib().oracle_build_regularization_layers.restype = C.c_int64
    e = lib().oracle_build_regularization_layers(_p(nodes), C.c_int(n), _p(radii), C.c_int(layer_count), C.c_int(max_vertex_degree),
                                                 _p(node_indices), _p(counts), _p(edges), _p(eli))
    if e < 0:
        raise RuntimeError(f"Attempting to generate a coarser layer (layer {-e}) out of a finer layer of the same size, which is not possible.")
    return dict(node_indices=node_indices, layer_node_counts=counts, edges=edges[:e].copy(), edge_layer_indices=eli[:e].copy(),
                layer_decimation_radii=radii)


# --- normals (cpp/geometry/functional/NormalsOperations.cpp) -------------------------------------------------------------------
def compute_triangle_normals(positions, triangles, normalized=True):
    positions, triangles = _f32(positions), _i64(triangles)
    out = np.empty((triangles.shape[0], 3), np.float32)
    lib().oracle_compute_triangle_normals(_p(positions), _p(triangles), C.c_int64(triangles.shape[0]), C.c_int(int(normalized)), _p(out))
    return out


def compute_vertex_normals(positions, triangles, normalized=True, triangle_normals=None):
    """As ComputeVertexNormals: un-normalised triangle normals are computed when the mesh has none."""
    positions, triangles = _f32(positions), _i64(triangles)
    tn = compute_triangle_normals(positions, triangles, False) if triangle_normals is None else _f32(triangle_normals)
    out = np.empty((positions.shape[0], 3), np.float32)
    lib().oracle_compute_vertex_normals(_p(triangles), C.c_int64(triangles.shape[0]), _p(tn), C.c_int64(positions.shape[0]), C.c_int(int(normalized)), _p(out))
    return out


def compute_ordered_point_cloud_normals(points, height, width):
    points = _f32(points)
    out = np.empty((height * width, 3), np.float32)
    lib().oracle_compute_ordered_point_cloud_normals(_p(points), C.c_int(height), C.c_int(width), _p(out))
    return out


def compute_anchors_shortest_path(points, nodes, edges, anchor_count, node_coverage=0.05, node_weights=None):
    """compute_anchors_and_weights_shortest_path_{fixed,variable}_node_weight; edges [N, degree] int32 adjacency (-1 = none)."""
    points, nodes, edges = _f32(points), _f32(nodes), _i32(edges)
    if nodes.shape[0] < anchor_count:
        raise RuntimeError("node count must be at least anchor_count")
    nw = None if node_weights is None else _f32(node_weights)
    anchors = np.empty((points.shape[0], anchor_count), np.int32)
    weights = np.empty((points.shape[0], anchor_count), np.float32)
    lib().oracle_compute_anchors_shortest_path(_p(points), C.c_int64(points.shape[0]), _p(nodes), C.c_int(nodes.shape[0]), _p(nw),
                                               C.c_float(node_coverage), _p(edges), C.c_int(edges.shape[1]), C.c_int(anchor_count),
                                               _p(anchors), _p(weights))
    return anchors, weights


def shade_edges_flat(pixel_face_indices, pixel_face_distances, pixel_line_width, line_color):
    faces, dist = _i64(pixel_face_indices), _f32(pixel_face_distances)
    h, w, k = faces.shape
    out = np.empty((h, w, 3), np.uint8)
    lib().oracle_shade_edges_flat(_p(faces), _p(dist), C.c_int(h), C.c_int(w), C.c_int(k), C.c_float(pixel_line_width), _p(_f32(line_color)), _p(out))
    return out


def shade_vertex_colors(pixel_face_indices, pixel_barycentric, face_vertex_colors):
    faces, bary, colors = _i64(pixel_face_indices), _f32(pixel_barycentric), _f32(face_vertex_colors)
    h, w, k = faces.shape
    out = np.empty((h, w, 3), np.uint8)
    lib().oracle_shade_vertex_colors(_p(faces), _p(bary), C.c_int(h), C.c_int(w), C.c_int(k), _p(colors), _p(out))
    return out


# --- non-rigid TSDF integration (cpp/geometry/kernel/NonRigidSurfaceVoxelBlockGridImpl.h:50-229) ---------------------------------
def integrate_non_rigid(block_indices, block_keys, tsdf, weight, voxel_color, block_resolution, voxel_size, sdf_truncation_distance, depth, color,
                        depth_normals, depth_intrinsics, color_intrinsics, extrinsics, nodes, node_rotations, node_translations, node_coverage,
                        anchor_count, minimum_valid_anchor_count, depth_scale, depth_max, node_coverage_weights=None):
    """Updates tsdf (and voxel_color) in place; returns cos_voxel_ray_to_normal [H,W]."""
    block_indices, block_keys = _i64(block_indices), _i32(block_keys)
    assert tsdf.dtype == np.float32 and tsdf.flags.c_contiguous
    weight_dtype = 0 if weight.dtype == np.float32 else 1
    depth = np.ascontiguousarray(depth)
    depth_dtype = 0 if depth.dtype == np.float32 else 1
    h, w = depth.shape
    ch = cw = 0
    color_dtype = 0
    if color is not None:
        color = np.ascontiguousarray(color)
        color_dtype = 0 if color.dtype == np.float32 else 1
        ch, cw = color.shape[:2]
    normals = _f32(depth_normals)
    nodes, R, t = _f32(nodes), _f32(node_rotations), _f32(node_translations)
    ncw = None if node_coverage_weights is None else _f32(node_coverage_weights)
    cos = np.empty((h, w), np.float32)
    lib().oracle_integrate_non_rigid(
        _p(block_indices), C.c_int64(block_indices.shape[0]), _p(block_keys), _p(tsdf), _p(weight), C.c_int(weight_dtype),
        None if voxel_color is None else _p(voxel_color), C.c_int(block_resolution), C.c_float(voxel_size), C.c_float(sdf_truncation_distance),
        _p(depth), C.c_int(depth_dtype), None if color is None else _p(color), C.c_int(color_dtype), _p(normals), C.c_int(h), C.c_int(w),
        C.c_int(ch), C.c_int(cw), _p(_f64(depth_intrinsics)), _p(_f64(color_intrinsics)), _p(_f64(extrinsics)), _p(nodes), _p(R), _p(t),
        C.c_int(nodes.shape[0]), _p(ncw), C.c_float(node_coverage), C.c_int(anchor_count), C.c_int(minimum_valid_anchor_count),
        C.c_float(depth_scale), C.c_float(depth_max), _p(cos))
    return cos


def integrate_rigid(block_indices, block_keys, tsdf, weight, voxel_color, block_resolution, voxel_size, sdf_truncation_distance, depth, color,
                    depth_intrinsics, color_intrinsics, extrinsics, depth_scale, depth_max):
    """Open3D's VoxelBlockGrid::Integrate restated (oracle_tsdf.cpp): updates tsdf, weight (and voxel_color) in place."""
    block_indices, block_keys = _i64(block_indices), _i32(block_keys)
    assert tsdf.dtype == np.float32 and tsdf.flags.c_contiguous and weight.flags.c_contiguous
    weight_dtype = 0 if weight.dtype == np.float32 else 1
    depth = np.ascontiguousarray(depth)
    depth_dtype = 0 if depth.dtype == np.float32 else 1
    h, w = depth.shape
    ch = cw = color_dtype = 0
    if color is not None:
        color = np.ascontiguousarray(color)
        color_dtype = 0 if color.dtype == np.float32 else 1
        ch, cw = color.shape[:2]
    lib().oracle_integrate_rigid(
        _p(block_indices), C.c_int64(block_indices.shape[0]), _p(block_keys), _p(tsdf), _p(weight), C.c_int(weight_dtype),
        None if voxel_color is None else _p(voxel_color), C.c_int(block_resolution), C.c_float(voxel_size), C.c_float(sdf_truncation_distance),
        _p(depth), C.c_int(depth_dtype), None if color is None else _p(color), C.c_int(color_dtype), C.c_int(h), C.c_int(w), C.c_int(ch), C.c_int(cw),
        _p(_f64(depth_intrinsics)), _p(_f64(depth_intrinsics if color_intrinsics is None else color_intrinsics)), _p(_f64(extrinsics)),
        C.c_float(depth_scale), C.c_float(depth_max))


def bounding_boxes_of_warped_blocks(block_keys, nodes, node_rotations, node_translations, node_coverage, anchor_count, minimum_valid_anchor_count,
                                    voxel_size, block_resolution, extrinsics):
    keys = _i32(block_keys)
    nodes, R, t = _f32(nodes), _f32(node_rotations), _f32(node_translations)
    boxes = np.empty((keys.shape[0], 6), np.float32)
    lib().oracle_bounding_boxes_of_warped_blocks(_p(keys), C.c_int64(keys.shape[0]), _p(nodes), _p(R), _p(t), C.c_int(nodes.shape[0]), C.c_float(node_coverage),
                                                 C.c_int(anchor_count), C.c_int(minimum_valid_anchor_count), C.c_float(voxel_size), C.c_int(block_resolution),
                                                 _p(_f64(extrinsics)), _p(boxes))
    return boxes


def boxes_intersecting_surface_mask(boxes, intrinsics, depth, depth_scale, depth_max, stride, truncation_distance):
    boxes = _f32(boxes)
    depth = np.ascontiguousarray(depth)
    mask = np.empty(boxes.shape[0], np.uint8)
    lib().oracle_boxes_intersecting_surface_mask(_p(boxes), C.c_int64(boxes.shape[0]), _p(_f64(intrinsics)), _p(depth), C.c_int(0 if depth.dtype == np.float32 else 1),
                                                 C.c_int(depth.shape[0]), C.c_int(depth.shape[1]), C.c_float(depth_scale), C.c_float(depth_max), C.c_int(stride),
                                                 C.c_float(truncation_distance), _p(mask))
    return mask.astype(bool)


def mean_grid_downsample_3d_points(points, grid_cell_size, grid_offset=0.0):
    points = _f32(points)
    out = np.empty((points.shape[0], 3), np.float32)
    lib().oracle_grid_mean.restype = C.c_int64
    bins = lib().oracle_grid_mean(_p(points), C.c_int64(points.shape[0]), C.c_float(grid_cell_size), C.c_float(grid_offset), C.c_int(0), _p(out), None)
    return out[:bins].copy()


def closest_to_mean_grid_subsample_3d_points(points, grid_cell_size):
    points = _f32(points)
    out = np.empty(points.shape[0], np.int64)
    lib().oracle_grid_mean.restype = C.c_int64
    bins = lib().oracle_grid_mean(_p(points), C.c_int64(points.shape[0]), C.c_float(grid_cell_size), C.c_float(0.0), C.c_int(1), None, _p(out))
    return out[:bins].copy()


def _extended_cell(radius):
    r = np.float32(radius)
    return float(np.sqrt(np.float32(2) * (r * r), dtype=np.float32) * np.float32(2))


def fast_mean_radius_downsample_3d_points(points, radius):
    """cpp/geometry/functional/kernel/GeometrySamplingImpl.h:47-59"""
    cell = _extended_cell(radius)
    return mean_grid_downsample_3d_points(mean_grid_downsample_3d_points(points, cell), cell, 0.5)


def fast_median_radius_subsample_3d_points(points, radius):
    """cpp/geometry/functional/kernel/GeometrySamplingImpl.h:78-91"""
    cell = _extended_cell(radius)
    points = _f32(points)
    stage1 = median_grid_subsample_3d_points(points, cell)
    stage2 = median_grid_subsample_3d_points(points[stage1], cell, 0.5)
    return stage1[stage2]


# --- un-fused assembly stages ------------------------------------------------------------------------
def pixel_vertex_anchor_jacobians(rasterized_vertex_position_jacobians, rasterized_vertex_normal_jacobians, warped_vertex_position_jacobians,
                                  warped_vertex_normal_jacobians, point_map_vectors, rasterized_normals, residual_mask, pixel_faces,
                                  face_anchors, face_anchor_counts, node_count, use_tukey=False, tukey_cutoff=0.01, max_pixels_per_node=0):
    """PixelVertexAnchorJacobiansAndNodeAssociations (cap lifted unless max_pixels_per_node > 0). Returns
    (pixel_jacobians [PX,3A,6], pixel_node_jacobian_counts [PX], node_pixel_lists [N,cap], node_pixel_counts [N], dropped)."""
    rv, rn = _f32(rasterized_vertex_position_jacobians), _f32(rasterized_vertex_normal_jacobians)
    wv, wn = _f32(warped_vertex_position_jacobians), _f32(warped_vertex_normal_jacobians)
    pm, nl = _f32(point_map_vectors).reshape(-1, 3), _f32(rasterized_normals).reshape(-1, 3)
    mask = np.ascontiguousarray(residual_mask, np.uint8).reshape(-1)
    pf = _i64(pixel_faces).reshape(mask.shape[0], -1)
    fa, fc = _i32(face_anchors), _i32(face_anchor_counts)
    PX, A, K = mask.shape[0], wv.shape[1], pf.shape[1]
    counts = np.zeros(node_count, np.int64)
    for px in np.nonzero(mask)[0]:
        f = pf[px, 0]
        for k in range(fc[f]):
            counts[fa[f, k, 0]] += 1
    cap = int(max(1, counts.max() if node_count else 1, max_pixels_per_node + 1 if max_pixels_per_node > 0 else 1))
    pj = np.empty((PX, 3 * A, 6), np.float32)
    pc = np.empty(PX, np.int32)
    lists = np.empty((node_count, cap), np.int32)
    nc = np.empty(node_count, np.int32)
    dropped = lib().oracle_pixel_vertex_anchor_jacobians(_p(rv), _p(rn), _p(wv), _p(wn), _p(pm), _p(nl), _p(mask), _p(pf), C.c_int(K), _p(fa), _p(fc),
                                                         C.c_int64(PX), C.c_int(A), C.c_int(node_count), C.c_int(use_tukey), C.c_float(tukey_cutoff),
                                                         C.c_int64(cap), C.c_int(max_pixels_per_node), _p(pj), _p(pc), _p(lists), _p(nc))
    return pj, pc, lists, nc, int(dropped)


def depth_hessian_and_gradient(pixel_jacobians, node_pixel_lists, node_pixel_counts, residuals, residual_mask, anchor_count):
    """ComputeDepthHessianApproximationBlocks_UnorderedNodePixels + ComputeNegativeDepthGradient_UnorderedNodePixels."""
    pj, lists, nc = _f32(pixel_jacobians), _i32(node_pixel_lists), _i32(node_pixel_counts)
    N, cap = lists.shape
    H = np.empty((N, 6, 6), np.float32)
    g = np.empty(N * 6, np.float32)
    res = _f32(residuals).reshape(-1)
    mask = np.ascontiguousarray(residual_mask, np.uint8).reshape(-1)
    lib().oracle_depth_hessian_blocks(_p(pj), _p(lists), _p(nc), C.c_int(N), C.c_int64(cap), C.c_int(0), _p(H))
    lib().oracle_negative_depth_gradient(_p(res), _p(mask), _p(pj), _p(lists), _p(nc), C.c_int(N), C.c_int64(cap), C.c_int(anchor_count), C.c_int(0), _p(g))
    return H, g


# --- generic block-sparse routines (oracle_blocks.cpp) ----------------------------------------------------------------------------
def matmul_block_sparse_row_wise(a_blocks, b_blocks, b_coordinates, padded=False):
    a, b, bc = _f32(a_blocks), _f32(b_blocks), _i32(b_coordinates)
    s = a.shape[1]
    c = np.empty_like(b)
    lib().oracle_matmul_block_sparse_row_wise_padded(_p(a), C.c_int(a.shape[0]), _p(b), _p(bc), C.c_int(b.shape[0]), C.c_int(s), _p(c))
    if padded:
        return c
    keep = bc[:, 0] < a.shape[0]
    return c[keep], bc[keep]


def matmul_block_sparse(a_blocks, a_breadboard, transpose_a, b_blocks, b_breadboard, transpose_b):
    a, b = _f32(a_blocks), _f32(b_blocks)
    ab, bb = np.ascontiguousarray(a_breadboard, np.int16), np.ascontiguousarray(b_breadboard, np.int16)
    s = a.shape[1]
    out_rows = ab.shape[1] if transpose_a else ab.shape[0]
    out_cols = bb.shape[0] if transpose_b else bb.shape[1]
    blocks = np.empty((out_rows * out_cols, s, s), np.float32)
    mask = np.empty(out_rows * out_cols, np.uint8)
    lib().oracle_matmul_block_sparse(_p(a), _p(ab), C.c_int(ab.shape[0]), C.c_int(ab.shape[1]), C.c_int(transpose_a), _p(b), _p(bb),
                                     C.c_int(bb.shape[0]), C.c_int(bb.shape[1]), C.c_int(transpose_b), C.c_int(s), _p(blocks), _p(mask))
    keep = mask.astype(bool)
    coords = np.stack(np.divmod(np.arange(out_rows * out_cols), out_cols), 1).astype(np.int32)
    return blocks[keep], coords[keep]


def block_sparse_and_vector_product(blocks, m, coordinates, offset, transpose, vector_b):
    a, co, b = _f32(blocks), _i32(coordinates), _f32(vector_b)
    c = np.empty(m, np.float32)
    lib().oracle_block_sparse_and_vector_product(_p(a), _p(co), C.c_int(a.shape[0]), C.c_int(a.shape[1]), C.c_int(m), C.c_int(transpose),
                                                 C.c_int(offset[0]), C.c_int(offset[1]), _p(b), _p(c))
    return c


def diagonal_block_sparse_and_vector_product(blocks, vector_b):
    a, b = _f32(blocks), _f32(vector_b)
    c = np.empty_like(b)
    lib().oracle_diagonal_block_sparse_and_vector_product(_p(a), C.c_int(a.shape[0]), C.c_int(a.shape[1]), _p(b), _p(c))
    return c


def invert_positive_semidefinite_blocks(blocks):
    a = _f32(blocks)
    out = np.zeros_like(a)
    fails = lib().oracle_invert_positive_semidefinite_blocks(_p(a), C.c_int(a.shape[0]), C.c_int(a.shape[1]), _p(out))
    return out, int(fails)


def invert_triangular_blocks(blocks, upper):
    a = _f32(blocks)
    out = np.empty_like(a)
    lib().oracle_invert_triangular_blocks(_p(a), C.c_int(a.shape[0]), C.c_int(a.shape[1]), C.c_int(upper), _p(out))
    return out


def transpose_blocks(blocks):
    a = _f32(blocks).copy()
    lib().oracle_transpose_blocks_in_place(_p(a), C.c_int(a.shape[0]), C.c_int(a.shape[1]))
    return a


def fill_in_diagonal_blocks(matrix, blocks):
    m, b = _f32(matrix).copy(), _f32(blocks)
    lib().oracle_fill_in_diagonal_blocks(_p(m), C.c_int(m.shape[0]), _p(b), C.c_int(b.shape[0]), C.c_int(b.shape[1]))
    return m


def get_diagonal_blocks(matrix, block_size):
    m = _f32(matrix)
    out = np.empty((m.shape[0] // block_size, block_size, block_size), np.float32)
    lib().oracle_get_diagonal_blocks(_p(m), C.c_int(m.shape[0]), C.c_int(block_size), _p(out))
    return out


def fill_in_sparse_blocks(matrix, blocks, coordinates, offset=(0, 0), transpose=False):
    m, b, co = _f32(matrix).copy(), _f32(blocks), _i32(coordinates)
    lib().oracle_fill_in_sparse_blocks(_p(m), C.c_int(m.shape[0]), C.c_int(m.shape[1]), _p(b), _p(co), C.c_int(b.shape[0]), C.c_int(b.shape[1]),
                                       C.c_int(offset[0]), C.c_int(offset[1]), C.c_int(transpose))
    return m


def get_sparse_blocks(matrix, block_size, coordinates):
    m, co = _f32(matrix), _i32(coordinates)
    out = np.empty((co.shape[0], block_size, block_size), np.float32)
    lib().oracle_get_sparse_blocks(_p(m), C.c_int(m.shape[1]), C.c_int(block_size), _p(co), C.c_int(co.shape[0]), _p(out))
    return out


def voxel_block_neighbors(block_keys):
    """VoxelBlockGrid::BufferRadiusNeighbors: [B, 27] indices of the blocks around every block, -1 where absent."""
    keys = np.ascontiguousarray(block_keys, np.int32)
    out = np.empty((keys.shape[0], 27), np.int32)
    lib().oracle_voxel_block_neighbors(_p(keys), C.c_int64(keys.shape[0]), _p(out))
    return out


def extract_triangle_mesh(block_keys, tsdf, weight, color=None, voxel_size=0.01, weight_threshold=3.0):
    """Open3D's ExtractTriangleMesh over voxel blocks [B,R,R,R]: (vertices, normals, colors or None, triangles)."""
    keys, tsdf, weight = np.ascontiguousarray(block_keys, np.int32), _f32(tsdf), _f32(weight)
    R = tsdf.shape[1]
    col = None if color is None else _f32(color)
    counts = np.zeros(2, np.int64)
    null = C.c_void_p(None)
    args = (_p(keys), C.c_int64(keys.shape[0]), _p(tsdf), _p(weight), null if col is None else _p(col), C.c_int(R), C.c_float(voxel_size), C.c_float(weight_threshold))
    lib().oracle_extract_triangle_mesh(*args, _p(counts), null, null, null, null)
    vertices, normals = np.empty((counts[0], 3), np.float32), np.empty((counts[0], 3), np.float32)
    colors = None if col is None else np.empty((counts[0], 3), np.float32)
    triangles = np.empty((counts[1], 3), np.int64)
    lib().oracle_extract_triangle_mesh(*args, _p(counts), _p(vertices), _p(normals), null if colors is None else _p(colors), _p(triangles))
    return vertices, normals, colors, triangles
