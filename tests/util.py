"""Shared helpers for the parity tests: scenes, oracle-rendered targets, tolerances."""
import numpy as np

from dynamicfuion_python_b200 import synthetic


def rel_close(a, b, rtol, floor_scale=1e-6):
    """|a-b| <= rtol * max(|a|, |b|, floor) with floor = floor_scale * ||b||_inf (SURVEY.md §8(d) parity gates)."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    floor = floor_scale * max(np.abs(b).max(), 1e-30)
    bound = rtol * np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)
    bad = np.abs(a - b) > bound
    return not bad.any(), float((np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)).max())


def scene_with_target(oracle, config, height, width, translation_scale=1.0):
    """Scene + anchors (oracle) + target reference points rendered by the oracle from the ground-truth motion."""
    s = synthetic.make_scene(config, height, width)
    s["translations_gt"] = (s["translations_gt"] * translation_scale).astype(s["translations_gt"].dtype)
    anchors, weights = oracle.compute_anchors(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    wp, wn = oracle.warp(s["positions"], s["normals"], s["nodes"], s["rotations_gt"], s["translations_gt"], anchors, weights)
    fv, mask = oracle.face_vertices_ndc(wp, s["triangles"], s["K"], height, width, 0.0, 10.0)
    idx, depth, _, _, _ = oracle.rasterize(fv, mask, height, width, 0.0, 1, perspective_correct=True)
    ref_pts, ref_mask = oracle.unproject_depth(depth[..., 0], s["K"], None, 1.0, 10.0)
    s.update(anchors=anchors, anchor_weights=weights, reference_points=ref_pts, reference_mask=ref_mask, target_depth=depth[..., 0])
    return s


def scene_with_gpu_target(nnrt, config, height, width):
    """As scene_with_target, with anchors and the target rendered by the CUDA path (already bit-exact against the oracle at test sizes:
    test_gpu_geometry / test_gpu_raster); for sizes at which the oracle's renderer takes minutes (3840 x 2160)."""
    import torch
    s = synthetic.make_scene(config, height, width)
    G, R = nnrt.geometry.functional, nnrt.rendering
    anchors, weights = G.compute_anchors_and_weights_euclidean_fixed_node_weight(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    warped = G.warp_triangle_mesh((s["positions"], s["normals"], s["triangles"]), s["nodes"], s["rotations_gt"], s["translations_gt"], anchors, weights)
    fv, mask = R.functional.get_mesh_ndc_face_vertices_and_clip_mask((warped.positions, None, s["triangles"]), s["K"], (height, width), 0.0, 10.0)
    _, depth, _, _ = R.rasterize_ndc_triangles(fv, mask, (height, width), 0.0, 1, -1, -1, True, False, True)
    depth = depth[..., 0]
    # unprojection as PerspectiveProjectionImpl.h:62-150 with depth_scale 1: X = (u - cx) d / fx, Y = (v - cy) d / fy, Z = d
    K = torch.as_tensor(s["K"], dtype=torch.float32, device=depth.device)
    v, u = torch.meshgrid(torch.arange(height, device=depth.device, dtype=torch.float32), torch.arange(width, device=depth.device, dtype=torch.float32), indexing="ij")
    valid = (depth > 0) & (depth < 10.0)
    points = torch.stack([(u - K[0, 2]) * depth / K[0, 0], (v - K[1, 2]) * depth / K[1, 1], depth], -1) * valid[..., None]
    s.update(anchors=anchors, anchor_weights=weights, reference_points=points.reshape(-1, 3).contiguous(), reference_mask=valid.reshape(-1), target_depth=depth)
    return s


def oracle_fit(oracle, s, iterations, **kw):
    args = dict(iteration_count=iterations, lm_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"],
                edge_layer_indices=s["edge_layer_indices"], layer_radii=s["layer_radii"], parallel=True)
    args.update(kw)
    return oracle.fit_to_image(s["positions"], s["normals"], s["triangles"], s["nodes"], s["edges"], s["arrow_base"], s["anchors"],
                               s["anchor_weights"], s["reference_points"], s["reference_mask"], s["height"], s["width"], s["K"], **args)


# ---- scenarios of the reference's enabled fitter tests (cpp/tests/test_deformable_mesh_fitter_{one_node,two_nodes}.cpp) -----------
# The meshes those tests read are behind a download; the expectations (tests/golden/fitter_kats.npz, parsed from the .cpp files) do not
# depend on the tessellation, so equivalent planes are regenerated here. World-to-camera: the reference flips its Blender meshes by
# 180 degrees about Y and moves them 1.2 m from the camera (mesh_transform), under which a world rotation Rx(a) becomes Rx(-a).
def rotation_x(angle_deg):
    a = np.deg2rad(angle_deg)
    return np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])


def _grid_plane(x0, x1, y0, y1, nx, ny, z):
    xs, ys = np.linspace(x0, x1, nx), np.linspace(y0, y1, ny)
    xx, yy = np.meshgrid(xs, ys)
    pos = np.stack([xx.ravel(), yy.ravel(), np.full(nx * ny, z)], 1)
    idx = np.arange(nx * ny).reshape(ny, nx)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel()
    tris = np.concatenate([np.stack([a, b, c], 1), np.stack([b, d, c], 1)], 0).astype(np.int64)
    return pos, tris


def reference_fitter_scenario(oracle, kats, scenario, angle_deg):
    """Source mesh, nodes, anchors and the reference observation of one scenario (see make_golden_r2.fitter_cases), plus the
    expected node state. Target depth is rendered by the oracle rasteriser exactly as the reference test does (blur 0.5 px for the
    translation scenario, 0 otherwise; perspective-corrected; background depth 0 and masked out)."""
    H, W = (int(v) for v in kats["fitter/one_node/image_resolution"])
    K = kats["fitter/one_node/projection_matrix"]
    blur = 0.5 if scenario == 0 else 0.0
    if scenario in (0, 1, 2):
        nodes = kats["fitter/one_node/node_positions"].astype(np.float32)
        coverage, anchor_count = float(kats["fitter/one_node/node_coverage"]), 1
        # a single quad of side 0.125 around the node: the face the reference's own single-pixel check script holds for this test
        # (math_check_scripts/dense_depth_jacobians.py:37-39: vertices (+-0.0625, +-0.0625, 1.2)); with it all nine enabled one-node
        # cases meet the reference's tolerances, larger planes do not (the undamped in-plane directions then diverge in mode ALL)
        pos, tris = _grid_plane(-0.0625, 0.0625, -0.0625, 0.0625, 2, 2, 1.2)
        g = nodes[0].astype(np.float64)
        Rt = rotation_x(-angle_deg) if scenario in (1, 2) else np.eye(3)
        tt = np.array([0, 0, 0.2]) if scenario in (0, 2) else np.zeros(3)
        target = (pos - g) @ Rt.T + g + tt
        expected_R = Rt[None]
        key = {0: "translation_case", 1: "rotation_case", 2: "combined_case"}[scenario]
        expected_t = kats["fitter/one_node/expected_translation_" + key]
        assert np.allclose(expected_t[0], tt)
    else:
        nodes = kats["fitter/two_nodes/node_positions"].astype(np.float32)
        coverage, anchor_count = float(kats["fitter/two_nodes/node_coverage"]), 2
        if scenario == 3:  # two separate planes, one per node, rotated +-angle about their node's x axis and pushed -+0.1 in z
            expected_t = kats["fitter/two_nodes/expected_translations_planes"]
            expected_R = np.stack([rotation_x(angle_deg), rotation_x(-angle_deg)])
            parts, tri_parts, targets, offset = [], [], [], 0
            for n in range(2):
                p, t = _grid_plane(-0.0625, 0.0625, nodes[n, 1] - 0.0625, nodes[n, 1] + 0.0625, 2, 2, 1.2)
                g = nodes[n].astype(np.float64)
                parts.append(p); tri_parts.append(t + offset); offset += len(p)
                targets.append((p - g) @ expected_R[n].T + g + expected_t[n])
            pos, tris, target = np.concatenate(parts), np.concatenate(tri_parts), np.concatenate(targets)
        else:  # one contiguous surface skinned to both nodes; the target is the source under the expected node motion
            expected_t = kats["fitter/two_nodes/expected_translations_surface"]
            expected_R = np.stack([np.eye(3), np.eye(3)])
            pos, tris = _grid_plane(-0.0625, 0.0625, -0.4375, 0.4375, 2, 8, 1.2)
            target = None
    pos = pos.astype(np.float32)
    nrm = np.tile(np.array([0, 0, -1], np.float32), (len(pos), 1))
    area = synthetic._front_facing(pos.astype(np.float64), tris, K, H, W)
    tris = np.where((area < 0)[:, None], tris[:, [0, 2, 1]], tris)  # front faces are clockwise in the reference's NDC
    anchors, weights = oracle.compute_anchors(pos, nodes, anchor_count, 0, coverage)
    if target is None:
        target, _ = oracle.warp(pos, nrm, nodes, expected_R.astype(np.float32), expected_t.astype(np.float32), anchors, weights)
    fv, mask = oracle.face_vertices_ndc(target.astype(np.float32), tris, K, H, W, 0.0, 10.0)
    _, depth, _, _, _ = oracle.rasterize(fv, mask, H, W, blur, 1, perspective_correct=True)
    depth = depth[..., 0].copy()
    depth[depth == -1.0] = 0.0
    ref_pts, ref_mask = oracle.unproject_depth(depth, K, None, 1.0, 10.0)
    ref_mask = np.logical_and(np.asarray(ref_mask).reshape(-1), (depth != 0).reshape(-1))
    return dict(positions=pos, normals=nrm, triangles=tris, nodes=nodes, node_coverage=coverage, anchor_count=anchor_count, anchors=anchors,
                anchor_weights=weights, depth=depth, depth_mask=depth != 0, reference_points=ref_pts, reference_mask=ref_mask, K=K, height=H, width=W,
                expected_rotations=expected_R, expected_translations=expected_t)


def reference_all_close(value, expected, atol):
    """Open3D AllClose(expected, rtol = 1, atol) exactly as the reference's fitter tests call it: |a - b| <= atol + 1 * |b|."""
    return bool(np.all(np.abs(np.asarray(value, np.float64) - expected) <= atol + np.abs(expected)))
