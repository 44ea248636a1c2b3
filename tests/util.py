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


def oracle_fit(oracle, s, iterations, **kw):
    args = dict(iteration_count=iterations, lm_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"],
                edge_layer_indices=s["edge_layer_indices"], layer_radii=s["layer_radii"], parallel=True)
    args.update(kw)
    return oracle.fit_to_image(s["positions"], s["normals"], s["triangles"], s["nodes"], s["edges"], s["arrow_base"], s["anchors"],
                               s["anchor_weights"], s["reference_points"], s["reference_mask"], s["height"], s["width"], s["K"], **args)
