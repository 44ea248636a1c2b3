"""GPU parity of the warp-field construction against the CPU oracle: every output is index work and must be
bit-exact — sampled indices and their order, neighbour slots and distances, virtual node order, edges, edge layers."""
import os

import numpy as np
import pytest
import torch

from test_warp_field_build import GOLDEN, check_hierarchy_against_kat, check_median_against_kat
from util import oracle_fit, rel_close
from dynamicfuion_python_b200 import synthetic

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


def test_median_grid_subsample_kat(nnrt):
    kat = np.load(GOLDEN)
    sample = nnrt.geometry.functional.median_grid_subsample_3d_points(kat["median/points"], float(kat["median/cell_size"]))
    assert sample.dtype == torch.int64
    check_median_against_kat(kat, _np(sample))


@pytest.mark.parametrize("count,extent,cell", [(1, 1.0, 0.3), (10, 1.0, 0.5), (1000, 1.0, 0.25), (5000, 1.0, 0.013), (3000, 0.05, 1.0),
                                               (60000, 2.0, 0.11), (300000, 3.0, 0.05), (40000, 0.05, 1.0)])
def test_median_grid_subsample_bit_exact(nnrt, oracle, count, extent, cell):
    """Singletons (cell much smaller than the spacing), eight bins of ~375 points each (3000 points around the origin, unit cells), negative coordinates,
    sizes across several scan tiles, eight bins of ~5000 points (beyond the shared-memory sort's capacity)."""
    rng = np.random.default_rng(count)
    pts = rng.uniform(-extent, extent, (count, 3)).astype(np.float32)
    got = _np(nnrt.geometry.functional.median_grid_subsample_3d_points(pts, cell))
    want = oracle.median_grid_subsample_3d_points(pts, cell)
    assert np.array_equal(got, want)


def test_median_grid_subsample_duplicates_and_ties(nnrt, oracle):
    rng = np.random.default_rng(5)
    base = rng.uniform(0, 1, (200, 3)).astype(np.float32)
    pts = np.concatenate([base, base, base[:50]], 0)  # exact duplicates: equal distance sums, the smallest index must win
    got = _np(nnrt.geometry.functional.median_grid_subsample_3d_points(pts, 0.2))
    assert np.array_equal(got, oracle.median_grid_subsample_3d_points(pts, 0.2))
    assert got.max() < 200
    empty = nnrt.geometry.functional.median_grid_subsample_3d_points(np.zeros((0, 3), np.float32), 0.2)
    assert empty.shape[0] == 0
    with pytest.raises(RuntimeError):
        nnrt.geometry.functional.median_grid_subsample_3d_points(pts, 0.0)


@pytest.mark.parametrize("indexed_count,k", [(1, 1), (3, 4), (250, 4), (2500, 8)])
def test_k_nearest_bit_exact(nnrt, oracle, indexed_count, k):
    rng = np.random.default_rng(indexed_count)
    indexed = rng.uniform(-1, 1, (indexed_count, 3)).astype(np.float32)
    queries = rng.uniform(-1, 1, (777, 3)).astype(np.float32)
    tree = nnrt.core.KDTree(indexed)
    idx, dist = tree.find_k_nearest_to_points(queries, k)
    idx_o, dist_o = oracle.find_k_nearest(indexed, queries, k)
    assert np.array_equal(_np(idx), idx_o) and np.array_equal(_np(dist), dist_o)  # slot order included
    idx_s, dist_s = tree.find_k_nearest_to_points(queries, k, sort_output=True)
    order = np.argsort(dist_o, axis=1, kind="stable")
    assert np.array_equal(_np(dist_s), np.take_along_axis(dist_o, order, 1))
    assert np.array_equal(_np(idx_s), np.take_along_axis(idx_o, order, 1))
    with pytest.raises(RuntimeError):
        tree.find_k_nearest_to_points(queries, 9)


def test_node_coverage_weights_bit_exact(nnrt, oracle):
    rng = np.random.default_rng(8)
    nodes = rng.uniform(-0.3, 0.3, (1500, 3)).astype(np.float32)
    wf = nnrt.geometry.GraphWarpField(nodes, 0.05)  # MINIMAL_K_NEIGHBOR_NODE_DISTANCE by default
    assert np.array_equal(_np(wf.node_coverage_weights), oracle.node_coverage_weights(nodes, 0.05))
    one = nnrt.geometry.GraphWarpField(nodes[:1], 0.07, anchor_count=1)
    assert _np(one.node_coverage_weights)[0] == np.float32(0.07)


def test_hierarchical_constructor_reference_kat(nnrt):
    """cpp/tests/test_graph_warp_field.cpp:33-339 through the reference's constructor signature."""
    kat = np.load(GOLDEN)
    G = nnrt.geometry
    wf = G.HierarchicalGraphWarpField(kat["hgwf/nodes"], 0.25, False, 4, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 3, 4,
                                      lambda i_layer, node_coverage: np.float32(2.0 ** i_layer) * node_coverage)
    check_hierarchy_against_kat(kat, _np(wf.get_virtual_node_indices()), wf.layer_node_counts, _np(wf.get_edges()))
    assert wf.get_regularization_level_count() == 3
    for i in range(3):
        level = wf.get_regularization_level(i)
        assert np.array_equal(np.sort(_np(level.node_indices)), np.sort(kat[f"hgwf/layer_{i}"]))
        assert np.array_equal(_np(level.node_positions), kat["hgwf/nodes"][_np(level.node_indices)])
    assert np.array_equal(_np(wf.get_node_positions()), kat["hgwf/nodes"])  # original order by default
    assert np.array_equal(_np(wf.get_node_positions(True)), kat["hgwf/nodes"][_np(wf.node_indices)])
    assert np.allclose(_np(wf.get_layer_decimation_radii()), [0.25, 0.5, 1.0])


@pytest.mark.parametrize("node_count,coverage,layers,degree", [(250, 0.02, 2, 4), (2000, 0.01, 3, 4), (40, 0.05, 2, 8), (3000, 0.004, 4, 3)])
def test_regularization_layers_bit_exact(nnrt, oracle, node_count, coverage, layers, degree):
    """Random node clouds: virtual order, per-layer counts, edges and edge layers equal the oracle's; (40 nodes, degree 8) leaves a
    coarse layer with fewer nodes than the degree (rows shortened, no -1 targets)."""
    rng = np.random.default_rng(node_count)
    nodes = rng.uniform(-0.1, 0.1, (node_count, 3)).astype(np.float32)
    want = oracle.build_regularization_layers(nodes, coverage, layers, degree)
    G = nnrt.geometry
    wf = G.HierarchicalGraphWarpField(nodes, coverage, False, 4, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, layers, degree)
    assert wf.layer_node_counts == list(want["layer_node_counts"])
    assert np.array_equal(_np(wf.node_indices), want["node_indices"])
    assert np.array_equal(_np(wf.edges), want["edges"])
    e = len(want["edges"])
    assert np.array_equal(_np(wf.edge_layer_indices)[:e], want["edge_layer_indices"])
    assert np.array_equal(_np(wf.layer_decimation_radii), want["layer_decimation_radii"])
    assert (_np(wf.edges) >= 0).all()


def test_layer_that_cannot_be_decimated_raises(nnrt):
    nodes = np.array([[0, 0, 0], [10, 0, 0], [0, 10, 0], [0, 0, 10]], np.float32)
    with pytest.raises(RuntimeError, match="coarser layer"):
        nnrt.geometry.HierarchicalGraphWarpField(nodes, 0.05, layer_count=2)
    with pytest.raises(RuntimeError, match="layer_count"):
        nnrt.geometry.HierarchicalGraphWarpField(nodes, 0.05, layer_count=0)


def test_orders_are_consistent(nnrt):
    """Original-order accessors and virtual-order storage describe the same field; anchors computed in either order warp alike."""
    s = synthetic.make_scene("tinyr", 120, 160)
    rng = np.random.default_rng(2)
    nodes = s["positions"][rng.choice(len(s["positions"]), 60, replace=False)]
    G = nnrt.geometry
    wf = G.HierarchicalGraphWarpField(nodes, 0.02, False, 4, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 2, 3)
    assert not wf._identity_order
    t = rng.normal(0, 0.01, (60, 3)).astype(np.float32)
    wf.set_node_translations(t)
    assert np.array_equal(_np(wf.get_node_translations()), t)
    assert np.array_equal(_np(wf.get_node_translations(True)), t[_np(wf.node_indices)])
    wf.translate_nodes(t)
    assert np.array_equal(_np(wf.get_node_translations()), t + t)
    mesh = (s["positions"], s["normals"], s["triangles"])
    a_orig, w_orig = wf.precompute_anchors_and_weights(mesh)
    a_virt, w_virt = wf.precompute_anchors_and_weights(mesh, use_virtual_ordering=True)
    assert np.array_equal(_np(a_orig), _np(wf.node_indices)[_np(a_virt)])
    flat = G.GraphWarpField(nodes, 0.02, False, 4, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    flat.set_node_translations(t + t)
    a_flat, w_flat = flat.precompute_anchors_and_weights(mesh)
    want = flat.warp_mesh(mesh, a_flat, w_flat)
    got = wf.warp_mesh(mesh, a_orig, w_orig)
    got_virtual = wf.warp_mesh(mesh, a_virt, w_virt, use_virtual_ordering=True)
    assert torch.allclose(got.positions, want.positions, atol=1e-6) and torch.equal(got.positions, got_virtual.positions)


def test_reference_ordering_quirks_flag(nnrt):
    """with `reference_ordering_quirks` set, the class reproduces the reference's inconsistencies between its two node
    orders (HierarchicalGraphWarpField.cpp:222-224, 249-253, 271-277; WarpField.cpp:129-144), each checked here against the formula
    the reference's code evaluates, written in numpy on original-order arrays; without the flag both orders describe one field."""
    s = synthetic.make_scene("tinyr", 120, 160)
    rng = np.random.default_rng(5)
    nodes = s["positions"][rng.choice(len(s["positions"]), 60, replace=False)]
    G = nnrt.geometry
    method = G.WarpNodeCoverageComputationMethod
    wf = G.HierarchicalGraphWarpField(nodes, 0.02, False, 4, 0, method.MINIMAL_K_NEIGHBOR_NODE_DISTANCE, 2, 3)
    assert not wf._identity_order and not wf.reference_ordering_quirks
    index = _np(wf.node_indices)  # virtual -> original
    rotations, _ = synthetic.smooth_node_motion(nodes, 20.0, 0.0, np.random.default_rng(1))
    deltas, _ = synthetic.smooth_node_motion(nodes, 15.0, 0.0, np.random.default_rng(2))
    rotations, deltas = rotations.astype(np.float32), deltas.astype(np.float32)
    weights_original = _np(wf.get_node_coverage_weights()).copy()
    mesh = (s["positions"], s["normals"], s["triangles"])

    # consistent behaviour (flag off): rotating in virtual order == rotating in original order with the permuted deltas
    wf.set_node_rotations(rotations)
    wf.rotate_nodes(deltas[index], use_virtual_ordering=True)
    assert np.allclose(_np(wf.get_node_rotations()), rotations @ deltas, atol=1e-6)

    wf.reference_ordering_quirks = True
    # (1) RotateNodes(deltas, true): new[i] = R_original[i] deltas[i], then R_original[index[i]] = new[i]
    wf.set_node_rotations(rotations)
    virtual_deltas = deltas[index]
    wf.rotate_nodes(virtual_deltas, use_virtual_ordering=True)
    want = np.empty_like(rotations)
    want[index] = rotations @ virtual_deltas
    assert np.allclose(_np(wf.get_node_rotations()), want, atol=1e-6)
    assert not np.allclose(want, rotations @ deltas, atol=1e-3)  # the quirk is visible on this field
    # original-order rotation stays consistent
    wf.set_node_rotations(rotations)
    wf.rotate_nodes(deltas)
    assert np.allclose(_np(wf.get_node_rotations()), rotations @ deltas, atol=1e-6)
    # (2) GetNodeCoverageWeights(false) returns the rotations
    assert np.array_equal(_np(wf.get_node_coverage_weights()), _np(wf.get_node_rotations()))
    assert np.array_equal(_np(wf.get_node_coverage_weights(True)), weights_original[index])
    # (3) PrecomputeAnchorsAndWeights(mesh, true), variable coverage: virtual-order positions with ORIGINAL-order weights
    a_quirk, w_quirk = wf.precompute_anchors_and_weights(mesh, use_virtual_ordering=True)
    a_want, w_want = G.functional.compute_anchors_and_weights_euclidean_variable_node_weight(s["positions"], nodes[index], weights_original, 4, 0)
    assert np.array_equal(_np(a_quirk), _np(a_want)) and np.array_equal(_np(w_quirk), _np(w_want))
    a_orig, w_orig = wf.precompute_anchors_and_weights(mesh)  # original order: consistent in the reference too
    a_flat, w_flat = G.functional.compute_anchors_and_weights_euclidean_variable_node_weight(s["positions"], nodes, weights_original, 4, 0)
    assert np.array_equal(_np(a_orig), _np(a_flat)) and np.allclose(_np(w_orig), _np(w_flat), atol=1e-7)
    # (4) WarpMesh reads anchors as original-order indices whatever order they were computed in
    wf.set_node_rotations(rotations)
    t = rng.normal(0, 0.01, (60, 3)).astype(np.float32)
    wf.set_node_translations(t)
    got = wf.warp_mesh(mesh, a_quirk, w_quirk, use_virtual_ordering=True)
    want_mesh = G.functional.warp_triangle_mesh(mesh, nodes, rotations, t, a_quirk, w_quirk)
    assert torch.allclose(got.positions, want_mesh.positions, atol=1e-6)
    wf.reference_ordering_quirks = False
    consistent = wf.warp_mesh(mesh, a_quirk, w_quirk, use_virtual_ordering=True)
    assert not torch.allclose(consistent.positions, want_mesh.positions, atol=1e-4)


def test_fit_with_constructed_warp_field_matches_oracle(nnrt, oracle):
    """End to end: nodes sampled from the mesh, layers + edges built on the device by the constructor, anchors computed by the fitter
    in virtual order, one GN iteration (the 1e-3 per-iteration delta bar) — against the oracle running on the oracle-built hierarchy."""
    H, W = 120, 160
    s = synthetic.make_scene("tinyr", H, W)
    coverage = 0.02
    nodes = s["positions"][oracle.median_grid_subsample_3d_points(s["positions"], 0.045)]
    built = oracle.build_regularization_layers(nodes, coverage, 2, 3)
    virtual_nodes = nodes[built["node_indices"]]
    n, e = len(nodes), len(built["edges"])
    rotations_gt, translations_gt = synthetic.smooth_node_motion(virtual_nodes, 3.0, 0.005, np.random.default_rng(7))
    anchors, weights = oracle.compute_anchors(s["positions"], virtual_nodes, 4, 0, coverage)
    wp, wn = oracle.warp(s["positions"], s["normals"], virtual_nodes, rotations_gt, translations_gt, anchors, weights)
    fv, mask = oracle.face_vertices_ndc(wp, s["triangles"], s["K"], H, W, 0.0, 10.0)
    _, depth, _, _, _ = oracle.rasterize(fv, mask, H, W, 0.0, 1, perspective_correct=True)
    ref_pts, ref_mask = oracle.unproject_depth(depth[..., 0], s["K"], None, 1.0, 10.0)
    eli = np.zeros(max(e, n), np.int8)
    eli[:e] = built["edge_layer_indices"]

    G = nnrt.geometry
    wf = G.HierarchicalGraphWarpField(nodes, coverage, False, 4, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 2, 3)
    assert np.array_equal(_np(wf.node_indices), built["node_indices"]) and np.array_equal(_np(wf.edges), built["edges"])
    # The fitter computes its own anchors (virtual order). Their indices equal the oracle's; the weights differ in the last bit (expf),
    # and one ulp of weight moves warped vertices enough to flip coverage of edge pixels, so the oracle fit is given the SAME weights.
    gpu_anchors, gpu_weights = wf.precompute_anchors_and_weights((s["positions"], s["normals"], s["triangles"]), use_virtual_ordering=True)
    assert np.array_equal(_np(gpu_anchors), anchors)
    np.testing.assert_allclose(_np(gpu_weights), weights, rtol=2e-5, atol=1e-7)
    scene = dict(s, nodes=virtual_nodes, edges=built["edges"], arrow_base=int(built["layer_node_counts"][0]), anchors=anchors,
                 anchor_weights=_np(gpu_weights), reference_points=ref_pts, reference_mask=ref_mask, edge_layer_indices=eli,
                 layer_radii=built["layer_decimation_radii"])
    ref = oracle_fit(oracle, scene, 1)
    fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=s["lm_factor"],
                                                        arap_term_weight=s["arap_term_weight"])
    fitter.fit_to_image(wf, (s["positions"], s["normals"], s["triangles"]), None, ref_pts, ref_mask, s["K"], np.eye(4), (H, W))
    fitter.check_status()
    t_virtual = _np(wf.get_node_translations(True))
    deltas = _np(fitter.diagnostics()["deltas"])
    ok, worst = rel_close(deltas, ref["deltas"][0], 1e-3, floor_scale=1e-2)  # same gates as test_gpu_fitter's single-iteration parity
    assert ok, f"deltas {worst}"
    assert np.abs(deltas - ref["deltas"][0]).max() <= 1e-3 * np.abs(ref["deltas"][0]).max()
    np.testing.assert_allclose(t_virtual, ref["translations"], rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose(_np(wf.get_node_rotations(True)), ref["rotations"], rtol=1e-3, atol=1e-6)
    assert np.array_equal(_np(wf.get_node_translations())[built["node_indices"]], t_virtual)


@pytest.mark.parametrize("count,extent,cell", [(30, 5.0, 5.0), (2000, 1.0, 0.2), (100000, 2.0, 0.07)])
def test_mean_grid_ops_bit_exact(nnrt, oracle, count, extent, cell):
    """mean_grid_downsample_3d_points / closest_to_mean_grid_subsample_3d_points: the reference's KATs, and random clouds against the
    oracle — means summed in ascending point order on both sides, so bit-exact."""
    kat = np.load(GOLDEN)
    f = nnrt.geometry.functional
    if count == 30:
        means = _np(f.mean_grid_downsample_3d_points(kat["mean/points"], float(kat["mean/cell_size"])))
        np.testing.assert_allclose(means[np.argsort(means[:, 0])], kat["mean/points_gt_sorted_by_x"], atol=1e-6)
        assert np.array_equal(_np(f.closest_to_mean_grid_subsample_3d_points(kat["closest/points"], float(kat["closest/cell_size"]))), kat["closest/sample_gt"])
    rng = np.random.default_rng(count)
    pts = rng.uniform(-extent, extent, (count, 3)).astype(np.float32)
    assert np.array_equal(_np(f.mean_grid_downsample_3d_points(pts, cell)), oracle.mean_grid_downsample_3d_points(pts, cell))
    assert np.array_equal(_np(f.closest_to_mean_grid_subsample_3d_points(pts, cell)), oracle.closest_to_mean_grid_subsample_3d_points(pts, cell))


def test_fast_radius_variants_bit_exact(nnrt, oracle):
    """fast_mean_radius_downsample_3d_points / fast_median_radius_subsample_3d_points (two grid passes, the second shifted by half a
    cell): the reference's cloud and property (cpp/tests/test_sampling.cpp:131-141), and random clouds against the oracle."""
    kat = np.load(GOLDEN)
    f = nnrt.geometry.functional
    rng = np.random.default_rng(17)
    for pts, radius in ((kat["radius/points"], float(kat["radius/radius"])), (rng.uniform(-1, 1, (20000, 3)).astype(np.float32), 0.05),
                        (rng.normal(0, 0.3, (5000, 3)).astype(np.float32), 0.011)):
        means = _np(f.fast_mean_radius_downsample_3d_points(pts, radius))
        assert np.array_equal(means, oracle.fast_mean_radius_downsample_3d_points(pts, radius))
        sample = _np(f.fast_median_radius_subsample_3d_points(pts, radius))
        assert np.array_equal(sample, oracle.fast_median_radius_subsample_3d_points(pts, radius))
    means = _np(f.fast_mean_radius_downsample_3d_points(kat["radius/points"], 10.0))
    d = np.linalg.norm(means[:, None].astype(np.float64) - means[None].astype(np.float64), axis=2)
    np.fill_diagonal(d, np.inf)
    assert d.min() >= 10.0


def test_fitter_accepts_a_constructed_three_layer_hierarchy(nnrt, oracle):
    """A hierarchy built by the reference constructor with more than two layers has edges between coarser layers (out of bounds in
    the reference's ARAP Hessian). The fitter routes them through corner_upper blocks
    (test_gpu_fitter.py::test_three_layer_hierarchy_uses_corner_upper_blocks checks the arithmetic): the two classes' defaults work
    together and a fit to the field's own rendering leaves it where it is."""
    s = synthetic.make_scene("tinyr", 120, 160)
    nodes = s["positions"][oracle.median_grid_subsample_3d_points(s["positions"], 0.03)]
    G = nnrt.geometry
    wf = G.HierarchicalGraphWarpField(nodes, 0.015, False, 4, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 3, 3)
    assert wf.get_regularization_level_count() == 3 and wf.layer_node_counts[2] > 0
    mesh = (s["positions"], s["normals"], s["triangles"])
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], 120, 160, 0.0, 10.0)
    _, depth, _, _, _ = oracle.rasterize(fv, mask, 120, 160, 0.0, 1, perspective_correct=True)
    depth = np.where(depth[..., 0] > 0, depth[..., 0], 0).astype(np.float32)
    fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=3, preconditioning_dampening_factor=1.0, arap_term_weight=20.0)
    fitter.fit_to_image(wf, mesh, None, depth, None, s["K"], np.eye(4), 1.0)
    fitter.check_status()
    assert bool(torch.isfinite(wf.get_node_translations()).all()) and float(wf.get_node_translations().abs().max()) < 2e-2
    with pytest.raises(RuntimeError, match="coarser layer"):  # an edge INTO the first layer is still refused
        bad = G.HierarchicalGraphWarpField.from_graph(nodes[:6], np.array([[3, 0]], np.int32), 3, 0.015, False, 4, 0,
                                                      G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4)
        fitter2 = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=1.0)
        fitter2.fit_to_image(bad, mesh, None, depth, None, s["K"], np.eye(4), 1.0)


def test_kdtree_reference_kat(nnrt):
    """cpp/tests/test_kdtree.cpp:243-276 through nnrt.core.KDTree(points).find_k_nearest_to_points(queries, 4, sort_output=True)."""
    kat = np.load(GOLDEN)
    idx, dist = nnrt.core.KDTree(kat["knn/points"]).find_k_nearest_to_points(kat["knn/queries"], 4, sort_output=True)
    assert np.array_equal(_np(idx), kat["knn/indices_gt"])
    np.testing.assert_allclose(_np(dist), kat["knn/distances_gt"], atol=1e-5)
