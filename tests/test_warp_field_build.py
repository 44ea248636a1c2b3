"""CPU suite: the oracle's warp-field construction (median-grid subsampling, layers, edges) against the reference's own known-answer
tests (cpp/tests/test_sampling.cpp:150-177, cpp/tests/test_graph_warp_field.cpp:33-339; vectors in tests/golden/warp_field_kats.npz) and
against plain numpy restatements."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "warp_field_kats.npz")


def check_hierarchy_against_kat(kat, node_indices, layer_node_counts, edges):
    """The assertions of TestHierarchicalGraphWarpFieldConstructor (layers as sets, edge sources exactly, edge node pairs sorted)."""
    offsets = np.concatenate([[0], np.cumsum(layer_node_counts)])
    assert list(layer_node_counts) == [19, 10, 4]
    for i in range(3):
        assert np.array_equal(np.sort(node_indices[offsets[i]:offsets[i + 1]]), np.sort(kat[f"hgwf/layer_{i}"]))
        assert np.array_equal(np.sort(node_indices[offsets[i]:offsets[i + 1]]), np.sort(kat["hgwf/virtual_node_indices"][offsets[i]:offsets[i + 1]]))
    assert np.array_equal(edges[:, 0], kat["hgwf/edge_sources_virtual"])
    pairs = node_indices[edges.astype(np.int64)]
    order = np.lexsort((pairs[:, 1], pairs[:, 0]))
    assert np.array_equal(pairs[order], kat["hgwf/edge_node_indices_sorted"])


def check_median_against_kat(kat, sample):
    chosen = set(int(i) for i in sample)
    assert len(chosen) == 4
    assert all(int(i) in chosen for i in kat["median/required"])
    assert all(int(a) in chosen or int(b) in chosen for a, b in kat["median/either"])


def test_oracle_median_grid_subsample_kat(oracle):
    kat = np.load(GOLDEN)
    check_median_against_kat(kat, oracle.median_grid_subsample_3d_points(kat["median/points"], float(kat["median/cell_size"])))


def test_oracle_hierarchical_constructor_kat(oracle):
    kat = np.load(GOLDEN)
    built = oracle.build_regularization_layers(kat["hgwf/nodes"], float(kat["hgwf/node_coverage"]), 3, 4, lambda i, c: np.float32(2.0 ** i) * c)
    check_hierarchy_against_kat(kat, built["node_indices"], built["layer_node_counts"], built["edges"])
    assert np.array_equal(built["edge_layer_indices"], np.repeat([2, 1], [40, 76]).astype(np.int8))
    assert np.allclose(built["layer_decimation_radii"], [0.25, 0.5, 1.0])


def test_oracle_median_matches_numpy():
    from oracle import oracle_py as oracle
    rng = np.random.default_rng(11)
    pts = rng.uniform(-1, 1, (700, 3)).astype(np.float32)
    sample = oracle.median_grid_subsample_3d_points(pts, 0.5)
    cells = np.floor(pts / np.float32(0.5)).astype(np.int32)
    _, first, inverse = np.unique(cells, axis=0, return_index=True, return_inverse=True)
    bin_order = np.argsort(first)  # bins in order of first occurrence
    assert len(sample) == len(first)
    for rank, b in enumerate(bin_order):
        members = np.flatnonzero(inverse.ravel() == b)
        d = np.linalg.norm(pts[members, None].astype(np.float64) - pts[None, members].astype(np.float64), axis=2).sum(1)
        best = d.min()
        assert sample[rank] in members
        assert d[list(members).index(sample[rank])] <= best * (1 + 1e-5) + 1e-6


def test_oracle_knn_and_coverage_weights_match_numpy(oracle):
    rng = np.random.default_rng(12)
    indexed = rng.uniform(-1, 1, (300, 3)).astype(np.float32)
    queries = rng.uniform(-1, 1, (500, 3)).astype(np.float32)
    idx, dist = oracle.find_k_nearest(indexed, queries, 4)
    d = np.linalg.norm(queries[:, None].astype(np.float64) - indexed[None].astype(np.float64), axis=2)
    expect = np.sort(np.argsort(d, axis=1)[:, :4], axis=1)
    assert np.array_equal(np.sort(idx, axis=1), expect)
    np.testing.assert_allclose(np.sort(dist, axis=1), np.sort(d, axis=1)[:, :4], rtol=1e-5)
    few, _ = oracle.find_k_nearest(indexed[:2], queries[:5], 4)
    assert np.array_equal(np.sort(few, axis=1), np.tile([-1, -1, 0, 1], (5, 1)))
    w = oracle.node_coverage_weights(indexed, 0.05)
    dd = np.linalg.norm(indexed[:, None].astype(np.float64) - indexed[None].astype(np.float64), axis=2)
    np.fill_diagonal(dd, np.inf)
    np.testing.assert_allclose(w, dd.min(1) ** 2, rtol=1e-5)
    assert oracle.node_coverage_weights(indexed[:1], 0.05)[0] == np.float32(0.05)


def test_oracle_layer_that_cannot_be_decimated_is_an_error(oracle):
    import pytest
    nodes = np.array([[0, 0, 0], [10, 0, 0], [0, 10, 0]], np.float32)  # every node alone in its cell
    with pytest.raises(RuntimeError):
        oracle.build_regularization_layers(nodes, 0.05, 2, 4)


def test_oracle_mean_grid_kats(oracle):
    """cpp/tests/test_sampling.cpp:70-118 (mean of 30 points in 5 m cells, compared after sorting by x) and :179-215 (closest to the cell
    mean: AllEqual to [2, 13], order included)."""
    kat = np.load(GOLDEN)
    means = oracle.mean_grid_downsample_3d_points(kat["mean/points"], float(kat["mean/cell_size"]))
    means = means[np.argsort(means[:, 0])]
    np.testing.assert_allclose(means, kat["mean/points_gt_sorted_by_x"], atol=1e-6)
    assert np.array_equal(oracle.closest_to_mean_grid_subsample_3d_points(kat["closest/points"], float(kat["closest/cell_size"])), kat["closest/sample_gt"])


def test_oracle_fast_radius_downsampling_reference_property(oracle):
    """cpp/tests/test_sampling.cpp:131-141 on the reference's own cloud: after FastMeanRadiusDownsample3dPoints(points, 10) no two
    points are closer than the radius."""
    kat = np.load(GOLDEN)
    out = oracle.fast_mean_radius_downsample_3d_points(kat["radius/points"], float(kat["radius/radius"]))
    d = np.linalg.norm(out[:, None].astype(np.float64) - out[None].astype(np.float64), axis=2)
    np.fill_diagonal(d, np.inf)
    assert 1 < len(out) < 100 and d.min() >= float(kat["radius/radius"])
    sample = oracle.fast_median_radius_subsample_3d_points(kat["radius/points"], float(kat["radius/radius"]))
    assert len(set(sample.tolist())) == len(sample) and 1 < len(sample) < 100


def test_oracle_knn_reference_kat(oracle):
    """cpp/tests/test_kdtree.cpp:243-276: neighbours of 3 queries among 15 points, k = 4, compared after sorting by distance."""
    kat = np.load(GOLDEN)
    idx, dist = oracle.find_k_nearest(kat["knn/points"], kat["knn/queries"], 4)
    order = np.argsort(dist, axis=1, kind="stable")
    assert np.array_equal(np.take_along_axis(idx, order, 1), kat["knn/indices_gt"])
    np.testing.assert_allclose(np.take_along_axis(dist, order, 1), kat["knn/distances_gt"], atol=1e-5)
