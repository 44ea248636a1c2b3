"""CPU tests (no GPU): the oracle against the reference's own known-answer vectors (tests/golden/reference_kats.npz, parsed from the
reference's Catch2 tests by tests/golden/make_golden.py), against its in-repo mesh-warping fixture, and against itself (regression pin)."""
import os

import numpy as np
import pytest

from util import oracle_fit, scene_with_target

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def kats():
    return np.load(os.path.join(GOLDEN, "reference_kats.npz"))


def test_anchor_kat(oracle, kats):
    """cpp/tests/test_anchor_computation.cpp:31-84: anchors equal and weights AllClose(1e-3, 1e-6) after sorting each row descending."""
    a, w = oracle.compute_anchors(kats["anchors/vertices"], kats["anchors/nodes"], 4, 0, 0.0, kats["anchors/node_weights"])
    assert np.array_equal(-np.sort(-a, axis=1), kats["anchors/anchors_gt_sorted_desc"].astype(np.int32))
    np.testing.assert_allclose(-np.sort(-w, axis=1), kats["anchors/weights_gt_sorted_desc"], rtol=1e-3, atol=1e-6)


def test_rodrigues_kat(oracle, kats):
    """cpp/tests/test_rodrigues.cpp:30-55, AllClose(1e-3, 1e-7)."""
    np.testing.assert_allclose(oracle.rodrigues(kats["rodrigues/axis_angle_vectors"]), kats["rodrigues/expected_matrices"], rtol=1e-3, atol=1e-6)


def test_unproject_kat(oracle, kats):
    """cpp/tests/test_unproject_3d_points.cpp:25-79: uint16 depth / 1000, depth_max 3, K = [100,0,2;0,100,2]."""
    pts, mask = oracle.unproject_depth(kats["unproject/depth_u16"].astype(np.float32), kats["unproject/intrinsics"], None, 1000.0, 3.0)
    assert np.array_equal(mask, kats["unproject/mask_gt"].astype(bool))
    np.testing.assert_allclose(pts, kats["unproject/points_gt"], rtol=1e-5, atol=1e-8)


def test_mesh_warp_9_vertex_kat(oracle, kats):
    """cpp/tests/test_mesh_warping.cpp:41-89: 9-vertex xy-plane (nodes = vertices), every node rotated 90 degrees about -Y around the
    origin, anchors computed on the fly (4 anchors, coverage 0.1). The canonical plane is recovered from the expected vertices
    (warped = (0, y, x)); vertices AllClose(1e-5, 1e-5), normals AllClose(1e-5, 1e-7)."""
    gt = kats["warp9/vertices_gt"]
    plane = np.stack([gt[:, 2], gt[:, 1], np.zeros(9)], 1).astype(np.float32)
    normals = np.tile(np.array([0, 0, 1], np.float32), (9, 1))
    R = np.array([[0, 0, -1], [0, 1, 0], [1, 0, 0]], np.float32)  # AngleAxis(-90 deg, Y)
    new_pos = plane @ R.T
    anchors, weights = oracle.compute_anchors(plane, plane, 4, 0, 0.1)
    wp, wn = oracle.warp(plane, normals, plane, np.tile(R, (9, 1, 1)), (new_pos - plane).astype(np.float32), anchors, weights)
    np.testing.assert_allclose(wp, gt, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(wn, kats["warp9/normals_gt"], rtol=1e-5, atol=1e-5)


def test_block_diagonal_cholesky_kat(oracle, kats):
    """cpp/tests/test_linalg_cholesky.cpp:36-110: three 6x6 SPD blocks, two right-hand sides, AllClose default (1e-5, 1e-8) loosened to
    the fp32 conditioning of the blocks (the reference runs LAPACK potrf/trsm, the oracle a hand-rolled Cholesky)."""
    A = kats["block_cholesky/A_blocks"].reshape(3, 6, 6)
    B = kats["block_cholesky/B"].reshape(18, 2)
    X = kats["block_cholesky/expected_X"].reshape(18, 2)
    for k in range(2):
        x, fails = oracle.solve_block_diagonal(A, B[:, k])
        assert fails == 0
        np.testing.assert_allclose(x, X[:, k], rtol=2e-3, atol=2e-3)
        A64 = np.zeros((18, 18)); [A64.__setitem__((slice(6 * i, 6 * i + 6), slice(6 * i, 6 * i + 6)), A[i]) for i in range(3)]
        np.testing.assert_allclose(x, np.linalg.solve(A64, B[:, k]), rtol=2e-3, atol=2e-3)


def test_mesh_warp_fixture_219_nodes(oracle):
    """tests/test_data/mesh_warping (tests/test_graph_warp_field.py:121-166): 47 382-vertex mesh (every 8th vertex kept), 219 nodes,
    node_coverage 0.05, 4 anchors, no thresholding; expected warped vertices allclose(atol=1e-6)."""
    g = np.load(os.path.join(GOLDEN, "mesh_warp_219.npz"))
    anchors, weights = oracle.compute_anchors(g["vertices"], g["nodes"], 4, 0, float(g["node_coverage"]))
    wp, _ = oracle.warp(g["vertices"], g["normals"], g["nodes"], g["rotations"], g["translations"], anchors, weights)
    err = np.abs(wp - g["warped_gt"]).max(axis=1)
    assert np.quantile(err, 0.999) <= 1e-6 + 2e-7 * np.abs(g["warped_gt"]).max(), float(err.max())  # fp32 warp vs the float64 fixture


def test_oracle_regression_pin(oracle):
    """One full GN iteration of the single-threaded oracle on the seeded 'tinyr' scene against the committed vectors: guards the
    checker itself against accidental edits (bit-exact indices, 1e-6 on floats)."""
    g = np.load(os.path.join(GOLDEN, "oracle_fit_tinyr.npz"))
    s = scene_with_target(oracle, "tinyr", 120, 160)
    ref = oracle_fit(oracle, s, 1, parallel=False)
    assert np.array_equal(ref["pixel_faces"].astype(np.int32), g["pixel_faces"])
    np.testing.assert_allclose(ref["hessian_diag"][0], g["hessian_diag"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(ref["deltas"][0], g["deltas"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(ref["translations"], g["translations"], rtol=1e-5, atol=1e-9)


def test_oracle_parallel_equals_serial(oracle):
    """The OpenMP variant (timed CPU baseline) visits pixels in the same per-node order as the serial one: identical results."""
    s = scene_with_target(oracle, "tinyr", 120, 160)
    a, b = oracle_fit(oracle, s, 2, parallel=False), oracle_fit(oracle, s, 2, parallel=True)
    assert np.array_equal(a["pixel_faces"], b["pixel_faces"])
    np.testing.assert_allclose(a["deltas"], b["deltas"], rtol=1e-5, atol=1e-9)
