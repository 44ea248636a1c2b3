"""Normals (nnrt.geometry.functional.compute_{triangle,vertex,ordered_point_cloud}_normals). CPU: the oracle against the reference's
own ground truth for ordered point-cloud normals (tests/golden/normals_kat.npz, from cpp/tests/test_normals_operations.cpp:118-140) and
against float64 numpy; GPU: bit-exact against the oracle (vertex normals are summed in ascending face order on both sides)."""
import os

import numpy as np
import pytest
import torch

from dynamicfuion_python_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "normals_kat.npz")


def _np(t):
    return t.detach().cpu().numpy()


def _kat_points():
    kat = np.load(GOLDEN)
    depth, (y0, x0), K = kat["depth_u16"], kat["origin"], kat["intrinsics"]
    h, w = depth.shape
    z = depth.astype(np.float64) / 1000.0  # open3d.geometry.PointCloud.CreateFromDepthImage: double precision, depth_scale 1000
    u, v = np.meshgrid(np.arange(x0, x0 + w), np.arange(y0, y0 + h))
    points = np.stack([(u - K[0, 2]) * z / K[0, 0], (v - K[1, 2]) * z / K[1, 1], z], -1).reshape(-1, 3).astype(np.float32)
    valid = depth > 0
    inner = np.zeros_like(valid)
    inner[1:-1, 1:-1] = valid[1:-1, 1:-1] & valid[1:-1, :-2] & valid[1:-1, 2:] & valid[:-2, 1:-1] & valid[2:, 1:-1]
    return points, h, w, kat["normals_gt"], inner


def test_oracle_ordered_point_cloud_normals_reference_ground_truth(oracle):
    """Where a pixel and its four neighbours have depth, the reference's stored normals are reproduced to fp32 accuracy of the
    unprojected points (the stored file zeroes pixels next to missing depth, which ComputeOrderedPointCloudNormals itself does not)."""
    points, h, w, gt, inner = _kat_points()
    normals = oracle.compute_ordered_point_cloud_normals(points, h, w).reshape(h, w, 3)
    assert inner.sum() > 7000
    assert np.abs(normals - gt)[inner].max() < 1e-4
    border = np.ones((h, w), bool)
    border[1:-1, 1:-1] = False
    assert not normals[border].any()
    assert (normals[..., 2] <= 0).all()


def test_oracle_mesh_normals_match_numpy(oracle):
    s = synthetic.make_scene("tinyr", 120, 160)
    p, t = s["positions"].astype(np.float64), s["triangles"]
    face = np.cross(p[t[:, 1]] - p[t[:, 0]], p[t[:, 2]] - p[t[:, 0]])
    np.testing.assert_allclose(oracle.compute_triangle_normals(s["positions"], t, False), face, rtol=1e-4, atol=1e-9)
    np.testing.assert_allclose(oracle.compute_triangle_normals(s["positions"], t, True), face / np.linalg.norm(face, axis=1, keepdims=True), atol=2e-6)
    vertex = np.zeros_like(p)
    for k in range(3):
        np.add.at(vertex, t[:, k], face)
    np.testing.assert_allclose(oracle.compute_vertex_normals(s["positions"], t, False), vertex, rtol=1e-4, atol=1e-9)
    np.testing.assert_allclose(oracle.compute_vertex_normals(s["positions"], t, True), vertex / np.linalg.norm(vertex, axis=1, keepdims=True), atol=2e-6)
    # degenerate triangle: zero normal stays zero un-normalised, becomes... stays zero when normalised (z > 0 test), an unreferenced vertex too
    positions = np.array([[0, 0, 0], [1, 0, 0], [2, 0, 0], [5, 5, 5]], np.float32)
    triangles = np.array([[0, 1, 2]], np.int64)
    assert not oracle.compute_triangle_normals(positions, triangles, True).any()
    assert not oracle.compute_vertex_normals(positions, triangles, True).any()


@pytest.mark.gpu
@pytest.mark.parametrize("config,size", [("tinyr", (120, 160)), ("C2r", (480, 640))])
def test_mesh_normals_bit_exact(nnrt, oracle, config, size):
    s = synthetic.make_scene(config, *size)
    f = nnrt.geometry.functional
    mesh = dict(positions=s["positions"], triangles=s["triangles"])
    for normalized in (False, True):
        got = f.compute_triangle_normals(mesh, normalized)
        assert np.array_equal(_np(got), oracle.compute_triangle_normals(s["positions"], s["triangles"], normalized))
        assert mesh["triangle_normals"] is got  # set in place, like the reference's mesh.SetTriangleNormals
    # vertex normals use the mesh's triangle normals when present (here: the normalised ones just stored) ...
    got = f.compute_vertex_normals(mesh, True)
    want = oracle.compute_vertex_normals(s["positions"], s["triangles"], True, oracle.compute_triangle_normals(s["positions"], s["triangles"], True))
    assert np.array_equal(_np(got), want) and mesh["normals"] is got
    # ... and compute un-normalised ones otherwise (a tuple mesh carries none)
    for normalized in (False, True):
        got = f.compute_vertex_normals((s["positions"], None, s["triangles"]), normalized)
        assert np.array_equal(_np(got), oracle.compute_vertex_normals(s["positions"], s["triangles"], normalized))


@pytest.mark.gpu
def test_vertex_normals_irregular_mesh_bit_exact(nnrt, oracle):
    """Random connectivity: high-degree vertices (a fan of 600 faces), repeated vertices inside a face, unreferenced vertices."""
    rng = np.random.default_rng(21)
    positions = rng.normal(0, 1, (500, 3)).astype(np.float32)
    triangles = rng.integers(0, 450, (4000, 3)).astype(np.int64)  # vertices 450.. are never referenced
    triangles[:600, 0] = 7
    triangles[600:620, 1] = triangles[600:620, 0]
    got = nnrt.geometry.functional.compute_vertex_normals((positions, None, triangles), True)
    want = oracle.compute_vertex_normals(positions, triangles, True)
    assert np.array_equal(_np(got), want, equal_nan=True)
    assert not _np(got)[450:].any()


@pytest.mark.gpu
def test_ordered_point_cloud_normals_bit_exact(nnrt, oracle):
    points, h, w, gt, inner = _kat_points()
    got = nnrt.geometry.functional.compute_ordered_point_cloud_normals(points, (h, w))
    assert np.array_equal(_np(got), oracle.compute_ordered_point_cloud_normals(points, h, w))
    assert np.abs(_np(got).reshape(h, w, 3) - gt)[inner].max() < 1e-4
    rng = np.random.default_rng(3)
    cloud = rng.normal(0, 1, (37 * 53, 3)).astype(np.float32)
    got = nnrt.geometry.functional.compute_ordered_point_cloud_normals(cloud, (37, 53))
    assert np.array_equal(_np(got), oracle.compute_ordered_point_cloud_normals(cloud, 37, 53))
    with pytest.raises(RuntimeError):
        nnrt.geometry.functional.compute_ordered_point_cloud_normals(cloud, (37, 54))
