"""Empty and degenerate inputs through both front ends: the entry points must return empty results (or the reference's error), not fault.
The reference's own tests exercise a few of these (empty face lists in test_rasterize.cpp, zero-size clouds in test_sampling.cpp)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_zero_point_inputs(nnrt):
    f = nnrt.geometry.functional
    nodes = np.array([[0.0, 0.0, 1.0], [0.1, 0.0, 1.0], [0.0, 0.1, 1.0], [0.1, 0.1, 1.1]], np.float32)
    none = np.zeros((0, 3), np.float32)
    anchors, weights = f.compute_anchors_and_weights_euclidean_fixed_node_weight(none, nodes, 4, 0, 0.05)
    assert tuple(anchors.shape) == (0, 4) and tuple(weights.shape) == (0, 4)
    rotations = np.broadcast_to(np.eye(3, dtype=np.float32), (4, 3, 3)).copy()
    warped = f.warp_point_cloud(none, nodes, rotations, np.zeros((4, 3), np.float32), 4, 0.05, 0)
    warped = warped if isinstance(warped, torch.Tensor) else getattr(warped, "positions", warped)
    assert tuple(torch.as_tensor(warped).shape) == (0, 3)
    for sample in (f.mean_grid_downsample_3d_points, f.median_grid_subsample_3d_points, f.closest_to_mean_grid_subsample_3d_points,
                   f.fast_mean_radius_downsample_3d_points, f.fast_median_radius_subsample_3d_points):
        assert sample(none, 0.05).shape[0] == 0
        assert sample(nodes[:1], 0.05).shape[0] == 1
    keys = np.zeros((0, 3), np.int32)
    assert tuple(f.voxel_block_neighbors(keys).shape) == (0, 27)
    mesh = f.extract_triangle_mesh(keys, dict(tsdf=torch.zeros((0, 8, 8, 8), device="cuda"), weight=torch.zeros((0, 8, 8, 8), device="cuda")), 0.01, 3.0)
    assert mesh.positions.shape[0] == 0 and mesh.triangles.shape[0] == 0


def test_mesh_without_faces_or_coverage(nnrt):
    """A mesh entirely outside the frustum renders nothing; normals of a mesh whose faces are all degenerate are zero, not NaN."""
    R = nnrt.rendering
    positions = np.array([[0.0, 0.0, -1.0], [0.1, 0.0, -1.0], [0.0, 0.1, -1.0]], np.float32)  # behind the camera
    triangles = np.array([[0, 1, 2]], np.int64)
    K = np.array([[100.0, 0, 32], [0, 100.0, 24], [0, 0, 1]])
    fv, mask = R.functional.get_mesh_ndc_face_vertices_and_clip_mask((positions, None, triangles), K, (48, 64), 0.0, 10.0)
    faces, depth, bary, dist = R.rasterize_ndc_triangles(fv, mask, (48, 64), 0.0, 1, -1, -1, True, False, True)
    assert (faces.cpu().numpy() == -1).all()
    flat = np.array([[0.0, 0.0, 1.0], [0.0, 0.0, 1.0], [0.0, 0.0, 1.0]], np.float32)
    normals = nnrt.geometry.functional.compute_vertex_normals((flat, None, triangles))
    normals = normals if isinstance(normals, torch.Tensor) else normals[0]
    assert torch.isfinite(torch.as_tensor(normals)).all()
