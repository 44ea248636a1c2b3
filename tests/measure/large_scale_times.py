"""C3-scale device times of the compute-shaped widened rows (no CPU baseline: minutes on one core): shortest-path anchors for a
2.2 M-vertex mesh against a 2000-node graph, and non-rigid TSDF integration of ~20 000 voxel blocks (10 M voxels) against 2000 nodes.
One JSON line per case."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from dynamicfuion_python_b200 import nnrt  # noqa: E402


def gpu_ms(fn, repeats=5):
    fn()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(repeats):
        fn()
    stop.record()
    torch.cuda.synchronize()
    return start.elapsed_time(stop) / repeats


def main():
    rng = np.random.default_rng(3)
    f = nnrt.geometry.functional
    G = nnrt.geometry
    # a 1.6 m x 0.9 m relief wall at z = 2 (the C3 stand-in of tools/c3_stage_times.py), 2000 nodes on it
    side = 1492
    x, y = np.meshgrid(np.linspace(-0.8, 0.8, side), np.linspace(-0.45, 0.45, side))
    z = 2.0 + 0.02 * np.sin(9 * x) * np.cos(7 * y)
    vertices = torch.as_tensor(np.stack([x, y, z], -1).reshape(-1, 3).astype(np.float32)).cuda()
    V = vertices.shape[0]
    node_ids = torch.as_tensor(rng.choice(V, 2000, replace=False)).cuda()
    nodes = vertices[node_ids].contiguous()
    tree = nnrt.core.KDTree(nodes)
    neighbours, _ = tree.find_k_nearest_to_points(nodes, 5, sort_output=True)
    edges = neighbours[:, 1:].contiguous()  # 4 nearest other nodes
    ms = gpu_ms(lambda: f.compute_anchors_and_weights_shortest_path_fixed_node_weight(vertices, nodes, edges, 4, 0.03))
    print(json.dumps(dict(op="compute_anchors_and_weights_shortest_path_fixed_node_weight", case=f"V={V}, 2000 nodes, degree 4, A=4", gpu_ms=round(ms, 3),
                          point_node_pairs_per_s=round(V * 2000 / ms * 1e3 / 1e9, 1))), flush=True)
    ms = gpu_ms(lambda: f.compute_anchors_and_weights_euclidean_fixed_node_weight(vertices, nodes, 4, 0, 0.03))
    print(json.dumps(dict(op="compute_anchors_and_weights_euclidean_fixed_node_weight (the per-fit anchors of the hot path, for scale)", case=f"V={V}, 2000 nodes, A=4",
                          gpu_ms=round(ms, 3), point_node_pairs_per_s=round(V * 2000 / ms * 1e3 / 1e9, 1))), flush=True)
    # voxel blocks of 8^3 voxels of 5 mm within 2 blocks of the wall
    R, voxel = 8, 0.005
    block = R * voxel
    gx, gy = np.meshgrid(np.arange(np.floor(-0.8 / block), np.ceil(0.8 / block)), np.arange(np.floor(-0.45 / block), np.ceil(0.45 / block)))
    layers = [np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, np.floor(2.0 / block) + dz)], -1) for dz in (-2, -1, 0, 1, 2)]
    keys = np.concatenate(layers).astype(np.int32)
    B = len(keys)
    H, W = 1080, 1920
    K = np.array([[1740.0, 0, 960.0], [0, 1740.0, 540.0], [0, 0, 1]])
    depth = torch.full((H, W), 2.0, dtype=torch.float32, device="cuda")
    normals = torch.zeros((H, W, 3), dtype=torch.float32, device="cuda")
    normals[..., 2] = 1.0
    wf = G.GraphWarpField(nodes, 0.03, True, 4, 1, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    values = dict(tsdf=torch.zeros((B, R, R, R), dtype=torch.float32, device="cuda"), weight=torch.zeros((B, R, R, R), dtype=torch.float32, device="cuda"))
    args = (torch.arange(B, device="cuda"), torch.as_tensor(keys).cuda(), values, R, voxel, wf, depth, None, normals, K, K, None, 1.0, 5.0, 2.0)
    ms = gpu_ms(lambda: f.integrate_non_rigid_voxel_blocks(*args))
    ms_full = gpu_ms(lambda: f.integrate_non_rigid_voxel_blocks(*args, _full_scan=True), 2)
    voxels = B * R ** 3
    print(json.dumps(dict(op="integrate_non_rigid_voxel_blocks", case=f"{B} blocks ({voxels} voxels), 2000 nodes, A=4, {W}x{H} depth", gpu_ms=round(ms, 3),
                          full_scan_ms=round(ms_full, 3), voxels_per_s=round(voxels / ms * 1e3 / 1e9, 2),
                          integrated_voxels=int((values["tsdf"] != 0).sum().item()))), flush=True)
    pts = vertices
    ms = gpu_ms(lambda: f.median_grid_subsample_3d_points(pts, 0.03))
    print(json.dumps(dict(op="median_grid_subsample_3d_points", case=f"V={V} -> nodes, cell 0.03", gpu_ms=round(ms, 3),
                          bins=int(f.median_grid_subsample_3d_points(pts, 0.03).shape[0]))), flush=True)


if __name__ == "__main__":
    main()
