"""Device time and algorithmic-byte rate of the rows built beyond the GN iteration (SURVEY.md §8(b) names and §8(f) ranks 2-4), next to
the CPU oracle (one thread) on the same inputs: shortest-path anchors, normals, fragment shaders, non-rigid TSDF integration.
One JSON line per case; `gbs` = algorithmic bytes (compulsory traffic of the API-visible tensors, stated per case) / device time."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from dynamicfuion_python_b200 import nnrt, synthetic  # noqa: E402
from oracle import oracle_py as oracle  # noqa: E402  (CPU baseline; measurement scripts live under tests/ because only tests may use the oracle)

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else None


def gpu_ms(fn, repeats=50):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(repeats):
        fn()
    stop.record()
    torch.cuda.synchronize()
    return start.elapsed_time(stop) / repeats


def cpu_ms(fn):
    t = time.perf_counter()
    fn()
    return (time.perf_counter() - t) * 1e3


def report(op, case, ms, algorithmic_bytes, oracle_ms, **extra):
    gbs = algorithmic_bytes / ms / 1e6
    line = dict(op=op, case=case, gpu_ms=round(ms, 4), algorithmic_mb=round(algorithmic_bytes / 1e6, 2), gbs=round(gbs, 1),
                frac_of_measured_hbm=None if PEAK is None else round(gbs / PEAK, 4), oracle_cpu_ms=None if oracle_ms != oracle_ms else round(oracle_ms, 2), **extra)
    print(json.dumps(line), flush=True)


def main():
    f = nnrt.geometry.functional
    H, W = 480, 640
    s = synthetic.make_scene("C2r", H, W)
    V, F, N = len(s["positions"]), len(s["triangles"]), len(s["nodes"])
    pos, tri, nodes = torch.as_tensor(s["positions"]).cuda(), torch.as_tensor(s["triangles"]).cuda(), torch.as_tensor(s["nodes"]).cuda()

    # shortest-path anchors: 12 B point in, 32 B anchors + weights out (A = 4); the node table (12 N + 16 N) is shared
    d = np.linalg.norm(s["nodes"][:, None].astype(np.float64) - s["nodes"][None].astype(np.float64), axis=2)
    np.fill_diagonal(d, np.inf)
    edges = np.argsort(d, axis=1, kind="stable")[:, :4].astype(np.int32)
    edges_t = torch.as_tensor(edges).cuda()
    ms = gpu_ms(lambda: f.compute_anchors_and_weights_shortest_path_fixed_node_weight(pos, nodes, edges_t, 4, s["node_coverage"]))
    report("compute_anchors_and_weights_shortest_path_fixed_node_weight", f"C2r mesh V={V}, {N} nodes, degree 4, A=4", ms, 44 * V + 28 * N,
           cpu_ms(lambda: oracle.compute_anchors_shortest_path(s["positions"], s["nodes"], edges, 4, s["node_coverage"])), bound="latency / issue (N distance evaluations + heap per point)")

    # normals: triangles 24 B + 3 gathered vertices 36 B in, 12 B out per face; vertex normals: 24 B + 12 B per face in, 12 B per vertex out
    mesh = (pos, None, tri)
    ms = gpu_ms(lambda: f.compute_triangle_normals(mesh, True))
    report("compute_triangle_normals", f"C2r mesh F={F}", ms, 72 * F, cpu_ms(lambda: oracle.compute_triangle_normals(s["positions"], s["triangles"], True)), bound="hbm")
    ms = gpu_ms(lambda: f.compute_vertex_normals(mesh, True))
    report("compute_vertex_normals (incl. un-normalised triangle normals)", f"C2r mesh V={V} F={F}", ms, 72 * F + 36 * F + 12 * V,
           cpu_ms(lambda: oracle.compute_vertex_normals(s["positions"], s["triangles"], True)), bound="latency (6 launches: CSR build + gather)")
    cloud = torch.as_tensor(np.random.default_rng(0).normal(0, 1, (H * W, 3)).astype(np.float32)).cuda()
    ms = gpu_ms(lambda: f.compute_ordered_point_cloud_normals(cloud, (H, W)))
    report("compute_ordered_point_cloud_normals", f"{W}x{H}", ms, 24 * H * W, cpu_ms(lambda: oracle.compute_ordered_point_cloud_normals(cloud.cpu().numpy(), H, W)), bound="hbm")

    # shaders over K = 1 fragments: edges 8 + 4 B in, 3 B out per pixel; vertex colours 8 + 12 B in + 36 B gathered per covered pixel, 3 B out
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], H, W, 0.0, 10.0)
    idx, depth, bary, dist, _ = oracle.rasterize(fv, mask, H, W, 0.0, 1, perspective_correct=True)
    t = [torch.as_tensor(a).cuda() for a in (idx, depth, bary, dist)]
    shader = nnrt.rendering.FlatEdgeShader(1.0, (1.0, 1.0, 1.0))
    ms = gpu_ms(lambda: shader.shade_meshes(*t))
    report("FlatEdgeShader.shade_meshes", f"{W}x{H}, K=1", ms, 15 * H * W, cpu_ms(lambda: oracle.shade_edges_flat(idx, dist, 1.0, (1.0, 1.0, 1.0))), bound="hbm")
    colors = np.random.default_rng(1).uniform(0, 1, (V, 3)).astype(np.float32)
    fvc = torch.as_tensor(colors[s["triangles"]]).cuda()
    covered = int((idx[..., 0] >= 0).sum())
    vcs = nnrt.rendering.VertexColorShader()
    mesh_colors = dict(colors=torch.as_tensor(colors).cuda(), triangles=tri)
    ms = gpu_ms(lambda: vcs.shade_meshes(*t, [mesh_colors]))
    report("VertexColorShader.shade_meshes (incl. the per-face colour gather)", f"{W}x{H}, K=1, {covered} covered pixels", ms, 23 * H * W + 36 * covered + 72 * F,
           cpu_ms(lambda: oracle.shade_vertex_colors(idx, bary, colors[s["triangles"]])), bound="hbm")
    del fvc

    # TSDF integration: per voxel tsdf read + write 8 B, weight 4 B (+ block key / index amortised); images are gathered (<= 4 + 12 B per voxel)
    R, voxel = 8, 0.004
    block = voxel * R
    lo = np.floor(s["positions"].min(0) / block).astype(np.int32) - 1
    hi = np.floor(s["positions"].max(0) / block).astype(np.int32) + 1
    keys = np.stack(np.meshgrid(*[np.arange(lo[k], hi[k] + 1) for k in range(3)], indexing="ij"), -1).reshape(-1, 3).astype(np.int32)
    B = len(keys)
    active = np.arange(B, dtype=np.int64)
    depth_img = np.where(depth[..., 0] > 0, depth[..., 0], 0).astype(np.float32)
    pts, _ = oracle.unproject_depth(depth_img, s["K"], None, 1.0, 10.0)
    normals = oracle.compute_ordered_point_cloud_normals(pts, H, W).reshape(H, W, 3)
    tsdf = np.zeros((B, R, R, R), np.float32)
    weight = np.zeros((B, R, R, R), np.float32)
    G = nnrt.geometry
    wf = G.GraphWarpField(s["nodes"], 0.02, True, 4, 1, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    wf.set_node_rotations(s["rotations_gt"])
    wf.set_node_translations(s["translations_gt"])
    values = dict(tsdf=torch.as_tensor(tsdf).cuda(), weight=torch.as_tensor(weight).cuda())
    args = (torch.as_tensor(active).cuda(), torch.as_tensor(keys).cuda(), values, R, voxel, wf, torch.as_tensor(depth_img).cuda(), None,
            torch.as_tensor(normals).cuda(), s["K"], s["K"], None, 1.0, 3.0, 2.0)
    ms = gpu_ms(lambda: f.integrate_non_rigid_voxel_blocks(*args), 20)
    ms_full = gpu_ms(lambda: f.integrate_non_rigid_voxel_blocks(*args, _full_scan=True), 20)
    voxels = B * R ** 3
    sample = active[:max(1, B // 50)]
    cpu = cpu_ms(lambda: oracle.integrate_non_rigid(sample, keys, tsdf.copy(), weight, None, R, voxel, 2 * voxel, depth_img, None, normals, s["K"], s["K"], np.eye(4),
                                                    s["nodes"], s["rotations_gt"], s["translations_gt"], 0.02, 4, 1, 1.0, 3.0)) * (B / len(sample))
    report("integrate_non_rigid_voxel_blocks", f"{B} blocks of 8^3 voxels ({voxels} voxels) around the C2r surface, {N} nodes, A=4, {W}x{H} depth", ms,
           12 * voxels + 16 * H * W, cpu, bound="issue (shortlisted node distances per voxel)", oracle_note=f"extrapolated from {len(sample)} blocks",
           full_scan_ms=round(ms_full, 4), full_scan_voxel_node_pairs_per_s=round(voxels * N / ms_full * 1e3 / 1e9, 1))
    # the same blocks against a 2000-node field (C3-scale graph): most nodes are far from any one block
    rng2 = np.random.default_rng(5)
    many = np.concatenate([s["positions"][rng2.choice(V, 1000, replace=False)], rng2.uniform(-1, 1, (1000, 3)) + [0, 0, 0.6]]).astype(np.float32)
    many = rng2.permutation(many)
    wf2 = G.GraphWarpField(many, 0.01, True, 4, 1, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    args2 = args[:5] + (wf2,) + args[6:]
    ms2 = gpu_ms(lambda: f.integrate_non_rigid_voxel_blocks(*args2), 20)
    ms2_full = gpu_ms(lambda: f.integrate_non_rigid_voxel_blocks(*args2, _full_scan=True), 20)
    report("integrate_non_rigid_voxel_blocks", f"{B} blocks ({voxels} voxels), 2000 nodes, A=4", ms2, 12 * voxels + 16 * H * W, float("nan"),
           bound="issue (shortlisted node distances per voxel)", full_scan_ms=round(ms2_full, 4),
           full_scan_voxel_node_pairs_per_s=round(voxels * 2000 / ms2_full * 1e3 / 1e9, 1))
    # block selection of FindBlocksIntersectingTruncationRegion: 27 neighbours of every block; 12 B key in, 24 B box out, then 24 B box in, 1 B out
    neighbours = np.unique((keys[:, None, :] + np.stack(np.meshgrid(*[np.arange(-1, 2)] * 3, indexing="ij"), -1).reshape(1, 27, 3)).reshape(-1, 3), axis=0).astype(np.int32)
    keys_t = torch.as_tensor(neighbours).cuda()
    ms = gpu_ms(lambda: f.bounding_boxes_of_warped_blocks(keys_t, wf, voxel, R))
    report("bounding_boxes_of_warped_blocks", f"{len(neighbours)} block keys x 8 corners, {N} nodes", ms, 36 * len(neighbours),
           cpu_ms(lambda: oracle.bounding_boxes_of_warped_blocks(neighbours, s["nodes"], s["rotations_gt"], s["translations_gt"], 0.02, 4, 1, voxel, R, np.eye(4))),
           bound="issue (N distance evaluations per corner)")
    centres = (neighbours.astype(np.float32) + 0.5) * block  # metric boxes of the same blocks (the reference's own boxes mix units)
    metric_boxes = np.concatenate([centres - block / 2, centres + block / 2], 1).astype(np.float32)
    boxes_t, depth_t = torch.as_tensor(metric_boxes).cuda(), torch.as_tensor(depth_img).cuda()
    ms = gpu_ms(lambda: f.axis_aligned_boxes_intersecting_surface_mask(boxes_t, depth_t, s["K"], 1.0, 3.0, 4, 2 * voxel))
    segments = int((depth_img[::4, ::4] > 0).sum())
    hits = int(oracle.boxes_intersecting_surface_mask(metric_boxes, s["K"], depth_img, 1.0, 3.0, 4, 2 * voxel).sum())
    report("axis_aligned_boxes_intersecting_surface_mask", f"{len(metric_boxes)} boxes x {segments} segments ({W}x{H}, stride 4), {hits} hits", ms,
           25 * len(metric_boxes) + 4 * (H // 4) * (W // 4),
           cpu_ms(lambda: oracle.boxes_intersecting_surface_mask(metric_boxes, s["K"], depth_img, 1.0, 3.0, 4, 2 * voxel)),
           bound="latency (a warp per box over its projected pixel rectangle)", reference_pairs=len(metric_boxes) * segments)


if __name__ == "__main__":
    main()
