"""Device time of the warp-field construction next to the CPU oracle: median-grid subsampling of mesh
vertices into nodes, and HierarchicalGraphWarpField's layers + edges. Prints one JSON line per case."""
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


def timed(fn, repeats=20):
    fn()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall = time.perf_counter()
    start.record()
    for _ in range(repeats):
        fn()
    stop.record()
    torch.cuda.synchronize()
    return start.elapsed_time(stop) / repeats, (time.perf_counter() - wall) * 1e3 / repeats


def main():
    G = nnrt.geometry
    rng = np.random.default_rng(1)
    s = synthetic.make_scene("C2r", 480, 640)
    vertices = torch.as_tensor(s["positions"]).cuda()
    for name, points, cell in (("C2r vertices -> nodes", vertices, 0.02), ("C2r vertices, fine grid", vertices, 0.004),
                               ("2.2M points", torch.as_tensor(rng.uniform(-1, 1, (2_200_000, 3)).astype(np.float32)).cuda(), 0.02)):
        gpu_ms, wall_ms = timed(lambda: G.functional.median_grid_subsample_3d_points(points, cell), 10)
        host = points.cpu().numpy()
        t = time.perf_counter()
        sample = oracle.median_grid_subsample_3d_points(host, cell)
        cpu_ms = (time.perf_counter() - t) * 1e3
        print(json.dumps(dict(op="median_grid_subsample_3d_points", case=name, points=int(points.shape[0]), cell=cell, bins=int(len(sample)),
                              gpu_ms=round(gpu_ms, 4), wall_ms=round(wall_ms, 4), oracle_cpu_ms=round(cpu_ms, 3))))
    for node_count, coverage, layers in ((249, 0.01, 2), (2000, 0.005, 2), (2000, 0.005, 3), (20000, 0.002, 3)):
        nodes = torch.as_tensor(rng.uniform(-0.15, 0.15, (node_count, 3)).astype(np.float32)).cuda()
        method = G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE
        gpu_ms, wall_ms = timed(lambda: G.HierarchicalGraphWarpField(nodes, coverage, False, 4, 0, method, layers, 4), 10)
        t = time.perf_counter()
        built = oracle.build_regularization_layers(nodes.cpu().numpy(), coverage, layers, 4)
        cpu_ms = (time.perf_counter() - t) * 1e3
        print(json.dumps(dict(op="HierarchicalGraphWarpField", nodes=node_count, layers=layers, layer_node_counts=[int(c) for c in built["layer_node_counts"]],
                              edges=int(len(built["edges"])), gpu_ms=round(gpu_ms, 4), wall_ms=round(wall_ms, 4), oracle_cpu_ms=round(cpu_ms, 3))))


if __name__ == "__main__":
    main()
