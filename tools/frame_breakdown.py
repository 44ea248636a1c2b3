"""Developer tool: splits a bench frame (bench.py, workload C2r) into per-fit setup and Gauss-Newton iterations by timing
fit_to_image with iteration_count = 0, 1 and 10 (CUDA events, resident depth frames)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from dynamicfuion_python_b200 import nnrt

dev = torch.device("cuda", 0)
s = bench.make_bench_scene()
G = nnrt.geometry
K, ext = s["K"], np.eye(4)
mesh = (torch.tensor(s["positions"], device=dev), torch.tensor(s["normals"], device=dev), torch.tensor(s["triangles"], device=dev))
N = s["nodes"].shape[0]
wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                  G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
anchors, weights = wf.precompute_anchors_and_weights(mesh)
R, t = bench.scaled_motion(s, s["frame_scales"][0])
wf.set_node_rotations(R); wf.set_node_translations(t)
warped = wf.warp_mesh(mesh, anchors, weights, True, ext)
fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(warped, K, (bench.HEIGHT, bench.WIDTH), 0.0, 10.0)
_, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (bench.HEIGHT, bench.WIDTH), 0.0, 1, -1, -1, True, False, True)
depth = torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous()
fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=10, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
eye = torch.eye(3, device=dev).repeat(N, 1, 1).contiguous()

def frame(iters):
    wf.node_rotations.copy_(eye); wf.node_translations.zero_()
    fitter.fit_to_image(wf, mesh, None, depth, None, K, ext, 1.0, iteration_count=iters)

for iters in (10, 0, 1, 10, 0, 1, 2):
    for _ in range(5): frame(iters)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(30): frame(iters)
    b.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"iterations {iters:2d}: {a.elapsed_time(b) / 30 * 1e3:8.1f} us/frame (events)  {(t1 - t0) / 30 * 1e6:8.1f} us/frame (wall)")
