"""Measurement tool for BASELINE.json configs[2] ("C3": ~4.45 M triangles, 1920x1080, ~2000 nodes): per-stage CUDA-event times of one
GN iteration (warp, face raster, band raster, assemble, system+stem, corner solve) and the achieved algorithmic GB/s of the data-term
stages. Synthetic: relief sheet n x n vertices, nodes on a regular grid of mesh vertices (fine layer + coarse layer), seeded."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dynamicfuion_python_b200 import nnrt, synthetic

H, W = 1080, 1920
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1492
fine_side, coarse_side = 44, 8
dev = torch.device("cuda", 0)
rng = np.random.default_rng(synthetic.SEED)
pos, nrm, tris = synthetic.relief_mesh(n, 0.3, 0.6, 0.02, rng=rng)
K = synthetic.default_intrinsics(H, W)
tris = synthetic.orient_front_faces(pos, nrm, tris, K, H, W)


def grid_nodes(side):
    idx = np.linspace(0, n - 1, side).round().astype(np.int64)
    return pos[(idx[:, None] * n + idx[None, :]).ravel()].astype(np.float64)


fine, coarse = grid_nodes(fine_side), grid_nodes(coarse_side) + rng.normal(scale=1e-4, size=(coarse_side ** 2, 3))
d = np.linalg.norm(fine[:, None] - coarse[None], axis=2)
nn = np.argsort(d, axis=1, kind="stable")[:, :4]
edges = np.stack([np.repeat(np.arange(len(fine)), 4), len(fine) + nn.ravel()], 1).astype(np.int32)
nodes = np.concatenate([fine, coarse], 0).astype(np.float32)
N, E, base = len(nodes), len(edges), len(fine)
coverage = 0.25 * 0.3 / fine_side
G = nnrt.geometry
ext = np.eye(4)
mesh = (torch.tensor(pos, device=dev), torch.tensor(nrm, device=dev), torch.tensor(tris, device=dev))


def field():
    return G.HierarchicalGraphWarpField.from_graph(nodes, edges, base, coverage, False, 4, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4,
                                        np.zeros(max(E, N), np.int8), np.array([0.015, 0.03], np.float32))


wf = field()
anchors, weights = wf.precompute_anchors_and_weights(mesh)
R, t = synthetic.smooth_node_motion(nodes, max_angle_deg=1.5, max_translation=0.002, rng=np.random.default_rng(synthetic.SEED + 1))
wf.set_node_rotations(R); wf.set_node_translations(t)
warped = wf.warp_mesh(mesh, anchors, weights, True, ext)
fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(warped, K, (H, W), 0.0, 10.0)
_, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (H, W), 0.0, 1, -1, -1, True, False, True)
depth = torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous()
wf = field()
fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=1.0, arap_term_weight=200.0 * (H * W / (480.0 * 640.0)) ** 0.5)
fitter.fit_to_image(wf, mesh, None, depth, None, K, ext, 1.0, anchors=anchors, anchor_weights=weights)
torch.cuda.synchronize()
try:
    fitter.check_status(); status = "ok"
except RuntimeError as error:
    status = str(error)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
stage = np.zeros(7)
reps = 10
for rep in range(reps):
    wf.set_node_rotations(np.tile(np.eye(3, dtype=np.float32), (N, 1, 1))); wf.set_node_translations(np.zeros((N, 3), np.float32))
    flush.fill_(rep)
    torch.cuda.synchronize()
    stage += np.array(fitter.profile_iteration(wf, K, ext))
stage /= reps
# whole iterations inside the captured graph: (10-iteration frame - 1-iteration frame) / 9
graph_ms = {}
for iters in (1, 10):
    def frame():
        wf.set_node_rotations(np.tile(np.eye(3, dtype=np.float32), (N, 1, 1))); wf.set_node_translations(np.zeros((N, 3), np.float32))
        fitter.fit_to_image(wf, mesh, None, depth, None, K, ext, 1.0, anchors=anchors, anchor_weights=weights, iteration_count=iters)
    for _ in range(2):
        frame()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        frame()
    b.record(); torch.cuda.synchronize()
    graph_ms[iters] = a.elapsed_time(b) / 5
try:
    fitter.check_status(); status10 = "ok"
except RuntimeError as error:
    status10 = str(error)
V, F, PX = pos.shape[0], tris.shape[0], H * W
names = ["warp_vertices", "face_raster", "band_raster", "assemble", "build_system", "stem_eliminate", "corner_solve_update"]
bytes_by = {"warp_vertices": 80 * V, "raster": 97 * F + 28 * PX, "assemble": 41 * PX + 168 * N}
data_ms = float(stage[:4].sum())
print(json.dumps({"config": "C3", "V": V, "F": F, "nodes": N, "stem": base, "corner": N - base, "edges": E, "image": f"{W}x{H}",
                  "coverage": float((depth > 0).float().mean()), "status_first_fit": status,
                  "stage_ms": dict(zip(names, [float(x) for x in stage])),
                  "ms_per_iteration_in_graph": (graph_ms[10] - graph_ms[1]) / 9, "ms_per_frame_10_iterations": graph_ms[10], "status_10_iterations": status10,
                  "iteration_algorithmic_bytes": int(80 * V + 97 * F + 69 * PX + 168 * N + 300 * E),
                  "data_term_ms": data_ms, "data_term_algorithmic_bytes": int(sum(bytes_by.values())),
                  "data_term_achieved_gbs": sum(bytes_by.values()) / (data_ms * 1e-3) / 1e9,
                  "warp_gbs": bytes_by["warp_vertices"] / (stage[0] * 1e-3) / 1e9,
                  "raster_gbs": bytes_by["raster"] / ((stage[1] + stage[2]) * 1e-3) / 1e9,
                  "assemble_gbs": bytes_by["assemble"] / (stage[3] * 1e-3) / 1e9}))
