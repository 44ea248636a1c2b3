"""Developer tool: cProfile of the host side of fit_to_image on the bench workload (where do the ~200 us of Python per frame go)."""
import cProfile, pstats, os, sys, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from dynamicfuion_python_b200 import nnrt
dev = torch.device("cuda", 0)
s = bench.make_bench_scene()
G = nnrt.geometry
K, ext = s["K"], np.eye(4)
mesh = (torch.tensor(s["positions"], device=dev), torch.tensor(s["normals"], device=dev), torch.tensor(s["triangles"], device=dev))
wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                  G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
depth = torch.rand((480, 640), device=dev) * 0 + 0.6
fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=1.0, arap_term_weight=200.0)
def frame():
    fitter.fit_to_image(wf, mesh, None, depth, None, K, ext, 1.0)
for _ in range(20): frame()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(300): frame()
pr.disable(); torch.cuda.synchronize()
out = io.StringIO(); pstats.Stats(pr, stream=out).sort_stats("tottime").print_stats(22); print(out.getvalue()[:6000])
