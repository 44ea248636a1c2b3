"""Developer script: hunt the intermittent NaN: 1-iteration fits carried forward, diagnostics after each."""
import sys, os
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle_py as oracle
from dynamicfuion_python_b200 import nnrt
from util import scene_with_target
s = scene_with_target(oracle, "C2r", 480, 640)
G = nnrt.geometry
f = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
for trial in range(6):
    wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                      G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
    for it in range(10):
        f.fit_to_image(wf, (s["positions"], s["normals"], s["triangles"]), None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4),
                       (480, 640), anchors=s["anchors"], anchor_weights=s["anchor_weights"])
        torch.cuda.synchronize()
        d = f.diagnostics()
        fin = {k: bool(torch.isfinite(v.float()).all()) for k, v in d.items()}
        Rf, tf = bool(torch.isfinite(wf.node_rotations).all()), bool(torch.isfinite(wf.node_translations).all())
        if not (all(fin.values()) and Rf and tf):
            dl = d["deltas"].cpu().numpy()
            badn = np.where(~np.isfinite(dl).all(1))[0]
            print("trial", trial, "iteration", it, "finite:", fin, "R", Rf, "t", tf, "bad delta nodes:", len(badn), badn[:12],
                  "max|H|", float(d["hessian_diagonal"].abs().max()), "max|g|", float(d["negative_gradient"].abs().max()))
            break
    else:
        print("trial", trial, "clean, max |delta| last", float(d["deltas"].abs().max()))
