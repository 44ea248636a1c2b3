"""Developer script: one C2r fit on the GPU vs the oracle, iteration by iteration."""
import sys, os
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle_py as oracle
from dynamicfuion_python_b200 import nnrt
from util import scene_with_target, oracle_fit

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2r"
H, W = (480, 640) if len(sys.argv) < 4 else (int(sys.argv[2]), int(sys.argv[3]))
s = scene_with_target(oracle, cfg, H, W)
G = nnrt.geometry
for iters in (1, 2, 3):
    wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                      G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
    fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=iters, preconditioning_dampening_factor=s["lm_factor"],
                                                        arap_term_weight=s["arap_term_weight"])
    fitter.fit_to_image(wf, (s["positions"], s["normals"], s["triangles"]), None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4),
                        (H, W), anchors=s["anchors"], anchor_weights=s["anchor_weights"])
    try:
        fitter.check_status()
    except RuntimeError as e:
        print("status:", e)
    ref = oracle_fit(oracle, s, iters)
    d = fitter.diagnostics()
    dl = d["deltas"].cpu().numpy(); rd = ref["deltas"][iters - 1]
    Hd = d["hessian_diagonal"].cpu().numpy(); rH = ref["hessian_diag"][iters - 1]
    print(iters, "delta err", np.abs(dl - rd).max(), "scale", np.abs(rd).max(), "H err", np.abs(Hd - rH).max() / np.abs(rH).max(),
          "finite", np.isfinite(dl).all(), "face mismatch", int((d["pixel_face_indices"].cpu().numpy() != ref["pixel_faces"]).sum()))
    t = wf.node_translations.cpu().numpy(); R = wf.node_rotations.cpu().numpy()
    bad = np.where(~np.isfinite(t).all(1) | (np.abs(t - ref["translations"]).max(1) > 1e-5))[0]
    print("   t err", np.abs(t - ref["translations"]).max(), "R err", np.abs(R - ref["rotations"]).max(), "bad nodes", bad[:20], len(bad))
    w = np.linalg.eigvalsh(rH.astype(np.float64)); print("   min eig of diag blocks", w.min())
