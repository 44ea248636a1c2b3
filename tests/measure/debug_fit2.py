"""Developer script: consecutive 1-iteration fits (state carried in the warp field) vs one 2-iteration fit vs the oracle."""
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
def field():
    return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                        G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
def fit(wf, fitter):
    fitter.fit_to_image(wf, (s["positions"], s["normals"], s["triangles"]), None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4),
                        (H, W), anchors=s["anchors"], anchor_weights=s["anchor_weights"])
    try:
        fitter.check_status()
    except RuntimeError as e:
        print("  status:", e)
ref1, ref2 = oracle_fit(oracle, s, 1), oracle_fit(oracle, s, 2)
wf = field()
f1 = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
fit(wf, f1)
print("A: after 1st single fit  t err", np.abs(wf.node_translations.cpu().numpy() - ref1["translations"]).max())
fit(wf, f1)
print("A: after 2nd single fit  t err", np.abs(wf.node_translations.cpu().numpy() - ref2["translations"]).max(), "(same fitter object)")
wf = field()
f2 = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=2, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
fit(wf, f2)
print("B: one 2-iteration fit   t err", np.abs(wf.node_translations.cpu().numpy() - ref2["translations"]).max())
d = f2.diagnostics()
print("   H finite", bool(torch.isfinite(d["hessian_diagonal"]).all()), "g finite", bool(torch.isfinite(d["negative_gradient"]).all()),
      "res finite", bool(torch.isfinite(d["residuals"]).all()), "covered px", int((d["pixel_face_indices"] >= 0).sum()))
print("stress: 2-iteration fits with a fresh fitter each time")
bad = 0
for k in range(12):
    wf = field()
    f = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=2, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
    fit(wf, f)
    err = np.abs(wf.node_translations.cpu().numpy() - ref2["translations"]).max()
    print("  fresh", k, err)
print("stress: same fitter, 10-iteration fits")
ref10 = oracle_fit(oracle, s, 10)
f = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=10, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
for k in range(6):
    wf = field()
    fit(wf, f)
    err = np.abs(wf.node_translations.cpu().numpy() - ref10["translations"]).max()
    print("  same", k, err)
