"""Developer/measurement tool: ms per GN iteration of the synthetic parity configurations (C1 plane, C1e plane with edges, C2r bench
scene) at 640x480, graph-captured 10-iteration fits, CUDA events. Prints one JSON line per configuration."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dynamicfuion_python_b200 import nnrt, synthetic

dev = torch.device("cuda", 0)
G = nnrt.geometry
for config in ("C1", "C1e", "C2r"):
    s = synthetic.make_scene(config, 480, 640)
    K, ext = s["K"], np.eye(4)
    mesh = (torch.tensor(s["positions"], device=dev), torch.tensor(s["normals"], device=dev), torch.tensor(s["triangles"], device=dev))
    N, E = s["nodes"].shape[0], len(s["edges"])

    def field():
        return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"] if E else None, s["arrow_base"], s["node_coverage"], False, 4, 0,
                                            G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
    wf = field()
    anchors, weights = wf.precompute_anchors_and_weights(mesh)
    wf.set_node_rotations(s["rotations_gt"]); wf.set_node_translations(s["translations_gt"])
    warped = wf.warp_mesh(mesh, anchors, weights, True, ext)
    fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(warped, K, (480, 640), 0.0, 10.0)
    _, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (480, 640), 0.0, 1, -1, -1, True, False, True)
    depth = torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous()
    wf = field()
    fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=10, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
    eye = torch.eye(3, device=dev).repeat(N, 1, 1).contiguous()
    times = {}
    for iters in (1, 10):
        def frame():
            wf.node_rotations.copy_(eye); wf.node_translations.zero_()
            fitter.fit_to_image(wf, mesh, None, depth, None, K, ext, 1.0, iteration_count=iters)
        for _ in range(5):
            frame()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(30):
            frame()
        b.record(); torch.cuda.synchronize()
        times[iters] = a.elapsed_time(b) / 30
    try:
        fitter.check_status(); status = "ok"
    except RuntimeError as error:
        status = str(error)
    per_iter = (times[10] - times[1]) / 9
    print(json.dumps({"config": config, "V": int(mesh[0].shape[0]), "F": int(mesh[2].shape[0]), "nodes": N, "edges": E, "coverage": float((depth > 0).float().mean()),
                      "ms_per_frame_10_iterations": times[10], "ms_per_iteration_in_graph": per_iter, "ms_setup_per_frame": times[1] - per_iter, "status": status}))
