"""Developer measurement: saturated cost of the two halves of an iteration (data term: warp + raster + assemble; solve: system + stem +
corner + update) when S sequences run concurrently, through the split-iteration C ABI (eager launches on S streams)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    from dynamicfuion_python_b200 import nnrt
    from dynamicfuion_python_b200.nnrt import _tensor as T
    dev = torch.device("cuda", 0)
    s = bench.make_bench_scene()
    G = nnrt.geometry
    K, ext = s["K"], np.eye(4)
    mesh = (torch.tensor(s["positions"], device=dev), torch.tensor(s["normals"], device=dev), torch.tensor(s["triangles"], device=dev))

    def new_field():
        return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                            G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])

    wf = new_field()
    anchors, weights = wf.precompute_anchors_and_weights(mesh)
    R, t = bench.scaled_motion(s, 1.0)
    wf.set_node_rotations(R); wf.set_node_translations(t)
    warped = wf.warp_mesh(mesh, anchors, weights, True, ext)
    fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(warped, K, (bench.HEIGHT, bench.WIDTH), 0.0, 10.0)
    _, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (bench.HEIGHT, bench.WIDTH), 0.0, 1, -1, -1, True, False, True)
    frame = torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous()
    Kh, Eh = T.host_f64(K, (3, 3)), T.host_f64(ext, (4, 4))
    for S in (1, 4, 16):
        seqs = []
        for k in range(S):
            q = dict(wf=new_field(), fitter=nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=s["lm_factor"],
                                                                                     arap_term_weight=s["arap_term_weight"]), stream=torch.cuda.Stream(dev))
            q["fitter"].fit_to_image(q["wf"], mesh, None, frame, None, K, ext, 1.0)
            seqs.append(q)
        torch.cuda.synchronize()

        def run(which, count):
            for r in range(count):
                for q in seqs:
                    h, Rp, tp, st = q["fitter"]._handle, T.ptr(q["wf"].node_rotations), T.ptr(q["wf"].node_translations), q["stream"].cuda_stream
                    if which in ("data", "both"):
                        T.call("nnrt_b200_fitter_iteration_data_term", h, Rp, tp, Kh.data_ptr(), Eh.data_ptr(), st)
                    if which in ("solve", "both"):
                        T.call("nnrt_b200_fitter_iteration_solve", h, Rp, tp, Kh.data_ptr(), Eh.data_ptr(), 0, st)
            sub = time.perf_counter()
            torch.cuda.synchronize()
            return sub

        out = {"sequences": S}
        for which in ("data", "solve", "both"):
            run(which, 5)
            t0 = time.perf_counter()
            sub = run(which, 40)
            dt = time.perf_counter() - t0
            out[which + "_us"] = dt / (40 * S) * 1e6
            out[which + "_host_us"] = (sub - t0) / (40 * S) * 1e6
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
