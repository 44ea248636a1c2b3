"""Developer measurement: S independent sequences (BASELINE.json configs[4]) fitted concurrently on ONE GPU, one fitter + warp field +
CUDA stream per sequence. Prints frames/s for device-resident and pinned-host depth frames as S grows.
  python tools/concurrent_sequences.py [rounds]
"""
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
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    from dynamicfuion_python_b200 import nnrt
    dev = torch.device("cuda", 0)
    s = bench.make_bench_scene()
    G = nnrt.geometry
    K, ext = s["K"], np.eye(4)
    mesh = (torch.tensor(s["positions"], device=dev), torch.tensor(s["normals"], device=dev), torch.tensor(s["triangles"], device=dev))
    N = s["nodes"].shape[0]

    def new_field():
        return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                            G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])

    wf = new_field()
    anchors, weights = wf.precompute_anchors_and_weights(mesh)
    frames_dev = []
    for scale in s["frame_scales"]:
        R, t = bench.scaled_motion(s, scale)
        wf.set_node_rotations(R); wf.set_node_translations(t)
        warped = wf.warp_mesh(mesh, anchors, weights, True, ext)
        fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(warped, K, (bench.HEIGHT, bench.WIDTH), 0.0, 10.0)
        _, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (bench.HEIGHT, bench.WIDTH), 0.0, 1, -1, -1, True, False, True)
        frames_dev.append(torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous())
    frames_host = [f.cpu().pin_memory() for f in frames_dev]
    eye = torch.eye(3, device=dev).repeat(N, 1, 1).contiguous()
    torch.cuda.synchronize()

    for S in (1, 2, 4, 8, 16, 32):
        seqs = []
        for k in range(S):
            seqs.append(dict(wf=new_field(), fitter=nnrt.alignment.DeformableMeshToImageFitter(
                max_iteration_count=bench.ITERATIONS_PER_FRAME, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"]),
                stream=torch.cuda.Stream(dev), host_R=torch.empty((N, 3, 3)).pin_memory(), host_t=torch.empty((N, 3)).pin_memory()))

        def run(frames, read_back, count):
            for r in range(count):
                for k, q in enumerate(seqs):
                    with torch.cuda.stream(q["stream"]):
                        if read_back:
                            q["stream"].synchronize()  # the previous frame's result of this sequence is on the host now
                        q["wf"].node_rotations.copy_(eye, non_blocking=True); q["wf"].node_translations.zero_()
                        q["fitter"].fit_to_image(q["wf"], mesh, None, frames[(r + k) % len(frames)], None, K, ext, 1.0)
                        if read_back:
                            q["host_R"].copy_(q["wf"].node_rotations, non_blocking=True)
                            q["host_t"].copy_(q["wf"].node_translations, non_blocking=True)
            submitted = time.perf_counter()
            torch.cuda.synchronize()
            return submitted

        out = {"sequences": S}
        for name, frames, rb in (("device", frames_dev, False), ("e2e", frames_host, True)):
            run(frames, rb, 3)
            t0 = time.perf_counter()
            submitted = run(frames, rb, rounds)
            dt = time.perf_counter() - t0
            out[name + "_frames_per_s"] = S * rounds / dt
            out[name + "_host_us_per_frame"] = (submitted - t0) / (S * rounds) * 1e6
        for q in seqs:
            q["fitter"].check_status()
        out["finite"] = bool(all(torch.isfinite(q["wf"].node_translations).all().item() for q in seqs))
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
