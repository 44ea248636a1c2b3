"""Per-frame device time of dynamicfuion_python_b200.fusion.FusionLoop (block selection -> fit -> integrate through the fitted field ->
extract -> warp) on the C2r relief at 640 x 480, and of its parts. Frames are rendered on the GPU from a smoothly growing node motion.
One JSON line. Host-side Python between the calls is inside the per-frame figure (CUDA events around whole frames, after a synchronise)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from dynamicfuion_python_b200 import nnrt, synthetic  # noqa: E402
from dynamicfuion_python_b200.fusion import FusionLoop  # noqa: E402


def main():
    H, W, frames = 480, 640, 8
    s = synthetic.make_scene("C2r", H, W)
    G, R = nnrt.geometry.functional, nnrt.rendering
    anchors, weights = G.compute_anchors_and_weights_euclidean_fixed_node_weight(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    identity = np.broadcast_to(np.eye(3, dtype=np.float32), s["rotations_gt"].shape).copy()

    def frame(k):
        a = k / float(frames)
        rotations = (identity + a * (s["rotations_gt"] - identity)).astype(np.float32)
        translations = (a * s["translations_gt"] + np.array([0, 0, 0.001 * k], np.float32)).astype(np.float32)
        warped = G.warp_triangle_mesh((s["positions"], s["normals"], s["triangles"]), s["nodes"], rotations, translations, anchors, weights)
        fv, mask = R.functional.get_mesh_ndc_face_vertices_and_clip_mask((warped.positions, None, s["triangles"]), s["K"], (H, W), 0.0, 10.0)
        _, depth, _, _ = R.rasterize_ndc_triangles(fv, mask, (H, W), 0.0, 1, -1, -1, True, False, True)
        return torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous()

    depths = [frame(k) for k in range(frames + 1)]
    if os.environ.get("NNRT_FUSION_SWEEP"):  # which tracker settings keep the reference iteration contractive on this scene
        for lm, arap, spacing in ((1.0, 20.0, 0.03), (1.0, 80.0, 0.03), (1.0, 200.0, 0.03), (1.0, 80.0, 0.04), (1.0, 80.0, 0.02), (0.1, 80.0, 0.03)):
            trial = FusionLoop(s["K"], voxel_size=0.004, block_resolution=8, block_count=8192, node_spacing=spacing, depth_scale=1.0, truncation_voxel_multiplier=3.0,
                               fitter=nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=10, preconditioning_dampening_factor=lm, arap_term_weight=arap))
            trial.first_frame(depths[0])
            out = []
            try:
                for k in range(1, frames + 1):
                    _, warped, _ = trial.next_frame(depths[k])
                    v = warped.positions
                    col, row = torch.round(v[:, 0] / v[:, 2] * s["K"][0, 0] + s["K"][0, 2]).long(), torch.round(v[:, 1] / v[:, 2] * s["K"][1, 1] + s["K"][1, 2]).long()
                    inside = (col >= 0) & (col < W) & (row >= 0) & (row < H)
                    d = depths[k][row[inside], col[inside]]
                    out.append(round(float((v[inside, 2] - d)[d > 0].abs().mean()) * 1e3, 3))
            except RuntimeError as error:
                out.append(str(error)[-50:])
            print("sweep", lm, arap, spacing, trial.graph.nodes.shape[0], out, flush=True)
        return
    loop = FusionLoop(s["K"], voxel_size=0.004, block_resolution=8, block_count=16384, node_spacing=0.02, depth_scale=1.0, truncation_voxel_multiplier=3.0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    canonical = loop.first_frame(depths[0])
    torch.cuda.synchronize()
    first_ms = (time.perf_counter() - t0) * 1e3
    per_frame, errors = [], []
    K = s["K"]
    for k in range(1, frames + 1):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        canonical, warped, _ = loop.next_frame(depths[k])
        torch.cuda.synchronize()
        per_frame.append((time.perf_counter() - t0) * 1e3)
        v = warped.positions
        col, row = torch.round(v[:, 0] / v[:, 2] * K[0, 0] + K[0, 2]).long(), torch.round(v[:, 1] / v[:, 2] * K[1, 1] + K[1, 2]).long()
        inside = (col >= 0) & (col < W) & (row >= 0) & (row < H)
        d = depths[k][row[inside], col[inside]]
        errors.append(float((v[inside, 2] - d)[d > 0].abs().mean()) * 1e3)
    print(json.dumps(dict(op="FusionLoop", scene=f"C2r relief, {W}x{H}, voxel 4 mm, blocks of 8^3", blocks=loop.volume.hashmap().size(),
                          nodes=int(loop.graph.nodes.shape[0]), canonical_vertices=int(canonical.positions.shape[0]), first_frame_ms=round(first_ms, 2),
                          frame_ms_median=round(float(np.median(per_frame[1:])), 3), frame_ms_all=[round(x, 3) for x in per_frame],
                          tracked_surface_error_mm=[round(e, 3) for e in errors],
                          note="wall clock around next_frame incl. host Python; 10 GN iterations per frame; every tensor stays on the device")), flush=True)


if __name__ == "__main__":
    main()
