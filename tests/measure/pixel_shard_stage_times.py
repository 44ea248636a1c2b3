"""Per-stage device times of one un-captured iteration of BASELINE.json configs[3] (the C2 mesh at 3840 x 2160) for the whole image
and for the share of one rank under the two pixel splits (contiguous bands, round-robin stripes): what part of the data term shrinks
with the number of GPUs and what part is replicated. One JSON line per case; runs on ONE GPU (the ranks' shares one after the other)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from dynamicfuion_python_b200 import nnrt, pixel_shard  # noqa: E402
from dynamicfuion_python_b200.nnrt import _tensor as T  # noqa: E402
from util import scene_with_gpu_target  # noqa: E402

STAGES = ["warp_vertices", "face_raster", "band_raster", "assemble", "build_system", "stem_eliminate", "corner_solve_update"]


def main():
    h, w = 2160, 3840
    s = scene_with_gpu_target(nnrt, "C2r", h, w)
    G = nnrt.geometry
    mesh = (s["positions"], s["normals"], s["triangles"])

    def field():
        return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                       G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
    fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
    fitter.fit_to_image(field(), mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (h, w), anchors=s["anchors"],
                        anchor_weights=s["anchor_weights"])
    fitter.check_status()

    def profile(label):
        best = None
        for _ in range(5):
            ms = fitter.profile_iteration(field(), s["K"])
            best = ms if best is None else [min(a, b) for a, b in zip(best, ms)]
        print(json.dumps(dict(case=label, image=f"{w}x{h}", data_term_ms=round(sum(best[:4]), 4), solve_ms=round(sum(best[4:]), 4),
                              stage_ms={k: round(v, 4) for k, v in zip(STAGES, best)})), flush=True)

    profile("whole image")
    for world in (2, 4, 8):
        for rank in sorted({0, world // 2}):
            begin, end = pixel_shard.pixel_rows(rank, world, h)
            T.call("nnrt_b200_fitter_set_pixel_rows", fitter._handle, begin, end)
            T.call("nnrt_b200_fitter_set_pixel_stripes", fitter._handle, 8, 0, 1)
            profile(f"band {rank} of {world} (rows {begin}-{end})")
            T.call("nnrt_b200_fitter_set_pixel_rows", fitter._handle, 0, h)
            for rows in (8, 16, 32, 64):
                T.call("nnrt_b200_fitter_set_pixel_stripes", fitter._handle, rows, rank, world)
                profile(f"{rows}-row stripes, rank {rank} of {world}")


if __name__ == "__main__":
    main()
