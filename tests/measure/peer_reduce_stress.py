"""Stress of the peer reduction (run under torchrun, one rank per GPU): many short fits in a row on a small frame, where an iteration is
a few tens of microseconds and the ranks drift against each other as much as they ever will; every fit must equal the single-GPU fit
bit for bit on every rank. Random host-side delays between fits de-synchronise the ranks on purpose. Prints one JSON line on rank 0."""
import json
import os
import random
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from dynamicfuion_python_b200 import nnrt
    from util import scene_with_gpu_target
    fits = int(os.environ.get("NNRT_STRESS_FITS", "300"))
    G = nnrt.geometry
    random.seed(1000 + rank)
    mismatches, total_iterations = 0, 0
    for config, h, w, iterations in (("tinyr", 120, 160, 7), ("C2r", 480, 640, 10)):
        s = scene_with_gpu_target(nnrt, config, h, w)
        mesh = (s["positions"], s["normals"], s["triangles"])

        def field():
            return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                           G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
        args = dict(max_iteration_count=iterations, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
        single = nnrt.alignment.DeformableMeshToImageFitter(**args)
        wf = field()
        single.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (h, w), anchors=s["anchors"], anchor_weights=s["anchor_weights"])
        single.check_status()
        want_R, want_t = wf.node_rotations.clone(), wf.node_translations.clone()
        sharded = nnrt.alignment.DeformableMeshToImageFitter(**args)
        sharded.shard_pixels(rank, world, stripe_rows=8, collective="peer")
        wf = field()
        eye = torch.eye(3, device="cuda").expand_as(wf.node_rotations).contiguous()
        for k in range(fits):
            wf.node_rotations.copy_(eye)
            wf.node_translations.zero_()
            if random.random() < 0.3:
                time.sleep(random.random() * 0.002)  # let this rank fall behind (or the others run ahead)
            sharded.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (h, w), anchors=s["anchors"],
                                 anchor_weights=s["anchor_weights"])
            if k % 25 == 24 or k == fits - 1:
                sharded.check_status()
            if not (torch.equal(wf.node_rotations, want_R) and torch.equal(wf.node_translations, want_t)):
                mismatches += 1
            total_iterations += iterations
        assert sharded._shard_collective == "peer" and sharded._peer_ready
        sharded.shard_pixels(0, 1)
    counts = torch.tensor([mismatches], device="cuda")
    dist.all_reduce(counts)
    if rank == 0:
        print(json.dumps(dict(op="peer reduction stress", ranks=world, fits_per_scene=fits, iterations_per_rank=total_iterations,
                              fits_that_differ_from_one_gpu=int(counts.item()))), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if int(counts.item()):
        sys.exit(1)


if __name__ == "__main__":
    main()
