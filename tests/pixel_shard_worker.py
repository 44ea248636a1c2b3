"""Worker of test_two_gpu_pixel_sharded_fit_matches_single_gpu (run under torchrun, one rank per GPU)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from dynamicfuion_python_b200 import nnrt
    from oracle import oracle_py as oracle
    from util import scene_with_target
    s = scene_with_target(oracle, "tinyr", 120, 160)
    G = nnrt.geometry
    mesh = (s["positions"], s["normals"], s["triangles"])

    def fit(shard, collective="nccl", stripe_rows=None):
        wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                          G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
        fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=3, preconditioning_dampening_factor=s["lm_factor"],
                                                            arap_term_weight=s["arap_term_weight"])
        if shard:
            fitter.shard_pixels(rank, world, stripe_rows=stripe_rows, collective=collective)
        fitter.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (s["height"], s["width"]),
                            anchors=s["anchors"], anchor_weights=s["anchor_weights"])
        fitter.check_status()
        first = (wf.node_rotations.clone(), wf.node_translations.clone())
        if shard:
            # the second fit of the same key records the loop (NCCL all-reduce included) into a CUDA graph and replays it, the third replays
            for _ in range(2):
                wf.node_rotations.copy_(torch.eye(3, device=wf.node_rotations.device).expand_as(wf.node_rotations))
                wf.node_translations.zero_()
                fitter.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (s["height"], s["width"]),
                                    anchors=s["anchors"], anchor_weights=s["anchor_weights"])
                fitter.check_status()
                assert torch.equal(wf.node_rotations, first[0]) and torch.equal(wf.node_translations, first[1]), "graph replay differs from the eager loop"
            if collective == "nccl":
                assert len(fitter._shard_graphs) == 1, "the sharded loop was not captured"
            else:
                assert fitter._shard_collective == "peer" and fitter._peer_ready, "the peer reduction fell back to the all-reduce"
        return first

    R1, t1 = fit(False)
    # bands + NCCL all-reduce, stripes + NCCL, stripes + the reduction fused into the stem kernel over peer memory (the default)
    for collective, stripe_rows in (("nccl", None), ("nccl", 8), ("peer", 8), ("peer", 32)):
        R2, t2 = fit(True, collective, stripe_rows)
        if os.environ.get("NNRT_B200_WORKER_TRACE"):
            torch.cuda.synchronize()
            print(f"[rank {rank}] {collective} stripe_rows={stripe_rows} done, probe {float(torch.zeros(1, device='cuda').sum())}", flush=True)
        # integer (fixed-point) sums added as int64 words: the sharded fit equals the single-GPU fit bit for bit
        assert torch.equal(t2, t1), (collective, stripe_rows, float((t2 - t1).abs().max()))
        assert torch.equal(R2, R1), (collective, stripe_rows, float((R2 - R1).abs().max()))
        # identical state on every rank (replicated solve on identical sums)
        gathered = [torch.zeros_like(t2) for _ in range(world)]
        dist.all_gather(gathered, t2)
        assert all(torch.equal(g, gathered[0]) for g in gathered)
    # the same through the C++ front end (pybind11 module `nnrt`): stripes + peer reduction, three fits in a row on one workspace
    from dynamicfuion_python_b200 import build
    cxx = build.import_pybind_module()
    GX = cxx.geometry
    for stripe_rows in (8, 32):
        wf = GX.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                      GX.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
        fitter = cxx.alignment.DeformableMeshToImageFitter(max_iteration_count=3, preconditioning_dampening_factor=s["lm_factor"],
                                                           arap_term_weight=s["arap_term_weight"])
        fitter.shard_pixels(rank, world, None, stripe_rows)
        for repeat in range(3):
            wf.set_node_rotations(torch.eye(3, device="cuda").repeat(len(s["nodes"]), 1, 1), True)
            wf.set_node_translations(torch.zeros((len(s["nodes"]), 3), device="cuda"), True)
            fitter.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (s["height"], s["width"]),
                                anchors=s["anchors"], anchor_weights=s["anchor_weights"])
            fitter.check_status()
            R3, t3 = wf.get_node_rotations(True), wf.get_node_translations(True)
            assert torch.equal(t3, t1) and torch.equal(R3, R1), ("pybind11 module", stripe_rows, repeat, float((t3 - t1).abs().max()))
        fitter.shard_pixels(0, 1)
        del fitter
    dist.barrier()
    if rank == 0:
        print("PIXEL_SHARD_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
