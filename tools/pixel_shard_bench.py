"""Developer/measurement tool for the pixel-sharded single-frame mode. Run under torchrun, one rank per GPU:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29621 tools/pixel_shard_bench.py
Every rank fits the SAME frame of the bench workload (C2r mesh and graph) at several image sizes; rank 0 prints one JSON line per size
with ms per GN iteration for the unsharded graph path (1 GPU) and for the sharded path over all ranks (max over ranks, CUDA events)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist

ITER = 10


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from dynamicfuion_python_b200 import nnrt, synthetic, frame_batch
    import bench
    G = nnrt.geometry
    for height, width in ((480, 640), (1080, 1920), (2160, 3840)):
        s = synthetic.make_scene("C2r", height, width)
        # the data term grows with the pixel count; keep its balance with the ARAP term and the damping of the 640x480 bench scene
        ratio = height * width / (480.0 * 640.0)
        s["arap_term_weight"] *= ratio ** 0.5
        s["lm_factor"] = min(1.0, s["lm_factor"] * ratio) if ratio <= 1.0 else 1.0
        K, ext = s["K"], np.eye(4)
        mesh = (torch.tensor(s["positions"], device=dev), torch.tensor(s["normals"], device=dev), torch.tensor(s["triangles"], device=dev))
        N = s["nodes"].shape[0]

        def field():
            return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
        wf = field()
        anchors, weights = wf.precompute_anchors_and_weights(mesh)
        R, t = bench.scaled_motion(s, 1.0)
        wf.set_node_rotations(R); wf.set_node_translations(t)
        warped = wf.warp_mesh(mesh, anchors, weights, True, ext)
        fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(warped, K, (height, width), 0.0, 10.0)
        _, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (height, width), 0.0, 1, -1, -1, True, False, True)
        depth = torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous()
        eye = torch.eye(3, device=dev).repeat(N, 1, 1).contiguous()
        results = {}
        for label, shard in (("graph_1gpu", False), (f"sharded_{world}gpu", True)):
            wf = field()
            fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=ITER, preconditioning_dampening_factor=s["lm_factor"],
                                                                arap_term_weight=s["arap_term_weight"])
            if shard:
                fitter.shard_pixels(rank, world)

            def frame():
                wf.node_rotations.copy_(eye); wf.node_translations.zero_()
                fitter.fit_to_image(wf, mesh, None, depth, None, K, ext, 1.0, anchors=anchors, anchor_weights=weights)
            for _ in range(3):
                frame()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10):
                frame()
            b.record(); torch.cuda.synchronize()
            try:
                fitter.check_status()
                status = "ok"
            except RuntimeError as error:
                status = str(error)
            ms = frame_batch.max_over_ranks([a.elapsed_time(b) / 10], dev)[0]
            results[label] = {"ms_per_frame": ms, "ms_per_iteration": ms / ITER, "status": status}
        if rank == 0:
            print(json.dumps({"image": f"{width}x{height}", "coverage": float((depth > 0).float().mean()), "iterations_per_frame": ITER,
                              "n_gpus": world, **results}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
