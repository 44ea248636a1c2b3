"""bench.py — dense-depth GN iteration benchmark (BASELINE.json metric: ms per GN iteration @640x480; frames/s box-wide).

A frame fit of configuration C2 (bunny-scale relief mesh, ~82k triangles, 249 nodes = 208 stem + 41 corner, 832 edges, 640x480
depth) is the reference's once-per-fit setup (anchors, face-anchor association, reference unprojection;
DeformableMeshToImageFitter.cpp:96-106,293-307) plus ITERATIONS_PER_FRAME Gauss-Newton iterations (:111-275). A "step" is one frame
of each of `--sequences` independent sequences resident on the GPU (BASELINE.json configs[4]; each sequence has its own fitter, warp
field and CUDA stream, frames inside a sequence are serial). The single-sequence latency (ms per GN iteration, the <= 1 ms target)
is measured in a separate leg and reported as `ms_per_iteration` / `latency`.

  python bench.py --gpus N --steps K --warmup W              our sm_100a path (one process per GPU; frames are sharded, no collective)
  python bench.py --impl reference --gpus N --steps K ...    the reference algorithm's CPU path (oracle restatement, OpenMP, all host threads)

Prints ONE JSON line on rank 0. `value` = frames/s over all ranks with inputs resident in HBM; `e2e` = the same through the public
nnrt.alignment.DeformableMeshToImageFitter.fit_to_image call with the depth frame in pinned host memory (H2D in, node state D2H out).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ITERATIONS_PER_FRAME = 10
HEIGHT, WIDTH = 480, 640
WORKLOAD = "C2r"
FRAME_POOL = 8  # distinct target depth frames cycled through (each rank offsets its start)


def _env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line)."""
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_index), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); smax.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for name, flag in zip(names, parts[3:7]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)), "samples": len(sm),
                "reasons": sorted(reasons)}


def algorithmic_bytes(V, F, PX, N, E):
    """compulsory traffic of the API-visible data per GN iteration, by stage."""
    return {
        "warp": 80 * V,                       # pos+normal+anchors+weights in (56 B), warped pos+normal out (24 B)
        "raster": 97 * F + 28 * PX,           # int64 indices + NDC face record write/read, fragments written
        "assemble": 41 * PX + 168 * N,        # fragments + reference point + mask read, node state + J^T J + J^T r
        "solve": 300 * E,                     # edge indices + wing block written and read
        "total": 80 * V + 97 * F + 69 * PX + 168 * N + 300 * E,
    }


def make_bench_scene():
    from dynamicfuion_python_b200 import synthetic
    s = synthetic.make_scene(WORKLOAD, HEIGHT, WIDTH)
    s["frame_scales"] = [0.5 + 0.5 * (k + 1) / FRAME_POOL for k in range(FRAME_POOL)]
    return s


def scaled_motion(s, scale):
    """Ground-truth node motion scaled towards the identity (axis-angle scaling through Rodrigues on the host)."""
    from dynamicfuion_python_b200 import synthetic
    R, t = synthetic.smooth_node_motion(s["nodes"], max_angle_deg=3.0 * scale, max_translation=0.005 * scale,
                                        rng=np.random.default_rng(synthetic.SEED + 1))
    return R, t


STAGES = ["warp_vertices", "face_raster", "band_raster", "assemble", "build_system", "stem_eliminate", "corner_solve_update"]
STAGE_KERNEL = {"warp_vertices": "warp_vertices_kernel<4>", "face_raster": "face_raster_kernel", "band_raster": "band_raster_kernel",
                "assemble": "assemble_kernel<4>", "build_system": "system_stem_kernel", "stem_eliminate": "(fused into system_stem_kernel)",
                "corner_solve_update": "corner_solve_regtile_kernel"}
SETUP_LAUNCHES = 5  # anchors, 3 pack kernels, reference unprojection


def run_b200(args):
    import torch
    import torch.distributed as dist
    rank, world, local = _env_int("RANK", 0), _env_int("WORLD_SIZE", 1), _env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl b200) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from dynamicfuion_python_b200 import nnrt
    dev = torch.device("cuda", local)
    s = make_bench_scene()
    G = nnrt.geometry
    K, ext = s["K"], np.eye(4)
    mesh = (torch.tensor(s["positions"], device=dev), torch.tensor(s["normals"], device=dev), torch.tensor(s["triangles"], device=dev))
    V, F, N, E, PX = s["positions"].shape[0], s["triangles"].shape[0], s["nodes"].shape[0], s["edges"].shape[0], HEIGHT * WIDTH

    def new_field():
        return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                            G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])

    # target depth frames rendered by the CUDA rasteriser from scaled ground-truth motion (synthetic data, seeded)
    wf = new_field()
    anchors, weights = wf.precompute_anchors_and_weights(mesh)
    frames_dev = []
    for scale in s["frame_scales"]:
        R, t = scaled_motion(s, scale)
        wf.set_node_rotations(R); wf.set_node_translations(t)
        warped = wf.warp_mesh(mesh, anchors, weights, True, ext)
        fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(warped, K, (HEIGHT, WIDTH), 0.0, 10.0)
        _, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (HEIGHT, WIDTH), 0.0, 1, -1, -1, True, False, True)
        frames_dev.append(torch.where(depth[..., 0] > 0, depth[..., 0], torch.zeros_like(depth[..., 0])).contiguous())
    frames_host = [f.cpu().pin_memory() for f in frames_dev]
    coverage = float((frames_dev[0] > 0).float().mean())

    from dynamicfuion_python_b200 import frame_batch
    S = args.sequences
    eye = torch.eye(3, device=dev).repeat(N, 1, 1).contiguous()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    # one warp field + fitter + stream + pinned result buffers per resident sequence (BASELINE.json configs[4]: independent sequences)
    seqs = [dict(wf=new_field(),
                 fitter=nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=ITERATIONS_PER_FRAME,
                                                                   preconditioning_dampening_factor=s["lm_factor"],
                                                                   arap_term_weight=s["arap_term_weight"]),
                 host_R=torch.empty((N, 3, 3), dtype=torch.float32).pin_memory(),
                 host_t=torch.empty((N, 3), dtype=torch.float32).pin_memory()) for _ in range(S)]
    pool = frame_batch.SequenceStreams(S, dev)
    wf, fitter = seqs[0]["wf"], seqs[0]["fitter"]

    def fit(q, depth, read_back):
        q["wf"].node_rotations.copy_(eye); q["wf"].node_translations.zero_()
        q["fitter"].fit_to_image(q["wf"], mesh, None, depth, None, K, ext, 1.0)
        if read_back:  # result of the frame back on the host: two asynchronous copies into pinned buffers on the sequence's stream
            q["host_R"].copy_(q["wf"].node_rotations, non_blocking=True)
            q["host_t"].copy_(q["wf"].node_translations, non_blocking=True)

    def step(frames, index, read_back):
        """One frame of every resident sequence, each on its own stream. With read_back the host holds a sequence's previous result
        (stream synchronised) before that sequence's next frame is enqueued."""
        for k, q in enumerate(seqs):
            if read_back:
                pool.synchronize(k)
            pool.run(k, fit, q, frames[(first + index + k) % FRAME_POOL], read_back)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    first = rank % FRAME_POOL
    for i in range(args.warmup):
        step(frames_dev, i, False)
    torch.cuda.synchronize()
    for q in seqs:
        q["fitter"].check_status()

    # ---- value: inputs resident in HBM, one CUDA event pair on the launching stream around the K steps (sequence streams forked from /
    # joined into it); the resident sequences' working sets (S x workspace) exceed L2, so a sequence finds its data evicted each frame
    sampler = ClockSampler(local)
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if rank == 0:
        sampler.start()
    start.record()
    pool.fork()
    for i in range(args.steps):
        step(frames_dev, i, False)
    pool.join()
    end.record()
    barrier()
    dev_ms = start.elapsed_time(end)
    for q in seqs:
        q["fitter"].check_status()
        if not bool(torch.isfinite(q["wf"].node_translations).all()):
            raise RuntimeError("bench: fit diverged (non-finite node translations)")

    # ---- e2e: public API call with the depth frame in pinned host memory + D2H of the node state, wall clock around K steps
    for i in range(min(args.warmup, 3)):
        step(frames_host, i, True)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(frames_host, i, True)
    pool.synchronize()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    # ---- latency: ONE sequence alone on the GPU, L2 flushed between frames, CUDA events per frame (the <= 1 ms / iteration target)
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    barrier()
    with torch.cuda.stream(pool.streams[0]):  # the stream this sequence's graph was captured for
        for i in range(args.steps):
            flush.fill_(i & 0xFF)
            starts[i].record()
            fit(seqs[0], frames_dev[(first + i) % FRAME_POOL], False)
            ends[i].record()
    barrier()
    latency_ms = sum(a.elapsed_time(b) for a, b in zip(starts, ends)) / args.steps

    dev_ms, e2e_ms, latency_ms = frame_batch.max_over_ranks([dev_ms, e2e_ms, latency_ms], dev)  # frames are sharded, no data-path collective: only the clock is reduced

    # ---- per-kernel stage times (CUDA events between the 7 launches of an iteration, outside the graph) for the roofline
    stage_ms = np.zeros(7)
    profiled = 0
    if rank == 0:
        for rep in range(4):
            wf.node_rotations.copy_(eye); wf.node_translations.zero_()
            for it in range(ITERATIONS_PER_FRAME):
                flush.fill_(it)
                torch.cuda.synchronize()
                stage_ms += np.array(fitter.profile_iteration(wf, K, ext))
                profiled += 1
        stage_ms /= profiled
    out = None
    if rank == 0:
        out = dict(dev_ms=dev_ms, e2e_ms=e2e_ms, latency_ms=latency_ms, sequences=S, stage_ms=stage_ms.tolist(), clocks=clocks, coverage=coverage, dims=(V, F, PX, N, E),
                   world=world, launches=int(fitter.launches_per_iteration()), workspace=int(fitter.workspace_bytes()) * S)
    if world > 1:
        dist.destroy_process_group()
    return out


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def oracle_frame(oracle, s, depth, iterations):
    """One frame of the reference algorithm on the host cores (oracle restatement, OpenMP): per-fit setup + `iterations` GN iterations.
    Returns (setup_seconds, iteration_seconds_total)."""
    t0 = time.perf_counter()
    anchors, weights = oracle.compute_anchors(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    ref_pts, ref_mask = oracle.unproject_depth(depth, s["K"], None, 1.0, 10.0)
    t1 = time.perf_counter()
    oracle.fit_to_image(s["positions"], s["normals"], s["triangles"], s["nodes"], s["edges"], s["arrow_base"], anchors, weights, ref_pts, ref_mask,
                        HEIGHT, WIDTH, s["K"], iteration_count=iterations, lm_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"],
                        edge_layer_indices=s["edge_layer_indices"], layer_radii=s["layer_radii"], parallel=True, trace=False)
    t2 = time.perf_counter()
    return t1 - t0, t2 - t1


def oracle_target_depth(oracle, s, scale):
    R, t = scaled_motion(s, scale)
    anchors, weights = oracle.compute_anchors(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    wp, _ = oracle.warp(s["positions"], s["normals"], s["nodes"], R, t, anchors, weights)
    fv, mask = oracle.face_vertices_ndc(wp, s["triangles"], s["K"], HEIGHT, WIDTH, 0.0, 10.0)
    _, depth, _, _, _ = oracle.rasterize(fv, mask, HEIGHT, WIDTH, 0.0, 1, perspective_correct=True)
    d = depth[..., 0].copy()
    d[d < 0] = 0
    return d


def cpu_baseline_sample(s, iterations):
    """Bounded CPU sample: one frame, `iterations` GN iterations timed, extrapolated to ITERATIONS_PER_FRAME."""
    from oracle import oracle_py as oracle
    oracle.lib()
    oracle.set_num_threads(host_cores())  # torchrun exports OMP_NUM_THREADS=1
    depth = oracle_target_depth(oracle, s, s["frame_scales"][0])
    oracle_frame(oracle, s, depth, 1)  # warm-up (page faults, OpenMP pool)
    setup_s, iter_s = oracle_frame(oracle, s, depth, iterations)
    per_iter = iter_s / iterations
    frame_s = setup_s + per_iter * ITERATIONS_PER_FRAME
    return dict(value=1.0 / frame_s, unit="frames/s", cores=host_cores(), kind="port",
                sample=f"1 frame of {WORKLOAD} @640x480 on '{cpu_model()}': per-fit setup {setup_s * 1e3:.1f} ms + {iterations} GN iterations timed "
                       f"({per_iter * 1e3:.1f} ms/iteration), extrapolated to {ITERATIONS_PER_FRAME} iterations/frame; oracle restatement of the "
                       "reference's Open3D-CPU path, g++ -O2 -fopenmp",
                ms_per_iteration=per_iter * 1e3)


def run_reference(args):
    """--impl reference: the reference algorithm's CPU path (oracle port; the reference itself needs Open3D 0.17 and cannot be built here)."""
    rank = _env_int("RANK", 0)
    if rank != 0:
        return None
    from oracle import oracle_py as oracle
    oracle.lib()
    threads = oracle.set_num_threads(host_cores())  # torchrun exports OMP_NUM_THREADS=1
    s = make_bench_scene()
    depth = oracle_target_depth(oracle, s, s["frame_scales"][0])
    _, probe = oracle_frame(oracle, s, depth, 1)
    budget = 150.0
    iters = int(max(1, min(ITERATIONS_PER_FRAME, budget / max(1e-3, (args.steps + args.warmup) * probe))))
    for _ in range(args.warmup):
        oracle_frame(oracle, s, depth, iters)
    total_setup = total_iter = 0.0
    for _ in range(args.steps):
        a, b = oracle_frame(oracle, s, depth, iters)
        total_setup += a; total_iter += b
    per_iter = total_iter / (args.steps * iters)
    frame_s = total_setup / args.steps + per_iter * ITERATIONS_PER_FRAME
    value = 1.0 / frame_s
    sample = (f"each step = per-fit setup + {iters} of {ITERATIONS_PER_FRAME} GN iterations of {WORKLOAD} @640x480, extrapolated to a full frame; "
              f"{threads} threads on '{cpu_model()}'")
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": frame_s * 1e3 * args.sequences * args.gpus, "frames_per_step": args.sequences * args.gpus, "ms_per_frame": frame_s * 1e3,
            "ms_per_iteration": per_iter * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": bench_config(args.gpus, args.sequences),
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": host_cores(), "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}


METRIC = "frames/s box-wide (frame = per-fit setup + 10 dense-depth GN iterations @640x480); ms per GN iteration in ms_per_iteration"


def bench_config(gpus, sequences):
    return {"workload": f"{WORKLOAD}: relief mesh V=41209 F=81608, 249 nodes (208 stem + 41 corner, 1494x1494 arrowhead system), 832 edges, "
                        f"{WIDTH}x{HEIGHT} depth, {ITERATIONS_PER_FRAME} GN iterations/frame (BASELINE.json configs[1]); a step = one frame of each "
                        f"of {sequences} independent sequences per GPU (configs[4])",
            "parallelism": f"frame-batch: {gpus} GPU(s) x {sequences} resident sequences, one fitter + CUDA stream per sequence, no collective",
            "sequences_per_gpu": sequences,
            "l2": f"throughput legs: the {sequences} resident sequences' working sets ({sequences} x 23 MB) exceed the 126 MB L2, every sequence "
                  "finds its data evicted at its next frame; latency leg: 256 MiB device fill between frames (outside the per-frame CUDA events)",
            "frames": f"{FRAME_POOL} distinct synthetic target depth frames cycled"}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sequences", type=int, default=16, help="independent sequences resident per GPU, each on its own stream")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the bounded CPU sample (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 0)
    if args.impl == "reference":
        line = run_reference(args)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    r = run_b200(args)
    if r is None:
        return
    V, F, PX, N, E = r["dims"]
    world = r["world"]
    S = r["sequences"]
    frames = args.steps * world * S
    value = frames / (r["dev_ms"] * 1e-3)
    e2e_value = frames / (r["e2e_ms"] * 1e-3)
    ms_per_step = r["dev_ms"] / args.steps
    stage_ms = dict(zip(STAGES, r["stage_ms"]))
    bytes_by = algorithmic_bytes(V, F, PX, N, E)
    stage_bytes = {"warp_vertices": bytes_by["warp"], "face_raster": bytes_by["raster"], "band_raster": 0, "assemble": bytes_by["assemble"],
                   "build_system": 168 * N + 144 * E, "stem_eliminate": 156 * E + 168 * N, "corner_solve_update": 4 * (6 * (N - 208)) ** 2 + 72 * N}
    dominant = max(STAGES, key=lambda k: stage_ms[k])
    peak, peak_kind = load_peaks()
    achieved = stage_bytes[dominant] / (stage_ms[dominant] * 1e-3) / 1e9 if stage_ms[dominant] > 0 else 0.0
    iter_ms_profiled = sum(stage_ms.values())
    traffic = None
    try:  # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture (profiles/ncu_traffic.json)
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            traffic = json.load(fh).get(STAGE_KERNEL[dominant].split("(")[0], {}).get("dram_bytes_per_launch")
    except (OSError, ValueError):
        pass
    roofline = {"bound": "hbm", "kernel": STAGE_KERNEL[dominant], "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_kind": peak_kind, "algorithmic_bytes_per_launch": stage_bytes[dominant],
                "avg_launch_ms": stage_ms[dominant], "share_of_iteration": stage_ms[dominant] / iter_ms_profiled if iter_ms_profiled else None,
                "stage_ms": stage_ms, "iteration_algorithmic_bytes": bytes_by["total"],
                "iteration_achieved_gbs": bytes_by["total"] / (iter_ms_profiled * 1e-3) / 1e9 if iter_ms_profiled else None}
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "frames_per_step": S * world, "ms_per_iteration": r["latency_ms"] / ITERATIONS_PER_FRAME,
            "latency": {"ms_per_frame": r["latency_ms"], "ms_per_iteration": r["latency_ms"] / ITERATIONS_PER_FRAME,
                        "frames_per_s_one_sequence": 1e3 / r["latency_ms"],
                        "note": "one sequence alone on the GPU, L2 flushed between frames, CUDA events per frame (per-fit setup included)"},
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": bench_config(world, S),
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": PX * 4 * S, "d2h_bytes_per_step": N * 12 * 4 * S,
                    "ms_per_step": r["e2e_ms"] / args.steps, "timing": "wall clock around K steps of fit_to_image calls (one per resident sequence), synchronize + barrier on both sides, max over ranks"},
            "gpu_launches": args.steps * S * (SETUP_LAUNCHES + ITERATIONS_PER_FRAME * r["launches"]),
            "clocks": r["clocks"], "roofline": roofline, "coverage": r["coverage"], "workspace_bytes": r["workspace"]}
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_sample(make_bench_scene(), 3)
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
