"""Pixel-sharded single-frame mode: two bands on one GPU summed by hand reproduce the single-GPU iteration bit for bit;
the two-process NCCL path is exercised when two GPUs are visible."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from util import rel_close, scene_with_target

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


def _field(nnrt, s):
    G = nnrt.geometry
    return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                        G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])


@pytest.mark.parametrize("config,h,w,bands,stripe_rows", [("tinyr", 120, 160, 2, None), ("C2r", 2160, 3840, 2, None), ("C2r", 2160, 3840, 3, None),
                                                          ("tinyr", 120, 160, 3, 8), ("tinyr", 120, 160, 2, 16), ("C2r", 480, 640, 8, 8),
                                                          ("C2r", 2160, 3840, 4, 16), ("C2r", 2160, 3840, 8, 16), ("C2r", 2160, 3840, 5, 40)])
def test_bands_on_one_gpu_equal_whole_image(nnrt_mirror, oracle, config, h, w, bands, stripe_rows):
    """The ranks' bands are run one after the other through the split-iteration C ABI with the accumulators summed in between (int64
    words): J^T J / J^T r and the update must equal the unsharded iteration exactly. The 3840 x 2160 cases are BASELINE.json configs[3]
    (the C2 mesh at 4K) split in two and in three uneven bands, and in round-robin stripes for 4, 8 and 5 ranks (40-row stripes leave a
    partial last stripe); their target is rendered on the GPU (the oracle's renderer takes minutes there). The printed per-rank times
    show the load balance of a split (the slowest rank sets the pace of a sharded iteration)."""
    nnrt = nnrt_mirror  # drives the split-iteration C ABI through the ctypes fitter handle
    from dynamicfuion_python_b200 import _lib, pixel_shard
    from dynamicfuion_python_b200.nnrt import _tensor as T
    from util import scene_with_gpu_target
    import ctypes as C
    s = scene_with_target(oracle, config, h, w) if h <= 480 else scene_with_gpu_target(nnrt, config, h, w)
    mesh = (s["positions"], s["normals"], s["triangles"])
    args = dict(max_iteration_count=1, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
    call = dict(anchors=s["anchors"], anchor_weights=s["anchor_weights"])
    whole = nnrt.alignment.DeformableMeshToImageFitter(**args)
    wf = _field(nnrt, s)
    whole.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (s["height"], s["width"]), **call)
    whole.check_status()
    ref = whole.diagnostics()

    fitter = nnrt.alignment.DeformableMeshToImageFitter(**args)
    wf2 = _field(nnrt, s)
    fitter.fit_to_image(wf2, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (s["height"], s["width"]),
                        iteration_count=1, **call)  # sets model + reference; result discarded
    wf2 = _field(nnrt, s)
    R, t = wf2.node_rotations, wf2.node_translations
    K, ext = T.host_f64(s["K"], (3, 3)), T.host_f64(np.eye(4), (4, 4))
    count = C.c_int64()
    pointer = _lib.lib().nnrt_b200_fitter_accumulators(fitter._handle, C.byref(count))
    acc = pixel_shard.DeviceBuffer(pointer, count.value).tensor(R.device)
    total = torch.zeros_like(acc)
    covered = np.zeros((s["height"],), bool)
    times = []
    for rank in range(bands):
        if stripe_rows is None:  # one contiguous band per rank
            ranges = [pixel_shard.pixel_rows(rank, bands, s["height"])]
            T.call("nnrt_b200_fitter_set_pixel_rows", fitter._handle, ranges[0][0], ranges[0][1])
        else:  # stripes of stripe_rows rows dealt round-robin
            ranges = pixel_shard.stripe_rows_of(rank, bands, s["height"], stripe_rows)
            T.call("nnrt_b200_fitter_set_pixel_stripes", fitter._handle, stripe_rows, rank, bands)
        for begin, end in ranges:
            assert not covered[begin:end].any()
            covered[begin:end] = True
        acc.zero_()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        T.call("nnrt_b200_fitter_iteration_data_term", fitter._handle, T.ptr(R), T.ptr(t), K.data_ptr(), ext.data_ptr(), T.stream())
        stop.record()
        total += acc
        stop.synchronize()
        times.append(start.elapsed_time(stop))
    assert covered.all()
    print(f"\n{config} {w}x{h}, {bands} {'bands' if stripe_rows is None else f'x {stripe_rows}-row stripes'}: data term per rank (ms, un-captured) "
          + " ".join(f"{x:.3f}" for x in times) + f"; max / mean = {max(times) / (sum(times) / len(times)):.2f}")
    T.call("nnrt_b200_fitter_set_pixel_rows", fitter._handle, 0, s["height"])
    T.call("nnrt_b200_fitter_set_pixel_stripes", fitter._handle, 8, 0, 1)
    acc.copy_(total)
    T.call("nnrt_b200_fitter_iteration_solve", fitter._handle, T.ptr(R), T.ptr(t), K.data_ptr(), ext.data_ptr(), 0, T.stream())
    fitter.check_status()
    d = fitter.diagnostics()
    # the per-node sums are integers (103-bit fixed point) and every warp of a band sees the same 32 pixels as in the whole image:
    # the sharded iteration is BIT-IDENTICAL to the unsharded one
    assert np.array_equal(_np(d["hessian_diagonal"]), _np(ref["hessian_diagonal"]))
    assert np.array_equal(_np(d["negative_gradient"]), _np(ref["negative_gradient"]))
    assert np.array_equal(_np(d["deltas"]), _np(ref["deltas"]))
    assert np.array_equal(_np(t), _np(wf.node_translations)) and np.array_equal(_np(R), _np(wf.node_rotations))


@pytest.mark.parametrize("config,h,w", [("tinyr", 120, 160), ("C2r", 480, 640)])
def test_fit_is_bit_identical_run_to_run(nnrt, oracle, config, h, w):
    """No floating-point atomics anywhere on the iteration: per-node sums are integer (fixed point), the Schur complement is gathered
    in a fixed order. Ten iterations repeated five times (fresh fitter each time) give the same bits."""
    s = scene_with_target(oracle, config, h, w)
    mesh = (s["positions"], s["normals"], s["triangles"])
    results = []
    for rep in range(5):
        fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=10, preconditioning_dampening_factor=s["lm_factor"],
                                                            arap_term_weight=s["arap_term_weight"])
        wf = _field(nnrt, s)
        fitter.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (h, w), anchors=s["anchors"],
                            anchor_weights=s["anchor_weights"])
        d = fitter.diagnostics()
        results.append((_np(wf.node_rotations).copy(), _np(wf.node_translations).copy(), _np(d["hessian_diagonal"]).copy(), _np(d["deltas"]).copy()))
    for r in results[1:]:
        for a, b in zip(r, results[0]):
            assert np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("config,h,w,ranks,path", [("tinyr", 120, 160, 2, "fit"), ("tinyr", 120, 160, 3, "split"), ("C2r", 480, 640, 2, "fit"),
                                                   ("C2r", 480, 640, 3, "split")])
def test_peer_reduce_protocol_on_one_gpu(nnrt_mirror, oracle, config, h, w, ranks, path):
    """The peer reduction (sums exchanged by the stem kernel itself: flags + peer loads, two buffers alternating by iteration) with the
    ranks played by fitters of ONE process on one GPU, each on its own stream -- the exchange buffers are then ordinary device pointers
    (nnrt_b200_fitter_enable_peer_reduce) where separate processes pass IPC handles. Ten iterations, submitted rank after rank, so the
    kernels of different ranks really wait for each other: every rank ends bit-identical to the unsharded fit. "fit" = the library's
    captured loop (one call per rank), "split" = the two halves of each iteration launched eagerly. (Two or three ranks only: the
    streams of one process share the device's eight hardware queues, and a rank queued behind another rank's waiting kernel would run
    into the 4 s wait limit -- an artefact of playing all ranks on one device, which one process per GPU does not have.)"""
    nnrt = nnrt_mirror
    from dynamicfuion_python_b200.nnrt import _tensor as T
    import ctypes as C
    s = scene_with_target(oracle, config, h, w)
    mesh = (s["positions"], s["normals"], s["triangles"])
    args = dict(max_iteration_count=10, preconditioning_dampening_factor=s["lm_factor"], arap_term_weight=s["arap_term_weight"])
    call = dict(anchors=s["anchors"], anchor_weights=s["anchor_weights"])
    whole = nnrt.alignment.DeformableMeshToImageFitter(**args)
    wf = _field(nnrt, s)
    whole.fit_to_image(wf, mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (h, w), **call)
    whole.check_status()
    K, ext = T.host_f64(s["K"], (3, 3)), T.host_f64(np.eye(4), (4, 4))
    fitters, fields, pointers = [], [], (C.c_void_p * ranks)()
    for rank in range(ranks):
        fitter = nnrt.alignment.DeformableMeshToImageFitter(**args)
        fitter.fit_to_image(_field(nnrt, s), mesh, None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (h, w), iteration_count=1, **call)
        fitter.check_status()
        pointer = C.c_void_p()
        T.call("nnrt_b200_fitter_peer_exchange", fitter._handle, C.byref(pointer), None, None)
        pointers[rank] = pointer.value
        fitters.append(fitter)
        fields.append(_field(nnrt, s))
    for rank, fitter in enumerate(fitters):
        T.call("nnrt_b200_fitter_enable_peer_reduce", fitter._handle, rank, ranks, pointers)
        T.call("nnrt_b200_fitter_set_pixel_stripes", fitter._handle, 8, rank, ranks)
    streams = [torch.cuda.Stream() for _ in range(ranks)]
    torch.cuda.synchronize()
    for repeat in range(2):  # the second pass replays each rank's captured graph; the iteration counters keep counting across fits
        for f in fields:
            f.node_rotations.copy_(torch.eye(3, device="cuda").expand_as(f.node_rotations))
            f.node_translations.zero_()
        torch.cuda.synchronize()
        if path == "fit":
            for fitter, f, stream in zip(fitters, fields, streams):
                T.call("nnrt_b200_fitter_fit", fitter._handle, T.ptr(f.node_rotations), T.ptr(f.node_translations), K.data_ptr(), ext.data_ptr(), 10,
                       C.c_void_p(stream.cuda_stream))
        else:
            for it in range(10):
                for fitter, f, stream in zip(fitters, fields, streams):
                    T.call("nnrt_b200_fitter_iteration_data_term", fitter._handle, T.ptr(f.node_rotations), T.ptr(f.node_translations), K.data_ptr(),
                           ext.data_ptr(), C.c_void_p(stream.cuda_stream))
                for fitter, f, stream in zip(fitters, fields, streams):
                    T.call("nnrt_b200_fitter_iteration_solve", fitter._handle, T.ptr(f.node_rotations), T.ptr(f.node_translations), K.data_ptr(),
                           ext.data_ptr(), it, C.c_void_p(stream.cuda_stream))
        torch.cuda.synchronize()
        for fitter, f in zip(fitters, fields):
            fitter.check_status()
            assert np.array_equal(_np(f.node_translations), _np(wf.node_translations)) and np.array_equal(_np(f.node_rotations), _np(wf.node_rotations))
    for fitter in fitters:
        T.call("nnrt_b200_fitter_disable_peer_reduce", fitter._handle)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpu_pixel_sharded_fit_matches_single_gpu():
    """torchrun, 2 ranks; bands and stripes, NCCL all-reduce and the peer reduction over CUDA IPC: both ranks end in the same node state,
    equal to the single-GPU fit bit for bit."""
    script = os.path.join(os.path.dirname(__file__), "pixel_shard_worker.py")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29611", script], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "PIXEL_SHARD_OK" in out.stdout
