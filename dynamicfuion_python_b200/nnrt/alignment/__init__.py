"""nnrt.alignment — DeformableMeshToImageFitter (cpp/alignment/DeformableMeshToImageFitter.h:30-126; C++-only in the reference,
exposed here under its natural name with the constructor's argument names and defaults)."""
import ctypes as C
import enum

import torch

from ... import _lib, pixel_shard
from .. import _tensor as T
from ..geometry import WarpNodeCoverageComputationMethod
from . import functional


class IterationMode(enum.IntEnum):
    ALL = 0
    TRANSLATION_ONLY = 1
    ROTATION_ONLY = 2


class DeformableMeshToImageFitter:
    def __init__(self, max_iteration_count=100, iteration_mode_sequence=(IterationMode.ALL,), minimal_update_threshold=1e-6,
                 use_perspective_correction=True, max_depth=10.0, use_tukey_penalty_for_depth_term=False, tukey_penalty_cutoff_cm=0.01,
                 preconditioning_dampening_factor=0.0, arap_term_weight=200.0, use_huber_penalty_for_arap_term=False,
                 huber_penalty_constant=0.0001):
        if preconditioning_dampening_factor < 0.0 or preconditioning_dampening_factor > 1.0:  # DeformableMeshToImageFitter.cpp:79-82
            raise RuntimeError("`preconditioning_dampening_factor` should be a small non-negative value between 0 and 1. "
                               f"Got {preconditioning_dampening_factor}.")
        modes = [IterationMode(m) for m in iteration_mode_sequence]
        self.max_iteration_count = int(max_iteration_count)
        self.iteration_mode_sequence = modes
        self.min_update_threshold = float(minimal_update_threshold)  # the reference never updates `maximum_update` (:100,111): unused
        self.options = _lib.FitterOptions(self.max_iteration_count, int(bool(use_perspective_correction)), float(max_depth),
                                          int(bool(use_tukey_penalty_for_depth_term)), float(tukey_penalty_cutoff_cm),
                                          float(preconditioning_dampening_factor), float(arap_term_weight),
                                          int(bool(use_huber_penalty_for_arap_term)), float(huber_penalty_constant))
        self._handle = None
        self._pixel_shard = None
        self._stripe_rows = None
        self._shard_collective, self._peer_ready = "nccl", False
        self._shard_graphs, self._shard_seen, self._shard_keep = {}, None, None
        self._shape = None
        self._capacity = (0, 0)
        self._model_key = None
        self._keep = None

    def __del__(self):
        self._release()

    def _release(self):
        if getattr(self, "_handle", None):
            try:
                _lib.lib().nnrt_b200_fitter_destroy(self._handle)
            except TypeError:  # interpreter shutdown: the module globals are already torn down, the process frees the workspace
                pass
            self._handle = None

    def _ensure(self, shape):
        if self._handle is not None and self._shape == shape:
            return
        v, f, n, e, a, h, w = shape
        same_problem = self._handle is not None and self._shape[2:] == shape[2:]
        if same_problem and v <= self._capacity[0] and f <= self._capacity[1]:
            # a canonical mesh re-extracted every frame changes its counts: the workspace is kept, only the launch sizes change
            T.call("nnrt_b200_fitter_set_mesh_size", self._handle, v, f)
            self._shape = shape
            return
        # a mesh that outgrew the workspace gets headroom, so that a slowly growing surface does not reallocate every frame; the first
        # workspace of a problem is exact (a fixed canonical mesh, the usual case, wastes nothing)
        capacity = (v + v // 2, f + f // 2) if same_problem else (v, f)
        self._release()
        handle = C.c_void_p()
        _lib.check(_lib.lib().nnrt_b200_fitter_create(C.byref(self.options), capacity[0], capacity[1], n, e, a, h, w, C.byref(handle)))
        self._handle, self._shape, self._model_key, self._capacity = handle, shape, None, capacity
        if capacity != (v, f):
            T.call("nnrt_b200_fitter_set_mesh_size", handle, v, f)
        self._peer_ready = False  # a new fitter object: its exchange buffer has not been mapped by the peers
        modes = (C.c_int32 * len(self.iteration_mode_sequence))(*[int(m) for m in self.iteration_mode_sequence])
        _lib.check(_lib.lib().nnrt_b200_fitter_set_iteration_modes(handle, modes, len(self.iteration_mode_sequence)))

    def shard_pixels(self, rank, world_size, group=None, stripe_rows=pixel_shard.STRIPE_ROWS, collective=None):
        """Pixel-sharded single-frame mode (dynamicfuion_python_b200.pixel_shard): this process fits its share of the image rows and the
        per-node sums are summed over the ranks of `group` once per iteration. The rows are dealt in stripes of `stripe_rows` (a multiple
        of 8), stripe s to rank s % world_size, so that a surface in the middle of the frame loads every GPU alike; stripe_rows=None
        gives one contiguous band per rank, pixel_rows(rank, world_size, height). `collective`: "peer" (default, NNRT_B200_SHARD_COLLECTIVE)
        = the consuming kernel reads the ranks' sums over NVLink peer memory itself and the whole fit is one call into the C library;
        "nccl" = an int64 all-reduce between the two halves of an iteration. world_size 1 turns the mode off."""
        if stripe_rows is not None and (int(stripe_rows) <= 0 or int(stripe_rows) % 8):
            raise ValueError(f"stripe_rows must be a positive multiple of 8 (the assembly tile height), got {stripe_rows}")
        collective = pixel_shard.COLLECTIVE if collective is None else collective
        if collective not in ("peer", "nccl"):
            raise ValueError(f"collective must be 'peer' or 'nccl', got {collective!r}")
        self._pixel_shard = None if world_size <= 1 else (int(rank), int(world_size), group)
        self._stripe_rows = None if stripe_rows is None else int(stripe_rows)
        self._shard_collective, self._peer_ready = collective, False
        if self._handle is not None:
            T.call("nnrt_b200_fitter_disable_peer_reduce", self._handle)
            if self._pixel_shard is None:
                T.call("nnrt_b200_fitter_set_pixel_rows", self._handle, 0, self._shape[5])
                T.call("nnrt_b200_fitter_set_pixel_stripes", self._handle, 8, 0, 1)

    def _enable_peer_reduce(self, rank, world, group):
        """Collective over `group`: every rank clears its exchange buffer, the CUDA IPC handles are gathered, every rank maps its peers.
        All ranks take the same path: if any of them cannot map a peer (no NVLink / IPC not permitted), all fall back to the all-reduce."""
        torch.cuda.synchronize()
        torch.distributed.barrier(group=group)  # no rank is still reading a buffer that is about to be cleared
        handle = (C.c_ubyte * 64)()
        status = _lib.lib().nnrt_b200_fitter_peer_exchange(self._handle, None, None, handle)
        gathered = [None] * world
        torch.distributed.all_gather_object(gathered, bytes(handle) if status == 0 else None, group=group)
        ok = status == 0 and all(h is not None for h in gathered)
        if ok:
            blob = (C.c_ubyte * (64 * world)).from_buffer_copy(b"".join(gathered))
            ok = _lib.lib().nnrt_b200_fitter_enable_peer_reduce_ipc(self._handle, rank, world, blob) == 0
        verdicts = [None] * world
        torch.distributed.all_gather_object(verdicts, bool(ok), group=group)  # also the barrier the protocol asks for before the first fit
        if not all(verdicts):
            T.call("nnrt_b200_fitter_disable_peer_reduce", self._handle)
            self._shard_collective = "nccl"
            return False
        return True

    def workspace_bytes(self):
        return 0 if self._handle is None else _lib.lib().nnrt_b200_fitter_workspace_bytes(self._handle)

    # --- setup -------------------------------------------------------------------------------------------------
    def _set_model(self, warp_field, canonical_mesh, image_size, anchors=None, anchor_weights=None):
        mesh = T.MeshView(canonical_mesh)
        if mesh.normals is None:
            raise RuntimeError("Input warped mesh needs to have normals defined; it doesn't.")
        if mesh.triangles is None:
            raise RuntimeError("Argument mesh needs to have both triangle vertex indices and vertex positions.")
        if anchors is None:  # warp_field.PrecomputeAnchorsAndWeights(canonical_mesh, true), DeformableMeshToImageFitter.cpp:96-97
            if hasattr(warp_field, "node_indices"):  # hierarchical field: anchors must index the virtual order the solver works in
                anchors, anchor_weights = warp_field.precompute_anchors_and_weights(mesh, use_virtual_ordering=True)
            else:
                anchors, anchor_weights = warp_field.precompute_anchors_and_weights(mesh)
        anchors, anchor_weights = T.cuda(anchors, torch.int32), T.cuda(anchor_weights, torch.float32)
        edges = getattr(warp_field, "edges", None)
        e = 0 if edges is None else int(edges.shape[0])
        n = warp_field.nodes.shape[0]
        self._ensure((mesh.positions.shape[0], mesh.triangles.shape[0], n, e, anchors.shape[1], int(image_size[0]), int(image_size[1])))
        fixed = warp_field.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE
        eli = radii = ncw = None
        layer_count = 0
        if e > 0:
            if fixed:
                eli, radii = warp_field.edge_layer_indices, warp_field.layer_decimation_radii
                layer_count = radii.shape[0]
                if eli.shape[0] < max(e, n):
                    raise RuntimeError("edge_layer_indices must have at least max(edge_count, node_count) entries")
            else:
                ncw = warp_field.node_coverage_weights
        base = getattr(warp_field, "first_layer_node_count", n)
        # The graph belongs to the warp field and is the same object frame after frame; its upload (host-side validation, one
        # stream synchronisation) is redone only when one of its tensors was replaced or written to (torch version counters).
        graph_key = (int(base), layer_count) + tuple(None if x is None else (x.data_ptr(), x._version, tuple(x.shape))
                                                     for x in (warp_field.nodes, edges if e > 0 else None, eli, radii, ncw))
        if graph_key != self._model_key:
            T.call("nnrt_b200_fitter_set_graph", self._handle, T.ptr(warp_field.nodes), T.ptr(edges) if e > 0 else None, int(base),
                   T.ptr(eli), T.ptr(radii), int(layer_count), T.ptr(ncw), T.stream())
            self._model_key, self._graph_tensors = graph_key, (warp_field.nodes, edges, eli, radii, ncw)  # keep the key's pointers alive
        T.call("nnrt_b200_fitter_set_mesh", self._handle, T.ptr(mesh.positions), T.ptr(mesh.normals), T.ptr(mesh.triangles),
               T.ptr(anchors), T.ptr(anchor_weights), T.stream())

    def _fit(self, warp_field, intrinsic_matrix, extrinsic_matrix, iteration_count=None):
        K = T.host_f64(intrinsic_matrix, (3, 3))
        ext = T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsic_matrix is None else extrinsic_matrix, (4, 4))
        R, t = warp_field.node_rotations, warp_field.node_translations
        if not (R.is_contiguous() and t.is_contiguous() and R.is_cuda and t.is_cuda):
            raise RuntimeError("warp field node rotations / translations must be contiguous CUDA tensors")
        iters = self.max_iteration_count if iteration_count is None else int(iteration_count)
        if iters <= 0:
            iters = self.max_iteration_count
        if self._pixel_shard is None:
            T.call("nnrt_b200_fitter_fit", self._handle, T.ptr(R), T.ptr(t), K.data_ptr(), ext.data_ptr(), iters, T.stream())
        else:
            # pixel-sharded: this rank's band of rows, one int64 all-reduce of the per-node sums per iteration, replicated solve.
            # The loop (2 launches groups + 1 NCCL collective per iteration) is recorded ONCE into a CUDA graph per (node state, camera,
            # iteration count, band) and replayed; the first call of a key runs eagerly (it also brings up the NCCL communicator, which
            # cannot happen under capture).
            rank, world, group = self._pixel_shard
            if self._stripe_rows is None:
                begin, end = pixel_shard.pixel_rows(rank, world, self._shape[5])
                T.call("nnrt_b200_fitter_set_pixel_rows", self._handle, begin, end)
                T.call("nnrt_b200_fitter_set_pixel_stripes", self._handle, 8, 0, 1)
            else:
                begin, end = 0, -self._stripe_rows  # (key only) striped: every rank walks the whole image height
                T.call("nnrt_b200_fitter_set_pixel_rows", self._handle, 0, self._shape[5])
                T.call("nnrt_b200_fitter_set_pixel_stripes", self._handle, self._stripe_rows, rank, world)
            if self._shard_collective == "peer" and not self._peer_ready:
                self._peer_ready = self._enable_peer_reduce(rank, world, group)
            if self._shard_collective == "peer":
                # the sums are exchanged inside the iteration's own kernels: the sharded fit is the library's captured loop
                T.call("nnrt_b200_fitter_fit", self._handle, T.ptr(R), T.ptr(t), K.data_ptr(), ext.data_ptr(), iters, T.stream())
                self._keep = (K, ext)
                return
            count = C.c_int64()
            pointer = _lib.lib().nnrt_b200_fitter_accumulators(self._handle, C.byref(count))
            acc = pixel_shard.DeviceBuffer(pointer, count.value).tensor(R.device)

            def loop():
                for it in range(iters):
                    T.call("nnrt_b200_fitter_iteration_data_term", self._handle, T.ptr(R), T.ptr(t), K.data_ptr(), ext.data_ptr(), T.stream())
                    pixel_shard.all_reduce_sum(acc, group)
                    T.call("nnrt_b200_fitter_iteration_solve", self._handle, T.ptr(R), T.ptr(t), K.data_ptr(), ext.data_ptr(), it, T.stream())
            key = (R.data_ptr(), t.data_ptr(), K.numpy().tobytes(), ext.numpy().tobytes(), iters, begin, end, world, id(group), self._handle.value)
            graph = self._shard_graphs.get(key)
            if graph is None:
                if self._shard_seen == key and pixel_shard.GRAPH_CAPTURE:
                    graph = torch.cuda.CUDAGraph()
                    Kc, extc = K, ext  # the captured launches read the host matrices at capture time only (by value)
                    with torch.cuda.graph(graph):
                        loop()
                    self._shard_graphs = {key: graph}  # one graph kept: a new key (new warp field, camera ...) replaces it
                    self._shard_keep = (Kc, extc, acc)
                    graph.replay()
                else:
                    self._shard_seen = key
                    loop()
            else:
                graph.replay()
        self._keep = (K, ext)

    def check_status(self):
        """Waits for the stream and raises if the last fit overflowed its band queue or met a non-positive-definite block."""
        T.call("nnrt_b200_fitter_status", self._handle, T.stream())

    # --- FitToImage overloads (cpp/alignment/DeformableMeshToImageFitter.cpp:85-329) --------------------------------
    def fit_to_image(self, warp_field, canonical_mesh, *args, **kwargs):
        """fit_to_image(warp_field, canonical_mesh, reference_color_image, reference_depth_image, reference_image_mask, intrinsic_matrix,
        extrinsic_matrix, depth_scale), or with a reference point cloud:
        fit_to_image(warp_field, canonical_mesh, reference_color_image, reference_point_cloud, reference_point_mask, intrinsic_matrix,
        extrinsic_matrix, rendering_image_size). The colour image is unused by the reference beyond its size and may be None."""
        names = ["reference_color_image", "reference", "reference_mask", "intrinsic_matrix", "extrinsic_matrix", "last"]
        rgbd = kwargs.pop("reference_image", None)
        if rgbd is None and args and (hasattr(args[0], "depth") or (isinstance(args[0], (tuple, list)) and len(args[0]) == 2 and len(args) <= 5)):
            rgbd, args = args[0], args[1:]  # RGB-D overload (.cpp:316-329): reference_image, reference_image_mask, K, extrinsics, depth_scale
            names = names[2:]
        if rgbd is not None:
            color, depth = (rgbd.color, rgbd.depth) if hasattr(rgbd, "depth") else rgbd
            if names[0] == "reference_color_image":  # passed by keyword: the remaining positionals start at the mask
                names = names[2:]
            kwargs.update(reference_color_image=color, reference_depth_image=depth)
        vals = dict(zip(names, args))
        vals.update(kwargs)
        ref = vals.get("reference", vals.get("reference_depth_image", vals.get("reference_point_cloud")))
        mask = vals.get("reference_mask", vals.get("reference_image_mask", vals.get("reference_point_mask")))
        K, ext = vals["intrinsic_matrix"], vals.get("extrinsic_matrix")
        last = vals.get("last", vals.get("depth_scale", vals.get("rendering_image_size")))
        anchors, weights = vals.get("anchors"), vals.get("anchor_weights")
        ref_t = ref.point["positions"] if hasattr(ref, "point") else (ref.as_tensor() if hasattr(ref, "as_tensor") else ref)
        ref_t = T.to_torch(ref_t)
        is_points = ref_t.dim() == 2 and ref_t.shape[1] == 3 and isinstance(last, (tuple, list))
        if is_points:
            image_size = (int(last[0]), int(last[1]))
            self._set_model(warp_field, canonical_mesh, image_size, anchors, weights)
            pts = T.cuda(ref_t, torch.float32)
            m = None if mask is None else T.cuda(T.to_torch(mask).reshape(-1), torch.uint8)
            T.call("nnrt_b200_fitter_set_reference_points", self._handle, T.ptr(pts), T.ptr(m), T.stream())
        else:
            depth = ref_t[..., 0] if ref_t.dim() == 3 else ref_t
            image_size = (int(depth.shape[0]), int(depth.shape[1]))
            self._set_model(warp_field, canonical_mesh, image_size, anchors, weights)
            if depth.dtype == torch.uint16:
                d, code = T.cuda(depth.view(torch.int16), torch.int16), 1
            else:
                d, code = T.cuda(depth, torch.float32), 0
            m = None if mask is None else T.cuda(T.to_torch(mask).reshape(-1), torch.uint8)
            Kh = T.host_f64(K, (3, 3))
            T.call("nnrt_b200_fitter_set_reference_depth", self._handle, T.ptr(d), code, T.ptr(m), float(1000.0 if last is None else last),
                   Kh.data_ptr(), T.stream())
        self._fit(warp_field, K, ext, vals.get("iteration_count"))

    def diagnostics(self):
        """Last iteration's system and per-pixel outputs: dict(hessian_diagonal [N,6,6], negative_gradient [N,6], deltas [N,6],
        pixel_face_indices [H,W] int64, residuals [H,W], residual_mask [H,W] bool)."""
        _, _, n, _, _, h, w = self._shape
        Hd, g, d = T.empty((n, 6, 6), torch.float32), T.empty((n, 6), torch.float32), T.empty((n, 6), torch.float32)
        faces, res, mask = T.empty((h, w), torch.int64), T.empty((h, w), torch.float32), T.empty((h, w), torch.uint8)
        T.call("nnrt_b200_fitter_get_diagnostics", self._handle, T.ptr(Hd), T.ptr(g), T.ptr(d), T.ptr(faces), T.ptr(res), T.ptr(mask), T.stream())
        return dict(hessian_diagonal=Hd, negative_gradient=g, deltas=d, pixel_face_indices=faces, residuals=res, residual_mask=mask.to(torch.bool))

    def profile_iteration(self, warp_field, intrinsic_matrix, extrinsic_matrix=None):
        """Runs ONE iteration outside the graph with CUDA events between stages; returns ms for
        [warp_vertices, face_raster, band_raster, assemble, build_system, stem_eliminate, corner_solve_update]."""
        K = T.host_f64(intrinsic_matrix, (3, 3))
        ext = T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsic_matrix is None else extrinsic_matrix, (4, 4))
        out = (C.c_float * 7)()
        T.call("nnrt_b200_fitter_profile_iteration", self._handle, T.ptr(warp_field.node_rotations), T.ptr(warp_field.node_translations),
               K.data_ptr(), ext.data_ptr(), out, T.stream())
        return list(out)

    def launches_per_iteration(self):
        return int(_lib.lib().nnrt_b200_fitter_launches_per_iteration(self._handle))
