"""Bitwise A/B of two builds of libnnrt_b200.so: one data-term pass (warp -> raster -> assemble) per scene, the raw 103-bit accumulator
words of J^T J / J^T r dumped to an .npz. Run once per library and compare the files word for word -- the check used for arithmetic
rewrites that claim identical bits (csrc/common.cuh::SharedDivisor: 0 of 32 896 words differ on C1 / C2r at 640 x 480 and C2r at 4K).

    python tests/measure/ab_accumulators.py <libnnrt_b200.so | tree> out.npz      # "tree": the in-tree build
    python tests/measure/ab_accumulators.py --compare a.npz b.npz
Symbols of _lib.py's table that an older library lacks are skipped."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np


def compare(a_path, b_path):
    a, b = np.load(a_path), np.load(b_path)
    worst = 0
    for key in a.files:
        differing = int((a[key] != b[key]).sum())
        worst = max(worst, differing)
        print(key, "accumulator words", a[key].size, "differing", differing)
    return worst


def dump(library, out_path):
    import torch
    from dynamicfuion_python_b200 import _lib
    if library != "tree":
        _lib.LIB_PATH = os.path.abspath(library)
        handle = C.CDLL(_lib.LIB_PATH)
        for name in [n for n in _lib._SIGNATURES if not hasattr(handle, n)]:
            _lib._SIGNATURES.pop(name)
    from dynamicfuion_python_b200 import nnrt, pixel_shard
    from dynamicfuion_python_b200.nnrt import _tensor as T
    from util import scene_with_gpu_target
    G = nnrt.geometry
    out = {}
    for config, h, w in (("C1", 480, 640), ("C2r", 480, 640), ("C2r", 2160, 3840)):
        s = scene_with_gpu_target(nnrt, config, h, w)
        field = lambda: G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                                G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4,
                                                                s["edge_layer_indices"], s["layer_radii"])
        fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=s["lm_factor"],
                                                            arap_term_weight=s["arap_term_weight"])
        fitter.fit_to_image(field(), (s["positions"], s["normals"], s["triangles"]), None, s["reference_points"], s["reference_mask"], s["K"],
                            np.eye(4), (h, w), iteration_count=1, anchors=s["anchors"], anchor_weights=s["anchor_weights"])  # model + reference
        wf = field()
        R, t = wf.node_rotations, wf.node_translations
        K, ext = T.host_f64(s["K"], (3, 3)), T.host_f64(np.eye(4), (4, 4))
        count = C.c_int64()
        pointer = _lib.lib().nnrt_b200_fitter_accumulators(fitter._handle, C.byref(count))
        acc = pixel_shard.DeviceBuffer(pointer, count.value).tensor(R.device)
        acc.zero_()
        T.call("nnrt_b200_fitter_iteration_data_term", fitter._handle, T.ptr(R), T.ptr(t), K.data_ptr(), ext.data_ptr(), T.stream())
        torch.cuda.synchronize()
        out[f"{config}_{h}x{w}"] = acc.cpu().numpy().copy()
    np.savez(out_path, **out)
    print("dumped", {k: (v.size, int(np.count_nonzero(v))) for k, v in out.items()})


if __name__ == "__main__":
    if sys.argv[1] == "--compare":
        sys.exit(1 if compare(sys.argv[2], sys.argv[3]) else 0)
    dump(sys.argv[1], sys.argv[2])
