"""Regenerates tests/golden/*.npz from the reference checkout (run in the build container, where /root/reference exists).

Nothing here copies reference source: the script PARSES the inline tensors of the reference's own Catch2 tests (known-answer
vectors) and sub-samples one of its binary fixtures, and stores them as numpy arrays with their provenance:
  reference_kats.npz   inline vectors of cpp/tests/test_{anchor_computation,rodrigues,unproject_3d_points,linalg_cholesky,matmul3d}.cpp
  mesh_warp_219.npz    every 8th vertex of tests/test_data/mesh_warping/test_{input,gt}_mesh.ply + nodes/rotations/translations.npy
  oracle_fit_tinyr.npz one full GN iteration of the ORACLE on the seeded 'tinyr' scene (regression pin of the oracle itself)
  shortest_path_kat.npz inline data + ground truth of tests/test_compute_anchors.py:57-185 (shortest-path anchors, 11 nodes, 2 vertices)
  normals_kat.npz      96 x 96 crop of cpp/tests/test_data/images/red_shorts_200_depth.png with the matching crop of
                       arrays/red_shorts_200_normals.npy (ground truth of cpp/tests/test_normals_operations.cpp:118-140)
  warp_field_kats.npz  inline vectors of cpp/tests/test_sampling.cpp (median grid subsampling) and
                       cpp/tests/test_graph_warp_field.cpp (HierarchicalGraphWarpField constructor: layers, edges, virtual order)
"""
import os
import re
import sys

import numpy as np

REF = os.environ.get("NNRT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

_NUM = re.compile(r"[-+]?(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?")


def inline_tensor(cpp_text, name, occurrence=0):
    """Values of `o3c::Tensor name(std::vector<T>{ ... }, ...)` as float64 (bools become 0/1)."""
    text = re.sub(r"//[^\n]*", "", cpp_text)
    hits = [m for m in re.finditer(r"\b" + re.escape(name) + r"\s*\(\s*std::vector<([^>]+)>\s*\{", text)]
    m = hits[occurrence]
    depth, i = 1, m.end()
    while depth:
        depth += {"{": 1, "}": -1}.get(text[i], 0)
        i += 1
    body = text[m.end():i - 1]
    if "bool" in m.group(1):
        return np.array([1.0 if t == "true" else 0.0 for t in re.findall(r"true|false", body)])
    return np.array([float(t) for t in _NUM.findall(re.sub(r"(?<=[\d.])f\b", "", body))])


def braces_after(cpp_text, name, occurrence=0):
    """Numbers of the first `{ ... }` initialiser that follows the given occurrence of `name` (comments stripped)."""
    text = re.sub(r"//[^\n]*", "", cpp_text)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    start = [m.end() for m in re.finditer(r"\b" + re.escape(name) + r"\b", text)][occurrence]
    i = text.index("{", start)
    depth, j = 1, i + 1
    while depth:
        depth += {"{": 1, "}": -1}.get(text[j], 0)
        j += 1
    return np.array([float(t) for t in _NUM.findall(re.sub(r"(?<=[\d.])f\b", "", text[i + 1:j - 1]))])


def read(path):
    with open(os.path.join(REF, path)) as fh:
        return fh.read()


def read_ply_vertices(path):
    with open(path, "rb") as fh:
        header = b""
        while not header.endswith(b"end_header\n"):
            header += fh.readline()
        n = int(re.search(rb"element vertex (\d+)", header).group(1))
        dt = np.dtype([("p", "<f8", 3), ("n", "<f8", 3), ("c", "u1", 3)])
        v = np.frombuffer(fh.read(n * dt.itemsize), dtype=dt)
    return v["p"].copy(), v["n"].copy()


def main():
    out = {}
    t = read("cpp/tests/test_anchor_computation.cpp")
    out["anchors/vertices"] = inline_tensor(t, "vertices").reshape(8, 3)
    out["anchors/nodes"] = inline_tensor(t, "nodes").reshape(2, 3)
    out["anchors/node_weights"] = inline_tensor(t, "node_weights")
    out["anchors/anchors_gt_sorted_desc"] = inline_tensor(t, "anchors_gt").reshape(8, 4)
    out["anchors/weights_gt_sorted_desc"] = inline_tensor(t, "weights_gt").reshape(8, 4)
    t = read("cpp/tests/test_rodrigues.cpp")
    out["rodrigues/axis_angle_vectors"] = inline_tensor(t, "axis_angle_vectors").reshape(3, 3)
    out["rodrigues/expected_matrices"] = inline_tensor(t, "expected_matrices").reshape(3, 3, 3)
    t = read("cpp/tests/test_unproject_3d_points.cpp")
    out["unproject/intrinsics"] = inline_tensor(t, "intrinsics").reshape(3, 3)
    out["unproject/depth_u16"] = inline_tensor(t, "depth").reshape(4, 4)
    out["unproject/points_gt"] = inline_tensor(t, "points_ground_truth").reshape(16, 3)
    out["unproject/mask_gt"] = inline_tensor(t, "mask_ground_truth").reshape(16)
    t = read("cpp/tests/test_mesh_warping.cpp")
    out["warp9/vertices_gt"] = inline_tensor(t, "warped_plane_vertices_ground_truth").reshape(9, 3)
    out["warp9/normals_gt"] = inline_tensor(t, "warped_plane_normals_ground_truth").reshape(9, 3)
    t = read("cpp/tests/test_linalg_cholesky.cpp")
    for key in ("A_blocks", "B", "expected_X"):
        try:
            out["block_cholesky/" + key] = inline_tensor(t, key)
        except IndexError:
            pass
    np.savez_compressed(os.path.join(HERE, "reference_kats.npz"), **out)
    print("reference_kats.npz:", {k: v.shape for k, v in out.items()})

    wf = {}
    t = read("cpp/tests/test_sampling.cpp")
    t = t[t.index("void TestGridMedianDownsampling"):]
    wf["median/points"] = braces_after(t, "point_data").reshape(10, 3).astype(np.float32)
    wf["median/cell_size"] = np.float32(1.0)
    wf["median/required"] = np.array([2, 4])                    # REQUIRE(find(2)), REQUIRE(find(4))
    wf["median/either"] = np.array([[6, 7], [8, 9]])            # REQUIRE(find(6) || find(7)), REQUIRE(find(8) || find(9))
    t = read("cpp/tests/test_sampling.cpp")
    g = t[t.index("void TestGridDownsampling_Generic"):]
    wf["mean/points"] = braces_after(g, "point_data").reshape(30, 3).astype(np.float32)
    wf["mean/cell_size"] = np.float32(5.0)                      # downsample(points, 5.0)
    wf["mean/points_gt_sorted_by_x"] = braces_after(g, "downsampled_points_gt_data").reshape(4, 3).astype(np.float32)
    g = t[t.index("void TestClosestToMeanSampling"):]
    wf["closest/points"] = braces_after(g, "point_data").reshape(17, 3).astype(np.float32)
    wf["closest/cell_size"] = np.float32(2.0)
    wf["closest/sample_gt"] = braces_after(g, "sample_gt").astype(np.int64)  # AllEqual: order included
    # Test3DKnnSearch (cpp/tests/test_kdtree.cpp:243-276): 15 points, 3 queries, k = 4, neighbours sorted by distance
    t = read("cpp/tests/test_kdtree.cpp")
    t = t[t.index("void Test3DKnnSearch"):]
    wf["knn/points"] = braces_after(t, "point_data").reshape(15, 3).astype(np.float32)
    wf["knn/queries"] = braces_after(t, "query_point_data").reshape(3, 3).astype(np.float32)
    wf["knn/indices_gt"] = braces_after(t, "gt_nn_indices_data").reshape(3, 4).astype(np.int32)
    wf["knn/distances_gt"] = braces_after(t, "gt_nn_distances_data").reshape(3, 4).astype(np.float32)
    # TestRadiusDownsampling (cpp/tests/test_sampling.cpp:131-141): the reference's 100-point cloud; REQUIRE(min pairwise distance >= radius)
    wf["radius/points"] = np.load(os.path.join(REF, "cpp/tests/test_data/arrays/downsampling_source.npy")).astype(np.float32)
    wf["radius/radius"] = np.float32(10.0)
    t = read("cpp/tests/test_graph_warp_field.cpp")
    t = t[t.index("void TestHierarchicalGraphWarpFieldConstructor"):]
    wf["hgwf/nodes"] = braces_after(t, "node_data").reshape(33, 3).astype(np.float32)
    wf["hgwf/node_coverage"] = np.float32(0.25)                 # hgwf(nodes, 0.25, false, 4, 0, FIXED_NODE_COVERAGE, 3, 4, 2^i * coverage)
    wf["hgwf/layer_count"], wf["hgwf/max_vertex_degree"] = np.int32(3), np.int32(4)
    for i in range(3):
        wf[f"hgwf/layer_{i}"] = braces_after(t, f"ground_truth_layer_{i}").astype(np.int64)
    wf["hgwf/edge_sources_virtual"] = braces_after(t, "edge_source_virtual_node_indices_gt").astype(np.int32)
    wf["hgwf/edge_node_indices_sorted"] = braces_after(t, "edge_node_indices_gt").reshape(116, 2).astype(np.int64)
    wf["hgwf/virtual_node_indices"] = braces_after(t, "virtual_node_indices_gt").astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "warp_field_kats.npz"), **wf)
    print("warp_field_kats.npz:", {k: v.shape for k, v in wf.items()})
    # shortest-path anchors: the inline data of the reference's python test, obtained by executing ONLY that data-building function
    # (tests/test_compute_anchors.py:57-146, numpy + math) -- nothing of it is stored as source
    import ast
    import math
    tree = ast.parse(read("tests/test_compute_anchors.py"))
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "prepare_test_data_anchor_computation_manual"][0]
    scope = {"np": np, "math": math}
    exec(compile(ast.Module([fn], []), "reference_kat", "exec"), scope)
    vertices, nodes, edges, anchors_gt, weights_s0, weights_s1 = scope["prepare_test_data_anchor_computation_manual"](1.0)
    np.savez_compressed(os.path.join(HERE, "shortest_path_kat.npz"), vertices=vertices, nodes=nodes, edges=edges, anchors_gt=anchors_gt,
                        weights_gt=np.stack([weights_s0, weights_s1]), anchor_count=np.int32(8), node_coverage=np.float32(1.0))
    print("shortest_path_kat.npz:", vertices.shape, nodes.shape, edges.shape)

    # ordered point-cloud normals: a 96 x 96 crop of the reference's depth frame with the ground truth its own test compares against
    # (cpp/tests/test_normals_operations.cpp:118-140: red_shorts_200_depth.png, red_shorts_intrinsics.txt, red_shorts_200_normals.npy)
    import cv2
    data = os.path.join(REF, "cpp/tests/test_data")
    depth = cv2.imread(os.path.join(data, "images/red_shorts_200_depth.png"), cv2.IMREAD_UNCHANGED)
    normals_gt = np.load(os.path.join(data, "arrays/red_shorts_200_normals.npy"))
    K = np.loadtxt(os.path.join(data, "intrinsics/red_shorts_intrinsics.txt"))
    y0, x0, size = 200, 280, 96
    np.savez_compressed(os.path.join(HERE, "normals_kat.npz"), depth_u16=depth[y0:y0 + size, x0:x0 + size], origin=np.array([y0, x0]),
                        normals_gt=normals_gt[y0:y0 + size, x0:x0 + size], intrinsics=K[:3, :3])
    print("normals_kat.npz:", depth[y0:y0 + size, x0:x0 + size].shape, "valid", int((depth[y0:y0 + size, x0:x0 + size] > 0).sum()))
    if "--warp-field-only" in sys.argv:
        return

    d = os.path.join(REF, "tests/test_data/mesh_warping")
    vin, nin = read_ply_vertices(os.path.join(d, "test_input_mesh.ply"))
    vgt, _ = read_ply_vertices(os.path.join(d, "test_gt_mesh.ply"))
    sel = slice(0, None, 8)
    np.savez_compressed(os.path.join(HERE, "mesh_warp_219.npz"), vertices=vin[sel].astype(np.float32), normals=nin[sel].astype(np.float32),
                        warped_gt=vgt[sel], nodes=np.load(os.path.join(d, "nodes.npy")), rotations=np.load(os.path.join(d, "rotations.npy")),
                        translations=np.load(os.path.join(d, "translations.npy")), node_coverage=np.float32(0.05))
    print("mesh_warp_219.npz:", vin[sel].shape)

    from oracle import oracle_py as oracle
    from util import scene_with_target, oracle_fit
    s = scene_with_target(oracle, "tinyr", 120, 160)
    ref = oracle_fit(oracle, s, 1, parallel=False)
    np.savez_compressed(os.path.join(HERE, "oracle_fit_tinyr.npz"), hessian_diag=ref["hessian_diag"][0], negative_gradient=ref["negative_gradient"][0],
                        deltas=ref["deltas"][0], pixel_faces=ref["pixel_faces"].astype(np.int32), residual_sum=np.float64(ref["residuals"].astype(np.float64).sum()),
                        translations=ref["translations"], rotations=ref["rotations"])
    print("oracle_fit_tinyr.npz written")


if __name__ == "__main__":
    main()
