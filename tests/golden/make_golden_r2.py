"""Round-2 golden vectors: pins for the Jacobian chain, the whole fitter loop and the block-sparse pieces of the arrowhead solve.
Run in the build container (where /root/reference exists): `python tests/golden/make_golden_r2.py`. Writes tests/golden/fitter_kats.npz.

Nothing here copies reference source. The script
  * IMPORTS the reference's own numpy check script math_check_scripts/dense_depth_jacobians.py and stores the inputs of its two
    single-pixel cases together with every intermediate its main() prints (distorted / corrected barycentrics, d rho / dV chain,
    d w_l / dV, d n_l / dV), evaluated in float64 by the reference's functions;
  * PARSES the inline expectations of the reference's enabled fitter tests (cpp/tests/test_deformable_mesh_fitter_one_node.cpp,
    ..._two_nodes.cpp): node positions, expected node translations, tolerances, iteration counts and modes per TEST_CASE. The meshes
    those tests read (plane_skin_1_node_*.ply ...) are behind a download; tests/test_reference_fitter_kats.py regenerates equivalent
    planes (the expectation does not depend on the tessellation);
  * PARSES the inline known-answer tensors of cpp/tests/test_linalg_matmul_block_sparse.cpp and test_linalg_block_routines.cpp;
  * LOADS the in-repo fixtures cpp/tests/test_data/arrays/extracted_face_{vertices,mask}{,_multiple_meshes}.npy.
"""
import importlib.util
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, braces_after, read  # noqa: E402


def dense_depth_cases(out):
    spec = importlib.util.spec_from_file_location("dense_depth_jacobians", os.path.join(REF, "math_check_scripts/dense_depth_jacobians.py"))
    m = importlib.util.module_from_spec(spec)
    stdout, sys.stdout = sys.stdout, open(os.devnull, "w")  # its helpers print
    try:
        spec.loader.exec_module(m)
        for name in ("translation_case_pixel4848", "rotation_case_pixel4646"):
            tp = getattr(m, name)
            K = m.intrinsics_ndc
            (d0, d1, d2), bary_distorted = m.jacobian_barycentrics_perspective_distorted_wrt_vertices(
                tp.face_vertex0, tp.face_vertex1, tp.face_vertex2, K, tp.tested_ray_point)
            j0 = m.jacobian_percpsective_projections_wrt_vertex(tp.face_vertex0, K)
            j1 = m.jacobian_percpsective_projections_wrt_vertex(tp.face_vertex1, K)
            j2 = m.jacobian_percpsective_projections_wrt_vertex(tp.face_vertex2, K)
            d_dist_dV = np.hstack((d0.dot(j0), d1.dot(j1), d2.dot(j2)))
            bary_corrected = m.perspective_correct_barycentrics(bary_distorted, tp.face_vertex0, tp.face_vertex1, tp.face_vertex2)
            d_corr = m.compute_jacobian_perspective_correction(bary_distorted, tp.face_vertex0, tp.face_vertex1, tp.face_vertex2)
            d_corr_dV = m.jacobian_barycentrics_corrected_wrt_vertices(d_dist_dV, d_corr)
            fv = np.stack([tp.face_vertex0, tp.face_vertex1, tp.face_vertex2], 1)
            fn = np.stack([tp.face_normal0, tp.face_normal1, tp.face_normal2], 1)
            p = f"dense_depth/{name}/"
            out[p + "face_vertices"] = fv.T.copy()
            out[p + "face_normals"] = fn.T.copy()
            out[p + "ray_point"] = np.asarray(tp.tested_ray_point)
            out[p + "ndc_intrinsics"] = np.array([K.fx, K.fy, K.cx, K.cy])
            out[p + "residual"] = np.float64(tp.residual)
            out[p + "rendered_point_position"] = np.asarray(tp.rendered_point_position)
            out[p + "rendered_point_normal"] = np.asarray(tp.rendered_point_normal)
            out[p + "reference_point_position"] = np.asarray(tp.reference_point_position)
            out[p + "bary_distorted"] = bary_distorted
            out[p + "d_bary_distorted_dV"] = d_dist_dV
            out[p + "bary_corrected"] = bary_corrected
            out[p + "d_bary_corrected_d_distorted_and_z"] = d_corr
            out[p + "d_bary_corrected_dV"] = d_corr_dV
            out[p + "d_wl_dV"] = fv.dot(d_corr_dV) + np.kron(bary_corrected, np.eye(3))
            out[p + "d_nl_dV"] = fn.dot(d_corr_dV)
    finally:
        sys.stdout.close()
        sys.stdout = stdout


MODES = {"ALL": 0, "TRANSLATION_ONLY": 1, "ROTATION_ONLY": 2}


def _calls(text, function):
    """Arguments of every `function(` call inside an ENABLED TEST_CASE with a CPU device."""
    text = re.sub(r"//[^\n]*", "", text)
    calls = []
    for m in re.finditer(r'TEST_CASE\("([^"]+)"\)\s*\{(.*?)\n\}', text, flags=re.S):
        title, body = m.group(1), m.group(2)
        c = re.search(re.escape(function) + r"\s*\((.*)\)\s*;", body, flags=re.S)
        if c and "CPU:0" in body:
            calls.append((title, re.sub(r"\s+", " ", c.group(1))))
    return calls


def fitter_cases(out):
    """One row per enabled CPU TEST_CASE: [scenario, angle_deg, mode, iterations, rotation_atol, translation_atol]; scenarios:
    0 = 1-node plane translation, 1 = 1-node plane rotation, 2 = 1-node plane (rotation 45 + translation), 3 = 2 node planes,
    4 = 2-node contiguous surface translation."""
    one = read("cpp/tests/test_deformable_mesh_fitter_one_node.cpp")
    two = read("cpp/tests/test_deformable_mesh_fitter_two_nodes.cpp")
    rows, titles = [], []

    def mode_of(args):
        return MODES[re.search(r"IterationMode::(\w+)", args).group(1)]

    def ints(args):
        return [int(t) for t in re.findall(r"(?<![\w.])-?\d+(?![\w.])", re.sub(r"\{[^}]*\}", "", args))]

    for title, args in _calls(one, "TestDeformableImageFitter_1NodePlaneTranslation"):
        rows.append([0, 0, mode_of(args), ints(args)[0], 1e-5, 1e-5]); titles.append(title)      # AllClose(expected, 1., 1e-5) x2 (:126-127)
    for title, args in _calls(one, "TestDeformableImageFitter_1NodePlaneRotation"):
        angle, iters = ints(args)[:2]
        rows.append([1, angle, mode_of(args), iters, 1e-4, 1e-3]); titles.append(title)          # :259-260
    for title, args in _calls(one, "TestDeformableImageFitter_1NodePlane"):
        rows.append([2, 45, mode_of(args), ints(args)[0], 1e-3, 1e-3]); titles.append(title)      # :416-417
    for title, args in _calls(two, "TestDeformableImageFitter_2NodePlanes"):
        angle, iters = ints(args)[:2]
        tol = [float(t) for t in re.findall(r"\d+e-\d+", args)]
        rows.append([3, angle, mode_of(args), iters, tol[0], tol[1]]); titles.append(title)
    for title, args in _calls(two, "TestDeformableImageFitter_2NodeSurfaceTranslation"):
        tol = [float(t) for t in re.findall(r"\d+e-\d+", args)]
        rows.append([4, 0, mode_of(args), ints(args)[0], tol[0], tol[1]]); titles.append(title)
    out["fitter/cases"] = np.array(rows, np.float64)
    out["fitter/case_titles"] = np.array(titles)
    # inline data of the scenarios
    t1 = re.sub(r"//[^\n]*", "", one)
    f = t1[t1.index("void TestDeformableImageFitter_1NodePlaneTranslation"):]
    out["fitter/one_node/node_positions"] = braces_after(f, "node_positions").reshape(1, 3)
    out["fitter/one_node/expected_translation_translation_case"] = braces_after(f, "expected_node_translations").reshape(1, 3)
    out["fitter/one_node/mesh_transform"] = braces_after(f, "mesh_transform").reshape(4, 4)
    out["fitter/one_node/projection_matrix"] = braces_after(f, "projection_matrix").reshape(3, 3)
    out["fitter/one_node/image_resolution"] = braces_after(f, "image_resolution").astype(np.int64)
    f = t1[t1.index("void TestDeformableImageFitter_1NodePlaneRotation"):]
    out["fitter/one_node/expected_translation_rotation_case"] = braces_after(f, "expected_node_translations").reshape(1, 3)
    f = t1[t1.index("void TestDeformableImageFitter_1NodePlane("):]
    out["fitter/one_node/expected_translation_combined_case"] = braces_after(f, "expected_node_translations").reshape(1, 3)
    t2 = re.sub(r"//[^\n]*", "", two)
    f = t2[t2.index("void TestDeformableImageFitter_2NodePlanes"):]
    out["fitter/two_nodes/node_positions"] = braces_after(f, "node_positions").reshape(2, 3)
    out["fitter/two_nodes/expected_translations_planes"] = braces_after(f, "expected_node_translations").reshape(2, 3)
    f = t2[t2.index("void TestDeformableImageFitter_2NodeSurfaceTranslation"):]
    out["fitter/two_nodes/expected_translations_surface"] = braces_after(f, "expected_node_translations").reshape(2, 3)
    # constructor arguments shared by the scenarios (one_node.cpp:117-118, two_nodes.cpp:133-134, :268-269)
    out["fitter/one_node/node_coverage"], out["fitter/two_nodes/node_coverage"] = np.float32(0.25), np.float32(0.1)
    out["fitter/lm_factor"], out["fitter/two_nodes/surface_lm_factor"] = np.float32(0.001), np.float32(0.1)


def _tensor(section, name, shape, occurrence=0, dtype=np.float32):
    return braces_after(section, name, occurrence).reshape(shape).astype(dtype)


def block_sparse_kats(out):
    t = re.sub(r"//[^\n]*", "", read("cpp/tests/test_linalg_matmul_block_sparse.cpp"))
    s = t[t.index("void TestMatmulBlockSparseRowWise_Small"):t.index("void TestMatmulBlockSparse_Small")]
    p = "bs/rowwise/"
    out[p + "a_blocks"] = _tensor(s, "a_blocks", (3, 3, 3))
    out[p + "b_blocks"] = _tensor(s, "b_blocks", (11, 3, 3))
    out[p + "b_block_coordinates"] = _tensor(s, "b_block_coordinates", (11, 2), dtype=np.int32)
    out[p + "c_blocks_gt"] = _tensor(s, "c_blocks_gt", (8, 3, 3))
    out[p + "c_block_coordinates_gt"] = _tensor(s, "c_block_coordinates_gt", (8, 2), dtype=np.int32)
    out[p + "c_blocks_padded_gt"] = _tensor(s, "c_blocks_padded_gt", (11, 3, 3))
    s = t[t.index("void TestMatmulBlockSparse_Small"):t.index("void TestBlockSparseAndVectorProduct_Small")]
    p = "bs/matmul/"
    out[p + "a_blocks"] = _tensor(s, "a_blocks", (3, 2, 2))
    out[p + "a_breadboard"] = _tensor(s, "a_breadboard", (2, 3), dtype=np.int16)
    out[p + "b_blocks"] = _tensor(s, "b_blocks", (3, 2, 2))
    out[p + "b_breadboard"] = _tensor(s, "b_breadboard", (3, 2), dtype=np.int16)
    for name in "cdef":  # c = A B, d = B^T A^T, e = B^T B, f = A A^T  (:295-383)
        blocks = braces_after(s, name + "_blocks_gt")
        out[p + name + "_blocks_gt"] = blocks.reshape(-1, 2, 2).astype(np.float32)
        out[p + name + "_block_coordinates_gt"] = braces_after(s, name + "_block_coordinates_gt").reshape(-1, 2).astype(np.int32)
    s = t[t.index("void TestBlockSparseAndVectorProduct_Small"):t.index("void TestDiagonalBlockSparseAndVectorProduct_Small")]
    p = "bs/spmv/"
    out[p + "m"] = np.int32(4)
    out[p + "a_blocks"] = _tensor(s, "a_blocks", (3, 2, 2))
    out[p + "a_coordinates"] = _tensor(s, "a_coordinates", (3, 2), dtype=np.int32)
    out[p + "vector_b"] = _tensor(s, "vector_b", (6,))
    out[p + "vector_c_gt"] = _tensor(s, "vector_c_gt", (4,))
    out[p + "b_blocks"] = _tensor(s, "b_blocks", (3, 2, 2))
    out[p + "b_coordinates"] = _tensor(s, "b_coordinates", (3, 2), dtype=np.int32)
    out[p + "vector_d_gt"] = _tensor(s, "vector_d_gt", (4,))
    s = t[t.index("void TestDiagonalBlockSparseAndVectorProduct_Small"):]
    p = "bs/diag_spmv/"
    out[p + "d_blocks"] = _tensor(s, "d_blocks", (3, 2, 2))
    out[p + "vector_b"] = _tensor(s, "vector_b", (6,))
    out[p + "vector_c_gt"] = _tensor(s, "vector_c_gt", (6,))

    t = re.sub(r"//[^\n]*", "", read("cpp/tests/test_linalg_block_routines.cpp"))

    def section(name, end):
        a = t.index("void " + name)
        return t[a:t.index("void " + end, a + 1)] if end else t[a:]

    s = section("TestInvertTriangularBlocks_Lower", "TestInvertTriangularBlocks_Upper")
    out["br/tri_lower/blocks"] = braces_after(s, "blocks_lower").astype(np.float32)
    out["br/tri_lower/inverted_gt"] = braces_after(s, "inverted_blocks_lower_gt").astype(np.float32)
    s = section("TestInvertTriangularBlocks_Upper", "TestTransposeBlocks")
    out["br/tri_upper/blocks"] = braces_after(s, "blocks_upper").astype(np.float32)
    out["br/tri_upper/inverted_gt"] = braces_after(s, "inverted_blocks_upper_gt").astype(np.float32)
    s = section("TestTransposeBlocks", "TestInvertPositiveSemidefiniteBlocks")
    out["br/transpose/blocks"] = braces_after(s, "blocks").astype(np.float32)
    out["br/transpose/transposed_gt"] = braces_after(s, "blocks_transposed_gt").astype(np.float32)
    s = section("TestInvertPositiveSemidefiniteBlocks", "TestFillInDiagonalBlocks")
    out["br/invert_psd/blocks"] = _tensor(s, "blocks", (3, 3, 3))
    out["br/invert_psd/inverted_gt"] = _tensor(s, "inverted_blocks_gt", (3, 3, 3))
    # the block lists of the four fill / get tests are Arange(0, 24).Reshape({6, 2, 2}) and the filled matrix starts as Zeros({12, 12}) (:208-209,266,284-285,395)
    s = section("TestFillInDiagonalBlocks", "TestGetDiagonalBlocks")
    out["br/fill_diag/matrix_filled_gt"] = braces_after(s, "matrix_filled_gt").astype(np.float32)
    s = section("TestGetDiagonalBlocks", "TestFillInSparseBlocks")
    out["br/get_diag/matrix"] = braces_after(s, "matrix").astype(np.float32)
    s = section("TestFillInSparseBlocks", "TestGetSparseBlocks")
    out["br/fill_sparse/coordinates"] = braces_after(s, "coordinates").astype(np.int32)
    out["br/fill_sparse/matrix_filled_gt"] = braces_after(s, "matrix_filled_gt").astype(np.float32)
    out["br/fill_sparse/blocks_to_transpose"] = braces_after(s, "blocks_to_transpose").astype(np.float32)
    out["br/fill_sparse/coordinates_to_transpose"] = braces_after(s, "coordinates_to_transpose").astype(np.int32)
    out["br/fill_sparse/matrix_filled_gt2"] = braces_after(s, "matrix_filled_gt2").astype(np.float32)
    s = section("TestGetSparseBlocks", None)
    out["br/get_sparse/matrix"] = braces_after(s, "matrix").astype(np.float32)
    out["br/get_sparse/coordinates"] = braces_after(s, "coordinates").astype(np.int32)


def extracted_faces(out):
    d = os.path.join(REF, "cpp/tests/test_data/arrays")
    for suffix in ("", "_multiple_meshes"):
        out["faces/vertices" + suffix] = np.load(os.path.join(d, f"extracted_face_vertices{suffix}.npy"))
        out["faces/mask" + suffix] = np.load(os.path.join(d, f"extracted_face_mask{suffix}.npy"))
    # cpp/tests/test_extract_face_vertices.cpp:32-41,72-84: GenerateXyPlane(1.2615, (0,0,1), 4 subdivisions); K; {480,640}; near 0, far 2;
    # second case adds GenerateSphere(0.4, (0,0,0.5), resolution 32)
    out["faces/plane_side"], out["faces/plane_center"], out["faces/plane_subdivisions"] = np.float32(1.2615), np.array([0, 0, 1], np.float32), np.int32(4)
    out["faces/sphere_radius"], out["faces/sphere_center"], out["faces/sphere_resolution"] = np.float32(0.4), np.array([0, 0, 0.5], np.float32), np.int32(32)
    out["faces/intrinsics"] = np.array([[580., 0, 320], [0, 580, 240], [0, 0, 1]])
    out["faces/image_size"] = np.array([480, 640])
    out["faces/near_far"] = np.array([0.0, 2.0])
    out["faces/kept_counts"] = np.array([334, 1994])


def main():
    out = {}
    dense_depth_cases(out)
    fitter_cases(out)
    block_sparse_kats(out)
    extracted_faces(out)
    np.savez_compressed(os.path.join(HERE, "fitter_kats.npz"), **out)
    for k, v in out.items():
        print(k, getattr(v, "shape", ()), getattr(v, "dtype", ""))


if __name__ == "__main__":
    main()
