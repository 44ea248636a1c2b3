"""CPU tests (no GPU): the oracle against the round-2 pins in tests/golden/fitter_kats.npz (tests/golden/make_golden_r2.py):
the reference's own single-pixel Jacobian check script, the inline expectations of its ENABLED one-/two-node fitter tests, the
inline known-answer tensors of its block-sparse routines and its in-repo extracted-face fixtures."""
import os

import numpy as np
import pytest

from util import reference_all_close, reference_fitter_scenario

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def kats():
    return np.load(os.path.join(GOLDEN, "fitter_kats.npz"))


DENSE_DEPTH_CASES = [("translation_case_pixel4848", (48, 48)), ("rotation_case_pixel4646", (46, 46))]


def dense_depth_inputs(kats, case, pixel):
    p = f"dense_depth/{case}/"
    v, u = pixel
    faces = np.full((100, 100, 1), -1, np.int64)
    faces[v, u, 0] = 0
    bary = np.zeros((100, 100, 1, 3), np.float32)
    bary[v, u, 0] = kats[p + "bary_corrected"]
    K = np.array([[100.0, 0, 50], [0, 100, 50], [0, 0, 1]])  # NDC intrinsics (2, -2, 0, 0) on 100 x 100, as the script states
    return kats[p + "face_vertices"].astype(np.float32), kats[p + "face_normals"].astype(np.float32), np.array([[0, 1, 2]], np.int64), faces, bary, K


@pytest.mark.parametrize("case,pixel", DENSE_DEPTH_CASES)
def test_jacobian_chain_against_reference_check_script(oracle, kats, case, pixel):
    """math_check_scripts/dense_depth_jacobians.py (float64, the reference author's own derivation check): d w_l / dV and d n_l / dV of
    its two single-pixel cases. The oracle works in fp32 and adds the reference's 1e-8 guards to the squared areas (2e-5 relative)."""
    p = f"dense_depth/{case}/"
    pos, nrm, tris, faces, bary, K = dense_depth_inputs(kats, case, pixel)
    ndc_K, _ = oracle.intrinsics_to_ndc(K, 100, 100)
    assert np.allclose([ndc_K[0, 0], ndc_K[1, 1], ndc_K[0, 2], ndc_K[1, 2]], kats[p + "ndc_intrinsics"])
    jv, jn = oracle.rasterized_surface_jacobians(pos, nrm, tris, faces, bary, ndc_K, True, False)
    v, u = pixel
    np.testing.assert_allclose(jv[v, u], kats[p + "d_wl_dV"], rtol=1e-4, atol=1e-5)
    flat = jn[v, u].reshape(30)  # 3 x 9 Jacobian, then the three barycentrics (RasterizedSurfaceJacobiansImpl.h:186-196)
    np.testing.assert_allclose(flat[:27].reshape(3, 9), kats[p + "d_nl_dV"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(flat[27:], kats[p + "bary_corrected"], rtol=1e-6, atol=1e-7)


def fitter_case_ids(kats_path=os.path.join(GOLDEN, "fitter_kats.npz")):
    g = np.load(kats_path)
    return list(range(len(g["fitter/cases"]))), [str(t).replace("Test DMI Fitter", "").strip(" -") for t in g["fitter/case_titles"]]


_CASES, _IDS = fitter_case_ids()


def fitter_case(kats, index):
    scenario, angle, mode, iterations, rotation_atol, translation_atol = kats["fitter/cases"][index]
    lm = float(kats["fitter/two_nodes/surface_lm_factor"] if int(scenario) == 4 else kats["fitter/lm_factor"])
    return int(scenario), float(angle), int(mode), int(iterations), float(rotation_atol), float(translation_atol), lm


@pytest.mark.parametrize("index", _CASES, ids=_IDS)
def test_enabled_reference_fitter_cases(oracle, kats, index):
    """cpp/tests/test_deformable_mesh_fitter_one_node.cpp:107-127,228-260,385-420 and ..._two_nodes.cpp: every enabled CPU TEST_CASE with
    its iteration mode, iteration count, damping and the reference's own check AllClose(expected, rtol = 1, atol). Pins the whole GN
    loop of the oracle (ALL, TRANSLATION_ONLY and ROTATION_ONLY; block-diagonal solve; node update) end to end."""
    scenario, angle, mode, iterations, rotation_atol, translation_atol, lm = fitter_case(kats, index)
    s = reference_fitter_scenario(oracle, kats, scenario, angle)
    r = oracle.fit_to_image(s["positions"], s["normals"], s["triangles"], s["nodes"], None, len(s["nodes"]), s["anchors"], s["anchor_weights"],
                            s["reference_points"], s["reference_mask"], s["height"], s["width"], s["K"], iteration_count=iterations, lm_factor=lm,
                            iteration_modes=[mode], parallel=False)
    assert reference_all_close(r["rotations"], s["expected_rotations"], rotation_atol)
    assert reference_all_close(r["translations"], s["expected_translations"], translation_atol)
    # beyond the reference's criterion (which is loose on non-zero entries): the fit lands on the expected state itself
    assert np.abs(r["rotations"] - s["expected_rotations"]).max() <= max(10 * rotation_atol, 1e-4)
    assert np.abs(r["translations"] - s["expected_translations"]).max() <= max(translation_atol, 1e-5) * (20 if scenario == 3 else 1)


def test_block_sparse_matmul_kats(oracle, kats):
    """cpp/tests/test_linalg_matmul_block_sparse.cpp:30-538 (AllClose defaults 1e-5 / 1e-8)."""
    a, b, bc = kats["bs/rowwise/a_blocks"], kats["bs/rowwise/b_blocks"], kats["bs/rowwise/b_block_coordinates"]
    c, cc = oracle.matmul_block_sparse_row_wise(a, b, bc)
    np.testing.assert_allclose(c, kats["bs/rowwise/c_blocks_gt"], rtol=1e-5, atol=1e-8)
    assert np.array_equal(cc, kats["bs/rowwise/c_block_coordinates_gt"])
    np.testing.assert_allclose(oracle.matmul_block_sparse_row_wise(a, b, bc, padded=True), kats["bs/rowwise/c_blocks_padded_gt"], rtol=1e-5, atol=1e-8)
    A, Ab, B, Bb = kats["bs/matmul/a_blocks"], kats["bs/matmul/a_breadboard"], kats["bs/matmul/b_blocks"], kats["bs/matmul/b_breadboard"]
    for name, args in dict(c=(A, Ab, 0, B, Bb, 0), d=(B, Bb, 1, A, Ab, 1), e=(B, Bb, 1, B, Bb, 0), f=(A, Ab, 0, A, Ab, 1)).items():
        blocks, coords = oracle.matmul_block_sparse(*args)
        np.testing.assert_allclose(blocks, kats[f"bs/matmul/{name}_blocks_gt"], rtol=1e-5, atol=1e-8)
        assert np.array_equal(coords, kats[f"bs/matmul/{name}_block_coordinates_gt"])
    m = int(kats["bs/spmv/m"])
    np.testing.assert_allclose(oracle.block_sparse_and_vector_product(kats["bs/spmv/a_blocks"], m, kats["bs/spmv/a_coordinates"], (0, 0), 0, kats["bs/spmv/vector_b"]),
                               kats["bs/spmv/vector_c_gt"], rtol=1e-5)
    np.testing.assert_allclose(oracle.block_sparse_and_vector_product(kats["bs/spmv/b_blocks"], m, kats["bs/spmv/b_coordinates"], (0, 0), 1, kats["bs/spmv/vector_b"]),
                               kats["bs/spmv/vector_d_gt"], rtol=1e-5)
    np.testing.assert_allclose(oracle.diagonal_block_sparse_and_vector_product(kats["bs/diag_spmv/d_blocks"], kats["bs/diag_spmv/vector_b"]),
                               kats["bs/diag_spmv/vector_c_gt"], rtol=1e-5)


def test_block_routine_kats(oracle, kats):
    """cpp/tests/test_linalg_block_routines.cpp:32-407."""
    inv, fails = oracle.invert_positive_semidefinite_blocks(kats["br/invert_psd/blocks"])
    assert fails == 0
    np.testing.assert_allclose(inv, kats["br/invert_psd/inverted_gt"], rtol=1e-4, atol=1e-6)
    for key, upper in (("tri_lower", 0), ("tri_upper", 1)):
        np.testing.assert_allclose(oracle.invert_triangular_blocks(kats[f"br/{key}/blocks"].reshape(3, 3, 3), upper),
                                   kats[f"br/{key}/inverted_gt"].reshape(3, 3, 3), rtol=1e-4, atol=1e-7)
    assert np.array_equal(oracle.transpose_blocks(kats["br/transpose/blocks"].reshape(3, 3, 3)), kats["br/transpose/transposed_gt"].reshape(3, 3, 3))
    blocks = np.arange(24, dtype=np.float32).reshape(6, 2, 2)
    assert np.array_equal(oracle.fill_in_diagonal_blocks(np.zeros((12, 12), np.float32), blocks), kats["br/fill_diag/matrix_filled_gt"].reshape(12, 12))
    assert np.array_equal(oracle.get_diagonal_blocks(kats["br/get_diag/matrix"].reshape(12, 12), 2), blocks)
    m = oracle.fill_in_sparse_blocks(np.zeros((12, 12), np.float32), blocks, kats["br/fill_sparse/coordinates"].reshape(6, 2))
    assert np.array_equal(m, kats["br/fill_sparse/matrix_filled_gt"].reshape(12, 12))
    m = oracle.fill_in_sparse_blocks(m, kats["br/fill_sparse/blocks_to_transpose"].reshape(3, 2, 2), kats["br/fill_sparse/coordinates_to_transpose"].reshape(3, 2),
                                     (0, 0), True)
    assert np.array_equal(m, kats["br/fill_sparse/matrix_filled_gt2"].reshape(12, 12))
    assert np.array_equal(oracle.get_sparse_blocks(kats["br/get_sparse/matrix"].reshape(12, 12), 2, kats["br/get_sparse/coordinates"].reshape(6, 2)), blocks)


def test_schur_complement_from_reference_pieces(oracle, kats):
    """The oracle's fused Schur complement S = C - B^T D^-1 B equals the reference's own composition of the pinned pieces
    (SchurComplement.cpp:44-77: InvertPositiveSemidefiniteBlocks -> MatmulBlockSparseRowWisePadded -> MatmulBlockSparse(B^T, .))."""
    rng = np.random.default_rng(5)
    n_stem, n_corner = 9, 4
    diag = np.stack([(lambda a: a @ a.T + 6 * np.eye(6))(rng.normal(size=(6, 6))) for _ in range(n_stem + n_corner)]).astype(np.float32)
    coords = np.array([(i, n_stem + j) for i in range(n_stem) for j in rng.choice(n_corner, 2, replace=False)], np.int32)
    wing = (rng.normal(size=(len(coords), 6, 6)) * 0.3).astype(np.float32)
    S = oracle.schur_complement(diag, n_stem, wing, coords)
    dinv, fails = oracle.invert_positive_semidefinite_blocks(diag[:n_stem])
    assert fails == 0
    X = oracle.matmul_block_sparse_row_wise(dinv, wing, coords, padded=True)
    board = np.full((n_stem, n_corner), -1, np.int16)
    board[coords[:, 0], coords[:, 1] - n_stem] = np.arange(len(coords))
    blocks, bc = oracle.matmul_block_sparse(wing, board, 1, X, board, 0)
    expect = np.zeros((6 * n_corner, 6 * n_corner), np.float32)
    expect = oracle.fill_in_diagonal_blocks(expect, diag[n_stem:])
    for blk, (i, j) in zip(blocks, bc):
        expect[6 * i:6 * i + 6, 6 * j:6 * j + 6] -= blk
    np.testing.assert_allclose(S, expect, rtol=1e-4, atol=1e-4)


def subdivided_plane(side, center, subdivisions):
    """test::GenerateXyPlane (cpp/tests/test_utils/geometry.cpp:32-72): a quad of two triangles subdivided by Open3D's legacy
    SubdivideMidpoint (each triangle -> (v0, m01, m20), (m01, v1, m12), (m12, v2, m20), (m01, m12, m20); midpoints in float64)."""
    h = side / 2
    pos = [tuple(p) for p in (np.array([[-h, -h, 0], [-h, h, 0], [h, -h, 0], [h, h, 0]], np.float32) + center).astype(np.float64)]
    tris = [(0, 1, 2), (2, 1, 3)]
    for _ in range(subdivisions):
        mid, new = {}, []

        def m(a, b):
            key = (min(a, b), max(a, b))
            if key not in mid:
                pos.append(tuple(0.5 * (np.array(pos[a]) + np.array(pos[b]))))
                mid[key] = len(pos) - 1
            return mid[key]
        for a, b, c in tris:
            ab, bc, ca = m(a, b), m(b, c), m(c, a)
            new += [(a, ab, ca), (ab, b, bc), (bc, c, ca), (ab, bc, ca)]
        tris = new
    return np.array(pos, np.float32), np.array(tris, np.int64)


def uv_sphere(radius, center, resolution):
    """open3d::geometry::TriangleMesh::CreateSphere (test::GenerateSphere, cpp/tests/test_utils/geometry.cpp:76-94)."""
    v = [(0, 0, radius), (0, 0, -radius)]
    step = np.pi / resolution
    for i in range(1, resolution):
        for j in range(2 * resolution):
            v.append((np.sin(step * i) * np.cos(step * j) * radius, np.sin(step * i) * np.sin(step * j) * radius, np.cos(step * i) * radius))
    t = []
    for j in range(2 * resolution):
        j1 = (j + 1) % (2 * resolution)
        t.append((0, 2 + j, 2 + j1))
        base = 2 + 2 * resolution * (resolution - 2)
        t.append((1, base + j1, base + j))
    for i in range(1, resolution - 1):
        base1 = 2 + 2 * resolution * (i - 1)
        base2 = base1 + 2 * resolution
        for j in range(2 * resolution):
            j1 = (j + 1) % (2 * resolution)
            t += [(base2 + j, base1 + j1, base1 + j), (base2 + j, base2 + j1, base1 + j1)]
    return np.array(v, np.float64).astype(np.float32) + center, np.array(t, np.int64)


def extracted_face_meshes(kats):
    plane = subdivided_plane(float(kats["faces/plane_side"]), kats["faces/plane_center"], int(kats["faces/plane_subdivisions"]))
    sphere = uv_sphere(float(kats["faces/sphere_radius"]), kats["faces/sphere_center"], int(kats["faces/sphere_resolution"]))
    return plane, sphere


def test_extracted_face_fixtures(oracle, kats):
    """cpp/tests/test_extract_face_vertices.cpp:32-110 against the in-repo extracted_face_{vertices,mask}{,_multiple_meshes}.npy: clip
    masks equal face for face (334 of 512 plane faces, 1994 of 4480 with the sphere appended), NDC vertices of kept faces within the
    fixture's float32 rounding (it was produced by a float64 Python script). The reference compares vertices for the first 334
    faces of the multi-mesh case only: the fixture's masked entries are uninitialised memory."""
    (ppos, ptri), (spos, stri) = extracted_face_meshes(kats)
    K, (H, W), (near, far) = kats["faces/intrinsics"], kats["faces/image_size"], kats["faces/near_far"]
    fv, mask = oracle.face_vertices_ndc(ppos, ptri, K, int(H), int(W), float(near), float(far))
    mask = mask.astype(bool)
    assert mask.sum() == kats["faces/kept_counts"][0] and np.array_equal(mask, kats["faces/mask"])
    np.testing.assert_allclose(fv[mask], kats["faces/vertices"][mask], rtol=1e-5, atol=3e-6)
    pos = np.concatenate([ppos, spos])
    tri = np.concatenate([ptri, stri + len(ppos)])
    fv, mask = oracle.face_vertices_ndc(pos, tri, K, int(H), int(W), float(near), float(far))
    mask = mask.astype(bool)
    assert mask.sum() == kats["faces/kept_counts"][1] and np.array_equal(mask, kats["faces/mask_multiple_meshes"])
    np.testing.assert_allclose(fv[mask], kats["faces/vertices_multiple_meshes"][mask], rtol=1e-5, atol=3e-6)
