"""GPU parity against vectors the REFERENCE holds (tests/golden/fitter_kats.npz; the CPU twin of this file, test_reference_kats.py,
pins the oracle with the same vectors): the reference author's single-pixel Jacobian check script, every enabled one-/two-node fitter
test case through the public depth-image FitToImage overload, the block-sparse known-answer tensors and the extracted-face fixtures.
Plus the un-fused assembly stages (per-pixel node Jacobians, per-edge ARAP quantities) against the oracle before any per-node sum."""
import os

import numpy as np
import pytest
import torch

from test_reference_kats import DENSE_DEPTH_CASES, _CASES, _IDS, dense_depth_inputs, extracted_face_meshes, fitter_case
from util import oracle_fit, reference_all_close, reference_fitter_scenario, rel_close, scene_with_target

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _np(t):
    return t.detach().cpu().numpy()


@pytest.fixture(scope="module")
def kats():
    return np.load(os.path.join(GOLDEN, "fitter_kats.npz"))


@pytest.mark.parametrize("case,pixel", DENSE_DEPTH_CASES)
def test_jacobian_chain_against_reference_check_script(nnrt, oracle, kats, case, pixel):
    """math_check_scripts/dense_depth_jacobians.py (float64): d w_l / dV, d n_l / dV of its two single-pixel cases on the CUDA path."""
    p = f"dense_depth/{case}/"
    pos, nrm, tris, faces, bary, K = dense_depth_inputs(kats, case, pixel)
    ndc_K, _ = oracle.intrinsics_to_ndc(K, 100, 100)
    jv, jn = nnrt.alignment.functional.rasterized_surface_jacobians((pos, nrm, tris), faces, bary, ndc_K, True)
    v, u = pixel
    np.testing.assert_allclose(_np(jv)[v, u], kats[p + "d_wl_dV"], rtol=1e-4, atol=1e-5)
    flat = _np(jn)[v, u].reshape(30)
    np.testing.assert_allclose(flat[:27].reshape(3, 9), kats[p + "d_nl_dV"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(flat[27:], kats[p + "bary_corrected"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("index", _CASES, ids=_IDS)
def test_enabled_reference_fitter_cases(nnrt, oracle, kats, index):
    """cpp/tests/test_deformable_mesh_fitter_{one_node,two_nodes}.cpp, every enabled case, through
    DeformableMeshToImageFitter.fit_to_image(warp_field, mesh, color, DEPTH IMAGE, mask, K, extrinsics, depth_scale = 1) with the
    reference's constructor arguments, checked with the reference's own AllClose(expected, 1, atol) and against the oracle's result."""
    scenario, angle, mode, iterations, rotation_atol, translation_atol, lm = fitter_case(kats, index)
    s = reference_fitter_scenario(oracle, kats, scenario, angle)
    G, A = nnrt.geometry, nnrt.alignment
    wf = G.HierarchicalGraphWarpField(s["nodes"], s["node_coverage"], False, s["anchor_count"], 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 1)
    fitter = A.DeformableMeshToImageFitter(iterations, [A.IterationMode(mode)], 1e-6, True, 10.0, False, 0.01, lm)
    fitter.fit_to_image(wf, (s["positions"], s["normals"], s["triangles"]), None, s["depth"], s["depth_mask"], s["K"], np.eye(4), 1.0)
    fitter.check_status()
    R, t = _np(wf.get_node_rotations()), _np(wf.get_node_translations())
    assert reference_all_close(R, s["expected_rotations"], rotation_atol)
    assert reference_all_close(t, s["expected_translations"], translation_atol)
    ref = oracle.fit_to_image(s["positions"], s["normals"], s["triangles"], s["nodes"], None, len(s["nodes"]), s["anchors"], s["anchor_weights"],
                              s["reference_points"], s["reference_mask"], s["height"], s["width"], s["K"], iteration_count=iterations, lm_factor=lm,
                              iteration_modes=[mode], parallel=False)
    # free-running over up to 6 iterations: two fp32 implementations of a (near-singular: lambda = 0.001) iteration
    np.testing.assert_allclose(R, ref["rotations"], rtol=0, atol=max(10 * rotation_atol, 2e-4))
    np.testing.assert_allclose(t, ref["translations"], rtol=0, atol=max(translation_atol, 2e-4))


def test_block_sparse_matmul_kats(nnrt, kats):
    """cpp/tests/test_linalg_matmul_block_sparse.cpp:30-538 on the CUDA library."""
    L = nnrt.core.linalg
    a, b, bc = kats["bs/rowwise/a_blocks"], kats["bs/rowwise/b_blocks"], kats["bs/rowwise/b_block_coordinates"]
    c, cc = L.matmul_block_sparse_row_wise(a, b, bc)
    np.testing.assert_allclose(_np(c), kats["bs/rowwise/c_blocks_gt"], rtol=1e-5, atol=1e-8)
    assert np.array_equal(_np(cc), kats["bs/rowwise/c_block_coordinates_gt"])
    np.testing.assert_allclose(_np(L.matmul_block_sparse_row_wise_padded(a, b, bc)), kats["bs/rowwise/c_blocks_padded_gt"], rtol=1e-5, atol=1e-8)
    A, Ab, B, Bb = kats["bs/matmul/a_blocks"], kats["bs/matmul/a_breadboard"], kats["bs/matmul/b_blocks"], kats["bs/matmul/b_breadboard"]
    N, TR = L.MatrixPreprocessingOperation.NONE, L.MatrixPreprocessingOperation.TRANSPOSE
    for name, args in dict(c=(A, Ab, N, B, Bb, N), d=(B, Bb, TR, A, Ab, TR), e=(B, Bb, TR, B, Bb, N), f=(A, Ab, N, A, Ab, TR)).items():
        blocks, coords = L.matmul_block_sparse(*args)
        np.testing.assert_allclose(_np(blocks), kats[f"bs/matmul/{name}_blocks_gt"], rtol=1e-5, atol=1e-8)
        assert np.array_equal(_np(coords), kats[f"bs/matmul/{name}_block_coordinates_gt"])
    m = int(kats["bs/spmv/m"])
    np.testing.assert_allclose(_np(L.block_sparse_and_vector_product(kats["bs/spmv/a_blocks"], m, kats["bs/spmv/a_coordinates"], (0, 0), N, kats["bs/spmv/vector_b"])),
                               kats["bs/spmv/vector_c_gt"], rtol=1e-5)
    np.testing.assert_allclose(_np(L.block_sparse_and_vector_product(kats["bs/spmv/b_blocks"], m, kats["bs/spmv/b_coordinates"], (0, 0), TR, kats["bs/spmv/vector_b"])),
                               kats["bs/spmv/vector_d_gt"], rtol=1e-5)
    np.testing.assert_allclose(_np(L.diagonal_block_sparse_and_vector_product(kats["bs/diag_spmv/d_blocks"], kats["bs/diag_spmv/vector_b"])),
                               kats["bs/diag_spmv/vector_c_gt"], rtol=1e-5)


def test_block_routine_kats(nnrt, kats):
    """cpp/tests/test_linalg_block_routines.cpp:32-407 on the CUDA library."""
    L = nnrt.core.linalg
    np.testing.assert_allclose(_np(L.invert_positive_semidefinite_blocks(kats["br/invert_psd/blocks"])), kats["br/invert_psd/inverted_gt"], rtol=1e-4, atol=1e-6)
    with pytest.raises(RuntimeError):
        L.invert_positive_semidefinite_blocks(-np.eye(3, dtype=np.float32)[None])
    for key, uplo in (("tri_lower", L.UpLoTriangular.LOWER), ("tri_upper", L.UpLoTriangular.UPPER)):
        np.testing.assert_allclose(_np(L.invert_triangular_blocks(kats[f"br/{key}/blocks"].reshape(3, 3, 3), uplo)),
                                   kats[f"br/{key}/inverted_gt"].reshape(3, 3, 3), rtol=1e-4, atol=1e-7)
    blocks = torch.tensor(kats["br/transpose/blocks"].reshape(3, 3, 3), device="cuda")
    assert np.array_equal(_np(L.transpose_blocks_in_place(blocks)), kats["br/transpose/transposed_gt"].reshape(3, 3, 3))
    arange = np.arange(24, dtype=np.float32).reshape(6, 2, 2)
    m = torch.zeros((12, 12), device="cuda")
    assert np.array_equal(_np(L.fill_in_diagonal_blocks(m, arange)), kats["br/fill_diag/matrix_filled_gt"].reshape(12, 12))
    assert np.array_equal(_np(L.get_diagonal_blocks(torch.tensor(kats["br/get_diag/matrix"].reshape(12, 12), device="cuda"), 2)), arange)
    m = torch.zeros((12, 12), device="cuda")
    L.fill_in_sparse_blocks(m, arange, kats["br/fill_sparse/coordinates"].reshape(6, 2))
    assert np.array_equal(_np(m), kats["br/fill_sparse/matrix_filled_gt"].reshape(12, 12))
    L.fill_in_sparse_blocks(m, kats["br/fill_sparse/blocks_to_transpose"].reshape(3, 2, 2), kats["br/fill_sparse/coordinates_to_transpose"].reshape(3, 2), (0, 0), True)
    assert np.array_equal(_np(m), kats["br/fill_sparse/matrix_filled_gt2"].reshape(12, 12))
    got = L.get_sparse_blocks(torch.tensor(kats["br/get_sparse/matrix"].reshape(12, 12), device="cuda"), 2, kats["br/get_sparse/coordinates"].reshape(6, 2))
    assert np.array_equal(_np(got), arange)
    with pytest.raises(RuntimeError):
        L.get_diagonal_blocks(torch.zeros((12, 12), device="cuda"), 5)


def test_schur_complement_from_reference_pieces(nnrt):
    """The fused Schur complement equals the reference's composition of the pinned pieces (SchurComplement.cpp:44-77)."""
    L = nnrt.core.linalg
    rng = np.random.default_rng(5)
    n_stem, n_corner = 9, 4
    diag = np.stack([(lambda a: a @ a.T + 6 * np.eye(6))(rng.normal(size=(6, 6))) for _ in range(n_stem + n_corner)]).astype(np.float32)
    coords = np.array([(i, n_stem + j) for i in range(n_stem) for j in rng.choice(n_corner, 2, replace=False)], np.int32)
    wing = (rng.normal(size=(len(coords), 6, 6)) * 0.3).astype(np.float32)
    A = L.BlockSparseArrowheadMatrix()
    A.arrow_base_block_index, A.diagonal_block_count = n_stem, n_stem + n_corner
    A.set_diagonal_blocks(diag)
    A.upper_wing_blocks, A.upper_wing_block_coordinates = wing, coords
    S = _np(L.compute_schur_complement_of_arrow_stem_dense(A))
    X = L.matmul_block_sparse_row_wise_padded(L.invert_positive_semidefinite_blocks(diag[:n_stem]), wing, coords)
    board = np.full((n_stem, n_corner), -1, np.int16)
    board[coords[:, 0], coords[:, 1] - n_stem] = np.arange(len(coords))
    blocks, bc = L.matmul_block_sparse(wing, board, L.MatrixPreprocessingOperation.TRANSPOSE, X, board, L.MatrixPreprocessingOperation.NONE)
    expect = torch.zeros((6 * n_corner, 6 * n_corner), device="cuda")
    L.fill_in_diagonal_blocks(expect, diag[n_stem:])
    product = torch.zeros_like(expect)
    L.fill_in_sparse_blocks(product, blocks, bc)
    np.testing.assert_allclose(S, _np(expect - product), rtol=1e-4, atol=1e-4)


def test_extracted_face_fixtures(nnrt, kats):
    """cpp/tests/test_extract_face_vertices.cpp:32-110 on the CUDA path, single mesh and the multi-mesh overload."""
    (ppos, ptri), (spos, stri) = extracted_face_meshes(kats)
    K, (H, W), (near, far) = kats["faces/intrinsics"], kats["faces/image_size"], kats["faces/near_far"]
    F = nnrt.rendering.functional
    fv, mask = F.get_mesh_ndc_face_vertices_and_clip_mask((ppos, None, ptri), K, (int(H), int(W)), float(near), float(far))
    fv, mask = _np(fv), _np(mask)
    assert mask.sum() == kats["faces/kept_counts"][0] and np.array_equal(mask, kats["faces/mask"])
    assert np.all(fv[~mask] == 0)  # the reference leaves masked faces uninitialised and its test zeroes them (:54-55)
    np.testing.assert_allclose(fv[mask], kats["faces/vertices"][mask], rtol=1e-5, atol=3e-6)
    fv, mask, counts = F.get_mesh_ndc_face_vertices_and_clip_mask([(ppos, None, ptri), (spos, None, stri)], K, (int(H), int(W)), float(near), float(far))
    fv, mask = _np(fv), _np(mask)
    assert _np(counts).tolist() == [len(ptri), len(stri)]
    assert mask.sum() == kats["faces/kept_counts"][1] and np.array_equal(mask, kats["faces/mask_multiple_meshes"])
    np.testing.assert_allclose(fv[mask], kats["faces/vertices_multiple_meshes"][mask], rtol=1e-5, atol=3e-6)


def test_interpolate_vertex_attributes(nnrt, oracle):
    """nnrt.rendering.functional.interpolate_vertex_attributes (InterpolateFaceAttributesImpl.h:31-86) against the oracle: K = 1 and
    K = 4 fragments, 3 and 5 channels, empty pixels give zeros and stop the fragment list."""
    from dynamicfuion_python_b200 import synthetic
    H, W = 120, 160
    s = synthetic.make_scene("tinyr", H, W)
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], H, W, 0.0, 10.0)
    rng = np.random.default_rng(8)
    for K, C in ((1, 3), (4, 5)):
        idx, _, bary, _, _ = oracle.rasterize(fv, mask, H, W, 1.0, K, perspective_correct=True, cull_back_faces=False)
        attrs = rng.normal(size=(len(s["triangles"]), 3, C)).astype(np.float32)
        got = _np(nnrt.rendering.functional.interpolate_vertex_attributes(idx, bary, attrs))
        want = oracle.interpolate_vertex_attributes(idx, bary, attrs)
        assert got.shape == (H, W, K, C)
        assert np.array_equal(got, want)
        assert np.all(got[idx < 0] == 0)


def _stage_inputs(nnrt, oracle, s, rotations, translations):
    H, W = s["height"], s["width"]
    wp, wn = oracle.warp(s["positions"], s["normals"], s["nodes"], rotations, translations, s["anchors"], s["anchor_weights"])
    fv, mask = oracle.face_vertices_ndc(wp, s["triangles"], s["K"], H, W, 0.0, 10.0)
    idx, depth, bary, _, _ = oracle.rasterize(fv, mask, H, W, 0.5, 1, perspective_correct=True)
    ndcK, _ = oracle.intrinsics_to_ndc(s["K"], H, W)
    rv, rn = oracle.rasterized_surface_jacobians(wp, wn, s["triangles"], idx, bary, ndcK, True)
    wvj, wnj = oracle.warped_surface_jacobians(s["positions"], s["normals"], s["nodes"], rotations, s["anchors"], s["anchor_weights"])
    face_normals = wn[s["triangles"]]
    normals = oracle.interpolate_vertex_attributes(idx, bary, face_normals).reshape(-1, 3)
    points, rendered_mask = oracle.unproject_depth(depth[..., 0], s["K"], None, 1.0, 10.0)
    residual_mask = np.logical_and(np.asarray(rendered_mask).reshape(-1), np.asarray(s["reference_mask"]).reshape(-1))
    point_map = (points.reshape(-1, 3) - np.asarray(s["reference_points"]).reshape(-1, 3)).astype(np.float32)
    return dict(idx=idx, rv=rv, rn=rn, wvj=wvj, wnj=wnj, normals=normals, point_map=point_map, residual_mask=residual_mask)


@pytest.mark.parametrize("config,h,w,tukey", [("tinyr", 120, 160, False), ("tinyr", 120, 160, True), ("C2r", 240, 320, False)])
def test_pixel_vertex_anchor_jacobians_per_pixel(nnrt, oracle, config, h, w, tukey):
    """nnrt_b200_pixel_vertex_anchor_jacobians against the oracle's PixelVertexAnchorJacobiansAndNodeAssociations per PIXEL and per
    face-anchor entry -- before any per-node sum, so nothing cancels: 1e-5 relative with a 1e-6 floor (identical inputs, identical
    arithmetic up to the order of two additions)."""
    s = scene_with_target(oracle, config, h, w)
    N = len(s["nodes"])
    R, t = s["rotations_gt"] if config == "tinyr" else np.tile(np.eye(3, dtype=np.float32), (N, 1, 1)), (0.5 * s["translations_gt"]).astype(np.float32)
    st = _stage_inputs(nnrt, oracle, s, R, t)
    fa, fc = oracle.associate_faces_with_anchors(s["triangles"], s["anchors"])
    pj_o, pc_o, lists, nc_o, dropped = oracle.pixel_vertex_anchor_jacobians(st["rv"], st["rn"], st["wvj"], st["wnj"], st["point_map"], st["normals"], st["residual_mask"],
                                                                            st["idx"], fa, fc, N, tukey, 0.01)
    assert dropped == 0
    F = nnrt.alignment.functional
    entries, counts = F.associate_faces_with_anchors(s["triangles"], s["anchors"])
    pj, pc, nc = F.pixel_vertex_anchor_jacobians_and_node_associations(st["rv"], st["rn"], st["wvj"], st["wnj"], st["point_map"], st["normals"], st["residual_mask"],
                                                                       st["idx"], s["triangles"], entries, counts, N, 0.01, 0, tukey)
    listed = st["residual_mask"].copy()
    if tukey:  # pixels beyond the cutoff are dropped on the Jacobian side (PixelVertexAnchorJacobiansImpl.h:200-204)
        listed &= np.abs((st["normals"] * st["point_map"]).sum(1)) <= 0.0101
    assert np.array_equal(_np(pc)[listed], pc_o[listed])
    assert np.array_equal(_np(nc), nc_o)
    ok, worst = rel_close(_np(pj), pj_o, 1e-5, floor_scale=1e-6)
    assert ok, worst
    for mode, cols in ((1, slice(3, 6)), (2, slice(0, 3))):  # TRANSLATION_ONLY / ROTATION_ONLY: the corresponding half, stride 3
        pj3, _, _ = F.pixel_vertex_anchor_jacobians_and_node_associations(st["rv"], st["rn"], st["wvj"], st["wnj"], st["point_map"], st["normals"], st["residual_mask"],
                                                                          st["idx"], s["triangles"], entries, counts, N, 0.01, mode, tukey)
        assert np.array_equal(_np(pj3), _np(pj)[:, :, cols])
    # the fitter's fused per-node sums against the sums of these per-pixel Jacobians in float64: what is left is fp32 summation only
    H64 = np.zeros((N, 6, 6))
    pjd = _np(pj).astype(np.float64)
    e = _np(entries)
    faces = st["idx"].reshape(-1)
    for px in np.nonzero(listed)[0]:
        for k in range(pc_o[px]):
            J = pjd[px, k]
            H64[e[faces[px], k, 0]] += np.outer(J, J)
    if not tukey:
        G = nnrt.geometry
        wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                     G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
        wf.set_node_rotations(R); wf.set_node_translations(t)
        fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=1, preconditioning_dampening_factor=0.0, arap_term_weight=0.0)
        fitter.fit_to_image(wf, (s["positions"], s["normals"], s["triangles"]), None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4),
                            (h, w), anchors=s["anchors"], anchor_weights=s["anchor_weights"])
        Hd = _np(fitter.diagnostics()["hessian_diagonal"]).astype(np.float64)
        scale = np.sqrt(np.einsum("nii,njj->nij", H64, H64))  # |H_ij| <= sqrt(H_ii H_jj): the natural size of an entry of a Gram matrix
        assert np.abs(Hd - H64).max() <= 2e-5 * scale.max()
        assert np.all(np.abs(Hd - H64) <= 1e-4 * np.maximum(scale, 1e-6 * scale.max()))


@pytest.mark.parametrize("fixed_coverage,huber", [(True, False), (True, True), (False, False)])
def test_arap_term_per_edge(nnrt, oracle, fixed_coverage, huber):
    """nnrt_b200_arap_term against the oracle's ComputeEdgeResiduals / ComputeDenseArapEdgeJacobians /
    ComputeArapBlockSparseHessianApproximation / ComputeNegativeArapGradient per EDGE and per node: bit-exact (same operations in
    the same order; per-node sums in ascending edge order)."""
    from dynamicfuion_python_b200 import synthetic
    s = synthetic.make_scene("C2r", 120, 160)
    N, E = len(s["nodes"]), len(s["edges"])
    R, t = s["rotations_gt"], s["translations_gt"]
    G = nnrt.geometry
    if fixed_coverage:
        wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                     G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])
        kw = dict(edge_layer_indices=s["edge_layer_indices"], layer_radii=s["layer_radii"])
    else:
        wf = G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                                     G.WarpNodeCoverageComputationMethod.MINIMAL_K_NEIGHBOR_NODE_DISTANCE, 4)
        kw = dict(node_coverage_weights=_np(wf.node_coverage_weights))
    wf.set_node_rotations(R); wf.set_node_translations(t)
    out = nnrt.alignment.functional.arap_term(wf, 200.0, huber, 1e-4)
    res, jac, diag, wing, coords, g = oracle.arap_terms(s["edges"], s["nodes"], R, t, 200.0, s["arrow_base"], **kw)
    if huber:
        dl, half = np.float32(1e-4), np.float32(0.5) * np.float32(1e-4) * np.float32(1e-4)
        res = np.where(res >= dl, np.abs(res) - half, np.float32(0.5) * res * res).astype(np.float32)
        from oracle import oracle_py
        import ctypes as C
        g = np.empty(6 * N, np.float32)
        oracle_py.lib().oracle_negative_arap_gradient(res.ctypes.data_as(C.c_void_p), jac.ctypes.data_as(C.c_void_p),
                                                      np.ascontiguousarray(s["edges"], np.int32).ctypes.data_as(C.c_void_p), C.c_int64(E), C.c_int(N),
                                                      g.ctypes.data_as(C.c_void_p))
    assert np.array_equal(_np(out["edge_residuals"]), res)
    assert np.array_equal(_np(out["edge_jacobians"]), jac)
    assert np.array_equal(_np(out["upper_wing_block_coordinates"]), coords)
    assert np.array_equal(_np(out["upper_wing_blocks"]), wing)
    assert np.array_equal(_np(out["hessian_diagonal_blocks"]), diag)
    assert np.array_equal(_np(out["negative_gradient"]), g)


def test_three_fit_to_image_overloads_agree(nnrt, oracle, kats):
    """FitToImage's RGB-D overload forwards to the depth-image overload, which unprojects and forwards to the point-cloud overload
    (DeformableMeshToImageFitter.cpp:278-329): the three must leave the warp field in the same state, bit for bit."""
    import types
    scenario, angle, mode, iterations, _, _, lm = fitter_case(kats, _CASES[0])
    s = reference_fitter_scenario(oracle, kats, scenario, angle)
    G, A = nnrt.geometry, nnrt.alignment
    mesh = (s["positions"], s["normals"], s["triangles"])
    color = np.zeros((s["height"], s["width"], 3), np.uint8)
    results = []
    for call in ("rgbd_object", "rgbd_pair", "depth", "points"):
        wf = G.HierarchicalGraphWarpField(s["nodes"], s["node_coverage"], False, s["anchor_count"], 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 1)
        fitter = A.DeformableMeshToImageFitter(iterations, [A.IterationMode(mode)], 1e-6, True, 10.0, False, 0.01, lm)
        if call == "rgbd_object":  # anything with .color / .depth, like open3d.t.geometry.RGBDImage
            fitter.fit_to_image(wf, mesh, types.SimpleNamespace(color=color, depth=s["depth"]), s["depth_mask"], s["K"], np.eye(4), 1.0)
        elif call == "rgbd_pair":
            fitter.fit_to_image(wf, mesh, (color, s["depth"]), s["depth_mask"], s["K"], np.eye(4), 1.0)
        elif call == "depth":
            fitter.fit_to_image(wf, mesh, color, s["depth"], s["depth_mask"], s["K"], np.eye(4), 1.0)
        else:
            fitter.fit_to_image(wf, mesh, color, s["reference_points"], s["reference_mask"], s["K"], np.eye(4), (s["height"], s["width"]))
        fitter.check_status()
        results.append((_np(wf.get_node_rotations()), _np(wf.get_node_translations())))
    for R, t in results[1:3]:
        assert np.array_equal(R, results[0][0]) and np.array_equal(t, results[0][1])
    # the point-cloud overload gets the oracle's unprojection instead of the device's: equal up to that
    np.testing.assert_allclose(results[3][1], results[0][1], rtol=0, atol=1e-6)
