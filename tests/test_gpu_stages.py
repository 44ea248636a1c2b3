"""GPU parity of the stage-level C ABI against the CPU oracle on identical seeded inputs."""
import numpy as np
import pytest
import torch

from util import rel_close
from dynamicfuion_python_b200 import synthetic

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("variant", ["fixed", "variable", "fixed_threshold", "variable_threshold"])
def test_anchors_bit_exact(nnrt, oracle, variant):
    rng = np.random.default_rng(3)
    pts = rng.uniform(-0.5, 0.5, (20000, 3)).astype(np.float32)
    nodes = rng.uniform(-0.5, 0.5, (250, 3)).astype(np.float32)
    nw = rng.uniform(0.002, 0.02, 250).astype(np.float32)
    min_valid = 2 if "threshold" in variant else 0
    f = nnrt.geometry.functional
    if variant.startswith("fixed"):
        a, w = f.compute_anchors_and_weights_euclidean_fixed_node_weight(pts, nodes, 4, min_valid, 0.05)
        ao, wo = oracle.compute_anchors(pts, nodes, 4, min_valid, 0.05)
    else:
        a, w = f.compute_anchors_and_weights_euclidean_variable_node_weight(pts, nodes, nw, 4, min_valid)
        ao, wo = oracle.compute_anchors(pts, nodes, 4, min_valid, 0.0, nw)
    assert np.array_equal(_np(a), ao)  # slot order included
    np.testing.assert_allclose(_np(w), wo, rtol=2e-5, atol=1e-7)


@pytest.mark.parametrize("anchor_count,node_count", [(1, 1), (4, 3), (8, 2000)])
def test_anchors_edge_cases(nnrt, oracle, anchor_count, node_count):
    rng = np.random.default_rng(4)
    pts = rng.uniform(-1, 1, (513, 3)).astype(np.float32)
    nodes = rng.uniform(-1, 1, (node_count, 3)).astype(np.float32)
    a, w = nnrt.geometry.functional.compute_anchors_and_weights_euclidean_fixed_node_weight(pts, nodes, anchor_count, 0, 0.3)
    ao, wo = oracle.compute_anchors(pts, nodes, anchor_count, 0, 0.3)
    assert np.array_equal(_np(a), ao)
    np.testing.assert_allclose(_np(w), wo, rtol=2e-5, atol=1e-7)


def test_warp_matches_oracle(nnrt, oracle):
    s = synthetic.make_scene("tinyr", 120, 160)
    anchors, weights = oracle.compute_anchors(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    ext = np.eye(4)
    ext[:3, 3] = [0.01, -0.02, 0.03]
    wf = nnrt.geometry.GraphWarpField(s["nodes"], s["node_coverage"], False, 4, 0, nnrt.geometry.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    wf.set_node_rotations(s["rotations_gt"]); wf.set_node_translations(s["translations_gt"])
    out = wf.warp_mesh((s["positions"], s["normals"], s["triangles"]), anchors, weights, True, ext)
    wp, wn = oracle.warp(s["positions"], s["normals"], s["nodes"], s["rotations_gt"], s["translations_gt"], anchors, weights, ext)
    assert np.array_equal(_np(out.positions), wp)  # same op order, no FMA: bit-exact
    assert np.array_equal(_np(out.normals), wn)


def test_face_vertices_and_raster_bit_exact(nnrt, oracle):
    H, W = 240, 320
    s = synthetic.make_scene("tiny", H, W)
    fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask((s["positions"], s["normals"], s["triangles"]), s["K"], (H, W), 0.0, 10.0)
    fvo, masko = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], H, W, 0.0, 10.0)
    assert np.array_equal(_np(mask), masko)
    assert np.array_equal(_np(fv), fvo)
    for K, blur, persp, clip, cull in [(1, 0.5, True, False, True), (4, 0.0, False, False, False), (8, 1.0, True, True, False)]:
        idx, depth, bary, dist = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (H, W), blur, K, -1, -1, persp, clip, cull)
        io, do, bo, so, _ = oracle.rasterize(fvo, masko, H, W, blur, K, perspective_correct=persp, clip_barycentric=clip, cull_back_faces=cull)
        assert np.array_equal(_np(idx), io), f"K={K}: {(_np(idx) != io).sum()} differing fragments"
        assert np.array_equal(_np(depth), do)
        assert np.array_equal(_np(bary), bo)
        assert np.array_equal(_np(dist), so)


def test_raster_large_faces_and_empty(nnrt, oracle):
    H, W = 96, 128
    s = synthetic.make_scene("C1", H, W)  # plane of ~2k triangles spans most of the image: band-queue path
    fvo, masko = oracle.face_vertices_ndc(s["positions"] * np.array([1, 1, 0.5], np.float32), s["triangles"], s["K"], H, W)
    idx, depth, bary, dist = nnrt.rendering.rasterize_ndc_triangles(fvo, None, (H, W), 0.0, 2)
    io, do, bo, so, _ = oracle.rasterize(fvo, None, H, W, 0.0, 2)
    assert np.array_equal(_np(idx), io) and np.array_equal(_np(depth), do)
    idx, depth, bary, dist = nnrt.rendering.rasterize_ndc_triangles(np.zeros((0, 3, 3), np.float32), None, (H, W), 0.0, 1)
    assert (_np(idx) == -1).all() and (_np(depth) == -1).all()
    with pytest.raises(RuntimeError):
        nnrt.rendering.rasterize_ndc_triangles(fvo, None, (H, W), 0.0, 9)
    with pytest.raises(RuntimeError):
        nnrt.rendering.rasterize_ndc_triangles(fvo, None, (H, W), 0.0, 1, bin_size=2)


def test_small_ops(nnrt, oracle):
    rng = np.random.default_rng(5)
    depth = rng.uniform(0.0, 4.0, (60, 80)).astype(np.float32)
    K = synthetic.default_intrinsics(60, 80)
    ext = np.eye(4); ext[:3, 3] = [0.1, 0.2, 0.3]
    pts, mask = nnrt.geometry.functional.unproject_raster_depth_without_filtering(depth, K, ext, 1.0, 3.0)
    po, mo = oracle.unproject_depth(depth, K, ext, 1.0, 3.0)
    assert np.array_equal(_np(mask), mo)
    np.testing.assert_allclose(_np(pts), po, rtol=1e-6, atol=1e-7)
    v = rng.normal(size=(100, 3)).astype(np.float32) * 0.3
    np.testing.assert_allclose(_np(nnrt.core.linalg.AxisAngleVectorsToMatricesRodrigues(v)), oracle.rodrigues(v), rtol=1e-5, atol=1e-6)
    a, b = rng.normal(size=(50, 3, 3)).astype(np.float32), rng.normal(size=(50, 3, 3)).astype(np.float32)
    np.testing.assert_allclose(_np(nnrt.core.matmul3d(a, b)), a @ b, rtol=1e-5, atol=1e-6)
    n1, v1, v2 = [rng.normal(size=(1000, 3)).astype(np.float32) for _ in range(3)]
    d = nnrt.geometry.functional.compute_point_to_plane_distances((v1, n1, None), v2)
    assert np.array_equal(_np(d), oracle.point_to_plane(n1, v1, v2))


def test_jacobian_stages(nnrt, oracle):
    H, W = 120, 160
    s = synthetic.make_scene("tinyr", H, W)
    anchors, weights = oracle.compute_anchors(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    wpo, wno = oracle.warp(s["positions"], s["normals"], s["nodes"], s["rotations_gt"], s["translations_gt"], anchors, weights)
    fvo, masko = oracle.face_vertices_ndc(wpo, s["triangles"], s["K"], H, W, 0.0, 10.0)
    io, do, bo, so, _ = oracle.rasterize(fvo, masko, H, W, 0.5, 1, perspective_correct=True)
    ndcK, _ = oracle.intrinsics_to_ndc(s["K"], H, W)
    for persp in (True, False):
        jv, jn = nnrt.alignment.functional.rasterized_surface_jacobians((wpo, wno, s["triangles"]), io, bo, ndcK, persp)
        jvo, jno = oracle.rasterized_surface_jacobians(wpo, wno, s["triangles"], io, bo, ndcK, persp)
        ok, worst = rel_close(_np(jv), jvo, 1e-4)
        assert ok, worst
        ok, worst = rel_close(_np(jn), jno, 1e-4)
        assert ok, worst
    wf = nnrt.geometry.GraphWarpField(s["nodes"], s["node_coverage"], False, 4, 0, nnrt.geometry.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    wf.set_node_rotations(s["rotations_gt"])
    jv, jn = nnrt.alignment.functional.warped_surface_jacobians((s["positions"], s["normals"], s["triangles"]), wf, anchors, weights)
    jvo, jno = oracle.warped_surface_jacobians(s["positions"], s["normals"], s["nodes"], s["rotations_gt"], anchors, weights)
    assert np.array_equal(_np(jv), jvo) and np.array_equal(_np(jn), jno)
    entries, counts = nnrt.alignment.functional.associate_faces_with_anchors(s["triangles"], anchors)
    fa, co = oracle.associate_faces_with_anchors(s["triangles"], anchors)
    assert np.array_equal(_np(counts), co)
    e = _np(entries)
    for f in range(0, len(co), 37):
        for k in range(co[f]):
            assert e[f, k, 0] == fa[f, k, 0]
            for fv in range(3):
                slot = (int(e[f, k, 1]) >> (8 * fv)) & 0xFF
                assert (slot if slot != 0xFF else -1) == fa[f, k, 4 + fv]


def test_shared_divisor_quotient_is_the_ieee_quotient():
    """csrc/common.cuh::SharedDivisor (the division of the barycentric Jacobian chain: one reciprocal per denominator, a corrected multiply
    per numerator) against the device's IEEE division, BIT FOR BIT, over 2^24 pairs: half spread over twenty decades, half drawn from
    the ranges the chain divides in (areas and squared areas of NDC triangles from 1e-8 up, depths 0.3 .. 5 m, g^2 up to z^4)."""
    from dynamicfuion_python_b200 import _lib
    lib = _lib.lib()
    n = 1 << 24
    g = torch.Generator(device="cuda").manual_seed(11)
    def spread(lo, hi, count, signed):
        magnitude = torch.exp(torch.empty(count, device="cuda").uniform_(float(np.log(lo)), float(np.log(hi)), generator=g))
        magnitude = magnitude * torch.empty(count, device="cuda").uniform_(1.0, 2.0, generator=g)
        if signed:
            magnitude = magnitude * (torch.randint(0, 2, (count,), device="cuda", generator=g).float() * 2 - 1)
        return magnitude
    half = n // 2
    numerators = torch.cat([spread(1e-10, 1e10, half, True), spread(1e-9, 1e3, half, True)])
    divisors = torch.cat([spread(1e-10, 1e10, half, True), spread(1e-8, 7e2, half, False)])
    numerators[:4] = torch.tensor([0.0, 1.0, -3.0, 1e-8], device="cuda")  # exact cases: zero numerator, exact quotients
    divisors[:4] = torch.tensor([2.5, 1.0, 3.0, 1e-8], device="cuda")
    shared, ieee = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    rc = lib.nnrt_b200_selfcheck_shared_divisor_quotients(numerators.data_ptr(), divisors.data_ptr(), n, shared.data_ptr(), ieee.data_ptr(),
                                                          torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.nnrt_b200_last_error()
    torch.cuda.synchronize()
    assert torch.equal(ieee, numerators / divisors)  # the comparison arm is the division torch (and the oracle's host code) performs
    mismatches = int((shared.view(torch.int32) != ieee.view(torch.int32)).sum())
    assert mismatches == 0, mismatches
