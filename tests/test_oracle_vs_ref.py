"""CPU tests: the oracle restatement against the reference's OWN device-function headers compiled verbatim (oracle/_ref/libnnrt_ref.so,
built by `make -C oracle ref` from /root/reference; skipped where that library does not exist)."""
import ctypes as C

import numpy as np
import pytest

from dynamicfuion_python_b200 import synthetic


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def ref(oracle):
    lib = oracle.ref_lib()
    if lib is None:
        pytest.skip("oracle/_ref/libnnrt_ref.so not built (needs /root/reference)")
    return lib


@pytest.mark.parametrize("variant", [0, 1, 2, 3])
def test_anchors_match_reference_headers(oracle, ref, variant):
    """WarpUtilities.h FindAnchorsAndWeightsForPoint_Euclidean_* + KnnUtilities.h: indices (slot order included) bit-exact, weights 1e-6."""
    rng = np.random.default_rng(21 + variant)
    pts = rng.uniform(-0.5, 0.5, (3000, 3)).astype(np.float32)
    nodes = rng.uniform(-0.5, 0.5, (97, 3)).astype(np.float32)
    nw = rng.uniform(0.002, 0.02, 97).astype(np.float32)
    min_valid = 2 if variant >= 2 else 0
    cov = 0.07
    a_ref = np.full((3000, 4), -1, np.int32); w_ref = np.zeros((3000, 4), np.float32)
    ref.ref_anchors(_p(pts), C.c_int64(3000), _p(nodes), C.c_int(97), _p(nw), C.c_float(cov * cov), C.c_int(4), C.c_int(min_valid), C.c_int(variant),
                    _p(a_ref), _p(w_ref))
    if variant % 2 == 0:
        a, w = oracle.compute_anchors(pts, nodes, 4, min_valid, cov)
    else:
        a, w = oracle.compute_anchors(pts, nodes, 4, min_valid, 0.0, nw)
    assert np.array_equal(a, a_ref)
    valid = a_ref >= 0
    np.testing.assert_allclose(w[valid], w_ref[valid], rtol=1e-6, atol=1e-9)


def test_blend_warp_matches_reference_headers(oracle, ref):
    """WarpUtilities.h BlendWarp: bit-exact (same operation order, no FMA contraction on either side)."""
    s = synthetic.make_scene("tinyr", 120, 160)
    anchors, weights = oracle.compute_anchors(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    wp, wn = oracle.warp(s["positions"], s["normals"], s["nodes"], s["rotations_gt"], s["translations_gt"], anchors, weights)
    V = s["positions"].shape[0]
    rp, rn = np.empty((V, 3), np.float32), np.empty((V, 3), np.float32)
    nodes, R, t = s["nodes"].copy(), s["rotations_gt"].copy(), s["translations_gt"].copy()
    ref.ref_blend_warp(_p(s["positions"]), _p(s["normals"]), C.c_int64(V), _p(nodes), _p(R), _p(t), _p(anchors), _p(weights), C.c_int(4), _p(rp), _p(rn))
    assert np.array_equal(wp, rp) and np.array_equal(wn, rn)


@pytest.mark.parametrize("K,blur,persp,clip,cull", [(1, 0.5, 1, 0, 1), (4, 0.0, 0, 0, 0), (8, 1.0, 1, 1, 0)])
def test_raster_matches_reference_per_pixel_routine(oracle, ref, K, blur, persp, clip, cull):
    """RayFaceIntersection.h UpdateQueueIfPixelInsideFace driven brute-force over all faces vs the oracle's binned rasteriser:
    face indices, depths, barycentrics and distances bit-exact."""
    H, W = 48, 64
    s = synthetic.make_scene("tiny", H, W)
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], H, W, 0.0, 10.0)
    io, do, bo, so, _ = oracle.rasterize(fv, mask, H, W, blur, K, perspective_correct=bool(persp), clip_barycentric=bool(clip), cull_back_faces=bool(cull))
    kept = np.ascontiguousarray(fv[mask])
    remap = np.nonzero(mask)[0]
    ri = np.empty((H, W, K), np.int64); rd = np.empty((H, W, K), np.float32); rb = np.empty((H, W, K, 3), np.float32); rs = np.empty((H, W, K), np.float32)
    blur_ndc = np.float32(blur) / (np.float32(min(H, W)) / np.float32(2.0))
    ref.ref_rasterize_bruteforce(_p(kept), C.c_int64(kept.shape[0]), C.c_int(H), C.c_int(W), C.c_float(blur_ndc), C.c_int(K), C.c_int(persp), C.c_int(clip),
                                 C.c_int(cull), _p(ri), _p(rd), _p(rb), _p(rs))
    ri = np.where(ri >= 0, remap[np.maximum(ri, 0)], -1)
    assert np.array_equal(io, ri)
    assert np.array_equal(do, rd) and np.array_equal(bo, rb) and np.array_equal(so, rs)


@pytest.mark.parametrize("persp", [1, 0])
def test_barycentric_jacobian_matches_reference_headers(oracle, ref, persp):
    """BarycentricCoordinateJacobians.h (3x9 d rho / d V) through the oracle's rasterized_surface_jacobians on real fragments: 1e-5 relative."""
    H, W = 48, 64
    s = synthetic.make_scene("tiny", H, W)
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], H, W, 0.0, 10.0)
    io, do, bo, so, _ = oracle.rasterize(fv, mask, H, W, 0.5, 1, perspective_correct=bool(persp))
    ndcK, _ = oracle.intrinsics_to_ndc(s["K"], H, W)
    jv, jn = oracle.rasterized_surface_jacobians(s["positions"], s["normals"], s["triangles"], io, bo, ndcK, bool(persp))
    ndcK64 = np.ascontiguousarray(ndcK, np.float64)
    checked = 0
    for v, u in zip(*np.nonzero(io[..., 0] >= 0)):
        if (v * W + u) % 7:
            continue
        f = io[v, u, 0]
        tri = s["triangles"][f]
        v0, v1, v2 = [np.ascontiguousarray(s["positions"][tri[k]], np.float32) for k in range(3)]
        rng_w, rng_h = (2.0 * W / H, 2.0) if W > H else (2.0, 2.0 * H / W)
        ray = np.array([-np.float32(rng_w) / 2 + (np.float32(rng_w) * u + np.float32(rng_w) / 2) / np.float32(W),
                        -np.float32(rng_h) / 2 + (np.float32(rng_h) * v + np.float32(rng_h) / 2) / np.float32(H)], np.float32)
        out = np.empty(27, np.float32)
        bary = np.ascontiguousarray(bo[v, u, 0], np.float32)
        ref.ref_barycentric_jacobian(_p(ray), _p(v0), _p(v1), _p(v2), _p(ndcK64), C.c_int(persp), _p(bary), _p(out))
        drho = out.reshape(3, 9)
        # d w_l / d V = [v0 v1 v2] d rho/dV + rho (x) I3  (RasterizedSurfaceJacobiansImpl.h:150-175)
        Vm = np.stack([v0, v1, v2], 1).astype(np.float64)
        expect = Vm @ drho.astype(np.float64) + np.kron(bary.astype(np.float64)[None, :], np.eye(3))
        np.testing.assert_allclose(jv[v, u], expect, rtol=2e-4, atol=2e-5 * max(1.0, np.abs(expect).max()))
        checked += 1
    assert checked > 20
