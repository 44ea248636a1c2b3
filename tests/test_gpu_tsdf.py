"""Non-rigid TSDF integration (cpp/geometry/kernel/NonRigidSurfaceVoxelBlockGridImpl.h:50-229): the voxel kernel
against the oracle on a seeded scene — a depth frame rendered from the relief mesh, voxel blocks around the surface, a moving warp
field. Index-like outcomes (which voxels are updated, which voxel owns a pixel of the cosine image) must agree exactly; values to the
fp32 rounding of expf amplified by the projective distance's cancellation (1e-4 of the truncated range)."""
import numpy as np
import pytest
import torch

from dynamicfuion_python_b200 import synthetic

pytestmark = pytest.mark.gpu


def _scene(oracle, depth_u16, with_color, variable_coverage):
    H, W = 120, 160
    s = synthetic.make_scene("tinyr", H, W)
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], H, W, 0.0, 10.0)
    _, depth, _, _, _ = oracle.rasterize(fv, mask, H, W, 0.0, 1, perspective_correct=True)
    depth = np.where(depth[..., 0] > 0, depth[..., 0], 0).astype(np.float32)
    points, _ = oracle.unproject_depth(depth, s["K"], None, 1.0, 10.0)
    normals = oracle.compute_ordered_point_cloud_normals(points, H, W).reshape(H, W, 3)
    depth_scale = 1.0
    if depth_u16:
        depth, depth_scale = np.round(depth * 1000.0).astype(np.uint16), 1000.0
    # voxel blocks (8^3 voxels of 4 mm) covering the surface's bounding box, plus empty capacity
    voxel_size, R = 0.004, 8
    block = voxel_size * R
    lo = np.floor(s["positions"].min(0) / block).astype(np.int32) - 1
    hi = np.floor(s["positions"].max(0) / block).astype(np.int32) + 1
    grid = np.stack(np.meshgrid(*[np.arange(lo[k], hi[k] + 1) for k in range(3)], indexing="ij"), -1).reshape(-1, 3).astype(np.int32)
    rng = np.random.default_rng(9)
    capacity = len(grid) + 7
    keys = np.zeros((capacity, 3), np.int32)
    slots = rng.permutation(capacity)[:len(grid)]
    keys[slots] = grid
    active = np.sort(slots)[::2].astype(np.int64)  # every other block is integrated
    tsdf = rng.uniform(-1, 1, (capacity, R, R, R)).astype(np.float32)
    weight = rng.integers(0, 5, (capacity, R, R, R)).astype(np.float32 if not depth_u16 else np.uint16)
    voxel_color = rng.uniform(0, 255, (capacity, R, R, R, 3)).astype(np.float32) if with_color else None
    color = None
    if with_color:
        color = rng.uniform(0, 1, (H, W, 3)).astype(np.float32) if not depth_u16 else rng.integers(0, 256, (H, W, 3)).astype(np.uint8)
    node_weights = rng.uniform(0.01, 0.03, len(s["nodes"])).astype(np.float32) ** 2 if variable_coverage else None
    return dict(s=s, H=H, W=W, depth=depth, depth_scale=depth_scale, normals=normals, voxel_size=voxel_size, R=R, keys=keys, active=active,
                tsdf=tsdf, weight=weight, voxel_color=voxel_color, color=color, node_weights=node_weights)


@pytest.mark.parametrize("depth_u16,with_color,variable_coverage,min_valid", [(False, False, False, 0), (True, True, False, 2), (False, True, True, 1)])
def test_integrate_non_rigid_matches_oracle(nnrt, oracle, depth_u16, with_color, variable_coverage, min_valid):
    c = _scene(oracle, depth_u16, with_color, variable_coverage)
    s = c["s"]
    coverage = 0.03
    ext = np.eye(4)
    ext[:3, 3] = [0.002, -0.001, 0.003]
    tsdf_o = c["tsdf"].copy()
    color_o = None if c["voxel_color"] is None else c["voxel_color"].copy()
    cos_o = oracle.integrate_non_rigid(c["active"], c["keys"], tsdf_o, c["weight"], color_o, c["R"], c["voxel_size"], 2.0 * c["voxel_size"], c["depth"],
                                       c["color"], c["normals"], s["K"], s["K"], ext, s["nodes"], s["rotations_gt"], s["translations_gt"], coverage, 4,
                                       min_valid, c["depth_scale"], 3.0, c["node_weights"])
    G = nnrt.geometry
    method = G.WarpNodeCoverageComputationMethod
    wf = G.GraphWarpField(s["nodes"], coverage, True, 4, min_valid,
                          method.MINIMAL_K_NEIGHBOR_NODE_DISTANCE if variable_coverage else method.FIXED_NODE_COVERAGE)
    if variable_coverage:
        wf.node_coverage_weights = torch.as_tensor(c["node_weights"]).cuda()
    wf.set_node_rotations(s["rotations_gt"])
    wf.set_node_translations(s["translations_gt"])
    weight_t = torch.as_tensor(c["weight"].view(np.int16) if depth_u16 else c["weight"]).cuda()
    values = dict(tsdf=torch.as_tensor(c["tsdf"]).cuda(), weight=weight_t)
    if with_color:
        values["color"] = torch.as_tensor(c["voxel_color"]).cuda()
    cos = G.functional.integrate_non_rigid_voxel_blocks(c["active"], c["keys"], values, c["R"], c["voxel_size"], wf, c["depth"], c["color"], c["normals"],
                                                        s["K"], s["K"], ext, c["depth_scale"], 3.0, 2.0)
    tsdf_g = values["tsdf"].cpu().numpy()
    changed_o, changed_g = tsdf_o != c["tsdf"], tsdf_g != c["tsdf"]
    assert changed_o.sum() > 1500  # the test integrates something
    # expf differs in the last bit between device and glibc. One ulp of a blend weight moves the warped voxel by ~1e-7 m: (i) the
    # projective distance (depth - z, a few mm) over the 8 mm truncation turns that into ~1e-5 of tsdf; (ii) a voxel projecting within
    # 1e-7 px of a rounding boundary, or sitting exactly at the truncation / cosine gate, can land on the other side. (ii) is counted,
    # not tolerated silently: at most 1 voxel in 1000 may differ in kind.
    flipped = (changed_o != changed_g).sum()
    assert flipped <= 1e-3 * changed_o.sum(), f"{flipped} of {changed_o.sum()} voxels gated differently"
    both = changed_o & changed_g
    off = np.abs(tsdf_g - tsdf_o)[both] > 1e-4
    assert off.sum() <= 1e-3 * both.sum(), f"{off.sum()} of {both.sum()} integrated voxels differ by more than 1e-4"
    cos_g = cos.cpu().numpy()
    same_pixels = (cos_g != 0) == (cos_o != 0)
    assert (~same_pixels).sum() <= 1e-3 * (cos_o != 0).sum()
    cos_off = np.abs(cos_g - cos_o)[same_pixels] > 1e-5
    assert cos_off.sum() <= 1e-3 * (cos_o != 0).sum(), f"{cos_off.sum()} cosine pixels differ"  # the same voxel owns (almost) every pixel
    inactive = np.ones(len(c["keys"]), bool)
    inactive[c["active"]] = False
    assert np.array_equal(tsdf_g[inactive], c["tsdf"][inactive])
    if with_color:
        color_off = np.abs(values["color"].cpu().numpy() - color_o).max(-1)[both] > 1e-3  # colours do not depend on the warped depth
        assert color_off.sum() <= 1e-3 * both.sum()
        assert (color_o != c["voxel_color"]).any()
    assert np.array_equal(values["weight"].cpu().numpy(), weight_t.cpu().numpy())  # weights are read only, as in the reference


def test_block_selection_matches_oracle(nnrt, oracle):
    """The device work of FindBlocksIntersectingTruncationRegion: bounding boxes of warped blocks (quirks of the reference included) and
    the box / truncation-segment mask, which visits only the pixels inside a box's projected rectangle yet must equal the reference's
    all-pairs result."""
    c = _scene(oracle, False, False, False)
    s = c["s"]
    G = nnrt.geometry
    f = G.functional
    ext = np.eye(4)
    ext[:3, 3] = [0.002, -0.001, 0.003]
    for min_valid, coverage in ((0, 0.3), (2, 0.2), (3, 0.05)):
        # the reference adds a side in metres to a key in blocks: keys near the surface's metric coordinates keep the test meaningful
        rng = np.random.default_rng(min_valid)
        keys = np.concatenate([rng.integers(-1, 1, (300, 3)), np.array([[0, 0, 0], [-1, -1, 0], [5, 5, 5]])]).astype(np.int32)
        wf = G.GraphWarpField(s["nodes"], coverage, True, 4, min_valid, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
        wf.set_node_rotations(s["rotations_gt"])
        wf.set_node_translations(s["translations_gt"])
        boxes = f.bounding_boxes_of_warped_blocks(keys, wf, c["voxel_size"], c["R"], ext)
        want = oracle.bounding_boxes_of_warped_blocks(keys, s["nodes"], s["rotations_gt"], s["translations_gt"], coverage, 4, min_valid, c["voxel_size"], c["R"], ext)
        np.testing.assert_allclose(boxes.cpu().numpy(), want, rtol=1e-5, atol=1e-6)
    # boxes around the surface, in front of the camera, behind it, inverted (max < min) and far off to the side
    rng = np.random.default_rng(5)
    centres = np.concatenate([s["positions"][rng.choice(len(s["positions"]), 400)] + rng.normal(0, 0.03, (400, 3)), rng.uniform(-1, 1, (200, 3)) * [1.0, 1.0, 2.0]])
    half = rng.uniform(0.002, 0.05, (600, 3))
    boxes = np.concatenate([centres - half, centres + half], 1).astype(np.float32)
    boxes[-20:] = np.concatenate([boxes[-20:, 3:], boxes[-20:, :3]], 1)  # inverted
    boxes[-40:-20, 2] = -0.5  # reach behind the camera
    for depth, scale, stride in ((c["depth"], 1.0, 4), (np.round(c["depth"] * 1000).astype(np.uint16), 1000.0, 1), (c["depth"], 1.0, 7)):
        got = f.axis_aligned_boxes_intersecting_surface_mask(boxes, depth, s["K"], scale, 3.0, stride, 2 * c["voxel_size"])
        want = oracle.boxes_intersecting_surface_mask(boxes, s["K"], depth, scale, 3.0, stride, 2 * c["voxel_size"])
        assert 20 < want.sum() < 580
        assert np.array_equal(got.cpu().numpy(), want)


@pytest.mark.parametrize("node_count,coverage,variable,min_valid", [(30, 0.03, False, 0), (2500, 0.012, False, 1), (2500, 0.01, True, 2), (1500, 0.5, False, 0)])
def test_block_shortlist_equals_full_scan_bit_for_bit(nnrt, oracle, node_count, coverage, variable, min_valid):
    """The default entry point scans a per-block shortlist of nodes; the full scan over every node is the yardstick. They must agree in
    every bit of tsdf, colour and cosine — the pruning argument (running A-th distance + block diameter) leaves no slot different.
    (1500 nodes packed into the surface's neighbourhood with a huge coverage overflow the 1024-entry shortlist: the fallback path.)"""
    c = _scene(oracle, False, True, variable)
    s = c["s"]
    rng = np.random.default_rng(node_count)
    if node_count == 30:
        nodes = s["nodes"]
    elif node_count == 1500:
        nodes = (s["positions"][rng.choice(len(s["positions"]), node_count, replace=False)] + rng.normal(0, 0.002, (node_count, 3))).astype(np.float32)
    else:  # many nodes: most of them far from any one block
        near = s["positions"][rng.choice(len(s["positions"]), node_count // 2, replace=False)] + rng.normal(0, 0.01, (node_count // 2, 3))
        far = rng.uniform(-1, 1, (node_count - node_count // 2, 3)) + [0, 0, 0.6]
        nodes = rng.permutation(np.concatenate([near, far])).astype(np.float32)
    rotations, translations = synthetic.smooth_node_motion(nodes, 3.0, 0.005, np.random.default_rng(1))
    G = nnrt.geometry
    method = G.WarpNodeCoverageComputationMethod
    outputs = []
    for full_scan in (False, True):
        wf = G.GraphWarpField(nodes, coverage, True, 4, min_valid, method.MINIMAL_K_NEIGHBOR_NODE_DISTANCE if variable else method.FIXED_NODE_COVERAGE)
        if variable:
            wf.node_coverage_weights = torch.as_tensor((rng.uniform(0.5, 1.5, len(nodes)) * coverage).astype(np.float32) ** 2).cuda() if not outputs else weights_used
            weights_used = wf.node_coverage_weights
        wf.set_node_rotations(rotations)
        wf.set_node_translations(translations)
        values = dict(tsdf=torch.as_tensor(c["tsdf"]).cuda(), weight=torch.as_tensor(c["weight"]).cuda(), color=torch.as_tensor(c["voxel_color"]).cuda())
        cos = G.functional.integrate_non_rigid_voxel_blocks(c["active"], c["keys"], values, c["R"], c["voxel_size"], wf, c["depth"], c["color"], c["normals"],
                                                            s["K"], s["K"], None, c["depth_scale"], 3.0, 2.0, _full_scan=full_scan)
        outputs.append((values["tsdf"].cpu().numpy(), values["color"].cpu().numpy(), cos.cpu().numpy()))
    (tsdf_a, color_a, cos_a), (tsdf_b, color_b, cos_b) = outputs
    assert (tsdf_b != c["tsdf"]).sum() > 500
    assert np.array_equal(tsdf_a, tsdf_b) and np.array_equal(color_a, color_b) and np.array_equal(cos_a, cos_b)
