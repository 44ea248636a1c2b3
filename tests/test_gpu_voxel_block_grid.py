"""nnrt.geometry.NonRigidSurfaceVoxelBlockGrid (cpp/geometry/NonRigidSurfaceVoxelBlockGrid.cpp) and its on-disk form
(cpp/io/VoxelBlockGridIO.cpp): block activation over the device block hash, the integrate -> extract -> fit loop of the fusion pipeline
 through the class, block selection against the oracle's pieces, and the file round trip."""
import gzip
import struct

import numpy as np
import pytest
import torch

from dynamicfuion_python_b200 import synthetic

pytestmark = pytest.mark.gpu


def _grid(nnrt, weight_dtype=torch.float32, with_color=True, voxel_size=0.004, resolution=8, capacity=64):
    names, dtypes, channels = ["tsdf", "weight"], [torch.float32, weight_dtype], [(1,), (1,)]
    if with_color:
        names, dtypes, channels = names + ["color"], dtypes + [torch.float32], channels + [(3,)]
    return nnrt.geometry.NonRigidSurfaceVoxelBlockGrid(names, dtypes, channels, voxel_size, resolution, capacity)


def test_block_hash_map_activation(nnrt_mirror):
    """HashMap::Activate / Find semantics on the append-only block map: a key is stored once, at the rank of its first appearance;
    masks mark the keys a call newly activated; the map grows past its capacity keeping the buffer indices of existing blocks."""
    grid = _grid(nnrt_mirror, capacity=16)
    table = grid.hashmap()
    rng = np.random.default_rng(3)
    stored = {}
    for batch in range(6):
        keys = rng.integers(-4, 5, (40, 3)).astype(np.int32)
        indices, masks = table.activate(keys)
        indices, masks = indices.cpu().numpy(), masks.cpu().numpy()
        for row, key in enumerate(map(tuple, keys)):
            if key not in stored:
                assert masks[row], (batch, row)
                stored[key] = len(stored)
            else:
                assert not masks[row], (batch, row)
            assert indices[row] == stored[key]
        assert table.size() == len(stored) and table.capacity() >= table.size()
        assert np.array_equal(table.key_tensor()[:table.size()].cpu().numpy(), np.array(list(stored), np.int32))
    queries = np.concatenate([np.array(list(stored), np.int32)[::3], np.array([[50, 0, 0], [0, -60, 1]], np.int32)])
    found, mask = table.find(queries)
    assert mask.cpu().numpy().tolist() == [True] * (len(queries) - 2) + [False, False]
    assert np.array_equal(found.cpu().numpy()[:-2], np.arange(len(stored))[::3])
    assert np.array_equal(table.active_buf_indices().cpu().numpy(), np.arange(len(stored)))


def test_inactive_neighbour_coordinates_order(nnrt_mirror):
    """BufferCoordinatesOfInactiveNeighborBlocks: (neighbour offset, block) order with repeats, offsets numbered dx fastest."""
    grid = _grid(nnrt_mirror)
    keys = np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0], [5, 5, 5]], np.int32)
    grid.hashmap().activate(keys)
    got = grid.buffer_coordinates_of_inactive_neighbor_blocks().cpu().numpy()
    active = set(map(tuple, keys))
    want = []
    for n in range(27):
        offset = np.array([n % 3 - 1, (n % 9) // 3 - 1, n // 9 - 1])
        for key in keys:
            if tuple(key + offset) not in active:
                want.append(key + offset)
    assert np.array_equal(got, np.array(want, np.int32))
    offered = grid.activate_sleeve_blocks()
    assert offered == len(want)
    assert grid.hashmap().size() == len(active | set(map(tuple, want)))


def _frame(oracle, s, H, W, rotations, translations, anchors, weights):
    """Depth (metres), ordered normals and colour of the warped relief mesh."""
    wp, _ = oracle.warp(s["positions"], s["normals"], s["nodes"], rotations, translations, anchors, weights)
    fv, mask = oracle.face_vertices_ndc(wp, s["triangles"], s["K"], H, W, 0.0, 10.0)
    _, depth, _, _, _ = oracle.rasterize(fv, mask, H, W, 0.0, 1, perspective_correct=True)
    depth = np.where(depth[..., 0] > 0, depth[..., 0], 0).astype(np.float32)
    points, _ = oracle.unproject_depth(depth, s["K"], None, 1.0, 10.0)
    # The reference's kernel skips a voxel when dot(direction to the camera, normal) > 0.5 (NonRigidSurfaceVoxelBlockGridImpl.h:192):
    # it integrates where the depth normals point AWAY from the camera; the ordered normals of this scene point towards it.
    normals = -oracle.compute_ordered_point_cloud_normals(points, H, W).reshape(H, W, 3)
    return depth, normals, points


def test_fusion_loop_through_the_grid_class(nnrt_mirror, oracle):
    """The loop of apps/fusion/pipeline.py on the device: blocks around the first depth frame are activated, the frame is integrated
    (identity warp), the canonical mesh is extracted, the surface moves, the blocks the moved surface reaches are found through the
    moved warp field and that frame is integrated THROUGH the field into the canonical volume. Checks: the class equals the functional
    entry points on the same tensors (those are tested against the oracle in test_gpu_tsdf / test_mesh_extraction), the extracted
    surface lies on the relief, and block selection equals the oracle's two pieces composed as NonRigidSurfaceVoxelBlockGrid.cpp:150-178
    composes them."""
    nnrt = nnrt_mirror
    G = nnrt.geometry
    H, W = 120, 160
    s = synthetic.make_scene("tinyr", H, W)
    node = s["positions"].mean(0, keepdims=True).astype(np.float32)
    s = dict(s, nodes=node)
    identity, zero = np.eye(3, dtype=np.float32)[None], np.zeros((1, 3), np.float32)
    anchors, weights = oracle.compute_anchors(s["positions"], node, 1, 0, 10.0)
    depth0, normals0, points0 = _frame(oracle, s, H, W, identity, zero, anchors, weights)
    color = np.random.default_rng(1).uniform(0, 1, (H, W, 3)).astype(np.float32)
    voxel_size, R = 0.004, 8
    grid = _grid(nnrt, torch.float32, True, voxel_size, R, 32)
    valid = depth0.reshape(-1) > 0
    seeds = np.unique(np.floor(points0[valid] / (voxel_size * R)).astype(np.int32), axis=0)
    wf = G.GraphWarpField(node, 10.0, False, 1, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    K = s["K"]
    for _ in range(4):
        cos = grid.integrate_non_rigid(seeds, wf, depth0, color, normals0, K, K, np.eye(4), 1.0, 3.0, 2.0)
    assert grid.extract_triangle_mesh(3.0).positions.shape[0] == 0  # no voxel has any weight yet (see below): an empty mesh, no error
    grid.activate_sleeve_blocks()
    for _ in range(2):
        grid.integrate_non_rigid(np.zeros((0, 3), np.int32), wf, depth0, color, normals0, K, K, np.eye(4), 1.0, 3.0, 2.0)
    size = grid.hashmap().size()
    assert size > len(seeds) and (cos != 0).sum().item() > 1000
    # The reference's non-rigid kernel reads the voxel weight and never writes it (NonRigidSurfaceVoxelBlockGridImpl.h:201-203; kept):
    # in its pipeline the weights come from the rigid Integrate of the first frame. The test plays that part for the voxels it touched.
    assert float(grid.attribute("weight").abs().max()) == 0.0
    # the class integrates into every active block: same as the functional call on its tensors
    check = _grid(nnrt, torch.float32, True, voxel_size, R, grid.hashmap().capacity())
    check.hashmap().activate(grid.hashmap().key_tensor()[:size])
    views = check._value_views()
    for _ in range(6):
        # blocks activated by the sleeve saw only the last two integrations
        n = len(seeds) if _ < 4 else size
        G.functional.integrate_non_rigid_voxel_blocks(np.arange(n), check.hashmap().key_tensor(), views, R, voxel_size, wf, depth0, color, normals0,
                                                      K, K, np.eye(4), 1.0, 3.0, 2.0)
    assert torch.equal(check.attribute("tsdf"), grid.attribute("tsdf")) and torch.equal(check.attribute("weight"), grid.attribute("weight"))
    assert torch.equal(check.attribute("color"), grid.attribute("color"))

    grid.attribute("weight")[grid.attribute("tsdf") != 0] = 5.0
    mesh = grid.extract_triangle_mesh(3.0)
    direct = G.functional.extract_triangle_mesh(grid.hashmap().key_tensor()[:size], {k: v[:size] for k, v in grid._value_views().items()}, voxel_size, 3.0)
    assert torch.equal(mesh.positions, direct.positions) and torch.equal(mesh.triangles, direct.triangles)
    vertices = mesh.positions.cpu().numpy()
    assert len(vertices) > 2000
    # the extracted surface is the relief: every vertex projects onto a depth pixel whose surface point is within a voxel
    u = np.clip(np.round(vertices[:, 0] / vertices[:, 2] * K[0, 0] + K[0, 2]).astype(int), 0, W - 1)
    v = np.clip(np.round(vertices[:, 1] / vertices[:, 2] * K[1, 1] + K[1, 2]).astype(int), 0, H - 1)
    seen = depth0[v, u] > 0
    assert seen.mean() > 0.95 and np.abs(vertices[seen, 2] - depth0[v, u][seen]).max() < 2.5 * voxel_size

    # the surface moves (the tracker's part is tested in test_fusion_loop_module); the field carries the motion into the volume's frame
    # (a generic direction: with no lateral component, box faces at x = 0 / y = 0 contain the rays of pixel column cx / row cy exactly,
    # and the slab test's 0 / 0 is decided differently by CUDA's and the host's min / max -- in the reference as well)
    shift = np.array([[0.0011, -0.0007, 0.006]], np.float32)
    depth1, normals1, points1 = _frame(oracle, s, H, W, identity, shift, anchors, weights)
    fitted = shift
    wf.set_node_translations(fitted)

    # blocks the moved surface reaches, through the fitted field: the reference's composition, piece by piece against the oracle
    want_coordinates = grid.buffer_coordinates_of_inactive_neighbor_blocks().cpu().numpy()
    boxes = oracle.bounding_boxes_of_warped_blocks(want_coordinates, node, identity, fitted, 10.0, 1, 0, voxel_size, R, np.eye(4))
    got_boxes = grid.get_bounding_boxes_of_warped_blocks(want_coordinates, wf, np.eye(4)).cpu().numpy()
    np.testing.assert_allclose(got_boxes, boxes, rtol=1e-5, atol=1e-6)
    # (the mask is decided on box faces: it is compared on the SAME boxes; the box kernel agrees with the oracle to rounding, above)
    mask = oracle.boxes_intersecting_surface_mask(got_boxes, K, depth1, 1.0, 3.0, 4, 8.0 * voxel_size)
    assert 0 < mask.sum() < len(mask), (int(mask.sum()), len(mask))
    found = grid.find_blocks_intersecting_truncation_region(depth1, wf, K, np.eye(4), 1.0, 3.0, 2.0)
    assert np.array_equal(found.cpu().numpy(), want_coordinates[mask])
    before = grid.attribute("tsdf").clone()
    grid.integrate_non_rigid(found, wf, depth1, color, normals1, K, K, np.eye(4), 1.0, 3.0, 2.0)
    assert grid.hashmap().size() == size + len(np.unique(want_coordinates[mask], axis=0))
    assert not torch.equal(before[:size], grid.attribute("tsdf")[:size])
    origins = grid.extract_voxel_block_coordinates().cpu().numpy()
    assert np.allclose(origins, grid.hashmap().key_tensor()[:grid.hashmap().size()].cpu().numpy() * (R * voxel_size))
    values = grid.extract_voxel_values_and_coordinates().cpu().numpy()
    assert values.shape == (grid.hashmap().size() * R ** 3, 3 + 1 + 1 + 3)
    first_key = grid.hashmap().key_tensor()[0].cpu().numpy()
    assert np.allclose(values[1, :3], (first_key * R + [1, 0, 0]) * voxel_size)  # x runs fastest
    assert np.array_equal(values[:R ** 3, 3], grid.attribute("tsdf")[0].reshape(-1).cpu().numpy())


@pytest.mark.parametrize("weight_dtype,with_color,compressed", [(torch.float32, True, True), (torch.uint16, False, False)])
def test_voxel_block_grid_files(nnrt_mirror, tmp_path, weight_dtype, with_color, compressed):
    """write_/read_(non_rigid_surface_)voxel_block_grid in the layout of cpp/io/VoxelBlockGridIO.cpp:26-60: f32 voxel size, i64 block
    resolution, i32 attribute count, (u32 length, name, i32 index) per attribute, device (i32, i32), i64 capacity, then the ACTIVE keys
    and value rows as tensors. Header fields are parsed by hand; the round trip restores keys and values bit for bit."""
    nnrt = nnrt_mirror
    grid = _grid(nnrt, weight_dtype, with_color, 0.005, 4, 40)
    rng = np.random.default_rng(11)
    keys = np.unique(rng.integers(-6, 7, (25, 3)).astype(np.int32), axis=0)
    grid.hashmap().activate(keys)
    size = grid.hashmap().size()
    grid.attribute("tsdf")[:size] = torch.as_tensor(rng.uniform(-1, 1, (size, 4, 4, 4, 1)).astype(np.float32)).cuda()
    weights = rng.integers(0, 50, (size, 4, 4, 4, 1))
    if weight_dtype == torch.uint16:
        grid.attribute("weight").view(torch.int16)[:size] = torch.as_tensor(weights.astype(np.int16)).cuda()
    else:
        grid.attribute("weight")[:size] = torch.as_tensor(weights.astype(np.float32)).cuda()
    if with_color:
        grid.attribute("color")[:size] = torch.as_tensor(rng.uniform(0, 255, (size, 4, 4, 4, 3)).astype(np.float32)).cuda()
    path = str(tmp_path / "volume.dat")
    nnrt.io.write_non_rigid_surface_voxel_block_grid(path, grid, compressed)
    raw = open(path, "rb").read()
    assert (raw[:2] == b"\x1f\x8b") == compressed
    if compressed:
        raw = gzip.decompress(raw)
    voxel_size, resolution, count = struct.unpack_from("<fqi", raw, 0)
    assert (np.float32(voxel_size), resolution, count) == (np.float32(0.005), 4, 3 if with_color else 2)
    offset, names = 16, []
    for index in range(count):
        (length,) = struct.unpack_from("<I", raw, offset)
        names.append(raw[offset + 4:offset + 4 + length].decode())
        assert struct.unpack_from("<i", raw, offset + 4 + length) == (index,)
        offset += 8 + length
    assert names == ["tsdf", "weight"] + (["color"] if with_color else [])
    device_type, device_id, capacity = struct.unpack_from("<iiq", raw, offset)
    assert (device_type, device_id, capacity) == (1, torch.cuda.current_device(), 40)
    assert struct.unpack_from("<Iqq", raw, offset + 16) == (2, size, 3)  # the key tensor's shape record: [active blocks, 3]
    for reader, cls in ((nnrt.io.read_non_rigid_surface_voxel_block_grid, nnrt.geometry.NonRigidSurfaceVoxelBlockGrid),
                        (nnrt.io.read_voxel_block_grid, nnrt.geometry.VoxelBlockGrid)):
        back = reader(path, compressed)
        assert type(back) is cls and back.attribute_names() == names
        assert back.voxel_size == pytest.approx(0.005) and back.block_resolution == 4 and back.hashmap().capacity() == 40
        assert back.hashmap().size() == size
        assert torch.equal(back.hashmap().key_tensor()[:size], grid.hashmap().key_tensor()[:size])
        for name in names:
            a, b = back.attribute(name)[:size], grid.attribute(name)[:size]
            assert a.dtype == b.dtype and a.shape == b.shape
            assert torch.equal(a.view(torch.int16) if a.dtype == torch.uint16 else a, b.view(torch.int16) if b.dtype == torch.uint16 else b)
    with pytest.raises(RuntimeError):
        nnrt.io.read_voxel_block_grid(str(tmp_path / "missing.dat"))
    with pytest.raises(RuntimeError):
        nnrt.io.write_voxel_block_grid(path, "not a grid")


@pytest.mark.parametrize("depth_u16,with_color", [(False, True), (True, False)])
def test_rigid_integration_matches_oracle(nnrt_mirror, oracle, depth_u16, with_color):
    """VoxelBlockGrid.integrate (Open3D's rigid integration, restated; csrc/tsdf.cu::integrate_rigid_kernel) against the oracle's
    restatement of the same published kernel: two frames in a row, weights counted per voxel, values to fp32 rounding; and
    compute_unique_block_coordinates against a numpy statement of Open3D's DepthTouch sampling (same block set)."""
    nnrt = nnrt_mirror
    H, W = 120, 160
    s = synthetic.make_scene("tinyr", H, W)
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], H, W, 0.0, 10.0)
    _, depth, _, _, _ = oracle.rasterize(fv, mask, H, W, 0.0, 1, perspective_correct=True)
    depth = np.where(depth[..., 0] > 0, depth[..., 0], 0).astype(np.float32)
    scale = 1.0
    if depth_u16:
        depth, scale = np.round(depth * 1000.0).astype(np.uint16), 1000.0
    rng = np.random.default_rng(4)
    color = rng.uniform(0, 1, (H, W, 3)).astype(np.float32) if with_color else None
    ext = np.eye(4)
    ext[:3, 3] = [0.002, -0.001, 0.003]
    voxel_size, R, multiplier = 0.004, 8, 3.0
    grid = _grid(nnrt, torch.uint16 if depth_u16 else torch.float32, with_color, voxel_size, R, 16)
    blocks = grid.compute_unique_block_coordinates(depth, s["K"], ext, scale, 3.0, multiplier)
    # DepthTouch in numpy: every 4th pixel, 4 samples along the ray over [d - truncation, d + truncation]
    d = depth[::4, ::4].astype(np.float32) / np.float32(scale)
    v, u = np.meshgrid(np.arange(d.shape[0], dtype=np.float32) * 4, np.arange(d.shape[1], dtype=np.float32) * 4, indexing="ij")
    valid = (d > 0) & (d < 3.0)
    K = s["K"].astype(np.float32)
    rays = np.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], np.ones_like(u)], -1)[valid].astype(np.float32)
    pose = np.linalg.inv(ext)
    directions = rays @ pose[:3, :3].T.astype(np.float32)
    t_min, t_max = np.maximum(d[valid] - voxel_size * multiplier, 0), np.minimum(d[valid] + voxel_size * multiplier, 3.0)
    t = t_min[:, None] + np.arange(4, dtype=np.float32)[None] * ((t_max - t_min) / 3.0)[:, None]
    points = pose[:3, 3].astype(np.float32) + t[..., None] * directions[:, None, :]
    want_blocks = np.unique(np.floor(points.astype(np.float32) / np.float32(voxel_size * R)).astype(np.int32).reshape(-1, 3), axis=0)
    got_blocks = blocks.cpu().numpy()
    assert len(np.unique(got_blocks, axis=0)) == len(got_blocks)
    got_set, want_set = set(map(tuple, got_blocks)), set(map(tuple, want_blocks))
    # a sample within an ulp of a block face may fall either way -- and with depth in whole millimetres, a 12 mm truncation and 32 mm
    # blocks, samples land exactly on block faces: a few blocks of the set depend on the last bit
    assert len(got_set ^ want_set) <= 0.05 * len(want_set) and len(want_set) > 100

    keys = np.zeros((grid.hashmap().capacity() + len(got_blocks), 3), np.int32)
    keys[:len(got_blocks)] = got_blocks
    tsdf_o = np.zeros((len(keys), R, R, R), np.float32)
    weight_o = np.zeros((len(keys), R, R, R), np.uint16 if depth_u16 else np.float32)
    color_o = np.zeros((len(keys), R, R, R, 3), np.float32) if with_color else None
    for frame in range(2):
        frame_depth = depth if frame == 0 else (depth + (2 if depth_u16 else 0.002)).astype(depth.dtype) * (depth > 0)
        if with_color:
            grid.integrate(blocks, frame_depth, color, s["K"], s["K"], ext, scale, 3.0, multiplier)
        else:
            grid.integrate(blocks, frame_depth, s["K"], ext, scale, 3.0, multiplier)
        oracle.integrate_rigid(np.arange(len(got_blocks)), keys, tsdf_o, weight_o, color_o, R, voxel_size, voxel_size * multiplier, frame_depth, color,
                               s["K"], s["K"], ext, scale, 3.0)
    size = grid.hashmap().size()
    assert size == len(got_blocks)
    weight_g = grid.attribute("weight")[:size]
    weight_g = (weight_g.view(torch.int16) if depth_u16 else weight_g).cpu().numpy().reshape(size, R, R, R)
    assert np.array_equal(weight_g.astype(np.float32), weight_o[:size].astype(np.float32)) and weight_o.max() == 2 and (weight_o[:size] == 2).sum() > 5000
    tsdf_g = grid.attribute("tsdf")[:size].cpu().numpy().reshape(size, R, R, R)
    assert np.abs(tsdf_g - tsdf_o[:size]).max() < 1e-5 and np.abs(tsdf_o).max() > 0.9
    if with_color:
        color_g = grid.attribute("color")[:size].cpu().numpy()
        assert np.abs(color_g - color_o[:size]).max() < 1e-3 and color_o.max() > 200
    mesh = grid.extract_triangle_mesh(1.5)
    assert mesh.positions.shape[0] > 2000


def test_fusion_loop_module(nnrt_mirror, oracle):
    """dynamicfuion_python_b200.fusion.FusionLoop: the frame loop of apps/fusion/pipeline.py with the dense-depth fitter as the tracker.
    A relief that recedes and bends a little more every frame; each tracked frame must leave the warped canonical mesh on the observed
    surface (rendered depth of the warped mesh against the input depth) and the fitter healthy."""
    from dynamicfuion_python_b200.fusion import FusionLoop
    nnrt = nnrt_mirror
    H, W = 120, 160
    s = synthetic.make_scene("tinyr", H, W)
    anchors, weights = oracle.compute_anchors(s["positions"], s["nodes"], 4, 0, s["node_coverage"])
    rotations, translations = synthetic.smooth_node_motion(s["nodes"], 1.0, 0.002, np.random.default_rng(3))
    identity = np.broadcast_to(np.eye(3, dtype=np.float32), rotations.shape).copy()

    def frame(k):
        a = k / 3.0  # motion grows linearly over the frames, plus 1.5 mm of recession per frame
        R = (identity + a * (rotations - identity)).astype(np.float32)
        t = (a * translations + np.array([0, 0, 0.0015 * k], np.float32)).astype(np.float32)
        wp, _ = oracle.warp(s["positions"], s["normals"], s["nodes"], R, t, anchors, weights)
        fv, mask = oracle.face_vertices_ndc(wp, s["triangles"], s["K"], H, W, 0.0, 10.0)
        _, depth, _, _, _ = oracle.rasterize(fv, mask, H, W, 0.0, 1, perspective_correct=True)
        return np.where(depth[..., 0] > 0, depth[..., 0], 0).astype(np.float32)

    loop = FusionLoop(s["K"], voxel_size=0.004, block_resolution=8, block_count=512, node_spacing=0.03, depth_scale=1.0, truncation_voxel_multiplier=3.0,
                      nnrt=nnrt)  # default tracker settings: coverage = spacing / 4, arap_term_weight 20, lambda 1 (the contractive ones)
    K = s["K"]

    def surface_error(positions, depth):
        """Mean |vertex z - depth at the vertex's pinhole pixel| and the share of vertices that land on the observed surface."""
        v = positions.cpu().numpy()
        col, row = np.round(v[:, 0] / v[:, 2] * K[0, 0] + K[0, 2]).astype(int), np.round(v[:, 1] / v[:, 2] * K[1, 1] + K[1, 2]).astype(int)
        inside = (col >= 0) & (col < W) & (row >= 0) & (row < H)
        inside[inside] &= depth[row[inside], col[inside]] > 0
        return float(np.abs(v[inside, 2] - depth[row[inside], col[inside]]).mean()), float(inside.mean())

    canonical = loop.first_frame(frame(0))
    assert canonical.positions.shape[0] > 2000 and loop.graph.nodes.shape[0] >= 30
    error, share = surface_error(canonical.positions, frame(0))
    assert error < 0.0003 and share > 0.99  # the canonical mesh is the first frame's surface to a fraction of a voxel
    untracked, _ = surface_error(canonical.positions, frame(3))
    assert untracked > 0.003
    for k in range(1, 4):
        depth = frame(k)
        canonical, warped, cosines = loop.next_frame(depth)
        assert canonical.positions.shape[0] > 2000 and torch.isfinite(warped.positions).all() and (cosines != 0).sum().item() > 1000
        error, share = surface_error(warped.positions, depth)
        # tracked: the warped canonical mesh stays on the observed surface (0.30 / 0.37 / 0.44 mm measured; untracked it would be 1.2 / 2.3 / 3.4 mm away)
        assert error < 0.0008 and share > 0.97, (k, error, share)
    assert loop.frame_index == 4
    assert abs(float(loop.graph.node_translations[:, 2].mean()) - 0.0045) < 0.0015  # 1.5 mm of recession per frame, recovered by the nodes
