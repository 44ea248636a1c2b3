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
    (identity warp) until the weights pass the extraction threshold, the canonical mesh is extracted, a one-node warp field is fitted to
    a frame of the translated surface, that frame is integrated THROUGH the fitted field into the canonical volume, and the blocks the
    moved surface reaches are found. Checks: the class equals the functional entry points on the same tensors (those are tested against
    the oracle in test_gpu_tsdf / test_mesh_extraction), the extracted surface lies on the relief, the fit recovers the translation,
    and block selection equals the oracle's two pieces composed as NonRigidSurfaceVoxelBlockGrid.cpp:150-178 composes them."""
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

    # the surface moves; fit the one-node field to the new frame with the extracted canonical mesh, translation only
    shift = np.array([[0.0, 0.0, 0.006]], np.float32)  # along the view axis: what a point-to-plane term on a shallow relief determines well
    depth1, normals1, points1 = _frame(oracle, s, H, W, identity, shift, anchors, weights)
    fitter = nnrt.alignment.DeformableMeshToImageFitter(max_iteration_count=8, preconditioning_dampening_factor=0.0, arap_term_weight=0.0,
                                                        iteration_mode_sequence=[nnrt.alignment.IterationMode.TRANSLATION_ONLY])
    field = G.HierarchicalGraphWarpField.from_graph(node, None, 1, 10.0, False, 1, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
    fitter.fit_to_image(field, mesh, None, depth1, None, K, np.eye(4), 1.0)
    fitter.check_status()
    fitted = field.node_translations.cpu().numpy()
    # (sliding the shallow relief sideways changes the point-to-plane residuals little -- the aperture direction of this scene -- so
    # only the axial component is held to the voxel scale)
    assert abs(fitted[0, 2] - shift[0, 2]) < 1e-3 and np.abs(fitted[0, :2]).max() < 2e-2, fitted
    wf.set_node_translations(fitted)

    # blocks the moved surface reaches, through the fitted field: the reference's composition, piece by piece against the oracle
    want_coordinates = grid.buffer_coordinates_of_inactive_neighbor_blocks().cpu().numpy()
    boxes = oracle.bounding_boxes_of_warped_blocks(want_coordinates, node, identity, fitted, 10.0, 1, 0, voxel_size, R, np.eye(4))
    mask = oracle.boxes_intersecting_surface_mask(boxes, K, depth1, 1.0, 3.0, 4, 8.0 * voxel_size)
    found = grid.find_blocks_intersecting_truncation_region(depth1, wf, K, np.eye(4), 1.0, 3.0, 2.0)
    assert np.array_equal(found.cpu().numpy(), want_coordinates[mask]) and 0 < mask.sum() < len(mask)
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
