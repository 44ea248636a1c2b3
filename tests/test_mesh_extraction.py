"""Triangle-mesh extraction from TSDF voxel blocks: the marching-cubes tables derived by
tools/generate_marching_cubes_tables.py, the oracle restatement of Open3D's ExtractTriangleMesh, and the CUDA path
(nnrt.geometry.functional.extract_triangle_mesh / voxel_block_neighbors) against the oracle, bit for bit. Parity is unpinned by the
reference (its extraction runs through Open3D and downloaded volumes); what can be checked independently is geometry and topology."""
import itertools

import numpy as np
import pytest
import torch

R = 8


def _blocks(keys, field, voxel_size, weight_value=10.0):
    """TSDF blocks [B,R,R,R] (z, y, x major to minor) sampling `field` (signed distance in metres / truncation) at the voxel centres."""
    keys = np.asarray(keys, np.int32)
    z, y, x = np.meshgrid(np.arange(R), np.arange(R), np.arange(R), indexing="ij")
    tsdf = np.empty((len(keys), R, R, R), np.float32)
    for b, k in enumerate(keys):
        p = np.stack([(k[0] * R + x) * voxel_size, (k[1] * R + y) * voxel_size, (k[2] * R + z) * voxel_size], -1)
        tsdf[b] = field(p)
    return keys, tsdf, np.full(tsdf.shape, weight_value, np.float32)


def _cube_of_blocks(lo, hi):
    return [k for k in itertools.product(range(lo, hi), repeat=3)]


def _sphere(center, radius, truncation):
    return lambda p: np.clip((np.linalg.norm(p - center, axis=-1) - radius) / truncation, -1, 1).astype(np.float32)


def _edge_use(triangles):
    """directed edge -> count; a closed, consistently oriented surface uses every directed edge once and its reverse once"""
    e = np.concatenate([triangles[:, [0, 1]], triangles[:, [1, 2]], triangles[:, [2, 0]]])
    keys, counts = np.unique(e, axis=0, return_counts=True)
    return {tuple(k): c for k, c in zip(keys, counts)}


def test_derived_tables_are_consistent():
    """Every configuration: triangles use exactly the sign-change edges, the complement configuration uses the same edges, and each
    triangle fan closes (every edge of the cube's polygon set is used once per direction or lies on the polygon boundary...)."""
    import re
    text = open(__file__.replace("tests/test_mesh_extraction.py", "dynamicfuion_python_b200/csrc/marching_cubes_tables.h")).read()
    edge_table = [int(v, 16) for v in re.findall(r"0x[0-9a-f]{3}", text.split("kMcEdgeTable[256]")[1].split("};")[0])]
    rows = re.findall(r"\{([-\d, ]+)\},", text.split("kMcTriangleTable[256][16]")[1])
    tri_table = [[int(v) for v in row.split(",")] for row in rows]
    assert len(edge_table) == 256 and len(tri_table) == 256
    assert edge_table[:8] == [0x000, 0x109, 0x203, 0x30a, 0x406, 0x50f, 0x605, 0x70c]  # the published edge table starts like this
    corners = [(0, 1), (1, 2), (2, 3), (3, 0), (4, 5), (5, 6), (6, 7), (7, 4), (0, 4), (1, 5), (2, 6), (3, 7)]
    for c in range(256):
        crossing = sum(1 << e for e, (a, b) in enumerate(corners) if ((c >> a) & 1) != ((c >> b) & 1))
        assert edge_table[c] == crossing == edge_table[255 - c]
        used = [v for v in tri_table[c] if v >= 0]
        assert len(used) % 3 == 0 and set(used) == {e for e in range(12) if crossing >> e & 1}
        assert all(v == -1 for v in tri_table[c][len(used):])


def test_oracle_sphere_is_closed_oriented_and_on_the_surface(oracle):
    voxel_size, radius, center = 0.01, 0.105, np.array([0.013, -0.004, 0.021])
    keys, tsdf, weight = _blocks(_cube_of_blocks(-2, 2), _sphere(center, radius, 0.04), voxel_size)
    vertices, normals, _, triangles = oracle.extract_triangle_mesh(keys, tsdf, weight, None, voxel_size, 3.0)
    assert len(vertices) > 2000 and len(triangles) > 4000
    # geometry: linear interpolation of a distance field puts the vertices on the sphere
    assert np.abs(np.linalg.norm(vertices - center, axis=1) - radius).max() < 0.05 * voxel_size
    # topology: closed (Euler characteristic 2) and consistently oriented
    use = _edge_use(triangles)
    assert all(c == 1 for c in use.values()) and all((b, a) in use for a, b in use)
    assert len(vertices) - len(use) // 2 + len(triangles) == 2
    # orientation: face normals and the interpolated TSDF gradients point outwards
    p0, p1, p2 = (vertices[triangles[:, k]] for k in range(3))
    face_normals = np.cross(p1 - p0, p2 - p0)
    assert (np.einsum("ij,ij->i", face_normals, (p0 + p1 + p2) / 3 - center) > 0).all()
    assert (np.einsum("ij,ij->i", normals, vertices - center) > 0.99 * np.linalg.norm(vertices - center, axis=1)).all()
    np.testing.assert_allclose(np.linalg.norm(normals, axis=1), 1.0, atol=2e-3)  # n / (|n| + 1e-5)


def test_oracle_weights_missing_blocks_and_neighbours(oracle):
    voxel_size = 0.01
    keys, tsdf, weight = _blocks(_cube_of_blocks(-2, 2), _sphere(np.zeros(3), 0.1, 0.04), voxel_size)
    full = oracle.extract_triangle_mesh(keys, tsdf, weight, None, voxel_size, 3.0)
    # light voxels (weight <= threshold) open the surface there: fewer vertices, nothing new
    light = weight.copy()
    light[tsdf.shape[0] // 2:] = 3.0
    part = oracle.extract_triangle_mesh(keys, tsdf, light, None, voxel_size, 3.0)
    assert 0 < len(part[0]) < len(full[0])
    assert {tuple(v) for v in part[0]} <= {tuple(v) for v in full[0]}
    # dropping blocks does the same; the neighbour table says which blocks exist
    keep = np.array([k[0] < 1 for k in keys])
    fewer = oracle.extract_triangle_mesh(keys[keep], tsdf[keep], weight[keep], None, voxel_size, 3.0)
    assert 0 < len(fewer[0]) < len(full[0]) and {tuple(v) for v in fewer[0]} <= {tuple(v) for v in full[0]}
    nb = oracle.voxel_block_neighbors(keys)
    lookup = {tuple(k): i for i, k in enumerate(keys)}
    for b in (0, 21, 63):
        for slot in range(27):
            d = (slot % 3 - 1, (slot // 3) % 3 - 1, slot // 9 - 1)
            assert nb[b, slot] == lookup.get(tuple(keys[b] + d), -1)
    assert nb[:, 13].tolist() == list(range(len(keys)))


def _np(t):
    return t.detach().cpu().numpy()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["sphere", "two_spheres_uint16_color", "holes"])
def test_gpu_extraction_equals_oracle_bit_for_bit(nnrt_mirror, oracle, case):
    """Vertices, normals, colours and triangles in the same (block, voxel, edge) order as the oracle, every bit equal."""
    f = nnrt_mirror.geometry.functional
    voxel_size, rng = 0.0075, np.random.default_rng(8)
    if case == "sphere":
        keys, tsdf, weight = _blocks(_cube_of_blocks(-2, 2), _sphere(np.array([0.004, 0.002, -0.003]), 0.081, 0.03), voxel_size)
        color = None
    elif case == "two_spheres_uint16_color":
        a, b = _sphere(np.array([-0.05, 0.0, 0.0]), 0.04, 0.03), _sphere(np.array([0.06, 0.01, 0.02]), 0.045, 0.03)
        keys, tsdf, weight = _blocks(rng.permutation(_cube_of_blocks(-2, 2)), lambda p: np.minimum(a(p), b(p)), voxel_size)
        color = rng.uniform(0, 255, tsdf.shape + (3,)).astype(np.float32)
    else:
        keys, tsdf, weight = _blocks(_cube_of_blocks(-2, 2), _sphere(np.zeros(3), 0.085, 0.03), voxel_size)
        keep = rng.random(len(keys)) > 0.3
        keys, tsdf, weight = keys[keep], tsdf[keep], weight[keep]
        weight[rng.random(weight.shape) < 0.05] = 1.0  # below the threshold
        color = None
    values = {"tsdf": tsdf, "weight": weight.astype(np.uint16).view(np.int16) if "uint16" in case else weight, "color": color}
    if "uint16" in case:
        values["weight"] = torch.from_numpy(weight.astype(np.uint16).view(np.int16)).view(torch.uint16)
    assert np.array_equal(_np(f.voxel_block_neighbors(keys)), oracle.voxel_block_neighbors(keys))
    mesh = f.extract_triangle_mesh(keys, values, voxel_size, 3.0)
    vertices, normals, colors, triangles = oracle.extract_triangle_mesh(keys, tsdf, weight, color, voxel_size, 3.0)
    assert len(vertices) > 500
    assert np.array_equal(_np(mesh.positions), vertices) and np.array_equal(_np(mesh.triangles), triangles)
    assert np.array_equal(_np(mesh.normals), normals)
    if color is not None:
        assert np.array_equal(_np(mesh.colors), colors)
    if case != "holes":
        use = _edge_use(triangles)
        assert all(c == 1 for c in use.values()) and all((b, a) in use for a, b in use)


@pytest.mark.gpu
def test_gpu_extract_then_fit_stays_on_the_device(nnrt_mirror):
    """integrate -> extract -> fit without leaving the device: the extracted mesh (positions, TSDF-gradient normals, triangles) goes
    straight into the fitter as its canonical mesh."""
    nnrt = nnrt_mirror
    f = nnrt.geometry.functional
    voxel_size = 0.01
    # a bump facing the camera at z ~ 1.2
    center = np.array([0.0, 0.0, 1.45])
    keys, tsdf, weight = _blocks([k for k in itertools.product(range(-3, 3), range(-3, 3), range(13, 17))], _sphere(center, 0.22, 0.04), voxel_size)
    mesh = f.extract_triangle_mesh(keys, {"tsdf": tsdf, "weight": weight}, voxel_size, 3.0)
    assert mesh.positions.is_cuda and mesh.positions.shape[0] > 2000
    G, A = nnrt.geometry, nnrt.alignment
    # one node carrying the whole cap, as in the reference's one-node fitter tests (translation only, lambda = 0.001)
    nodes = torch.tensor([[0.0, 0.0, 1.25]], device="cuda")
    wf = G.HierarchicalGraphWarpField(nodes, 1.0, False, 1, 0, G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 1)
    K = np.array([[300.0, 0, 160], [0, 300.0, 120], [0, 0, 1]])
    # target: the same surface moved 1 cm towards the camera, rendered by the CUDA rasteriser
    moved = (mesh.positions - torch.tensor([0.0, 0.0, 0.01], device="cuda"), None, mesh.triangles)
    fv, mask = nnrt.rendering.functional.get_mesh_ndc_face_vertices_and_clip_mask(moved, K, (240, 320), 0.0, 10.0)
    _, depth, _, _ = nnrt.rendering.rasterize_ndc_triangles(fv, mask, (240, 320), 0.0, 1, -1, -1, True, False, True)
    assert (depth[..., 0] > 0).float().mean() > 0.05  # the extracted triangles face the camera for the rasteriser's culling rule
    fitter = A.DeformableMeshToImageFitter(5, [A.IterationMode.TRANSLATION_ONLY], 1e-6, True, 10.0, False, 0.01, 0.001)
    fitter.fit_to_image(wf, mesh, None, depth[..., 0].contiguous(), None, K, None, 1.0)
    fitter.check_status()
    t = _np(wf.get_node_translations())
    np.testing.assert_allclose(t[0], [0.0, 0.0, -0.01], atol=5e-4)


@pytest.mark.gpu
def test_pybind_module_extraction_equals_the_mirror(nnrt_mirror):
    from dynamicfuion_python_b200 import build
    module = build.import_pybind_module()
    voxel_size, rng = 0.0075, np.random.default_rng(3)
    keys, tsdf, weight = _blocks(_cube_of_blocks(-2, 2), _sphere(np.array([0.004, 0.002, -0.003]), 0.081, 0.03), voxel_size)
    values = {"tsdf": tsdf, "weight": weight, "color": rng.uniform(0, 255, tsdf.shape + (3,)).astype(np.float32)}
    a = nnrt_mirror.geometry.functional.extract_triangle_mesh(keys, values, voxel_size, 3.0)
    b = module.geometry.functional.extract_triangle_mesh(keys, values, voxel_size, 3.0)
    assert torch.equal(a.positions, b.vertex["positions"]) and torch.equal(a.normals, b.vertex["normals"])
    assert torch.equal(a.triangles, b.triangle["indices"]) and torch.equal(a.colors, b.vertex["colors"])
    assert torch.equal(nnrt_mirror.geometry.functional.voxel_block_neighbors(keys), module.geometry.functional.voxel_block_neighbors(keys))
