"""Deterministic synthetic scenes for the dense-depth fitter (SURVEY.md §8(d) C1/C2/C3 shapes).

Pure numpy input generation: meshes, two-layer deformation graphs, ground-truth node motion.
Rendering a target depth image is left to the caller (tests use the CPU oracle, bench.py uses the
CUDA rasteriser) through the `rasterize` callable so this module depends on neither.
"""
import numpy as np

SEED = 20240711


def default_intrinsics(height=480, width=640):
    """K of cpp/tests/test_rasterize.cpp:176-180 scaled to the image size (580 px focal at 640x480)."""
    s = width / 640.0
    return np.array([[580.0 * s, 0, width / 2.0], [0, 580.0 * s, height / 2.0], [0, 0, 1]], np.float64)


def _front_facing(positions, triangles, K, height, width):
    """Flip triangle winding where needed so faces whose normal points to the camera are front (CW in the
    reference's NDC, cpp/rendering/functional/kernel/BarycentricCoordinates.h:36-48)."""
    m = min(height, width)
    fx, fy = 2 * K[0, 0] / m, -2 * K[1, 1] / m
    p = positions[triangles]  # F,3,3
    x = fx * p[..., 0] / p[..., 2]
    y = fy * p[..., 1] / p[..., 2]
    area = (x[:, 0] - x[:, 1]) * (y[:, 1] - y[:, 2]) - (y[:, 0] - y[:, 1]) * (x[:, 1] - x[:, 2])
    return area


def plane_mesh(n=33, side=1.0, z=1.2, jitter=1e-4, rng=None):
    rng = rng or np.random.default_rng(SEED)
    lin = np.linspace(-side / 2, side / 2, n, dtype=np.float64)
    xx, yy = np.meshgrid(lin, lin)
    pos = np.stack([xx.ravel(), yy.ravel(), np.full(n * n, z)], 1)
    pos[:, 2] += rng.uniform(-jitter, jitter, n * n)
    idx = np.arange(n * n).reshape(n, n)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel()
    tris = np.concatenate([np.stack([a, b, c], 1), np.stack([b, d, c], 1)], 0).astype(np.int64)
    nrm = np.tile(np.array([0, 0, -1.0]), (n * n, 1))
    return pos.astype(np.float32), nrm.astype(np.float32), tris


def relief_mesh(n=203, side=0.3, z=0.6, amplitude=0.02, rng=None):
    """Camera-facing bumpy sheet ("bunny-scale relief"): V = n*n, F = 2*(n-1)^2 (n=203: 41209 / 81608).
    Every degree of freedom of a node is observable through point-to-plane residuals and there are no grazing
    silhouettes inside the sheet, so the reference's Gauss-Newton iteration is well conditioned on it."""
    rng = rng or np.random.default_rng(SEED)
    lin = np.linspace(-side / 2, side / 2, n, dtype=np.float64)
    xx, yy = np.meshgrid(lin, lin)
    h = np.zeros_like(xx)
    hx = np.zeros_like(xx)
    hy = np.zeros_like(xx)
    for octave in range(3):
        for _ in range(4):
            k = rng.normal(size=2) * (2.0 ** octave) * 18.0
            ph = rng.uniform(0, 2 * np.pi)
            a = amplitude * (0.5 ** octave) / 4.0
            arg = k[0] * xx + k[1] * yy + ph
            h += a * np.sin(arg)
            hx += a * k[0] * np.cos(arg)
            hy += a * k[1] * np.cos(arg)
    pos = np.stack([xx.ravel(), yy.ravel(), (z + h).ravel()], 1)
    nrm = np.stack([hx.ravel(), hy.ravel(), -np.ones(n * n)], 1)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    idx = np.arange(n * n).reshape(n, n)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, :-1].ravel(), idx[1:, 1:].ravel()
    tris = np.concatenate([np.stack([a, b, c], 1), np.stack([b, d, c], 1)], 0).astype(np.int64)
    return pos.astype(np.float32), nrm.astype(np.float32), tris


def icosphere(subdivisions):
    t = (1 + 5 ** 0.5) / 2
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
                  [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                  [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(subdivisions):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]], 0)
        e.sort(axis=1)
        ue, inv = np.unique(e, axis=0, return_inverse=True)
        mid = v[ue[:, 0]] + v[ue[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        nv = len(v)
        v = np.concatenate([v, mid], 0)
        F = len(f)
        m01, m12, m20 = nv + inv[:F], nv + inv[F:2 * F], nv + inv[2 * F:]
        f = np.concatenate([np.stack([f[:, 0], m01, m20], 1), np.stack([f[:, 1], m12, m01], 1),
                            np.stack([f[:, 2], m20, m12], 1), np.stack([m01, m12, m20], 1)], 0)
    return v, f


def bumpy_sphere_mesh(subdivisions=6, radius=0.15, center=(0.0, 0.0, 0.6), amplitude=0.1, rng=None):
    """Genus-0 displaced icosphere: V = 10*4^s + 2, F = 20*4^s (s=6: 40962 / 81920)."""
    rng = rng or np.random.default_rng(SEED)
    v, f = icosphere(subdivisions)
    disp = np.zeros(len(v))
    for octave in range(3):
        k = rng.normal(size=(6, 3)) * (2.0 ** octave) * 2.0
        ph = rng.uniform(0, 2 * np.pi, 6)
        disp += (0.5 ** octave) * np.sin(v @ k.T + ph).sum(1) / 6.0
    r = radius * (1 + amplitude * disp)
    pos = v * r[:, None] + np.asarray(center)
    # area-weighted vertex normals
    fn = np.cross(pos[f[:, 1]] - pos[f[:, 0]], pos[f[:, 2]] - pos[f[:, 0]])
    nrm = np.zeros_like(pos)
    for k in range(3):
        np.add.at(nrm, f[:, k], fn)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True) + 1e-20
    if np.mean(np.sum(nrm * (pos - np.asarray(center)), 1)) < 0:
        nrm = -nrm
    return pos.astype(np.float32), nrm.astype(np.float32), f.astype(np.int64)


def orient_front_faces(pos, nrm, tris, K, height, width):
    """Make outward-normal faces front-facing under the reference's clockwise-in-NDC rule."""
    area = _front_facing(pos.astype(np.float64), tris, K, height, width)
    fn = nrm[tris].mean(1)
    toward_camera = np.sum(fn * pos[tris].mean(1), 1) < 0
    flip = (area > 0) != toward_camera
    tris = tris.copy()
    tris[flip] = tris[flip][:, [0, 2, 1]]
    return tris


def _farthest_point_subsample(points, count, rng):
    idx = [int(rng.integers(len(points)))]
    d = np.linalg.norm(points - points[idx[0]], axis=1)
    for _ in range(count - 1):
        i = int(np.argmax(d))
        idx.append(i)
        d = np.minimum(d, np.linalg.norm(points - points[i], axis=1))
    return np.array(idx)


def two_layer_graph(points, fine_count, coarse_count, degree=4, rng=None):
    """Nodes in virtual order (fine layer first, then coarse); every fine node gets `degree` edges to its
    nearest coarse nodes: the regime in which the reference's ARAP Hessian is defined."""
    rng = rng or np.random.default_rng(SEED)
    fine = points[_farthest_point_subsample(points, fine_count, rng)]
    if coarse_count <= 0:
        return fine.astype(np.float32), np.zeros((0, 2), np.int32), fine_count
    coarse = fine[_farthest_point_subsample(fine, coarse_count, rng)] + rng.normal(scale=1e-3, size=(coarse_count, 3))
    d = np.linalg.norm(fine[:, None] - coarse[None], axis=2)
    nn = np.argsort(d, axis=1, kind="stable")[:, :min(degree, coarse_count)]
    edges = np.stack([np.repeat(np.arange(fine_count), nn.shape[1]), fine_count + nn.ravel()], 1).astype(np.int32)
    nodes = np.concatenate([fine, coarse], 0).astype(np.float32)
    return nodes, edges, fine_count


def smooth_node_motion(nodes, max_angle_deg=5.0, max_translation=0.01, rng=None):
    rng = rng or np.random.default_rng(SEED + 1)
    c = nodes.mean(0)
    k = rng.normal(size=(3, 3)) * 3.0
    ph = rng.uniform(0, 2 * np.pi, 3)
    field = np.sin((nodes - c) @ k.T + ph)
    axis_angle = field * np.deg2rad(max_angle_deg) / np.sqrt(3)
    k2 = rng.normal(size=(3, 3)) * 3.0
    tr = np.cos((nodes - c) @ k2.T + ph) * max_translation / np.sqrt(3)
    theta = np.linalg.norm(axis_angle, axis=1, keepdims=True) + 1e-12
    kx = axis_angle / theta
    K = np.zeros((len(nodes), 3, 3))
    K[:, 0, 1], K[:, 0, 2], K[:, 1, 0], K[:, 1, 2], K[:, 2, 0], K[:, 2, 1] = -kx[:, 2], kx[:, 1], kx[:, 2], -kx[:, 0], -kx[:, 1], kx[:, 0]
    R = np.eye(3)[None] + np.sin(theta)[..., None] * K + (1 - np.cos(theta))[..., None] * (K @ K)
    return R.astype(np.float32), tr.astype(np.float32)


def make_scene(config="C1", height=480, width=640, rng=None):
    """Returns a dict of numpy inputs for one fit.

    Solver settings note: with the reference's defaults (arap_term_weight 200, no damping) and overlapping
    anchor weights the reference's own iteration diverges on every multi-node graph (block-diagonal data term
    overshoots by ~anchor_count; see DESIGN.md "reference behaviour"). The scenes therefore use concentrated
    anchor weights (node_coverage ~ quarter node spacing), arap_term_weight 20 and lambda 1.0, for which the
    reference iteration is contractive, so 10 free-running iterations stay comparable.
 config: 'C1' plane/16 nodes, 'C1e' plane 12+4 two-layer,
    'C2' bumpy sphere ~250 nodes (208+41), 'tiny' 64x48-friendly sphere for fast CPU tests."""
    rng = rng or np.random.default_rng(SEED)
    K = default_intrinsics(height, width)
    arap, radii = 20.0, None
    if config in ("C1", "C1e"):
        pos, nrm, tris = plane_mesh(33, 1.0, 1.2, rng=rng)
        g = np.linspace(-0.375, 0.375, 4)
        gx, gy = np.meshgrid(g, g)
        fine = np.stack([gx.ravel(), gy.ravel(), np.full(16, 1.2)], 1).astype(np.float32)
        if config == "C1":
            nodes, edges, base = fine, np.zeros((0, 2), np.int32), 16
        else:
            nodes, edges, base = two_layer_graph(fine, 16, 0)[0], None, 16
            coarse = np.array([[-0.25, -0.25, 1.2], [0.25, -0.25, 1.2], [-0.25, 0.25, 1.2], [0.25, 0.25, 1.2]], np.float32)
            d = np.linalg.norm(fine[:, None] - coarse[None], axis=2)
            nn = np.argsort(d, axis=1, kind="stable")[:, :2]
            edges = np.stack([np.repeat(np.arange(16), 2), 16 + nn.ravel()], 1).astype(np.int32)
            nodes = np.concatenate([fine, coarse], 0)
        # node spacing 0.25: with coverage 0.0625 (a quarter of it) the reference's own iteration does not contract on this plane (its
        # block-diagonal data term lets every node explain the whole residual of the vertices it shares: translation updates go
        # 0.035 -> 0.12 -> 0.10 -> 0.19 and the two-layer variant reaches NaN in 7 iterations at 640 x 480, measured with the oracle);
        # at an eighth of the spacing ten iterations reduce the translation error 0.029 -> 0.007 with decaying updates
        coverage = 0.03
        motion = dict(max_angle_deg=3.0, max_translation=0.05)
    elif config == "C2":
        pos, nrm, tris = bumpy_sphere_mesh(6, 0.15, (0, 0, 0.6), rng=rng)
        nodes, edges, base = two_layer_graph(pos.astype(np.float64), 208, 41, 4, rng)
        coverage = 0.025
        motion = dict(max_angle_deg=5.0, max_translation=0.01)
    elif config == "C2r":
        pos, nrm, tris = relief_mesh(203, 0.3, 0.6, 0.02, rng=rng)
        nodes, edges, base = two_layer_graph(pos.astype(np.float64), 208, 41, 4, rng)
        coverage = 0.005
        motion = dict(max_angle_deg=3.0, max_translation=0.005)
        # 640x480 bench scene: the reference's default arap_term_weight with edge weight 0.015 keeps its iteration contractive
        # for 10 iterations (translation error 0.0029 -> 0.0008, residual rms 2.2e-3 -> 1.4e-4; see DESIGN.md "bench scene")
        arap, radii = 200.0, np.array([0.015, 0.03], np.float32)
    elif config == "C3":
        # BASELINE.json configs[2] scale: 4.45 M triangles at 1920 x 1080, ~2000 nodes with a ~300-node coarse layer.
        # A relief sheet of 1492 x 1492 vertices (V = 2 226 064, F = 4 446 162); nodes are mesh vertices on regular grids
        # (41 x 41 = 1681 fine + 17 x 17 = 289 coarse -> a 867-unknown reduced corner), farthest-point sampling being quadratic here.
        n = 1492
        pos, nrm, tris = relief_mesh(n, 0.3, 0.6, 0.02, rng=rng)

        def grid_nodes(side):
            idx = np.linspace(0, n - 1, side).round().astype(np.int64)
            return pos[(idx[:, None] * n + idx[None, :]).ravel()].astype(np.float64)
        fine, coarse = grid_nodes(41), grid_nodes(17) + rng.normal(scale=1e-4, size=(289, 3))
        d = np.linalg.norm(fine[:, None] - coarse[None], axis=2)
        nn = np.argsort(d, axis=1, kind="stable")[:, :4]
        edges = np.stack([np.repeat(np.arange(len(fine)), 4), len(fine) + nn.ravel()], 1).astype(np.int32)
        nodes, base = np.concatenate([fine, coarse], 0).astype(np.float32), len(fine)
        coverage = 0.25 * 0.3 / 41
        motion = dict(max_angle_deg=1.5, max_translation=0.002)
        arap, radii = 200.0 * (height * width / (480.0 * 640.0)) ** 0.5, np.array([0.015, 0.03], np.float32)
    elif config == "tinyr":
        pos, nrm, tris = relief_mesh(41, 0.3, 0.6, 0.02, rng=rng)
        nodes, edges, base = two_layer_graph(pos.astype(np.float64), 24, 6, 3, rng)
        coverage = 0.015
        motion = dict(max_angle_deg=3.0, max_translation=0.005)
    elif config == "tiny":
        pos, nrm, tris = bumpy_sphere_mesh(3, 0.15, (0, 0, 0.6), rng=rng)
        nodes, edges, base = two_layer_graph(pos.astype(np.float64), 24, 6, 3, rng)
        coverage = 0.06
        motion = dict(max_angle_deg=4.0, max_translation=0.008)
    else:
        raise ValueError(config)
    tris = orient_front_faces(pos, nrm, tris, K, height, width)
    R_gt, t_gt = smooth_node_motion(nodes, rng=np.random.default_rng(SEED + 1), **motion)
    E = len(edges)
    return dict(config=config, positions=pos, normals=nrm, triangles=tris, nodes=nodes, edges=edges, arrow_base=base,
                node_coverage=coverage, arap_term_weight=arap, lm_factor=1.0, K=K, height=height, width=width, rotations_gt=R_gt, translations_gt=t_gt,
                edge_layer_indices=np.zeros(max(E, len(nodes), 1), np.int8), layer_radii=np.array([coverage, 2 * coverage], np.float32) if radii is None else radii)
