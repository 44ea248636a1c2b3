"""CPU: the oracle's non-rigid TSDF integration against a float64 numpy restatement in the one regime with a closed form — a single
node with identity motion and one anchor (the warp is the identity), flat depth, no colour."""
import numpy as np


def test_oracle_integrate_identity_warp_matches_numpy(oracle):
    H, W, R, voxel = 48, 64, 8, 0.01
    K = np.array([[60.0, 0, 32.0], [0, 60.0, 24.0], [0, 0, 1]])
    depth = np.full((H, W), 0.5, np.float32)
    depth[:, :8] = 0.0  # a band without measurements
    normals = np.zeros((H, W, 3), np.float32)
    normals[..., 2] = 1.0  # cosine = view . normal = -1 head on: the reference integrates only where it is <= 0.5
    keys = np.array([[0, 0, 5], [-1, 0, 6], [0, -1, 5], [3, 3, 3]], np.int32)  # the last block is not listed as active
    active = np.array([0, 1, 2], np.int64)
    rng = np.random.default_rng(2)
    tsdf = rng.uniform(-1, 1, (4, R, R, R)).astype(np.float32)
    weight = rng.integers(0, 4, (4, R, R, R)).astype(np.float32)
    nodes = np.array([[0.0, 0.0, 0.5]], np.float32)
    got = tsdf.copy()
    cos = oracle.integrate_non_rigid(active, keys, got, weight, None, R, voxel, 3 * voxel, depth, None, normals, K, K, np.eye(4), nodes,
                                     np.eye(3, dtype=np.float32)[None], np.zeros((1, 3), np.float32), 1.0, 1, 0, 1.0, 3.0)
    want = tsdf.astype(np.float64).copy()
    safe = np.zeros(tsdf.shape, bool)  # voxels whose gates are not decided by the last bits of the projection
    cos_want = np.zeros((H, W))
    trunc = 3 * voxel
    for b in active:
        zl, yl, xl = np.meshgrid(np.arange(R), np.arange(R), np.arange(R), indexing="ij")  # x fastest in memory
        p = (keys[b][None, None, None, :] * R + np.stack([xl, yl, zl], -1)) * voxel
        u, v = K[0, 0] * p[..., 0] / p[..., 2] + K[0, 2], K[1, 1] * p[..., 1] / p[..., 2] + K[1, 2]
        inside = (u >= 0) & (v >= 0) & (u <= W - 1) & (v <= H - 1)
        ui, vi = np.floor(np.abs(u) + 0.5).astype(int) * np.sign(u).astype(int), np.floor(np.abs(v) + 0.5).astype(int) * np.sign(v).astype(int)
        frac_u, frac_v = np.abs(u - np.floor(u) - 0.5), np.abs(v - np.floor(v) - 0.5)
        margin = (u > 1e-3) & (v > 1e-3) & (u < W - 1 - 1e-3) & (v < H - 1 - 1e-3)
        safe[b] = (frac_u > 1e-3) & (frac_v > 1e-3) & (margin | ~((u > -1e-3) & (v > -1e-3) & (u < W - 1 + 1e-3) & (v < H - 1 + 1e-3)))
        safe[b] &= np.abs(0.5 - p[..., 2] + trunc) > 1e-5
        for idx in np.argwhere(inside):
            z_, y_, x_ = idx
            d = depth[vi[z_, y_, x_], ui[z_, y_, x_]]
            if d <= 0:
                continue
            view = -p[z_, y_, x_] / np.linalg.norm(p[z_, y_, x_])
            cosine = view @ normals[vi[z_, y_, x_], ui[z_, y_, x_]]
            cos_want[vi[z_, y_, x_], ui[z_, y_, x_]] = cosine
            psdf = d - p[z_, y_, x_, 2]
            if psdf <= -trunc or cosine > 0.5:
                continue
            wgt = weight[b, z_, y_, x_]
            want[b, z_, y_, x_] = (wgt * want[b, z_, y_, x_] + min(psdf, trunc) / trunc) / (wgt + 1)
    assert (want != tsdf).sum() > 300
    assert np.array_equal(got[3], tsdf[3])
    assert safe[:3].mean() > 0.8
    assert np.abs(got - want)[safe].max() < 1e-4
    assert (np.abs(cos - cos_want) > 1e-5).mean() < 0.02  # a pixel's last writer can be an unsafe voxel
    assert not cos[:, :8].any()
