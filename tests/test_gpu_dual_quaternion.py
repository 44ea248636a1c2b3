"""Dual-quaternion blended warp. The reference has no implementation of it (BASELINE.json names it; cpp/geometry/TransformationMode.h only
selects a storage mode), so there is no reference oracle: the kernel is checked against an independent float64 restatement of Kavan et
al.'s dual-quaternion linear blending built on scipy's quaternion conversions, and through the algorithm's defining properties."""
import numpy as np
import pytest
from scipy.spatial.transform import Rotation

from dynamicfuion_python_b200 import synthetic

pytestmark = pytest.mark.gpu


def _qmul(a, b):  # (w, x, y, z)
    return np.array([a[0] * b[0] - a[1:] @ b[1:], *(a[0] * b[1:] + b[0] * a[1:] + np.cross(a[1:], b[1:]))])


def dqb_numpy(points, normals, nodes, R, t, anchors, weights, extrinsics):
    out_p, out_n = np.zeros_like(points, dtype=np.float64), np.zeros_like(points, dtype=np.float64)
    E = np.asarray(extrinsics, np.float64)
    for i, (p, n) in enumerate(zip(points.astype(np.float64), normals.astype(np.float64))):
        p, n = E[:3, :3] @ p + E[:3, 3], E[:3, :3] @ n
        real, dual, pivot = np.zeros(4), np.zeros(4), None
        for a, w in zip(anchors[i], weights[i].astype(np.float64)):
            if a < 0:
                continue
            x, y, z, s = Rotation.from_matrix(R[a].astype(np.float64)).as_quat()
            q = np.array([s, x, y, z])
            tau = nodes[a] + t[a] - R[a].astype(np.float64) @ nodes[a]
            d = 0.5 * _qmul(np.array([0.0, *tau]), q)
            pivot = q if pivot is None else pivot
            if q @ pivot < 0:
                w = -w
            real, dual = real + w * q, dual + w * d
        if pivot is None:
            continue
        norm = np.linalg.norm(real)
        q, d = real / norm, dual / norm
        Rb = Rotation.from_quat([q[1], q[2], q[3], q[0]]).as_matrix()
        tb = 2.0 * _qmul(d, np.array([q[0], -q[1], -q[2], -q[3]]))[1:]
        out_p[i], out_n[i] = Rb @ p + tb, Rb @ n
    return out_p, out_n


def test_dual_quaternion_warp(nnrt, oracle):
    s = synthetic.make_scene("tinyr", 120, 160)
    rng = np.random.default_rng(13)
    nodes = s["nodes"]
    R = Rotation.from_rotvec(rng.normal(0, 0.8, (len(nodes), 3))).as_matrix().astype(np.float32)  # large, distinct rotations (all quaternion branches)
    R[0] = Rotation.from_rotvec([np.pi * 0.999, 0, 0]).as_matrix()
    R[1] = Rotation.from_rotvec([0, np.pi * 0.999, 0]).as_matrix()
    R[2] = Rotation.from_rotvec([0, 0, -np.pi * 0.999]).as_matrix()
    t = rng.normal(0, 0.02, (len(nodes), 3)).astype(np.float32)
    anchors, weights = oracle.compute_anchors(s["positions"], nodes, 4, 2, 0.03)  # threshold variant: some anchors are -1, some points invalid
    ext = np.eye(4)
    ext[:3, :3] = Rotation.from_rotvec([0.1, -0.2, 0.05]).as_matrix()
    ext[:3, 3] = [0.01, -0.02, 0.03]
    f = nnrt.geometry.functional
    mesh = (s["positions"], s["normals"], s["triangles"])
    got = f.warp_triangle_mesh_dual_quaternion(mesh, nodes, R, t, anchors, weights, 0, ext)
    want_p, want_n = dqb_numpy(s["positions"], s["normals"], nodes, R, t, anchors, weights, ext)
    assert (anchors < 0).any() and (anchors >= 0).all(1).any()
    np.testing.assert_allclose(got.positions.cpu().numpy(), want_p, atol=2e-5)
    np.testing.assert_allclose(got.normals.cpu().numpy(), want_n, atol=2e-5)
    cloud = f.warp_point_cloud_dual_quaternion(s["positions"], nodes, R, t, anchors, weights, 0, ext)
    assert np.array_equal(cloud.cpu().numpy(), got.positions.cpu().numpy())

    # properties: (i) one motion shared by all nodes is reproduced exactly as that rigid motion, whatever the weights (linear blending
    # does too); (ii) lengths of normals are preserved (linear blending of distinct rotations shrinks them); (iii) with a single
    # anchor both blends coincide
    Rg = Rotation.from_rotvec([0.3, 0.2, -0.4]).as_matrix()
    tg = np.array([0.05, -0.01, 0.02])
    Rs = np.repeat(Rg[None], len(nodes), 0).astype(np.float32)
    ts = (Rg @ nodes.T).T - nodes + tg  # R (x - g) + g + t == Rg x + tg for every node
    a0, w0 = oracle.compute_anchors(s["positions"], nodes, 4, 0, 0.03)
    rigid = f.warp_triangle_mesh_dual_quaternion(mesh, nodes, Rs, ts.astype(np.float32), a0, w0)
    np.testing.assert_allclose(rigid.positions.cpu().numpy(), (Rg @ s["positions"].T).T + tg, atol=2e-6)
    dq = f.warp_triangle_mesh_dual_quaternion(mesh, nodes, R, t, a0, w0)
    lb = f.warp_triangle_mesh(mesh, nodes, R, t, a0, w0)
    lengths = np.linalg.norm(s["normals"], axis=1)
    np.testing.assert_allclose(np.linalg.norm(dq.normals.cpu().numpy(), axis=1), lengths, rtol=1e-5)
    assert (np.linalg.norm(lb.normals.cpu().numpy(), axis=1) < 0.99 * lengths).any()
    a1, w1 = oracle.compute_anchors(s["positions"], nodes, 1, 0, 0.03)
    one_dq = f.warp_triangle_mesh_dual_quaternion(mesh, nodes, R, t, a1, w1)
    one_lb = f.warp_triangle_mesh(mesh, nodes, R, t, a1, w1)
    np.testing.assert_allclose(one_dq.positions.cpu().numpy(), one_lb.positions.cpu().numpy(), atol=2e-6)
