"""Shortest-path anchors (nnrt.geometry.functional.compute_anchors_and_weights_shortest_path_*). CPU: the oracle against the reference's
known-answer test (tests/test_compute_anchors.py:57-185, vectors in tests/golden/shortest_path_kat.npz) and an independent heapq
Dijkstra; GPU: bit-exact anchors (slot order included) against the oracle, including tied path lengths and a full queue."""
import heapq
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "shortest_path_kat.npz")


def _np(t):
    return t.detach().cpu().numpy()


def knn_graph(nodes, degree):
    d = np.linalg.norm(nodes[:, None].astype(np.float64) - nodes[None].astype(np.float64), axis=2)
    np.fill_diagonal(d, np.inf)
    return np.argsort(d, axis=1, kind="stable")[:, :degree].astype(np.int32)


def grid_graph(side):
    """side x side lattice with 4-neighbour edges: many equal path lengths."""
    ys, xs = np.meshgrid(np.arange(side), np.arange(side), indexing="ij")
    nodes = np.stack([xs.ravel(), ys.ravel(), np.zeros(side * side)], 1).astype(np.float32)
    edges = -np.ones((side * side, 4), np.int32)
    for y in range(side):
        for x in range(side):
            k = 0
            for dy, dx in ((0, 1), (1, 0), (0, -1), (-1, 0)):
                if 0 <= y + dy < side and 0 <= x + dx < side:
                    edges[y * side + x, k] = (y + dy) * side + x + dx
                    k += 1
    return nodes, edges


def test_oracle_reference_kat(oracle):
    kat = np.load(GOLDEN)
    anchors, weights = oracle.compute_anchors_shortest_path(kat["vertices"], kat["nodes"], kat["edges"], int(kat["anchor_count"]), float(kat["node_coverage"]))
    assert np.array_equal(anchors, kat["anchors_gt"])
    np.testing.assert_allclose(weights.sum(1), 1.0, rtol=1e-6)
    np.testing.assert_allclose(weights, kat["weights_gt"], rtol=1e-5, atol=1e-8)  # np.allclose defaults of the reference's test


def test_oracle_matches_heapq_dijkstra(oracle):
    rng = np.random.default_rng(31)
    nodes = rng.uniform(-1, 1, (120, 3)).astype(np.float32)
    edges = knn_graph(nodes, 4)
    points = rng.uniform(-1, 1, (200, 3)).astype(np.float32)
    anchors, weights = oracle.compute_anchors_shortest_path(points, nodes, edges, 4, 0.3)
    n64 = nodes.astype(np.float64)
    for p, got in zip(points.astype(np.float64), anchors):
        settled = []
        while len(settled) < 4:
            rest = [n for n in range(len(nodes)) if n not in settled]
            seed = min(rest, key=lambda n: np.linalg.norm(n64[n] - p))
            heap = [(np.linalg.norm(n64[seed] - p), seed)]
            while heap and len(settled) < 4:
                dist, n = heapq.heappop(heap)
                if n in settled:
                    continue
                settled.append(n)
                for t in edges[n]:
                    if t > -1:
                        heapq.heappush(heap, (dist + np.linalg.norm(n64[t] - n64[n]), int(t)))
        assert list(got) == settled
    np.testing.assert_allclose(weights.sum(1), 1.0, rtol=1e-5)


@pytest.mark.gpu
def test_reference_kat_on_device(nnrt):
    kat = np.load(GOLDEN)
    anchors, weights = nnrt.geometry.functional.compute_anchors_and_weights_shortest_path_fixed_node_weight(
        kat["vertices"], kat["nodes"], kat["edges"], int(kat["anchor_count"]), float(kat["node_coverage"]))
    assert np.array_equal(_np(anchors), kat["anchors_gt"])
    np.testing.assert_allclose(_np(weights), kat["weights_gt"], rtol=1e-5, atol=1e-8)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["knn4", "lattice_ties", "dense_full_queue", "variable_weights", "image_shaped", "many_nodes"])
def test_bit_exact_against_oracle(nnrt, oracle, case):
    rng = np.random.default_rng(len(case))
    f = nnrt.geometry.functional
    node_weights = None
    if case == "lattice_ties":  # points at cell centres and on nodes: equal seeds and equal path lengths everywhere
        nodes, edges = grid_graph(12)
        points = np.concatenate([nodes[:60] + np.float32([0.5, 0.5, 0.0]), nodes[60:90]], 0).astype(np.float32)
        anchor_count, coverage = 8, 1.5
    elif case == "dense_full_queue":  # degree 8, 8 anchors: up to 57 inserts into the 40-entry queue, the overflow is dropped
        nodes = rng.uniform(-1, 1, (300, 3)).astype(np.float32)
        edges = knn_graph(nodes, 8)
        points = rng.uniform(-1, 1, (3000, 3)).astype(np.float32)
        anchor_count, coverage = 8, 0.4
    elif case == "many_nodes":  # more nodes than the kernel's shared-memory cache: tiled first seeding scan, re-seeds from global memory
        nodes = rng.uniform(-0.5, 0.5, (4000, 3)).astype(np.float32)
        edges = knn_graph(nodes, 3)
        edges[rng.random(edges.shape) < 0.5] = -1  # sparse: components get used up, forcing re-seeds
        points = rng.uniform(-0.5, 0.5, (3000, 3)).astype(np.float32)
        anchor_count, coverage = 6, 0.05
    else:
        nodes = rng.uniform(-0.5, 0.5, (250, 3)).astype(np.float32)
        edges = knn_graph(nodes, 4)
        edges[rng.random(edges.shape) < 0.1] = -1  # missing neighbours, disconnected parts
        points = rng.uniform(-0.5, 0.5, (20000, 3)).astype(np.float32)
        anchor_count, coverage = 4, 0.1
        if case == "variable_weights":
            node_weights = rng.uniform(0.005, 0.05, 250).astype(np.float32)
    if case == "image_shaped":
        points = points[:96 * 128].reshape(96, 128, 3)
    if node_weights is None:
        anchors, weights = f.compute_anchors_and_weights_shortest_path_fixed_node_weight(points, nodes, edges, anchor_count, coverage)
    else:
        anchors, weights = f.compute_anchors_and_weights_shortest_path_variable_node_weight(points, nodes, node_weights, edges, anchor_count)
    want_a, want_w = oracle.compute_anchors_shortest_path(points.reshape(-1, 3), nodes, edges, anchor_count, coverage, node_weights)
    assert tuple(anchors.shape) == points.shape[:-1] + (anchor_count,)
    assert np.array_equal(_np(anchors).reshape(-1, anchor_count), want_a)
    np.testing.assert_allclose(_np(weights).reshape(-1, anchor_count), want_w, rtol=2e-5, atol=1e-7)  # expf: device vs glibc


@pytest.mark.gpu
def test_argument_errors(nnrt):
    f = nnrt.geometry.functional
    nodes = np.zeros((3, 3), np.float32)
    with pytest.raises(RuntimeError):
        f.compute_anchors_and_weights_shortest_path_fixed_node_weight(np.zeros((5, 3), np.float32), nodes, np.zeros((3, 2), np.int64), 2, 0.1)
    with pytest.raises(RuntimeError, match="at least anchor_count"):
        f.compute_anchors_and_weights_shortest_path_fixed_node_weight(np.zeros((5, 3), np.float32), nodes, np.zeros((3, 2), np.int32), 4, 0.1)
    with pytest.raises(RuntimeError):
        f.compute_anchors_and_weights_shortest_path_fixed_node_weight(np.zeros((5, 3), np.float32), nodes, np.zeros((4, 2), np.int32), 2, 0.1)
