"""nnrt.geometry — GraphWarpField (pybind name of nnrt::geometry::WarpField, cpp/pybind/geometry/geometry.cpp:278-316) and
HierarchicalGraphWarpField (cpp/geometry/HierarchicalGraphWarpField.h:35-96; not bound in the reference, added under its natural name)."""
import enum

import torch

from .. import _tensor as T
from . import functional


class AnchorComputationMethod(enum.Enum):
    EUCLIDEAN = 0
    SHORTEST_PATH = 1


class WarpNodeCoverageComputationMethod(enum.Enum):
    FIXED_NODE_COVERAGE = 0
    MINIMAL_K_NEIGHBOR_NODE_DISTANCE = 1


class GraphWarpField:
    """Node store + linear-blend warp. Defaults follow cpp/geometry/WarpField.h:33-41."""

    def __init__(self, nodes, node_coverage=0.05, threshold_nodes_by_distance=False, anchor_count=4, minimum_valid_anchor_count=0,
                 warp_node_coverage_computation_method=WarpNodeCoverageComputationMethod.MINIMAL_K_NEIGHBOR_NODE_DISTANCE,
                 warp_node_coverage_neighbor_count=4):
        self.nodes = T.cuda(nodes, torch.float32)
        if self.nodes.dim() != 2 or self.nodes.shape[1] != 3:
            raise RuntimeError(f"nodes must have shape [N, 3], got {tuple(self.nodes.shape)}")
        n = self.nodes.shape[0]
        if n < anchor_count:  # cpp/geometry/WarpField.cpp:57-61
            raise RuntimeError(f"Anchor count for warp field, {anchor_count}, exceeds node count, which is {n}. Anchors are nodes closest to a "
                               "specific point using some distance metric, hence there cannot be more anchors than nodes.")
        self.node_coverage = float(node_coverage)
        self.threshold_nodes_by_distance_by_default = bool(threshold_nodes_by_distance)
        self.anchor_count = int(anchor_count)
        self.minimum_valid_anchor_count = int(minimum_valid_anchor_count)
        self.warp_node_coverage_computation_method = warp_node_coverage_computation_method
        self.warp_node_coverage_neighbor_count = warp_node_coverage_neighbor_count
        self.node_rotations = torch.eye(3, dtype=torch.float32, device=self.nodes.device).repeat(n, 1, 1).contiguous()
        self.node_translations = torch.zeros((n, 3), dtype=torch.float32, device=self.nodes.device)
        self.node_coverage_weights = None
        if warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod.MINIMAL_K_NEIGHBOR_NODE_DISTANCE:
            self.recompute_node_coverage_weights()

    # --- RecomputeNodeCoverageWeights, cpp/geometry/WarpField.cpp:249-262: squared distance to the nearest other node
    def recompute_node_coverage_weights(self):
        n = self.nodes.shape[0]
        if n == 1:
            self.node_coverage_weights = torch.full((1,), self.node_coverage, dtype=torch.float32, device=self.nodes.device)
        else:
            d = torch.cdist(self.nodes.double(), self.nodes.double())
            d.fill_diagonal_(float("inf"))
            nearest = d.min(dim=1).values.float()
            self.node_coverage_weights = (nearest * nearest).contiguous()

    def get_warped_nodes(self):
        return self.nodes + self.node_translations

    def get_node_extent(self):
        return torch.stack([self.nodes.min(dim=0).values, self.nodes.max(dim=0).values])

    def precompute_anchors_and_weights(self, input_mesh):
        mesh = T.MeshView(input_mesh)
        if self.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE:
            return functional.compute_anchors_and_weights_euclidean_fixed_node_weight(
                mesh.positions, self.nodes, self.anchor_count, self.minimum_valid_anchor_count, self.node_coverage)
        return functional.compute_anchors_and_weights_euclidean_variable_node_weight(
            mesh.positions, self.nodes, self.node_coverage_weights, self.anchor_count, self.minimum_valid_anchor_count)

    def warp_mesh(self, input_mesh, *args, disable_neighbor_thresholding=True, extrinsics=None, anchors=None, anchor_weights=None):
        """warp_mesh(mesh, disable_neighbor_thresholding=True, extrinsics=eye4) or
        warp_mesh(mesh, anchors, anchor_weights, disable_neighbor_thresholding=True, extrinsics=eye4) (cpp/geometry/WarpField.cpp:101-146)."""
        args = list(args)
        if args and not isinstance(args[0], bool):
            anchors, anchor_weights = args[0], args[1]
            args = args[2:]
        if args:
            disable_neighbor_thresholding = args[0]
        if len(args) > 1:
            extrinsics = args[1]
        mesh = T.MeshView(input_mesh)
        if anchors is None:
            anchors, anchor_weights = self.precompute_anchors_and_weights(mesh)
        min_valid = 0 if disable_neighbor_thresholding else self.minimum_valid_anchor_count
        wp, wn = functional._warp(mesh.positions, mesh.normals, self.nodes, self.node_rotations, self.node_translations, anchors,
                                  anchor_weights, min_valid, extrinsics)
        return mesh.with_vertices(wp, wn)

    def clone(self):
        other = GraphWarpField.__new__(GraphWarpField)
        other.__dict__.update(self.__dict__)
        other.node_rotations = self.node_rotations.clone()
        other.node_translations = self.node_translations.clone()
        if self.node_coverage_weights is not None:
            other.node_coverage_weights = self.node_coverage_weights.clone()
        return other

    def reset_rotations(self):
        self.node_rotations.copy_(torch.eye(3, dtype=torch.float32, device=self.nodes.device).expand_as(self.node_rotations))

    def apply_transformations(self):
        """Bakes translations into node positions and resets the motion (WarpField::ApplyTransformations)."""
        out = self.clone()
        out.nodes = (self.nodes + self.node_translations).contiguous()
        out.node_translations = torch.zeros_like(self.node_translations)
        out.reset_rotations()
        return out

    def get_node_rotations(self):
        return self.node_rotations

    def get_node_translations(self):
        return self.node_translations

    def set_node_rotations(self, node_rotations):
        r = T.cuda(node_rotations, torch.float32)
        if r.shape != self.node_rotations.shape:
            raise RuntimeError(f"node_rotations must have shape {tuple(self.node_rotations.shape)}")
        self.node_rotations = r.clone()

    def set_node_translations(self, node_translations):
        t = T.cuda(node_translations, torch.float32)
        if t.shape != self.node_translations.shape:
            raise RuntimeError(f"node_translations must have shape {tuple(self.node_translations.shape)}")
        self.node_translations = t.clone()

    def translate_nodes(self, node_translation_deltas):
        self.node_translations += T.cuda(node_translation_deltas, torch.float32)

    def rotate_nodes(self, node_rotation_deltas):
        from ..core import matmul3d
        self.node_rotations = matmul3d(self.node_rotations, T.cuda(node_rotation_deltas, torch.float32))


class HierarchicalGraphWarpField(GraphWarpField):
    """Warp field with regularisation layers. This build takes the graph in *virtual* node order (layer 0 nodes first, then coarser
    layers; identity node_indices permutation) with explicit `edges` [E,2] (fine node -> coarser node) — the ordering quirks of the
    reference's re-indexing are no-ops in that regime. Automatic layer construction (median-grid subsampling)
    is a "next" row of SURVEY.md §8(f)."""

    def __init__(self, nodes, edges=None, first_layer_node_count=None, node_coverage=0.05, threshold_nodes_by_distance=False,
                 anchor_count=4, minimum_valid_anchor_count=0,
                 warp_node_coverage_computation_method=WarpNodeCoverageComputationMethod.MINIMAL_K_NEIGHBOR_NODE_DISTANCE,
                 warp_node_coverage_neighbor_count=4, edge_layer_indices=None, layer_decimation_radii=None):
        super().__init__(nodes, node_coverage, threshold_nodes_by_distance, anchor_count, minimum_valid_anchor_count,
                         warp_node_coverage_computation_method, warp_node_coverage_neighbor_count)
        n = self.nodes.shape[0]
        self.edges = torch.zeros((0, 2), dtype=torch.int32, device=self.nodes.device) if edges is None else T.cuda(edges, torch.int32)
        e = self.edges.shape[0]
        self.first_layer_node_count = n if first_layer_node_count is None else int(first_layer_node_count)
        if edge_layer_indices is None:
            edge_layer_indices = torch.zeros((max(e, n),), dtype=torch.int8)
        self.edge_layer_indices = T.cuda(edge_layer_indices, torch.int8)
        if layer_decimation_radii is None:
            layer_decimation_radii = [self.node_coverage * (i + 1) for i in range(2)]  # radius(i) = (i+1) * node_coverage
        self.layer_decimation_radii = T.cuda(torch.as_tensor(layer_decimation_radii, dtype=torch.float32), torch.float32)

    def get_edges(self):
        return self.edges

    def clone(self):
        other = super().clone()
        other.__class__ = HierarchicalGraphWarpField
        return other
