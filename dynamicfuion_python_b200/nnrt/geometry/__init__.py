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
            self.node_coverage_weights = T.empty((n,), torch.float32)
            T.call("nnrt_b200_node_coverage_weights", T.ptr(self.nodes), n, self.node_coverage, T.ptr(self.node_coverage_weights), T.stream())

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


class RegularizationLayer:
    """cpp/geometry/HierarchicalGraphWarpField.h:27-32: one layer of the hierarchy. node_indices are ORIGINAL node indices; edges
    ([E_l, 2] int32, virtual indices, this layer's nodes -> the next coarser layer) is None for the coarsest layer."""

    def __init__(self, decimation_radius, node_indices, node_positions, edges):
        self.decimation_radius = decimation_radius
        self.node_indices = node_indices
        self.node_positions = node_positions
        self.edges = edges


class HierarchicalGraphWarpField(GraphWarpField):
    """Warp field with regularisation layers (cpp/geometry/HierarchicalGraphWarpField.h:35-96, .cpp:36-195): the constructor
    median-grid-subsamples the nodes into `layer_count` layers and connects every node of a layer to its `max_vertex_degree`
    nearest nodes of the next coarser layer, on the device (nnrt_b200_build_regularization_layers).

    Storage is in VIRTUAL order (layer 0 first, then coarser layers; `node_indices` maps virtual -> original index), because that is
    the order the fitter and the edge list work in; the reference keeps original order and re-gathers on every access. The
    accessors take the reference's `use_virtual_ordering` flag and default to the original order. The reference's own
    inconsistencies between the two orders are not reproduced by default: both orders always describe the
    same field. Setting `reference_ordering_quirks = True` reproduces them for callers that depend on the
    reference's exact behaviour with more than one layer (they are no-ops when the order is the identity):
      * rotate_nodes(deltas, use_virtual_ordering=True) multiplies the ORIGINAL-order rotations by the virtual-order deltas, row by
        row, and scatters the products through node_indices (HierarchicalGraphWarpField.cpp:271-277);
      * get_node_coverage_weights(use_virtual_ordering=False) returns the node rotations (.cpp:222-224);
      * precompute_anchors_and_weights(mesh, use_virtual_ordering=True) with MINIMAL_K_NEIGHBOR_NODE_DISTANCE pairs virtual-order
        node positions with ORIGINAL-order coverage weights (.cpp:249-253);
      * warp_mesh always reads its anchors as original-order indices, also those computed in virtual order (WarpField.cpp:129-144).
    `from_graph` takes an explicit graph that is already in virtual order."""

    reference_ordering_quirks = False

    def __init__(self, nodes, node_coverage=0.05, threshold_nodes_by_distance_by_default=False, anchor_count=4, minimum_valid_anchor_count=0,
                 warp_node_coverage_computation_method=WarpNodeCoverageComputationMethod.MINIMAL_K_NEIGHBOR_NODE_DISTANCE,
                 layer_count=4, max_vertex_degree=4, compute_layer_decimation_radius=None):
        import ctypes as C
        import numpy as np
        super().__init__(nodes, node_coverage, threshold_nodes_by_distance_by_default, anchor_count, minimum_valid_anchor_count,
                         warp_node_coverage_computation_method)
        if layer_count < 1:  # HierarchicalGraphWarpField.cpp:59-61
            raise RuntimeError(f"layer_count must be a positive integer, got {layer_count}.")
        coverage = np.float32(self.node_coverage)
        radius = compute_layer_decimation_radius or (lambda i_layer, c: np.float32(i_layer + 1) * np.float32(c))  # .h:46-47
        radii = np.array([coverage] + [np.float32(radius(i, coverage)) for i in range(1, layer_count)], np.float32)
        n, degree = self.nodes.shape[0], int(max_vertex_degree)
        node_indices = T.empty((n,), torch.int64)
        edges = T.empty((max(n * degree, 1), 2), torch.int32)
        edge_layers = T.empty((max(n * degree, 1),), torch.int8)
        counts = (C.c_int32 * layer_count)()
        edge_count = C.c_int64()
        T.call("nnrt_b200_build_regularization_layers", T.ptr(self.nodes), n, radii.ctypes.data, int(layer_count), degree, T.ptr(node_indices),
               counts, T.ptr(edges), T.ptr(edge_layers), C.byref(edge_count), T.stream())
        e = edge_count.value
        self.layer_node_counts = list(counts)
        self.node_indices = node_indices
        original_nodes = self.nodes
        self._set_order(node_indices, identity=layer_count == 1)
        self.nodes = original_nodes[node_indices].contiguous()
        if self.node_coverage_weights is not None:
            self.node_coverage_weights = self.node_coverage_weights[node_indices].contiguous()
        self.edges = edges[:e].contiguous()
        # the reference's ARAP weight reads edge_layer_indices at NODE positions, so the tensor is kept at least
        # node_count long (zero padded) instead of letting that read run off the end
        self.edge_layer_indices = torch.zeros((max(e, n),), dtype=torch.int8, device=self.nodes.device)
        self.edge_layer_indices[:e] = edge_layers[:e]
        self.first_layer_node_count = self.layer_node_counts[0]
        self.layer_decimation_radii = T.cuda(torch.from_numpy(radii), torch.float32)
        self.regularization_layers = []
        first, edge_first = 0, e
        for i_layer, count in enumerate(self.layer_node_counts):
            layer_edges = None
            if i_layer < layer_count - 1:  # edges are stored coarsest pair first
                layer_edge_count = count * min(degree, self.layer_node_counts[i_layer + 1])
                edge_first -= layer_edge_count
                layer_edges = self.edges[edge_first:edge_first + layer_edge_count]
            self.regularization_layers.append(RegularizationLayer(float(radii[i_layer]), node_indices[first:first + count],
                                                                  self.nodes[first:first + count], layer_edges))
            first += count

    def _set_order(self, node_indices, identity):
        """`identity`: the virtual order is the original order (one layer, or an explicit graph); with more layers the sampled nodes
        move behind the finer layer and the accessors gather."""
        self._identity_order = bool(identity)
        if identity:
            self._original_to_virtual = node_indices
        else:
            self._original_to_virtual = torch.empty_like(node_indices)
            self._original_to_virtual[node_indices] = torch.arange(node_indices.shape[0], device=node_indices.device)

    @classmethod
    def from_graph(cls, nodes, edges=None, first_layer_node_count=None, node_coverage=0.05, threshold_nodes_by_distance=False,
                   anchor_count=4, minimum_valid_anchor_count=0,
                   warp_node_coverage_computation_method=WarpNodeCoverageComputationMethod.MINIMAL_K_NEIGHBOR_NODE_DISTANCE,
                   warp_node_coverage_neighbor_count=4, edge_layer_indices=None, layer_decimation_radii=None):
        """An explicit graph in virtual node order (identity node_indices): layer-0 nodes first, `edges` [E,2] from a layer-0 node to a
        coarser node."""
        self = cls.__new__(cls)
        GraphWarpField.__init__(self, nodes, node_coverage, threshold_nodes_by_distance, anchor_count, minimum_valid_anchor_count,
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
        self.node_indices = torch.arange(n, device=self.nodes.device)
        self._set_order(self.node_indices, identity=True)
        self.layer_node_counts = [self.first_layer_node_count, n - self.first_layer_node_count] if self.first_layer_node_count < n else [n]
        self.regularization_layers = None
        return self

    # --- order handling ----------------------------------------------------------------------------------------------------------
    def _to_original(self, virtual_tensor, use_virtual_ordering):
        if use_virtual_ordering or self._identity_order:
            return virtual_tensor
        return virtual_tensor[self._original_to_virtual]

    def _to_virtual(self, tensor, use_virtual_ordering):
        if use_virtual_ordering or self._identity_order:
            return tensor
        return tensor[self.node_indices]

    def get_regularization_level(self, i_layer):
        return self.regularization_layers[i_layer]

    def get_regularization_level_count(self):
        return len(self.layer_node_counts)

    def get_node_positions(self, use_virtual_ordering=False):
        return self._to_original(self.nodes, use_virtual_ordering)

    def get_node_rotations(self, use_virtual_ordering=False):
        return self._to_original(self.node_rotations, use_virtual_ordering)

    def get_node_translations(self, use_virtual_ordering=False):
        return self._to_original(self.node_translations, use_virtual_ordering)

    def get_node_coverage_weights(self, use_virtual_ordering=False):
        if self.reference_ordering_quirks and not use_virtual_ordering:
            return self._to_original(self.node_rotations, False)  # the reference returns node_rotations here (.cpp:222-224)
        return self._to_original(self.node_coverage_weights, use_virtual_ordering)

    def get_warped_nodes(self, use_virtual_ordering=False):
        return self._to_original(self.nodes + self.node_translations, use_virtual_ordering)

    def get_edges(self):
        return self.edges

    def get_virtual_node_indices(self):
        return self.node_indices

    def get_edge_layer_indices(self):
        return self.edge_layer_indices

    def get_layer_decimation_radii(self):
        return self.layer_decimation_radii

    def set_node_rotations(self, node_rotations, use_virtual_ordering=False):
        super().set_node_rotations(self._to_virtual(T.cuda(node_rotations, torch.float32), use_virtual_ordering))

    def set_node_translations(self, node_translations, use_virtual_ordering=False):
        super().set_node_translations(self._to_virtual(T.cuda(node_translations, torch.float32), use_virtual_ordering))

    def translate_nodes(self, node_translation_deltas, use_virtual_ordering=False):
        super().translate_nodes(self._to_virtual(T.cuda(node_translation_deltas, torch.float32), use_virtual_ordering))

    def rotate_nodes(self, node_rotation_deltas, use_virtual_ordering=False):
        deltas = T.cuda(node_rotation_deltas, torch.float32)
        if self.reference_ordering_quirks and use_virtual_ordering and not self._identity_order:
            # new[i] = R_original[i] deltas_virtual[i], scattered to original index node_indices[i], i.e. virtual slot i (.cpp:271-277)
            from ..core import matmul3d
            self.node_rotations = matmul3d(self._to_original(self.node_rotations, False).contiguous(), deltas.contiguous())
            return
        super().rotate_nodes(self._to_virtual(deltas, use_virtual_ordering).contiguous())

    def precompute_anchors_and_weights(self, input_mesh, use_virtual_ordering=False):
        """Anchors index the virtual order if `use_virtual_ordering` (what the fitter asks for, DeformableMeshToImageFitter.cpp:96-97),
        else the original order (HierarchicalGraphWarpField.cpp:228-256)."""
        quirk = (self.reference_ordering_quirks and use_virtual_ordering and not self._identity_order
                 and self.warp_node_coverage_computation_method != WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE)
        if quirk:  # virtual-order positions against original-order coverage weights (.cpp:249-253)
            return functional.compute_anchors_and_weights_euclidean_variable_node_weight(
                T.MeshView(input_mesh).positions, self.nodes, self._to_original(self.node_coverage_weights, False).contiguous(), self.anchor_count,
                self.minimum_valid_anchor_count)
        if self.reference_ordering_quirks and not use_virtual_ordering and not self._identity_order:
            # the reference searches the original-order node array: same anchor sets as below, but in ITS slot order (the K-nearest
            # search keeps candidates in the order of the scan, KnnUtilities.h), which the blending sums follow
            positions, nodes = T.MeshView(input_mesh).positions, self._to_original(self.nodes, False).contiguous()
            if self.warp_node_coverage_computation_method == WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE:
                return functional.compute_anchors_and_weights_euclidean_fixed_node_weight(positions, nodes, self.anchor_count,
                                                                                          self.minimum_valid_anchor_count, self.node_coverage)
            return functional.compute_anchors_and_weights_euclidean_variable_node_weight(
                positions, nodes, self._to_original(self.node_coverage_weights, False).contiguous(), self.anchor_count, self.minimum_valid_anchor_count)
        anchors, weights = super().precompute_anchors_and_weights(input_mesh)
        if use_virtual_ordering or self._identity_order:
            return anchors, weights
        mapped = self.node_indices.to(torch.int32)[anchors.clamp(min=0).long()]
        return torch.where(anchors >= 0, mapped, anchors), weights

    def warp_mesh(self, input_mesh, *args, use_virtual_ordering=False, **kwargs):
        """As GraphWarpField.warp_mesh; supplied anchors index the original order unless `use_virtual_ordering`."""
        args = list(args)
        if args and not isinstance(args[0], bool):
            kwargs["anchors"], kwargs["anchor_weights"] = args[0], args[1]
            args = args[2:]
        anchors = kwargs.get("anchors")
        if self.reference_ordering_quirks:
            use_virtual_ordering = False  # WarpField::WarpMesh has no such flag: anchors always index the original order
        if anchors is not None and not (use_virtual_ordering or self._identity_order):
            anchors = T.cuda(anchors, torch.int32)
            mapped = self._original_to_virtual.to(torch.int32)[anchors.clamp(min=0).long()]
            kwargs["anchors"] = torch.where(anchors >= 0, mapped, anchors)
        return super().warp_mesh(input_mesh, *args, **kwargs)

    def clone(self):
        other = super().clone()
        other.__class__ = HierarchicalGraphWarpField
        return other


from .voxel_block_grid import BlockHashMap, NonRigidSurfaceVoxelBlockGrid, VoxelBlockGrid  # noqa: E402
